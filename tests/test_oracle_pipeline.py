"""CPU test of the whole stage chain on the two oracles: compiled reference stage functions against the C port,
with and without intra coding units in the picture."""
import numpy as np
import pytest

from ffmpeg_b200 import synth
from oracle import refapi
from tests import pipeline


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed,w,h,ctb,lmcs,frac", [(1, 416, 240, 7, False, 0.0), (2, 256, 128, 5, True, 0.0),
                                                    (21, 416, 240, 7, False, 0.3), (22, 416, 240, 6, True, 0.5),
                                                    (23, 208, 120, 5, False, 1.0)])
def test_port_chain_matches_reference_chain(seed, w, h, ctb, lmcs, frac, monkeypatch):
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, intra_fraction=frac)
    want = pipeline.oracle_picture(pic)
    monkeypatch.setattr(refapi, "available", lambda: False)
    got = pipeline.oracle_picture(pic)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed,w,h,ctb,lmcs,frac", [(41, 256, 128, 6, False, 0.0), (42, 256, 128, 5, True, 0.0),
                                                    (43, 208, 120, 6, True, 0.4), (44, 208, 120, 7, False, 1.0)])
def test_port_chain_matches_reference_chain_12_bit(seed, w, h, ctb, lmcs, frac, monkeypatch):
    """The same chain at bit depth 12 (the reference instantiates its templates for 8, 10 and 12 bits, vvcdsp.c:228-253;
    planes stay 16-bit containers)."""
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, lmcs=lmcs, intra_fraction=frac, bd=12)
    want = pipeline.oracle_picture(pic)
    assert max(int(p.max()) for p in want) > 1023          # really 12-bit content
    monkeypatch.setattr(refapi, "available", lambda: False)
    got = pipeline.oracle_picture(pic)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
