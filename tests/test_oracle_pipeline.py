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
