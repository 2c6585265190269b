"""CPU tests: the oracle's C port of SAO / ALF (+CC-ALF) / LMCS against the compiled reference's own stage
functions (ff_vvc_sao_filter, ff_vvc_alf_filter, ff_vvc_lmcs_filter run CTB by CTB with their saved-border
buffers) and against committed golden digests.  Parameter ranges follow tests/checkasm/vvc_sao.c:37 and
tests/checkasm/vvc_alf.c:52-79; slice boundaries with pps_loop_filter_across_slices_enabled_flag = 0 exercise
edge_restore_1 and the ALF edge clamps, which checkasm does not cover."""
import os

import numpy as np
import pytest

from ffmpeg_b200 import synth
from oracle import portapi, refapi
from tests import util

needs_ref = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
CASES = [(7, 1, 1), (6, 3, 0), (5, 4, 0), (7, 2, 0), (6, 1, 1), (5, 9, 0)]


@needs_ref
@pytest.mark.parametrize("i,case", list(enumerate(CASES)))
def test_port_sao_matches_reference(i, case):
    ctb, ns, lfase = case
    lf = synth.loopfilter_ctbs(400 + i, 208, 136, ctb_log2=ctb, n_slices=ns, lfase=lfase)
    base = synth.smooth_planes(i, lf["dims"])
    a = [p.copy() for p in base]
    refapi.sao_picture(a, lf)
    b = [np.zeros_like(p) for p in base]
    portapi.sao_picture(b, base, lf)
    assert sum(int((x != y).sum()) for x, y in zip(a, base)) > 1000
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
@pytest.mark.parametrize("i,case", list(enumerate(CASES)))
def test_port_alf_matches_reference(i, case):
    ctb, ns, lfase = case
    lf = synth.loopfilter_ctbs(500 + i, 208, 136, ctb_log2=ctb, n_slices=ns, lfase=lfase)
    base = synth.smooth_planes(i, lf["dims"])
    a = [p.copy() for p in base]
    refapi.alf_picture(a, lf)
    b = [np.zeros_like(p) for p in base]
    portapi.alf_picture(b, base, lf)
    assert sum(int((x != y).sum()) for x, y in zip(a, base)) > 1000
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
def test_port_alf_virtual_boundary_on_ctb_multiple_height():
    """Height a multiple of the CTB size: the virtual boundary at ctb_size - 4 also acts in the last CTB row."""
    lf = synth.loopfilter_ctbs(77, 128, 128, ctb_log2=6, alf_on=1.0)
    base = synth.smooth_planes(77, lf["dims"])
    a = [p.copy() for p in base]
    refapi.alf_picture(a, lf)
    b = [np.zeros_like(p) for p in base]
    portapi.alf_picture(b, base, lf)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
def test_port_lmcs_matches_reference():
    _, inv = synth.lmcs_luts(3)
    base = util.noise_planes(3, synth.plane_dims(200, 72))
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    refapi.lmcs_picture(a, inv)
    portapi.lmcs_picture(b, inv)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], base[1])


def test_port_loopfilter_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "loopfilter.npz"))
    for name in ("a", "b"):
        seed, w, h, ctb, ns, lfase = (int(v) for v in g[name + "_args"])
        lf = synth.loopfilter_ctbs(seed, w, h, ctb_log2=ctb, n_slices=ns, lfase=lfase)
        base = synth.smooth_planes(seed, lf["dims"])
        s = [np.zeros_like(p) for p in base]
        portapi.sao_picture(s, base, lf)
        assert util.plane_md5(s) == list(g[name + "_sao_md5"])
        o = [np.zeros_like(p) for p in base]
        portapi.alf_picture(o, s, lf)
        assert util.plane_md5(o) == list(g[name + "_alf_md5"])


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("lfate", [0, 1])
def test_port_matches_reference_on_a_tile_grid(lfate):
    """2x2 tiles, two slices: the flags synth derives on the host (what a real emitter hands the engine) against the
    reference's own derivation from ctb_to_col_bd / ctb_to_row_bd (vvc_filter.c:163-212, 1264-1280)."""
    lf = synth.loopfilter_ctbs(77, 416, 240, ctb_log2=5, n_slices=2, lfase=0, tile_cols=(0, 6), tile_rows=(0, 4), lfate=lfate)
    base = synth.smooth_planes(7, lf["dims"])
    a = [p.copy() for p in base]
    refapi.sao_picture(a, lf)
    b = [np.zeros_like(p) for p in base]
    portapi.sao_picture(b, base, lf)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    c = [p.copy() for p in a]
    refapi.alf_picture(c, lf)
    d = [np.zeros_like(p) for p in a]
    portapi.alf_picture(d, a, lf)
    assert all(np.array_equal(x, y) for x, y in zip(c, d))
