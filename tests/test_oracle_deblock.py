"""CPU tests: the oracle's C port of deblocking against the compiled reference's CTU-ordered
ff_vvc_deblock_vertical/horizontal (oracle/_ref) and against committed golden digests; plus the host-side
edge packer.  The reference has no checkasm test for deblocking, so shapes come from the
driver's own rules (vvc_filter.c:374-397: lengths 1/2/3/5/7, luma grid 4, chroma grid 8)."""
import os

import numpy as np
import pytest

from ffmpeg_b200 import lib, synth
from oracle import portapi, refapi
from tests import util

needs_ref = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")


def port_deblock(planes, t):
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], t["params"].ctb_log2_size)
    portapi.deblock_picture(planes, edge, t["qp"], t["ctb_offsets"], t["stride4"], t["params"])
    return edge


@needs_ref
@pytest.mark.parametrize("seed,ctb_log2", [(0, 7), (1, 7), (2, 6), (3, 5), (4, 7), (5, 6)])
def test_port_deblock_matches_reference(seed, ctb_log2):
    t = synth.deblock_picture(300 + seed, 208, 136, ctb_log2=ctb_log2)
    base = synth.smooth_planes(seed, t["dims"])
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    refapi.deblock_picture(a, t)
    port_deblock(b, t)
    assert sum(int((x != y).sum()) for x, y in zip(a, base)) > 1000      # the filter did something
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_packer_chroma_lengths():
    t = synth.deblock_picture(7, 128, 128)
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], 7)
    assert np.array_equal(edge[:, 0] & 3, t["bs"][:, 0])
    assert np.array_equal((edge[:, 0] >> 2) & 7, t["len_p"])
    assert np.array_equal(edge[:, 0] >> 5, t["len_q"])
    lp, lq, b = (edge[:, 1:] >> 2) & 7, edge[:, 1:] >> 5, edge[:, 1:] & 3
    assert set(np.unique(lq)) <= {0, 1, 3} and set(np.unique(lp)) <= {0, 1, 3}
    assert np.all((lq == 3) | (lq == (b == 2)))
    # horizontal edges on a CTU row boundary never use more than one P-side sample (vvc_filter.c:806-807)
    assert np.all(lp[1][:, ::32, :] <= 1)


def test_port_deblock_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "deblock.npz"))
    for name in ("a", "b"):
        seed, w, h, ctb = (int(v) for v in g[name + "_args"])
        t = synth.deblock_picture(seed, w, h, ctb_log2=ctb)
        planes = synth.smooth_planes(seed, t["dims"])
        port_deblock(planes, t)
        assert util.plane_md5(planes) == list(g[name + "_plane_md5"])


def _norm(e):
    """Edge bytes with bs == 0 carry no information (the filter skips them)."""
    return np.where((e & 3) == 0, 0, e)


@pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("seed,w,h,ctb,frac,ns", [(40, 416, 240, 6, 0.3, 1), (41, 256, 192, 7, 0.5, 1), (42, 208, 120, 5, 0.1, 1),
                                                 (43, 416, 240, 5, 0.3, 3)])
def test_bs_port_matches_the_references_vvc_deblock_bs(seed, w, h, ctb, frac, ns):
    """boundary strengths and filter lengths from a CONSISTENT picture state (the CU / TU / MV tables of a
    synthetic picture with affine, GPM, CIIP, ISP and intra units) — the reference's own vvc_deblock_bs over fc->tab.*
    against the restatement over the engine's compact per-4x4 tables."""
    from oracle import port_bs
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, intra_fraction=frac)
    st = synth.deblock_state(pic)
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    slices = synth.slice_map(np.random.default_rng(seed), cw, ch, ns)
    lfase = 1 if ns == 1 else 0
    bs, lp, lq = refapi.deblock_bs(st, w, h, ctb, slices, lfase)
    assert (bs == 2).any() and (bs == 1).any() and (lp == 7).any() and (lp == 2).any()
    want = lib.pack_deblock_edges(bs, lp, lq, np.ascontiguousarray(st["tb_width"][1]), np.ascontiguousarray(st["tb_height"][1]), ctb)
    edges = (synth.alf_edge_flags(slices, cw, ch, lfase) & 3).astype(np.uint8)
    got = port_bs.derive_edges(st["motion"], st["part"], w, h, ctb, edges)
    assert np.array_equal(_norm(got), _norm(want))
