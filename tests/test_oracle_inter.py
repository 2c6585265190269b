"""CPU tests: the oracle's C port of translational inter prediction against the compiled reference's own
predict_inter() (oracle/_ref) and against committed golden digests.  Block shapes, fractions and weights
follow tests/checkasm/vvc_mc.c:68-324 (all widths/heights, random mx/my, random denom/w/o)."""
import os

import numpy as np
import pytest

from ffmpeg_b200 import abi, synth
from oracle import portapi, refapi
from tests import util

needs_ref = pytest.mark.skipif(not refapi.available(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("seed", range(4))
def test_port_inter_matches_reference(seed):
    pic = synth.inter_picture(200 + seed, 208, 120, n_refs=3)
    base = util.noise_planes(seed, pic["dims"])
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    refapi.inter_batch(a, pic["refs"], pic["blocks"], pic["params"])
    portapi.inter_batch(b, pic["refs"], pic["blocks"], pic["params"])
    assert not np.array_equal(a[0], base[0])
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
def test_port_inter_lmcs_and_far_vectors():
    pic = synth.inter_picture(77, 128, 64, far_mv=0.5)
    pic["blocks"]["flags"] |= abi.PB_LMCS
    lut = np.sort(np.random.default_rng(5).integers(0, 1024, 1024)).astype(np.uint16)
    base = util.noise_planes(3, pic["dims"])
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    refapi.inter_batch(a, pic["refs"], pic["blocks"], pic["params"], lut)
    portapi.inter_batch(b, pic["refs"], pic["blocks"], pic["params"], lut)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
def test_port_put_matches_reference_put():
    """inter.put (14-bit intermediate): port's tmp14 view against the table entry itself, every width."""
    import ctypes as C
    rng = np.random.default_rng(9)
    dims = synth.plane_dims(64, 64)
    ref = util.noise_planes(4, dims)
    for w in (4, 8, 16):
        for h in (4, 8, 16):
            for _ in range(6):
                mvx, mvy = int(rng.integers(-64, 64)), int(rng.integers(-64, 64))
                b = np.zeros(1, abi.PB_DTYPE)
                b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"] = 16, 16, w, h, 1
                b["mv"][0, 0] = (mvx, mvy)
                t14 = [np.zeros((hh, ww), np.int16) for (ww, hh) in dims]
                portapi.inter_batch([p.copy() for p in ref], [ref], b, abi.inter_params(), tmp14=t14)
                for c, (is_c, sh) in enumerate(((0, 0), (1, 1), (1, 1))):
                    pad = np.pad(ref[c], 16, mode="edge")
                    xi, yi = (16 >> sh) + (mvx >> (4 + sh)), (16 >> sh) + (mvy >> (4 + sh))
                    mx, my = (mvx & 31, mvy & 31) if is_c else (mvx & 15, mvy & 15)
                    out = np.zeros((16, 128), np.int16)
                    refapi.lib().vvcref_put(10, is_c, C.c_void_p(out.ctypes.data),
                                            C.c_void_p(pad.ctypes.data + (yi + 16) * pad.strides[0] + (xi + 16) * 2),
                                            C.c_ssize_t(pad.strides[0]), w >> sh, h >> sh, mx, my, 0)
                    got = t14[c][16 >> sh:(16 >> sh) + (h >> sh), 16 >> sh:(16 >> sh) + (w >> sh)]
                    assert np.array_equal(got, out[:h >> sh, :w >> sh]), (w, h, mvx, mvy, c)


def test_port_inter_matches_golden():
    g = np.load(os.path.join(util.GOLDEN, "inter.npz"))
    for name in ("a", "b"):
        seed, w, h, n_refs = (int(v) for v in g[name + "_args"])
        pic = synth.inter_picture(seed, w, h, n_refs=n_refs)
        planes = util.noise_planes(seed, pic["dims"])
        portapi.inter_batch(planes, pic["refs"], pic["blocks"], pic["params"])
        assert util.plane_md5(planes) == list(g[name + "_plane_md5"])


@needs_ref
@pytest.mark.parametrize("seed", range(6))
def test_port_dmvr_bdof_match_reference(seed):
    """DMVR (bilinear search + parametric refinement, window-confined final MC) and BDOF through the reference's own
    predict_inter(); seeds 4-5 use identical references with mirrored vectors to hit the early exits
    (vvc_inter.c:716, 745-747).  The refined vectors (tab_dmvr_mvf) must match too."""
    tools = ("bi", "wp", "bcw", "hpel") + (("dmvr",) if seed % 3 != 1 else ()) + (("bdof",) if seed % 3 != 0 else ())
    pic = synth.inter_picture(300 + seed, 208, 120, tools=tools, sizes=(8, 16), mv_range=(64 * 16 if seed < 4 else 24))
    if seed >= 4:
        pic["refs"][1] = [p.copy() for p in pic["refs"][0]]
        pic["blocks"]["mv"][:, 1, :] = -pic["blocks"]["mv"][:, 0, :]
    base = util.noise_planes(seed, pic["dims"])
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    mva = refapi.inter_batch(a, pic["refs"], pic["blocks"], pic["params"], want_refined_mv=True)
    mvb = portapi.inter_batch(b, pic["refs"], pic["blocks"], pic["params"], want_refined_mv=True)
    assert np.array_equal(mva, mvb)
    if "dmvr" in tools and seed < 4:
        assert (mva != pic["blocks"]["mv"]).any()
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


@needs_ref
@pytest.mark.parametrize("seed,kinds", [(0, (0, 2)), (1, (0, 1)), (2, (0, 1, 2)), (3, (2,)), (4, (1,))])
def test_port_affine_prof_gpm_match_reference(seed, kinds):
    """Affine (4x4 sub-blocks, filter set 2, PROF, averaged chroma vectors) and GPM coding units: the reference expands
    CU-level descriptions itself (pred_affine_blk / pred_gpm_blk); the port consumes the host-expanded records."""
    pic = synth.inter_cu_picture(600 + seed, 208, 120, kinds=kinds)
    base = util.noise_planes(seed, pic["dims"])
    a, b = [p.copy() for p in base], [p.copy() for p in base]
    refapi.inter_cus(a, pic["refs"], pic["cus"], pic["sb_mv"], pic["params"])
    portapi.inter_batch(b, pic["refs"], pic["blocks"], pic["params"], prof=pic["prof"])
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
