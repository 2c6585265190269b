"""GPU parity: CUDA inter prediction (through the C ABI) against the oracle, bit-exact."""
import numpy as np
import pytest

from ffmpeg_b200 import abi, lib, synth
from oracle import portapi, refapi
from tests import util

pytestmark = pytest.mark.gpu


def oracle_inter(dst, pic, lut=None):
    if refapi.available():
        refapi.inter_batch(dst, pic["refs"], pic["blocks"], pic["params"], lut)
    else:
        portapi.inter_batch(dst, pic["refs"], pic["blocks"], pic["params"], lut)


def _params_for(pic):
    """InterParams with the width-class counts of exactly the blocks being submitted (subsets stay width-ordered)."""
    import ctypes as C
    p = abi.InterParams.from_buffer_copy(bytes(pic["params"]))
    b = pic["blocks"]
    assert np.all(np.diff(b["w"].astype(np.int32)) <= 0)
    p.n_w16, p.n_w8 = int(np.sum(b["w"] == 16)), int(np.sum(b["w"] == 8))
    return p


def gpu_inter(pic, base, lut=None, tmp14=False):
    import torch
    pic = dict(pic, params=_params_for(pic))
    dst = util.to_device_planes(base, torch)
    refs = [util.to_device_planes(r, torch) for r in pic["refs"]]
    blocks = torch.from_numpy(pic["blocks"].view(np.uint8).reshape(-1)).cuda()
    lut_d = None if lut is None else torch.from_numpy(lut.view(np.int16)).cuda()
    if tmp14:
        t = [torch.zeros_like(d) for d in dst]
        lib.inter_batch(None, [lib.device_planes(r) for r in refs], blocks, len(pic["blocks"]), pic["params"],
                        tmp14=lib.device_planes(t))
        torch.cuda.synchronize()
        return [x.cpu().numpy() for x in t]
    lib.inter_batch(lib.device_planes(dst), [lib.device_planes(r) for r in refs], blocks, len(pic["blocks"]),
                    pic["params"], lmcs_fwd_lut=lut_d)
    torch.cuda.synchronize()
    return util.from_device_planes(dst)


@pytest.mark.parametrize("seed,w,h,n_refs", [(1, 416, 240, 2), (2, 416, 240, 4), (3, 832, 480, 3), (4, 136, 72, 1)])
def test_inter_batch_matches_oracle(seed, w, h, n_refs):
    pic = synth.inter_picture(seed, w, h, n_refs=n_refs)
    base = util.noise_planes(seed, pic["dims"])
    want = [p.copy() for p in base]
    oracle_inter(want, pic)
    got = gpu_inter(pic, base)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
    assert util.plane_md5(got) == util.plane_md5(want)


def test_inter_far_vectors_and_lmcs():
    pic = synth.inter_picture(5, 256, 128, far_mv=0.6)
    pic["blocks"]["flags"] |= abi.PB_LMCS
    lut = np.sort(np.random.default_rng(5).integers(0, 1024, 1024)).astype(np.uint16)
    base = util.noise_planes(5, pic["dims"])
    want = [p.copy() for p in base]
    oracle_inter(want, pic, lut)
    got = gpu_inter(pic, base, lut)
    assert util.plane_md5(got) == util.plane_md5(want)


def test_inter_put_intermediate_matches_port():
    """inter.put view (14-bit int16 intermediate of list 0)."""
    pic = synth.inter_picture(6, 256, 128)
    base = util.noise_planes(6, pic["dims"])
    t14 = [np.zeros(p.shape, np.int16) for p in base]
    portapi.inter_batch([p.copy() for p in base], pic["refs"], pic["blocks"], pic["params"], tmp14=t14)
    got = gpu_inter(pic, base, tmp14=True)
    for c in range(3):
        assert np.array_equal(got[c], t14[c])


def test_inter_full_size_properties_4k():
    """4K: (a) an all-zero-MV uni prediction with default weights copies the reference picture exactly,
    (b) predicting two disjoint halves of the batch composes to the whole, (c) oracle agrees on a sample."""
    import torch
    pic = synth.inter_picture(8, 3840, 2160, n_refs=2, tools=("bi",))
    base = util.noise_planes(8, pic["dims"])
    ident = dict(pic, blocks=pic["blocks"].copy())
    ident["blocks"]["pred_flag"] = 1
    ident["blocks"]["mv"] = 0
    ident["blocks"]["ref_idx"] = 0
    got = gpu_inter(ident, base)
    slot = pic["params"].ref_slot[0][0]
    assert util.plane_md5(got) == util.plane_md5(pic["refs"][slot])
    whole = gpu_inter(pic, base)
    n = len(pic["blocks"])
    first = gpu_inter(dict(pic, blocks=pic["blocks"][: n // 2]), base)
    both = gpu_inter(dict(pic, blocks=pic["blocks"][n // 2:]), first)
    assert util.plane_md5(both) == util.plane_md5(whole)
    sample = dict(pic, blocks=pic["blocks"][:3000])
    want = [p.copy() for p in base]
    oracle_inter(want, sample)
    got = gpu_inter(sample, base)
    assert util.plane_md5(got) == util.plane_md5(want)


@pytest.mark.parametrize("seed", range(6))
def test_dmvr_bdof_match_oracle(seed):
    """DMVR + BDOF on the GPU: pixels and the refined vectors (the host's tab_dmvr_mvf feedback) against the oracle."""
    import torch
    tools = ("bi", "wp", "bcw", "hpel") + (("dmvr",) if seed % 3 != 1 else ()) + (("bdof",) if seed % 3 != 0 else ())
    pic = synth.inter_picture(40 + seed, 416, 240, tools=tools, sizes=(8, 16) if seed & 1 else (4, 8, 16),
                              mv_range=(64 * 16 if seed < 4 else 24))
    if seed >= 4:
        pic["refs"][1] = [p.copy() for p in pic["refs"][0]]
        pic["blocks"]["mv"][:, 1, :] = -pic["blocks"]["mv"][:, 0, :]
    base = util.noise_planes(seed, pic["dims"])
    want = [p.copy() for p in base]
    orc = refapi if refapi.available() else portapi
    want_mv = orc.inter_batch(want, pic["refs"], pic["blocks"], pic["params"], want_refined_mv=True)
    dst = util.to_device_planes(base, torch)
    refs = [util.to_device_planes(r, torch) for r in pic["refs"]]
    blocks = torch.from_numpy(pic["blocks"].view(np.uint8).reshape(-1)).cuda()
    mv_d = torch.zeros((len(pic["blocks"]), 2, 2), dtype=torch.int32, device="cuda")
    lib.inter_batch(lib.device_planes(dst), [lib.device_planes(r) for r in refs], blocks, len(pic["blocks"]),
                    _params_for(pic), refined_mv=mv_d)
    torch.cuda.synchronize()
    got = util.from_device_planes(dst)
    assert np.array_equal(mv_d.cpu().numpy(), want_mv)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c


@pytest.mark.parametrize("seed,kinds,lmcs", [(0, (0, 2), False), (1, (0, 1), False), (2, (0, 1, 2), True), (3, (2,), False),
                                             (4, (1,), True)])
def test_affine_prof_gpm_match_oracle(seed, kinds, lmcs):
    """Affine + PROF and GPM coding units on the GPU (expanded records) against the reference's CU-level predict_inter()."""
    import torch
    pic = synth.inter_cu_picture(70 + seed, 416, 240, kinds=kinds, lmcs=lmcs)
    lut = np.sort(np.random.default_rng(seed).integers(0, 1024, 1024)).astype(np.uint16) if lmcs else None
    base = util.noise_planes(seed, pic["dims"])
    want = [p.copy() for p in base]
    if refapi.available():
        refapi.inter_cus(want, pic["refs"], pic["cus"], pic["sb_mv"], pic["params"], lut)
    else:
        portapi.inter_batch(want, pic["refs"], pic["blocks"], pic["params"], lut, prof=pic["prof"])
    dst = util.to_device_planes(base, torch)
    refs = [util.to_device_planes(r, torch) for r in pic["refs"]]
    blocks = torch.from_numpy(pic["blocks"].view(np.uint8).reshape(-1)).cuda()
    prof = torch.from_numpy(pic["prof"].view(np.uint8).reshape(-1)).cuda() if len(pic["prof"]) else None
    lib.inter_batch(lib.device_planes(dst), [lib.device_planes(r) for r in refs], blocks, len(pic["blocks"]),
                    _params_for(pic), lmcs_fwd_lut=None if lut is None else torch.from_numpy(lut.view(np.int16)).cuda(),
                    prof=prof)
    torch.cuda.synchronize()
    got = util.from_device_planes(dst)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
