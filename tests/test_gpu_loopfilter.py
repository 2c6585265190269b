"""GPU parity: CUDA SAO / ALF (+CC-ALF) / LMCS picture passes against the oracle, bit-exact."""
import numpy as np
import pytest

from ffmpeg_b200 import lib, synth
from oracle import portapi, refapi
from tests import util

pytestmark = pytest.mark.gpu
CASES = [(7, 1, 1), (6, 3, 0), (5, 4, 0), (7, 2, 0)]


def oracle_sao(base, lf):
    if refapi.available():
        out = [p.copy() for p in base]
        refapi.sao_picture(out, lf)
        return out
    out = [np.zeros_like(p) for p in base]
    portapi.sao_picture(out, base, lf)
    return out


def oracle_alf(base, lf):
    if refapi.available():
        out = [p.copy() for p in base]
        refapi.alf_picture(out, lf)
        return out
    out = [np.zeros_like(p) for p in base]
    portapi.alf_picture(out, base, lf)
    return out


def gpu_sao(base, lf):
    import torch
    src = util.to_device_planes(base, torch)
    dst = [torch.zeros_like(t) for t in src]
    ctbs = torch.from_numpy(lf["sao"].view(np.uint8).reshape(-1)).cuda()
    lib.sao_picture(lib.device_planes(dst), lib.device_planes(src), ctbs, lf["ctb_log2"])
    torch.cuda.synchronize()
    return util.from_device_planes(dst)


def gpu_alf(base, lf):
    import torch
    src = util.to_device_planes(base, torch)
    dst = [torch.zeros_like(t) for t in src]
    ctbs = torch.from_numpy(lf["alf"].view(np.uint8).reshape(-1)).cuda()
    params = torch.from_numpy(lf["alf_params"].view(np.uint8).reshape(-1)).cuda()
    lib.alf_picture(lib.device_planes(dst), lib.device_planes(src), ctbs, params, lf["ctb_log2"])
    torch.cuda.synchronize()
    return util.from_device_planes(dst)


def gpu_sao_alf(base, lf, sao=True, alf=True, table=True):
    """The fused pass: vvcb200_sao_alf_picture (table: with the per-picture coefficient table in caller scratch)."""
    import torch
    scratch = torch.full((lib.sao_alf_scratch_bytes(),), 0x5a, dtype=torch.uint8, device="cuda") if table else None
    src = util.to_device_planes(base, torch)
    dst = util.to_device_planes([np.full_like(p, 0x5555) for p in base], torch)
    sao_d = torch.from_numpy(lf["sao"].view(np.uint8).reshape(-1)).cuda() if sao else None
    alf_d = torch.from_numpy(lf["alf"].view(np.uint8).reshape(-1)).cuda() if alf else None
    params = torch.from_numpy(lf["alf_params"].view(np.uint8).reshape(-1)).cuda() if alf else None
    lib.sao_alf_picture(lib.device_planes(dst), lib.device_planes(src), sao_d, alf_d, params, lf["ctb_log2"], lf.get("bd", 10),
                        scratch=scratch)
    torch.cuda.synchronize()
    return util.from_device_planes(dst)


@pytest.mark.parametrize("w,h", [(416, 240), (136, 72)])
@pytest.mark.parametrize("i,case", list(enumerate(CASES)))
def test_fused_sao_alf_matches_oracle_chain(i, case, w, h):
    ctb, ns, lfase = case
    lf = synth.loopfilter_ctbs(40 + i, w, h, ctb_log2=ctb, n_slices=ns, lfase=lfase)
    base = synth.smooth_planes(40 + i, lf["dims"])
    sao = oracle_sao(base, lf)
    want = oracle_alf(sao, lf)
    for table in (True, False):
        got = gpu_sao_alf(base, lf, table=table)
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "plane %d differs (table %d)" % (c, table)
    # either stage alone
    got = gpu_sao_alf(base, lf, alf=False)
    for c in range(3):
        assert np.array_equal(got[c], sao[c]), "SAO only: plane %d differs" % c
    want = oracle_alf(base, lf)
    got = gpu_sao_alf(base, lf, sao=False)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "ALF only: plane %d differs" % c


def test_fused_sao_alf_1080p_and_12_bit():
    for (w, h, ctb, bd, ns, lfase) in ((1920, 1080, 7, 10, 3, 0), (832, 480, 6, 12, 2, 0)):
        lf = synth.loopfilter_ctbs(51, w, h, ctb_log2=ctb, n_slices=ns, lfase=lfase, alf_on=0.9, bd=bd)
        base = synth.smooth_planes(51, lf["dims"], bd=bd)
        want = oracle_alf(oracle_sao(base, lf), lf)
        assert util.plane_md5(gpu_sao_alf(base, lf)) == util.plane_md5(want)


@pytest.mark.parametrize("i,case", list(enumerate(CASES)))
def test_sao_matches_oracle(i, case):
    ctb, ns, lfase = case
    lf = synth.loopfilter_ctbs(10 + i, 416, 240, ctb_log2=ctb, n_slices=ns, lfase=lfase)
    base = synth.smooth_planes(i, lf["dims"])
    want, got = oracle_sao(base, lf), gpu_sao(base, lf)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c


@pytest.mark.parametrize("i,case", list(enumerate(CASES)))
def test_alf_matches_oracle(i, case):
    ctb, ns, lfase = case
    lf = synth.loopfilter_ctbs(20 + i, 416, 240, ctb_log2=ctb, n_slices=ns, lfase=lfase)
    base = synth.smooth_planes(i, lf["dims"])
    want, got = oracle_alf(base, lf), gpu_alf(base, lf)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c


def test_alf_ctb_multiple_height_and_1080p():
    for (w, h, ctb) in ((256, 256, 7), (1920, 1080, 7)):
        lf = synth.loopfilter_ctbs(31, w, h, ctb_log2=ctb, alf_on=1.0)
        base = synth.smooth_planes(31, lf["dims"])
        assert util.plane_md5(gpu_alf(base, lf)) == util.plane_md5(oracle_alf(base, lf))
        assert util.plane_md5(gpu_sao(base, lf)) == util.plane_md5(oracle_sao(base, lf))


def test_filters_off_are_identity_4k():
    lf = synth.loopfilter_ctbs(5, 3840, 2160)
    lf["sao"]["type_idx"] = 0
    lf["alf"]["ctb_flag"] = 0
    lf["alf"]["cc_idc"] = 0
    base = synth.smooth_planes(5, lf["dims"])
    assert util.plane_md5(gpu_sao(base, lf)) == util.plane_md5(base)
    assert util.plane_md5(gpu_alf(base, lf)) == util.plane_md5(base)
    assert util.plane_md5(gpu_sao_alf(base, lf)) == util.plane_md5(base)


def test_lmcs_lut_pass():
    import torch
    fwd, inv = synth.lmcs_luts(2)
    base = util.noise_planes(2, synth.plane_dims(1920, 1080))
    dev = util.to_device_planes(base, torch)
    lib.lmcs_picture(lib.device_planes(dev), torch.from_numpy(inv.view(np.int16)).cuda())
    torch.cuda.synchronize()
    got = util.from_device_planes(dev)
    assert np.array_equal(got[0], inv[base[0]])
    assert np.array_equal(got[1], base[1])


@pytest.mark.parametrize("lfate", [0, 1])
@pytest.mark.parametrize("ctb,w,h", [(5, 416, 240), (6, 416, 240), (7, 832, 480)])
def test_two_by_two_tiles_match_reference(ctb, w, h, lfate):
    """pps_loop_filter_across_tiles_enabled_flag 0 and 1 on a 2x2 tile grid with two slices (vvc_filter.c:163-212,
    1264-1280): the host derives the per-CTB flags, the reference derives its own from ctb_to_col_bd / ctb_to_row_bd."""
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    lf = synth.loopfilter_ctbs(60 + ctb, w, h, ctb_log2=ctb, n_slices=2, lfase=0, tile_cols=(0, cw // 2), tile_rows=(0, ch // 2),
                               lfate=lfate)
    assert lf["tiles"]["num_tiles"] == 4
    base = synth.smooth_planes(60, lf["dims"])
    sao = oracle_sao(base, lf)
    want = oracle_alf(sao, lf)
    assert util.plane_md5(gpu_sao(base, lf)) == util.plane_md5(sao)
    assert util.plane_md5(gpu_alf(sao, lf)) == util.plane_md5(want)
    assert util.plane_md5(gpu_sao_alf(base, lf)) == util.plane_md5(want)
