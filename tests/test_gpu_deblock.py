"""GPU parity: CUDA deblocking against the CTU-ordered oracle, bit-exact — the two picture-level passes
(vvcb200_deblock_picture) and the tile-resident single pass (vvcb200_deblock_lmcs_picture)."""
import numpy as np
import pytest

from ffmpeg_b200 import lib, synth
from oracle import portapi, refapi
from tests import util

pytestmark = pytest.mark.gpu


def oracle_deblock(planes, t, edge):
    if refapi.available():
        refapi.deblock_picture(planes, t)
    else:
        portapi.deblock_picture(planes, edge, t["qp"], t["ctb_offsets"], t["stride4"], t["params"])


def gpu_deblock(base, t, edge, fused=False, inv_lut=None):
    import torch
    dev = util.to_device_planes(base, torch)
    edge_d = torch.from_numpy(edge).cuda()
    qp_d = torch.from_numpy(t["qp"]).cuda()
    off_d = torch.from_numpy(t["ctb_offsets"]).cuda()
    tables = lib.deblock_tables(edge_d, qp_d, off_d, t["stride4"])
    if fused:
        out = util.to_device_planes([np.full_like(p, 0x5555) for p in base], torch)
        lut = None if inv_lut is None else torch.from_numpy(inv_lut.view(np.int16)).cuda()
        lib.deblock_lmcs_picture(lib.device_planes(out), lib.device_planes(dev), tables, t["params"], lut)
        torch.cuda.synchronize()
        return util.from_device_planes(out)
    lib.deblock_picture(lib.device_planes(dev), tables, t["params"])
    torch.cuda.synchronize()
    return util.from_device_planes(dev)


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("seed,w,h,ctb_log2", [(1, 416, 240, 7), (2, 416, 240, 6), (3, 832, 480, 7), (4, 136, 72, 5),
                                               (5, 1920, 1080, 7)])
def test_deblock_matches_oracle(seed, w, h, ctb_log2, fused):
    t = synth.deblock_picture(seed, w, h, ctb_log2=ctb_log2)
    base = synth.smooth_planes(seed, t["dims"])
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], ctb_log2)
    want = [p.copy() for p in base]
    oracle_deblock(want, t, edge)
    got = gpu_deblock(base, t, edge, fused)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
    assert util.plane_md5(got) == util.plane_md5(want)


@pytest.mark.parametrize("bd", [10, 12])
def test_tile_pass_with_inverse_lmcs_matches_lut_then_oracle(bd):
    t = synth.deblock_picture(21, 832, 480, bd=bd, ladf=True)
    base = synth.smooth_planes(21, t["dims"], bd=bd)
    _, inv = synth.lmcs_luts(3, bd=bd)
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], 7)
    want = [inv[base[0]], base[1].copy(), base[2].copy()]       # lmcs.filter on luma, then the reference's deblocking
    oracle_deblock(want, t, edge)
    got = gpu_deblock(base, t, edge, True, inv)
    assert util.plane_md5(got) == util.plane_md5(want)


@pytest.mark.parametrize("fused", [False, True])
def test_deblock_no_edges_is_identity_4k(fused):
    t = synth.deblock_picture(9, 3840, 2160)
    base = synth.smooth_planes(9, t["dims"])
    edge = np.zeros((2, 3, t["rows4"], t["stride4"]), np.uint8)
    got = gpu_deblock(base, t, edge, fused)
    assert util.plane_md5(got) == util.plane_md5(base)


def test_tile_pass_matches_two_pass_4k():
    t = synth.deblock_picture(11, 3840, 2160)
    base = synth.smooth_planes(11, t["dims"])
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], 7)
    assert util.plane_md5(gpu_deblock(base, t, edge, True)) == util.plane_md5(gpu_deblock(base, t, edge, False))
