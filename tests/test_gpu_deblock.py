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


def _norm(e):
    return np.where((e & 3) == 0, 0, e)


@pytest.mark.parametrize("seed,w,h,ctb,frac,ns", [(40, 416, 240, 6, 0.3, 1), (43, 416, 240, 5, 0.3, 3), (44, 832, 480, 7, 0.2, 2),
                                                 (45, 1920, 1080, 7, 0.1, 1)])
def test_device_derived_edges_match_the_reference_and_filter_the_same_picture(seed, w, h, ctb, frac, ns):
    """vvcb200_deblock_derive_edges over a consistent picture state against the reference's vvc_deblock_bs
    (oracle/_ref), every 4x4 unit; then the picture deblocked from the device-derived edges against the reference's
    deblocking driven by ITS tables."""
    import torch
    from oracle import port_bs
    pic = synth.full_picture(seed, w, h, ctb_log2=ctb, intra_fraction=frac)
    st = synth.deblock_state(pic)
    cw, ch = (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb
    slices = synth.slice_map(np.random.default_rng(seed), cw, ch, ns)
    lfase = 1 if ns == 1 else 0
    edges = (synth.alf_edge_flags(slices, cw, ch, lfase) & 3).astype(np.uint8)
    d_motion = torch.from_numpy(st["motion"].view(np.uint8).reshape(-1)).cuda()
    d_part = torch.from_numpy(st["part"].view(np.uint8).reshape(-1)).cuda()
    d_edges = torch.from_numpy(edges).cuda()
    out = torch.full((2, 3, st["rows4"], st["stride4"]), 0x77, dtype=torch.uint8, device="cuda")
    lib.deblock_derive_edges(out, d_motion, d_part, d_edges, w, h, st["stride4"], ctb)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    if w * h <= 416 * 240:
        assert np.array_equal(got, port_bs.derive_edges(st["motion"], st["part"], w, h, ctb, edges))
    t = synth.deblock_picture(seed, w, h, ctb_log2=ctb)          # qp tables, per-CTB offsets, LADF parameters
    if refapi.available():
        bs, lp, lq = refapi.deblock_bs(st, w, h, ctb, slices, lfase)
        want = lib.pack_deblock_edges(bs, lp, lq, np.ascontiguousarray(st["tb_width"][1]), np.ascontiguousarray(st["tb_height"][1]), ctb)
        assert np.array_equal(_norm(got), _norm(want))
        t = dict(t, bs=bs, len_p=lp, len_q=lq, tb_w_c=np.ascontiguousarray(st["tb_width"][1]), tb_h_c=np.ascontiguousarray(st["tb_height"][1]))
        base = synth.smooth_planes(seed, t["dims"])
        ref_pic = [p.copy() for p in base]
        refapi.deblock_picture(ref_pic, t)
        for fused in (False, True):
            assert util.plane_md5(gpu_deblock(base, t, got, fused)) == util.plane_md5(ref_pic)


def test_luma_and_chroma_halves_of_the_tile_pass_equal_the_whole():
    """vvcb200_deblock_lmcs_planes: luma alone (which may start when luma is final) + chroma alone = the picture call."""
    import torch
    seed, w, h, ctb = 47, 832, 480, 7
    t = synth.deblock_picture(seed, w, h, ctb_log2=ctb)
    base = synth.smooth_planes(seed, t["dims"])
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], ctb)
    whole = gpu_deblock(base, t, edge, fused=True)
    src = util.to_device_planes(base, torch)
    dst = [torch.zeros_like(p) for p in src]
    d_edge, d_qp, d_off = torch.from_numpy(edge).cuda(), torch.from_numpy(t["qp"]).cuda(), torch.from_numpy(t["ctb_offsets"]).cuda()
    tabs = lib.deblock_tables(d_edge, d_qp, d_off, t["stride4"])
    second = torch.cuda.Stream()
    second.wait_stream(torch.cuda.current_stream())
    lib.deblock_lmcs_planes(lib.device_planes(dst), lib.device_planes(src), tabs, t["params"], planes=1, stream=second.cuda_stream)
    lib.deblock_lmcs_planes(lib.device_planes(dst), lib.device_planes(src), tabs, t["params"], planes=2)
    torch.cuda.synchronize()
    assert util.plane_md5(util.from_device_planes(dst)) == util.plane_md5(whole)
