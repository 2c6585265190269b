"""GPU parity: CUDA deblocking (picture-level V then H passes) against the CTU-ordered oracle, bit-exact."""
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


def gpu_deblock(base, t, edge):
    import torch
    dev = util.to_device_planes(base, torch)
    edge_d = torch.from_numpy(edge).cuda()
    qp_d = torch.from_numpy(t["qp"]).cuda()
    off_d = torch.from_numpy(t["ctb_offsets"]).cuda()
    tables = lib.deblock_tables(edge_d, qp_d, off_d, t["stride4"])
    lib.deblock_picture(lib.device_planes(dev), tables, t["params"])
    torch.cuda.synchronize()
    return util.from_device_planes(dev)


@pytest.mark.parametrize("seed,w,h,ctb_log2", [(1, 416, 240, 7), (2, 416, 240, 6), (3, 832, 480, 7), (4, 136, 72, 5),
                                               (5, 1920, 1080, 7)])
def test_deblock_matches_oracle(seed, w, h, ctb_log2):
    t = synth.deblock_picture(seed, w, h, ctb_log2=ctb_log2)
    base = synth.smooth_planes(seed, t["dims"])
    edge = lib.pack_deblock_edges(t["bs"], t["len_p"], t["len_q"], t["tb_w_c"], t["tb_h_c"], ctb_log2)
    want = [p.copy() for p in base]
    oracle_deblock(want, t, edge)
    got = gpu_deblock(base, t, edge)
    for c in range(3):
        assert np.array_equal(got[c], want[c]), "plane %d differs" % c
    assert util.plane_md5(got) == util.plane_md5(want)


def test_deblock_no_edges_is_identity_4k():
    t = synth.deblock_picture(9, 3840, 2160)
    base = synth.smooth_planes(9, t["dims"])
    edge = np.zeros((2, 3, t["rows4"], t["stride4"]), np.uint8)
    got = gpu_deblock(base, t, edge)
    assert util.plane_md5(got) == util.plane_md5(base)
