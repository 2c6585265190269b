"""GPU test of vvcb200_intra_picture_static, the intra entry point that keeps its marks in the library's built-in
scratch and allocates nothing.  Collected last on purpose (file name): it must pass after every other GPU test has
used the device, and on its own.  Through the C ABI; compared with vvcb200_intra_picture on two calls in a row."""
import ctypes as C

import numpy as np
import pytest

from ffmpeg_b200 import abi, lib, synth
from tests import util

pytestmark = pytest.mark.gpu


def _stage(pic):
    """INTER + RESIDUAL on the device; returns the planes before INTRA, the residual planes and the intra records."""
    import torch
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    dev = util.to_device_planes([np.zeros((h, w), np.uint16) for (w, h) in pic["dims"]], torch)
    dpl = lib.device_planes(dev)
    if len(it["blocks"]):
        refs = [util.to_device_planes(r, torch) for r in it["refs"]]
        blocks = torch.from_numpy(it["blocks"].view(np.uint8).reshape(-1)).cuda()
        lib.inter_batch(dpl, [lib.device_planes(r) for r in refs], blocks, len(it["blocks"]), it["params"])
    res = [torch.zeros_like(t) for t in dev]
    tbs = torch.from_numpy(rs["tbs"].view(np.uint8).reshape(-1)).cuda()
    levels = torch.from_numpy(rs["levels"]).cuda()
    lib.residual_batch(dpl, levels, tbs, len(rs["tbs"]), rs["params"], residual=lib.device_planes(res))
    w, h, ctb = pic["width"], pic["height"], pic["ctb_log2"]
    recs = lib.intra_schedule(ia["blocks"], ia["ctu_first"], (w + (1 << ctb) - 1) >> ctb, (h + (1 << ctb) - 1) >> ctb, ctb, w, h)
    torch.cuda.synchronize()
    return dev, res, torch.from_numpy(recs.view(np.uint8).reshape(-1)).cuda(), len(recs)


@pytest.mark.parametrize("seed,w,h,ctb", [(41, 416, 240, 6), (42, 1920, 1080, 7)])
def test_static_scratch_matches_caller_scratch_twice_in_a_row(seed, w, h, ctb):
    import torch
    pic = synth.recon_picture(seed, w, h, ctb_log2=ctb, intra_fraction=0.7)
    dev, res, recs, n = _stage(pic)
    base = [t.clone() for t in dev]
    dpl, rpl = lib.device_planes(dev), lib.device_planes(res)
    scratch = torch.zeros(lib.intra_scratch_bytes(w, h), dtype=torch.uint8, device="cuda")
    lib.intra_picture(dpl, rpl, recs, n, pic["intra"]["params"], scratch)
    torch.cuda.synchronize()
    want = util.from_device_planes(dev)
    for call in range(2):
        for t, b in zip(dev, base):
            t.copy_(b)
        lib.intra_picture_static(dpl, rpl, recs, n, pic["intra"]["params"])
        torch.cuda.synchronize()
        got = util.from_device_planes(dev)
        for c in range(3):
            assert np.array_equal(got[c], want[c]), "call %d plane %d differs" % (call, c)


def test_static_scratch_rejects_pictures_beyond_its_size():
    l = lib.load()
    pl = abi.Planes()
    pl.width[0], pl.height[0] = 3840, 2160
    ip = abi.IntraParams()
    ip.bit_depth, ip.ctb_log2_size = 10, 7
    dummy = C.c_void_p(256)                      # never dereferenced: the size check comes first
    rc = l.vvcb200_intra_picture_static(C.byref(pl), None, dummy, 1, C.byref(ip), None)
    assert rc == -1
    msg = l.vvcb200_last_error()
    assert b"1920x1088" in msg and b"3840x2160" in msg
