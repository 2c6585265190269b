"""Shared helpers for the parity tests."""
import hashlib
import os

import numpy as np

from ffmpeg_b200 import abi
from ffmpeg_b200.hostutil import from_device_planes, md5, plane_md5, to_device_planes  # noqa: F401
from oracle import portapi, refapi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def noise_planes(seed, dims, bd=10):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, 1 << bd, (h, w)).astype(np.uint16) for (w, h) in dims]


def oracle_residual(planes, pic, want_residual=False):
    """The checker: compiled reference when present (oracle/_ref), else the C port."""
    if refapi.available():
        return refapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], pic["intra_mode"], want_residual)
    return portapi.residual_batch(planes, pic["levels"], pic["tbs"], pic["params"], want_residual)


def oracle_inter(planes, it, fwd=None, use_ref=None, threads=1):
    """Stage INTER of a picture's `inter` dict on the oracle.  Reference: the block-level harness on the regular
    sub-blocks, the CU-level harness (predict_inter on whole units) on the affine / GPM units; port: every record."""
    if use_ref is None:
        use_ref = refapi.available()
    blocks = it["blocks"]
    if not len(blocks):
        return
    if not use_ref:
        prof = it.get("prof")
        portapi.inter_batch(planes, it["refs"], blocks, it["params"], fwd, prof=prof if prof is not None and len(prof) else None)
        return
    special = abi.PB_GPM | abi.PB_AFFINE | abi.PB_LUMA_ONLY | abi.PB_CHROMA_ONLY | abi.PB_CIIP   # CIIP: predicted in RECON
    regular = blocks[(blocks["flags"] & special) == 0]
    if len(regular):
        refapi.inter_batch(planes, it["refs"], regular, it["params"], fwd)
    cus = it.get("cus")
    if cus is not None and len(cus):
        refapi.inter_cus(planes, it["refs"], cus, it["sb_mv"], it["params"], fwd)


def oracle_recon(planes, pic, order=None):
    """Stage INTER + RECON of a synth.recon_picture(): the compiled reference's ff_vvc_reconstruct when present,
    else the C port's inter -> residual planes -> intra chain.  `order`: intra records to use with the port
    (e.g. the wavefront order) instead of the decoding order."""
    it, rs, ia = pic["inter"], pic["residual"], pic["intra"]
    use_ref = refapi.available() and order is None
    oracle_inter(planes, it, None, use_ref)
    if use_ref:
        refapi.recon_picture(planes, pic["ref"], ia["params"], inter=it)
        return planes
    res = [np.zeros(p.shape, np.int16) for p in planes]
    portapi.residual_batch(planes, rs["levels"], rs["tbs"], rs["params"], residual_planes=res)
    portapi.intra_picture(planes, res, ia["blocks"] if order is None else order, ia["params"])
    return planes


def read_device_planes(planes):
    """Host copies of the planes an abi.Planes struct points at in DEVICE memory (cudaMemcpy2D through the CUDA runtime,
    which shares the primary context with the engine); used to look at exported windows."""
    import ctypes as C
    import glob
    paths = glob.glob("/usr/local/cuda/lib64/libcudart.so*") + ["libcudart.so"]
    rt = C.CDLL(sorted(paths)[0])
    out = []
    for c in range(3):
        w, h = planes.width[c], planes.height[c]
        a = np.zeros((h, w), np.uint16)
        rc = rt.cudaMemcpy2D(C.c_void_p(a.ctypes.data), C.c_size_t(w * 2), C.c_void_p(planes.plane[c]), C.c_size_t(planes.stride[c]),
                             C.c_size_t(w * 2), C.c_size_t(h), C.c_int(2))
        assert rc == 0, "cudaMemcpy2D failed: %d" % rc
        out.append(a)
    return out
