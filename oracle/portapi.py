"""TEST INFRASTRUCTURE: ctypes loader for oracle/_build/libvvcport.so (the plain-C restatement)."""
import ctypes as C
import os
import subprocess

import numpy as np

from ffmpeg_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_build", "libvvcport.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    return SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            build()
        _lib = C.CDLL(SO)
    return _lib


def itx(coeffs, trh, trv, log2w, log2h, nzw, nzh, bd=10, log2_range=15):
    assert coeffs.dtype == np.int32 and coeffs.flags.c_contiguous
    lib().port_itx(C.c_void_p(coeffs.ctypes.data), trh, trv, log2w, log2h, nzw, nzh, log2_range, bd)
    return True


def residual_batch(planes, levels, tbs, params, want_residual=False, residual_planes=None):
    """residual_planes: int16 planes receiving the residual of TBs flagged TB_TO_RESIDUAL (else they are added too)."""
    pl = abi.host_planes(planes)
    rp = None if residual_planes is None else abi.host_planes([r.view(np.uint16) for r in residual_planes])
    levels = np.ascontiguousarray(levels, np.int16)
    tbs = np.ascontiguousarray(tbs, abi.TB_DTYPE)
    res = None
    if want_residual:
        total = int(np.sum(1 << (tbs["log2_w"].astype(np.int64) + tbs["log2_h"])))
        res = np.zeros(total, np.int32)
    lib().port_residual_batch(C.byref(pl), C.c_void_p(levels.ctypes.data), C.c_void_p(tbs.ctypes.data), len(tbs),
                              C.byref(params), C.c_void_p(res.ctypes.data) if res is not None else None,
                              C.byref(rp) if rp is not None else None)
    return res


def inter_batch(dst, refs, blocks, params, lmcs_fwd_lut=None, tmp14=None, want_refined_mv=False, prof=None):
    pl = abi.host_planes(dst)
    ra = abi.planes_array([abi.host_planes(r) for r in refs])
    blocks = np.ascontiguousarray(blocks, abi.PB_DTYPE)
    lut = None if lmcs_fwd_lut is None else np.ascontiguousarray(lmcs_fwd_lut, np.uint16)
    t14 = None
    if tmp14 is not None:
        t14 = abi.host_planes([t.view(np.uint16) for t in tmp14])
    mv = np.zeros((len(blocks), 2, 2), np.int32) if want_refined_mv else None
    lib().port_inter_batch(C.byref(pl), ra, len(refs), C.c_void_p(blocks.ctypes.data), len(blocks), C.byref(params),
                           C.c_void_p(lut.ctypes.data) if lut is not None else None,
                           C.byref(t14) if t14 is not None else None,
                           C.c_void_p(mv.ctypes.data) if mv is not None else None,
                           C.c_void_p(np.ascontiguousarray(prof).ctypes.data) if prof is not None and len(prof) else None)
    return mv


def deblock_picture(planes, edge, qp, ctb_offsets, stride4, params):
    """Port of both deblocking passes; edge = packed bytes [2][3][r4][s4] (host numpy)."""
    from ffmpeg_b200 import lib as product
    pl = abi.host_planes(planes)
    t = product.deblock_tables(edge, qp, ctb_offsets, stride4)
    lib().port_deblock_picture(C.byref(pl), C.byref(t), C.byref(params))


def sao_picture(dst, src, lf, bit_depth=None):
    bit_depth = lf.get("bd", 10) if bit_depth is None else bit_depth
    d, s = abi.host_planes(dst), abi.host_planes(src)
    sao = np.ascontiguousarray(lf["sao"])
    lib().port_sao_picture(C.byref(d), C.byref(s), C.c_void_p(sao.ctypes.data), int(lf["ctb_log2"]), bit_depth)


def alf_picture(dst, src, lf):
    d, s = abi.host_planes(dst), abi.host_planes(src)
    alf = np.ascontiguousarray(lf["alf"])
    p = np.ascontiguousarray(lf["alf_params"])
    lib().port_alf_picture(C.byref(d), C.byref(s), C.c_void_p(alf.ctypes.data), C.c_void_p(p.ctypes.data))


def lmcs_picture(planes, lut):
    pl = abi.host_planes(planes)
    lut = np.ascontiguousarray(lut, np.uint16)
    lib().port_lmcs_picture(C.byref(pl), C.c_void_p(lut.ctypes.data))


def intra_picture(planes, residual_planes, blocks, params):
    pl = abi.host_planes(planes)
    rp = abi.host_planes([r.view(np.uint16) for r in residual_planes])
    blocks = np.ascontiguousarray(blocks, abi.IB_DTYPE)
    lib().port_intra_picture(C.byref(pl), C.byref(rp), C.c_void_p(blocks.ctypes.data), len(blocks), C.byref(params))
