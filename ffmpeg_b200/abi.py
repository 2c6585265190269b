"""ctypes / numpy mirrors of the POD structs declared in include/vvc_b200.h.

These are shared by the product binding (ffmpeg_b200.lib), by bench.py and by the test oracle
loaders (oracle/refapi.py, oracle/portapi.py); sizes are asserted against the C side at load time.
"""
import ctypes as C

import numpy as np

# enum TxType (reference: libavcodec/vvc/vvcdsp.h:30-35)
DCT2, DST7, DCT8 = 0, 1, 2

TB_TS, TB_BDPCM, TB_BDPCM_VERT, TB_DEP_QUANT = 1, 2, 4, 8
TB_SCALING, TB_INTER, TB_JOINT, TB_JOINT_SHIFT, TB_LMCS_SCALE = 16, 32, 64, 128, 256

# VVCB200TransformBlock
TB_DTYPE = np.dtype([
    ("x0", "<u2"), ("y0", "<u2"), ("c_idx", "u1"), ("log2_w", "u1"), ("log2_h", "u1"),
    ("nzw", "u1"), ("nzh", "u1"), ("trh", "u1"), ("trv", "u1"), ("qp", "u1"),
    ("flags", "<u2"), ("lfnst", "u1"), ("reserved", "u1"), ("lmcs_scale", "<u2"),
    ("reserved2", "<u2"), ("coeff_off", "<u4"),
], align=False)
assert TB_DTYPE.itemsize == 24


class ResidualParams(C.Structure):
    _fields_ = [
        ("bit_depth", C.c_int32), ("log2_transform_range", C.c_int32),
        ("joint_c_sign", C.c_int32), ("levels_sparse", C.c_int32), ("n_class", C.c_int32 * 4),
        ("scaling_matrix_rec", (C.c_uint8 * 64) * 28),
        ("scaling_matrix_dc_rec", C.c_uint8 * 14),
        ("pad", C.c_uint8 * 2),
    ]


class Planes(C.Structure):
    """VVCB200Planes (device pointers) and oracle VVCRefPlanes (host pointers) share this layout."""
    _fields_ = [
        ("plane", C.c_void_p * 3), ("stride", C.c_int32 * 3),
        ("width", C.c_int32 * 3), ("height", C.c_int32 * 3),
    ]


def residual_params(bit_depth=10, joint_c_sign=1, scaling=None, dc=None):
    p = ResidualParams()
    p.bit_depth, p.log2_transform_range, p.joint_c_sign = bit_depth, 15, joint_c_sign
    m = np.full((28, 64), 16, np.uint8) if scaling is None else np.ascontiguousarray(scaling, np.uint8)
    d = np.full(14, 16, np.uint8) if dc is None else np.ascontiguousarray(dc, np.uint8)
    C.memmove(p.scaling_matrix_rec, m.ctypes.data, 28 * 64)
    C.memmove(p.scaling_matrix_dc_rec, d.ctypes.data, 14)
    return p


def residual_class(tbs):
    """Vectorised vvcb200_residual_class() (include/vvc_b200.h) for a TB_DTYPE array."""
    fast = ((tbs["flags"] & (TB_TS | TB_BDPCM)) == 0) & (tbs["lfnst"] == 0) & (tbs["log2_w"] >= 1) & (tbs["log2_h"] >= 1)
    m = np.maximum(tbs["log2_w"], tbs["log2_h"]).astype(np.int32)
    cls = np.where(m >= 5, 0, np.where(m == 4, 1, np.where(m == 3, 2, 3)))
    return np.where(fast, cls, 4).astype(np.int32)


def tb_order(tbs):
    """Permutation into class order; inside a class larger areas first, then grouped by block shape, transform pair
    and flags, so that warps running side by side execute the same code with the same matrices (not required for
    correctness: the kernel only needs the class order)."""
    cls = residual_class(tbs)
    area = tbs["log2_w"].astype(np.int32) + tbs["log2_h"]
    shape = tbs["log2_w"].astype(np.int32) * 8 + tbs["log2_h"]
    tx = tbs["trh"].astype(np.int32) * 4 + tbs["trv"]
    flags = tbs["flags"].astype(np.int32) & ~np.int32(TB_INTER)
    return np.lexsort((flags, tx, shape, -area, cls)), cls


def order_tbs_by_class(tbs, params, extra=()):
    """Sort a TB batch into class order (see tb_order) and record the class sizes in `params`.
    `extra` arrays (one entry per TB) are permuted alongside.  Returns (tbs, *extra)."""
    order, cls = tb_order(tbs)
    for c in range(4):
        params.n_class[c] = int(np.sum(cls == c))
    return (np.ascontiguousarray(tbs[order]),) + tuple(np.ascontiguousarray(e[order]) for e in extra)


def host_planes(arrays):
    """Planes struct over a list of three 2-D uint16 numpy arrays (row stride taken from the array)."""
    p = Planes()
    for c, a in enumerate(arrays):
        assert a.dtype == np.uint16 and a.ndim == 2 and a.strides[1] == 2
        p.plane[c] = a.ctypes.data
        p.stride[c] = a.strides[0]
        p.width[c] = a.shape[1]
        p.height[c] = a.shape[0]
    return p


# ---- inter prediction -----------------------------------------------------------------------------------
PB_WP, PB_HPEL_IF, PB_AFFINE, PB_LUMA_ONLY, PB_CHROMA_ONLY = 1, 2, 4, 8, 16
PB_BDOF, PB_DMVR, PB_CIIP, PB_LMCS = 32, 64, 128, 256
PB_GPM, PB_PROF0, PB_PROF1 = 512, 1024, 2048


def gpm_aux(partition_idx, log2_cb_w, log2_cb_h, off_x, off_y, list0, list1):
    """VVCB200_GPM_AUX of include/vvc_b200.h"""
    return (partition_idx | ((log2_cb_w - 3) << 6) | ((log2_cb_h - 3) << 8) | ((off_x >> 2) << 10) |
            ((off_y >> 2) << 14) | (list0 << 18) | (list1 << 19))


PROF_DTYPE = np.dtype([("diff_mv_x", "i1", (2, 16)), ("diff_mv_y", "i1", (2, 16))])
assert PROF_DTYPE.itemsize == 64
PF_L0, PF_L1, PF_BI = 1, 2, 3

PB_DTYPE = np.dtype([
    ("x0", "<u2"), ("y0", "<u2"), ("w", "u1"), ("h", "u1"), ("pred_flag", "u1"), ("bcw_idx", "u1"),
    ("ref_idx", "i1", (2,)), ("flags", "<u2"), ("mv", "<i4", (2, 2)), ("aux", "<u4"),
], align=False)
assert PB_DTYPE.itemsize == 32


class InterParams(C.Structure):
    _fields_ = [
        ("bit_depth", C.c_int32), ("log2_denom", C.c_int32 * 2), ("n_w16", C.c_int32), ("n_w8", C.c_int32),
        ("num_ref_idx_active", C.c_int32 * 2), ("dmvr_first", C.c_int32),
        ("weight", ((C.c_int16 * 16) * 3) * 2), ("offset", ((C.c_int16 * 16) * 3) * 2),
        ("ref_slot", (C.c_uint8 * 16) * 2),
    ]


def inter_params(bit_depth=10, log2_denom=(0, 0), weight=None, offset=None, ref_slot=None):
    p = InterParams()
    p.bit_depth = bit_depth
    p.log2_denom[0], p.log2_denom[1] = log2_denom
    w = np.zeros((2, 3, 16), np.int16) if weight is None else np.ascontiguousarray(weight, np.int16)
    if weight is None:
        w[:, 0, :] = 1 << log2_denom[0]
        w[:, 1:, :] = 1 << log2_denom[1]
    o = np.zeros((2, 3, 16), np.int16) if offset is None else np.ascontiguousarray(offset, np.int16)
    r = np.zeros((2, 16), np.uint8) if ref_slot is None else np.ascontiguousarray(ref_slot, np.uint8)
    C.memmove(p.weight, w.ctypes.data, w.nbytes)
    C.memmove(p.offset, o.ctypes.data, o.nbytes)
    C.memmove(p.ref_slot, r.ctypes.data, r.nbytes)
    return p


def order_blocks_by_width(blocks, params):
    """Sort a prediction-block batch into the width order the engine wants and record the class sizes.
    Inside a width class records are grouped by the tools they use (GPM, affine, DMVR, BDOF, weighted, plain; uni / bi):
    not required for correctness, but warps that run at the same time then execute the same code path, which is what
    the instruction cache of an SM needs (the kernel is ~120 KB of code, see profiles/)."""
    f = blocks["flags"].astype(np.int64)
    kind = (f & (PB_GPM | PB_AFFINE | PB_LUMA_ONLY | PB_CHROMA_ONLY | PB_DMVR | PB_BDOF | PB_WP | PB_PROF0 | PB_PROF1)) * 4 + \
        blocks["pred_flag"].astype(np.int64) + 1024 * 1024 * (blocks["bcw_idx"] != 0)
    # the records with the longest latency first (DMVR search + BDOF, BDOF, GPM): a warp spends ten times as long on one of
    # those as on a plain block, and at the end of the grid they would be its tail
    slow = np.where(f & PB_DMVR, 0, np.where(f & PB_BDOF, 1, np.where(f & PB_GPM, 2, 3)))
    order = np.lexsort((kind, slow, -blocks["w"].astype(np.int32)))
    blocks = np.ascontiguousarray(blocks[order])
    params.n_w16 = int(np.sum(blocks["w"] == 16))
    params.n_w8 = int(np.sum(blocks["w"] == 8))
    params.dmvr_first = 1                           # slow rank 0 = DMVR: a prefix of each width class
    return blocks


def planes_array(list_of_planes):
    """Contiguous C array of Planes structs (for `refs`)."""
    arr = (Planes * len(list_of_planes))()
    for i, p in enumerate(list_of_planes):
        arr[i] = p
    return arr


# ---- deblocking ----------------------------------------------------------------------------------------------
class DeblockParams(C.Structure):
    _fields_ = [
        ("bit_depth", C.c_int32), ("ctb_log2_size", C.c_int32), ("ladf_enabled", C.c_int32),
        ("ladf_lowest_interval_qp_offset", C.c_int32), ("num_ladf_intervals", C.c_int32),
        ("ladf_qp_offset", C.c_int32 * 4), ("ladf_interval_lower_bound", C.c_int32 * 5),
    ]


class DeblockTables(C.Structure):
    _fields_ = [
        ("edge", (C.c_void_p * 3) * 2), ("qp", C.c_void_p * 3), ("ctb_offsets", C.c_void_p),
        ("stride4", C.c_int32), ("reserved", C.c_int32),
    ]


# ---- SAO / ALF -------------------------------------------------------------------------------------------------
SAO_DTYPE = np.dtype([
    ("type_idx", "u1", (3,)), ("band_position", "u1", (3,)), ("eo_class", "u1", (3,)), ("no_filter", "u1"),
    ("offset_val", "<i2", (3, 5)),
], align=False)
assert SAO_DTYPE.itemsize == 40

ALF_CTB_DTYPE = np.dtype([
    ("ctb_flag", "u1", (3,)), ("filt_set_idx_y", "u1"), ("alt_idx", "u1", (2,)), ("cc_idc", "u1", (2,)),
    ("edges", "u1"), ("reserved", "u1", (3,)),
], align=False)
assert ALF_CTB_DTYPE.itemsize == 12

ALF_PARAMS_DTYPE = np.dtype([
    ("bit_depth", "<i4"), ("ctb_log2_size", "<i4"),
    ("luma_coeff", "<i2", (7, 25, 12)), ("luma_clip_idx", "u1", (7, 25, 12)),
    ("chroma_coeff", "<i2", (8, 6)), ("chroma_clip_idx", "u1", (8, 6)),
    ("cc_coeff", "<i2", (2, 4, 7)), ("cc_enabled", "<i4", (2,)),
], align=True)


# ---- intra --------------------------------------------------------------------------------------------------------
IB_MIP, IB_MIP_TRANSPOSED, IB_ISP, IB_BDPCM, IB_RESIDUAL, IB_CIIP, IB_ADD_ONLY, IB_LMCS_SCALE = 1, 2, 4, 8, 16, 32, 64, 128
TB_TO_RESIDUAL = 512
TB_LEVELS8 = 1024

IB_DTYPE = np.dtype([
    ("x0", "<u2"), ("y0", "<u2"), ("w", "u1"), ("h", "u1"), ("c_idx", "u1"), ("mode", "i1"), ("ref_idx", "u1"),
    ("flags", "u1"), ("mip_mode", "u1"), ("up_left", "u1"), ("avail_top", "<u2"), ("avail_left", "<u2"),
    ("cb_w", "u1"), ("cb_h", "u1"), ("ciip_weight", "u1"), ("res_planes", "u1"),
    ("vpdu_x", "u1"), ("vpdu_y", "u1"), ("vpdu_avail", "u1"), ("isp_coded", "u1"), ("reserved2", "<u4"),
], align=False)
assert IB_DTYPE.itemsize == 28


class IntraParams(C.Structure):
    _fields_ = [("bit_depth", C.c_int32), ("ctb_log2_size", C.c_int32), ("chroma_vertical_collocated", C.c_int32),
                ("lmcs_chroma_scale", C.c_int32), ("lmcs_pivot", C.c_uint16 * 17), ("lmcs_chroma_scale_coeff", C.c_uint16 * 16),
                ("lmcs_min_bin_idx", C.c_uint8), ("lmcs_max_bin_idx", C.c_uint8), ("prepared", C.c_uint8), ("pad", C.c_uint8 * 3),
                ("n_luma", C.c_int32), ("n_add_only", C.c_int32), ("d_phase_counts", C.c_void_p), ("luma_done_event", C.c_void_p)]


class Window(C.Structure):
    _fields_ = [("crop_left", C.c_int32), ("crop_right", C.c_int32), ("crop_top", C.c_int32), ("crop_bottom", C.c_int32)]


class Export(C.Structure):
    _fields_ = [("planes", Planes), ("ready", C.c_void_p)]


# ---- boundary-strength derivation on the device (vvcb200_deblock_derive_edges) --------------------------------------
MOTION_DTYPE = np.dtype([("mv", "<i4", (2, 2)), ("ref_id", "i1", (2,)), ("pred_flag", "u1"), ("ciip_flag", "u1")])   # 20 bytes
PARTITION_DTYPE = np.dtype([("tb_off", "u1"), ("tb_size", "u1"), ("cu_off_x", "u1"), ("cu_off_y", "u1"),
                            ("ctb_off", "u1"), ("ctb_size", "u1"), ("cflags", "u1"), ("reserved", "u1")])          # 8 bytes
REF_MVF_DTYPE = np.dtype([("mv", "<i4", (2, 2)), ("ref_idx", "i1", (2,)), ("hpel_if_idx", "u1"), ("bcw_idx", "u1"),
                          ("pred_flag", "u1"), ("ciip_flag", "u1"), ("pad", "u1", (2,))])   # MvField, vvc_ctu.h:195-202 (24 bytes)
