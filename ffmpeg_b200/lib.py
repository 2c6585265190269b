"""ctypes binding of libvvcb200.so (the C ABI in include/vvc_b200.h).

There is deliberately no fallback: if the shared library is missing, or CUDA is unusable, calls raise.
Device memory in tests/bench is plain torch CUDA tensors; only raw pointers cross the ABI.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "libvvcb200.so")
_lib = None


class VVCB200Error(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            raise VVCB200Error("%s is missing: run `python -m ffmpeg_b200.build` (no CPU fallback exists)" % SO)
        _lib = C.CDLL(SO)
        _lib.vvcb200_last_error.restype = C.c_char_p
        _lib.vvcb200_launch_count.restype = C.c_uint64
        if _lib.vvcb200_abi_version() != 2:
            raise VVCB200Error("ABI version mismatch")
    return _lib


def check(rc):
    if rc != 0:
        raise VVCB200Error("vvcb200 error %d: %s" % (rc, load().vvcb200_last_error().decode()))


def launch_count():
    return int(load().vvcb200_launch_count())


def _ptr(t):
    """Raw device/host pointer of a torch tensor, numpy array or int."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    if isinstance(t, np.ndarray):
        return C.c_void_p(t.ctypes.data)
    return C.c_void_p(int(t))


def _stream(stream):
    if stream is None:
        import torch
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return C.c_void_p(int(stream))


def device_planes(tensors):
    """VVCB200Planes over three 2-D torch CUDA tensors of dtype int16/uint16 (row stride from the tensor)."""
    p = abi.Planes()
    for c, t in enumerate(tensors):
        assert t.is_cuda and t.dim() == 2 and t.element_size() == 2 and t.stride(1) == 1
        p.plane[c] = t.data_ptr()
        p.stride[c] = t.stride(0) * 2
        p.width[c] = t.shape[1]
        p.height[c] = t.shape[0]
    return p


def pack_levels_sparse(levels, tbs):
    """vvcb200_pack_levels_sparse (host, no CUDA): dense int16 level buffer + TB records -> (sparse buffer, TB records with
    coeff_off into it).  Pair with a ResidualParams whose levels_sparse is 1 (sparse_params)."""
    levels = np.ascontiguousarray(levels, np.int16)
    out_tbs = np.ascontiguousarray(tbs, abi.TB_DTYPE).copy()
    cap = levels.size + 2 * len(out_tbs) + 2 * (levels.size // 32) + 64
    out = np.zeros(cap, np.int16)
    n = C.c_size_t(0)
    check(load().vvcb200_pack_levels_sparse(_ptr(levels), _ptr(out_tbs), len(out_tbs), _ptr(out), C.c_size_t(cap), C.byref(n)))
    return out[:n.value].copy(), out_tbs


def pack_levels(levels, tbs, sparse=False, narrow=True):
    """vvcb200_pack_levels (host, no CUDA): dense int16 level buffer + TB records -> (packed buffer, TB records with coeff_off
    into it and TB_LEVELS8 on the blocks stored as int8).  sparse: pair with sparse_params()."""
    levels = np.ascontiguousarray(levels, np.int16)
    out_tbs = np.ascontiguousarray(tbs, abi.TB_DTYPE).copy()
    cap = levels.size + 4 * len(out_tbs) + 2 * (levels.size // 32) + 64
    out = np.zeros(cap, np.int16)
    n = C.c_size_t(0)
    check(load().vvcb200_pack_levels(_ptr(levels), _ptr(out_tbs), len(out_tbs), int(bool(sparse)), int(bool(narrow)), _ptr(out),
                                     C.c_size_t(cap), C.byref(n)))
    return out[:n.value].copy(), out_tbs


def sparse_params(params):
    p = abi.ResidualParams.from_buffer_copy(bytes(params))
    p.levels_sparse = 1
    return p


def residual_batch(dst, levels, tbs, n, params, residual=None, stream=None):
    """vvcb200_residual_batch: dst/residual are device_planes() structs (or None), levels/tbs device tensors."""
    check(load().vvcb200_residual_batch(
        C.byref(dst) if dst is not None else None, C.byref(residual) if residual is not None else None,
        _ptr(levels), _ptr(tbs), int(n), C.byref(params), _stream(stream)))


def dsp_itx_batch(coeffs, block_off, tbs, n, log2_transform_range=15, bit_depth=10, stream=None):
    check(load().vvcb200_dsp_itx_batch(_ptr(coeffs), _ptr(block_off), _ptr(tbs), int(n),
                                       int(log2_transform_range), int(bit_depth), _stream(stream)))


def dsp_add_residual_batch(dst, res, block_off, tbs, n, bit_depth=10, stream=None):
    check(load().vvcb200_dsp_add_residual_batch(C.byref(dst), _ptr(res), _ptr(block_off), _ptr(tbs), int(n),
                                                int(bit_depth), _stream(stream)))


def intra_scratch_bytes(width, height):
    f = load().vvcb200_intra_scratch_bytes
    f.restype = C.c_size_t
    return int(f(int(width), int(height)))


INTRA_ORDER_DIAGONAL, INTRA_ORDER_DEPTH = 0, 1


def intra_schedule(blocks, ctu_first, ctus_w, ctus_h, ctb_log2, width, height, want_levels=False, order=INTRA_ORDER_DEPTH,
                   phases=False):
    """vvcb200_intra_schedule: per-CTU decoding-order records -> wavefront / dependency-depth order (host helper).
    phases=True: the records come back grouped luma / chroma add-only / other chroma, with (n_luma, n_add_only) as the
    last element of the result (what VVCB200IntraParams.n_luma / n_add_only want)."""
    blocks = np.ascontiguousarray(blocks, abi.IB_DTYPE)
    ctu_first = np.ascontiguousarray(ctu_first, np.uint32)
    assert len(ctu_first) == ctus_w * ctus_h + 1 and ctu_first[-1] == len(blocks)
    out = np.zeros(len(blocks), abi.IB_DTYPE)
    levels = C.c_int32(0)
    counts = (C.c_int32 * 2)(0, 0)
    check(load().vvcb200_intra_schedule(_ptr(blocks), _ptr(ctu_first), int(ctus_w), int(ctus_h), int(ctb_log2), int(width),
                                        int(height), int(order), _ptr(out), C.byref(levels), counts if phases else None))
    res = (out, int(levels.value)) if want_levels else (out,)
    if phases:
        res = res + ((int(counts[0]), int(counts[1])),)
    return res if len(res) > 1 else res[0]


def intra_schedule_device(d_in, d_ctu_first, n, ctus_w, ctus_h, d_out, d_counts, d_scratch, stream=None):
    """vvcb200_intra_schedule_device: all arguments device tensors (records as emitted + CTU ranges in, phased diagonal
    order + the three phase counts out)."""
    check(load().vvcb200_intra_schedule_device(_ptr(d_in), _ptr(d_ctu_first), int(n), int(ctus_w), int(ctus_h), _ptr(d_out),
                                               _ptr(d_counts), _ptr(d_scratch), _stream(stream)))


def intra_schedule_scratch_bytes(n, ctus_w, ctus_h):
    f = load().vvcb200_intra_schedule_scratch_bytes
    f.restype = C.c_size_t
    return int(f(int(n), int(ctus_w), int(ctus_h)))


def phased_params(params, counts):
    """Copy of an IntraParams with the phase counts of intra_schedule(..., phases=True)."""
    p = abi.IntraParams.from_buffer_copy(bytes(params))
    p.n_luma, p.n_add_only = counts
    return p


def intra_picture(pic, residual, blocks, n, params, scratch, stream=None, shadow=None):
    """vvcb200_intra_picture: pic/residual/shadow are device_planes() structs, blocks/scratch device tensors."""
    check(load().vvcb200_intra_picture(C.byref(pic), C.byref(residual) if residual is not None else None,
                                       C.byref(shadow) if shadow is not None else None,
                                       _ptr(blocks), int(n), C.byref(params), _ptr(scratch), _stream(stream)))


def intra_prepare(pic, residual, blocks, n, params, scratch, stream=None, shadow=None):
    """vvcb200_intra_prepare: the pre-pass of intra_picture alone (then call intra_picture with params.prepared = 1)."""
    check(load().vvcb200_intra_prepare(C.byref(pic), C.byref(residual) if residual is not None else None,
                                       C.byref(shadow) if shadow is not None else None,
                                       _ptr(blocks), int(n), C.byref(params), _ptr(scratch), _stream(stream)))


def intra_picture_static(pic, residual, blocks, n, params, stream=None):
    """vvcb200_intra_picture_static: as intra_picture, marks in the library's built-in scratch (<= 1920x1088)."""
    check(load().vvcb200_intra_picture_static(C.byref(pic), C.byref(residual) if residual is not None else None,
                                              _ptr(blocks), int(n), C.byref(params), _stream(stream)))


# ---- Level B: picture engine ---------------------------------------------------------------------------
class Arena:
    """Bump allocator over a block from Context.arena()."""

    def __init__(self, base, nbytes):
        self.base, self.nbytes, self.used = base, nbytes, 0

    def take(self, shape, dtype):
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        off = (self.used + 255) & ~255
        if off + n > self.nbytes:
            raise VVCB200Error("arena of %d bytes is full" % self.nbytes)
        self.used = off + n
        buf = (C.c_uint8 * max(n, 1)).from_address(self.base + off)
        return np.frombuffer(buf, dtype=np.uint8, count=n).view(dtype).reshape(shape)

    def copy(self, a):
        out = self.take(a.shape, a.dtype)
        out[...] = a
        return out


class Context:
    """VVCB200Ctx wrapper: one per GPU per stream of pictures.  Host buffers are numpy arrays; use
    pinned() to allocate page-locked ones (needed for asynchronous copies)."""

    def __init__(self, device, width, height, bit_depth=10, chroma_format_idc=1, dpb_size=3, validate=2):
        self._h = C.c_void_p()
        check(load().vvcb200_ctx_create(C.byref(self._h), int(device), int(width), int(height),
                                        int(chroma_format_idc), int(bit_depth), int(dpb_size)))
        check(load().vvcb200_ctx_set_validation(self._h, int(validate)))    # tests check every record; see the header
        self.width, self.height, self.bit_depth, self.dpb_size = width, height, bit_depth, dpb_size
        self._pinned = []
        self._keep = []

    def close(self):
        if self._h:
            load().vvcb200_ctx_destroy(self._h)
            self._h = C.c_void_p()
        for p in self._pinned:
            load().vvcb200_host_free(C.c_void_p(p))
        self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def pinned(self, shape, dtype):
        """Page-locked numpy array owned by this context."""
        l = load()
        l.vvcb200_host_alloc.restype = C.c_void_p
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = l.vvcb200_host_alloc(C.c_size_t(max(n, 1)))
        if not p:
            raise VVCB200Error(l.vvcb200_last_error().decode())
        self._pinned.append(p)
        buf = (C.c_uint8 * max(n, 1)).from_address(p)
        return np.frombuffer(buf, dtype=np.uint8, count=n).view(dtype).reshape(shape)

    def arena(self, nbytes):
        """vvcb200_ctx_arena_alloc: a pinned block mirrored on the device (freed with the context); returns an Arena whose
        take() / copy() carve 256-byte aligned numpy arrays out of it in call order."""
        l = load()
        l.vvcb200_ctx_arena_alloc.restype = C.c_void_p
        p = l.vvcb200_ctx_arena_alloc(self._h, C.c_size_t(int(nbytes)))
        if not p:
            raise VVCB200Error(l.vvcb200_last_error().decode())
        return Arena(p, int(nbytes))

    def pinned_frame(self):
        """The three planes of a 4:2:0 frame back to back in ONE pinned allocation (a download is then one copy)."""
        w, h = self.width, self.height
        flat = self.pinned((w * h * 3 // 2,), np.uint16)
        y, cb = flat[:w * h].reshape(h, w), flat[w * h:w * h * 5 // 4].reshape(h // 2, w // 2)
        return [y, cb, flat[w * h * 5 // 4:].reshape(h // 2, w // 2)]

    def pinned_copy(self, a):
        out = self.pinned(a.shape, a.dtype)
        out[...] = a
        return out

    def pinned_planes(self):
        dims = [(self.width, self.height), (self.width // 2, self.height // 2), (self.width // 2, self.height // 2)]
        return [self.pinned((h, w), np.uint16) for (w, h) in dims]

    def stream(self):
        l = load()
        l.vvcb200_ctx_stream.restype = C.c_void_p
        return l.vvcb200_ctx_stream(self._h)

    def planes(self, slot=-1):
        p = abi.Planes()
        check(load().vvcb200_ctx_planes(self._h, int(slot), C.byref(p)))
        return p

    def sync(self):
        check(load().vvcb200_ctx_sync(self._h))
        self._keep = []

    def begin(self, slot=0):
        check(load().vvcb200_pic_begin(self._h, int(slot)))

    @staticmethod
    def _plane_args(planes):
        ptrs = (C.c_void_p * 3)(*[p.ctypes.data for p in planes])
        strides = (C.c_int32 * 3)(*[p.strides[0] for p in planes])
        return ptrs, strides

    def upload(self, planes, slot=-1):
        ptrs, strides = self._plane_args(planes)
        check(load().vvcb200_pic_upload(self._h, int(slot), ptrs, strides))

    def download(self, planes, slot=-1):
        ptrs, strides = self._plane_args(planes)
        check(load().vvcb200_pic_download(self._h, int(slot), ptrs, strides))

    def export(self, slot=-1, window=None):
        """vvcb200_pic_export: (Planes of the window in device memory, cudaEvent_t handle)."""
        w = abi.Window(*window) if window is not None else None
        out = abi.Export()
        check(load().vvcb200_pic_export(self._h, int(slot), C.byref(w) if w is not None else None, C.byref(out)))
        return out.planes, out.ready

    def release(self, slot=-1, done_event=None):
        check(load().vvcb200_pic_release(self._h, int(slot), C.c_void_p(done_event) if done_event else None))

    def download_window(self, planes, window, slot=-1):
        ptrs, strides = self._plane_args(planes)
        w = abi.Window(*window)
        check(load().vvcb200_pic_download_window(self._h, int(slot), C.byref(w), ptrs, strides))

    def inter(self, blocks, params, lmcs_fwd_lut=None, refined_mv=None, prof=None):
        assert blocks.dtype == abi.PB_DTYPE
        check(load().vvcb200_pic_inter(self._h, _ptr(blocks), len(blocks), C.byref(params), _ptr(lmcs_fwd_lut),
                                       _ptr(refined_mv), _ptr(prof) if prof is not None and len(prof) else None,
                                       len(prof) if prof is not None else 0))

    def residual(self, levels, tbs, params):
        assert levels.dtype == np.int16 and tbs.dtype == abi.TB_DTYPE
        check(load().vvcb200_pic_residual(self._h, _ptr(levels), C.c_size_t(levels.size), _ptr(tbs), len(tbs),
                                          C.byref(params)))

    def intra(self, blocks, ctu_first, ctus_w, ctus_h, params):
        blocks = np.ascontiguousarray(blocks, abi.IB_DTYPE)
        ctu_first = np.ascontiguousarray(ctu_first, np.uint32)
        check(load().vvcb200_pic_intra(self._h, _ptr(blocks), _ptr(ctu_first), int(ctus_w), int(ctus_h), C.byref(params)))

    def lmcs(self, inv_lut):
        check(load().vvcb200_pic_lmcs(self._h, _ptr(inv_lut)))

    def deblock(self, edge, qp, ctb_offsets, stride4, params):
        t = deblock_tables(edge, qp, ctb_offsets, stride4)
        self._keep.append(t)
        check(load().vvcb200_pic_deblock(self._h, C.byref(t), C.byref(params)))

    def sao_alf(self, sao, alf, alf_params, ctb_log2):
        check(load().vvcb200_pic_sao_alf(self._h, _ptr(sao), _ptr(alf), _ptr(alf_params), int(ctb_log2)))


def inter_batch(dst, refs, blocks, n, params, lmcs_fwd_lut=None, tmp14=None, refined_mv=None, prof=None, stream=None):
    """vvcb200_inter_batch: dst/tmp14 device_planes() structs, refs list of device_planes(), blocks device tensor."""
    ra = abi.planes_array(refs)
    check(load().vvcb200_inter_batch(
        C.byref(dst) if dst is not None else None, ra, len(refs), _ptr(blocks), int(n), C.byref(params),
        _ptr(lmcs_fwd_lut), C.byref(tmp14) if tmp14 is not None else None, _ptr(refined_mv), _ptr(prof), _stream(stream)))


def _ptr_grid(tabs, shape):
    if len(shape) == 1:
        return (C.c_void_p * shape[0])(*[tabs[i].ctypes.data for i in range(shape[0])])
    inner = C.c_void_p * shape[1]
    return (inner * shape[0])(*[inner(*[tabs[i, j].ctypes.data for j in range(shape[1])]) for i in range(shape[0])])


def pack_deblock_edges(bs, len_p, len_q, tb_w_c, tb_h_c, ctb_log2, out=None):
    """vvcb200_pack_deblock_edges (host, no CUDA): reference-format tables -> packed edge bytes [2][3][rows4][stride4]."""
    rows4, stride4 = bs.shape[-2:]
    if out is None:
        out = np.zeros((2, 3, rows4, stride4), np.uint8)
    check(load().vvcb200_pack_deblock_edges(_ptr_grid(out, (2, 3)), _ptr_grid(bs, (2, 3)), _ptr_grid(len_p, (2,)),
                                            _ptr_grid(len_q, (2,)), _ptr(tb_w_c), _ptr(tb_h_c), stride4, rows4, ctb_log2))
    return out


def deblock_tables(edge, qp, ctb_offsets, stride4):
    """DeblockTables struct over device tensors: edge [2][3][r4][s4] uint8, qp [3][r4][s4] int8, ctb_offsets [n][6] int8."""
    t = abi.DeblockTables()
    for d in range(2):
        for c in range(3):
            t.edge[d][c] = edge[d][c].data_ptr() if hasattr(edge, "data_ptr") else edge[d, c].ctypes.data
    for c in range(3):
        t.qp[c] = qp[c].data_ptr() if hasattr(qp, "data_ptr") else qp[c].ctypes.data
    t.ctb_offsets = ctb_offsets.data_ptr() if hasattr(ctb_offsets, "data_ptr") else ctb_offsets.ctypes.data
    t.stride4 = stride4
    return t


def deblock_derive_edges(edge_out, motion, part, ctb_edges, width, height, stride4, ctb_log2, stream=None):
    """vvcb200_deblock_derive_edges: edge_out a device tensor [2][3][r4][s4] uint8; motion / part device tensors of
    VVCB200MotionUnit / VVCB200PartitionUnit records; ctb_edges a device tensor (one byte per CTB) or None."""
    inner = C.c_void_p * 3
    ptrs = (inner * 2)(*[inner(*[edge_out[d][c].data_ptr() for c in range(3)]) for d in range(2)])
    check(load().vvcb200_deblock_derive_edges(ptrs, _ptr(motion), _ptr(part), _ptr(ctb_edges), int(width), int(height),
                                              int(stride4), int(ctb_log2), _stream(stream)))


def deblock_picture(pic, tables, params, stream=None):
    check(load().vvcb200_deblock_picture(C.byref(pic), C.byref(tables), C.byref(params), _stream(stream)))


def deblock_lmcs_picture(dst, src, tables, params, inv_lut=None, stream=None):
    """vvcb200_deblock_lmcs_picture: both deblocking passes (and the inverse LMCS mapping of luma) in one out-of-place pass."""
    check(load().vvcb200_deblock_lmcs_picture(C.byref(dst), C.byref(src), C.byref(tables), C.byref(params), _ptr(inv_lut),
                                              _stream(stream)))


def deblock_lmcs_planes(dst, src, tables, params, inv_lut=None, planes=3, stream=None):
    """vvcb200_deblock_lmcs_planes: the tile pass for luma (1), chroma (2) or all planes (3)."""
    check(load().vvcb200_deblock_lmcs_planes(C.byref(dst), C.byref(src), C.byref(tables), C.byref(params), _ptr(inv_lut),
                                             int(planes), _stream(stream)))


def sao_picture(dst, src, ctbs, ctb_log2, bit_depth=10, stream=None):
    """vvcb200_sao_picture: dst/src device_planes(); ctbs device tensor of VVCB200SaoCtb records."""
    check(load().vvcb200_sao_picture(C.byref(dst), C.byref(src), _ptr(ctbs), int(ctb_log2), int(bit_depth), _stream(stream)))


def alf_picture(dst, src, ctbs, params, ctb_log2, bit_depth=10, stream=None):
    """vvcb200_alf_picture: ctbs / params are DEVICE tensors (VVCB200AlfCtb records, one VVCB200AlfParams)."""
    check(load().vvcb200_alf_picture(C.byref(dst), C.byref(src), _ptr(ctbs), _ptr(params), int(ctb_log2), int(bit_depth),
                                     _stream(stream)))


def sao_alf_scratch_bytes():
    f = load().vvcb200_sao_alf_scratch_bytes
    f.restype = C.c_size_t
    return int(f())


def sao_alf_picture(dst, src, sao, alf, params, ctb_log2, bit_depth=10, scratch=None, stream=None):
    """vvcb200_sao_alf_picture: SAO + ALF (+ CC-ALF) in one pass; sao / alf / params device tensors or None; scratch: a
    device tensor of sao_alf_scratch_bytes() bytes (the picture's coefficient table) or None."""
    check(load().vvcb200_sao_alf_picture(C.byref(dst), C.byref(src), _ptr(sao), _ptr(alf), _ptr(params), int(ctb_log2),
                                         int(bit_depth), _ptr(scratch), _stream(stream)))


def lmcs_picture(pic, lut, bit_depth=10, stream=None):
    check(load().vvcb200_lmcs_picture(C.byref(pic), _ptr(lut), int(bit_depth), _stream(stream)))
