"""TEST INFRASTRUCTURE: ctypes loader for oracle/_ref/libvvcref.so (the reference's own C code)."""
import ctypes as C
import os

import numpy as np

from ffmpeg_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libvvcref.so")


def available():
    return os.path.exists(SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(SO)
        _lib.vvcref_table.restype = C.c_void_p
        _lib.vvcref_table.argtypes = [C.c_char_p, C.POINTER(C.c_size_t)]
    return _lib


def table(name, dtype, shape):
    """Copy of a constant table exported by the reference's vvc_data.c."""
    n = C.c_size_t(0)
    p = lib().vvcref_table(name.encode(), C.byref(n))
    if not p:
        raise KeyError(name)
    a = np.frombuffer((C.c_uint8 * n.value).from_address(p), dtype=dtype).copy()
    return a.reshape(shape)


def itx(coeffs, trh, trv, log2w, log2h, nzw, nzh, bd=10, log2_range=15):
    """itx.itx[trh][trv][log2w][log2h] in place on an int32 array of w*h; returns False if slot empty."""
    assert coeffs.dtype == np.int32 and coeffs.flags.c_contiguous
    r = lib().vvcref_itx(bd, trh, trv, log2w, log2h, C.c_void_p(coeffs.ctypes.data), nzw, nzh, log2_range)
    return r == 0


def residual_batch(planes, levels, tbs, params, intra_mode=None, want_residual=False):
    """Reference itransform() over a TB batch on host planes (list of 3 uint16 arrays), in place."""
    pl = abi.host_planes(planes)
    levels = np.ascontiguousarray(levels, np.int16)
    tbs = np.ascontiguousarray(tbs, abi.TB_DTYPE)
    res = None
    if want_residual:
        total = int(np.sum(1 << (tbs["log2_w"].astype(np.int64) + tbs["log2_h"])))
        res = np.zeros(total, np.int32)
    im = None if intra_mode is None else np.ascontiguousarray(intra_mode, np.int8)
    r = lib().vvcref_residual_batch(
        C.byref(pl), C.c_void_p(levels.ctypes.data), C.c_void_p(tbs.ctypes.data), len(tbs), C.byref(params),
        C.c_void_p(im.ctypes.data) if im is not None else None,
        C.c_void_p(res.ctypes.data) if res is not None else None)
    if r != 0:
        raise RuntimeError("vvcref_residual_batch rejected a descriptor")
    return res


def inter_batch(dst, refs, blocks, params, lmcs_fwd_lut=None, want_refined_mv=False):
    """Reference predict_inter() over a batch of sub-block descriptors; dst/refs are lists of 3 uint16 arrays."""
    pl = abi.host_planes(dst)
    ra = abi.planes_array([abi.host_planes(r) for r in refs])
    blocks = np.ascontiguousarray(blocks, abi.PB_DTYPE)
    lut = None if lmcs_fwd_lut is None else np.ascontiguousarray(lmcs_fwd_lut, np.uint16)
    mv = np.zeros((len(blocks), 2, 2), np.int32) if want_refined_mv else None
    r = lib().vvcref_inter_batch(C.byref(pl), ra, len(refs), C.c_void_p(blocks.ctypes.data), len(blocks),
                                 C.byref(params), C.c_void_p(lut.ctypes.data) if lut is not None else None,
                                 C.c_void_p(mv.ctypes.data) if mv is not None else None)
    if r != 0:
        raise RuntimeError("vvcref_inter_batch rejected a descriptor")
    return mv


def _ptr_grid(tabs, shape):
    """ctypes array-of-pointers (nested) over a numpy array of tables whose last two axes are one table."""
    if len(shape) == 1:
        return (C.c_void_p * shape[0])(*[tabs[i].ctypes.data for i in range(shape[0])])
    inner = C.c_void_p * shape[1]
    return (inner * shape[0])(*[inner(*[tabs[i, j].ctypes.data for j in range(shape[1])]) for i in range(shape[0])])


def _row_chunks(n_rows, threads):
    threads = max(1, min(threads, n_rows))
    edges = [round(i * n_rows / threads) for i in range(threads + 1)]
    return [(edges[i], edges[i + 1]) for i in range(threads) if edges[i + 1] > edges[i]]


def _parallel(fn, chunks):
    """Run fn(ry0, ry1) for every CTB-row chunk on its own thread (ctypes drops the GIL) and join."""
    import threading
    if len(chunks) == 1:
        fn(*chunks[0])
        return
    ths = [threading.Thread(target=fn, args=c) for c in chunks]
    for t in ths:
        t.start()
    for t in ths:
        t.join()


def _ctb_rows(planes, ctb_log2):
    return (planes[0].shape[0] + (1 << ctb_log2) - 1) >> ctb_log2


def deblock_picture(planes, t, threads=1):
    """Reference ff_vvc_deblock_vertical/horizontal over the picture; `t` is synth.deblock_picture().
    threads > 1: CTB rows in parallel, with a join between the vertical-edge and the horizontal-edge pass."""
    pl = abi.host_planes(planes)
    for k in ("bs", "len_p", "len_q", "tb_w_c", "tb_h_c", "qp", "ctb_offsets"):
        assert t[k].flags.c_contiguous
    args = (C.byref(pl), _ptr_grid(t["bs"], (2, 3)), _ptr_grid(t["len_p"], (2,)),
            _ptr_grid(t["len_q"], (2,)), C.c_void_p(t["tb_w_c"].ctypes.data),
            C.c_void_p(t["tb_h_c"].ctypes.data), _ptr_grid(t["qp"], (3,)),
            C.c_void_p(t["ctb_offsets"].ctypes.data), C.byref(t["params"]))
    chunks = _row_chunks(_ctb_rows(planes, t["params"].ctb_log2_size), threads)
    for mask in (1, 2):
        _parallel(lambda a, b: lib().vvcref_deblock_picture(*args, mask, a, b), chunks)


def _plant_tiles(lf):
    """Tile grid of synth.loopfilter_ctbs(tile_cols=..., tile_rows=...) for the next SAO / ALF calls (or one tile)."""
    t = lf.get("tiles")
    if t is None:
        lib().vvcref_set_tiles(None, 0, None, 0, 1, 1)
        return
    col, row = np.ascontiguousarray(t["col_bd"], np.uint16), np.ascontiguousarray(t["row_bd"], np.uint16)
    lib().vvcref_set_tiles(C.c_void_p(col.ctypes.data), len(col), C.c_void_p(row.ctypes.data), len(row), int(t["num_tiles"]),
                           int(t["lfate"]))


def sao_picture(planes, lf, bit_depth=None, threads=1):
    """Reference SAO stage in place; `lf` is synth.loopfilter_ctbs().  Phase 1 saves every CTB's pre-SAO borders,
    phase 2 filters; each phase may run CTB rows on several threads."""
    bit_depth = lf.get("bd", 10) if bit_depth is None else bit_depth
    _plant_tiles(lf)
    pl = abi.host_planes(planes)
    sao = np.ascontiguousarray(lf["sao"])
    sl = np.ascontiguousarray(lf["slice_idx"], np.int16)
    l = lib()
    l.vvcref_sao_scratch_bytes.restype = C.c_size_t
    scratch = np.zeros(l.vvcref_sao_scratch_bytes(C.byref(pl), int(lf["ctb_log2"])), np.uint8)
    chunks = _row_chunks(_ctb_rows(planes, lf["ctb_log2"]), threads)
    for phase in (1, 2):
        _parallel(lambda a, b: l.vvcref_sao_picture(C.byref(pl), C.c_void_p(sao.ctypes.data), C.c_void_p(sl.ctypes.data),
                                                    int(lf["lfase"]), int(lf["ctb_log2"]), bit_depth,
                                                    C.c_void_p(scratch.ctypes.data), phase, a, b), chunks)


def alf_picture(planes, lf, threads=1):
    _plant_tiles(lf)
    pl = abi.host_planes(planes)
    alf = np.ascontiguousarray(lf["alf"])
    sl = np.ascontiguousarray(lf["slice_idx"], np.int16)
    p = np.ascontiguousarray(lf["alf_params"])
    l = lib()
    l.vvcref_alf_scratch_bytes.restype = C.c_size_t
    scratch = np.zeros(l.vvcref_alf_scratch_bytes(C.byref(pl), int(lf["ctb_log2"])), np.uint8)
    chunks = _row_chunks(_ctb_rows(planes, lf["ctb_log2"]), threads)
    for phase in (1, 2):
        _parallel(lambda a, b: l.vvcref_alf_picture(C.byref(pl), C.c_void_p(alf.ctypes.data), C.c_void_p(p.ctypes.data),
                                                    C.c_void_p(sl.ctypes.data), int(lf["lfase"]),
                                                    C.c_void_p(scratch.ctypes.data), phase, a, b), chunks)


def lmcs_picture(planes, inv_lut, ctb_log2=7, bit_depth=10, threads=1):
    pl = abi.host_planes(planes)
    lut = np.ascontiguousarray(inv_lut, np.uint16)
    chunks = _row_chunks(_ctb_rows(planes, ctb_log2), threads)
    _parallel(lambda a, b: lib().vvcref_lmcs_picture(C.byref(pl), C.c_void_p(lut.ctypes.data), ctb_log2, bit_depth, a, b),
              chunks)


def inter_cus(dst, refs, cus, sb_mv, params, lmcs_fwd_lut=None):
    """Reference predict_inter() over CU-level descriptions (synth.inter_cu_picture)."""
    pl = abi.host_planes(dst)
    ra = abi.planes_array([abi.host_planes(r) for r in refs])
    cus = np.ascontiguousarray(cus)
    sb = np.ascontiguousarray(sb_mv, np.int32)
    lut = None if lmcs_fwd_lut is None else np.ascontiguousarray(lmcs_fwd_lut, np.uint16)
    r = lib().vvcref_inter_cus(C.byref(pl), ra, len(refs), C.c_void_p(cus.ctypes.data), len(cus),
                               C.c_void_p(sb.ctypes.data) if sb.size else None, C.byref(params),
                               C.c_void_p(lut.ctypes.data) if lut is not None else None)
    if r != 0:
        raise RuntimeError("vvcref_inter_cus failed")


def recon_picture(planes, ref, intra_params, threads=1, inter=None, lmcs_fwd_lut=None):
    """Reference ff_vvc_reconstruct() over the whole picture; `ref` is synth.recon_picture()["ref"].
    threads > 1: CTU rows on that many host threads in the reference's wavefront order.
    inter: the picture's `inter` dict (refs, params), needed when the picture has CIIP units."""
    pl = abi.host_planes(planes)
    cus, tus = np.ascontiguousarray(ref["cus"]), np.ascontiguousarray(ref["tus"])
    tbs = np.ascontiguousarray(ref["tbs"], abi.TB_DTYPE)
    lv = np.ascontiguousarray(ref["levels"], np.int16)
    ra, n_refs = None, 0
    if inter is not None:
        ra, n_refs = abi.planes_array([abi.host_planes(r) for r in inter["refs"]]), len(inter["refs"])
    lut = None if lmcs_fwd_lut is None else np.ascontiguousarray(lmcs_fwd_lut, np.uint16)
    r = lib().vvcref_recon_picture(C.byref(pl), C.c_void_p(cus.ctypes.data), len(cus), C.c_void_p(tus.ctypes.data),
                               C.c_void_p(tbs.ctypes.data), C.c_void_p(lv.ctypes.data), C.byref(ref["params"]),
                               C.byref(intra_params), int(threads), ra, n_refs,
                               C.byref(inter["params"]) if inter is not None else None,
                               C.c_void_p(lut.ctypes.data) if lut is not None else None)
    if r != 0:
        raise RuntimeError("vvcref_recon_picture failed")


def deblock_bs(st, width, height, ctb_log2, slice_idx=None, lfase=1):
    """Reference vvc_deblock_bs (vvc_filter.c:756-782 and everything it calls) over synth.deblock_state() tables.
    Returns bs [2][3][r4][s4], len_p [2][r4][s4], len_q [2][r4][s4] as the reference leaves fc->tab.*_bs / *_p / *_q."""
    r4, s4 = st["rows4"], st["stride4"]
    bs = np.zeros((2, 3, r4, s4), np.uint8)
    len_p, len_q = np.zeros((2, r4, s4), np.uint8), np.zeros((2, r4, s4), np.uint8)
    keep = {k: np.ascontiguousarray(st[k]) for k in ("mvf", "tb_pos_x0", "tb_pos_y0", "tb_width", "tb_height", "cb_pos_x", "cb_pos_y",
                                                      "cb_width", "cb_height", "tu_coded_flag", "tu_joint", "pcmf", "msf", "iaf", "ref_poc")}
    sl = None if slice_idx is None else np.ascontiguousarray(slice_idx, np.int16)
    p = lambda a: C.c_void_p(a.ctypes.data)
    _plant_tiles({})
    r = lib().vvcref_deblock_bs(int(width), int(height), int(ctb_log2), p(keep["mvf"]),
                                _ptr_grid(keep["tb_pos_x0"], (2,)), _ptr_grid(keep["tb_pos_y0"], (2,)),
                                _ptr_grid(keep["tb_width"], (2,)), _ptr_grid(keep["tb_height"], (2,)),
                                p(keep["cb_pos_x"]), p(keep["cb_pos_y"]), p(keep["cb_width"]), p(keep["cb_height"]),
                                _ptr_grid(keep["tu_coded_flag"], (3,)), p(keep["tu_joint"]), _ptr_grid(keep["pcmf"], (2,)),
                                p(keep["msf"]), p(keep["iaf"]), p(keep["ref_poc"]), p(sl) if sl is not None else None, int(lfase),
                                _ptr_grid(bs, (2, 3)), _ptr_grid(len_p, (2,)), _ptr_grid(len_q, (2,)))
    if r != 0:
        raise RuntimeError("vvcref_deblock_bs failed")
    return bs, len_p, len_q


def emit_inter_cus(width, height, cus, sb_mv, params, order=False):
    """synth's CU-level descriptions -> the reference's CodingUnit / PredictionUnit / MvField state (as
    inter_cus builds it) -> the product's emitter (ffmpeg_b200/emit/vvcb200_emit.c) -> (blocks, prof, params copy)."""
    cus = np.ascontiguousarray(cus)
    sb = np.ascontiguousarray(sb_mv, np.int32)
    cap = int(sum((int(c["w"]) // 4) * (int(c["h"]) // 4) * 2 for c in cus)) + 16
    blocks = np.zeros(cap, abi.PB_DTYPE)
    prof = np.zeros(len(cus) + 1, abi.PROF_DTYPE)
    nb, npf = C.c_int(0), C.c_int(0)
    p = abi.InterParams.from_buffer_copy(bytes(params))
    r = lib().vvcref_emit_inter_cus(int(width), int(height), C.c_void_p(cus.ctypes.data), len(cus),
                                    C.c_void_p(sb.ctypes.data) if sb.size else None, C.byref(p),
                                    C.c_void_p(blocks.ctypes.data), cap, C.byref(nb), C.c_void_p(prof.ctypes.data), len(prof),
                                    C.byref(npf), int(order))
    if r != 0:
        raise RuntimeError("vvcref_emit_inter_cus failed: %d" % r)
    return blocks[:nb.value].copy(), prof[:npf.value].copy(), p


def emit_recon_picture(width, height, ref, intra_params, inter=None, fixed=False, shrink=0):
    """synth's CU / TU arrays -> the reference's CTU / CodingUnit / TransformUnit / TransformBlock lists and
    per-min-CB tables (as recon_picture builds them) -> the product's RECON emitter (ffmpeg_b200/emit/vvcb200_emit_recon.h)
    -> (tbs, levels, intra blocks, ctu_first).  `inter`: the picture's inter dict when it has CIIP units."""
    cus, tus = np.ascontiguousarray(ref["cus"]), np.ascontiguousarray(ref["tus"])
    in_tbs = np.ascontiguousarray(ref["tbs"], abi.TB_DTYPE)
    in_lv = np.ascontiguousarray(ref["levels"], np.int16)
    n_ctus = ((width + (1 << intra_params.ctb_log2_size) - 1) >> intra_params.ctb_log2_size) * \
             ((height + (1 << intra_params.ctb_log2_size) - 1) >> intra_params.ctb_log2_size)
    # fixed: the emitter writes straight into these arrays (VVCB200EmitBatch.fixed) instead of growing its own; shrink: make
    # the TB array that much too small
    tbs = np.zeros(len(in_tbs) + 1 - shrink, abi.TB_DTYPE)
    lv = np.zeros(len(in_lv) + 1, np.int16)
    ibs = np.zeros(4 * len(cus) + 16, abi.IB_DTYPE)
    first = np.zeros(n_ctus + 2, np.uint32)
    nt, nl, ni, nc = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
    p = lambda a: C.c_void_p(a.ctypes.data)
    r = lib().vvcref_emit_recon_picture(int(width), int(height), p(cus), len(cus), p(tus), p(in_tbs), p(in_lv),
                                        C.byref(ref["params"]), C.byref(intra_params), int(inter is not None) | (2 if fixed else 0),
                                        C.byref(inter["params"]) if inter is not None else None,
                                        p(tbs), len(tbs), C.byref(nt), p(lv), len(lv), C.byref(nl),
                                        p(ibs), len(ibs), C.byref(ni), p(first), len(first), C.byref(nc))
    if r != 0:
        raise RuntimeError("vvcref_emit_recon_picture failed: %d" % r)
    return tbs[:nt.value].copy(), lv[:nl.value].copy(), ibs[:ni.value].copy(), first[:nc.value + 1].copy()


def emit_deblock_units(st, width, height, ctb_log2, slice_idx=None, lfase=1):
    """synth.deblock_state()'s tables in the reference's format (fc->tab.mvf, tb_pos_*, cb_pos_*, ...) ->
    vvcb200_emit_filters_ctu -> (motion units, partition units, per-CTB no-deblock flags)."""
    r4, s4 = st["rows4"], st["stride4"]
    cw, ch = (width + (1 << ctb_log2) - 1) >> ctb_log2, (height + (1 << ctb_log2) - 1) >> ctb_log2
    motion = np.zeros((r4, s4), abi.MOTION_DTYPE)
    part = np.zeros((r4, s4), abi.PARTITION_DTYPE)
    edges = np.zeros(cw * ch, np.uint8)
    keep = {k: np.ascontiguousarray(st[k]) for k in ("mvf", "tb_pos_x0", "tb_pos_y0", "tb_width", "tb_height", "cb_pos_x", "cb_pos_y",
                                                      "cb_width", "cb_height", "tu_coded_flag", "tu_joint", "pcmf", "msf", "iaf", "ref_poc")}
    sl = None if slice_idx is None else np.ascontiguousarray(slice_idx, np.int16)
    p = lambda a: C.c_void_p(a.ctypes.data)
    _plant_tiles({})
    r = lib().vvcref_emit_deblock_units(int(width), int(height), int(ctb_log2), p(keep["mvf"]),
                                        _ptr_grid(keep["tb_pos_x0"], (2,)), _ptr_grid(keep["tb_pos_y0"], (2,)),
                                        _ptr_grid(keep["tb_width"], (2,)), _ptr_grid(keep["tb_height"], (2,)),
                                        p(keep["cb_pos_x"]), p(keep["cb_pos_y"]), p(keep["cb_width"]), p(keep["cb_height"]),
                                        _ptr_grid(keep["tu_coded_flag"], (3,)), p(keep["tu_joint"]), _ptr_grid(keep["pcmf"], (2,)),
                                        p(keep["msf"]), p(keep["iaf"]), p(keep["ref_poc"]), p(sl) if sl is not None else None,
                                        int(lfase), p(motion), p(part), p(edges))
    if r != 0:
        raise RuntimeError("vvcref_emit_deblock_units failed")
    return motion, part, edges


def emit_loopfilter_ctbs(t, width, height, ctb_offsets, deblock_params):
    """synth.loopfilter_ctbs() parameters planted as SAOParams / ALFParams / VVCALF APS list / DBParams ->
    vvcb200_emit_filters_ctu + vvcb200_emit_alf_params + vvcb200_emit_deblock_params -> the engine's records."""
    n = len(t["sao"])
    sao_in, alf_in = np.ascontiguousarray(t["sao"]), np.ascontiguousarray(t["alf"])
    ap_in = np.ascontiguousarray(t["alf_params"])
    off_in = np.ascontiguousarray(ctb_offsets, np.int8)
    sl = np.ascontiguousarray(t["slice_idx"], np.int16)
    sao, alf = np.zeros(n, abi.SAO_DTYPE), np.zeros(n, abi.ALF_CTB_DTYPE)
    ap = np.zeros(1, abi.ALF_PARAMS_DTYPE)
    off, edges = np.zeros((n, 6), np.int8), np.zeros(n, np.uint8)
    dp = abi.DeblockParams()
    p = lambda a: C.c_void_p(a.ctypes.data)
    _plant_tiles(t)
    r = lib().vvcref_emit_loopfilter_ctbs(int(width), int(height), int(t["ctb_log2"]), int(t["bd"]), p(sao_in), p(alf_in), p(ap_in),
                                          p(sl), int(t["lfase"]), p(off_in), C.byref(deblock_params),
                                          p(sao), p(alf), p(ap), p(off), p(edges), C.byref(dp))
    _plant_tiles({})
    if r != 0:
        raise RuntimeError("vvcref_emit_loopfilter_ctbs failed")
    return sao, alf, ap, off, edges, dp


def emit_store_refined_check(dst, refs, blocks, params):
    """the reference predicts the batch (filling fc->ref->tab_dmvr_mvf through set_dmvr_info), then
    vvcb200_emit_store_refined_mv writes the same refined vectors into a fresh table; returns the number of differing units."""
    pl = abi.host_planes(dst)
    ra = abi.planes_array([abi.host_planes(r) for r in refs])
    blocks = np.ascontiguousarray(blocks, abi.PB_DTYPE)
    mv = np.zeros((len(blocks), 2, 2), np.int32)
    return lib().vvcref_emit_store_refined_check(C.byref(pl), ra, len(refs), C.c_void_p(blocks.ctypes.data), len(blocks),
                                                 C.byref(params), C.c_void_p(mv.ctypes.data)), mv


def emit_last_ms():
    """Host milliseconds the last emit_* call spent inside the product's emitter (not in the harness around it)."""
    f = lib().vvcref_emit_last_ns
    f.restype = C.c_int64
    return f() / 1e6


# ---- the checkasm-style shim (oracle/shimlib/dsp_shim.c -> oracle/_ref/libvvcshim.so) ---------------------------------
SHIM_FAMILIES = ("itx.itx", "itx.add_residual", "lmcs.filter", "inter.put", "inter.put_uni", "inter.put_uni_w")
_shim = None


def shim_available():
    return os.path.exists(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libvvcshim.so"))


def shim():
    global _shim
    if _shim is None:
        _shim = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libvvcshim.so"))
        _shim.vvcshim_report.restype = C.c_char_p
    return _shim


def shim_check(bit_depth, seed=1):
    """vvcshim_check: the reference's table against vvcb200_dsp_init()'s over the reference's own checkasm shapes.
    Returns ({family: (members, calls, differing calls)}, device errors, calls the device cannot represent)."""
    out = (C.c_int * (3 * len(SHIM_FAMILIES)))()
    r = shim().vvcshim_check(int(bit_depth), int(seed), out)
    res = {f: (out[3 * i], out[3 * i + 1], out[3 * i + 2]) for i, f in enumerate(SHIM_FAMILIES)}
    return res, r & 0xffff, r >> 16
