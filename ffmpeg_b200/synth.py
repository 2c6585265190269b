"""Synthetic "parsed picture" generator (host side, numpy): the descriptor batches a VVC parser would emit.

There are no VVC bitstreams offline, so BASELINE.json's configs are exercised with seeded
synthetic pictures: random CU/TB partitions, levels, QPs, transform types, motion, filter parameters.
The same arrays feed the CUDA engine, the oracle port and the compiled reference, so parity is checked on
identical inputs.  PRNG: numpy PCG64 with the seed given by the caller (printed by tests/bench).
"""
import numpy as np

from . import abi


def plane_dims(width, height, chroma_shift=1):
    return [(width, height), (width >> chroma_shift, height >> chroma_shift), (width >> chroma_shift, height >> chroma_shift)]


def alloc_planes_np(width, height, align=64):
    """Three uint16 planes with rows padded to `align` samples."""
    out = []
    for w, h in plane_dims(width, height):
        stride = (w + align - 1) // align * align
        out.append(np.zeros((h, stride), np.uint16)[:, :w])
    return out


def _split_region(rng, x, y, w, h, pw, ph, min_size, p_split, out):
    """Random quad/binary partition of (x, y, w, h) clipped to the pw x ph plane."""
    if x >= pw or y >= ph:
        return
    over_x, over_y = x + w > pw, y + h > ph
    can_h, can_v = h > min_size, w > min_size
    if over_x or over_y:
        if over_x and over_y and can_h and can_v:
            mode = 0
        elif over_x and can_v:
            mode = 2
        elif over_y and can_h:
            mode = 1
        else:
            return  # cannot fit (plane dims are multiples of min_size, so not reached)
    else:
        if not (can_h or can_v) or rng.random() >= p_split(w, h):
            out.append((x, y, w, h))
            return
        r = rng.random()
        if can_h and can_v:
            mode = 0 if r < 0.4 else (1 if r < 0.7 else 2)
        else:
            mode = 1 if can_h else 2
    if mode == 0:
        for dy in (0, h // 2):
            for dx in (0, w // 2):
                _split_region(rng, x + dx, y + dy, w // 2, h // 2, pw, ph, min_size, p_split, out)
    elif mode == 1:
        _split_region(rng, x, y, w, h // 2, pw, ph, min_size, p_split, out)
        _split_region(rng, x, y + h // 2, w, h // 2, pw, ph, min_size, p_split, out)
    else:
        _split_region(rng, x, y, w // 2, h, pw, ph, min_size, p_split, out)
        _split_region(rng, x + w // 2, y, w // 2, h, pw, ph, min_size, p_split, out)


def random_tiling(rng, pw, ph, cell, min_size, mean_log2=4.0):
    """List of (x, y, w, h) tiles covering the plane; tile sides are powers of two in [min_size, cell]."""
    def p_split(w, h):
        return min(0.95, max(0.05, 0.5 + 0.35 * (np.log2(max(w, h)) - mean_log2)))
    out = []
    for y in range(0, ph, cell):
        for x in range(0, pw, cell):
            _split_region(rng, x, y, cell, cell, pw, ph, min_size, p_split, out)
    return np.array(out, np.int32).reshape(-1, 4)


_DIAG4 = [(0, 0), (0, 1), (1, 0), (0, 2), (1, 1), (2, 0), (0, 3), (1, 2),
          (2, 1), (3, 0), (1, 3), (2, 2), (3, 1), (2, 3), (3, 2), (3, 3)]   # (x, y), up-right diagonal 4x4


def residual_picture(seed, width, height, coded_fraction=1.0, levels="realistic", tools=("mts", "ts", "bdpcm", "lfnst",
                     "scaling", "joint", "dep_quant"), mean_log2=4.0, max_qp=63, sort=True, bd=10):
    """One picture's residual batch.

    Returns dict(tbs=structured array TB_DTYPE, levels=int16 array, params=abi.ResidualParams,
                 intra_mode=int8 per TB (raw CU mode for LFNST TBs, reference side only), seed=...).
    """
    rng = np.random.default_rng(seed)
    tools = set(tools)
    dims = plane_dims(width, height)
    recs = []
    # luma tiles, then chroma tiles (one tiling shared by Cb and Cr so joint TBs can cover both)
    luma = random_tiling(rng, dims[0][0], dims[0][1], 64, 4, mean_log2)
    chroma = random_tiling(rng, dims[1][0], dims[1][1], 32, 2, mean_log2 - 1)
    for c_set, tiles in (((0,), luma), ((1, 2), chroma)):
        keep = rng.random(len(tiles)) < coded_fraction
        for (x, y, w, h) in tiles[keep]:
            if c_set == (0,):
                recs.append((x, y, 0, w, h, 0))
            else:
                jm = rng.integers(0, 6) if "joint" in tools and w * h > 4 else 0
                if jm in (1, 2, 3):      # joint: 1 = Cb carries (mode 2: both coded), 2 = Cb carries + shift, 3 = Cr carries + shift
                    recs.append((x, y, 1 if jm < 3 else 2, w, h, jm))
                else:
                    recs.append((x, y, 1, w, h, 0))
                    if rng.random() < 0.8:
                        recs.append((x, y, 2, w, h, 0))
    n = len(recs)
    tbs = np.zeros(n, abi.TB_DTYPE)
    intra_mode = np.zeros(n, np.int8)
    lv_chunks, off = [], 0
    for i, (x, y, c, w, h, jm) in enumerate(recs):
        l2w, l2h = int(np.log2(w)), int(np.log2(h))
        flags, lf, trh, trv = 0, 0, abi.DCT2, abi.DCT2
        inter = rng.random() < 0.6
        if inter:
            flags |= abi.TB_INTER
        if "dep_quant" in tools and rng.random() < 0.5:
            flags |= abi.TB_DEP_QUANT
        if "scaling" in tools and rng.random() < 0.4:
            flags |= abi.TB_SCALING
        if jm:
            flags |= abi.TB_JOINT | (abi.TB_JOINT_SHIFT if jm >= 2 else 0)
        r = rng.random()
        mask = None
        if "ts" in tools and w <= 32 and h <= 32 and w >= 4 and h >= 4 and r < 0.08:
            flags |= abi.TB_TS
            if "bdpcm" in tools and not inter and rng.random() < 0.5:
                flags |= abi.TB_BDPCM | (abi.TB_BDPCM_VERT if rng.random() < 0.5 else 0)
            nzw, nzh = int(rng.integers(1, w + 1)), int(rng.integers(1, h + 1))
        elif "lfnst" in tools and not inter and w >= 4 and h >= 4 and r < 0.2:
            idx = int(rng.integers(1, 3))
            mode = int(rng.integers(0, 67))
            intra_mode[i] = mode
            lf = -1          # resolved below by lfnst_code()
            nz = 8 if (w, h) in ((4, 4), (8, 8)) else 16
            k = int(rng.integers(1, nz + 1))
            pos = _DIAG4[:k]
            nzw, nzh = max(p[0] for p in pos) + 1, max(p[1] for p in pos) + 1
            mask = np.zeros((nzh, nzw), bool)
            for (px, py) in pos:
                mask[py, px] = True
            lf = lfnst_code(idx, mode, w, h, c)
        else:
            if c == 0 and "mts" in tools and 4 <= w <= 32 and 4 <= h <= 32 and r < 0.5:
                if r < 0.35:     # explicit MTS
                    trh, trv = [(abi.DST7, abi.DST7), (abi.DCT8, abi.DST7), (abi.DST7, abi.DCT8), (abi.DCT8, abi.DCT8)][int(rng.integers(0, 4))]
                else:            # implicit MTS
                    trh = abi.DST7 if w <= 16 else abi.DCT2
                    trv = abi.DST7 if h <= 16 else abi.DCT2
            nzw = int(rng.integers(1, min(w, 32 if trh == abi.DCT2 else 16) + 1))
            nzh = int(rng.integers(1, min(h, 32 if trv == abi.DCT2 else 16) + 1))
        if levels == "realistic":
            lv = np.clip(np.rint(rng.laplace(0, 12, (nzh, nzw))), -256, 255).astype(np.int16)
            lv[rng.random((nzh, nzw)) < 0.3] = 0
        else:
            lv = rng.integers(-32768, 32768, (nzh, nzw)).astype(np.int16)
        if mask is not None:
            lv[~mask] = 0
        lv[nzh - 1, nzw - 1] = lv[nzh - 1, nzw - 1] or 1       # the scan box corner is significant by construction
        if mask is not None and not mask[nzh - 1, nzw - 1]:
            lv[nzh - 1, nzw - 1] = 0
        qp = int(rng.integers(12, max_qp + 1))
        t = tbs[i]
        t["x0"], t["y0"], t["c_idx"], t["log2_w"], t["log2_h"] = x, y, c, l2w, l2h
        t["nzw"], t["nzh"], t["trh"], t["trv"], t["qp"] = nzw, nzh, trh, trv, qp
        t["flags"], t["lfnst"], t["coeff_off"] = flags, lf, off
        lv_chunks.append(lv.reshape(-1))
        off += lv.size
    level_buf = np.concatenate(lv_chunks) if lv_chunks else np.zeros(0, np.int16)
    sl = rng.integers(4, 65, (28, 64)).astype(np.uint8)
    dc = rng.integers(4, 65, 14).astype(np.uint8)
    params = abi.residual_params(bd, 1 if rng.random() < 0.5 else -1, sl, dc)
    tbs, intra_mode = abi.order_tbs_by_class(tbs, params, (intra_mode,))
    return dict(tbs=tbs, levels=level_buf, params=params, intra_mode=intra_mode, seed=seed, dims=dims)


# ---- host-side derivations the parser performs before emitting a TB descriptor -------------------------
_LFNST_TR_SET = None


def wide_angle_mode(mode, w, h, c_idx=0):
    """ff_vvc_wide_angle_mode_mapping (reference vvc_intra.c:693-714) for non-ISP blocks, raw mode 0..66."""
    if w == h or mode < 2 or mode > 66:
        return mode
    ratio = abs(int(np.log2(w)) - int(np.log2(h)))
    if w > h and mode >= 2 and mode < (8 + 2 * ratio if ratio > 1 else 8):
        return mode + 65
    if h > w and mode <= 66 and mode > (60 - 2 * ratio if ratio > 1 else 60):
        return mode - 67
    return mode


def lfnst_code(lfnst_idx, raw_mode, w, h, c_idx):
    """The `lfnst` byte of VVCB200TransformBlock from the CU's raw intra mode (vvc_intra.c:34-77, vvc_itx_1d.c:711)."""
    from .tables import LFNST_TR_SET
    m = wide_angle_mode(raw_mode, w, h, c_idx)
    tr_set = 1 if m < 0 else int(LFNST_TR_SET[m])
    return lfnst_idx | (tr_set << 2) | ((1 if m > 34 else 0) << 4)


# ---- inter prediction ------------------------------------------------------------------------------------
def smooth_planes(seed, dims, bd=10):
    """Reference pictures: low-pass filtered noise so that filters/decisions see texture and flat areas."""
    rng = np.random.default_rng(seed)
    out = []
    for (w, h) in dims:
        a = rng.integers(0, 1 << bd, (h // 8 + 2, w // 8 + 2)).astype(np.float64)
        a = np.kron(a, np.ones((8, 8)))[:h, :w]
        a = a + rng.normal(0, 24, (h, w))
        out.append(np.clip(np.rint(a), 0, (1 << bd) - 1).astype(np.uint16))
    return out


def _decorate_blocks(rng, b, tools, n_refs, mv_range=64 * 16, far_mv=0.02, keep_mv=False, bd=10):
    """Fill prediction mode, vectors and tool flags of the sub-blocks `b` (geometry already set) and build the
    slice's InterParams.  keep_mv: pred_flag / ref_idx / mv are already there (one vector per coding unit)."""
    n = len(b)
    if keep_mv:
        pf = b["pred_flag"].astype(np.int64)
    else:
        pf = rng.integers(1, 4 if "bi" in tools else 3, n)
        b["pred_flag"] = pf
        b["ref_idx"] = rng.integers(0, n_refs, (n, 2))
        mv = rng.integers(-mv_range, mv_range + 1, (n, 2, 2))
        far = rng.random(n) < far_mv                      # a few vectors pointing far outside the picture
        mv[far] = rng.integers(-(1 << 17), 1 << 17, (int(far.sum()), 2, 2))
        integer = rng.random((n, 2, 2)) < 0.2             # plenty of integer / half-pel components
        mv[integer] &= ~15
        half = rng.random((n, 2, 2)) < 0.15
        mv[half] = (mv[half] & ~15) | 8
        b["mv"] = mv
    flags = np.zeros(n, np.uint16)
    if "wp" in tools:
        flags |= np.where(rng.random(n) < 0.3, abi.PB_WP, 0).astype(np.uint16)
    if "hpel" in tools:
        flags |= np.where(rng.random(n) < 0.3, abi.PB_HPEL_IF, 0).astype(np.uint16)
    b["flags"] = flags
    if "bcw" in tools:
        b["bcw_idx"] = np.where((pf == 3) & (rng.random(n) < 0.3), rng.integers(1, 5, n), 0)
    # decoder-side tools on plain bi-predicted blocks of at least 8x8 (the conditions the parser enforces,
    # vvc_ctu.c derive_dmvr_bdof_flag: no BCW, no weighted prediction)
    big = (pf == 3) & (b["w"] >= 8) & (b["h"] >= 8)
    for tool, flag, prob in (("dmvr", abi.PB_DMVR, 0.5), ("bdof", abi.PB_BDOF, 0.5)):
        if tool in tools:
            pick = big & (rng.random(n) < prob)
            if tool == "dmvr":
                pick &= b["w"].astype(np.int32) * b["h"] >= 128
            b["flags"] = np.where(pick, (b["flags"] | flag) & ~np.uint16(abi.PB_WP), b["flags"])
            b["bcw_idx"] = np.where(pick, 0, b["bcw_idx"])
    ld = (int(rng.integers(0, 8)), int(rng.integers(0, 8)))
    weight = np.zeros((2, 3, 16), np.int16)
    offset = rng.integers(-64, 64, (2, 3, 16)).astype(np.int16)
    weight[:, 0, :] = (1 << ld[0]) + rng.integers(-128, 128, (2, 16))
    weight[:, 1:, :] = (1 << ld[1]) + rng.integers(-128, 128, (2, 2, 16))
    ref_slot = np.zeros((2, 16), np.uint8)
    ref_slot[:, :n_refs] = np.stack([rng.permutation(n_refs), rng.permutation(n_refs)])
    params = abi.inter_params(bd, ld, weight, offset, ref_slot)
    params.num_ref_idx_active[0] = params.num_ref_idx_active[1] = n_refs
    return abi.order_blocks_by_width(b, params), params


def inter_picture(seed, width, height, n_refs=2, tools=("bi", "wp", "bcw", "hpel"), mv_range=64 * 16, far_mv=0.02,
                  inter_fraction=1.0, sizes=(4, 8, 16), bd=10):
    """Prediction sub-block batch (<= 16x16) tiling the picture + per-slice InterParams."""
    rng = np.random.default_rng(seed)
    tools = set(tools)
    tiles = random_tiling(rng, width, height, 16, min(sizes), mean_log2=3.6)
    if inter_fraction < 1.0:
        tiles = tiles[rng.random(len(tiles)) < inter_fraction]
    n = len(tiles)
    b = np.zeros(n, abi.PB_DTYPE)
    b["x0"], b["y0"], b["w"], b["h"] = tiles[:, 0], tiles[:, 1], tiles[:, 2], tiles[:, 3]
    b, params = _decorate_blocks(rng, b, tools, n_refs, mv_range, far_mv, bd=bd)
    dims = plane_dims(width, height)
    refs = [smooth_planes(seed * 31 + i, dims, bd) for i in range(n_refs)]
    return dict(blocks=b, params=params, refs=refs, dims=dims, seed=seed)


# ---- deblocking --------------------------------------------------------------------------------------------
def deblock_picture(seed, width, height, ctb_log2=7, ladf=None, bs_prob=(0.25, 0.4, 0.35), bd=10):
    """Per-4x4 deblocking tables in the REFERENCE's format (fc->tab.vertical_bs & co, vvcdec.h:158-163), derived
    from a random TB partition with the reference's own length rules (vvc_filter.c:374-397, 418-427) so that
    neighbouring edges never overlap, plus QP tables, per-CTB offsets and LADF parameters."""
    rng = np.random.default_rng(seed)
    s4, r4 = (width + 3) // 4, (height + 3) // 4
    bs = np.zeros((2, 3, r4, s4), np.uint8)
    len_p = np.zeros((2, r4, s4), np.uint8)
    len_q = np.zeros((2, r4, s4), np.uint8)
    tb = np.zeros((2, r4, s4), np.int32)            # luma TB width / height per unit
    sb = np.zeros((r4, s4), bool)                   # unit belongs to a sub-block (affine / sbTMVP) CU
    qp = np.zeros((3, r4, s4), np.int8)
    luma = random_tiling(rng, width, height, 64, 4, 4.0)
    for (x, y, w, h) in luma:
        ux, uy, uw, uh = x >> 2, y >> 2, w >> 2, h >> 2
        tb[0, uy:uy + uh, ux:ux + uw] = w
        tb[1, uy:uy + uh, ux:ux + uw] = h
        sb[uy:uy + uh, ux:ux + uw] = w > 8 and h > 8 and rng.random() < 0.25
        qp[0, uy:uy + uh, ux:ux + uw] = rng.integers(10, 56)
    pick = lambda n: rng.choice(3, n, p=bs_prob).astype(np.uint8)
    for (x, y, w, h) in luma:
        ux, uy, uw, uh = x >> 2, y >> 2, w >> 2, h >> 2
        for d, (pos, size, ext, uext) in enumerate(((x, w, h, uh), (y, h, w, uw))):
            sl = (slice(uy, uy + uh), ux) if d == 0 else (uy, slice(ux, ux + uw))
            if pos > 0:                              # transform edge on the left / top side of this TB
                size_p = tb[d][uy:uy + uh, ux - 1] if d == 0 else tb[d][uy - 1, ux:ux + uw]
                sb_p = sb[uy:uy + uh, ux - 1] if d == 0 else sb[uy - 1, ux:ux + uw]
                small = (size_p <= 4) | (size <= 4)
                lp = np.where(small, 1, np.where(size_p >= 32, 7, 3))
                lq = np.where(small, 1, 7 if size >= 32 else 3)
                if sb[uy, ux]:
                    lq = np.minimum(lq, 5)
                lp = np.where(sb_p, np.minimum(lp, 5), lp)
                bs[d, 0][sl] = pick(uext)
                len_p[d][sl] = lp
                len_q[d][sl] = lq
            if sb[uy, ux]:                           # sub-block edges every 8 samples inside the TB
                for i in range(8, size, 8):
                    ln = 2 if i in (8, size - 8) else 3
                    sl2 = (slice(uy, uy + uh), ux + (i >> 2)) if d == 0 else (uy + (i >> 2), slice(ux, ux + uw))
                    bs[d, 0][sl2] = rng.integers(0, 2, uext)
                    len_p[d][sl2] = ln
                    len_q[d][sl2] = ln
    # chroma: its own TB partition (chroma samples), edges only on the 8-chroma-sample grid
    tbc = np.zeros((2, r4, s4), np.uint8)
    chroma = random_tiling(rng, width // 2, height // 2, 32, 4, 3.2)
    for (x, y, w, h) in chroma:
        ux, uy, uw, uh = x >> 1, y >> 1, max(w >> 1, 1), max(h >> 1, 1)
        tbc[0, uy:uy + uh, ux:ux + uw] = w
        tbc[1, uy:uy + uh, ux:ux + uw] = h
        q = rng.integers(22, 68, 2)
        qp[1, uy:uy + uh, ux:ux + uw] = q[0]
        qp[2, uy:uy + uh, ux:ux + uw] = q[1]
        if x > 0 and x % 8 == 0:
            for c in (1, 2):
                bs[0, c, uy:uy + uh, ux] = pick(uh)
        if y > 0 and y % 8 == 0:
            for c in (1, 2):
                bs[1, c, uy, ux:ux + uw] = pick(uw)
    n_ctb = ((width + (1 << ctb_log2) - 1) >> ctb_log2) * ((height + (1 << ctb_log2) - 1) >> ctb_log2)
    ctb_offsets = rng.integers(-12, 13, (n_ctb, 6)).astype(np.int8)
    p = abi.DeblockParams()
    p.bit_depth, p.ctb_log2_size = bd, ctb_log2
    use_ladf = bool(rng.random() < 0.5) if ladf is None else ladf
    if use_ladf:
        p.ladf_enabled, p.num_ladf_intervals = 1, 4
        p.ladf_lowest_interval_qp_offset = int(rng.integers(-6, 7))
        for i in range(4):
            p.ladf_qp_offset[i] = int(rng.integers(-6, 7))
        bounds = [0] + sorted(int(v) for v in rng.integers(100, 900, 3)) + [0]
        for i in range(5):
            p.ladf_interval_lower_bound[i] = bounds[i]
    return dict(bs=bs, len_p=len_p, len_q=len_q, tb_w_c=tbc[0], tb_h_c=tbc[1], qp=qp, ctb_offsets=ctb_offsets,
                params=p, stride4=s4, rows4=r4, dims=plane_dims(width, height), seed=seed)


# ---- SAO / ALF / LMCS ----------------------------------------------------------------------------------------
def slice_map(rng, ctb_w, ctb_h, n_slices):
    """Per-CTB slice index: raster-scan slices (consecutive CTB runs), like VVC raster-scan slice mode."""
    n = ctb_w * ctb_h
    if n_slices <= 1:
        return np.zeros(n, np.int16)
    cuts = np.sort(rng.choice(np.arange(1, n), min(n_slices - 1, n - 1), replace=False))
    return np.searchsorted(cuts, np.arange(n), side="right").astype(np.int16)


def tile_grid(ctb_w, ctb_h, tile_cols, tile_rows, lfate):
    """ctb_to_col_bd / ctb_to_row_bd (VVCPPS, first CTB column / row of the tile a CTB column / row belongs to, one entry
    past the end included) for tile boundaries given as CTB indices, e.g. tile_cols=(0, 2): two tile columns."""
    def bd(n, starts):
        starts = sorted(set(int(v) for v in starts) | {0})
        out = np.zeros(n + 1, np.uint16)
        for i in range(n + 1):
            out[i] = n if i == n else max(v for v in starts if v <= i)
        return out
    col, row = bd(ctb_w, tile_cols), bd(ctb_h, tile_rows)
    return dict(col_bd=col, row_bd=row, num_tiles=len(set(col[:-1])) * len(set(row[:-1])), lfate=int(lfate))


def sao_no_filter_flags(slice_idx, ctb_w, ctb_h, lfase, tiles=None):
    """Host-side derivation of VVCB200SaoCtb.no_filter (vert/horiz/diag_edge of ff_vvc_sao_filter,
    reference vvc_filter.c:163-212) from the slice map (pps_loop_filter_across_slices_enabled_flag = lfase) and the tile
    grid (tile_grid(); its sides count when pps_loop_filter_across_tiles_enabled_flag is 0 and there is more than one tile)."""
    out = np.zeros(ctb_w * ctb_h, np.uint8)
    no_tile = tiles is not None and tiles["num_tiles"] > 1 and not tiles["lfate"]
    if lfase and not no_tile:
        return out
    s = slice_idx.reshape(ctb_h, ctb_w)
    col = tiles["col_bd"] if tiles is not None else None
    row = tiles["row_bd"] if tiles is not None else None
    for y in range(ctb_h):
        for x in range(ctb_w):
            L, T, R, B = x == 0, y == 0, x == ctb_w - 1, y == ctb_h - 1
            lt = no_tile and not L and col[x] == x
            rt = no_tile and not R and col[x] != col[x + 1]
            ut = no_tile and not T and row[y] == y
            bt = no_tile and not B and row[y] != row[y + 1]
            sl = (lambda yy, xx: not lfase and s[y, x] != s[yy, xx])
            f = 0
            if not L and (sl(y, x - 1) or lt): f |= 1
            if not R and (sl(y, x + 1) or rt): f |= 2
            if not T and (sl(y - 1, x) or ut): f |= 4
            if not B and (sl(y + 1, x) or bt): f |= 8
            if not L and not T and (sl(y - 1, x - 1) or lt or ut): f |= 16
            if not T and not R and (sl(y - 1, x + 1) or rt or ut): f |= 32
            if not R and not B and (sl(y + 1, x + 1) or rt or bt): f |= 64
            if not L and not B and (sl(y + 1, x - 1) or lt or bt): f |= 128
            out[y * ctb_w + x] = f
    return out


def alf_edge_flags(slice_idx, ctb_w, ctb_h, lfase, tiles=None):
    """Host-side derivation of VVCB200AlfCtb.edges beyond the picture borders (reference vvc_filter.c:1264-1280)."""
    out = np.zeros(ctb_w * ctb_h, np.uint8)
    no_tile = tiles is not None and not tiles["lfate"]
    if lfase and not no_tile:
        return out
    s = slice_idx.reshape(ctb_h, ctb_w)
    col = tiles["col_bd"] if tiles is not None else None
    row = tiles["row_bd"] if tiles is not None else None
    for y in range(ctb_h):
        for x in range(ctb_w):
            f = 0
            if not lfase:
                if x > 0 and s[y, x] != s[y, x - 1]: f |= 1
                if y > 0 and s[y, x] != s[y - 1, x]: f |= 2
                if x < ctb_w - 1 and s[y, x] != s[y, x + 1]: f |= 4
                if y < ctb_h - 1 and s[y, x] != s[y + 1, x]: f |= 8
            if no_tile:
                if x > 0 and col[x] != col[x - 1]: f |= 1
                if y > 0 and row[y] != row[y - 1]: f |= 2
                if x < ctb_w - 1 and col[x] != col[x + 1]: f |= 4
                if y < ctb_h - 1 and row[y] != row[y + 1]: f |= 8
            out[y * ctb_w + x] = f
    return out


def loopfilter_ctbs(seed, width, height, ctb_log2=7, n_slices=1, lfase=1, alf_on=0.85, cc_on=0.5, bd=10,
                    tile_cols=None, tile_rows=None, lfate=1):
    """Per-CTB SAO and ALF parameters + the slice's ALF APS content (random, int8-range coefficients like
    tests/checkasm/vvc_alf.c:52-66; SAO offsets < 1 << (bd - 5) like tests/checkasm/vvc_sao.c:37)."""
    rng = np.random.default_rng(seed)
    ctb_w = (width + (1 << ctb_log2) - 1) >> ctb_log2
    ctb_h = (height + (1 << ctb_log2) - 1) >> ctb_log2
    n = ctb_w * ctb_h
    slices = slice_map(rng, ctb_w, ctb_h, n_slices)
    tiles = tile_grid(ctb_w, ctb_h, tile_cols or (0,), tile_rows or (0,), lfate) if (tile_cols or tile_rows) else None
    sao = np.zeros(n, abi.SAO_DTYPE)
    sao["type_idx"] = rng.integers(0, 3, (n, 3))
    sao["band_position"] = rng.integers(0, 32, (n, 3))
    sao["eo_class"] = rng.integers(0, 4, (n, 3))
    off = rng.integers(-(1 << (bd - 5)) + 1, 1 << (bd - 5), (n, 3, 5)).astype(np.int16)
    off[:, :, 0] = 0
    sao["offset_val"] = off
    sao["no_filter"] = sao_no_filter_flags(slices, ctb_w, ctb_h, lfase, tiles)
    alf = np.zeros(n, abi.ALF_CTB_DTYPE)
    alf["ctb_flag"] = rng.random((n, 3)) < alf_on
    alf["filt_set_idx_y"] = np.where(rng.random(n) < 0.4, rng.integers(0, 16, n), 16 + rng.integers(0, 6, n))
    alf["alt_idx"] = rng.integers(0, 8, (n, 2))
    alf["cc_idc"] = np.where(rng.random((n, 2)) < cc_on, rng.integers(1, 5, (n, 2)), 0)
    alf["edges"] = alf_edge_flags(slices, ctb_w, ctb_h, lfase, tiles)
    p = np.zeros(1, abi.ALF_PARAMS_DTYPE)
    p["bit_depth"], p["ctb_log2_size"] = bd, ctb_log2
    p["luma_coeff"] = rng.integers(-128, 128, (7, 25, 12))
    p["luma_clip_idx"] = rng.integers(0, 4, (7, 25, 12))
    p["chroma_coeff"] = rng.integers(-128, 128, (8, 6))
    p["chroma_clip_idx"] = rng.integers(0, 4, (8, 6))
    p["cc_coeff"] = rng.integers(-64, 64, (2, 4, 7))
    p["cc_enabled"] = 1
    return dict(sao=sao, alf=alf, alf_params=p, slice_idx=slices, lfase=lfase, ctb_log2=ctb_log2,
                dims=plane_dims(width, height), seed=seed, bd=bd, tiles=tiles)


def lmcs_luts(seed, bd=10):
    """A monotone forward LUT and its (approximate) inverse, both 1 << bd entries of uint16."""
    rng = np.random.default_rng(seed)
    n = 1 << bd
    steps = rng.integers(1, 4, n).astype(np.float64)
    fwd = np.clip(np.rint(np.cumsum(steps) / steps.sum() * (n - 1)), 0, n - 1).astype(np.uint16)
    inv = np.clip(np.searchsorted(fwd, np.arange(n)), 0, n - 1).astype(np.uint16)
    return fwd, inv


def lmcs_model(seed, fwd, bd=10):
    """Pivots of the 16 bins in the mapped domain (taken from the forward LUT) and per-bin chroma residual scale factors
    (11 fractional bits), shaped like VVCLMCS (reference vvc_ps.h:192-202)."""
    rng = np.random.default_rng(seed)
    org_cw = (1 << bd) // 16
    pivot = np.array([int(fwd[min(i * org_cw, (1 << bd) - 1)]) for i in range(16)] + [(1 << bd) - 1], np.uint16)
    coeff = rng.integers(1 << 10, 1 << 12, 16).astype(np.uint16)
    lo = int(rng.integers(0, 3))
    hi = int(rng.integers(13, 16))
    return dict(pivot=pivot, coeff=coeff, min_bin=lo, max_bin=hi)


# ---- a whole picture --------------------------------------------------------------------------------------------
def full_picture(seed, width, height, ctb_log2=7, n_refs=2, coded_fraction=1.0, lmcs=False, n_slices=1, lfase=1,
                 residual_tools=("mts", "ts", "lfnst", "scaling", "joint", "dep_quant"), mean_log2=4.3, intra_fraction=0.0, bd=10):
    """All descriptor batches of one inter picture (BASELINE config 2/3 shape): prediction sub-blocks, residual TBs,
    deblocking tables, SAO/ALF CTB parameters.  Stage inputs are generated independently of each other (the
    arithmetic of every stage takes its tables as given); see DESIGN.md for what a real parser adds."""
    if intra_fraction > 0:
        return _full_picture_with_intra(seed, width, height, ctb_log2, n_refs, lmcs, n_slices, lfase, mean_log2,
                                        intra_fraction, bd)
    inter = inter_picture(seed, width, height, n_refs=n_refs, tools=("bi", "wp", "bcw", "hpel", "dmvr", "bdof"), bd=bd)
    res = residual_picture(seed + 1, width, height, coded_fraction=coded_fraction, tools=residual_tools,
                           mean_log2=mean_log2, bd=bd)
    # every residual TB of an inter picture belongs to an inter CU
    res["tbs"]["flags"] |= abi.TB_INTER
    res["tbs"]["flags"] &= ~np.uint16(abi.TB_BDPCM | abi.TB_BDPCM_VERT)
    res["tbs"]["lfnst"] = 0          # LFNST is intra-only; its TBs stay as plain DCT2 blocks
    res["tbs"], res["intra_mode"] = abi.order_tbs_by_class(res["tbs"], res["params"], (res["intra_mode"],))
    db = deblock_picture(seed + 2, width, height, ctb_log2=ctb_log2, bd=bd)
    lf = loopfilter_ctbs(seed + 3, width, height, ctb_log2=ctb_log2, n_slices=n_slices, lfase=lfase, bd=bd)
    out = dict(inter=inter, residual=res, deblock=db, lf=lf, dims=inter["dims"], ctb_log2=ctb_log2, seed=seed,
               width=width, height=height, lmcs=None, bd=bd)
    if lmcs:
        out["lmcs"] = lmcs_luts(seed + 4, bd)
        inter["blocks"]["flags"] |= abi.PB_LMCS
    return out


def _full_picture_with_intra(seed, width, height, ctb_log2, n_refs, lmcs, n_slices, lfase, mean_log2, intra_fraction, bd=10):
    """full_picture() over ONE coding-unit partition: every CU is inter (prediction sub-blocks, TBs added in place) or
    intra (records for the wavefront, TBs routed through the residual planes), as in a P/B picture with intra CUs."""
    luts = lmcs_luts(seed + 4, bd) if lmcs else None
    rp = recon_picture(seed, width, height, ctb_log2=ctb_log2, intra_fraction=intra_fraction, n_refs=n_refs,
                       lmcs_chroma=lmcs_model(seed + 5, luts[0], bd) if lmcs else None, bd=bd,
                       tools=("mrl", "mip", "cclm", "pdpc", "isp", "bdpcm", "ciip", "mts", "ts", "lfnst", "joint",
                              "dep_quant", "scaling"), mean_log2=mean_log2,
                       inter_tools=("bi", "wp", "bcw", "hpel", "dmvr", "bdof"),
                       inter_kinds=(CU_REGULAR,) * 7 + (CU_AFFINE,) * 2 + (CU_GPM,), lmcs=lmcs, ciip_share=0.1)
    db = deblock_picture(seed + 2, width, height, ctb_log2=ctb_log2, bd=bd)
    lf = loopfilter_ctbs(seed + 3, width, height, ctb_log2=ctb_log2, n_slices=n_slices, lfase=lfase, bd=bd)
    out = dict(inter=rp["inter"], residual=rp["residual"], intra=rp["intra"], ref=rp["ref"], deblock=db, lf=lf,
               dims=rp["dims"], ctb_log2=ctb_log2, seed=seed, width=width, height=height, lmcs=None, bd=bd)
    out["residual"]["intra_mode"] = np.zeros(len(rp["residual"]["tbs"]), np.int32)
    if lmcs:
        out["lmcs"] = luts
    return out


# ---- CU-level inter description (affine + PROF, GPM) and its expansion into engine records ------------------------
CU_DTYPE = np.dtype([
    ("x0", "<i4"), ("y0", "<i4"), ("w", "<i4"), ("h", "<i4"), ("kind", "<i4"), ("flags", "<i4"), ("pred_flag", "<i4"),
    ("bcw_idx", "<i4"), ("ref_idx", "<i4", (2,)), ("mv", "<i4", (2, 2)), ("gpm_partition_idx", "<i4"),
    ("gpm_list", "<i4", (2,)), ("cb_prof_flag", "<i4", (2,)), ("diff_mv_x", "i1", (2, 16)), ("diff_mv_y", "i1", (2, 16)),
    ("sb_mv_off", "<i4"),
])                                  # mirrors VVCRefInterCU (oracle/ref/api_common.h)
CU_REGULAR, CU_AFFINE, CU_GPM = 0, 1, 2


def _round_mv_half(v):
    """ff_vvc_round_mv(mv, 0, 1): halve, rounding half away from zero (reference vvc_mvs.c:1739-1749)."""
    return (v + 1 - (v >= 0)) >> 1


def _emit_inter_cu(rng, c, kind, n_refs, lmcs, sb_mv, blocks, prof):
    """Fill the CU-level record `c` (geometry already set) for an inter CU of `kind` and append its engine records:
    sub-blocks to `blocks`, 4x4 sub-block vectors to `sb_mv`, PROF tables to `prof`."""
    x, y, w, h = int(c["x0"]), int(c["y0"]), int(c["w"]), int(c["h"])
    c["kind"] = kind
    pf = int(rng.integers(1, 4))
    c["pred_flag"] = pf if kind != CU_GPM else 3
    c["ref_idx"] = rng.integers(0, n_refs, 2)
    c["mv"] = rng.integers(-48 * 16, 48 * 16 + 1, (2, 2))
    flags = (abi.PB_WP if rng.random() < 0.3 and kind != CU_GPM else 0) | (abi.PB_LMCS if lmcs else 0)
    if kind == CU_REGULAR and rng.random() < 0.3:
        flags |= abi.PB_HPEL_IF
    c["flags"] = flags
    if pf == 3 and kind != CU_GPM and rng.random() < 0.3:
        c["bcw_idx"] = int(rng.integers(1, 5))
    if kind == CU_REGULAR:
        if w > 16 or h > 16:        # only reached when no other kind is allowed: keep the reference's one-sub-block view legal
            c["w"], c["h"] = w, h = min(w, 16), min(h, 16)
        b = np.zeros(1, abi.PB_DTYPE)
        b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"], b["bcw_idx"] = x, y, w, h, pf, c["bcw_idx"]
        b["ref_idx"], b["mv"], b["flags"] = c["ref_idx"], c["mv"], flags
        blocks.append(b)
    elif kind == CU_GPM:
        c["gpm_partition_idx"] = int(rng.integers(0, 64))
        c["gpm_list"] = rng.integers(0, 2, 2)
        for ty in range(0, h, 16):
            for tx in range(0, w, 16):
                b = np.zeros(1, abi.PB_DTYPE)
                b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"] = x + tx, y + ty, min(16, w), min(16, h), 3
                b["ref_idx"], b["mv"], b["flags"] = c["ref_idx"], c["mv"], flags | abi.PB_GPM
                b["aux"] = abi.gpm_aux(int(c["gpm_partition_idx"]), int(np.log2(w)), int(np.log2(h)), tx, ty,
                                       int(c["gpm_list"][0]), int(c["gpm_list"][1]))
                blocks.append(b)
    else:
        nsx, nsy = w // 4, h // 4
        c["sb_mv_off"] = len(sb_mv)
        base = c["mv"].astype(np.int64)
        gx = rng.integers(-6, 7, (2, 2))
        gy = rng.integers(-6, 7, (2, 2))
        prof_idx = len(prof)
        c["cb_prof_flag"] = rng.integers(0, 2, 2)
        pt = np.zeros(1, abi.PROF_DTYPE)
        pt["diff_mv_x"] = c["diff_mv_x"] = rng.integers(-31, 32, (2, 16))
        pt["diff_mv_y"] = c["diff_mv_y"] = rng.integers(-31, 32, (2, 16))
        prof.append(pt)
        mvs = np.zeros((nsy, nsx, 2, 2), np.int64)
        for sy in range(nsy):
            for sx in range(nsx):
                mvs[sy, sx] = np.clip(base + gx * sx + gy * sy, -(1 << 17), (1 << 17) - 1)
                sb_mv.append(mvs[sy, sx].astype(np.int32))
        pflags = (abi.PB_PROF0 if c["cb_prof_flag"][0] else 0) | (abi.PB_PROF1 if c["cb_prof_flag"][1] else 0)
        for sy in range(nsy):
            for sx in range(nsx):
                b = np.zeros(1, abi.PB_DTYPE)
                b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"], b["bcw_idx"] = x + 4 * sx, y + 4 * sy, 4, 4, pf, c["bcw_idx"]
                b["ref_idx"], b["mv"] = c["ref_idx"], mvs[sy, sx]
                b["flags"] = flags | abi.PB_AFFINE | abi.PB_LUMA_ONLY | pflags
                b["aux"] = prof_idx
                blocks.append(b)
                if not (sx & 1) and not (sy & 1):       # derive_affine_mvc: average with the diagonal neighbour
                    b = np.zeros(1, abi.PB_DTYPE)
                    b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"], b["bcw_idx"] = x + 4 * sx, y + 4 * sy, 8, 8, pf, c["bcw_idx"]
                    b["ref_idx"], b["mv"] = c["ref_idx"], _round_mv_half(mvs[sy, sx] + mvs[sy + 1, sx + 1])
                    b["flags"] = (flags & ~abi.PB_LMCS) | abi.PB_CHROMA_ONLY
                    blocks.append(b)


def inter_cu_picture(seed, width, height, n_refs=2, kinds=(CU_REGULAR, CU_AFFINE, CU_GPM), lmcs=False):
    """Random coding units of all three inter kinds tiling the picture; returns the CU-level arrays the reference
    harness consumes (`cus`, `sb_mv`) and the expanded engine batch (`blocks`, `prof`), plus params / refs."""
    rng = np.random.default_rng(seed)
    tiles = random_tiling(rng, width, height, 64, 8, mean_log2=4.2)
    cus = np.zeros(len(tiles), CU_DTYPE)
    sb_mv, blocks, prof = [], [], []
    ld = (int(rng.integers(0, 8)), int(rng.integers(0, 8)))
    for i, (x, y, w, h) in enumerate(tiles):
        c = cus[i]
        c["x0"], c["y0"], c["w"], c["h"] = x, y, w, h
        ratio_ok = max(w, h) // min(w, h) < 8
        kind = int(rng.choice(kinds))
        if kind == CU_GPM and not ratio_ok:
            kind = CU_AFFINE if CU_AFFINE in kinds else CU_REGULAR
        if kind == CU_REGULAR and (w > 16 or h > 16):
            kind = CU_AFFINE if CU_AFFINE in kinds else CU_GPM
            if kind == CU_GPM and not ratio_ok:
                kind = CU_REGULAR
        _emit_inter_cu(rng, c, kind, n_refs, lmcs, sb_mv, blocks, prof)
    weight = np.zeros((2, 3, 16), np.int16)
    offset = rng.integers(-64, 64, (2, 3, 16)).astype(np.int16)
    weight[:, 0, :] = (1 << ld[0]) + rng.integers(-128, 128, (2, 16))
    weight[:, 1:, :] = (1 << ld[1]) + rng.integers(-128, 128, (2, 2, 16))
    ref_slot = np.zeros((2, 16), np.uint8)
    ref_slot[:, :n_refs] = np.stack([rng.permutation(n_refs), rng.permutation(n_refs)])
    params = abi.inter_params(10, ld, weight, offset, ref_slot)
    params.num_ref_idx_active[0] = params.num_ref_idx_active[1] = n_refs
    blocks = abi.order_blocks_by_width(np.concatenate(blocks), params)
    dims = plane_dims(width, height)
    refs = [smooth_planes(seed * 31 + i, dims) for i in range(n_refs)]
    return dict(cus=cus, sb_mv=np.array(sb_mv, np.int32).reshape(-1, 2, 2), blocks=blocks,
                prof=np.concatenate(prof) if prof else np.zeros(0, abi.PROF_DTYPE), params=params, refs=refs, dims=dims, seed=seed)


# ---- pictures with intra coding units (CTU / CU / TU structure in decoding order) ---------------------------------------
REF_CU_DTYPE = np.dtype([
    ("x0", "<i4"), ("y0", "<i4"), ("w", "<i4"), ("h", "<i4"), ("ctu", "<i4"), ("pred_mode", "<i4"),
    ("intra_pred_mode_y", "<i4"), ("intra_pred_mode_c", "<i4"), ("intra_luma_ref_idx", "<i4"),
    ("mip_flag", "<i4"), ("mip_transposed", "<i4"), ("mip_mode", "<i4"), ("mip_chroma_direct", "<i4"),
    ("isp_split_type", "<i4"), ("num_intra_subpartitions", "<i4"), ("bdpcm_flag", "<i4", (3,)), ("ciip_flag", "<i4"),
    ("coded_flag", "<i4"), ("first_tu", "<i4"), ("n_tu", "<i4"), ("mts_idx", "<i4"), ("lfnst_idx", "<i4"),
    ("inter_flags", "<i4"), ("pred_flag", "<i4"), ("ref_idx", "<i4", (2,)), ("mv", "<i4", (2, 2)), ("tree_type", "<i4"),
])                                  # mirrors VVCRefCU (oracle/ref/api_common.h)
REF_TU_DTYPE = np.dtype([
    ("x0", "<i4"), ("y0", "<i4"), ("w", "<i4"), ("h", "<i4"), ("coded_flag", "<i4", (3,)), ("joint", "<i4"),
    ("tb", "<i4", (3,)),
])                                  # mirrors VVCRefTU
MODE_INTER, MODE_INTRA = 0, 1
CCLM_LT, CCLM_L, CCLM_T = 81, 82, 83
_MTS_OF = {(0, 0): 0, (1, 1): 1, (2, 1): 2, (1, 2): 3, (2, 2): 4}      # (trh, trv) -> mts_idx (vvc_intra.c:133-134)


def _split_cu(rng, x, y, w, h, pw, ph, min_size, mean_log2, out):
    """Recursive quad / binary partition of one CTU region in decoding (z) order, clipped to the picture."""
    if x >= pw or y >= ph:
        return
    over = x + w > pw or y + h > ph
    can_h, can_v = h > min_size, w > min_size
    p = min(0.95, max(0.05, 0.5 + 0.35 * (np.log2(max(w, h)) - mean_log2)))
    if not over and (max(w, h) <= 64) and (not (can_h or can_v) or rng.random() >= p):
        out.append((x, y, w, h))
        return
    r = rng.random()
    if w == h and can_h and can_v and (r < 0.5 or over or max(w, h) > 64):
        for dy in (0, h // 2):
            for dx in (0, w // 2):
                _split_cu(rng, x + dx, y + dy, w // 2, h // 2, pw, ph, min_size, mean_log2, out)
    elif can_h and (r < 0.75 or not can_v):
        _split_cu(rng, x, y, w, h // 2, pw, ph, min_size, mean_log2, out)
        _split_cu(rng, x, y + h // 2, w, h // 2, pw, ph, min_size, mean_log2, out)
    else:
        _split_cu(rng, x, y, w // 2, h, pw, ph, min_size, mean_log2, out)
        _split_cu(rng, x + w // 2, y, w // 2, h, pw, ph, min_size, mean_log2, out)


def _avail(mask, x, y, horizontal, limit):
    """Run of reconstructed 4x4 units starting at unit (x, y) going right (horizontal) or down, in units, < limit."""
    n = 0
    if horizontal:
        while x + n < limit and mask[y, x + n]:
            n += 1
    else:
        while y + n < limit and mask[y + n, x]:
            n += 1
    return n


def recon_picture(seed, width, height, ctb_log2=7, intra_fraction=0.5, n_refs=2, tools=("mrl", "mip", "cclm", "pdpc"),
                  mean_log2=4.3, inter_tools=None, inter_kinds=None, lmcs=False, lmcs_chroma=None, ciip_share=0.3, bd=10,
                  dual_tree=False):
    """A picture with intra and inter coding units.  Returns
         ref:   CU / TU arrays + TBs for the reference harness (vvcref_recon_picture), decoding order per CTU,
         inter: prediction sub-block batch of the inter CUs (stage INTER),
         residual: TB batch (TBs of intra CUs flagged TB_TO_RESIDUAL),
         intra: VVCB200IntraBlock records + per-CTU ranges (stage RECON wavefront).
    tools: intra side "mrl", "mip", "cclm", "isp", "bdpcm", "ciip" (inter units blended with a planar intra prediction);
    residual side "mts", "ts", "lfnst", "joint", "dep_quant",
    "scaling" ("pdpc" is always on, it is not signalled).  inter_tools: block-level tool mix of inter_picture() on the
    regular inter CUs; inter_kinds: CU kinds to draw from (CU_REGULAR, CU_AFFINE, CU_GPM) — affine / GPM units come with
    their CU-level description (`inter.cus`, `inter.sb_mv`) for the reference harness and PROF tables (`inter.prof`).
    dual_tree: an I picture with separate luma and chroma partitions (every unit intra; per 64x64 region the luma
    units come first, then the chroma units, vvc_ctu.c dual_tree_implicit_qt_split); intra_fraction is ignored.
    lmcs_chroma: an lmcs_model() -> luma-dependent chroma residual scaling: every chroma TB goes through the residual
    planes and every chroma block with a residual gets a record (intra, or add-only for inter units).
    Neighbour availability follows the reference's rules (vvc_intra.c:593-648) on a 4x4-unit reconstruction mask."""
    rng = np.random.default_rng(seed)
    tools = set(tools)
    ctb = 1 << ctb_log2
    cw, ch = (width + ctb - 1) >> ctb_log2, (height + ctb - 1) >> ctb_log2
    u4w, u4h = (width + 3) >> 2, (height + 3) >> 2
    done_l = np.zeros((u4h, u4w), bool)             # reconstructed luma / chroma, per 4x4 luma unit (one and the same
    done_c = done_l if not dual_tree else np.zeros((u4h, u4w), bool)   # map in single-tree pictures)
    luma_mode = np.zeros((u4h, u4w), np.int16)      # IntraPredModeY per unit (0 for MIP), for the chroma DM mode
    cus, tus, tbs, lv_chunks = [], [], [], []
    pblocks, iblocks, ctu_first = [], [], [0]
    cu_blocks, cu_recs, sb_mv, prof = [], [], [], []     # affine / GPM coding units
    is_intra = np.zeros((u4h, u4w), bool)
    ciip_flags = (abi.PB_WP if rng.random() < 0.5 else 0) | (abi.PB_LMCS if lmcs else 0) if "ciip" in tools else 0
    lv_off = 0
    pic_flags = (abi.TB_DEP_QUANT if "dep_quant" in tools and rng.random() < 0.5 else 0) | \
                (abi.TB_SCALING if "scaling" in tools and rng.random() < 0.6 else 0)

    def new_tb(x, y, c, w, h, inter, trh=0, trv=0, qp=30, to_res=False, ts=False, bdpcm=0, lfnst=0, joint=0):
        nonlocal lv_off
        flags = pic_flags | (abi.TB_INTER if inter else 0) | (abi.TB_TO_RESIDUAL if to_res else 0)
        mask = None
        if ts:
            flags |= abi.TB_TS
            if bdpcm:
                flags |= abi.TB_BDPCM | (abi.TB_BDPCM_VERT if bdpcm == 2 else 0)
            nzw, nzh = int(rng.integers(1, w + 1)), int(rng.integers(1, h + 1))
        elif lfnst:
            nz = 8 if (w, h) in ((4, 4), (8, 8)) else 16
            pos = _DIAG4[:int(rng.integers(1, nz + 1))]
            nzw, nzh = max(q[0] for q in pos) + 1, max(q[1] for q in pos) + 1
            mask = np.zeros((nzh, nzw), bool)
            for (px, py) in pos:
                mask[py, px] = True
        else:
            nzw = int(rng.integers(1, min(w, 32 if trh == 0 else 16) + 1))
            nzh = int(rng.integers(1, min(h, 32 if trv == 0 else 16) + 1))
        lv = np.clip(np.rint(rng.laplace(0, 10, (nzh, nzw))), -200, 200).astype(np.int16)
        if mask is not None:
            lv[~mask] = 0
            if not lv.any():
                lv[0, 0] = 3
        else:
            lv[nzh - 1, nzw - 1] = lv[nzh - 1, nzw - 1] or 1
        if joint:
            flags |= abi.TB_JOINT | (abi.TB_JOINT_SHIFT if joint >= 2 else 0)
        t = np.zeros(1, abi.TB_DTYPE)
        t["x0"], t["y0"], t["c_idx"], t["log2_w"], t["log2_h"] = x, y, c, int(np.log2(w)), int(np.log2(h))
        t["nzw"], t["nzh"], t["trh"], t["trv"], t["qp"] = nzw, nzh, trh, trv, qp
        t["flags"], t["lfnst"], t["coeff_off"] = flags, lfnst, lv_off
        lv_chunks.append(lv.reshape(-1))
        lv_off += lv.size
        tbs.append(t)
        return len(tbs) - 1

    def chroma_tbs(tu, x, y, w, h, inter, qp_c, coded, to_res, bdpcm_c):
        """Chroma TBs of a TU (luma geometry in): separate Cb / Cr, or one joint TB standing for both."""
        to_res = to_res or lmcs_chroma is not None
        jm = 0
        if "joint" in tools and (coded[1] or coded[2]) and not bdpcm_c and rng.random() < 0.3:
            jm = int(rng.integers(1, 4))     # 1: Cb carries, both coded; 2: Cb carries, Cr halved; 3: Cr carries, Cb halved
        if jm:
            carrier = 2 if jm == 3 else 1
            tu["joint"] = 1
            tu["coded_flag"][0, 1], tu["coded_flag"][0, 2] = int(jm != 3), int(jm != 2)
            ts = "ts" in tools and rng.random() < 0.1
            tu["tb"][0, carrier] = new_tb(x >> 1, y >> 1, carrier, w >> 1, h >> 1, inter, 0, 0, qp_c, to_res, ts=ts, joint=jm)
            return 3
        got = 0
        for c in (1, 2):
            if coded[c]:
                ts = bool(bdpcm_c) or ("ts" in tools and rng.random() < 0.1)
                tu["tb"][0, c] = new_tb(x >> 1, y >> 1, c, w >> 1, h >> 1, inter, 0, 0, qp_c, to_res, ts=ts, bdpcm=bdpcm_c)
                got |= 1 << (c - 1)
        return got

    vsize = min(ctb, 64)

    def set_vpdu(ib, x, y, ctb_left, ctb_up):
        vx, vy = x & ~(vsize - 1), y & ~(vsize - 1)
        ib["vpdu_x"], ib["vpdu_y"] = vx // vsize, vy // vsize
        ib["vpdu_avail"] = int((vx & (ctb - 1)) > 0 or ctb_left) | (int((vy & (ctb - 1)) > 0 or ctb_up) << 1)
        ib["flags"] |= abi.IB_LMCS_SCALE

    def add_only_records(x, y, w, h, res_c, ctb_left, ctb_up):
        """Chroma blocks of an inter unit whose residual waits for the luma-dependent scale."""
        for c in (1, 2):
            if res_c & (1 << (c - 1)):
                ib = np.zeros(1, abi.IB_DTYPE)
                ib["x0"], ib["y0"], ib["w"], ib["h"], ib["c_idx"] = x >> 1, y >> 1, w >> 1, h >> 1, c
                ib["flags"] = abi.IB_ADD_ONLY | abi.IB_RESIDUAL
                ib["cb_w"], ib["cb_h"] = w, h
                set_vpdu(ib, x, y, ctb_left, ctb_up)
                iblocks.append(ib)

    for ry in range(ch):
        for rx in range(cw):
            rs = ry * cw + rx
            regions = []
            if not dual_tree:
                tmp = []
                _split_cu(rng, rx * ctb, ry * ctb, ctb, ctb, width, height, 8, mean_log2, tmp)
                regions = [r + (0,) for r in tmp]
            else:                                   # tree 1 = luma units, 2 = chroma units, per region of at most 64x64
                rsz = min(ctb, 64)
                for qy in range(0, ctb, rsz):
                    for qx in range(0, ctb, rsz):
                        for tree in (1, 2):
                            tmp = []
                            _split_cu(rng, rx * ctb + qx, ry * ctb + qy, rsz, rsz, width, height, 8, mean_log2, tmp)
                            regions += [r + (tree,) for r in tmp]
            ctb_left, ctb_up = rx > 0, ry > 0
            for (x, y, w, h, tree) in regions:
                done = done_c if tree == 2 else done_l
                cu = np.zeros(1, REF_CU_DTYPE)
                cu["tree_type"] = tree
                intra = dual_tree or rng.random() < intra_fraction
                cu["x0"], cu["y0"], cu["w"], cu["h"], cu["ctu"] = x, y, w, h, rs
                cu["pred_mode"] = MODE_INTRA if intra else MODE_INTER
                cu["first_tu"], cu["n_tu"] = len(tus), 1
                cu["coded_flag"] = 1                    # the reference only skips reconstruct() for uncoded CUs
                coded = rng.random(3) < 0.8
                if tree == 1:
                    coded[1:] = False               # a luma unit of a dual tree has no chroma
                elif tree == 2:
                    coded[0] = False
                qp_y, qp_c = int(rng.integers(14, 60)), int(rng.integers(14, 60))
                ux, uy, uw, uh = x >> 2, y >> 2, w >> 2, h >> 2
                if not intra:
                    ciip = "ciip" in tools and rng.random() < ciip_share
                    tu = np.zeros(1, REF_TU_DTYPE)
                    tu["x0"], tu["y0"], tu["w"], tu["h"] = x, y, w, h
                    tu["coded_flag"] = coded
                    tu["tb"] = -1
                    trh = trv = 0
                    ts = "ts" in tools and w <= 32 and h <= 32 and rng.random() < 0.08
                    if not ts and "mts" in tools and w <= 32 and h <= 32 and rng.random() < 0.3:
                        trh, trv = [(1, 1), (2, 1), (1, 2), (2, 2)][int(rng.integers(0, 4))]
                    cu["mts_idx"] = _MTS_OF[(trh, trv)]
                    if coded[0]:
                        tu["tb"][0, 0] = new_tb(x, y, 0, w, h, True, trh, trv, qp_y, ciip, ts=ts)
                    res_c = chroma_tbs(tu, x, y, w, h, True, qp_c, coded, ciip, 0)
                    tus.append(tu)
                    kind = int(rng.choice(inter_kinds)) if inter_kinds else CU_REGULAR
                    if kind == CU_GPM and max(w, h) // min(w, h) >= 8:
                        kind = CU_REGULAR
                    if ciip:
                        # merge motion, default weighting unless the slice has explicit weights; the inter prediction is
                        # made in stage INTER (records flagged PB_CIIP), the planar prediction and the blend in stage RECON
                        pf = int(rng.integers(1, 4))
                        mv = rng.integers(-40 * 16, 40 * 16 + 1, (2, 2))
                        ridx = rng.integers(0, n_refs, 2)
                        fl = ciip_flags | (abi.PB_HPEL_IF if rng.random() < 0.2 else 0)
                        cu["ciip_flag"], cu["inter_flags"], cu["pred_flag"], cu["ref_idx"], cu["mv"] = 1, fl, pf, ridx, mv
                        for ty in range(0, h, 16):
                            for tx in range(0, w, 16):
                                b = np.zeros(1, abi.PB_DTYPE)
                                b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"] = x + tx, y + ty, min(16, w), min(16, h), pf
                                b["ref_idx"], b["mv"], b["flags"] = ridx, mv, fl | abi.PB_CIIP
                                cu_blocks.append(b)
                        x0b, y0b = x & (ctb - 1), y & (ctb - 1)
                        cand_up, cand_left = ctb_up or y0b > 0, ctb_left or x0b > 0
                        up_left = (cand_left and cand_up) if (x0b or y0b) else (ctb_left and ctb_up)
                        end_x4 = min(width, (rx + 1) * ctb) >> 2
                        end_y4 = min(height, (ry + 1) * ctb) >> 2
                        top4 = ((u4w - ux) if ctb_up else 0) if y0b == 0 else _avail(done, ux, uy - 1, True, end_x4)
                        left4 = ((end_y4 - uy) if ctb_left else 0) if x0b == 0 else _avail(done, ux - 1, uy, False, end_y4)
                        wt = 1 + int(cand_up and is_intra[(y - 1) >> 2, (x - 1 + w) >> 2]) + \
                            int(cand_left and is_intra[(y - 1 + h) >> 2, (x - 1) >> 2])      # vvc_inter.c:523-543
                        for c in (0, 1, 2):
                            hs = 1 if c else 0
                            ib = np.zeros(1, abi.IB_DTYPE)
                            ib["x0"], ib["y0"], ib["w"], ib["h"], ib["c_idx"] = x >> hs, y >> hs, w >> hs, h >> hs, c
                            has_res = bool(coded[0]) if c == 0 else bool(res_c & (1 << (c - 1)))
                            ib["flags"] = abi.IB_CIIP | (abi.IB_RESIDUAL if has_res else 0)
                            ib["ciip_weight"], ib["up_left"] = wt, up_left
                            ib["avail_top"], ib["avail_left"] = (top4 * 4) >> hs, (left4 * 4) >> hs
                            ib["cb_w"], ib["cb_h"] = w, h
                            if c and has_res and lmcs_chroma is not None:
                                set_vpdu(ib, x, y, ctb_left, ctb_up)
                            iblocks.append(ib)
                    elif kind != CU_REGULAR:
                        c = np.zeros(1, CU_DTYPE)
                        c["x0"], c["y0"], c["w"], c["h"] = x, y, w, h
                        _emit_inter_cu(rng, c[0], kind, n_refs, lmcs, sb_mv, cu_blocks, prof)
                        cu_recs.append(c)
                    else:
                        pf = int(rng.integers(1, 4))
                        mv = rng.integers(-40 * 16, 40 * 16 + 1, (2, 2))
                        ridx = rng.integers(0, n_refs, 2)
                        for ty in range(0, h, 16):
                            for tx in range(0, w, 16):
                                b = np.zeros(1, abi.PB_DTYPE)
                                b["x0"], b["y0"], b["w"], b["h"], b["pred_flag"] = x + tx, y + ty, min(16, w), min(16, h), pf
                                b["ref_idx"], b["mv"] = ridx, mv
                                pblocks.append(b)
                    if lmcs_chroma is not None and not ciip:
                        add_only_records(x, y, w, h, res_c, ctb_left, ctb_up)
                    done[uy:uy + uh, ux:ux + uw] = True
                    cus.append(cu)
                    continue
                # ---- intra CU: neighbour availability of the whole coding block
                x0b, y0b = x & (ctb - 1), y & (ctb - 1)
                cand_up, cand_left = ctb_up or y0b > 0, ctb_left or x0b > 0
                up_left = (cand_left and cand_up) if (x0b or y0b) else (ctb_left and ctb_up)
                end_x4 = min(width, (rx + 1) * ctb) >> 2
                end_y4 = min(height, (ry + 1) * ctb) >> 2
                top4 = ((u4w - ux) if ctb_up else 0) if y0b == 0 else _avail(done, ux, uy - 1, True, end_x4)
                left4 = ((end_y4 - uy) if ctb_left else 0) if x0b == 0 else _avail(done, ux - 1, uy, False, end_y4)
                # ---- luma tool: MIP | MRL | ISP | BDPCM | plain
                mode_y = int(rng.choice([0, 1, 18, 50] + list(range(2, 67))))
                ref_idx, isp, bdpcm_y, lfnst_idx = 0, 0, 0, 0
                mip = tree != 2 and "mip" in tools and rng.random() < 0.15
                r = rng.random()
                if tree == 2:       # chroma unit: its DM mode is the luma mode at the centre of the co-located area
                    mode_y = int(luma_mode[(y + h // 2) >> 2, (x + w // 2) >> 2])
                elif mip:
                    cu["mip_flag"], cu["mip_transposed"] = 1, int(rng.integers(0, 2))
                    size_id = 0 if (w == 4 and h == 4) else 1 if (w == 4 or h == 4 or (w == 8 and h == 8)) else 2
                    cu["mip_mode"] = int(rng.integers(0, (16, 8, 6)[size_id]))
                    mode_y = 0
                elif "mrl" in tools and y0b > 0 and mode_y != 0 and r < 0.2:
                    ref_idx = int(rng.choice([1, 2]))      # what ff_vvc_intra_luma_ref_idx() can return (vvc_cabac.c:1282-1290)
                elif "isp" in tools and w * h > 16 and r < 0.4:
                    isp = int(rng.integers(1, 3))            # 1 horizontal, 2 vertical split (vvc_ctu.h:113-117)
                elif "bdpcm" in tools and w <= 32 and h <= 32 and r < 0.55:
                    bdpcm_y = int(rng.integers(1, 3))        # 1 horizontal, 2 vertical
                    mode_y = 18 if bdpcm_y == 1 else 50
                cu["intra_pred_mode_y"], cu["intra_luma_ref_idx"] = mode_y, ref_idx
                cu["bdpcm_flag"][0, 0] = int(bdpcm_y > 0)
                # ---- chroma mode: planar / DC / hor / ver / DM / CCLM, or chroma BDPCM
                choices = [0, 1, 18, 50, mode_y] + ([CCLM_LT, CCLM_L, CCLM_T] if "cclm" in tools else [])
                mode_c = int(rng.choice(choices))
                bdpcm_c = 0
                if tree != 1 and "bdpcm" in tools and w <= 64 and h <= 64 and rng.random() < 0.12:
                    bdpcm_c = int(rng.integers(1, 3))
                    mode_c = 18 if bdpcm_c == 1 else 50
                cu["intra_pred_mode_c"] = mode_c
                cu["bdpcm_flag"][0, 1] = cu["bdpcm_flag"][0, 2] = int(bdpcm_c > 0)
                # ---- luma transform: TS | LFNST | explicit MTS | DCT2 (ISP: implicit MTS per sub-partition)
                trh = trv = 0
                ts_y = bool(bdpcm_y) or ("ts" in tools and not isp and w <= 32 and h <= 32 and rng.random() < 0.08)
                if not ts_y and not isp and tree != 2:
                    if "lfnst" in tools and (not mip or min(w, h) >= 16) and rng.random() < 0.3:
                        lfnst_idx = int(rng.integers(1, 3))
                    elif "mts" in tools and w <= 32 and h <= 32 and rng.random() < 0.3:
                        trh, trv = [(1, 1), (2, 1), (1, 2), (2, 2)][int(rng.integers(0, 4))]
                cu["mts_idx"] = _MTS_OF[(trh, trv)]
                cu["lfnst_idx"] = lfnst_idx
                # ---- transform units: one, or the ISP sub-partitions (chroma rides on the last one)
                if isp:
                    n_parts = 2 if (w, h) in ((4, 8), (8, 4)) else 4
                    pw, ph = (w, h // n_parts) if isp == 1 else (w // n_parts, h)
                    cu["isp_split_type"], cu["num_intra_subpartitions"] = isp, n_parts
                    cu["n_tu"] = n_parts
                    trh = 1 if 4 <= pw <= 16 else 0          # implicit MTS (vvc_intra.c:143-160)
                    trv = 1 if 4 <= ph <= 16 else 0
                else:
                    n_parts, pw, ph = 1, w, h
                res_c = 0
                isp_coded = 0
                for k in range(n_parts):
                    tu = np.zeros(1, REF_TU_DTYPE)
                    tx, ty = (x, y + k * ph) if isp != 2 else (x + k * pw, y)
                    tu["x0"], tu["y0"], tu["w"], tu["h"] = tx, ty, pw, ph
                    tu["tb"] = -1
                    c0 = bool(coded[0]) if n_parts == 1 else bool(rng.random() < 0.7)
                    tu["coded_flag"][0, 0] = int(c0)
                    isp_coded |= int(c0) << k
                    if c0:
                        code = lfnst_code(lfnst_idx, mode_y, w, h, 0) if lfnst_idx else 0
                        tu["tb"][0, 0] = new_tb(tx, ty, 0, pw, ph, False, trh, trv, qp_y, True, ts=ts_y, bdpcm=bdpcm_y, lfnst=code)
                    if k == n_parts - 1 and tree != 1:
                        tu["coded_flag"][0, 1], tu["coded_flag"][0, 2] = int(coded[1]), int(coded[2])
                        # chroma TBs of an ISP unit cover the whole coding block (vvc_ctu.c hls_transform_unit)
                        res_c = chroma_tbs(tu, x, y, w, h, False, qp_c, coded, True, bdpcm_c)
                    tus.append(tu)
                # ---- records for the engine
                for c in ((0, 1, 2), (0,), (1, 2))[tree]:
                    hs = 1 if c else 0
                    bw, bh = w >> hs, h >> hs
                    raw = mode_y if c == 0 else mode_c
                    if c == 2 and raw >= CCLM_LT:
                        continue                      # the CCLM record of c_idx 1 covers Cr too
                    ib = np.zeros(1, abi.IB_DTYPE)
                    ib["x0"], ib["y0"], ib["w"], ib["h"], ib["c_idx"] = x >> hs, y >> hs, bw, bh, c
                    ib["mode"] = raw if raw >= CCLM_LT else wide_angle_mode(raw, bw, bh, c)
                    ib["ref_idx"] = ref_idx if c == 0 else 0
                    fl = 0
                    if c == 0 and mip:
                        fl |= abi.IB_MIP | (abi.IB_MIP_TRANSPOSED if cu["mip_transposed"][0] else 0)
                        ib["mip_mode"] = cu["mip_mode"][0]
                    if c == 0 and isp:                # one record per coding block; w x h = the prediction unit
                        fl |= abi.IB_ISP
                        ib["w"], ib["h"] = max(pw, 4), ph
                        ib["isp_coded"] = isp_coded
                    if (c == 0 and bdpcm_y) or (c and bdpcm_c):
                        fl |= abi.IB_BDPCM
                    if raw >= CCLM_LT:
                        fl |= abi.IB_RESIDUAL if res_c else 0
                        ib["res_planes"] = res_c
                    elif c == 0:
                        fl |= abi.IB_RESIDUAL if (isp or coded[0]) else 0
                    elif res_c & (1 << (c - 1)):
                        fl |= abi.IB_RESIDUAL
                    ib["flags"] = fl
                    ib["up_left"] = up_left
                    ib["avail_top"], ib["avail_left"] = (top4 * 4) >> hs, (left4 * 4) >> hs
                    ib["cb_w"], ib["cb_h"] = w, h
                    if c and (fl & abi.IB_RESIDUAL) and lmcs_chroma is not None:
                        set_vpdu(ib, x, y, ctb_left, ctb_up)
                    iblocks.append(ib)
                done[uy:uy + uh, ux:ux + uw] = True
                is_intra[uy:uy + uh, ux:ux + uw] = True
                if tree != 2:
                    luma_mode[uy:uy + uh, ux:ux + uw] = mode_y
                cus.append(cu)
            ctu_first.append(len(iblocks))
    tbs = np.concatenate(tbs) if tbs else np.zeros(0, abi.TB_DTYPE)
    levels = np.concatenate(lv_chunks) if lv_chunks else np.zeros(0, np.int16)
    if pic_flags & abi.TB_SCALING:
        rparams = abi.residual_params(bd, 1 if rng.random() < 0.5 else -1, rng.integers(4, 65, (28, 64)).astype(np.uint8),
                                      rng.integers(4, 65, 14).astype(np.uint8))
    else:
        rparams = abi.residual_params(bd, 1 if rng.random() < 0.5 else -1)
    order, order_cls = abi.tb_order(tbs)
    for c in range(4):
        rparams.n_class[c] = int(np.sum(order_cls == c))
    ld = (0, 0)
    ref_slot = np.zeros((2, 16), np.uint8)
    ref_slot[:, :n_refs] = np.arange(n_refs)
    iparams = abi.inter_params(bd, ld, None, None, ref_slot)
    iparams.num_ref_idx_active[0] = iparams.num_ref_idx_active[1] = n_refs
    if pblocks and inter_tools is not None:     # the tool mix of inter_picture() on this picture's inter CUs
        pb, iparams = _decorate_blocks(np.random.default_rng(seed + 77), np.concatenate(pblocks), set(inter_tools), n_refs,
                                       keep_mv=True, bd=bd)
    else:
        pb = abi.order_blocks_by_width(np.concatenate(pblocks), iparams) if pblocks else np.zeros(0, abi.PB_DTYPE)
    if lmcs:
        pb["flags"] |= abi.PB_LMCS
    n_regular = len(pb)
    if cu_blocks:                               # one batch for the engine: regular sub-blocks + affine / GPM records
        pb = abi.order_blocks_by_width(np.concatenate([pb] + cu_blocks), iparams)
    dims = plane_dims(width, height)
    refs = [smooth_planes(seed * 31 + i, dims, bd) for i in range(n_refs)]
    ip = abi.IntraParams()
    ip.bit_depth, ip.ctb_log2_size, ip.chroma_vertical_collocated = bd, ctb_log2, int(rng.integers(0, 2))
    if lmcs_chroma is not None:
        ip.lmcs_chroma_scale = 1
        ip.lmcs_pivot[:] = [int(v) for v in lmcs_chroma["pivot"]]
        ip.lmcs_chroma_scale_coeff[:] = [int(v) for v in lmcs_chroma["coeff"]]
        ip.lmcs_min_bin_idx, ip.lmcs_max_bin_idx = lmcs_chroma["min_bin"], lmcs_chroma["max_bin"]
    return dict(
        ref=dict(cus=np.concatenate(cus), tus=np.concatenate(tus), tbs=tbs, levels=levels, params=rparams),
        inter=dict(blocks=pb, params=iparams, refs=refs, n_regular=n_regular,
                   cus=np.concatenate(cu_recs) if cu_recs else np.zeros(0, CU_DTYPE),
                   sb_mv=np.array(sb_mv, np.int32).reshape(-1, 2, 2),
                   prof=np.concatenate(prof) if prof else np.zeros(0, abi.PROF_DTYPE)),
        residual=dict(tbs=np.ascontiguousarray(tbs[order]), levels=levels, params=rparams),
        intra=dict(blocks=np.concatenate(iblocks) if iblocks else np.zeros(0, abi.IB_DTYPE),
                   ctu_first=np.array(ctu_first, np.uint32), params=ip),
        dims=dims, ctb_log2=ctb_log2, width=width, height=height, seed=seed, bd=bd)


# ---- the per-4x4 state boundary strengths are derived from ------------------------------------------------------------
def deblock_state(pic):
    """Per-4x4 tables of a full_picture(intra_fraction > 0) / recon_picture(): the reference's own set (fc->tab.mvf, tb_pos_*,
    tb_width / tb_height, cb_pos_*, cb_width / cb_height, tu_coded_flag, tu_joint_cbcr_residual_flag, pcmf, msf, iaf — written
    the way set_tb_pos / set_tb_tab / set_cb_pos do, vvc_ctu.c:40-74, 1145-1166) from the picture's CU / TU lists and
    prediction records, and the engine's compact form of the same (VVCB200MotionUnit, VVCB200PartitionUnit).
    Motion stored for a GPM unit is its first partition's (what the BS rules see is whatever the table holds)."""
    w, h = pic["width"], pic["height"]
    s4, r4 = (w + 3) // 4, (h + 3) // 4
    cus, tus = pic["ref"]["cus"], pic["ref"]["tus"]
    it = pic["inter"]
    mvf = np.zeros((r4, s4), abi.REF_MVF_DTYPE)                  # pred_flag 0 = PF_INTRA
    blocks = it["blocks"]
    for b in blocks[(blocks["flags"] & abi.PB_CHROMA_ONLY) == 0]:
        ux, uy, uw, uh = b["x0"] >> 2, b["y0"] >> 2, max(b["w"] >> 2, 1), max(b["h"] >> 2, 1)
        m = mvf[uy:uy + uh, ux:ux + uw]
        if b["flags"] & abi.PB_GPM:
            lst = (int(b["aux"]) >> 18) & 1
            m["pred_flag"] = 1 + lst
            m["mv"][..., lst, :] = b["mv"][0]
            m["ref_idx"][..., lst] = b["ref_idx"][0]
        else:
            m["pred_flag"], m["mv"], m["ref_idx"] = b["pred_flag"], b["mv"], b["ref_idx"]
        m["ciip_flag"] = 1 if b["flags"] & abi.PB_CIIP else 0
    tb_x0 = np.zeros((2, r4, s4), np.int32)
    tb_y0 = np.zeros((2, r4, s4), np.int32)
    tb_w = np.zeros((2, r4, s4), np.uint8)
    tb_h = np.zeros((2, r4, s4), np.uint8)
    cb_x, cb_y = np.zeros((r4, s4), np.int32), np.zeros((r4, s4), np.int32)
    cb_w, cb_h = np.zeros((r4, s4), np.uint8), np.zeros((r4, s4), np.uint8)
    cbf = np.zeros((3, r4, s4), np.uint8)
    joint = np.zeros((r4, s4), np.uint8)
    pcmf = np.zeros((2, r4, s4), np.uint8)
    msf, iaf = np.zeros((r4, s4), np.uint8), np.zeros((r4, s4), np.uint8)
    for c in it.get("cus", ()):
        if c["kind"] == CU_AFFINE:
            iaf[c["y0"] >> 2:(c["y0"] + c["h"]) >> 2, c["x0"] >> 2:(c["x0"] + c["w"]) >> 2] = 1

    def put_tb(tree, x, y, tw, th):                                # set_tb_pos: geometry in luma samples, tw / th in the tree's samples
        sh = 1 if tree else 0
        ux, uy = x >> 2, y >> 2
        uw, uh = max(1, tw >> (2 - sh)), max(1, th >> (2 - sh))
        tb_x0[tree, uy:uy + uh, ux:ux + uw], tb_y0[tree, uy:uy + uh, ux:ux + uw] = x, y
        tb_w[tree, uy:uy + uh, ux:ux + uw], tb_h[tree, uy:uy + uh, ux:ux + uw] = tw, th

    for cu in cus:
        x, y, cw_, ch_ = int(cu["x0"]), int(cu["y0"]), int(cu["w"]), int(cu["h"])
        tree = int(cu["tree_type"])
        ux, uy, uw, uh = x >> 2, y >> 2, cw_ >> 2, ch_ >> 2
        if tree != 2:
            cb_x[uy:uy + uh, ux:ux + uw], cb_y[uy:uy + uh, ux:ux + uw] = x, y
            cb_w[uy:uy + uh, ux:ux + uw], cb_h[uy:uy + uh, ux:ux + uw] = cw_, ch_
        isp = int(cu["isp_split_type"]) != 0
        for k in range(int(cu["n_tu"])):
            tu = tus[int(cu["first_tu"]) + k]
            tx, ty, tw, th = int(tu["x0"]), int(tu["y0"]), int(tu["w"]), int(tu["h"])
            if tree != 2:
                put_tb(0, tx, ty, tw, th)
                cbf[0, ty >> 2:(ty + max(th, 4)) >> 2, tx >> 2:(tx + max(tw, 4)) >> 2] = tu["coded_flag"][0]
            if tree != 1 and (not isp or k == int(cu["n_tu"]) - 1):
                gx, gy, gw, gh = (x, y, cw_, ch_) if isp else (tx, ty, tw, th)   # chroma TBs of an ISP unit cover the coding block
                put_tb(1, gx, gy, gw >> 1, gh >> 1)
                sl = (slice(gy >> 2, (gy + gh) >> 2), slice(gx >> 2, (gx + gw) >> 2))
                cbf[1][sl], cbf[2][sl], joint[sl] = tu["coded_flag"][1], tu["coded_flag"][2], tu["joint"]
    ref_poc = np.zeros((2, 16), np.int32)
    for l in range(2):
        for i in range(16):
            ref_poc[l, i] = 100 + int(it["params"].ref_slot[l][i])
    # ---- the engine's compact form
    uxs, uys = np.meshgrid(np.arange(s4), np.arange(r4))
    log2 = lambda a: np.where(a > 0, np.log2(np.maximum(a, 1)).astype(np.int64), 0)
    motion = np.zeros((r4, s4), abi.MOTION_DTYPE)
    motion["mv"], motion["pred_flag"], motion["ciip_flag"] = mvf["mv"], mvf["pred_flag"], mvf["ciip_flag"]
    for l in range(2):
        motion["ref_id"][..., l] = ref_poc[l][np.clip(mvf["ref_idx"][..., l], 0, 15)] - 100
    part = np.zeros((r4, s4), abi.PARTITION_DTYPE)
    part["tb_off"] = ((uxs - (tb_x0[0] >> 2)) & 15) | (((uys - (tb_y0[0] >> 2)) & 15) << 4)
    part["tb_size"] = log2(tb_w[0]) | (log2(tb_h[0]) << 3) | ((cbf[0] != 0) << 6) | ((pcmf[0] != 0) << 7)
    part["cu_off_x"] = ((uxs - (cb_x >> 2)) & 31) | (((msf | iaf) != 0) << 5) | ((cb_w > 8) << 6) | ((cb_h > 8) << 7)
    part["cu_off_y"] = (uys - (cb_y >> 2)) & 31
    part["ctb_off"] = ((uxs - (tb_x0[1] >> 2)) & 15) | (((uys - (tb_y0[1] >> 2)) & 15) << 4)
    part["ctb_size"] = log2(tb_w[1]) | (log2(tb_h[1]) << 3) | ((cbf[1] != 0) << 6) | ((cbf[2] != 0) << 7)
    part["cflags"] = (joint != 0) | ((pcmf[1] != 0) << 1)
    return dict(mvf=mvf, tb_pos_x0=tb_x0, tb_pos_y0=tb_y0, tb_width=tb_w, tb_height=tb_h, cb_pos_x=cb_x, cb_pos_y=cb_y,
                cb_width=cb_w, cb_height=cb_h, tu_coded_flag=cbf, tu_joint=joint, pcmf=pcmf, msf=msf, iaf=iaf, ref_poc=ref_poc,
                motion=motion, part=part, stride4=s4, rows4=r4)
