/*
 * oracle/ref/glue.c — TEST INFRASTRUCTURE, not product code.
 *
 * The reference's hot-path sources (vvcdsp.c, vvc_itx_1d.c, vvc_data.c, vvc_intra.c, vvc_filter.c,
 * vvc_inter.c, videodsp.c) are compiled where they lie under /root/reference by oracle/Makefile.
 * They reference a handful of symbols that live in the host-side (parser / DPB / libavutil) half of the
 * decoder, which the oracle does not build.  This file supplies those few symbols so the path links
 * stand-alone.  Each is a short restatement with the reference location it mirrors.
 */
#include <stdarg.h>
#include <stdio.h>
#include "libavcodec/vvc/vvc_ctu.h"
#include "libavcodec/vvc/vvc_mvs.h"
#include "libavcodec/vvc/vvc_refs.h"

/* libavutil/log.c — the path only logs diagnostics; route them to stderr. */
void av_log(void *avcl, int level, const char *fmt, ...)
{
    va_list ap;
    (void)avcl;
    if (level > AV_LOG_WARNING)
        return;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
}

/* vvc_mvs.c:1739-1749 — 8.5.2.14: round half away from zero by rshift, then scale by 1<<lshift. */
void ff_vvc_round_mv(Mv *mv, const int lshift, const int rshift)
{
    int *c[2] = { &mv->x, &mv->y };
    for (int i = 0; i < 2; i++) {
        int v = *c[i];
        if (rshift)
            v = (v + (1 << (rshift - 1)) - (v >= 0)) >> rshift;
        *c[i] = v * (1 << lshift);
    }
}

/* vvc_mvs.c:1751-1755 — clamp both components to 18 bits signed. */
void ff_vvc_clip_mv(Mv *mv)
{
    mv->x = av_clip(mv->x, -(1 << 17), (1 << 17) - 1);
    mv->y = av_clip(mv->y, -(1 << 17), (1 << 17) - 1);
}

/* vvc_mvs.c:1796-1802 — MvField of the 4x4 unit covering luma (x0, y0). */
MvField *ff_vvc_get_mvf(const VVCFrameContext *fc, const int x0, const int y0)
{
    return fc->tab.mvf + (y0 >> MIN_PU_LOG2) * fc->ps.pps->min_pu_width + (x0 >> MIN_PU_LOG2);
}

/* vvc_refs.c:66-74 — reference picture lists in force at the CTB covering (x0, y0) of `ref`. */
const RefPicList *ff_vvc_get_ref_list(const VVCFrameContext *fc, const VVCFrame *ref, int x0, int y0)
{
    const int sh = fc->ps.sps->ctb_log2_size_y;
    return (const RefPicList *)ref->rpl_tab[(y0 >> sh) * fc->ps.pps->ctb_width + (x0 >> sh)];
}

/* vvc_ctu.c:2532-2538 — QpY of the min-CB covering luma (xc, yc). */
int ff_vvc_get_qPy(const VVCFrameContext *fc, const int xc, const int yc)
{
    const int sh = fc->ps.sps->min_cb_log2_size_y;
    return fc->tab.qp[LUMA][(yc >> sh) * fc->ps.pps->min_cb_width + (xc >> sh)];
}

/* vvc_ctu.c:2468-2495 — tile / slice boundary flags and CTB neighbour availability of CTB (rx, ry). */
void ff_vvc_decode_neighbour(VVCLocalContext *lc, const int x_ctb, const int y_ctb,
    const int rx, const int ry, const int rs)
{
    const VVCFrameContext *fc = lc->fc;
    const VVCPPS *pps         = fc->ps.pps;
    const uint16_t *colbd     = pps->ctb_to_col_bd;
    const uint16_t *rowbd     = pps->ctb_to_row_bd;
    const int16_t *slice      = fc->tab.slice_idx;
    const int ctb             = fc->ps.sps->ctb_size_y;
    int flags = 0;

    lc->end_of_tiles_x = fc->ps.sps->width;
    lc->end_of_tiles_y = fc->ps.sps->height;
    if (colbd[rx] != colbd[rx + 1])
        lc->end_of_tiles_x = FFMIN(x_ctb + ctb, lc->end_of_tiles_x);
    if (rowbd[ry] != rowbd[ry + 1])
        lc->end_of_tiles_y = FFMIN(y_ctb + ctb, lc->end_of_tiles_y);

    if (rx > 0) {
        if (colbd[rx] != colbd[rx - 1])
            flags |= BOUNDARY_LEFT_TILE;
        if (slice[rs] != slice[rs - 1])
            flags |= BOUNDARY_LEFT_SLICE;
    }
    if (ry > 0) {
        if (rowbd[ry] != rowbd[ry - 1])
            flags |= BOUNDARY_UPPER_TILE;
        if (slice[rs] != slice[rs - pps->ctb_width])
            flags |= BOUNDARY_UPPER_SLICE;
    }
    lc->boundary_flags    = flags;
    lc->ctb_left_flag     = rx > 0 && !(flags & BOUNDARY_LEFT_TILE);
    lc->ctb_up_flag       = ry > 0 && !(flags & (BOUNDARY_UPPER_TILE | BOUNDARY_UPPER_SLICE));
    lc->ctb_up_right_flag = lc->ctb_up_flag && colbd[rx] == colbd[rx + 1] && rowbd[ry] == rowbd[ry - 1];
    lc->ctb_up_left_flag  = lc->ctb_left_flag && lc->ctb_up_flag;
}

/* vvc_ctu.c:2497-2511 — availability of the neighbours of block (x0, y0, w, h) inside the current CTB. */
void ff_vvc_set_neighbour_available(VVCLocalContext *lc, const int x0, const int y0, const int w, const int h)
{
    const int log2_ctb = lc->fc->ps.sps->ctb_log2_size_y;
    const int xb = x0 & ((1 << log2_ctb) - 1);
    const int yb = y0 & ((1 << log2_ctb) - 1);
    NeighbourAvailable *na = &lc->na;
    (void)h;

    na->cand_up   = lc->ctb_up_flag || yb;
    na->cand_left = lc->ctb_left_flag || xb;
    na->cand_up_left = (xb || yb) ? (na->cand_left && na->cand_up) : lc->ctb_up_left_flag;
    na->cand_up_right_sap = (xb + w == (1 << log2_ctb)) ? (lc->ctb_up_right_flag && !yb) : na->cand_up;
    na->cand_up_right = na->cand_up_right_sap && (x0 + w) < lc->end_of_tiles_x;
}

/* vvc_ctu.c:2513-2530 frees the refcounted CU/TU lists.  The oracle harness owns CU/TU storage in its
 * own arena, so "free" here only detaches the list. */
void ff_vvc_ctu_free_cus(CTU *ctu)
{
    ctu->cus = NULL;
}

/* vvcdec.c:927-930 (init_default_scale_m, run once by the decoder's static init): flat scaling factor 16. */
#include "libavcodec/vvc/vvc_data.h"
__attribute__((constructor)) static void oracle_static_init(void)
{
    memset(ff_vvc_default_scale_m, 16, sizeof(ff_vvc_default_scale_m));
}

#include "api_common.h"
int64_t vvcref_emit_ns;
VVCREF_EXPORT int64_t vvcref_emit_last_ns(void)
{
    return vvcref_emit_ns;
}
