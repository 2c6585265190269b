/*
 * oracle/ref/api_inter.c — TEST INFRASTRUCTURE.
 * Pulls the reference's inter-prediction stage driver (libavcodec/vvc/vvc_inter.c, read in place) into
 * this translation unit so that its file-static helpers (DMVR search, weight derivation, edge
 * emulation: vvc_inter.c:33-110, 642-748) are reachable from the flat test API below.
 */
#include "libavcodec/vvc/vvc_inter.c"
#include "api_common.h"
#include "vvcb200_emit.h"

#include <stdlib.h>

typedef struct InterEnv {
    VVCFrameContext      fc;
    SliceContext         sc;
    AVFrame              frame;
    VVCSPS               sps;
    VVCPPS               pps;
    H266RawSPS           rsps;
    H266RawPPS           rpps;
    H266RawPictureHeader rph;
    H266RawSliceHeader   rsh;
    RefPicList           rpl[2];
    VVCFrame             cur;
    VVCFrame             refs[16];
    AVFrame              ref_frames[16];
    CodingUnit           cu;
} InterEnv;

/*
 * INTER stage for a batch of prediction sub-blocks on host planes through the reference's own
 * predict_inter() (vvc_inter.c:875-892): each descriptor becomes a one-sub-block CodingUnit whose MvField
 * is planted in fc->tab.mvf, so the reference's MV -> address arithmetic, edge emulation, weight
 * derivation, DMVR search and BDOF all run unmodified.  Arguments mirror vvcb200_inter_batch().
 * refined_mv (optional, n x 2 x 2 ints): MvField.mv after DMVR as stored in tab_dmvr_mvf.
 */
static int g_store_check;

VVCREF_EXPORT int vvcref_inter_batch(const VVCRefPlanes *dst, const VVCRefPlanes *refs, int n_refs,
                                     const VVCB200PredBlock *blocks, int n, const VVCB200InterParams *p,
                                     const uint16_t *lmcs_fwd_lut, int32_t *refined_mv)
{
    InterEnv *e         = calloc(1, sizeof(*e));
    VVCLocalContext *lc = NULL;
    int ret             = -1;
    int min_pu_width, min_pu_height;

    if (!e || n_refs > 16 || !(lc = aligned_alloc(64, (sizeof(*lc) + 63) & ~(size_t)63)))
        return -1;
    memset(lc, 0, sizeof(*lc));
    e->fc.frame   = &e->frame;
    e->fc.ps.sps  = &e->sps;
    e->fc.ps.pps  = &e->pps;
    e->fc.ps.ph.r = &e->rph;
    e->fc.ref     = &e->cur;
    e->sps.r      = &e->rsps;
    e->pps.r      = &e->rpps;
    e->sc.sh.r    = &e->rsh;
    e->sc.rpl     = e->rpl;
    lc->fc        = &e->fc;
    lc->sc        = &e->sc;
    lc->cu        = &e->cu;
    for (int c = 0; c < 3; c++) {
        e->frame.data[c]     = dst->plane[c];
        e->frame.linesize[c] = dst->stride[c];
    }
    e->sps.width  = e->pps.width  = dst->width[0];
    e->sps.height = e->pps.height = dst->height[0];
    e->sps.bit_depth        = p->bit_depth;
    e->sps.pixel_shift      = 1;
    e->sps.hshift[1] = e->sps.hshift[2] = 1;
    e->sps.vshift[1] = e->sps.vshift[2] = 1;
    e->sps.ctb_log2_size_y  = 7;
    e->sps.ctb_size_y       = 128;
    e->rsps.sps_chroma_format_idc = 1;
    e->rsh.sh_slice_type    = VVC_SLICE_TYPE_B;
    min_pu_width  = e->pps.min_pu_width  = (dst->width[0]  + 3) >> 2;
    min_pu_height = e->pps.min_pu_height = (dst->height[0] + 3) >> 2;
    e->fc.tab.mvf        = calloc((size_t)min_pu_width * min_pu_height, sizeof(MvField));
    e->cur.tab_dmvr_mvf  = calloc((size_t)min_pu_width * min_pu_height, sizeof(MvField));
    if (!e->fc.tab.mvf || !e->cur.tab_dmvr_mvf)
        goto end;
    ff_vvc_dsp_init(&e->fc.vvcdsp, p->bit_depth);
    ff_videodsp_init(&e->fc.vdsp, p->bit_depth);
    for (int i = 0; i < n_refs; i++) {
        for (int c = 0; c < 3; c++) {
            e->ref_frames[i].data[c]     = refs[i].plane[c];
            e->ref_frames[i].linesize[c] = refs[i].stride[c];
        }
        e->refs[i].frame = &e->ref_frames[i];
    }
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < 15; i++)
            e->rpl[l].ref[i] = p->ref_slot[l][i] < n_refs ? &e->refs[p->ref_slot[l][i]] : NULL;
    {
        PredWeightTable *w = &e->sc.sh.pwt;
        w->log2_denom[0] = p->log2_denom[0];
        w->log2_denom[1] = p->log2_denom[1];
        for (int l = 0; l < 2; l++)
            for (int c = 0; c < 3; c++)
                for (int i = 0; i < MAX_WEIGHTS; i++) {
                    w->weight[l][c][i] = p->weight[l][c][i];
                    w->offset[l][c][i] = p->offset[l][c][i];
                }
    }
    if (lmcs_fwd_lut)
        memcpy(e->fc.ps.lmcs.fwd_lut, lmcs_fwd_lut, sizeof(uint16_t) << p->bit_depth);

    for (int i = 0; i < n; i++) {
        const VVCB200PredBlock *b = &blocks[i];
        CodingUnit *cu = &e->cu;
        MvField mvf    = { 0 };

        if (b->flags & (VVCB200_PB_LUMA_ONLY | VVCB200_PB_CHROMA_ONLY | VVCB200_PB_AFFINE | VVCB200_PB_CIIP))
            goto end;                       /* affine / CIIP records are driven by the CU-level harness */
        memset(cu, 0, sizeof(*cu));
        cu->x0 = b->x0; cu->y0 = b->y0; cu->cb_width = b->w; cu->cb_height = b->h;
        cu->pred_mode = MODE_INTER;
        cu->pu.mi.num_sb_x = cu->pu.mi.num_sb_y = 1;
        cu->pu.mi.hpel_if_idx = !!(b->flags & VVCB200_PB_HPEL_IF);
        cu->pu.mi.pred_flag   = b->pred_flag;
        cu->pu.bdof_flag      = !!(b->flags & VVCB200_PB_BDOF);
        cu->pu.dmvr_flag      = !!(b->flags & VVCB200_PB_DMVR);
        e->rpps.pps_weighted_bipred_flag = !!(b->flags & VVCB200_PB_WP);
        e->rsh.sh_lmcs_used_flag         = !!(b->flags & VVCB200_PB_LMCS);
        for (int l = 0; l < 2; l++) {
            mvf.mv[l].x    = b->mv[l][0];
            mvf.mv[l].y    = b->mv[l][1];
            mvf.ref_idx[l] = b->ref_idx[l];
        }
        mvf.pred_flag   = b->pred_flag;
        mvf.bcw_idx     = b->bcw_idx;
        mvf.hpel_if_idx = cu->pu.mi.hpel_if_idx;
        for (int y = b->y0 >> 2; y < (b->y0 + b->h) >> 2; y++)
            for (int x = b->x0 >> 2; x < (b->x0 + b->w) >> 2; x++)
                e->fc.tab.mvf[y * min_pu_width + x] = mvf;
        predict_inter(lc);
        if (refined_mv) {
            const MvField *r = cu->pu.dmvr_flag ? &e->cur.tab_dmvr_mvf[(b->y0 >> 2) * min_pu_width + (b->x0 >> 2)] : &mvf;
            for (int l = 0; l < 2; l++) {
                refined_mv[i * 4 + l * 2 + 0] = r->mv[l].x;
                refined_mv[i * 4 + l * 2 + 1] = r->mv[l].y;
            }
        }
    }
    ret = 0;
    if (g_store_check && refined_mv) {
        /* f1: the emitter's write-back of the refined vectors (vvcb200_emit_store_refined_mv) into a fresh table must
         * leave what the reference's own set_dmvr_info calls left in tab_dmvr_mvf */
        const size_t units = (size_t)min_pu_width * min_pu_height;
        MvField *want = e->cur.tab_dmvr_mvf;
        e->cur.tab_dmvr_mvf = calloc(units, sizeof(MvField));
        vvcb200_emit_store_refined_mv(&e->fc, blocks, n, refined_mv);
        for (size_t u = 0; u < units; u++) {
            const MvField *a = &want[u], *b = &e->cur.tab_dmvr_mvf[u];
            ret += a->mv[0].x != b->mv[0].x || a->mv[0].y != b->mv[0].y || a->mv[1].x != b->mv[1].x || a->mv[1].y != b->mv[1].y ||
                   a->ref_idx[0] != b->ref_idx[0] || a->ref_idx[1] != b->ref_idx[1] || a->pred_flag != b->pred_flag ||
                   a->bcw_idx != b->bcw_idx || a->hpel_if_idx != b->hpel_if_idx || a->ciip_flag != b->ciip_flag;
        }
        free(want);
    }
end:
    free(e->fc.tab.mvf);
    free(e->cur.tab_dmvr_mvf);
    free(lc);
    free(e);
    return ret;
}

/* Number of 4x4 units whose tab_dmvr_mvf entry differs between the reference's run and the emitter's write-back. */
VVCREF_EXPORT int vvcref_emit_store_refined_check(const VVCRefPlanes *dst, const VVCRefPlanes *refs, int n_refs,
                                                  const VVCB200PredBlock *blocks, int n, const VVCB200InterParams *p,
                                                  int32_t *refined_mv)
{
    int ret;
    g_store_check = 1;
    ret = vvcref_inter_batch(dst, refs, n_refs, blocks, n, p, NULL, refined_mv);
    g_store_check = 0;
    return ret;
}

/* inter.put[is_chroma][log2(w)-1][my != 0][mx != 0] (vvcdsp.h:49-51): src points at the block origin inside a
 * caller-padded buffer; filters are the reference's own tables, set `filter_set`. dst stride is MAX_PB_SIZE. */
VVCREF_EXPORT void vvcref_put(int bd, int is_chroma, int16_t *dst, const uint8_t *src, ptrdiff_t src_stride,
                              int w, int h, int mx, int my, int filter_set)
{
    VVCDSPContext c;
    const int8_t *hf = is_chroma ? ff_vvc_inter_chroma_filters[filter_set][mx] : ff_vvc_inter_luma_filters[filter_set][mx];
    const int8_t *vf = is_chroma ? ff_vvc_inter_chroma_filters[filter_set][my] : ff_vvc_inter_luma_filters[filter_set][my];
    ff_vvc_dsp_init(&c, bd);
    c.inter.put[is_chroma][av_log2(w) - 1][!!my][!!mx](dst, src, src_stride, h, hf, vf, w);
}

/*
 * INTER stage for CU-level descriptions (regular, affine + PROF, GPM) through the reference's own predict_inter():
 * the reference splits affine units into 4x4 sub-blocks, averages vectors for chroma (derive_affine_mvc) and blends
 * GPM partitions itself.  sb_mv: affine sub-block vectors, see VVCRefInterCU.sb_mv_off.
 */
/* When set, vvcref_inter_cus runs the product's emitter (ffmpeg_b200/emit/vvcb200_emit.c) on every coding unit it has
 * built, in place of the reference's predict_inter(): the same reference structs in, engine records out. */
static VVCB200EmitBatch *g_emit;

VVCREF_EXPORT int vvcref_inter_cus(const VVCRefPlanes *dst, const VVCRefPlanes *refs, int n_refs,
                                   const VVCRefInterCU *cus, int n, const int32_t *sb_mv,
                                   const VVCB200InterParams *p, const uint16_t *lmcs_fwd_lut)
{
    InterEnv *e         = calloc(1, sizeof(*e));
    VVCLocalContext *lc = NULL;
    int ret             = -1;
    int min_pu_width, min_pu_height;

    if (!e || n_refs > 16 || !(lc = aligned_alloc(64, (sizeof(*lc) + 63) & ~(size_t)63)))
        return -1;
    memset(lc, 0, sizeof(*lc));
    e->fc.frame   = &e->frame;
    e->fc.ps.sps  = &e->sps;
    e->fc.ps.pps  = &e->pps;
    e->fc.ps.ph.r = &e->rph;
    e->fc.ref     = &e->cur;
    e->sps.r      = &e->rsps;
    e->pps.r      = &e->rpps;
    e->sc.sh.r    = &e->rsh;
    e->sc.rpl     = e->rpl;
    lc->fc        = &e->fc;
    lc->sc        = &e->sc;
    lc->cu        = &e->cu;
    for (int c = 0; c < 3; c++) {
        e->frame.data[c]     = dst->plane[c];
        e->frame.linesize[c] = dst->stride[c];
    }
    e->sps.width  = e->pps.width  = dst->width[0];
    e->sps.height = e->pps.height = dst->height[0];
    e->sps.bit_depth        = p->bit_depth;
    e->sps.pixel_shift      = 1;
    e->sps.hshift[1] = e->sps.hshift[2] = 1;
    e->sps.vshift[1] = e->sps.vshift[2] = 1;
    e->sps.ctb_log2_size_y  = 7;
    e->sps.ctb_size_y       = 128;
    e->rsps.sps_chroma_format_idc = 1;
    e->rsh.sh_slice_type    = VVC_SLICE_TYPE_B;
    min_pu_width  = e->pps.min_pu_width  = (dst->width[0]  + 3) >> 2;
    min_pu_height = e->pps.min_pu_height = (dst->height[0] + 3) >> 2;
    e->fc.tab.mvf        = calloc((size_t)min_pu_width * min_pu_height, sizeof(MvField));
    e->cur.tab_dmvr_mvf  = calloc((size_t)min_pu_width * min_pu_height, sizeof(MvField));
    if (!e->fc.tab.mvf || !e->cur.tab_dmvr_mvf)
        goto end;
    ff_vvc_dsp_init(&e->fc.vvcdsp, p->bit_depth);
    ff_videodsp_init(&e->fc.vdsp, p->bit_depth);
    for (int i = 0; i < n_refs; i++) {
        for (int c = 0; c < 3; c++) {
            e->ref_frames[i].data[c]     = refs[i].plane[c];
            e->ref_frames[i].linesize[c] = refs[i].stride[c];
        }
        e->refs[i].frame = &e->ref_frames[i];
    }
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < 15; i++)
            e->rpl[l].ref[i] = p->ref_slot[l][i] < n_refs ? &e->refs[p->ref_slot[l][i]] : NULL;
    {
        PredWeightTable *w = &e->sc.sh.pwt;
        w->log2_denom[0] = p->log2_denom[0];
        w->log2_denom[1] = p->log2_denom[1];
        for (int l = 0; l < 2; l++)
            for (int c = 0; c < 3; c++)
                for (int i = 0; i < MAX_WEIGHTS; i++) {
                    w->weight[l][c][i] = p->weight[l][c][i];
                    w->offset[l][c][i] = p->offset[l][c][i];
                }
    }
    if (lmcs_fwd_lut)
        memcpy(e->fc.ps.lmcs.fwd_lut, lmcs_fwd_lut, sizeof(uint16_t) << p->bit_depth);

    for (int i = 0; i < n; i++) {
        const VVCRefInterCU *d = &cus[i];
        CodingUnit *cu = &e->cu;
        PredictionUnit *pu = &cu->pu;
        MvField mvf = { 0 };

        memset(cu, 0, sizeof(*cu));
        cu->x0 = d->x0; cu->y0 = d->y0; cu->cb_width = d->w; cu->cb_height = d->h;
        cu->pred_mode = MODE_INTER;
        pu->mi.hpel_if_idx = !!(d->flags & VVCB200_PB_HPEL_IF);
        pu->mi.pred_flag   = d->pred_flag;
        pu->bdof_flag      = !!(d->flags & VVCB200_PB_BDOF);
        pu->dmvr_flag      = !!(d->flags & VVCB200_PB_DMVR);
        cu->ciip_flag      = !!(d->flags & VVCB200_PB_CIIP);   /* predict_inter leaves such a unit to RECON; the emitter flags it */
        e->rpps.pps_weighted_bipred_flag = !!(d->flags & VVCB200_PB_WP);
        e->rsh.sh_lmcs_used_flag         = !!(d->flags & VVCB200_PB_LMCS);
        mvf.pred_flag   = d->pred_flag;
        mvf.ciip_flag   = cu->ciip_flag;
        mvf.bcw_idx     = d->bcw_idx;
        mvf.hpel_if_idx = pu->mi.hpel_if_idx;
        for (int l = 0; l < 2; l++) {
            mvf.mv[l].x    = d->mv[l][0];
            mvf.mv[l].y    = d->mv[l][1];
            mvf.ref_idx[l] = d->ref_idx[l];
        }
        if (d->kind == 1) {
            const int32_t *sb = sb_mv + (size_t)d->sb_mv_off * 4;
            pu->inter_affine_flag = 1;
            pu->mi.num_sb_x = d->w >> 2;
            pu->mi.num_sb_y = d->h >> 2;
            for (int l = 0; l < 2; l++) {
                pu->cb_prof_flag[l] = d->cb_prof_flag[l];
                for (int k = 0; k < 16; k++) {
                    pu->diff_mv_x[l][k] = d->diff_mv_x[l][k];
                    pu->diff_mv_y[l][k] = d->diff_mv_y[l][k];
                }
            }
            for (int sy = 0; sy < pu->mi.num_sb_y; sy++)
                for (int sx = 0; sx < pu->mi.num_sb_x; sx++, sb += 4) {
                    MvField m = mvf;
                    m.mv[0].x = sb[0]; m.mv[0].y = sb[1]; m.mv[1].x = sb[2]; m.mv[1].y = sb[3];
                    e->fc.tab.mvf[((d->y0 >> 2) + sy) * min_pu_width + (d->x0 >> 2) + sx] = m;
                }
        } else {
            pu->mi.num_sb_x = pu->mi.num_sb_y = 1;
            if (d->kind == 2) {
                pu->merge_gpm_flag    = 1;
                pu->gpm_partition_idx = d->gpm_partition_idx;
                for (int k = 0; k < 2; k++) {
                    const int lx = d->gpm_list[k];
                    memset(&pu->gpm_mv[k], 0, sizeof(pu->gpm_mv[k]));
                    pu->gpm_mv[k].pred_flag   = lx + PF_L0;
                    pu->gpm_mv[k].ref_idx[lx] = d->ref_idx[k];
                    pu->gpm_mv[k].mv[lx].x    = d->mv[k][0];
                    pu->gpm_mv[k].mv[lx].y    = d->mv[k][1];
                }
            }
            for (int y = d->y0 >> 2; y < (d->y0 + d->h) >> 2; y++)
                for (int x = d->x0 >> 2; x < (d->x0 + d->w) >> 2; x++)
                    e->fc.tab.mvf[y * min_pu_width + x] = mvf;
        }
        if (g_emit) {
            const int64_t t0 = vvcref_now_ns();
            vvcb200_emit_inter_cu(g_emit, lc);              /* the emitter in predict_inter's place */
            vvcref_emit_ns += vvcref_now_ns() - t0;
        } else
            predict_inter(lc);
    }
    ret = g_emit ? g_emit->error : 0;
end:
    free(e->fc.tab.mvf);
    free(e->cur.tab_dmvr_mvf);
    free(lc);
    free(e);
    return ret;
}


/*
 * f1 test entry: the CU-level descriptions are turned into the reference's CodingUnit / PredictionUnit / MvField state
 * exactly as vvcref_inter_cus does, and the emitter writes the engine's records from THAT.  blocks / prof: caller
 * arrays of cap_blocks / cap_prof records; returns the counts through n_blocks / n_prof.  order != 0: also apply the
 * batch order and return n_w16 / n_w8 in `p`.
 */
VVCREF_EXPORT int vvcref_emit_inter_cus(int width, int height, const VVCRefInterCU *cus, int n, const int32_t *sb_mv,
                                        VVCB200InterParams *p, VVCB200PredBlock *blocks, int cap_blocks, int *n_blocks,
                                        VVCB200ProfTable *prof, int cap_prof, int *n_prof, int order)
{
    VVCB200EmitBatch batch = { 0 };
    VVCRefPlanes pl = { { NULL, NULL, NULL }, { 0, 0, 0 }, { width, width / 2, width / 2 }, { height, height / 2, height / 2 } };
    int ret;
    g_emit = &batch;
    vvcref_emit_ns = 0;
    ret = vvcref_inter_cus(&pl, &pl, 0, cus, n, sb_mv, p, NULL);
    g_emit = NULL;
    {   /* the per-unit timer calls are not the emitter's: take their own cost (measured here) off again */
        int64_t t0 = vvcref_now_ns(), t1 = t0;
        for (int i = 0; i < 2000; i++)
            t1 = vvcref_now_ns();
        vvcref_emit_ns -= (t1 - t0) * n / 2000;
        if (vvcref_emit_ns < 0)
            vvcref_emit_ns = 0;
    }
    if (!ret && order) {
        const int64_t t0 = vvcref_now_ns();
        vvcb200_emit_order_blocks(&batch, p);
        vvcref_emit_ns += vvcref_now_ns() - t0;
    }
    if (!ret && (batch.n_blocks > cap_blocks || batch.n_prof > cap_prof))
        ret = -2;
    if (!ret) {
        memcpy(blocks, batch.blocks, (size_t)batch.n_blocks * sizeof(*blocks));
        memcpy(prof, batch.prof, (size_t)batch.n_prof * sizeof(*prof));
        *n_blocks = batch.n_blocks;
        *n_prof = batch.n_prof;
    }
    vvcb200_emit_free(&batch);
    return ret;
}
