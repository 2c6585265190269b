/*
 * vvcb200_emit.c — descriptor producer inside the reference decoder (see vvcb200_emit.h).
 * Every function cites the reference code whose traversal it follows; the pixel work of that code is what the
 * engine does from the records written here.
 */
#include <stdlib.h>
#include <string.h>

#include "libavcodec/vvc/vvc_ctu.h"
#include "libavcodec/vvc/vvc_mvs.h"
#include "libavcodec/vvc/vvcdec.h"

#include "vvcb200_emit.h"

/* ---- growable arrays ------------------------------------------------------------------------------------- */
static void *grow(VVCB200EmitBatch *b, void *p, int *cap, int need, size_t elem)
{
    if (need <= *cap)
        return p;
    if (b->fixed) {                          /* the arrays are the caller's (an arena region): they do not move */
        b->error = VVCB200_ERR_TOO_BIG;
        return p;
    }
    int want = *cap ? *cap * 2 : 1024;
    while (want < need)
        want *= 2;
    void *q = realloc(p, (size_t)want * elem);
    if (!q) {
        b->error = VVCB200_ERR_NOMEM;
        return p;
    }
    *cap = want;
    return q;
}

void vvcb200_emit_reset(VVCB200EmitBatch *b)
{
    b->n_blocks = b->n_prof = b->n_tbs = b->n_iblocks = b->n_ctus = 0;
    b->n_levels = 0;
    b->error = 0;
}

void vvcb200_emit_free(VVCB200EmitBatch *b)
{
    if (b->fixed) {
        memset(b, 0, sizeof(*b));
        return;
    }
    free(b->blocks);
    free(b->prof);
    free(b->tbs);
    free(b->levels);
    free(b->iblocks);
    free(b->ctu_first);
    memset(b, 0, sizeof(*b));
}

static VVCB200PredBlock *new_block(VVCB200EmitBatch *b)
{
    b->blocks = grow(b, b->blocks, &b->cap_blocks, b->n_blocks + 1, sizeof(*b->blocks));
    if (b->error)
        return NULL;
    VVCB200PredBlock *r = &b->blocks[b->n_blocks++];
    memset(r, 0, sizeof(*r));
    return r;
}

/* ---- stage INTER -------------------------------------------------------------------------------------------- */
/* flags every record of the unit carries: the slice-level switches derive_weight_uni / derive_weight look at
 * (vvc_inter.c:125-150) and the forward mapping predict_inter applies to non-CIIP units (vvc_inter.c:888-891) */
static int unit_flags(const VVCLocalContext *lc, const CodingUnit *cu)
{
    const VVCSH *sh   = &lc->sc->sh;
    const VVCPPS *pps = lc->fc->ps.pps;
    int f = 0;
    if ((IS_P(sh->r) && pps->r->pps_weighted_pred_flag) || (IS_B(sh->r) && pps->r->pps_weighted_bipred_flag))
        f |= VVCB200_PB_WP;
    if (sh->r->sh_lmcs_used_flag)
        f |= VVCB200_PB_LMCS;
    if (cu->ciip_flag)
        f |= VVCB200_PB_CIIP;
    return f;
}

static void fill_motion(VVCB200PredBlock *r, const MvField *mvf)
{
    r->pred_flag = mvf->pred_flag;
    r->bcw_idx   = mvf->bcw_idx;
    for (int l = 0; l < 2; l++) {
        r->ref_idx[l] = mvf->ref_idx[l];
        r->mv[l][0]   = mvf->mv[l].x;
        r->mv[l][1]   = mvf->mv[l].y;
    }
}

/* pred_regular_blk (vvc_inter.c:787-815): one sub-block per (sbx, sby); a sub-block larger than 16x16 (a unit without
 * DMVR has one sub-block, num_sb_x == num_sb_y == 1) is cut into 16x16 records — translational MC is separable per
 * sample, and 16x16 is the reference's own DMVR / BDOF unit */
static void emit_regular(VVCB200EmitBatch *b, const VVCLocalContext *lc, const CodingUnit *cu)
{
    const PredictionUnit *pu = &cu->pu;
    const MotionInfo *mi     = &pu->mi;
    const int sbw = cu->cb_width / mi->num_sb_x, sbh = cu->cb_height / mi->num_sb_y;
    int flags = unit_flags(lc, cu);

    if (mi->hpel_if_idx)
        flags |= VVCB200_PB_HPEL_IF;
    if (pu->bdof_flag)
        flags |= VVCB200_PB_BDOF;
    if (pu->dmvr_flag)
        flags |= VVCB200_PB_DMVR;
    for (int sby = 0; sby < mi->num_sb_y; sby++)
        for (int sbx = 0; sbx < mi->num_sb_x; sbx++) {
            const int x0 = cu->x0 + sbx * sbw, y0 = cu->y0 + sby * sbh;
            const MvField *mvf = ff_vvc_get_mvf(lc->fc, x0, y0);
            for (int ty = 0; ty < sbh; ty += 16)
                for (int tx = 0; tx < sbw; tx += 16) {
                    VVCB200PredBlock *r = new_block(b);
                    if (!r)
                        return;
                    r->x0 = x0 + tx;
                    r->y0 = y0 + ty;
                    r->w  = FFMIN(16, sbw);
                    r->h  = FFMIN(16, sbh);
                    fill_motion(r, mvf);
                    r->flags = flags;
                }
        }
}

/* pred_affine_blk (vvc_inter.c:828-873): a luma-only record per 4x4 sub-block (filter set 2, PROF per list where
 * cb_prof_flag says so) and, on every second sub-block in x and y, a chroma-only record with derive_affine_mvc's
 * vector (the mean of this sub-block's and its lower-right neighbour's, rounded by ff_vvc_round_mv(.., 0, 1)) */
static void emit_affine(VVCB200EmitBatch *b, const VVCLocalContext *lc, const CodingUnit *cu)
{
    const VVCFrameContext *fc = lc->fc;
    const PredictionUnit *pu  = &cu->pu;
    const MotionInfo *mi      = &pu->mi;
    const int sbw = cu->cb_width / mi->num_sb_x, sbh = cu->cb_height / mi->num_sb_y;
    const int hs = fc->ps.sps->hshift[1], vs = fc->ps.sps->vshift[1];
    const int flags = unit_flags(lc, cu);
    int prof_idx;

    if (sbw != 4 || sbh != 4 || hs != 1 || vs != 1) {
        b->error = VVCB200_ERR_UNSUP;
        return;
    }
    b->prof = grow(b, b->prof, &b->cap_prof, b->n_prof + 1, sizeof(*b->prof));
    if (b->error)
        return;
    prof_idx = b->n_prof++;
    for (int l = 0; l < 2; l++)
        for (int k = 0; k < 16; k++) {
            b->prof[prof_idx].diff_mv_x[l][k] = pu->diff_mv_x[l][k];
            b->prof[prof_idx].diff_mv_y[l][k] = pu->diff_mv_y[l][k];
        }
    for (int sby = 0; sby < mi->num_sb_y; sby++)
        for (int sbx = 0; sbx < mi->num_sb_x; sbx++) {
            const int x = cu->x0 + sbx * sbw, y = cu->y0 + sby * sbh;
            const MvField *mv = ff_vvc_get_mvf(fc, x, y);
            VVCB200PredBlock *r = new_block(b);
            if (!r)
                return;
            r->x0 = x; r->y0 = y; r->w = sbw; r->h = sbh;
            fill_motion(r, mv);
            r->pred_flag = mi->pred_flag;
            r->flags = flags | VVCB200_PB_AFFINE | VVCB200_PB_LUMA_ONLY |
                       (pu->cb_prof_flag[0] ? VVCB200_PB_PROF0 : 0) | (pu->cb_prof_flag[1] ? VVCB200_PB_PROF1 : 0);
            r->aux = prof_idx;
            if (!(sby & 1) && !(sbx & 1)) {
                const MvField *mv2 = ff_vvc_get_mvf(fc, x + hs * sbw, y + vs * sbh);
                MvField mvc = *mv;
                mvc.mv[0].x += mv2->mv[0].x; mvc.mv[0].y += mv2->mv[0].y;
                mvc.mv[1].x += mv2->mv[1].x; mvc.mv[1].y += mv2->mv[1].y;
                ff_vvc_round_mv(mvc.mv + 0, 0, 1);
                ff_vvc_round_mv(mvc.mv + 1, 0, 1);
                r = new_block(b);
                if (!r)
                    return;
                r->x0 = x; r->y0 = y; r->w = sbw << hs; r->h = sbh << vs;
                fill_motion(r, &mvc);
                r->pred_flag = mi->pred_flag;
                r->flags = (flags & ~VVCB200_PB_LMCS) | VVCB200_PB_CHROMA_ONLY;
            }
        }
}

/* pred_gpm_blk (vvc_inter.c:466-521): both partitions' predictions over the whole unit, blended by position; records
 * of at most 16x16 that carry the unit's partition, size and their own offset (VVCB200_GPM_AUX) */
static void emit_gpm(VVCB200EmitBatch *b, const VVCLocalContext *lc, const CodingUnit *cu)
{
    const PredictionUnit *pu = &cu->pu;
    const int flags = (unit_flags(lc, cu) & ~VVCB200_PB_WP) | VVCB200_PB_GPM;    /* GPM is never explicitly weighted */
    const int lw = av_log2(cu->cb_width), lh = av_log2(cu->cb_height);
    int list[2];

    for (int k = 0; k < 2; k++)
        list[k] = pu->gpm_mv[k].pred_flag - PF_L0;
    for (int ty = 0; ty < cu->cb_height; ty += 16)
        for (int tx = 0; tx < cu->cb_width; tx += 16) {
            VVCB200PredBlock *r = new_block(b);
            if (!r)
                return;
            r->x0 = cu->x0 + tx; r->y0 = cu->y0 + ty;
            r->w = FFMIN(16, cu->cb_width); r->h = FFMIN(16, cu->cb_height);
            r->pred_flag = PF_BI;
            for (int k = 0; k < 2; k++) {
                r->ref_idx[k] = pu->gpm_mv[k].ref_idx[list[k]];
                r->mv[k][0]   = pu->gpm_mv[k].mv[list[k]].x;
                r->mv[k][1]   = pu->gpm_mv[k].mv[list[k]].y;
            }
            r->flags = flags;
            r->aux = VVCB200_GPM_AUX(pu->gpm_partition_idx, lw, lh, tx, ty, list[0], list[1]);
        }
}

int vvcb200_emit_inter_cu(VVCB200EmitBatch *b, const VVCLocalContext *lc)
{
    const CodingUnit *cu     = lc->cu;
    const PredictionUnit *pu = &cu->pu;

    if (b->error)
        return b->error;
    if (pu->merge_gpm_flag)
        emit_gpm(b, lc, cu);
    else if (pu->inter_affine_flag)
        emit_affine(b, lc, cu);
    else
        emit_regular(b, lc, cu);
    return b->error;
}

/* ---- batch order --------------------------------------------------------------------------------------------
 * Width classes 16 / 8 / 4 (required by the kernels); inside a class the records with the longest latency first (DMVR,
 * BDOF, GPM), then grouped by the tools they use so that warps running side by side share a code path.  Stable. */
static long long block_key(const VVCB200PredBlock *r)
{
    const int f = r->flags;
    const long long kind = (long long)(f & (VVCB200_PB_GPM | VVCB200_PB_AFFINE | VVCB200_PB_LUMA_ONLY | VVCB200_PB_CHROMA_ONLY |
                                            VVCB200_PB_DMVR | VVCB200_PB_BDOF | VVCB200_PB_WP | VVCB200_PB_PROF0 | VVCB200_PB_PROF1)) * 4 +
                           r->pred_flag + 1024 * 1024 * (r->bcw_idx != 0);
    const int slow = (f & VVCB200_PB_DMVR) ? 0 : (f & VVCB200_PB_BDOF) ? 1 : (f & VVCB200_PB_GPM) ? 2 : 3;
    const int wcls = r->w == 16 ? 0 : r->w == 8 ? 1 : 2;
    return ((long long)wcls << 40) | ((long long)slow << 36) | kind;
}

void vvcb200_emit_order_blocks(VVCB200EmitBatch *b, VVCB200InterParams *params)
{
    const int n = b->n_blocks;
    VVCB200PredBlock *tmp = malloc((size_t)(n ? n : 1) * sizeof(*tmp));
    long long *key = malloc((size_t)(n ? n : 1) * sizeof(*key)), *ktmp = malloc((size_t)(n ? n : 1) * sizeof(*ktmp));
    if (!tmp || !key || !ktmp) {
        b->error = VVCB200_ERR_NOMEM;
        free(tmp); free(key); free(ktmp);
        return;
    }
    for (int i = 0; i < n; i++)
        key[i] = block_key(&b->blocks[i]);
    /* A picture has a few dozen distinct keys (width class x tool set x prediction direction): count per key and scatter —
     * stable, two passes over the records.  (The general merge sort below takes over if there are more than kMaxKeys.) */
    {
        enum { kMaxKeys = 1024 };
        long long dk[kMaxKeys];
        int count[kMaxKeys + 1], nd = 0, ok = 1;
        for (int i = 0; i < n && ok; i++) {
            int lo = 0, hi = nd;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (dk[mid] < key[i]) lo = mid + 1; else hi = mid;
            }
            if (lo == nd || dk[lo] != key[i]) {
                if (nd == kMaxKeys) {
                    ok = 0;
                    break;
                }
                memmove(&dk[lo + 1], &dk[lo], (size_t)(nd - lo) * sizeof(dk[0]));
                dk[lo] = key[i];
                nd++;
            }
        }
        if (ok) {
            int *bucket = (int *)ktmp;                       /* (ktmp is as large as the keys: room for n ints) */
            memset(count, 0, sizeof(int) * (size_t)(nd + 1));
            for (int i = 0; i < n; i++) {
                int lo = 0, hi = nd;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (dk[mid] < key[i]) lo = mid + 1; else hi = mid;
                }
                bucket[i] = lo;
                count[lo + 1]++;
            }
            for (int k = 0; k < nd; k++)
                count[k + 1] += count[k];
            for (int i = 0; i < n; i++)
                tmp[count[bucket[i]]++] = b->blocks[i];
            memcpy(b->blocks, tmp, (size_t)n * sizeof(*tmp));
            goto sorted;
        }
    }
    /* bottom-up merge sort: stable, O(n log n), no recursion */
    VVCB200PredBlock *src = b->blocks, *dst = tmp;
    long long *ks = key, *kd = ktmp;
    for (int width = 1; width < n; width *= 2) {
        for (int lo = 0; lo < n; lo += 2 * width) {
            const int mid = FFMIN(lo + width, n), hi = FFMIN(lo + 2 * width, n);
            int i = lo, j = mid, o = lo;
            while (i < mid || j < hi) {
                const int take_left = j >= hi || (i < mid && ks[i] <= ks[j]);
                const int s = take_left ? i++ : j++;
                dst[o] = src[s];
                kd[o++] = ks[s];
            }
        }
        VVCB200PredBlock *t = src; src = dst; dst = t;
        long long *kt = ks; ks = kd; kd = kt;
    }
    if (src != b->blocks)
        memcpy(b->blocks, src, (size_t)n * sizeof(*src));
sorted:
    params->n_w16 = params->n_w8 = 0;
    for (int i = 0; i < n; i++) {
        params->n_w16 += b->blocks[i].w == 16;
        params->n_w8  += b->blocks[i].w == 8;
    }
    params->dmvr_first = 1;                  /* slow rank 0: the DMVR records lead their width class */
    free(tmp); free(key); free(ktmp);
}

/* ---- array helpers shared with vvcb200_emit_recon.h ------------------------------------------------------------- */
VVCB200TransformBlock *vvcb200_emit_new_tb(VVCB200EmitBatch *b)
{
    b->tbs = grow(b, b->tbs, &b->cap_tbs, b->n_tbs + 1, sizeof(*b->tbs));
    if (b->error)
        return NULL;
    VVCB200TransformBlock *t = &b->tbs[b->n_tbs++];
    memset(t, 0, sizeof(*t));
    return t;
}

VVCB200IntraBlock *vvcb200_emit_new_iblock(VVCB200EmitBatch *b)
{
    b->iblocks = grow(b, b->iblocks, &b->cap_iblocks, b->n_iblocks + 1, sizeof(*b->iblocks));
    if (b->error)
        return NULL;
    VVCB200IntraBlock *r = &b->iblocks[b->n_iblocks++];
    memset(r, 0, sizeof(*r));
    return r;
}

/* room for n more levels; returns their place and the offset they start at */
int16_t *vvcb200_emit_levels(VVCB200EmitBatch *b, int n, uint32_t *off)
{
    if (b->n_levels + (size_t)n > b->cap_levels) {
        if (b->fixed) {
            b->error = VVCB200_ERR_TOO_BIG;
            return NULL;
        }
        size_t want = b->cap_levels ? b->cap_levels * 2 : 1 << 16;
        while (want < b->n_levels + (size_t)n)
            want *= 2;
        int16_t *q = realloc(b->levels, want * sizeof(int16_t));
        if (!q) {
            b->error = VVCB200_ERR_NOMEM;
            return NULL;
        }
        b->levels = q;
        b->cap_levels = want;
    }
    *off = (uint32_t)b->n_levels;
    b->n_levels += n;
    return b->levels + *off;
}

/* CTUs are emitted in raster order: ctu_first[rs] = records so far, ctu_first[rs + 1] is set when the CTU is done */
void vvcb200_emit_ctu_mark(VVCB200EmitBatch *b, int rs, int end)
{
    b->ctu_first = grow(b, b->ctu_first, &b->cap_ctus, rs + 2, sizeof(*b->ctu_first));
    if (b->error)
        return;
    b->ctu_first[rs + end] = (uint32_t)b->n_iblocks;
    if (rs + 1 > b->n_ctus)
        b->n_ctus = rs + 1;
}

/* ---- loop filters --------------------------------------------------------------------------------------------- */
#define EMIT_CTB(tab, x, y) ((tab)[(y) * pps->ctb_width + (x)])

/* The unfilterable sides / corners ff_vvc_sao_filter works out before it touches a sample (vvc_filter.c:163-211). */
static uint8_t sao_no_filter(const VVCFrameContext *fc, const int x_ctb, const int y_ctb)
{
    const VVCPPS *pps        = fc->ps.pps;
    const int16_t *slice_idx = fc->tab.slice_idx;
    const int lfase          = pps->r->pps_loop_filter_across_slices_enabled_flag;
    const int no_tile_filter = pps->r->num_tiles_in_pic > 1 && !pps->r->pps_loop_filter_across_tiles_enabled_flag;
    const int left = x_ctb == 0, top = y_ctb == 0, right = x_ctb == pps->ctb_width - 1, bottom = y_ctb == pps->ctb_height - 1;
    const int here = EMIT_CTB(slice_idx, x_ctb, y_ctb);
    int lt = 0, rt = 0, ut = 0, bt = 0, flags = 0;
#define OTHER_SLICE(dx, dy) (!lfase && here != EMIT_CTB(slice_idx, x_ctb + (dx), y_ctb + (dy)))

    if (!no_tile_filter && lfase)
        return 0;
    if (!left) {
        lt = no_tile_filter && pps->ctb_to_col_bd[x_ctb] == x_ctb;
        flags |= (OTHER_SLICE(-1, 0) || lt) << 0;
    }
    if (!right) {
        rt = no_tile_filter && pps->ctb_to_col_bd[x_ctb] != pps->ctb_to_col_bd[x_ctb + 1];
        flags |= (OTHER_SLICE(1, 0) || rt) << 1;
    }
    if (!top) {
        ut = no_tile_filter && pps->ctb_to_row_bd[y_ctb] == y_ctb;
        flags |= (OTHER_SLICE(0, -1) || ut) << 2;
    }
    if (!bottom) {
        bt = no_tile_filter && pps->ctb_to_row_bd[y_ctb] != pps->ctb_to_row_bd[y_ctb + 1];
        flags |= (OTHER_SLICE(0, 1) || bt) << 3;
    }
    if (!left && !top)
        flags |= (OTHER_SLICE(-1, -1) || lt || ut) << 4;
    if (!top && !right)
        flags |= (OTHER_SLICE(1, -1) || rt || ut) << 5;
    if (!right && !bottom)
        flags |= (OTHER_SLICE(1, 1) || rt || bt) << 6;
    if (!left && !bottom)
        flags |= (OTHER_SLICE(-1, 1) || lt || bt) << 7;
#undef OTHER_SLICE
    return flags;
}

/* fc->tab.mvf / tb_pos / cb_pos / flags of one 4x4 unit -> the two compact records vvcb200_deblock_derive_edges reads.
 * What vvc_deblock_bs looks at: vvc_filter.c:308-371 (motion), 399-475 (TB sizes), 477-782 (TB / CU / sub-block edges). */
static void emit_unit(const VVCB200EmitFilters *out, const VVCFrameContext *fc, const RefPicList *rpl, const int ux, const int uy)
{
    const int o = uy * fc->ps.pps->min_pu_width + ux;
    const MvField *m = &fc->tab.mvf[o];
    VVCB200MotionUnit *mu = &out->motion[o];
    VVCB200PartitionUnit *pu = &out->part[o];

    memset(mu, 0, sizeof(*mu));
    memset(pu, 0, sizeof(*pu));
    mu->pred_flag = m->pred_flag;
    mu->ciip_flag = m->ciip_flag;
    for (int l = 0; l < 2; l++) {
        mu->mv[l][0] = m->mv[l].x;
        mu->mv[l][1] = m->mv[l].y;
        /* the reference compares the POCs rpl->list[ref_idx] (vvc_filter.c:320-352); the picture behind one is the same
         * question, and the DPB slot answers it without a search */
        if (rpl && m->ref_idx[l] >= 0 && rpl[l].ref[m->ref_idx[l]])
            mu->ref_id[l] = (int8_t)(rpl[l].ref[m->ref_idx[l]] - fc->DPB);
    }
    pu->tb_off   = ((ux - (fc->tab.tb_pos_x0[LUMA][o] >> 2)) & 15) | (((uy - (fc->tab.tb_pos_y0[LUMA][o] >> 2)) & 15) << 4);
    pu->tb_size  = av_log2(FFMAX(fc->tab.tb_width[LUMA][o], 1)) | (av_log2(FFMAX(fc->tab.tb_height[LUMA][o], 1)) << 3) |
                   (!!fc->tab.tu_coded_flag[LUMA][o] << 6) | (!!fc->tab.pcmf[LUMA][o] << 7);
    pu->cu_off_x = ((ux - (fc->tab.cb_pos_x[LUMA][o] >> 2)) & 31) | ((fc->tab.msf[o] || fc->tab.iaf[o]) << 5) |
                   ((fc->tab.cb_width[LUMA][o] > 8) << 6) | ((fc->tab.cb_height[LUMA][o] > 8) << 7);
    pu->cu_off_y = (uy - (fc->tab.cb_pos_y[LUMA][o] >> 2)) & 31;
    pu->ctb_off  = ((ux - (fc->tab.tb_pos_x0[CHROMA][o] >> 2)) & 15) | (((uy - (fc->tab.tb_pos_y0[CHROMA][o] >> 2)) & 15) << 4);
    pu->ctb_size = av_log2(FFMAX(fc->tab.tb_width[CHROMA][o], 1)) | (av_log2(FFMAX(fc->tab.tb_height[CHROMA][o], 1)) << 3) |
                   (!!fc->tab.tu_coded_flag[CB][o] << 6) | (!!fc->tab.tu_coded_flag[CR][o] << 7);
    pu->cflags   = (fc->tab.tu_joint_cbcr_residual_flag[o] != 0) | (!!fc->tab.pcmf[CHROMA][o] << 1);
}

void vvcb200_emit_filters_ctu(const VVCB200EmitFilters *out, const VVCLocalContext *lc, const int rx, const int ry)
{
    const VVCFrameContext *fc = lc->fc;
    const VVCSPS *sps         = fc->ps.sps;
    const VVCPPS *pps         = fc->ps.pps;
    const int rs              = ry * pps->ctb_width + rx;
    const int lfase           = pps->r->pps_loop_filter_across_slices_enabled_flag;
    const int lfate           = pps->r->pps_loop_filter_across_tiles_enabled_flag;

    if (out->sao) {
        const SAOParams *s = &fc->tab.sao[rs];
        VVCB200SaoCtb *d = &out->sao[rs];
        memset(d, 0, sizeof(*d));
        for (int c = 0; c < 3; c++) {
            d->type_idx[c]      = s->type_idx[c];
            d->band_position[c] = s->band_position[c];
            d->eo_class[c]      = s->eo_class[c];
            for (int k = 0; k < 5; k++)
                d->offset_val[c][k] = s->offset_val[c][k];
        }
        d->no_filter = sao_no_filter(fc, rx, ry);
    }
    if (out->alf) {
        const ALFParams *a = &fc->tab.alf[rs];
        VVCB200AlfCtb *d = &out->alf[rs];
        /* ff_vvc_alf_filter's edges[] (vvc_filter.c:1266-1280) without the picture borders, which the device knows */
        const int last_x = rx == pps->ctb_width - 1, last_y = ry == pps->ctb_height - 1;
        int left = 0, top = 0, right = 0, bottom = 0;
        if (!lfate) {
            left   |= !!(lc->boundary_flags & BOUNDARY_LEFT_TILE);
            top    |= !!(lc->boundary_flags & BOUNDARY_UPPER_TILE);
            right  |= !last_x && pps->ctb_to_col_bd[rx] != pps->ctb_to_col_bd[rx + 1];
            bottom |= !last_y && pps->ctb_to_row_bd[ry] != pps->ctb_to_row_bd[ry + 1];
        }
        if (!lfase) {
            left   |= !!(lc->boundary_flags & BOUNDARY_LEFT_SLICE);
            top    |= !!(lc->boundary_flags & BOUNDARY_UPPER_SLICE);
            right  |= !last_x && EMIT_CTB(fc->tab.slice_idx, rx, ry) != EMIT_CTB(fc->tab.slice_idx, rx + 1, ry);
            bottom |= !last_y && EMIT_CTB(fc->tab.slice_idx, rx, ry) != EMIT_CTB(fc->tab.slice_idx, rx, ry + 1);
        }
        memset(d, 0, sizeof(*d));
        for (int c = 0; c < 3; c++)
            d->ctb_flag[c] = a->ctb_flag[c];
        d->filt_set_idx_y = a->ctb_filt_set_idx_y;
        for (int c = 0; c < 2; c++) {
            d->alt_idx[c] = a->alf_ctb_filter_alt_idx[c];
            d->cc_idc[c]  = a->ctb_cc_idc[c];
        }
        d->edges = left | (top << 1) | (right << 2) | (bottom << 3);
    }
    if (out->ctb_offsets) {
        const DBParams *p = &fc->tab.deblock[rs];
        for (int c = 0; c < 3; c++) {
            out->ctb_offsets[rs * 6 + c]     = p->beta_offset[c];
            out->ctb_offsets[rs * 6 + 3 + c] = p->tc_offset[c];
        }
    }
    if (out->ctb_edges)             /* vvc_filter.c:505-512, 572-579: CTB sides the deblocking filter must not cross */
        out->ctb_edges[rs] = ((!lfase && (lc->boundary_flags & BOUNDARY_LEFT_SLICE)) || (!lfate && (lc->boundary_flags & BOUNDARY_LEFT_TILE))) |
                             (((!lfase && (lc->boundary_flags & BOUNDARY_UPPER_SLICE)) || (!lfate && (lc->boundary_flags & BOUNDARY_UPPER_TILE))) << 1);
    if (out->motion && out->part) {
        const int x0 = rx << sps->ctb_log2_size_y, y0 = ry << sps->ctb_log2_size_y;
        const int ux1 = FFMIN(x0 + sps->ctb_size_y, pps->width) >> 2, uy1 = FFMIN(y0 + sps->ctb_size_y, pps->height) >> 2;
        const RefPicList *rpl = lc->sc->rpl;                    /* the lists of the slice this CTB belongs to */
        for (int uy = y0 >> 2; uy < uy1; uy++)
            for (int ux = x0 >> 2; ux < ux1; ux++)
                emit_unit(out, fc, rpl, ux, uy);
    }
}

void vvcb200_emit_alf_params(VVCB200AlfParams *p, const VVCFrameContext *fc, const SliceContext *sc)
{
    const H266RawSliceHeader *rsh = sc->sh.r;
    const VVCALF *chroma = fc->ps.alf_list[rsh->sh_alf_aps_id_chroma];

    memset(p, 0, sizeof(*p));
    p->bit_depth     = fc->ps.sps->bit_depth;
    p->ctb_log2_size = fc->ps.sps->ctb_log2_size_y;
    for (int k = 0; k < rsh->sh_num_alf_aps_ids_luma && k < 7; k++) {          /* alf_get_coeff_and_clip, vvc_filter.c:1155-1160 */
        const VVCALF *aps = fc->ps.alf_list[rsh->sh_alf_aps_id_luma[k]];
        if (!aps)
            continue;
        memcpy(p->luma_coeff[k], aps->luma_coeff, sizeof(p->luma_coeff[k]));
        memcpy(p->luma_clip_idx[k], aps->luma_clip_idx, sizeof(p->luma_clip_idx[k]));
    }
    if (chroma) {                                                                 /* alf_filter_chroma, 1196-1211 */
        memcpy(p->chroma_coeff, chroma->chroma_coeff, sizeof(p->chroma_coeff));
        memcpy(p->chroma_clip_idx, chroma->chroma_clip_idx, sizeof(p->chroma_clip_idx));
    }
    for (int i = 0; i < 2; i++) {                                                 /* alf_filter_cc, 1213-1228 */
        const VVCALF *cc = fc->ps.alf_list[i ? rsh->sh_alf_cc_cr_aps_id : rsh->sh_alf_cc_cb_aps_id];   /* NULL: no such APS */
        p->cc_enabled[i] = cc != NULL;
        if (cc)
            for (int f = 0; f < 4; f++)
                for (int k = 0; k < 7; k++)
                    p->cc_coeff[i][f][k] = cc->cc_coeff[i][f][k];
    }
}

void vvcb200_emit_deblock_params(VVCB200DeblockParams *p, const VVCFrameContext *fc)
{
    const VVCSPS *sps = fc->ps.sps;

    memset(p, 0, sizeof(*p));
    p->bit_depth     = sps->bit_depth;
    p->ctb_log2_size = sps->ctb_log2_size_y;
    p->ladf_enabled  = sps->r->sps_ladf_enabled_flag;                            /* get_qp_y / ladf, vvc_filter.c:829-859 */
    p->ladf_lowest_interval_qp_offset = sps->r->sps_ladf_lowest_interval_qp_offset;
    p->num_ladf_intervals = sps->num_ladf_intervals;
    for (int i = 0; i < 4; i++)
        p->ladf_qp_offset[i] = sps->r->sps_ladf_qp_offset[i];
    for (int i = 0; i < 5; i++)
        p->ladf_interval_lower_bound[i] = sps->ladf_interval_lower_bound[i];
}

void vvcb200_emit_store_refined_mv(const VVCFrameContext *fc, const VVCB200PredBlock *blocks, const int n, const int32_t *refined_mv)
{
    const int min_pu_width = fc->ps.pps->min_pu_width;

    for (int i = 0; i < n; i++) {
        const VVCB200PredBlock *b = &blocks[i];
        if (!(b->flags & VVCB200_PB_DMVR))
            continue;
        MvField m = fc->tab.mvf[(b->y0 >> MIN_PU_LOG2) * min_pu_width + (b->x0 >> MIN_PU_LOG2)];
        for (int l = 0; l < 2; l++) {
            m.mv[l].x = refined_mv[i * 4 + l * 2];
            m.mv[l].y = refined_mv[i * 4 + l * 2 + 1];
        }
        for (int y = b->y0; y < b->y0 + b->h; y += MIN_PU_SIZE)              /* set_dmvr_info, vvc_inter.c:749-761 */
            for (int x = b->x0; x < b->x0 + b->w; x += MIN_PU_SIZE)
                fc->ref->tab_dmvr_mvf[(y >> MIN_PU_LOG2) * min_pu_width + (x >> MIN_PU_LOG2)] = m;
    }
}
