/*
 * vvcb200_emit_recon.h — stage RECON of the descriptor emitter.
 *
 * To be #included at the END of libavcodec/vvc/vvc_intra.c (the maintainer's patch adds that one line): the code below
 * calls that file's static helpers — derive_qp, derive_transform_type, derive_ilfnst_pred_mode_intra,
 * add_reconstructed_area, get_luma_predict_unit, get_chroma_predict_unit, is_cclm — so that every derived field is the
 * reference's own derivation, not a restatement.  vvcb200_emit_recon_ctu() walks a CTU exactly as ff_vvc_reconstruct()
 * does (vvc_intra.c:498-527 -> reconstruct 480-496 -> predict_intra 245-275 + itransform 432-478) and writes records
 * where the reference does pixel work.  In this repository the including translation unit is oracle/ref/api_recon.c,
 * which already pulls vvc_intra.c in for the parity harness.
 */
#ifndef VVCB200_EMIT_RECON_H
#define VVCB200_EMIT_RECON_H

#include "vvcb200_emit.h"

/* sh_lmcs_used_flag && ph_chroma_residual_scale_flag: every chroma residual waits for the luma-dependent factor */
static int emit_chroma_scaling(const VVCLocalContext *lc)
{
    return lc->sc->sh.r->sh_lmcs_used_flag && lc->fc->ps.ph.r->ph_chroma_residual_scale_flag;
}

/* ---- transform blocks: what itransform() (vvc_intra.c:432-478) does for one coded TB ------------------------ */
static void emit_tb(VVCB200EmitBatch *b, VVCLocalContext *lc, const TransformUnit *tu, TransformBlock *tb)
{
    const VVCFrameContext *fc     = lc->fc;
    const VVCSPS *sps             = fc->ps.sps;
    const H266RawSliceHeader *rsh = lc->sc->sh.r;
    const CodingUnit *cu          = lc->cu;
    const int c_idx = tb->c_idx, hs = sps->hshift[c_idx], vs = sps->vshift[c_idx];
    const int w = tb->tb_width;
    VVCB200TransformBlock *t = vvcb200_emit_new_tb(b);
    int flags = 0, nzw, nzh;
    uint32_t off;
    int16_t *lv;

    if (!t)
        return;
    t->x0 = tb->x0 >> hs;
    t->y0 = tb->y0 >> vs;
    t->c_idx = c_idx;
    t->log2_w = tb->log2_tb_width;
    t->log2_h = tb->log2_tb_height;
    t->nzw = nzw = tb->max_scan_x + 1;
    t->nzh = nzh = tb->max_scan_y + 1;
    derive_qp(lc, tu, tb);                                   /* vvc_intra.c:277-310 */
    t->qp = tb->qp;
    if (tb->ts) {
        flags |= VVCB200_TB_TS;
    } else {
        enum TxType trh, trv;
        derive_transform_type(fc, lc, tb, &trh, &trv);      /* vvc_intra.c:130-164 */
        t->trh = trh;
        t->trv = trv;
        if (cu->apply_lfnst_flag[c_idx]) {
            /* ilfnst_transform (vvc_intra.c:65-127) + ff_vvc_inv_lfnst_1d (vvc_itx_1d.c:708-721): set and transposition
             * follow from the wide-angle mapped intra mode */
            const int m = derive_ilfnst_pred_mode_intra(lc, tb);
            const int set = m < 0 ? 1 : ff_vvc_lfnst_tr_set_index[m];
            t->lfnst = cu->lfnst_idx | (set << 2) | ((m > 34) << 4);
        }
    }
    if (cu->bdpcm_flag[c_idx]) {                             /* transform_bdpcm, vvc_intra.c:419-430 */
        const IntraPredMode mode = c_idx ? cu->intra_pred_mode_c : cu->intra_pred_mode_y;
        flags |= VVCB200_TB_BDPCM | (mode == INTRA_VERT ? VVCB200_TB_BDPCM_VERT : 0);
    }
    if (rsh->sh_dep_quant_used_flag)
        flags |= VVCB200_TB_DEP_QUANT;
    if (rsh->sh_explicit_scaling_list_used_flag &&
        !(sps->r->sps_scaling_matrix_for_lfnst_disabled_flag && cu->apply_lfnst_flag[c_idx]))
        flags |= VVCB200_TB_SCALING;                         /* derive_scale_m, vvc_intra.c:367-369 (TS is the kernel's rule) */
    if (cu->pred_mode != MODE_INTRA)
        flags |= VVCB200_TB_INTER;
    if (tu->joint_cbcr_residual_flag && c_idx) {             /* add_residual_for_joint_coding_chroma, vvc_intra.c:166-186 */
        flags |= VVCB200_TB_JOINT;
        if (tu->coded_flag[1] ^ tu->coded_flag[2])
            flags |= VVCB200_TB_JOINT_SHIFT;
    }
    /* the residual of intra and CIIP units meets its prediction in the intra stage (CIIP chroma only where the reference
     * blends it: w_c > 2, vvc_inter.c:590); so does every chroma residual of a picture with luma-dependent chroma scaling */
    if (cu->pred_mode == MODE_INTRA || (cu->ciip_flag && (!c_idx || (cu->cb_width >> hs) > 2)) ||
        (c_idx && emit_chroma_scaling(lc)))
        flags |= VVCB200_TB_TO_RESIDUAL;
    t->flags = flags;
    lv = vvcb200_emit_levels(b, nzw * nzh, &off);
    if (!lv)
        return;
    t->coeff_off = off;
    for (int y = 0; y < nzh; y++)
        for (int x = 0; x < nzw; x++)
            lv[y * nzw + x] = (int16_t)tb->coeffs[y * w + x];
}

/* ---- intra records: the derivations at the head of intra_pred() (vvc_intra_template.c:595-632) ------------- */
#define EMIT_UNBOUNDED 0x7fff   /* "as many as there are": the device takes min(needed, available) */

/* Geometry in luma samples, as the reference passes it; lc->na must be current (ff_vvc_set_neighbour_available). */
static VVCB200IntraBlock *emit_pred_record(VVCB200EmitBatch *b, const VVCLocalContext *lc,
                                           const int x0, const int y0, const int width, const int height, const int c_idx)
{
    const VVCFrameContext *fc = lc->fc;
    const VVCSPS *sps         = fc->ps.sps;
    const CodingUnit *cu      = lc->cu;
    const int hs = sps->hshift[c_idx], vs = sps->vshift[c_idx];
    const int x = x0 >> hs, y = y0 >> vs, w = width >> hs, h = height >> vs;
    const int min_cb_width = fc->ps.pps->min_cb_width;          /* SAMPLE_CTB reads it */
    const int x_cb = x0 >> sps->min_cb_log2_size_y, y_cb = y0 >> sps->min_cb_log2_size_y;
    const int pred_mode = c_idx ? cu->intra_pred_mode_c : cu->intra_pred_mode_y;
    VVCB200IntraBlock *ib = vvcb200_emit_new_iblock(b);

    if (!ib)
        return NULL;
    ib->x0 = x;
    ib->y0 = y;
    ib->w = w;
    ib->h = h;
    ib->c_idx = c_idx;
    ib->cb_w = cu->cb_width;
    ib->cb_h = cu->cb_height;
    ib->up_left = lc->na.cand_up_left;
    ib->avail_top  = ff_vvc_get_top_available(lc, x, y, EMIT_UNBOUNDED, c_idx);    /* vvc_intra.c:591-620 */
    ib->avail_left = ff_vvc_get_left_available(lc, x, y, EMIT_UNBOUNDED, c_idx);   /* vvc_intra.c:622-648 */
    if (is_cclm(pred_mode) && c_idx) {
        ib->mode = pred_mode;                                   /* 81 LT, 82 L, 83 T */
        return ib;
    }
    ib->mode = ff_vvc_wide_angle_mode_mapping(cu, w, h, c_idx, pred_mode);
    ib->ref_idx = c_idx ? 0 : cu->intra_luma_ref_idx;
    if (SAMPLE_CTB(fc->tab.imf, x_cb, y_cb) && (!c_idx || cu->mip_chroma_direct_flag)) {
        ib->flags |= VVCB200_IB_MIP | (SAMPLE_CTB(fc->tab.imtf, x_cb, y_cb) ? VVCB200_IB_MIP_TRANSPOSED : 0);
        ib->mip_mode = SAMPLE_CTB(fc->tab.imm, x_cb, y_cb);
    }
    if (cu->bdpcm_flag[c_idx])
        ib->flags |= VVCB200_IB_BDPCM;
    return ib;
}

/* The VPDU lmcs_derive_chroma_scale (vvc_intra_template.c:390-428) would look at for this coding unit, and whether the
 * luma column / row next to it counts — asked of the reference's own availability functions at this point of the walk. */
static void emit_set_vpdu(VVCB200IntraBlock *ib, const VVCLocalContext *lc)
{
    const VVCSPS *sps = lc->fc->ps.sps;
    const int size_y = FFMIN(sps->ctb_size_y, 64);
    const int x = lc->cu->x0 & ~(size_y - 1), y = lc->cu->y0 & ~(size_y - 1);

    ib->vpdu_x = x / size_y;
    ib->vpdu_y = y / size_y;
    ib->vpdu_avail = (ff_vvc_get_left_available(lc, x, y, 1, 0) ? 1 : 0) | (ff_vvc_get_top_available(lc, x, y, 1, 0) ? 2 : 0);
    ib->flags |= VVCB200_IB_LMCS_SCALE;
}

/* Which chroma planes of this TU end up with a residual: their own coded TB, or the other plane's joint CbCr TB. */
static int emit_chroma_residual_planes(const TransformUnit *tu)
{
    int planes = 0;
    for (int i = 0; i < tu->nb_tbs; i++) {
        const TransformBlock *tb = &tu->tbs[i];
        if (tb->c_idx && tb->has_coeffs)
            planes |= tu->joint_cbcr_residual_flag ? 3 : 1 << (tb->c_idx - 1);
    }
    return planes;
}

static int emit_luma_coded(const TransformUnit *tu)
{
    for (int i = 0; i < tu->nb_tbs; i++)
        if (!tu->tbs[i].c_idx && tu->tbs[i].has_coeffs)
            return 1;
    return 0;
}

/* predict_intra() (vvc_intra.c:245-275) for one transform unit and channel type */
static void emit_predict_intra(VVCB200EmitBatch *b, VVCLocalContext *lc, const TransformUnit *tu, const int idx,
                               const int target_ch_type)
{
    const CodingUnit *cu        = lc->cu;
    const VVCTreeType tree_type = cu->tree_type;
    const int scaling           = emit_chroma_scaling(lc);
    int x0, y0, w, h;

    if (cu->pred_mode != MODE_INTRA) {
        /* the prediction is already there (stage INTER, or the CIIP records); what is left for this stage is a chroma
         * residual that had to wait for the luma-dependent factor */
        const int hs = lc->fc->ps.sps->hshift[1], vs = lc->fc->ps.sps->vshift[1];
        const int ciip_chroma = cu->ciip_flag && (cu->cb_width >> hs) > 2;
        if (target_ch_type && scaling && !ciip_chroma) {
            const int planes = emit_chroma_residual_planes(tu);
            for (int c = 1; c <= 2; c++) {
                VVCB200IntraBlock *ib;
                if (!(planes & (1 << (c - 1))) || !(ib = vvcb200_emit_new_iblock(b)))
                    continue;
                ib->x0 = tu->x0 >> hs;
                ib->y0 = tu->y0 >> vs;
                ib->w = tu->width >> hs;
                ib->h = tu->height >> vs;
                ib->c_idx = c;
                ib->cb_w = cu->cb_width;
                ib->cb_h = cu->cb_height;
                ib->flags = VVCB200_IB_ADD_ONLY | VVCB200_IB_RESIDUAL;
                emit_set_vpdu(ib, lc);
            }
        }
        add_reconstructed_area(lc, target_ch_type, tu->x0, tu->y0, tu->width, tu->height);
        return;
    }
    if (!target_ch_type && tree_type != DUAL_TREE_CHROMA) {
        if (get_luma_predict_unit(cu, tu, idx, &x0, &y0, &w, &h)) {
            ff_vvc_set_neighbour_available(lc, x0, y0, w, h);
            if (cu->isp_split_type == ISP_NO_SPLIT) {
                VVCB200IntraBlock *ib = emit_pred_record(b, lc, x0, y0, w, h, 0);
                if (ib && emit_luma_coded(tu))
                    ib->flags |= VVCB200_IB_RESIDUAL;
            } else if (!idx) {
                /* one record for the whole coding block: the device walks the partitions (their prediction unit is
                 * w x h, the first one's), and needs to know which of them carry a residual */
                VVCB200IntraBlock *ib = emit_pred_record(b, lc, x0, y0, w, h, 0);
                const TransformUnit *t = cu->tus.head;
                if (ib) {
                    ib->flags |= VVCB200_IB_ISP | VVCB200_IB_RESIDUAL;
                    for (int k = 0; t; k++, t = t->next)
                        ib->isp_coded |= emit_luma_coded(t) << k;
                }
            }
            add_reconstructed_area(lc, 0, x0, y0, w, h);
        }
    }
    if (target_ch_type && tree_type != DUAL_TREE_LUMA) {
        if (get_chroma_predict_unit(cu, tu, idx, &x0, &y0, &w, &h)) {
            const int planes = emit_chroma_residual_planes(tu);
            ff_vvc_set_neighbour_available(lc, x0, y0, w, h);
            if (is_cclm(cu->intra_pred_mode_c)) {
                /* intra_cclm_pred predicts Cb and Cr in one go from the same luma: one record, c_idx 1 */
                VVCB200IntraBlock *ib = emit_pred_record(b, lc, x0, y0, w, h, 1);
                if (ib && planes) {
                    ib->flags |= VVCB200_IB_RESIDUAL;
                    ib->res_planes = planes;
                    if (scaling)
                        emit_set_vpdu(ib, lc);
                }
            } else {
                for (int c = 1; c <= 2; c++) {
                    VVCB200IntraBlock *ib = emit_pred_record(b, lc, x0, y0, w, h, c);
                    if (ib && (planes & (1 << (c - 1)))) {
                        ib->flags |= VVCB200_IB_RESIDUAL;
                        if (scaling)
                            emit_set_vpdu(ib, lc);
                    }
                }
            }
            add_reconstructed_area(lc, 1, x0, y0, w, h);
        }
    }
}

/* ciip_derive_intra_weight (vvc_inter.c:523-543; static in another file, so restated): 1 + the number of intra
 * neighbours among the above-right-most and the left-bottom-most one */
static int emit_ciip_weight(const VVCLocalContext *lc)
{
    const VVCFrameContext *fc = lc->fc;
    const CodingUnit *cu      = lc->cu;
    const int log2_ctb        = fc->ps.sps->ctb_log2_size_y;
    const int min_pu_width    = fc->ps.pps->min_pu_width;
    const int x0 = cu->x0, y0 = cu->y0;
    const int avail_l = lc->ctb_left_flag || av_mod_uintp2(x0, log2_ctb);
    const int avail_u = lc->ctb_up_flag || av_mod_uintp2(y0, log2_ctb);
    int w = 1;

    if (avail_u && fc->tab.mvf[((y0 - 1) >> MIN_PU_LOG2) * min_pu_width + ((x0 - 1 + cu->cb_width) >> MIN_PU_LOG2)].pred_flag == PF_INTRA)
        w++;
    if (avail_l && fc->tab.mvf[((y0 - 1 + cu->cb_height) >> MIN_PU_LOG2) * min_pu_width + ((x0 - 1) >> MIN_PU_LOG2)].pred_flag == PF_INTRA)
        w++;
    return w;
}

/* ff_vvc_predict_ciip() (vvc_inter.c:915-921 -> pred_regular_blk 782-811 with the intra halves of pred_regular_luma
 * 570-576 / pred_regular_chroma 631-637): a planar prediction of the coding block per plane, blended by the device with
 * the inter prediction stage INTER left there.  The residual (if any) is added by the same record. */
static void emit_ciip(VVCB200EmitBatch *b, VVCLocalContext *lc)
{
    const CodingUnit *cu = lc->cu;
    const TransformUnit *tu = cu->tus.head;
    const int hs = lc->fc->ps.sps->hshift[1];
    const int weight = emit_ciip_weight(lc);
    const int coded = cu->coded_flag && tu;
    const int planes = coded ? emit_chroma_residual_planes(tu) : 0;

    if (coded && tu->next) {            /* one record adds the residual over the whole block: it must be one TU */
        b->error = VVCB200_ERR_UNSUP;
        return;
    }
    ff_vvc_set_neighbour_available(lc, cu->x0, cu->y0, cu->cb_width, cu->cb_height);
    for (int c = 0; c < 3; c++) {
        VVCB200IntraBlock *ib;
        if (c && (!lc->fc->ps.sps->r->sps_chroma_format_idc || (cu->cb_width >> hs) <= 2))
            break;                      /* pred_regular_chroma: do_ciip needs w_c > 2 */
        ib = emit_pred_record(b, lc, cu->x0, cu->y0, cu->cb_width, cu->cb_height, c);
        if (!ib)
            return;
        ib->flags |= VVCB200_IB_CIIP;
        ib->ciip_weight = weight;
        if (c ? planes & (1 << (c - 1)) : coded && emit_luma_coded(tu)) {
            ib->flags |= VVCB200_IB_RESIDUAL;
            if (c && emit_chroma_scaling(lc))
                emit_set_vpdu(ib, lc);
        }
    }
}

/* reconstruct() (vvc_intra.c:480-496) */
static void emit_reconstruct(VVCB200EmitBatch *b, VVCLocalContext *lc)
{
    CodingUnit *cu  = lc->cu;
    const int start = cu->tree_type == DUAL_TREE_CHROMA;
    const int end   = lc->fc->ps.sps->r->sps_chroma_format_idc && (cu->tree_type != DUAL_TREE_LUMA);

    for (int ch_type = start; ch_type <= end; ch_type++) {
        TransformUnit *tu = cu->tus.head;
        for (int i = 0; tu; i++) {
            emit_predict_intra(b, lc, tu, i, ch_type);
            for (int k = 0; k < tu->nb_tbs; k++) {                 /* itransform(): the coded TBs of this channel type */
                TransformBlock *tb = &tu->tbs[k];
                if ((tb->c_idx > 0) == ch_type && tb->has_coeffs)
                    emit_tb(b, lc, tu, tb);
            }
            tu = tu->next;
        }
    }
}

int vvcb200_emit_recon_ctu(VVCB200EmitBatch *b, VVCLocalContext *lc, const int rs, const int rx, const int ry)
{
    const VVCFrameContext *fc = lc->fc;
    const VVCSPS *sps         = fc->ps.sps;
    CodingUnit *cu            = fc->tab.ctus[rs].cus;

    lc->num_ras[0] = lc->num_ras[1] = 0;
    ff_vvc_decode_neighbour(lc, rx << sps->ctb_log2_size_y, ry << sps->ctb_log2_size_y, rx, ry, rs);
    vvcb200_emit_ctu_mark(b, rs, 0);
    while (cu) {
        lc->cu = cu;
        if (cu->ciip_flag)
            emit_ciip(b, lc);
        if (cu->coded_flag) {
            emit_reconstruct(b, lc);
        } else {
            add_reconstructed_area(lc, LUMA, cu->x0, cu->y0, cu->cb_width, cu->cb_height);
            add_reconstructed_area(lc, CHROMA, cu->x0, cu->y0, cu->cb_width, cu->cb_height);
        }
        cu = cu->next;
    }
    vvcb200_emit_ctu_mark(b, rs, 1);
    return b->error ? -1 : 0;
}

#endif /* VVCB200_EMIT_RECON_H */
