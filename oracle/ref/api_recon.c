/*
 * oracle/ref/api_recon.c — TEST INFRASTRUCTURE.
 *
 * Drives the reference's residual path exactly as the decoder does: the translation unit #includes
 * the reference's vvc_intra.c (from /root/reference, never copied) so that its file-static
 * itransform()/dequant()/ilfnst_transform() (vvc_intra.c:65-127, 277-478) are reachable, builds the
 * minimal VVCFrameContext/VVCLocalContext state those functions dereference
 * from the same flat descriptors the CUDA engine consumes, and calls them.
 */
#include "libavcodec/vvc/vvc_intra.c"
#include "vvcb200_emit_recon.h"     /* the product's RECON emitter: it belongs at the end of vvc_intra.c */
#include "api_common.h"

#include <stdlib.h>
#include <pthread.h>
#include <stdatomic.h>

typedef struct ReconEnv {
    VVCFrameContext     fc;
    VVCLocalContext     lc;
    SliceContext        sc;
    AVFrame             frame;
    VVCSPS              sps;
    VVCPPS              pps;
    H266RawSPS          rsps;
    H266RawPPS          rpps;
    H266RawPictureHeader rph;
    H266RawSliceHeader  rsh;
    VVCScalingList      sl;
    CodingUnit          cu;
    TransformUnit       tu;
    /* CIIP: ff_vvc_predict_ciip() runs the inter path from inside RECON */
    RefPicList          rpl[2];
    VVCFrame            cur;
    VVCFrame            refs[16];
    AVFrame             ref_frames[16];
} ReconEnv;

static ReconEnv *env_new(const VVCRefPlanes *pl, const VVCB200ResidualParams *p)
{
    ReconEnv *e = calloc(1, sizeof(*e));
    if (!e)
        return NULL;
    e->fc.frame   = &e->frame;
    e->fc.ps.sps  = &e->sps;
    e->fc.ps.pps  = &e->pps;
    e->fc.ps.ph.r = &e->rph;
    e->fc.ps.sl   = &e->sl;
    e->sps.r      = &e->rsps;
    e->pps.r      = &e->rpps;
    e->sc.sh.r    = &e->rsh;
    e->lc.fc      = &e->fc;
    e->lc.sc      = &e->sc;
    e->lc.cu      = &e->cu;
    for (int c = 0; c < 3; c++) {
        e->frame.data[c]     = pl->plane[c];
        e->frame.linesize[c] = pl->stride[c];
    }
    e->sps.width                = pl->width[0];
    e->sps.height               = pl->height[0];
    e->sps.bit_depth            = p->bit_depth;
    e->sps.pixel_shift          = 1;
    e->sps.qp_bd_offset         = 6 * (p->bit_depth - 8);
    e->sps.log2_transform_range = p->log2_transform_range;
    e->sps.min_cb_log2_size_y   = 2;
    e->sps.hshift[1] = e->sps.hshift[2] = 1;
    e->sps.vshift[1] = e->sps.vshift[2] = 1;
    e->sps.ctb_log2_size_y      = 7;
    e->sps.ctb_size_y           = 128;
    e->pps.min_cb_width         = (pl->width[0] + 3) >> 2;
    e->rsps.sps_chroma_format_idc = 1;
    e->rsps.sps_mts_enabled_flag  = 1;
    e->rsps.sps_explicit_mts_intra_enabled_flag = 1;
    e->rph.ph_joint_cbcr_sign_flag = p->joint_c_sign < 0;
    memcpy(e->sl.scaling_matrix_rec, p->scaling_matrix_rec, sizeof(e->sl.scaling_matrix_rec));
    memcpy(e->sl.scaling_matrix_dc_rec, p->scaling_matrix_dc_rec, sizeof(e->sl.scaling_matrix_dc_rec));
    ff_vvc_dsp_init(&e->fc.vvcdsp, p->bit_depth);
    return e;
}

/*
 * Residual stage for a batch of TBs on host planes, through the reference's own itransform().
 * `levels`/`tbs`/`params` are exactly what vvcb200_residual_batch() takes (include/vvc_b200.h).
 * `intra_mode[i]` (may be NULL) is the CU's raw intra prediction mode for TBs with LFNST; the
 * reference derives the LFNST set/transpose from it (vvc_intra.c:34-62, 693-714).
 * `residual_out` (may be NULL): int32 w*h residual of each TB appended back-to-back (what add_residual saw).
 * Returns 0, or -1 on a descriptor the reference cannot express.
 */
VVCREF_EXPORT int vvcref_residual_batch(const VVCRefPlanes *pl, const int16_t *levels,
                                        const VVCB200TransformBlock *tbs, int n,
                                        const VVCB200ResidualParams *params, const int8_t *intra_mode,
                                        int32_t *residual_out)
{
    ReconEnv *e = env_new(pl, params);
    int *coeffs = malloc(sizeof(int) * 64 * 64);
    uint8_t *imf;
    if (!e || !coeffs)
        return -1;
    /* derive_ilfnst_pred_mode_intra() peeks at fc->tab.imf / ipm / cpm[0]; all-zero tables = "no MIP, MODE_INTER==0". */
    imf = calloc((size_t)e->pps.min_cb_width * ((pl->height[0] + 3) >> 2) + 64, 1);
    e->fc.tab.imf = imf;
    e->fc.tab.ipm = imf;
    e->fc.tab.cpm[0] = imf;

    for (int i = 0; i < n; i++) {
        const VVCB200TransformBlock *d = &tbs[i];
        CodingUnit *cu     = &e->cu;
        TransformUnit *tu  = &e->tu;
        TransformBlock *tb = &tu->tbs[0];
        const int w = 1 << d->log2_w, h = 1 << d->log2_h;
        const int hs = d->c_idx ? 1 : 0, vs = d->c_idx ? 1 : 0;
        const int16_t *lv = levels + d->coeff_off;

        memset(cu, 0, sizeof(*cu));
        memset(tu, 0, sizeof(*tu));
        memset(coeffs, 0, sizeof(int) * w * h);
        for (int y = 0; y < d->nzh; y++)
            for (int x = 0; x < d->nzw; x++)
                coeffs[y * w + x] = lv[y * d->nzw + x];

        e->rsh.sh_dep_quant_used_flag              = !!(d->flags & VVCB200_TB_DEP_QUANT);
        e->rsh.sh_explicit_scaling_list_used_flag  = !!(d->flags & VVCB200_TB_SCALING);
        e->rsh.sh_lmcs_used_flag                   = 0;   /* chroma scaling is exercised by the picture-level harness */

        cu->pred_mode   = (d->flags & VVCB200_TB_INTER) ? MODE_INTER : MODE_INTRA;
        cu->tree_type   = SINGLE_TREE;
        cu->x0 = d->x0 << hs; cu->y0 = d->y0 << vs;
        cu->cb_width = w << hs; cu->cb_height = h << vs;
        /* derive_qp(): luma adds qp_bd_offset, chroma takes cu->qp[idx] as is (vvc_intra.c:285-295). */
        cu->qp[LUMA] = (int)d->qp - e->sps.qp_bd_offset;
        cu->qp[CB] = cu->qp[CR] = cu->qp[JCBCR] = d->qp;
        cu->bdpcm_flag[d->c_idx] = !!(d->flags & VVCB200_TB_BDPCM);
        if (d->flags & VVCB200_TB_BDPCM) {
            const IntraPredMode m = (d->flags & VVCB200_TB_BDPCM_VERT) ? INTRA_VERT : INTRA_HORZ;
            cu->intra_pred_mode_y = cu->intra_pred_mode_c = m;
        }
        if (d->lfnst) {
            cu->lfnst_idx = d->lfnst & 3;
            cu->apply_lfnst_flag[d->c_idx] = 1;
            cu->intra_pred_mode_y = cu->intra_pred_mode_c = intra_mode ? intra_mode[i] : 0;
        }
        /* derive_transform_type(): explicit MTS index reproduces any (trh, trv) pair the host derived. */
        if (d->c_idx == 0) {
            static const int8_t mts_of[3][3] = { /* [trh][trv] */
                { MTS_DCT2_DCT2, -1, -1 }, { -1, MTS_DST7_DST7, MTS_DCT8_DST7 }, { -1, MTS_DST7_DCT8, MTS_DCT8_DCT8 } };
            /* NB the enum reads MTS_<trv>_<trh>: mts_to_trh[2] = DCT8, mts_to_trv[2] = DST7 (vvc_intra.c:133-134) */
            if (mts_of[d->trh][d->trv] >= 0) {
                cu->mts_idx = mts_of[d->trh][d->trv];
            } else {
                /* mixed DCT2/DSTx pairs only arise from implicit MTS (vvc_intra.c:152-153); call the table directly below */
                cu->mts_idx = MTS_DCT2_DCT2;
            }
        }

        tu->x0 = cu->x0; tu->y0 = cu->y0; tu->width = cu->cb_width; tu->height = cu->cb_height;
        tu->nb_tbs = 1;
        tu->joint_cbcr_residual_flag = !!(d->flags & VVCB200_TB_JOINT);
        if (d->flags & VVCB200_TB_JOINT) {
            /* c_idx 1 carries the residual when tu_cb_coded_flag is set, c_idx 2 otherwise (vvc_intra.c:172-173). */
            tu->coded_flag[1] = d->c_idx == 1;
            tu->coded_flag[2] = (d->flags & VVCB200_TB_JOINT_SHIFT) ? !tu->coded_flag[1] : tu->coded_flag[1];
            if (d->c_idx == 2 && !(d->flags & VVCB200_TB_JOINT_SHIFT))
                return -1;
        } else {
            tu->coded_flag[d->c_idx] = 1;
        }
        tb->has_coeffs = 1;
        tb->c_idx = d->c_idx;
        tb->ts = !!(d->flags & VVCB200_TB_TS);
        tb->x0 = cu->x0; tb->y0 = cu->y0;
        tb->tb_width = w; tb->tb_height = h;
        tb->log2_tb_width = d->log2_w; tb->log2_tb_height = d->log2_h;
        tb->min_scan_x = tb->min_scan_y = 0;
        tb->max_scan_x = d->nzw - 1; tb->max_scan_y = d->nzh - 1;
        tb->coeffs = coeffs;

        if (d->c_idx == 0 && !tb->ts && !d->lfnst &&
            ((d->trh == DCT2) != (d->trv == DCT2))) {
            /* implicit-MTS style mixed pair: same steps as itransform() with the pair forced */
            uint8_t *dst = &e->frame.data[0][d->y0 * e->frame.linesize[0] + (d->x0 << 1)];
            if (cu->bdpcm_flag[0])
                return -1;
            dequant(&e->lc, tu, tb);
            e->fc.vvcdsp.itx.itx[d->trh][d->trv][d->log2_w][d->log2_h](coeffs, tb->max_scan_x + 1, tb->max_scan_y + 1,
                e->sps.log2_transform_range, e->sps.bit_depth);
            e->fc.vvcdsp.itx.add_residual(dst, coeffs, w, h, e->frame.linesize[0]);
        } else {
            itransform(&e->lc, tu, 0, d->c_idx > 0);
        }
        if (residual_out) {
            memcpy(residual_out, coeffs, sizeof(int) * w * h);
            residual_out += w * h;
        }
    }
    free(imf);
    free(coeffs);
    free(e);
    return 0;
}

/*
 * Stage RECON of a whole picture through ff_vvc_reconstruct() (vvc_intra.c:498-527): for every CTU in raster order the
 * CU / TU lists are rebuilt from the flat arrays, the per-min-CB tables the intra code peeks at (imf / imtf / imm / ipm /
 * cpm) are filled, and the reference predicts every intra block from its reconstructed neighbours and adds every
 * residual, exactly as the decoder does after parsing.  Inter prediction must already be in the planes (stage INTER).
 */
/* CTU rows dealt round-robin to `threads` workers; a CTU starts when the row above is two CTUs ahead, the wavefront of
 * the reference's own scheduler (vvc_thread.c:156-184). */
/* When set, vvcref_recon_picture runs the product's emitter (ffmpeg_b200/emit/vvcb200_emit_recon.h) on every CTU in place
 * of ff_vvc_reconstruct(): the same reference structs in, engine records out, no pixel touched. */
static VVCB200EmitBatch *g_emit_recon;

typedef struct ReconWorker {
    ReconEnv *e;
    int first_row, step, ctb_w, ctb_h;
    atomic_int *row_done;
    pthread_t th;
} ReconWorker;

static void *recon_worker(void *arg)
{
    ReconWorker *w = arg;
    VVCLocalContext *lc = aligned_alloc(64, (sizeof(*lc) + 63) & ~(size_t)63);
    memset(lc, 0, sizeof(*lc));
    lc->fc = &w->e->fc;
    lc->sc = &w->e->sc;
    for (int ry = w->first_row; ry < w->ctb_h; ry += w->step)
        for (int rx = 0; rx < w->ctb_w; rx++) {
            if (ry > 0) {
                const int need = rx + 2 < w->ctb_w ? rx + 2 : w->ctb_w;
                while (atomic_load_explicit(&w->row_done[ry - 1], memory_order_acquire) < need)
                    ;
            }
            ff_vvc_reconstruct(lc, ry * w->ctb_w + rx, rx, ry);
            atomic_store_explicit(&w->row_done[ry], rx + 1, memory_order_release);
        }
    free(lc);
    return NULL;
}

VVCREF_EXPORT int vvcref_recon_picture(const VVCRefPlanes *pl, const VVCRefCU *cus, int n_cus, const VVCRefTU *tus,
                                       const VVCB200TransformBlock *tbs, const int16_t *levels,
                                       const VVCB200ResidualParams *params, const VVCB200IntraParams *ip, int threads,
                                       const VVCRefPlanes *refs, int n_refs, const VVCB200InterParams *inter,
                                       const uint16_t *lmcs_fwd_lut)
{
    ReconEnv *e = env_new(pl, params);
    const int ctb_log2 = ip->ctb_log2_size, ctb = 1 << ctb_log2;
    const int ctb_w = (pl->width[0] + ctb - 1) >> ctb_log2, ctb_h = (pl->height[0] + ctb - 1) >> ctb_log2;
    const int mcw = (pl->width[0] + 3) >> 2, mch = (pl->height[0] + 3) >> 2;
    CTU *ctus = calloc((size_t)ctb_w * ctb_h, sizeof(*ctus));
    CodingUnit *cu_store = calloc(n_cus, sizeof(*cu_store));
    uint8_t *tabs = calloc((size_t)mcw * mch, 5);
    uint16_t *col_bd = calloc(ctb_w + 2, sizeof(*col_bd)), *row_bd = calloc(ctb_h + 2, sizeof(*row_bd));
    int16_t *slice_idx = calloc((size_t)ctb_w * ctb_h, sizeof(*slice_idx));
    VVCLocalContext *lc = aligned_alloc(64, (sizeof(*lc) + 63) & ~(size_t)63);
    int n_tu_total = 0, n_tb_total = 0;
    TransformUnit *tu_store;
    int *coeff_store;
    size_t coeff_total = 0;

    for (int i = 0; i < n_cus; i++)
        n_tu_total += cus[i].n_tu;
    tu_store = calloc(n_tu_total ? n_tu_total : 1, sizeof(*tu_store));
    for (int i = 0; i < n_tu_total; i++)
        for (int c = 0; c < 3; c++)
            if (tus[i].tb[c] >= 0) {
                const VVCB200TransformBlock *d = &tbs[tus[i].tb[c]];
                coeff_total += (size_t)1 << (d->log2_w + d->log2_h);
                n_tb_total++;
            }
    coeff_store = calloc(coeff_total ? coeff_total : 1, sizeof(int));

    memset(lc, 0, sizeof(*lc));
    lc->fc = &e->fc;
    lc->sc = &e->sc;
    e->sps.ctb_log2_size_y = ctb_log2;
    e->sps.ctb_size_y      = ctb;
    e->pps.width = pl->width[0];
    e->pps.height = pl->height[0];
    e->pps.ctb_width = ctb_w;
    e->pps.ctb_height = ctb_h;
    e->pps.min_cb_width = e->pps.min_pu_width = e->pps.min_tu_width = mcw;
    e->pps.ctb_to_col_bd = col_bd;
    e->pps.ctb_to_row_bd = row_bd;
    e->rsps.sps_chroma_vertical_collocated_flag = ip->chroma_vertical_collocated;
    if (ip->lmcs_chroma_scale) {                            /* luma-dependent chroma residual scaling */
        e->rsh.sh_lmcs_used_flag = 1;
        e->rph.ph_chroma_residual_scale_flag = 1;
        e->fc.ps.lmcs.min_bin_idx = ip->lmcs_min_bin_idx;
        e->fc.ps.lmcs.max_bin_idx = ip->lmcs_max_bin_idx;
        memcpy(e->fc.ps.lmcs.pivot, ip->lmcs_pivot, sizeof(e->fc.ps.lmcs.pivot));
        memcpy(e->fc.ps.lmcs.chroma_scale_coeff, ip->lmcs_chroma_scale_coeff, sizeof(e->fc.ps.lmcs.chroma_scale_coeff));
    }
    e->fc.tab.slice_idx = slice_idx;
    e->fc.tab.ctus = ctus;
    e->fc.tab.imf    = tabs;
    e->fc.tab.imtf   = tabs + (size_t)mcw * mch;
    e->fc.tab.imm    = tabs + (size_t)mcw * mch * 2;
    e->fc.tab.ipm    = tabs + (size_t)mcw * mch * 3;
    e->fc.tab.cpm[0] = tabs + (size_t)mcw * mch * 4;
    /* motion field: PF_INTRA on intra units (read by ciip_derive_intra_weight), the merge motion on CIIP units */
    e->fc.tab.mvf = calloc((size_t)mcw * mch, sizeof(MvField));
    e->pps.min_pu_height = mch;
    if (refs && inter && n_refs <= 16) {
        PredWeightTable *w = &e->sc.sh.pwt;
        e->fc.ref = &e->cur;
        e->sc.rpl = e->rpl;
        e->rsh.sh_slice_type = VVC_SLICE_TYPE_B;
        ff_videodsp_init(&e->fc.vdsp, params->bit_depth);
        for (int i = 0; i < n_refs; i++) {
            for (int c = 0; c < 3; c++) {
                e->ref_frames[i].data[c]     = refs[i].plane[c];
                e->ref_frames[i].linesize[c] = refs[i].stride[c];
            }
            e->refs[i].frame = &e->ref_frames[i];
        }
        for (int l = 0; l < 2; l++)
            for (int i = 0; i < 15; i++)
                e->rpl[l].ref[i] = inter->ref_slot[l][i] < n_refs ? &e->refs[inter->ref_slot[l][i]] : NULL;
        w->log2_denom[0] = inter->log2_denom[0];
        w->log2_denom[1] = inter->log2_denom[1];
        for (int l = 0; l < 2; l++)
            for (int c = 0; c < 3; c++)
                for (int i = 0; i < MAX_WEIGHTS; i++) {
                    w->weight[l][c][i] = inter->weight[l][c][i];
                    w->offset[l][c][i] = inter->offset[l][c][i];
                }
        if (lmcs_fwd_lut)
            memcpy(e->fc.ps.lmcs.fwd_lut, lmcs_fwd_lut, sizeof(uint16_t) << params->bit_depth);
    }

    {
        int *cp = coeff_store;
        CodingUnit **tail = NULL;
        int last_ctu = -1, tu_i = 0;
        for (int i = 0; i < n_cus; i++) {
            const VVCRefCU *d = &cus[i];
            CodingUnit *cu = &cu_store[i];
            TransformUnit **tu_tail = &cu->tus.head;
            if (d->ctu != last_ctu) {
                tail = &ctus[d->ctu].cus;
                last_ctu = d->ctu;
            }
            *tail = cu;
            tail = &cu->next;
            cu->tree_type = d->tree_type == 1 ? DUAL_TREE_LUMA : d->tree_type == 2 ? DUAL_TREE_CHROMA : SINGLE_TREE;
            cu->x0 = d->x0; cu->y0 = d->y0; cu->cb_width = d->w; cu->cb_height = d->h;
            cu->pred_mode = d->pred_mode ? MODE_INTRA : MODE_INTER;
            cu->coded_flag = d->coded_flag;
            cu->intra_pred_mode_y = d->intra_pred_mode_y;
            cu->intra_pred_mode_c = d->intra_pred_mode_c;
            cu->intra_luma_ref_idx = d->intra_luma_ref_idx;
            cu->intra_mip_flag = d->mip_flag;
            cu->mip_chroma_direct_flag = d->mip_chroma_direct;
            cu->isp_split_type = d->isp_split_type;
            cu->num_intra_subpartitions = d->num_intra_subpartitions;
            cu->ciip_flag = d->ciip_flag;
            cu->mts_idx = d->mts_idx;
            cu->lfnst_idx = d->lfnst_idx;
            cu->apply_lfnst_flag[0] = d->lfnst_idx > 0;      /* single tree: LFNST on luma only */
            for (int c = 0; c < 3; c++)
                cu->bdpcm_flag[c] = d->bdpcm_flag[c];
            if (d->ciip_flag) {
                if (!e->fc.ref)
                    return -1;                                  /* CIIP units need the reference pictures */
                cu->intra_pred_mode_y = cu->intra_pred_mode_c = INTRA_PLANAR;
                cu->pu.mi.num_sb_x = cu->pu.mi.num_sb_y = 1;
                cu->pu.mi.hpel_if_idx = !!(d->inter_flags & VVCB200_PB_HPEL_IF);
                cu->pu.mi.pred_flag = d->pred_flag;
                /* slice-level switches, the same for every CIIP unit of the picture */
                e->rpps.pps_weighted_pred_flag = e->rpps.pps_weighted_bipred_flag = !!(d->inter_flags & VVCB200_PB_WP);
                e->rsh.sh_lmcs_used_flag |= !!(d->inter_flags & VVCB200_PB_LMCS);
            }
            for (int y = d->y0 >> 2; y < (d->y0 + d->h) >> 2; y++)
                for (int x = d->x0 >> 2; x < (d->x0 + d->w) >> 2; x++) {
                    const size_t o = (size_t)y * mcw + x;
                    MvField *mvf = &e->fc.tab.mvf[o];
                    if (d->tree_type == 2)
                        continue;                               /* the per-unit tables describe the luma tree */
                    mvf->pred_flag = d->pred_mode ? PF_INTRA : d->ciip_flag ? d->pred_flag : PF_L0;
                    mvf->ciip_flag = d->ciip_flag;
                    if (d->ciip_flag) {
                        for (int l = 0; l < 2; l++) {
                            mvf->mv[l].x = d->mv[l][0];
                            mvf->mv[l].y = d->mv[l][1];
                            mvf->ref_idx[l] = d->ref_idx[l];
                        }
                        mvf->hpel_if_idx = cu->pu.mi.hpel_if_idx;
                    }
                    e->fc.tab.imf[o]    = d->mip_flag;
                    e->fc.tab.imtf[o]   = d->mip_transposed;
                    e->fc.tab.imm[o]    = d->mip_mode;
                    e->fc.tab.ipm[o]    = d->intra_pred_mode_y;
                    e->fc.tab.cpm[0][o] = cu->pred_mode;
                }
            for (int k = 0; k < d->n_tu; k++, tu_i++) {
                const VVCRefTU *t = &tus[d->first_tu + k];
                TransformUnit *tu = &tu_store[tu_i];
                *tu_tail = tu;
                tu_tail = &tu->next;
                cu->tus.tail = tu;
                tu->x0 = t->x0; tu->y0 = t->y0; tu->width = t->w; tu->height = t->h;
                tu->joint_cbcr_residual_flag = t->joint;
                tu->nb_tbs = 3;
                for (int c = 0; c < 3; c++) {
                    TransformBlock *tb = &tu->tbs[c];
                    tu->coded_flag[c] = t->coded_flag[c];
                    tb->c_idx = c;
                    tb->x0 = t->x0; tb->y0 = t->y0;
                    tb->tb_width = t->w >> (c ? 1 : 0);
                    tb->tb_height = t->h >> (c ? 1 : 0);
                    if (c && d->isp_split_type) {           /* chroma of an ISP unit: the whole coding block */
                        tb->x0 = d->x0; tb->y0 = d->y0;
                        tb->tb_width = d->w >> 1; tb->tb_height = d->h >> 1;
                    }
                    tb->log2_tb_width = av_log2(tb->tb_width);
                    tb->log2_tb_height = av_log2(tb->tb_height);
                    if (t->tb[c] >= 0) {
                        const VVCB200TransformBlock *s = &tbs[t->tb[c]];
                        const int w = 1 << s->log2_w;
                        const int16_t *lv = levels + s->coeff_off;
                        tb->has_coeffs = 1;
                        tb->ts = !!(s->flags & VVCB200_TB_TS);
                        tb->max_scan_x = s->nzw - 1; tb->max_scan_y = s->nzh - 1;
                        tb->coeffs = cp;
                        for (int y = 0; y < s->nzh; y++)
                            for (int x = 0; x < s->nzw; x++)
                                cp[y * w + x] = lv[y * s->nzw + x];
                        cp += (size_t)1 << (s->log2_w + s->log2_h);
                        /* derive_qp(): luma adds qp_bd_offset, chroma takes cu->qp[idx] as is */
                        if (c == 0)
                            cu->qp[LUMA] = (int)s->qp - e->sps.qp_bd_offset;
                        else
                            cu->qp[c] = cu->qp[JCBCR] = s->qp;
                    }
                }
            }
        }
    }
    e->rsh.sh_dep_quant_used_flag = 0;
    e->rsh.sh_explicit_scaling_list_used_flag = 0;
    for (int i = 0; i < n_tu_total; i++)                    /* slice-level switches, carried by every TB of the picture */
        for (int c = 0; c < 3; c++)
            if (tus[i].tb[c] >= 0) {
                e->rsh.sh_dep_quant_used_flag |= !!(tbs[tus[i].tb[c]].flags & VVCB200_TB_DEP_QUANT);
                e->rsh.sh_explicit_scaling_list_used_flag |= !!(tbs[tus[i].tb[c]].flags & VVCB200_TB_SCALING);
            }
    if (g_emit_recon) {
        /* twice: the second pass, into arrays that already have their size (a decoder keeps its batch from picture to
         * picture, vvcb200_emit_reset), is the one that is timed */
        for (int pass = 0; pass < 2; pass++) {
            const int64_t t0 = vvcref_now_ns();
            vvcb200_emit_reset(g_emit_recon);
            for (int ry = 0; ry < ctb_h; ry++)
                for (int rx = 0; rx < ctb_w; rx++)
                    vvcb200_emit_recon_ctu(g_emit_recon, lc, ry * ctb_w + rx, rx, ry);
            vvcref_emit_ns = vvcref_now_ns() - t0;
        }
    } else if (threads <= 1) {
        for (int ry = 0; ry < ctb_h; ry++)
            for (int rx = 0; rx < ctb_w; rx++)
                ff_vvc_reconstruct(lc, ry * ctb_w + rx, rx, ry);
    } else {
        ReconWorker *ws = calloc(threads, sizeof(*ws));
        atomic_int *row_done = calloc(ctb_h, sizeof(*row_done));
        if (threads > ctb_h)
            threads = ctb_h;
        for (int t = 0; t < threads; t++) {
            ws[t] = (ReconWorker){ .e = e, .first_row = t, .step = threads, .ctb_w = ctb_w, .ctb_h = ctb_h, .row_done = row_done };
            pthread_create(&ws[t].th, NULL, recon_worker, &ws[t]);
        }
        for (int t = 0; t < threads; t++)
            pthread_join(ws[t].th, NULL);
        free(row_done);
        free(ws);
    }

    free(e->fc.tab.mvf);
    free(coeff_store); free(tu_store); free(lc); free(slice_idx); free(col_bd); free(row_bd); free(tabs);
    free(cu_store); free(ctus); free(e);
    (void)n_tb_total;
    return 0;
}

/*
 * f1 test entry: the CU / TU arrays are turned into the reference's CTU -> CodingUnit -> TransformUnit -> TransformBlock
 * lists and per-min-CB tables exactly as vvcref_recon_picture does, and the emitter writes the engine's RECON records
 * from THAT.  `in_tbs` / `in_levels` only feed the TransformBlock fields residual parsing would have left (scan box,
 * levels, ts, qp); everything else in the emitted records is derived from the reference structs.  Caller-allocated
 * outputs; returns 0, -1 (emitter error) or -2 (an output array is too small).
 */
VVCREF_EXPORT int vvcref_emit_recon_picture(int width, int height, const VVCRefCU *cus, int n_cus, const VVCRefTU *tus,
                                            const VVCB200TransformBlock *in_tbs, const int16_t *in_levels,
                                            const VVCB200ResidualParams *params, const VVCB200IntraParams *ip,
                                            int with_refs, const VVCB200InterParams *inter,
                                            VVCB200TransformBlock *tbs, int cap_tbs, int *n_tbs,
                                            int16_t *levels, int cap_levels, int *n_levels,
                                            VVCB200IntraBlock *iblocks, int cap_iblocks, int *n_iblocks,
                                            uint32_t *ctu_first, int cap_ctus, int *n_ctus)
{
    VVCB200EmitBatch batch = { 0 };
    VVCRefPlanes pl = { { NULL, NULL, NULL }, { 0, 0, 0 }, { width, width / 2, width / 2 }, { height, height / 2, height / 2 } };
    int ret;

    if (with_refs & 2) {
        /* the caller's arrays as the batch (VVCB200EmitBatch.fixed): what a decoder does with regions of a context arena */
        batch.fixed = 1;
        batch.tbs = tbs; batch.cap_tbs = cap_tbs;
        batch.levels = levels; batch.cap_levels = cap_levels;
        batch.iblocks = iblocks; batch.cap_iblocks = cap_iblocks;
        batch.ctu_first = ctu_first; batch.cap_ctus = cap_ctus;
    }
    g_emit_recon = &batch;
    /* CIIP units want a reference picture list to exist; the emitter never follows it */
    ret = vvcref_recon_picture(&pl, cus, n_cus, tus, in_tbs, in_levels, params, ip, 1, (with_refs & 1) ? &pl : NULL, 0, inter, NULL);
    g_emit_recon = NULL;
    if (!ret && batch.error)
        ret = batch.error == VVCB200_ERR_TOO_BIG ? -2 : -1;
    if (batch.fixed) {
        if (!ret) {
            *n_tbs = batch.n_tbs;
            *n_levels = (int)batch.n_levels;
            *n_iblocks = batch.n_iblocks;
            *n_ctus = batch.n_ctus;
        }
        return ret;
    }
    if (!ret && (batch.n_tbs > cap_tbs || batch.n_levels > (size_t)cap_levels || batch.n_iblocks > cap_iblocks ||
                 batch.n_ctus + 1 > cap_ctus))
        ret = -2;
    if (!ret) {
        memcpy(tbs, batch.tbs, (size_t)batch.n_tbs * sizeof(*tbs));
        memcpy(levels, batch.levels, batch.n_levels * sizeof(*levels));
        memcpy(iblocks, batch.iblocks, (size_t)batch.n_iblocks * sizeof(*iblocks));
        memcpy(ctu_first, batch.ctu_first, (size_t)(batch.n_ctus + 1) * sizeof(*ctu_first));
        *n_tbs = batch.n_tbs;
        *n_levels = (int)batch.n_levels;
        *n_iblocks = batch.n_iblocks;
        *n_ctus = batch.n_ctus;
    }
    vvcb200_emit_free(&batch);
    return ret;
}
