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
#include "api_common.h"

#include <stdlib.h>

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

        e->sps.hshift[1] = e->sps.hshift[2] = 1;
        e->sps.vshift[1] = e->sps.vshift[2] = 1;
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
