/*
 * oracle/port/port_residual.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's residual path for one picture's batch of transform blocks:
 *   transform_bdpcm        libavcodec/vvc/vvcdsp_template.c:76-95, vvc_intra.c:419-430
 *   dequant / scale_coeff  libavcodec/vvc/vvc_intra.c:277-417
 *   inverse LFNST          libavcodec/vvc/vvc_intra.c:65-127, vvc_itx_1d.c:708-721
 *   inverse DCT-II/DST-VII/DCT-VIII, 2-D and 1-D drivers   libavcodec/vvc/vvcdsp.c:67-138, vvc_itx_1d.c:88-706
 *   add_residual (+ joint CbCr)                             libavcodec/vvc/vvcdsp_template.c:32-74, vvc_intra.c:166-186
 * The 1-D transforms are written as plain matrix products over the non-zero inputs; this is
 * arithmetically identical (mod 2^32) to the reference's partial butterflies because those are exact
 * integer factorizations of the same matrices with the same zero-out guards.
 */
#include "port_common.h"

static const int8_t *tx_matrix(int type, int log2n)
{
    if (type == VVCB200_DCT2)
        return VVC_DCT2_MAT + VVC_DCT2_OFF[log2n - 1];
    if (type == VVCB200_DST7)
        return VVC_DST7_MAT + VVC_DST7_OFF[log2n - 2];
    return VVC_DCT8_MAT + VVC_DCT8_OFF[log2n - 2];
}

/* one 1-D inverse transform: out[n] = sum_{k<nz} in[k] * M[k][n]; in/out may alias (strided). */
static void inv_1d(int32_t *v, int stride, int type, int log2n, int nz)
{
    const int n = 1 << log2n;
    const int8_t *m = tx_matrix(type, log2n);
    uint32_t tmp[64];
    if (n == 1)
        return;
    for (int j = 0; j < n; j++) {
        uint32_t acc = 0;                       /* unsigned: wraps like the reference's int does in practice */
        for (int k = 0; k < nz; k++)
            acc += (uint32_t)v[k * stride] * (uint32_t)(int32_t)m[k * n + j];
        tmp[j] = acc;
    }
    for (int j = 0; j < n; j++)
        v[j * stride] = (int32_t)tmp[j];
}

/* itx.itx[trh][trv][log2w][log2h] (vvcdsp.c:94-138) on a w*h int32 block, in place. */
PORT_EXPORT void port_itx(int32_t *c, int trh, int trv, int log2w, int log2h, int nzw, int nzh,
                          int log2_range, int bd)
{
    const int w = 1 << log2w, h = 1 << log2h;
    const int dconly = trh == VVCB200_DCT2 && trv == VVCB200_DCT2;

    if (w > 1 && h > 1) {
        const int sh0 = 7, sh1 = 5 + log2_range - bd;
        if (w == h && nzw == 1 && nzh == 1 && dconly) {
            const int t  = (c[0] * 64 + (1 << (sh0 - 1))) >> sh0;
            const int dc = (t * 64 + (1 << (sh1 - 1))) >> sh1;
            for (int i = 0; i < w * h; i++)
                c[i] = dc;
            return;
        }
        for (int x = 0; x < nzw; x++)
            inv_1d(c + x, w, trv, log2h, nzh);
        for (int y = 0; y < h; y++) {
            for (int x = 0; x < nzw; x++)
                c[y * w + x] = port_clip_intp2((c[y * w + x] + (1 << (sh0 - 1))) >> sh0, log2_range);
            for (int x = nzw; x < w; x++)
                c[y * w + x] = 0;
        }
        for (int y = 0; y < h; y++)
            inv_1d(c + y * w, 1, trh, log2w, nzw);
        for (int i = 0; i < w * h; i++)
            c[i] = (c[i] + (1 << (sh1 - 1))) >> sh1;
    } else {
        const int sh = 6 + log2_range - bd;
        if (nzw == 1 && nzh == 1 && dconly) {
            const int dc = (c[0] * 64 + (1 << (sh - 1))) >> sh;
            for (int i = 0; i < w * h; i++)
                c[i] = dc;
            return;
        }
        if (w > 1)
            inv_1d(c, 1, trh, log2w, nzw);
        else
            inv_1d(c, 1, trv, log2h, nzh);
        for (int i = 0; i < w * h; i++)
            c[i] = (c[i] + (1 << (sh - 1))) >> sh;
    }
}

/* Table 38: scaling matrix id from (inter?, c_idx, max(log2w, log2h)) — vvc_intra.c:342-362 */
static int scaling_id(int inter, int c_idx, int log2max)
{
    static const uint8_t ids[2][3][6] = {
        { { 0, 2, 8, 14, 20, 26 }, { 0, 3, 9, 15, 21, 21 }, { 0, 4, 10, 16, 22, 22 } },
        { { 0, 5, 11, 17, 23, 27 }, { 0, 6, 12, 18, 24, 24 }, { 1, 7, 13, 19, 25, 25 } },
    };
    return ids[inter][c_idx][log2max - 1];
}

static void bdpcm(int32_t *c, int w, int h, int vertical, int log2_range)
{
    /* running sum down the columns (vertical) or along the rows, clipped each step */
    if (vertical) {
        for (int y = 1; y < h; y++)
            for (int x = 0; x < w; x++)
                c[y * w + x] = port_clip_intp2(c[y * w + x] + c[(y - 1) * w + x], log2_range);
    } else {
        for (int y = 0; y < h; y++)
            for (int x = 1; x < w; x++)
                c[y * w + x] = port_clip_intp2(c[y * w + x] + c[y * w + x - 1], log2_range);
    }
}

static void lfnst(int32_t *c, int w, int h, int code, int log2_range, int *nzw, int *nzh)
{
    const int idx = (code & 3) - 1, set = (code >> 2) & 3, transpose = (code >> 4) & 1;
    const int big = w >= 8 && h >= 8;
    const int n_out = big ? 48 : 16, size = big ? 8 : 4;
    const int nz = ((w == 8 && h == 8) || (w == 4 && h == 4)) ? 8 : 16;
    const int8_t *m = big ? VVC_LFNST_8X8 + (set * 2 + idx) * 16 * 48 : VVC_LFNST_4X4 + (set * 2 + idx) * 16 * 16;
    int32_t u[16], v[48];

    for (int i = 0; i < nz; i++) {
        const int p = VVC_DIAG_SCAN_4X4[i];
        u[i] = c[(p >> 2) * w + (p & 3)];
    }
    for (int j = 0; j < n_out; j++) {
        int t = 0;
        for (int i = 0; i < nz; i++)
            t += u[i] * m[i * n_out + j];
        v[j] = port_clip_intp2((t + 64) >> 7, log2_range);
    }
    /* v is the top-left size x size region in raster order, with the bottom-right 4x4 of the 8x8 case left out */
    for (int j = 0, y = 0; y < size; y++) {
        const int len = y < 4 ? size : 4;
        for (int x = 0; x < len; x++, j++) {
            if (transpose)
                c[x * w + y] = v[j];
            else
                c[y * w + x] = v[j];
        }
    }
    *nzw = *nzh = size;
}

/* Residual of one TB into res[w*h] (what add_residual receives).  Returns nothing; res fully written. */
static void tb_residual(int32_t *res, const VVCB200TransformBlock *d, const int16_t *levels,
                        const VVCB200ResidualParams *p)
{
    const int w = 1 << d->log2_w, h = 1 << d->log2_h;
    const int ts = d->flags & VVCB200_TB_TS;
    const int range = p->log2_transform_range;
    int nzw = d->nzw, nzh = d->nzh;
    const int16_t *lv = levels + d->coeff_off;

    memset(res, 0, sizeof(*res) * w * h);
    for (int y = 0; y < nzh; y++)
        for (int x = 0; x < nzw; x++)
            res[y * w + x] = lv[y * d->nzw + x];

    if (d->flags & VVCB200_TB_BDPCM) {
        bdpcm(res, w, h, !!(d->flags & VVCB200_TB_BDPCM_VERT), range);
        if (d->flags & VVCB200_TB_BDPCM_VERT) nzh = h; else nzw = w;
    }
    {   /* dequant */
        static const int level_scale[2][6] = { { 40, 45, 51, 57, 64, 72 }, { 57, 64, 72, 80, 90, 102 } };
        const int log_sum = d->log2_w + d->log2_h;
        const int rect = ts ? 0 : (log_sum & 1);
        const int dq = (d->flags & VVCB200_TB_DEP_QUANT) ? 1 : 0;
        const int bd_shift = ts ? 10 : p->bit_depth + rect + log_sum / 2 + 10 - range + dq;
        const int bd_off = (1 << bd_shift) >> 1;
        const int qp = d->qp + (ts ? 0 : dq);
        const int scale = level_scale[rect][qp % 6] << (qp / 6);
        const int use_sl = (d->flags & VVCB200_TB_SCALING) && !ts;
        const int id = scaling_id(!!(d->flags & VVCB200_TB_INTER), d->c_idx, d->log2_w > d->log2_h ? d->log2_w : d->log2_h);
        const int l2m = id < 2 ? 1 : id < 8 ? 2 : 3;
        for (int y = 0; y < nzh; y++) {
            for (int x = 0; x < nzw; x++) {
                int m = 16;
                if (use_sl) {
                    const int off = ((y << l2m) >> d->log2_h) << l2m;
                    m = p->scaling_matrix_rec[id][off + ((x << l2m) >> d->log2_w)];
                    if (id >= 14 && !x && !y)
                        m = p->scaling_matrix_dc_rec[id - 14];
                }
                if (res[y * w + x]) {
                    const int v = (int)((uint32_t)res[y * w + x] * (uint32_t)scale * (uint32_t)m + (uint32_t)bd_off);
                    res[y * w + x] = port_clip_intp2(v >> bd_shift, range);
                }
            }
        }
    }
    if (!ts) {
        if (d->lfnst)
            lfnst(res, w, h, d->lfnst, range, &nzw, &nzh);
        port_itx(res, d->trh, d->trv, d->log2_w, d->log2_h, nzw, nzh, range, p->bit_depth);
    }
}

static void add_block(const PortPlanes *pl, int c_idx, int x0, int y0, int w, int h, const int32_t *res,
                      int bd, int c_sign, int shift, int joint)
{
    for (int y = 0; y < h; y++) {
        uint16_t *dst = (uint16_t *)(pl->plane[c_idx] + (size_t)(y0 + y) * pl->stride[c_idx]) + x0;
        for (int x = 0; x < w; x++) {
            const int r = joint ? (res[y * w + x] * c_sign) >> shift : res[y * w + x];
            dst[x] = port_clip_pixel(dst[x] + r, bd);
        }
    }
}

/* LMCS chroma residual scaling, vvc_intra_template.c:431-448 */
static int lmcs_scale_one(int c, int scale, int bd)
{
    const int v = port_clip_intp2(c, bd);
    const int a = v < 0 ? -v : v;
    const int s = (a * scale + (1 << 10)) >> 11;
    return v < 0 ? -s : s;
}

/*
 * The whole residual stage on host planes; same arguments as vvcb200_residual_batch.
 * residual_out (optional): int32 residual blocks appended back to back (pre-LMCS-scaling, pre-joint).
 */
/* the TO_RESIDUAL variant of add_block: store the int16-clipped residual instead of adding it */
static void store_block(const PortPlanes *rp, int c_idx, int x0, int y0, int w, int h, const int32_t *res,
                        int c_sign, int shift, int joint)
{
    for (int y = 0; y < h; y++) {
        int16_t *dst = (int16_t *)(rp->plane[c_idx] + (size_t)(y0 + y) * rp->stride[c_idx]) + x0;
        for (int x = 0; x < w; x++) {
            const int r = joint ? (res[y * w + x] * c_sign) >> shift : res[y * w + x];
            dst[x] = port_clip_intp2(r, 15);
        }
    }
}

PORT_EXPORT int port_residual_batch(const PortPlanes *pl, const int16_t *levels, const VVCB200TransformBlock *tbs,
                                    int n, const VVCB200ResidualParams *p, int32_t *residual_out, const PortPlanes *rplanes)
{
    int32_t res[64 * 64], tmp[64 * 64];
    for (int i = 0; i < n; i++) {
        const VVCB200TransformBlock *d = &tbs[i];
        const int w = 1 << d->log2_w, h = 1 << d->log2_h;
        const int32_t *src = res;
        tb_residual(res, d, levels, p);
        if (residual_out) {
            memcpy(residual_out, res, sizeof(*res) * w * h);
            residual_out += w * h;
        }
        if (d->flags & VVCB200_TB_LMCS_SCALE) {
            for (int k = 0; k < w * h; k++)
                tmp[k] = lmcs_scale_one(res[k], d->lmcs_scale, p->bit_depth);
            src = tmp;
        }
        if ((d->flags & VVCB200_TB_TO_RESIDUAL) && rplanes) {
            store_block(rplanes, d->c_idx, d->x0, d->y0, w, h, src, 1, 0, 0);
            if ((d->flags & VVCB200_TB_JOINT) && d->c_idx)
                store_block(rplanes, 3 - d->c_idx, d->x0, d->y0, w, h, res, p->joint_c_sign,
                            !!(d->flags & VVCB200_TB_JOINT_SHIFT), 1);
            continue;
        }
        add_block(pl, d->c_idx, d->x0, d->y0, w, h, src, p->bit_depth, 1, 0, 0);
        if ((d->flags & VVCB200_TB_JOINT) && d->c_idx) {
            const int shift = !!(d->flags & VVCB200_TB_JOINT_SHIFT);
            if (d->flags & VVCB200_TB_LMCS_SCALE) {
                /* pred_residual_joint then scale then add (vvc_intra.c:180-183) */
                for (int k = 0; k < w * h; k++)
                    tmp[k] = lmcs_scale_one((res[k] * p->joint_c_sign) >> shift, d->lmcs_scale, p->bit_depth);
                add_block(pl, 3 - d->c_idx, d->x0, d->y0, w, h, tmp, p->bit_depth, 1, 0, 0);
            } else {
                add_block(pl, 3 - d->c_idx, d->x0, d->y0, w, h, res, p->bit_depth, p->joint_c_sign, shift, 1);
            }
        }
    }
    return 0;
}
