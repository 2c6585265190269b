/*
 * oracle/port/port_inter.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's translational inter prediction for a batch of sub-blocks:
 *   8-tap luma / 4-tap chroma put, put_uni, put_uni_w      libavcodec/h26x/h2656_inter_template.c:29-577
 *   avg / w_avg (+BCW weights)                              libavcodec/vvc/vvc_inter_template.c:25-58, vvc_inter.c:31,145-174
 *   MV -> integer offset + fraction, filter-set selection   libavcodec/vvc/vvc_inter.c:176-368
 *   replicate-edge reference fetch                          libavcodec/videodsp_template.c:27-105 (== coordinate clamp)
 *   forward LMCS on the luma prediction                     libavcodec/vvc/vvc_filter_template.c:25-36, vvc_inter.c:888-891
 */
#include "port_common.h"

#define TMP_STRIDE 128          /* MAX_PB_SIZE */

static inline int ref_sample(const PortPlanes *r, int c, int x, int y)
{
    x = port_clip(x, 0, r->width[c] - 1);
    y = port_clip(y, 0, r->height[c] - 1);
    return ((const uint16_t *)(r->plane[c] + (size_t)y * r->stride[c]))[x];
}

/* inter.put[luma|chroma][..][my != 0][mx != 0]: 14-bit intermediate of a w x h block whose integer
 * position in the reference is (xi, yi); taps = 8 (luma) or 4 (chroma). */
static void put_block(int16_t *dst, const PortPlanes *ref, int c, int xi, int yi, int w, int h,
                      const int8_t *hf, const int8_t *vf, int mx, int my, int taps, int bd)
{
    const int before = taps / 2 - 1;
    if (!mx && !my) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                dst[y * TMP_STRIDE + x] = ref_sample(ref, c, xi + x, yi + y) << (14 - bd);
    } else if (!my) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += hf[t] * ref_sample(ref, c, xi + x + t - before, yi + y);
                dst[y * TMP_STRIDE + x] = s >> (bd - 8);
            }
    } else if (!mx) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += vf[t] * ref_sample(ref, c, xi + x, yi + y + t - before);
                dst[y * TMP_STRIDE + x] = s >> (bd - 8);
            }
    } else {
        int16_t tmp[(16 + 7) * 16];
        for (int y = 0; y < h + taps - 1; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += hf[t] * ref_sample(ref, c, xi + x + t - before, yi + y - before);
                tmp[y * 16 + x] = s >> (bd - 8);
            }
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += vf[t] * tmp[(y + t) * 16 + x];
                dst[y * TMP_STRIDE + x] = s >> 6;
            }
    }
}

static const int bcw_w_lut[5] = { 4, 5, 3, 10, -2 };

static void predict_component(const PortPlanes *dst, const PortPlanes *refs, const VVCB200PredBlock *b,
                              const VVCB200InterParams *p, int c, const PortPlanes *tmp14)
{
    const int bd = p->bit_depth;
    const int hs = c ? 1 : 0, vs = c ? 1 : 0;
    const int w = b->w >> hs, h = b->h >> vs;
    const int x0 = b->x0 >> hs, y0 = b->y0 >> vs;
    const int taps = c ? 4 : 8;
    const int set = c ? 0 : (b->flags & VVCB200_PB_AFFINE) ? 2 : (b->flags & VVCB200_PB_HPEL_IF) ? 1 : 0;
    int16_t tmp[2][16 * TMP_STRIDE];
    uint16_t *out = (uint16_t *)(dst->plane[c] + (size_t)y0 * dst->stride[c]) + x0;
    const int ostride = dst->stride[c] / 2;

    for (int l = 0; l < 2; l++) {
        if (!(b->pred_flag & (1 << l)))
            continue;
        const int mvx = b->mv[l][0], mvy = b->mv[l][1];
        const int mx = c ? (mvx & 31) : (mvx & 15);
        const int my = c ? (mvy & 31) : (mvy & 15);
        const int xi = x0 + (mvx >> (4 + hs)), yi = y0 + (mvy >> (4 + vs));
        const int8_t *hf = c ? VVC_CHROMA_FILTERS + (set * 32 + mx) * 4 : VVC_LUMA_FILTERS + (set * 16 + mx) * 8;
        const int8_t *vf = c ? VVC_CHROMA_FILTERS + (set * 32 + my) * 4 : VVC_LUMA_FILTERS + (set * 16 + my) * 8;
        const PortPlanes *ref = &refs[p->ref_slot[l][b->ref_idx[l]]];
        put_block(tmp[l], ref, c, xi, yi, w, h, hf, vf, mx, my, taps, bd);
    }
    if (tmp14) {
        const int l = (b->pred_flag & 1) ? 0 : 1;
        for (int y = 0; y < h; y++) {
            int16_t *o = (int16_t *)(tmp14->plane[c] + (size_t)(y0 + y) * tmp14->stride[c]) + x0;
            for (int x = 0; x < w; x++)
                o[x] = tmp[l][y * TMP_STRIDE + x];
        }
        return;
    }
    if (b->pred_flag != 3) {
        const int l = b->pred_flag - 1;
        const int shift = 14 - bd;
        if (b->flags & VVCB200_PB_WP) {
            const int denom = p->log2_denom[c > 0];
            const int wx = p->weight[l][c][b->ref_idx[l]];
            const int ox = p->offset[l][c][b->ref_idx[l]] * (1 << (bd - 8));
            const int sh = denom + shift, off = 1 << (sh - 1);
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++)
                    out[y * ostride + x] = port_clip_pixel(((tmp[l][y * TMP_STRIDE + x] * wx + off) >> sh) + ox, bd);
        } else {
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++)
                    out[y * ostride + x] = port_clip_pixel((tmp[l][y * TMP_STRIDE + x] + (1 << (shift - 1))) >> shift, bd);
        }
    } else {
        const int wp = (b->flags & VVCB200_PB_WP) && !(b->flags & VVCB200_PB_DMVR);
        const int bcw = b->bcw_idx;
        const int base_shift = 15 - bd > 3 ? 15 - bd : 3;
        if ((!wp && !bcw) || (bcw && (b->flags & VVCB200_PB_CIIP))) {
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++)
                    out[y * ostride + x] = port_clip_pixel((tmp[0][y * TMP_STRIDE + x] + tmp[1][y * TMP_STRIDE + x] +
                                                            (1 << (base_shift - 1))) >> base_shift, bd);
        } else {
            int denom, w0, w1, o0, o1;
            if (bcw) {
                denom = 2; w1 = bcw_w_lut[bcw]; w0 = 8 - w1; o0 = o1 = 0;
            } else {
                denom = p->log2_denom[c > 0];
                w0 = p->weight[0][c][b->ref_idx[0]]; w1 = p->weight[1][c][b->ref_idx[1]];
                o0 = p->offset[0][c][b->ref_idx[0]]; o1 = p->offset[1][c][b->ref_idx[1]];
            }
            const int sh = denom + base_shift;
            const int off = (((o0 + o1) << (bd - 8)) + 1) << (sh - 1);
            for (int y = 0; y < h; y++)
                for (int x = 0; x < w; x++)
                    out[y * ostride + x] = port_clip_pixel((tmp[0][y * TMP_STRIDE + x] * w0 + tmp[1][y * TMP_STRIDE + x] * w1 + off) >> sh, bd);
        }
    }
}

PORT_EXPORT int port_inter_batch(const PortPlanes *dst, const PortPlanes *refs, int n_refs,
                                 const VVCB200PredBlock *blocks, int n, const VVCB200InterParams *p,
                                 const uint16_t *lmcs_fwd_lut, const PortPlanes *tmp14)
{
    (void)n_refs;
    for (int i = 0; i < n; i++) {
        const VVCB200PredBlock *b = &blocks[i];
        if (!(b->flags & VVCB200_PB_CHROMA_ONLY)) {
            predict_component(dst, refs, b, p, 0, tmp14);
            if ((b->flags & VVCB200_PB_LMCS) && lmcs_fwd_lut && !tmp14) {
                for (int y = 0; y < b->h; y++) {
                    uint16_t *o = (uint16_t *)(dst->plane[0] + (size_t)(b->y0 + y) * dst->stride[0]) + b->x0;
                    for (int x = 0; x < b->w; x++)
                        o[x] = lmcs_fwd_lut[o[x]];
                }
            }
        }
        if (!(b->flags & VVCB200_PB_LUMA_ONLY)) {
            predict_component(dst, refs, b, p, 1, tmp14);
            predict_component(dst, refs, b, p, 2, tmp14);
        }
    }
    return 0;
}
