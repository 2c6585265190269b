/*
 * oracle/port/port_inter.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's translational inter prediction for a batch of sub-blocks:
 *   8-tap luma / 4-tap chroma put, put_uni, put_uni_w      libavcodec/h26x/h2656_inter_template.c:29-577
 *   avg / w_avg (+BCW weights)                              libavcodec/vvc/vvc_inter_template.c:25-58, vvc_inter.c:31,145-174
 *   MV -> integer offset + fraction, filter-set selection   libavcodec/vvc/vvc_inter.c:176-368
 *   replicate-edge reference fetch                          libavcodec/videodsp_template.c:27-105 (== coordinate clamp)
 *   forward LMCS on the luma prediction                     libavcodec/vvc/vvc_filter_template.c:25-36, vvc_inter.c:888-891
 */
#define VVC_TABLES_WITH_GPM_WEIGHTS
#include "port_common.h"

#define TMP_STRIDE 128          /* MAX_PB_SIZE */

/* Fetch window of one (block, list, component): the whole picture for ordinary blocks; for DMVR blocks the
 * window the un-refined vector would have needed (emulated_edge_dmvr, vvc_inter.c:60-89), samples outside it are
 * replicated from its edge. */
typedef struct Win { int x0, x1, y0, y1; } Win;

static Win full_window(const PortPlanes *r, int c)
{
    Win w = { 0, r->width[c] - 1, 0, r->height[c] - 1 };
    return w;
}

static Win dmvr_window(const PortPlanes *r, int c, int x_sb, int y_sb, int bw, int bh, int before, int after)
{
    const int pw = r->width[c], ph = r->height[c];
    Win w;
    w.x0 = port_clip(x_sb - before, 0, pw - 1);
    w.y0 = port_clip(y_sb - before, 0, ph - 1);
    int ww = (pw < x_sb + bw + after ? pw : x_sb + bw + after) - w.x0;
    int hh = (ph < y_sb + bh + after ? ph : y_sb + bh + after) - w.y0;
    w.x1 = w.x0 + (ww > 1 ? ww : 1) - 1;
    w.y1 = w.y0 + (hh > 1 ? hh : 1) - 1;
    return w;
}

static inline int ref_sample(const PortPlanes *r, int c, int x, int y, const Win *w)
{
    x = port_clip(x, w->x0, w->x1);
    y = port_clip(y, w->y0, w->y1);
    return ((const uint16_t *)(r->plane[c] + (size_t)y * r->stride[c]))[x];
}

/* inter.put[luma|chroma][..][my != 0][mx != 0]: 14-bit intermediate of a w x h block whose integer
 * position in the reference is (xi, yi); taps = 8 (luma) or 4 (chroma). */
static void put_block(int16_t *dst, const PortPlanes *ref, int c, int xi, int yi, int w, int h,
                      const int8_t *hf, const int8_t *vf, int mx, int my, int taps, int bd, const Win *win)
{
    const int before = taps / 2 - 1;
    if (!mx && !my) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                dst[y * TMP_STRIDE + x] = ref_sample(ref, c, xi + x, yi + y, win) << (14 - bd);
    } else if (!my) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += hf[t] * ref_sample(ref, c, xi + x + t - before, yi + y, win);
                dst[y * TMP_STRIDE + x] = s >> (bd - 8);
            }
    } else if (!mx) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += vf[t] * ref_sample(ref, c, xi + x, yi + y + t - before, win);
                dst[y * TMP_STRIDE + x] = s >> (bd - 8);
            }
    } else {
        int16_t tmp[(16 + 7) * 16];
        for (int y = 0; y < h + taps - 1; y++)
            for (int x = 0; x < w; x++) {
                int s = 0;
                for (int t = 0; t < taps; t++)
                    s += hf[t] * ref_sample(ref, c, xi + x + t - before, yi + y - before, win);
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

/* ---- DMVR: 8.5.3, dmvr_mv_refine / parametric_mv_refine (vvc_inter.c:642-748), bilinear dmvr[][] (vvc_inter_template.c:324-413),
 * SAD (vvcdsp.c:49-65) ---------------------------------------------------------------------------------------------- */
static void dmvr_bilinear(int16_t *dst, const PortPlanes *ref, int xi, int yi, int w, int h, int mx, int my, int bd)
{
    const Win full = full_window(ref, 0);
    const int8_t *fx = VVC_DMVR_FILTERS + mx * 2, *fy = VVC_DMVR_FILTERS + my * 2;
    const int sh1 = bd - 6, of1 = 1 << (sh1 - 1);
#define S(x, y) ref_sample(ref, 0, xi + (x), yi + (y), &full)
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int v;
            if (!mx && !my)
                v = bd > 10 ? (S(x, y) + (1 << (bd - 11))) >> (bd - 10) : S(x, y) << (10 - bd);
            else if (!my)
                v = (fx[0] * S(x, y) + fx[1] * S(x + 1, y) + of1) >> sh1;
            else if (!mx)
                v = (fy[0] * S(x, y) + fy[1] * S(x, y + 1) + of1) >> sh1;
            else {
                const int t0 = (fx[0] * S(x, y) + fx[1] * S(x + 1, y) + of1) >> sh1;
                const int t1 = (fx[0] * S(x, y + 1) + fx[1] * S(x + 1, y + 1) + of1) >> sh1;
                v = (fy[0] * (int16_t)t0 + fy[1] * (int16_t)t1 + 8) >> 4;
            }
            dst[y * TMP_STRIDE + x] = v;
        }
#undef S
}

static int dmvr_sad(const int16_t *p0, const int16_t *p1, int dx, int dy, int w, int h)
{
    int sad = 0;
    dx -= 2;
    dy -= 2;
    p0 += (2 + dy) * TMP_STRIDE + 2 + dx;
    p1 += (2 - dy) * TMP_STRIDE + 2 - dx;
    for (int y = 0; y < h; y += 2)
        for (int x = 0; x < w; x++) {
            const int d = p0[y * TMP_STRIDE + x] - p1[y * TMP_STRIDE + x];
            sad += d < 0 ? -d : d;
        }
    return sad;
}

/* 8.5.3.5: sub-sample offset from three SADs on a line, in 1/16 units */
static int dmvr_parabola(int sm, int sc, int sp)
{
    int denom = ((sm + sp) - (sc << 1)) << 3;
    if (!denom)
        return 0;
    if (sm == sc)
        return -8;
    if (sp == sc)
        return 8;
    int num = (sm - sp) * 16, neg = num < 0, q = 0;
    if (neg)
        num = -num;
    for (int i = 0; i < 3; i++) {
        q <<= 1;
        if (num >= denom) {
            num -= denom;
            q++;
        }
        denom >>= 1;
    }
    return neg ? -q : q;
}

/* refines mv[2][2] in place; returns 0 if BDOF must be switched off for this sub-block (vvc_inter.c:745-747) */
static int dmvr_refine(const PortPlanes *refs, const VVCB200PredBlock *b, const VVCB200InterParams *p, int mv[2][2])
{
    int16_t local[2][20 * TMP_STRIDE];
    int sad[5][5];
    const int w = b->w, h = b->h, bd = p->bit_depth;
    for (int l = 0; l < 2; l++) {
        const PortPlanes *ref = &refs[p->ref_slot[l][b->ref_idx[l]]];
        dmvr_bilinear(local[l], ref, b->x0 + (mv[l][0] >> 4) - 2, b->y0 + (mv[l][1] >> 4) - 2, w + 4, h + 4,
                      mv[l][0] & 15, mv[l][1] & 15, bd);
    }
    int min_sad = dmvr_sad(local[0], local[1], 2, 2, w, h);
    int min_dx = 2, min_dy = 2;
    min_sad -= min_sad >> 2;
    sad[2][2] = min_sad;
    if (min_sad >= w * h) {
        for (int dy = 0; dy < 5; dy++)
            for (int dx = 0; dx < 5; dx++) {
                if (dx == 2 && dy == 2)
                    continue;
                sad[dy][dx] = dmvr_sad(local[0], local[1], dx, dy, w, h);
                if (sad[dy][dx] < min_sad) {
                    min_sad = sad[dy][dx];
                    min_dx = dx;
                    min_dy = dy;
                }
            }
        int dmv[2] = { (min_dx - 2) * 16, (min_dy - 2) * 16 };
        if (min_dx != 0 && min_dx != 4 && min_dy != 0 && min_dy != 4) {
            dmv[0] += dmvr_parabola(sad[min_dy][min_dx - 1], sad[min_dy][min_dx], sad[min_dy][min_dx + 1]);
            dmv[1] += dmvr_parabola(sad[min_dy - 1][min_dx], sad[min_dy][min_dx], sad[min_dy + 1][min_dx]);
        }
        for (int l = 0; l < 2; l++)
            for (int k = 0; k < 2; k++)
                mv[l][k] = port_clip(mv[l][k] + (1 - 2 * l) * dmv[k], -(1 << 17), (1 << 17) - 1);
    }
    return min_sad >= 2 * w * h;
}

/* ---- BDOF: bdof_fetch_samples / prof_grad_filter / derive_bdof_vx_vy / apply_bdof (vvc_inter_template.c:101-128, 135-158,
 * 237-317).  t0/t1: the two 14-bit predictions with a one-sample ring (element (y, x) at [(y + 1) * S + x + 1]), ring =
 * integer reference samples << (14 - bd).  Gradients use that ring; afterwards ring and gradient border are replaced by
 * replicated block-edge values, which is what the clamped indices below read. */
#define BS 18
static int sign3(int v) { return v < 0 ? -1 : v > 0; }

static void bdof_block(uint16_t *out, int ostride, const int16_t t[2][BS * BS], int w, int h, int bd)
{
    int16_t gh[2][16 * 16], gv[2][16 * 16];
    for (int l = 0; l < 2; l++)
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                const int16_t *c = &t[l][(y + 1) * BS + x + 1];
                gh[l][y * 16 + x] = (c[1] >> 6) - (c[-1] >> 6);
                gv[l][y * 16 + x] = (c[BS] >> 6) - (c[-BS] >> 6);
            }
    const int sh4 = 15 - bd, of4 = 1 << (sh4 - 1);
    for (int by = 0; by < h; by += 4)
        for (int bx = 0; bx < w; bx += 4) {
            int sgx2 = 0, sgy2 = 0, sgxgy = 0, sgxdi = 0, sgydi = 0;
            for (int y = by - 1; y < by + 5; y++)
                for (int x = bx - 1; x < bx + 5; x++) {
                    const int cy = port_clip(y, 0, h - 1), cx = port_clip(x, 0, w - 1);
                    const int i = cy * 16 + cx;
                    const int diff = (t[0][(cy + 1) * BS + cx + 1] >> 4) - (t[1][(cy + 1) * BS + cx + 1] >> 4);
                    const int th = (gh[0][i] + gh[1][i]) >> 1, tv = (gv[0][i] + gv[1][i]) >> 1;
                    sgx2 += th < 0 ? -th : th;
                    sgy2 += tv < 0 ? -tv : tv;
                    sgxgy += sign3(tv) * th;
                    sgxdi += -sign3(th) * diff;
                    sgydi += -sign3(tv) * diff;
                }
            int vx = 0, vy = 0;
            if (sgx2 > 0) {
                int lg = 0;
                while ((sgx2 >> (lg + 1)) > 0)
                    lg++;
                vx = port_clip((sgxdi * 4) >> lg, -15, 15);
            }
            if (sgy2 > 0) {
                int lg = 0;
                while ((sgy2 >> (lg + 1)) > 0)
                    lg++;
                vy = port_clip(((sgydi * 4) - ((vx * sgxgy) >> 1)) >> lg, -15, 15);
            }
            for (int y = by; y < by + 4; y++)
                for (int x = bx; x < bx + 4; x++) {
                    const int i = y * 16 + x;
                    const int off = vx * (gh[0][i] - gh[1][i]) + vy * (gv[0][i] - gv[1][i]);
                    out[y * ostride + x] = port_clip_pixel((t[0][(y + 1) * BS + x + 1] + of4 + t[1][(y + 1) * BS + x + 1] + off) >> sh4, bd);
                }
        }
}

static const int bcw_w_lut[5] = { 4, 5, 3, 10, -2 };

static void predict_component(const PortPlanes *dst, const PortPlanes *refs, const VVCB200PredBlock *b,
                              const VVCB200InterParams *p, int c, const PortPlanes *tmp14,
                              const int mv[2][2], int dmvr, int bdof, const VVCB200ProfTable *prof)
{
    const int bd = p->bit_depth;
    const int hs = c ? 1 : 0, vs = c ? 1 : 0;
    const int w = b->w >> hs, h = b->h >> vs;
    const int x0 = b->x0 >> hs, y0 = b->y0 >> vs;
    const int taps = c ? 4 : 8;
    const int gpm = !!(b->flags & VVCB200_PB_GPM);
    const int set = (c || gpm) ? 0 : (b->flags & VVCB200_PB_AFFINE) ? 2 : (b->flags & VVCB200_PB_HPEL_IF) ? 1 : 0;
    int16_t tmp[2][16 * TMP_STRIDE];
    uint16_t *out = (uint16_t *)(dst->plane[c] + (size_t)y0 * dst->stride[c]) + x0;
    const int ostride = dst->stride[c] / 2;
    int16_t ring[2][BS * BS];

    for (int l = 0; l < 2; l++) {
        if (!(b->pred_flag & (1 << l)))
            continue;
        const int mvx = mv[l][0], mvy = mv[l][1];
        const int mx = c ? (mvx & 31) : (mvx & 15);
        const int my = c ? (mvy & 31) : (mvy & 15);
        const int xi = x0 + (mvx >> (4 + hs)), yi = y0 + (mvy >> (4 + vs));
        const int8_t *hf = c ? VVC_CHROMA_FILTERS + (set * 32 + mx) * 4 : VVC_LUMA_FILTERS + (set * 16 + mx) * 8;
        const int8_t *vf = c ? VVC_CHROMA_FILTERS + (set * 32 + my) * 4 : VVC_LUMA_FILTERS + (set * 16 + my) * 8;
        /* GPM: prediction l comes from gpm_mv[l], which names its own reference list (vvc_inter.c:507-509) */
        const int list = gpm ? (b->aux >> (18 + l)) & 1 : l;
        const PortPlanes *ref = &refs[p->ref_slot[list][b->ref_idx[l]]];
        Win win = full_window(ref, c);
        if (dmvr)       /* fetches are confined to what the un-refined vector would have read */
            win = dmvr_window(ref, c, x0 + (b->mv[l][0] >> (4 + hs)), y0 + (b->mv[l][1] >> (4 + vs)), w, h,
                              taps / 2 - 1, taps / 2);
        put_block(tmp[l], ref, c, xi, yi, w, h, hf, vf, mx, my, taps, bd, &win);
        if (!c && prof && (b->flags & (l ? VVCB200_PB_PROF1 : VVCB200_PB_PROF0))) {
            /* PROF (fetch_samples + apply_prof*, vvc_inter_template.c:130-235): the 4x4 prediction gets a ring of integer
             * samples, gradients of the ringed block, and a per-sample offset gh * dmv_x + gv * dmv_y */
            const VVCB200ProfTable *pt = &prof[b->aux];
            const int ox = xi + (mx >> 3) - 1, oy = yi + (my >> 3) - 1;
            const int limit = 1 << (bd + 1 > 13 ? bd + 1 : 13);
            int16_t r[6 * 6];
            for (int y = -1; y <= 4; y++)
                for (int x = -1; x <= 4; x++)
                    r[(y + 1) * 6 + x + 1] = (x < 0 || y < 0 || x >= 4 || y >= 4)
                        ? ref_sample(ref, 0, ox + x + 1, oy + y + 1, &win) << (14 - bd) : tmp[l][y * TMP_STRIDE + x];
            for (int y = 0; y < 4; y++)
                for (int x = 0; x < 4; x++) {
                    const int16_t *q = &r[(y + 1) * 6 + x + 1];
                    const int gh = (q[1] >> 6) - (q[-1] >> 6), gv = (q[6] >> 6) - (q[-6] >> 6);
                    const int di = gh * pt->diff_mv_x[l][y * 4 + x] + gv * pt->diff_mv_y[l][y * 4 + x];
                    tmp[l][y * TMP_STRIDE + x] = q[0] + port_clip(di, -limit, limit - 1);
                }
        }
        if (bdof && !c) {
            /* the 14-bit prediction plus a ring of integer samples at the rounded position (bdof_fetch_samples) */
            const int ox = xi + (mx >> 3) - 1, oy = yi + (my >> 3) - 1;
            for (int y = -1; y <= h; y++)
                for (int x = -1; x <= w; x++)
                    ring[l][(y + 1) * BS + x + 1] = (x < 0 || y < 0 || x >= w || y >= h)
                        ? ref_sample(ref, 0, ox + x + 1, oy + y + 1, &win) << (14 - bd) : tmp[l][y * TMP_STRIDE + x];
        }
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
    if (bdof && !c) {
        bdof_block(out, ostride, (const int16_t (*)[BS * BS])ring, w, h, bd);
        return;
    }
    if (gpm) {
        /* put_gpm (vvc_inter_template.c:78-98) with the weight addressing of pred_gpm_blk (vvc_inter.c:466-505) */
        const int pidx = b->aux & 63, lw = (b->aux >> 6) & 3, lh = (b->aux >> 8) & 3;
        const int cu_x = ((b->aux >> 10) & 15) << 2, cu_y = ((b->aux >> 14) & 15) << 2;     /* luma offset inside the CU */
        const int angle = VVC_GPM_ANGLE_IDX[pidx];
        const uint8_t *wt = VVC_GPM_WEIGHTS + VVC_GPM_ANGLE_TO_WEIGHTS_IDX[angle] * 112 * 112;
        const int off_x = VVC_GPM_WEIGHTS_OFFSET_X[(pidx * 4 + lh) * 4 + lw], off_y = VVC_GPM_WEIGHTS_OFFSET_Y[(pidx * 4 + lh) * 4 + lw];
        const int mirror = VVC_GPM_ANGLE_TO_MIRROR[angle];
        const int shift = 17 - bd > 5 ? 17 - bd : 5, offset = 1 << (shift - 1);
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                const int X = cu_x + (x << hs), Y = cu_y + (y << vs);              /* luma position inside the CU */
                const int wx = mirror == 1 ? 111 - off_x - X : off_x + X;
                const int wy = mirror == 2 ? 111 - off_y - Y : off_y + Y;
                const int wgt = wt[wy * 112 + wx];
                out[y * ostride + x] = port_clip_pixel((tmp[0][y * TMP_STRIDE + x] * wgt + tmp[1][y * TMP_STRIDE + x] * (8 - wgt) + offset) >> shift, bd);
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

/* refined_mv (optional, n x 2 x 2): the vectors actually used, i.e. what the reference stores in tab_dmvr_mvf */
PORT_EXPORT int port_inter_batch(const PortPlanes *dst, const PortPlanes *refs, int n_refs,
                                 const VVCB200PredBlock *blocks, int n, const VVCB200InterParams *p,
                                 const uint16_t *lmcs_fwd_lut, const PortPlanes *tmp14, int32_t *refined_mv,
                                 const VVCB200ProfTable *prof)
{
    (void)n_refs;
    for (int i = 0; i < n; i++) {
        const VVCB200PredBlock *b = &blocks[i];
        int mv[2][2] = { { b->mv[0][0], b->mv[0][1] }, { b->mv[1][0], b->mv[1][1] } };
        const int dmvr = (b->flags & VVCB200_PB_DMVR) && b->pred_flag == 3;
        int bdof = (b->flags & VVCB200_PB_BDOF) && b->pred_flag == 3;
        if (dmvr && !dmvr_refine(refs, b, p, mv))
            bdof = 0;
        if (refined_mv)
            for (int k = 0; k < 4; k++)
                refined_mv[i * 4 + k] = mv[k >> 1][k & 1];
        if (!(b->flags & VVCB200_PB_CHROMA_ONLY)) {
            predict_component(dst, refs, b, p, 0, tmp14, mv, dmvr, bdof, prof);
            if ((b->flags & VVCB200_PB_LMCS) && lmcs_fwd_lut && !tmp14) {
                for (int y = 0; y < b->h; y++) {
                    uint16_t *o = (uint16_t *)(dst->plane[0] + (size_t)(b->y0 + y) * dst->stride[0]) + b->x0;
                    for (int x = 0; x < b->w; x++)
                        o[x] = lmcs_fwd_lut[o[x]];
                }
            }
        }
        if (!(b->flags & VVCB200_PB_LUMA_ONLY)) {
            predict_component(dst, refs, b, p, 1, tmp14, mv, dmvr, 0, prof);
            predict_component(dst, refs, b, p, 2, tmp14, mv, dmvr, 0, prof);
        }
    }
    return 0;
}
