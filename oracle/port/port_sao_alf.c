/*
 * oracle/port/port_sao_alf.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's SAO, ALF (luma / chroma / cross-component) and LMCS picture passes:
 *   SAO band / edge / border + slice-tile restore     libavcodec/h26x/h2656_sao_template.c:24-215, vvc/vvc_filter.c:154-298
 *   ALF buffer padding rules                           libavcodec/vvc/vvc_filter.c:1005-1142
 *   ALF classification, coefficient set-up, filters    libavcodec/vvc/vvc_filter_template.c:38-408, vvc_filter.c:1147-1228
 *   LMCS LUT pass                                      libavcodec/vvc/vvc_filter_template.c:25-36
 * Out-of-place: dst = filter(src), which equals the reference's in-place CTB loop with saved borders.
 */
#include <stdlib.h>
#include "port_common.h"

static inline int pix(const PortPlanes *p, int c, int x, int y)
{
    return ((const uint16_t *)(p->plane[c] + (size_t)y * p->stride[c]))[x];
}
static inline void put(const PortPlanes *p, int c, int x, int y, int v)
{
    ((uint16_t *)(p->plane[c] + (size_t)y * p->stride[c]))[x] = (uint16_t)v;
}

/* ------------------------------------------------------------------ SAO */
PORT_EXPORT int port_sao_picture(const PortPlanes *dst, const PortPlanes *src, const VVCB200SaoCtb *ctbs,
                                 int ctb_log2, int bd)
{
    const int ctb_width = (src->width[0] + (1 << ctb_log2) - 1) >> ctb_log2;
    for (int c = 0; c < 3; c++) {
        const int W = src->width[c], H = src->height[c], l2 = ctb_log2 - (c ? 1 : 0);
        for (int y = 0; y < H; y++) {
            for (int x = 0; x < W; x++) {
                const VVCB200SaoCtb *p = &ctbs[(y >> l2) * ctb_width + (x >> l2)];
                const int s = pix(src, c, x, y);
                const int16_t *ov = p->offset_val[c];
                int out = s;
                if (p->type_idx[c] == VVCB200_SAO_BAND) {
                    const int k = (((s >> (bd - 5)) & 31) - p->band_position[c]) & 31;
                    out = port_clip_pixel(s + (k < 4 ? ov[k + 1] : 0), bd);
                } else if (p->type_idx[c] == VVCB200_SAO_EDGE) {
                    static const int8_t dir[4][2] = { { -1, 0 }, { 0, -1 }, { -1, -1 }, { 1, -1 } };
                    static const uint8_t cat[5] = { 1, 2, 0, 3, 4 };
                    const int eo = p->eo_class[c];
                    const int cx0 = (x >> l2) << l2, cy0 = (y >> l2) << l2;
                    const int cw = W - cx0 < (1 << l2) ? W - cx0 : 1 << l2, ch = H - cy0 < (1 << l2) ? H - cy0 : 1 << l2;
                    const int lx = x - cx0, ly = y - cy0;
                    const int bl = cx0 == 0, bt = cy0 == 0, br = cx0 + cw == W, bb = cy0 + ch == H;
                    const int nf = p->no_filter;
                    int keep = 0, border = 0;
                    if (eo != 1 && ((bl && lx == 0) || (br && lx == cw - 1)))
                        border = 1;
                    if (eo != 0 && ((bt && ly == 0) || (bb && ly == ch - 1)))
                        border = 1;
                    if (!border && nf) {
                        const int ix = eo != 1 && bl, iy = eo != 0 && bt;
                        const int w = cw - (eo != 1 && br), h = ch - (eo != 0 && bb);
                        const int sul = !(nf & 16) && eo == 2 && !bl && !bt, sur = !(nf & 32) && eo == 3 && !bt && !br;
                        const int slr = !(nf & 64) && eo == 2 && !br && !bb, sll = !(nf & 128) && eo == 3 && !bl && !bb;
                        keep |= (nf & 1) && eo != 1 && lx == 0 && ly >= iy + sul && ly < h - sll;
                        keep |= (nf & 2) && eo != 1 && lx == w - 1 && ly >= iy + sur && ly < h - slr;
                        keep |= (nf & 4) && eo != 0 && ly == 0 && lx >= ix + sul && lx < w - sur;
                        keep |= (nf & 8) && eo != 0 && ly == h - 1 && lx >= ix + sll && lx < w - slr;
                        keep |= (nf & 16) && eo == 2 && lx == 0 && ly == 0;
                        keep |= (nf & 32) && eo == 3 && lx == w - 1 && ly == 0;
                        keep |= (nf & 64) && eo == 2 && lx == w - 1 && ly == h - 1;
                        keep |= (nf & 128) && eo == 3 && lx == 0 && ly == h - 1;
                    }
                    if (border) {
                        out = port_clip_pixel(s + ov[0], bd);
                    } else if (!keep) {
                        const int a = pix(src, c, port_clip(x + dir[eo][0], 0, W - 1), port_clip(y + dir[eo][1], 0, H - 1));
                        const int b = pix(src, c, port_clip(x - dir[eo][0], 0, W - 1), port_clip(y - dir[eo][1], 0, H - 1));
                        const int d = (s > a) - (s < a) + (s > b) - (s < b);
                        out = port_clip_pixel(s + ov[cat[2 + d]], bd);
                    }
                }
                put(dst, c, x, y, out);
            }
        }
    }
    return 0;
}

/* ------------------------------------------------------------------ LMCS */
PORT_EXPORT int port_lmcs_picture(const PortPlanes *pic, const uint16_t *lut)
{
    for (int y = 0; y < pic->height[0]; y++)
        for (int x = 0; x < pic->width[0]; x++)
            put(pic, 0, x, y, lut[pix(pic, 0, x, y)]);
    return 0;
}

/* ------------------------------------------------------------------ ALF */
typedef struct Ctb { int x0, y0, w, h, edges; } Ctb;

static Ctb ctb_of(const PortPlanes *src, int c, int x, int y, int ctb_log2, int edges)
{
    Ctb g;
    const int l2 = ctb_log2 - (c ? 1 : 0);
    g.x0 = (x >> l2) << l2;
    g.y0 = (y >> l2) << l2;
    g.w = src->width[c] - g.x0 < (1 << l2) ? src->width[c] - g.x0 : 1 << l2;
    g.h = src->height[c] - g.y0 < (1 << l2) ? src->height[c] - g.y0 : 1 << l2;
    g.edges = edges | (g.x0 == 0) | (g.y0 == 0) << 1 | (g.x0 + g.w >= src->width[c]) << 2 | (g.y0 + g.h >= src->height[c]) << 3;
    return g;
}

/* the sample CTB g sees at (x, y): neighbours' samples, or its own edge samples replicated at a limit */
static int padded(const PortPlanes *src, int c, const Ctb *g, int x, int y)
{
    if (x < g->x0 && (g->edges & 1)) x = g->x0;
    if (x >= g->x0 + g->w && (g->edges & 4)) x = g->x0 + g->w - 1;
    if (y < g->y0 && (g->edges & 2)) y = g->y0;
    if (y >= g->y0 + g->h && (g->edges & 8)) y = g->y0 + g->h - 1;
    return pix(src, c, x, y);
}

static void vb_offsets(int yy, int vb, int reach, int o[4], int *near_vb)
{
    o[1] = 1; o[2] = 2; o[3] = 3;
    if (yy < vb && yy >= vb - reach - 1) {
        o[1] = yy == vb - 1 ? 0 : 1;
        o[2] = yy >= vb - 2 ? o[1] : 2;
        o[3] = yy >= vb - 3 ? o[2] : 3;
    } else if (yy >= vb && yy <= vb + reach) {
        o[1] = yy == vb ? 0 : 1;
        o[2] = yy <= vb + 1 ? o[1] : 2;
        o[3] = yy <= vb + 2 ? o[2] : 3;
    }
    *near_vb = yy == vb - 1 || yy == vb;
}

static int clip2(int cur, int a, int b, int c) { return port_clip(a - cur, -c, c) + port_clip(b - cur, -c, c); }

/* one gradient cell of alf_classify: reference loop indices (yg, xg) = even offsets into the (h+4) x (w+4) window */
static void gradient(const PortPlanes *src, const Ctb *g, int yg, int xg, int vb, int out[4])
{
    int r[4] = { g->y0 + yg - 3, g->y0 + yg - 2, g->y0 + yg - 1, g->y0 + yg };
    const int x = g->x0 + xg - 2;
    if (yg == vb)
        r[3] = r[2];
    else if (yg == vb + 2)
        r[0] = r[1];
#define S(k, dx) padded(src, 0, g, x + (dx), r[k])
    const int v0 = S(1, 0) << 1, v1 = S(2, 1) << 1;
    out[0] = abs(v0 - S(0, 0) - S(2, 0)) + abs(v1 - S(1, 1) - S(3, 1));
    out[1] = abs(v0 - S(1, -1) - S(1, 1)) + abs(v1 - S(2, 0) - S(2, 2));
    out[2] = abs(v0 - S(0, -1) - S(2, 1)) + abs(v1 - S(1, 0) - S(3, 2));
    out[3] = abs(v0 - S(0, 1) - S(2, -1)) + abs(v1 - S(1, 2) - S(3, 0));
#undef S
}

PORT_EXPORT int port_alf_picture(const PortPlanes *dst, const PortPlanes *src, const VVCB200AlfCtb *ctbs,
                                 const VVCB200AlfParams *p)
{
    static const uint8_t tr_idx[4][12] = {
        { 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11 }, { 9, 4, 10, 8, 1, 5, 11, 7, 3, 0, 2, 6 },
        { 0, 3, 2, 1, 8, 7, 6, 5, 4, 9, 10, 11 }, { 9, 8, 10, 4, 3, 7, 11, 5, 1, 0, 2, 6 },
    };
    static const uint8_t act_map[16] = { 0, 1, 2, 2, 2, 2, 2, 3, 3, 3, 3, 3, 3, 3, 3, 4 };
    static const uint8_t clip_sh[4] = { 0, 3, 5, 7 };
    const int bd = p->bit_depth, l2 = p->ctb_log2_size;
    const int ctb_width = (src->width[0] + (1 << l2) - 1) >> l2;

    /* luma, 4x4 block by 4x4 block */
    for (int by = 0; by < src->height[0]; by += 4) {
        for (int bx = 0; bx < src->width[0]; bx += 4) {
            const VVCB200AlfCtb *cp = &ctbs[(by >> l2) * ctb_width + (bx >> l2)];
            if (!cp->ctb_flag[0]) {
                for (int i = 0; i < 16; i++)
                    put(dst, 0, bx + (i & 3), by + (i >> 2), pix(src, 0, bx + (i & 3), by + (i >> 2)));
                continue;
            }
            const Ctb g = ctb_of(src, 0, bx, by, l2, cp->edges);
            const int vb = (1 << l2) - 4, yb = by - g.y0, xb = bx - g.x0;
            int start = 0, end = 4, ac = 2, sum[4] = { 0 };
            if (yb + 4 == vb) { end = 3; ac = 3; } else if (yb == vb) { start = 1; ac = 3; }
            for (int i = start; i < end; i++)
                for (int j = 0; j < 4; j++) {
                    int gr[4];
                    gradient(src, &g, yb + 2 * i, xb + 2 * j, vb, gr);
                    for (int k = 0; k < 4; k++)
                        sum[k] += gr[k];
                }
            const int dir_hv = sum[0] <= sum[1], hv1 = sum[dir_hv], hv0 = sum[!dir_hv];
            const int dir_d = sum[2] <= sum[3], d1 = sum[2 + dir_d], d0 = sum[2 + !dir_d];
            const int dir1 = (uint64_t)d1 * hv0 <= (uint64_t)hv1 * d0;
            const int hi = dir1 ? hv1 : d1, lo = dir1 ? hv0 : d0;
            int cls = act_map[port_clip(((sum[0] + sum[1]) * ac) >> (bd - 1), 0, 15)];
            if (hi * 2 > 9 * lo)
                cls += (2 * dir1 + 2) * 5;
            else if (hi > 2 * lo)
                cls += (2 * dir1 + 1) * 5;
            const int tr = dir_d * 2 + dir_hv;
            int f[12], cl[12];
            for (int j = 0; j < 12; j++) {
                const int idx = tr_idx[tr][j];
                if (cp->filt_set_idx_y < 16) {
                    f[j] = VVC_ALF_FIX_FILT_COEFF[VVC_ALF_CLASS_TO_FILT_MAP[cp->filt_set_idx_y * 25 + cls] * 12 + idx];
                    cl[j] = 1 << bd;
                } else {
                    const int k = cp->filt_set_idx_y - 16;
                    f[j] = p->luma_coeff[k][VVC_ALF_APS_CLASS_TO_FILT_MAP[cls]][idx];
                    cl[j] = 1 << (bd - clip_sh[p->luma_clip_idx[k][cls][idx]]);
                }
            }
            for (int i = 0; i < 4; i++) {
                int o[4], near_vb;
                vb_offsets(yb + i, vb, 3, o, &near_vb);
                for (int j = 0; j < 4; j++) {
                    const int x = bx + j, y = by + i;
                    const int cur = pix(src, 0, x, y);
#define T(dx, dy) padded(src, 0, &g, x + (dx), y + (dy))
                    int s = 0;
                    s += f[0] * clip2(cur, T(0, o[3]), T(0, -o[3]), cl[0]);
                    s += f[1] * clip2(cur, T(1, o[2]), T(-1, -o[2]), cl[1]);
                    s += f[2] * clip2(cur, T(0, o[2]), T(0, -o[2]), cl[2]);
                    s += f[3] * clip2(cur, T(-1, o[2]), T(1, -o[2]), cl[3]);
                    s += f[4] * clip2(cur, T(2, o[1]), T(-2, -o[1]), cl[4]);
                    s += f[5] * clip2(cur, T(1, o[1]), T(-1, -o[1]), cl[5]);
                    s += f[6] * clip2(cur, T(0, o[1]), T(0, -o[1]), cl[6]);
                    s += f[7] * clip2(cur, T(-1, o[1]), T(1, -o[1]), cl[7]);
                    s += f[8] * clip2(cur, T(-2, o[1]), T(2, -o[1]), cl[8]);
                    s += f[9] * clip2(cur, T(3, 0), T(-3, 0), cl[9]);
                    s += f[10] * clip2(cur, T(2, 0), T(-2, 0), cl[10]);
                    s += f[11] * clip2(cur, T(1, 0), T(-1, 0), cl[11]);
#undef T
                    s = near_vb ? (s + 512) >> 10 : (s + 64) >> 7;
                    put(dst, 0, x, y, port_clip_pixel(s + cur, bd));
                }
            }
        }
    }
    /* chroma + cross-component */
    for (int c = 1; c < 3; c++) {
        for (int y = 0; y < src->height[c]; y++) {
            for (int x = 0; x < src->width[c]; x++) {
                const VVCB200AlfCtb *cp = &ctbs[(y >> (l2 - 1)) * ctb_width + (x >> (l2 - 1))];
                const Ctb g = ctb_of(src, c, x, y, l2, cp->edges);
                int v = pix(src, c, x, y);
                if (cp->ctb_flag[c]) {
                    const int alt = cp->alt_idx[c - 1], vb = (1 << (l2 - 1)) - 2;
                    const int16_t *f = p->chroma_coeff[alt];
                    int cl[6], o[4], near_vb, s = 0;
                    for (int j = 0; j < 6; j++)
                        cl[j] = 1 << (bd - clip_sh[p->chroma_clip_idx[alt][j]]);
                    vb_offsets(y - g.y0, vb, 1, o, &near_vb);
#define T(dx, dy) padded(src, c, &g, x + (dx), y + (dy))
                    s += f[0] * clip2(v, T(0, o[2]), T(0, -o[2]), cl[0]);
                    s += f[1] * clip2(v, T(1, o[1]), T(-1, -o[1]), cl[1]);
                    s += f[2] * clip2(v, T(0, o[1]), T(0, -o[1]), cl[2]);
                    s += f[3] * clip2(v, T(-1, o[1]), T(1, -o[1]), cl[3]);
                    s += f[4] * clip2(v, T(2, 0), T(-2, 0), cl[4]);
                    s += f[5] * clip2(v, T(1, 0), T(-1, 0), cl[5]);
#undef T
                    s = near_vb ? (s + 512) >> 10 : (s + 64) >> 7;
                    v = port_clip_pixel(s + v, bd);
                }
                if (cp->cc_idc[c - 1] && p->cc_enabled[c - 1]) {
                    const int16_t *f = p->cc_coeff[c - 1][cp->cc_idc[c - 1] - 1];
                    const Ctb gl = ctb_of(src, 0, 2 * x, 2 * y, l2, cp->edges);
                    const int vb = (1 << l2) - 4, pos = 2 * (y - g.y0), lx = 2 * x, ly = 2 * y;
                    int r0 = ly - 1, r2 = ly + 1, r3 = ly + 2, s = 0;
                    if (pos == vb - 2 || pos == vb + 1)
                        r3 = r2;
                    else if (pos == vb - 1 || pos == vb)
                        r3 = r2 = r0 = ly;
                    const int cur = padded(src, 0, &gl, lx, ly);
                    s += f[0] * (padded(src, 0, &gl, lx, r0) - cur);
                    s += f[1] * (padded(src, 0, &gl, lx - 1, ly) - cur);
                    s += f[2] * (padded(src, 0, &gl, lx + 1, ly) - cur);
                    s += f[3] * (padded(src, 0, &gl, lx - 1, r2) - cur);
                    s += f[4] * (padded(src, 0, &gl, lx, r2) - cur);
                    s += f[5] * (padded(src, 0, &gl, lx + 1, r2) - cur);
                    s += f[6] * (padded(src, 0, &gl, lx, r3) - cur);
                    s = port_clip((s + 64) >> 7, -(1 << (bd - 1)), (1 << (bd - 1)) - 1);
                    v = port_clip_pixel(v + s, bd);
                }
                put(dst, c, x, y, v);
            }
        }
    }
    return 0;
}
