/*
 * oracle/port/port_intra.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's intra prediction for a picture's prediction blocks in decoding order:
 *   reference sample preparation (substitution, [1 2 1] smoothing, MRL, angular extension)   vvc_intra_template.c:450-598
 *   planar / DC / vertical / horizontal / angular + PDPC                                       vvc_intra_template.c:600-705, 838-1001
 *   MIP (boundary down-sampling, matrix product, up-sampling)                                  vvc_intra_template.c:707-836
 *   CCLM (luma down-sampling, 4-point min/max linear model)                                    vvc_intra_template.c:29-375
 *   helpers: PDPC / nscale / reference-filter decisions                                        vvc_intra.c:538-660
 * The host states how many neighbour samples are reconstructed (VVCB200IntraBlock.avail_*); prediction is followed by
 * the addition of the block's residual from the int16 residual planes when the record says so.
 */
#include <stdlib.h>
#include "port_common.h"

#define EDGE 64
typedef struct Edge {
    uint16_t left_a[6 * EDGE + 8], top_a[6 * EDGE + 8], fleft_a[6 * EDGE + 8], ftop_a[6 * EDGE + 8];
    uint16_t *left, *top;
    int filter_flag;
} Edge;

static int ilog2(int v) { int l = 0; while (v >> (l + 1)) l++; return l; }
static int angle_of(int mode) { return VVC_INTRA_ANGLE[mode + 14]; }
static int inv_angle_of(int mode) { return VVC_INTRA_INV_ANGLE[mode + 14]; }

static int ref_filter_flag(int mode)
{
    static const int8_t modes[] = { -14, -12, -10, -6, 0, 2, 34, 66, 72, 76, 78, 80 };
    for (unsigned i = 0; i < sizeof(modes); i++)
        if (modes[i] == mode)
            return 1;
    return 0;
}

static int nscale_of(int w, int h, int mode)
{
    if (mode == 0 || mode == 1 || mode == 18 || mode == 50)
        return (ilog2(w) + ilog2(h) - 2) >> 2;
    const int side = mode >= 50 ? h : w;                  /* callers never pass 18 < mode < 50 */
    const int ns = ilog2(side) - ilog2(3 * inv_angle_of(mode) - 2) + 8;
    return ns < 2 ? ns : 2;
}

static int need_pdpc(int w, int h, int bdpcm, int mode, int ref_idx)
{
    if (w >= 4 && h >= 4 && !ref_idx && !bdpcm) {
        if (mode == 0 || mode == 1 || mode == 18 || mode == 50)
            return 1;
        if (mode > 18 && mode < 50)
            return 0;
        return nscale_of(w, h, mode) >= 0;
    }
    return 0;
}

#define PIX(dx, dy) src[(dx) + (ptrdiff_t)stride * (dy)]

static void prepare_edge(Edge *e, const uint16_t *src, int stride, int w, int h, const VVCB200IntraBlock *b, int pdpc, int bd)
{
    const int c_idx = b->c_idx, mode = b->mode, ref_idx = b->ref_idx;
    const int is_mip = b->flags & VVCB200_IB_MIP, isp = b->flags & VVCB200_IB_ISP;
    const int rff = is_mip ? 0 : ref_filter_flag(mode);
    const int filter_flag = !ref_idx && w * h > 32 && !c_idx && !isp && rff;
    const int ref_line = ref_idx == 3 ? -4 : -1 - ref_idx;
    uint16_t *left = e->left_a + EDGE + 4, *top = e->top_a + EDGE + 4;
    int left_size, top_size, ul, ut, refw = 0, refh = 0, angle = 0, inv = 0;

    if (is_mip || mode == 0) {
        left_size = h + 1; top_size = w + 1;
        ul = left_size + filter_flag; ut = top_size + filter_flag;
    } else if (mode == 1) {
        ul = left_size = h; ut = top_size = w;
    } else if (mode == 50) {
        ul = left_size = pdpc ? h : 1; ut = top_size = w;
    } else if (mode == 18) {
        ul = left_size = h; ut = top_size = pdpc ? w : 1;
    } else {
        if (!isp || c_idx) { refw = w * 2; refh = h * 2; }
        else { refw = b->cb_w + w; refh = b->cb_h + h; }
        angle = angle_of(mode);
        inv = inv_angle_of(mode);
        ut = top_size = refw; ul = left_size = refh;
    }
    const int la = ul < b->avail_left ? ul : b->avail_left;
    const int ta = ut < b->avail_top ? ut : b->avail_top;
    for (int i = 0; i < la; i++)
        left[i] = PIX(ref_line, i);
    for (int i = 0; i < ta; i++)
        top[i] = PIX(i, ref_line);
    for (int i = -1; i >= ref_line; i--) {
        if (b->up_left) { left[i] = PIX(ref_line, i); top[i] = PIX(i, ref_line); }
        else if (la)    { left[i] = top[i] = left[0]; }
        else if (ta)    { left[i] = top[i] = top[0]; }
        else            { left[i] = top[i] = 1 << (bd - 1); }
    }
    for (int i = ta; i < ut; i++)
        top[i] = top[ta - 1];
    for (int i = la; i < ul; i++)
        left[i] = left[la - 1];
    if (rff && filter_flag) {
        uint16_t *fl = e->fleft_a + EDGE + 4, *ft = e->ftop_a + EDGE + 4;
        const int keep_last = left_size == ul;
        fl[-1] = ft[-1] = (left[0] + 2 * left[-1] + top[0] + 2) >> 2;
        for (int i = 0; i < ul - keep_last; i++)
            fl[i] = (left[i - 1] + 2 * left[i] + left[i + 1] + 2) >> 2;
        for (int i = 0; i < ut - keep_last; i++)
            ft[i] = (top[i - 1] + 2 * top[i] + top[i + 1] + 2) >> 2;
        if (keep_last) {
            ft[ut - 1] = top[ut - 1];
            fl[ul - 1] = left[ul - 1];
        }
        left = fl;
        top = ft;
    }
    e->filter_flag = 0;
    if (!is_mip && mode != 0 && mode != 1) {
        if (!(rff || ref_idx || isp)) {
            static const int thres[] = { 24, 14, 2, 0, 0 };
            const int d50 = abs(mode - 50), d18 = abs(mode - 18);
            e->filter_flag = (d50 < d18 ? d50 : d18) > thres[((ilog2(w) + ilog2(h)) >> 1) - 2];
        }
        if (mode != 50 && mode != 18) {
            if (mode >= 34) {
                if (angle < 0) {
                    uint16_t *p = top - (ref_idx + 1);
                    for (int x = -h; x < 0; x++) {
                        const int t = (x * inv + 256) >> 9;
                        p[x] = left[-1 - ref_idx + (t < h ? t : h)];
                    }
                } else {
                    for (int i = refw; i <= refw + (w / h > 1 ? w / h : 1) * ref_idx + 1; i++)
                        top[i] = top[refw - 1];
                }
            } else {
                if (angle < 0) {
                    uint16_t *p = left - (ref_idx + 1);
                    for (int x = -w; x < 0; x++) {
                        const int t = (x * inv + 256) >> 9;
                        p[x] = top[-1 - ref_idx + (t < w ? t : w)];
                    }
                } else {
                    for (int i = refh; i <= refh + (h / w > 1 ? h / w : 1) * ref_idx + 1; i++)
                        left[i] = left[refh - 1];
                }
            }
        }
    }
    e->left = left;
    e->top = top;
}

static void pred_mip(uint16_t *src, int stride, const uint16_t *top, const uint16_t *left, int w, int h, int mode_id,
                     int transposed, int bd)
{
    const int size_id = (w == 4 && h == 4) ? 0 : ((w == 4 || h == 4) || (w == 8 && h == 8)) ? 1 : 2;
    const int bsz = size_id ? 4 : 2, psz = size_id == 2 ? 8 : 4;
    const int in_size = 2 * bsz - (size_id == 2);
    const int nin = size_id == 0 ? 4 : size_id == 1 ? 8 : 7, nout = size_id == 2 ? 64 : 16;
    const uint8_t *m = VVC_MIP_MAT + VVC_MIP_OFF[size_id] + mode_id * nout * nin;
    const int up_hor = w / psz, up_ver = h / psz;
    int red[16], *rt = red, *rl = red + bsz;
    if (transposed) { int *t = rt; rt = rl; rl = t; }
    for (int pass = 0; pass < 2; pass++) {                      /* boundary down-sampling */
        const uint16_t *ref = pass ? left : top;
        int *out = pass ? rl : rt;
        const int n = pass ? h : w, dwn = n / bsz;
        for (int i = 0; i < bsz; i++) {
            int r = 0;
            for (int j = 0; j < dwn; j++)
                r += ref[i * dwn + j];
            out[i] = dwn == 1 ? r : (r + (dwn >> 1)) >> ilog2(dwn);
        }
    }
    const int temp0 = red[0];
    int off = 1, ow;
    if (size_id != 2) { off = 0; ow = (1 << (bd - 1)) - temp0; }
    else ow = red[1] - temp0;
    red[0] = ow;
    for (int i = 1; i < in_size; i++) {
        red[i] = red[i + off] - temp0;
        ow += red[i];
    }
    ow = 32 - 32 * ow;
    for (int y = 0; y < psz; y++)
        for (int x = 0; x < psz; x++) {
            int p = 0;
            for (int i = 0; i < in_size; i++)
                p += red[i] * m[i];
            m += in_size;
            p = port_clip_pixel(((p + ow) >> 6) + temp0, bd);
            if (transposed)
                src[(y * up_hor + up_hor - 1) + (ptrdiff_t)stride * (x * up_ver + up_ver - 1)] = p;
            else
                src[(x * up_hor + up_hor - 1) + (ptrdiff_t)stride * (y * up_ver + up_ver - 1)] = p;
        }
    if (up_hor > 1)                                              /* horizontal, on the rows that hold predictions */
        for (int i = 0; i < psz; i++) {
            uint16_t *row = src + (ptrdiff_t)stride * (i * up_ver + up_ver - 1);
            int before = left[i * up_ver + up_ver - 1];
            for (int j = 0; j < psz; j++) {
                const int after = row[j * up_hor + up_hor - 1];
                for (int k = 1; k < up_hor; k++)
                    row[j * up_hor + k - 1] = ((up_hor - k) * before + k * after + up_hor / 2) / up_hor;
                before = after;
            }
        }
    if (up_ver > 1)                                              /* vertical, every column */
        for (int x = 0; x < w; x++) {
            int before = top[x];
            for (int j = 0; j < psz; j++) {
                const int after = src[x + (ptrdiff_t)stride * (j * up_ver + up_ver - 1)];
                for (int k = 1; k < up_ver; k++)
                    src[x + (ptrdiff_t)stride * (j * up_ver + k - 1)] = ((up_ver - k) * before + k * after + up_ver / 2) / up_ver;
                before = after;
            }
        }
}

static void intra_pred_block(uint16_t *src, int stride, const VVCB200IntraBlock *b, int bd)
{
    const int w = b->w, h = b->h, c_idx = b->c_idx, mode = b->mode, ref_idx = b->ref_idx;
    const int is_mip = b->flags & VVCB200_IB_MIP;
    const int pdpc = need_pdpc(w, h, !!(b->flags & VVCB200_IB_BDPCM), mode, ref_idx);
    static Edge edge_storage;
    Edge *e = &edge_storage;
    prepare_edge(e, src, stride, w, h, b, pdpc, bd);
    const uint16_t *left = e->left, *top = e->top;

    if (is_mip) {
        pred_mip(src, stride, top, left, w, h, b->mip_mode, !!(b->flags & VVCB200_IB_MIP_TRANSPOSED), bd);
    } else if (mode == 0) {
        const int lw = ilog2(w), lh = ilog2(h);
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                const int pv = ((h - 1 - y) * top[x] + (y + 1) * left[h]) << lw;
                const int ph = ((w - 1 - x) * left[y] + (x + 1) * top[w]) << lh;
                PIX(x, y) = (pv + ph + w * h) >> (lw + lh + 1);
            }
    } else if (mode == 1) {
        int sum = 0;
        unsigned offset = w == h ? w << 1 : (w > h ? w : h);
        const int shift = ilog2(offset);
        if (w >= h) for (int i = 0; i < w; i++) sum += top[i];
        if (w <= h) for (int i = 0; i < h; i++) sum += left[i];
        const int dc = (sum + (offset >> 1)) >> shift;
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                PIX(x, y) = dc;
    } else if (mode == 50) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                PIX(x, y) = top[x];
    } else if (mode == 18) {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                PIX(x, y) = left[y];
    } else {
        const int angle = angle_of(mode), vertical = mode >= 34;
        const int inv = pdpc ? inv_angle_of(mode) : 0, nscale = pdpc ? nscale_of(w, h, mode) : 0;
        /* main = the reference row/column the angle walks along; side = the other one (PDPC) */
        const uint16_t *main_ref = (vertical ? top : left) - (1 + ref_idx), *side = vertical ? left : top;
        const int n_main = vertical ? h : w, n_cross = vertical ? w : h;
        for (int a = 0; a < n_main; a++) {                    /* a: position along the angle's axis */
            const int pos = (1 + ref_idx + a) * angle;
            const int idx = (pos >> 5) + ref_idx, fact = pos & 31;
            for (int c = 0; c < n_cross; c++) {
                const uint16_t *p = main_ref + c + idx;
                int pred;
                if (!fact && (c_idx || !e->filter_flag)) {
                    pred = p[1];
                } else if (!c_idx) {
                    const int8_t *f = VVC_INTRA_LUMA_FILTER + (e->filter_flag * 32 + fact) * 4;
                    pred = port_clip_pixel((p[0] * f[0] + p[1] * f[1] + p[2] * f[2] + p[3] * f[3] + 32) >> 6, bd);
                } else {
                    pred = ((32 - fact) * p[1] + fact * p[2] + 16) >> 5;
                }
                if (pdpc && c < (3 << nscale)) {
                    const int s = side[a + ((256 + (c + 1) * inv) >> 9)];
                    int wgt = 32 >> ((c << 1) >> nscale);
                    pred = port_clip_pixel(pred + (((s - pred) * wgt + 32) >> 6), bd);
                }
                if (vertical)
                    PIX(c, a) = pred;
                else
                    PIX(a, c) = pred;
            }
        }
    }
    if (pdpc && !is_mip && (mode == 0 || mode == 1 || mode == 18 || mode == 50)) {
        const int scale = (ilog2(w) + ilog2(h) - 2) >> 2;
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) {
                const int val = PIX(x, y);
                int l, t, wl, wt;
                const int sx = (x << 1) >> scale, sy = (y << 1) >> scale;
                if (mode == 0 || mode == 1) {
                    l = left[y]; t = top[x];
                    wl = 32 >> (sx < 31 ? sx : 31); wt = 32 >> (sy < 31 ? sy : 31);
                } else {
                    l = left[y] - left[-1] + val; t = top[x] - top[-1] + val;
                    wl = mode == 50 ? 32 >> (sx < 31 ? sx : 31) : 0;
                    wt = mode == 18 ? 32 >> (sy < 31 ? sy : 31) : 0;
                }
                PIX(x, y) = port_clip_pixel(val + ((wl * (l - val) + wt * (t - val) + 32) >> 6), bd);
            }
    }
}

/* ---- CCLM ------------------------------------------------------------------------------------------------------ */
static void cclm_block(const PortPlanes *pic, const VVCB200IntraBlock *b, const VVCB200IntraParams *ip)
{
    const int bd = ip->bit_depth, w = b->w, h = b->h, x = b->x0, y = b->y0, x0 = x * 2, y0 = y * 2;
    const int avail_t = b->avail_top > 0, avail_l = b->avail_left > 0;
    const int mode = b->mode;
    const ptrdiff_t ls = pic->stride[0] / 2;
    const uint16_t *luma = (const uint16_t *)pic->plane[0] + x0 + y0 * ls;
    uint16_t *dst[2] = { (uint16_t *)pic->plane[1] + x + (ptrdiff_t)y * (pic->stride[1] / 2),
                         (uint16_t *)pic->plane[2] + x + (ptrdiff_t)y * (pic->stride[2] / 2) };
    const ptrdiff_t cs[2] = { pic->stride[1] / 2, pic->stride[2] / 2 };
    int a[2] = { 0, 0 }, bb[2] = { 1 << (bd - 1), 1 << (bd - 1) }, k[2] = { 0, 0 };
    static uint16_t dsy[64 * 64];

    if (!avail_t && !avail_l) {
        for (int c = 0; c < 2; c++)
            for (int j = 0; j < h; j++)
                for (int i = 0; i < w; i++)
                    dst[c][i + j * cs[c]] = 1 << (bd - 1);
        return;
    }
#define L(dx, dy) luma[(dx) + (ptrdiff_t)ls * (dy)]
    /* down-sampled co-located luma (cclm_get_luma_rec_pixels); unavailable left / top neighbours repeat the block edge */
    for (int j = 0; j < h; j++)
        for (int i = 0; i < w; i++) {
            const int lx = 2 * i, ly = 2 * j;
            const int xl = (i || avail_l) ? lx - 1 : lx;
            if (ip->chroma_vertical_collocated) {
                const int yt = (j || avail_t) ? ly - 1 : ly;
                dsy[j * w + i] = (L(xl, ly) + L(lx, yt) + 4 * L(lx, ly) + L(lx + 1, ly) + L(lx, ly + 1) + 4) >> 3;
            } else {
                dsy[j * w + i] = (L(xl, ly) + L(xl, ly + 1) + 2 * L(lx, ly) + 2 * L(lx, ly + 1) + L(lx + 1, ly) + L(lx + 1, ly + 1) + 4) >> 3;
            }
        }
    /* sample positions (cclm_get_select_pos) */
    int num[2], cnt[2] = { 0, 0 }, pos[2][4];
    const int mn = w < h ? w : h;
    const int is4 = !avail_t || !avail_l || mode != 81;
    if (mode == 81) {
        num[0] = avail_t ? w : 0;
        num[1] = avail_l ? h : 0;
    } else {
        const int tt = w + mn < b->avail_top ? w + mn : b->avail_top, tl = h + mn < b->avail_left ? h + mn : b->avail_left;
        num[0] = (avail_t && mode == 83) ? tt : 0;
        num[1] = (avail_l && mode == 82) ? tl : 0;
    }
    if (num[0] || num[1]) {
        uint16_t sel[3][8];
        for (int i = 0; i < 2; i++) {
            const int start = num[i] >> (2 + is4), step = (num[i] >> (1 + is4)) > 1 ? num[i] >> (1 + is4) : 1;
            cnt[i] = num[i] < ((1 + is4) << 1) ? num[i] : (1 + is4) << 1;
            for (int c = 0; c < cnt[i]; c++)
                pos[i][c] = start + c * step;
        }
        const int ctu_boundary = !(y0 & ((1 << ip->ctb_log2_size) - 1));
        for (int i = 0; i < cnt[0]; i++) {                       /* top neighbours */
            const int lx = pos[0][i] * 2;
            const int xl = (lx || avail_l) ? lx - 1 : lx;
            if (!ctu_boundary) {
                if (ip->chroma_vertical_collocated)
                    sel[0][i] = (L(lx, -3) + L(xl, -2) + 4 * L(lx, -2) + L(lx + 1, -2) + L(lx, -1) + 4) >> 3;
                else
                    sel[0][i] = (L(xl, -2) + L(xl, -1) + 2 * (L(lx, -2) + L(lx, -1)) + L(lx + 1, -2) + L(lx + 1, -1) + 4) >> 3;
            } else {
                sel[0][i] = (L(xl, -1) + 2 * L(lx, -1) + L(lx + 1, -1) + 2) >> 2;
            }
            sel[1][i] = dst[0][pos[0][i] - cs[0]];
            sel[2][i] = dst[1][pos[0][i] - cs[1]];
        }
        for (int i = 0; i < cnt[1]; i++) {                       /* left neighbours (avail_l is 1 whenever cnt[1] > 0) */
            const int ly = pos[1][i] * 2;
            if (ip->chroma_vertical_collocated) {
                const int yt = (ly || avail_t) ? ly - 1 : ly;
                sel[0][cnt[0] + i] = (L(-3, ly) + L(-2, yt) + 4 * L(-2, ly) + L(-1, ly) + L(-2, ly + 1) + 4) >> 3;
            } else {
                sel[0][cnt[0] + i] = (L(-3, ly) + L(-3, ly + 1) + 2 * L(-2, ly) + 2 * L(-2, ly + 1) + L(-1, ly) + L(-1, ly + 1) + 4) >> 3;
            }
            sel[1][cnt[0] + i] = dst[0][-1 + pos[1][i] * cs[0]];
            sel[2][cnt[0] + i] = dst[1][-1 + pos[1][i] * cs[1]];
        }
        if (cnt[0] + cnt[1] == 2)
            for (int c = 0; c < 3; c++) {
                sel[c][3] = sel[c][0]; sel[c][2] = sel[c][1]; sel[c][0] = sel[c][1]; sel[c][1] = sel[c][3];
            }
        int mi[2] = { 0, 2 }, ma[2] = { 1, 3 }, t;
#define SWAP(p, q) (t = p, p = q, q = t)
        if (sel[0][mi[0]] > sel[0][mi[1]]) SWAP(mi[0], mi[1]);
        if (sel[0][ma[0]] > sel[0][ma[1]]) SWAP(ma[0], ma[1]);
        if (sel[0][mi[0]] > sel[0][ma[1]]) { SWAP(mi[0], ma[0]); SWAP(mi[1], ma[1]); }
        if (sel[0][mi[1]] > sel[0][ma[0]]) SWAP(mi[1], ma[0]);
        int mx[3], mnv[3];
        for (int c = 0; c < 3; c++) {
            mx[c] = (sel[c][ma[0]] + sel[c][ma[1]] + 1) >> 1;
            mnv[c] = (sel[c][mi[0]] + sel[c][mi[1]] + 1) >> 1;
        }
        const int diff = mx[0] - mnv[0];
        if (!diff) {
            for (int i = 0; i < 2; i++) { a[i] = k[i] = 0; bb[i] = mnv[i + 1]; }
        } else {
            static const int div_sig[] = { 0, 7, 6, 5, 5, 4, 4, 3, 3, 2, 2, 1, 1, 1, 1, 0 };
            for (int i = 0; i < 2; i++) {
                const int diffc = mx[i + 1] - mnv[i + 1];
                int xx = ilog2(diff);
                const int norm = ((diff << 4) >> xx) & 15;
                xx += norm ? 1 : 0;
                const int yy = abs(diffc) > 0 ? ilog2(abs(diffc)) + 1 : 0;
                const int v = div_sig[norm] | 8;
                a[i] = (diffc * v + (1 << yy >> 1)) >> yy;
                k[i] = 3 + xx - yy > 1 ? 3 + xx - yy : 1;
                if (3 + xx - yy < 1)
                    a[i] = (a[i] < 0 ? -1 : a[i] > 0) * 15;
                bb[i] = mnv[i + 1] - ((a[i] * mnv[0]) >> k[i]);
            }
        }
    }
#undef L
    for (int c = 0; c < 2; c++)
        for (int j = 0; j < h; j++)
            for (int i = 0; i < w; i++)
                dst[c][i + j * cs[c]] = port_clip_pixel(((dsy[j * w + i] * a[c]) >> k[c]) + bb[c], bd);
}

/* lmcs_derive_chroma_scale (vvc_intra_template.c:376-428): the factor of the VPDU from the mean of the reconstructed luma
 * samples in the column left of it and the row above it; 0 = no scaling */
static int vpdu_chroma_scale(const PortPlanes *pic, const VVCB200IntraBlock *b, const VVCB200IntraParams *ip)
{
    if (!(b->flags & VVCB200_IB_LMCS_SCALE))
        return 0;
    const int size = (1 << ip->ctb_log2_size) < 64 ? 1 << ip->ctb_log2_size : 64;
    const int x = b->vpdu_x * size, y = b->vpdu_y * size;
    const int rows = size < pic->height[0] - y ? size : pic->height[0] - y;
    const int cols = size < pic->width[0] - x ? size : pic->width[0] - x;
    const ptrdiff_t ls = pic->stride[0] / 2;
    const uint16_t *luma = (const uint16_t *)pic->plane[0] + (ptrdiff_t)y * ls + x;
    int sum = 0, cnt = 0;
    if (b->vpdu_avail & 1) {
        for (int i = 0; i < size; i++)
            sum += luma[-1 + ls * (i < rows ? i : rows - 1)];
        cnt = size;
    }
    if (b->vpdu_avail & 2) {
        for (int i = 0; i < size; i++)
            sum += luma[-ls + (i < cols ? i : cols - 1)];
        cnt += size;
    }
    const int mean = cnt ? (sum + (cnt >> 1)) >> ilog2(cnt) : 1 << (ip->bit_depth - 1);
    int i = ip->lmcs_min_bin_idx;
    while (i <= ip->lmcs_max_bin_idx && i < 16 && mean >= ip->lmcs_pivot[i + 1])
        i++;
    return ip->lmcs_chroma_scale_coeff[i < 15 ? i : 15];
}

static void add_residual_plane(const PortPlanes *pic, const PortPlanes *res, int c, int x, int y, int w, int h, int bd, int scale)
{
    for (int j = 0; j < h; j++) {
        uint16_t *d = (uint16_t *)(pic->plane[c] + (size_t)(y + j) * pic->stride[c]) + x;
        const int16_t *r = (const int16_t *)(res->plane[c] + (size_t)(y + j) * res->stride[c]) + x;
        for (int i = 0; i < w; i++) {
            int v = r[i];
            if (scale) {                                        /* lmcs_scale_chroma, vvc_intra_template.c:431-449 */
                const int cl = v < -(1 << bd) ? -(1 << bd) : v > (1 << bd) - 1 ? (1 << bd) - 1 : v;
                v = cl > 0 ? (cl * scale + (1 << 10)) >> 11 : -((-cl * scale + (1 << 10)) >> 11);
            }
            d[i] = port_clip_pixel(d[i] + v, bd);
        }
    }
}

/* blocks are in decoding order (CTU raster order, decoding order inside each CTU), so a plain loop honours every
 * dependency the wavefront has to */
PORT_EXPORT int port_intra_picture(const PortPlanes *pic, const PortPlanes *residual, const VVCB200IntraBlock *blocks, int n,
                                   const VVCB200IntraParams *ip)
{
    for (int i = 0; i < n; i++) {
        const VVCB200IntraBlock *b = &blocks[i];
        const int c = b->c_idx;
        const int scale = vpdu_chroma_scale(pic, b, ip);
        if (b->flags & VVCB200_IB_ADD_ONLY) {
            if (b->flags & VVCB200_IB_RESIDUAL)
                add_residual_plane(pic, residual, c, b->x0, b->y0, b->w, b->h, ip->bit_depth, scale);
            continue;
        }
        if (b->mode >= 81) {
            cclm_block(pic, b, ip);
            if (b->flags & VVCB200_IB_RESIDUAL)
                for (int k = 1; k <= 2; k++)
                    if (b->res_planes & k)
                        add_residual_plane(pic, residual, k, b->x0, b->y0, b->w, b->h, ip->bit_depth, scale);
            continue;
        }
        if ((b->flags & VVCB200_IB_ISP) && !c) {
            /* one record per ISP coding block: the sub-partitions are predicted and reconstructed one after the other,
             * each from its predecessor (vvc_intra.c:217-226, 244-252); w x h is the prediction unit */
            const int hor = b->w == b->cb_w, n_parts = hor ? b->cb_h / b->h : b->cb_w / b->w;
            for (int k = 0; k < n_parts; k++) {
                VVCB200IntraBlock s = *b;
                if (hor) {
                    s.y0 += k * b->h;
                    s.avail_left = b->avail_left > k * b->h ? b->avail_left - k * b->h : 0;
                    if (k) { s.avail_top = b->w; s.up_left = b->avail_left > 0; }
                } else {
                    s.x0 += k * b->w;
                    s.avail_top = b->avail_top > k * b->w ? b->avail_top - k * b->w : 0;
                    if (k) { s.avail_left = b->h; s.up_left = b->avail_top > 0; }
                }
                uint16_t *src = (uint16_t *)(pic->plane[0] + (size_t)s.y0 * pic->stride[0]) + s.x0;
                intra_pred_block(src, pic->stride[0] / 2, &s, ip->bit_depth);
                if (b->flags & VVCB200_IB_RESIDUAL)
                    add_residual_plane(pic, residual, 0, s.x0, s.y0, s.w, s.h, ip->bit_depth, 0);
            }
            continue;
        }
        uint16_t *src = (uint16_t *)(pic->plane[c] + (size_t)b->y0 * pic->stride[c]) + b->x0;
        if (b->flags & VVCB200_IB_CIIP) {
            /* the inter prediction is in the picture; blend it with the intra prediction (put_ciip, vvc_inter_template.c:60-76) */
            static uint16_t inter[64 * 64];
            const int stride = pic->stride[c] / 2, wt = b->ciip_weight;
            for (int j = 0; j < b->h; j++)
                for (int i2 = 0; i2 < b->w; i2++)
                    inter[j * b->w + i2] = src[i2 + (ptrdiff_t)stride * j];
            intra_pred_block(src, stride, b, ip->bit_depth);
            for (int j = 0; j < b->h; j++)
                for (int i2 = 0; i2 < b->w; i2++)
                    src[i2 + (ptrdiff_t)stride * j] = (src[i2 + (ptrdiff_t)stride * j] * wt + inter[j * b->w + i2] * (4 - wt) + 2) >> 2;
        } else
        intra_pred_block(src, pic->stride[c] / 2, b, ip->bit_depth);
        if (b->flags & VVCB200_IB_RESIDUAL)
            add_residual_plane(pic, residual, c, b->x0, b->y0, b->w, b->h, ip->bit_depth, scale);
    }
    return 0;
}
