/*
 * oracle/port/port_deblock.c — TEST INFRASTRUCTURE (see port_common.h).
 * Restatement of the reference's deblocking filter for a whole picture:
 *   drivers: tc/beta derivation, QP averaging, LADF      libavcodec/vvc/vvc_filter.c:38-52, 819-1003
 *   luma decisions + weak/strong/long filters            libavcodec/vvc/vvc_filter_template.c:466-652,
 *                                                         libavcodec/h26x/h2656_deblock_template.c:25-83
 *   chroma decisions + filters                            vvc_filter_template.c:654-771, h2656_deblock_template.c:85-99
 * Input tables are the packed per-4x4 edge bytes of include/vvc_b200.h (bs | lenP << 2 | lenQ << 5).
 * Edges are visited in raster order, vertical edges of the whole picture first, then horizontal ones.
 */
#include <stdlib.h>
#include "port_common.h"

static const uint16_t tc_tab[66] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 3, 4, 4, 4, 4, 5, 5, 5, 5, 7, 7, 8, 9, 10,
    10, 11, 13, 14, 15, 17, 19, 21, 24, 25, 29, 33, 36, 41, 45, 51, 57, 64, 71, 80, 89, 100, 112, 125,
    141, 157, 177, 198, 222, 250, 280, 314, 352, 395,
};
static const uint8_t beta_tab[64] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 20, 22, 24,
    26, 28, 30, 32, 34, 36, 38, 40, 42, 44, 46, 48, 50, 52, 54, 56, 58, 60, 62, 64, 66, 68, 70, 72, 74, 76, 78,
    80, 82, 84, 86, 88,
};

typedef struct Seg { uint16_t *pix; ptrdiff_t xs, ys; } Seg;
#define P_(l, i) s->pix[-((i) + 1) * s->xs + (l) * s->ys]
#define Q_(l, i) s->pix[(i) * s->xs + (l) * s->ys]

static int scale_tc(int tc, int bd) { return bd < 10 ? (tc + (1 << (9 - bd))) >> (10 - bd) : tc << (bd - 10); }

static void luma_long(const Seg *s, int tc, int lp, int lq)
{
    static const int8_t w3[] = { 53, 32, 11 }, w5[] = { 58, 45, 32, 19, 6 }, w7[] = { 59, 50, 41, 32, 23, 14, 5 };
    static const int8_t c3[] = { 6, 4, 2 }, c5[] = { 6, 5, 4, 3, 2 }, c7[] = { 6, 5, 4, 3, 2, 1, 1 };
    const int8_t *wp = lp == 3 ? w3 : lp == 5 ? w5 : w7, *cp = lp == 3 ? c3 : lp == 5 ? c5 : c7;
    const int8_t *wq = lq == 3 ? w3 : lq == 5 ? w5 : w7, *cq = lq == 3 ? c3 : lq == 5 ? c5 : c7;
    for (int l = 0; l < 4; l++) {
        int p[7], q[7], m;
        for (int i = 0; i < 7; i++) {
            p[i] = P_(l, i);
            q[i] = Q_(l, i);
        }
        if (lp == 5 && lq == 5)
            m = (p[4] + p[3] + 2 * (p[2] + p[1] + p[0] + q[0] + q[1] + q[2]) + q[3] + q[4] + 8) >> 4;
        else if (lp == lq)
            m = (p[6] + p[5] + p[4] + p[3] + p[2] + p[1] + 2 * (p[0] + q[0]) + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + 8) >> 4;
        else if (lp + lq == 12)
            m = (p[5] + p[4] + p[3] + p[2] + 2 * (p[1] + p[0] + q[0] + q[1]) + q[2] + q[3] + q[4] + q[5] + 8) >> 4;
        else if (lp + lq == 8)
            m = (p[3] + p[2] + p[1] + p[0] + q[0] + q[1] + q[2] + q[3] + 4) >> 3;
        else if (lq == 7)
            m = (2 * (p[2] + p[1] + p[0] + q[0]) + p[0] + p[1] + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + 8) >> 4;
        else
            m = (p[6] + p[5] + p[4] + p[3] + p[2] + p[1] + 2 * (q[2] + q[1] + q[0] + p[0]) + q[0] + q[1] + 8) >> 4;
        const int refp = (P_(l, lp) + P_(l, lp - 1) + 1) >> 1;
        const int refq = (Q_(l, lq) + Q_(l, lq - 1) + 1) >> 1;
        for (int i = 0; i < lp; i++) {
            const int lim = (tc * cp[i]) >> 1;
            P_(l, i) = p[i] + port_clip(((m * wp[i] + refp * (64 - wp[i]) + 32) >> 6) - p[i], -lim, lim);
        }
        for (int i = 0; i < lq; i++) {
            const int lim = (tc * cq[i]) >> 1;
            Q_(l, i) = q[i] + port_clip(((m * wq[i] + refq * (64 - wq[i]) + 32) >> 6) - q[i], -lim, lim);
        }
    }
}

static void luma_segment(const Seg *s, int beta_in, int tc_in, int lp, int lq, int hor_ctu_edge, int bd)
{
    const int tc = scale_tc(tc_in, bd);
    if (!tc)
        return;
    const int dp0 = abs(P_(0, 2) - 2 * P_(0, 1) + P_(0, 0)), dq0 = abs(Q_(0, 2) - 2 * Q_(0, 1) + Q_(0, 0));
    const int dp3 = abs(P_(3, 2) - 2 * P_(3, 1) + P_(3, 0)), dq3 = abs(Q_(3, 2) - 2 * Q_(3, 1) + Q_(3, 0));
    const int d0 = dp0 + dq0, d3 = dp3 + dq3;
    const int tc25 = (tc * 5 + 1) >> 1;
    const int large_p = lp > 3 && !hor_ctu_edge, large_q = lq > 3;
    const int beta = beta_in << (bd - 8);

    if (large_p || large_q) {
        const int dp0l = large_p ? (dp0 + abs(P_(0, 5) - 2 * P_(0, 4) + P_(0, 3)) + 1) >> 1 : dp0;
        const int dq0l = large_q ? (dq0 + abs(Q_(0, 5) - 2 * Q_(0, 4) + Q_(0, 3)) + 1) >> 1 : dq0;
        const int dp3l = large_p ? (dp3 + abs(P_(3, 5) - 2 * P_(3, 4) + P_(3, 3)) + 1) >> 1 : dp3;
        const int dq3l = large_q ? (dq3 + abs(Q_(3, 5) - 2 * Q_(3, 4) + Q_(3, 3)) + 1) >> 1 : dq3;
        const int d0l = dp0l + dq0l, d3l = dp3l + dq3l;
        lp = large_p ? lp : 3;
        lq = large_q ? lq : 3;
        if (d0l + d3l < beta) {
            int sp[2], sq[2], ok = 1;
            for (int k = 0; k < 2; k++) {
                const int l = k * 3;
                sp[k] = abs(P_(l, 3) - P_(l, 0)) + (lp == 7 ? abs(P_(l, 7) - P_(l, 6) - P_(l, 5) + P_(l, 4)) : 0);
                sq[k] = abs(Q_(l, 0) - Q_(l, 3)) + (lq == 7 ? abs(Q_(l, 4) - Q_(l, 5) - Q_(l, 6) + Q_(l, 7)) : 0);
                if (large_p)
                    sp[k] = (sp[k] + abs(P_(l, 3) - P_(l, lp)) + 1) >> 1;
                if (large_q)
                    sq[k] = (sq[k] + abs(Q_(l, 3) - Q_(l, lq)) + 1) >> 1;
                ok = ok && sp[k] + sq[k] < (beta * 3 >> 5) && abs(P_(l, 0) - Q_(l, 0)) < tc25;
            }
            if (ok && (d0l << 1) < (beta >> 4) && (d3l << 1) < (beta >> 4)) {
                luma_long(s, tc, lp, lq);
                return;
            }
        }
    }
    if (d0 + d3 >= beta)
        return;
    if (lp > 2 && lq > 2 &&
        abs(P_(0, 3) - P_(0, 0)) + abs(Q_(0, 3) - Q_(0, 0)) < (beta >> 3) && abs(P_(0, 0) - Q_(0, 0)) < tc25 &&
        abs(P_(3, 3) - P_(3, 0)) + abs(Q_(3, 3) - Q_(3, 0)) < (beta >> 3) && abs(P_(3, 0) - Q_(3, 0)) < tc25 &&
        (d0 << 1) < (beta >> 2) && (d3 << 1) < (beta >> 2)) {
        for (int l = 0; l < 4; l++) {
            const int p3 = P_(l, 3), p2 = P_(l, 2), p1 = P_(l, 1), p0 = P_(l, 0);
            const int q0 = Q_(l, 0), q1 = Q_(l, 1), q2 = Q_(l, 2), q3 = Q_(l, 3);
            P_(l, 0) = p0 + port_clip(((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3) - p0, -3 * tc, 3 * tc);
            P_(l, 1) = p1 + port_clip(((p2 + p1 + p0 + q0 + 2) >> 2) - p1, -2 * tc, 2 * tc);
            P_(l, 2) = p2 + port_clip(((2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3) - p2, -tc, tc);
            Q_(l, 0) = q0 + port_clip(((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3) - q0, -3 * tc, 3 * tc);
            Q_(l, 1) = q1 + port_clip(((p0 + q0 + q1 + q2 + 2) >> 2) - q1, -2 * tc, 2 * tc);
            Q_(l, 2) = q2 + port_clip(((2 * q3 + 3 * q2 + q1 + q0 + p0 + 4) >> 3) - q2, -tc, tc);
        }
    } else {
        const int side_thr = (beta + (beta >> 1)) >> 3;
        const int two_p = lp > 1 && lq > 1 && dp0 + dp3 < side_thr;
        const int two_q = lp > 1 && lq > 1 && dq0 + dq3 < side_thr;
        for (int l = 0; l < 4; l++) {
            const int p2 = P_(l, 2), p1 = P_(l, 1), p0 = P_(l, 0), q0 = Q_(l, 0), q1 = Q_(l, 1), q2 = Q_(l, 2);
            int delta = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
            if (abs(delta) >= 10 * tc)
                continue;
            delta = port_clip(delta, -tc, tc);
            P_(l, 0) = port_clip_pixel(p0 + delta, bd);
            Q_(l, 0) = port_clip_pixel(q0 - delta, bd);
            if (two_p)
                P_(l, 1) = port_clip_pixel(p1 + port_clip((((p2 + p0 + 1) >> 1) - p1 + delta) >> 1, -(tc >> 1), tc >> 1), bd);
            if (two_q)
                Q_(l, 1) = port_clip_pixel(q1 + port_clip((((q2 + q0 + 1) >> 1) - q1 - delta) >> 1, -(tc >> 1), tc >> 1), bd);
        }
    }
}

static void chroma_segment(const Seg *s, int beta_in, int tc_in, int lp, int lq, int shift, int bd)
{
    const int tc = scale_tc(tc_in, bd);
    const int size = shift ? 2 : 4, n = shift ? 1 : 3;
    const int beta = beta_in << (bd - 8);
    const int tc25 = (tc * 5 + 1) >> 1;
    if (!tc || !lp || !lq)
        return;
    if (lq == 3) {
        int d[2], dsam[2];
        for (int k = 0; k < 2; k++) {
            const int l = k ? n : 0;
            const int p1 = P_(l, 1), p0 = P_(l, 0);
            const int p2 = lp == 1 ? p1 : P_(l, 2), p3 = lp == 1 ? p1 : P_(l, 3);
            d[k] = abs(p2 - 2 * p1 + p0) + abs(Q_(l, 2) - 2 * Q_(l, 1) + Q_(l, 0));
            dsam[k] = (d[k] << 1) < (beta >> 2) && abs(p3 - p0) + abs(Q_(l, 0) - Q_(l, 3)) < (beta >> 3) &&
                      abs(p0 - Q_(l, 0)) < tc25;
        }
        if (!(d[0] + d[1] < beta && dsam[0] && dsam[1]))
            lp = lq = 1;
    }
    for (int l = 0; l < size; l++) {
        const int p3 = P_(l, 3), p2 = P_(l, 2), p1 = P_(l, 1), p0 = P_(l, 0);
        const int q0 = Q_(l, 0), q1 = Q_(l, 1), q2 = Q_(l, 2), q3 = Q_(l, 3);
        if (lp == 3 && lq == 3) {
            P_(l, 0) = port_clip((p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3, p0 - tc, p0 + tc);
            P_(l, 1) = port_clip((2 * p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3, p1 - tc, p1 + tc);
            P_(l, 2) = port_clip((3 * p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3, p2 - tc, p2 + tc);
        } else if (lq == 3) {
            P_(l, 0) = port_clip((3 * p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3, p0 - tc, p0 + tc);
        }
        if (lq == 3) {
            const int a = lp == 3 ? p2 + p1 + p0 : 2 * p1 + p0;       /* P-side taps of the Q0 filter */
            Q_(l, 0) = port_clip((a + 2 * q0 + q1 + q2 + q3 + 4) >> 3, q0 - tc, q0 + tc);
            Q_(l, 1) = port_clip((p1 + p0 + q0 + 2 * q1 + q2 + 2 * q3 + 4) >> 3, q1 - tc, q1 + tc);
            Q_(l, 2) = port_clip((p0 + q0 + q1 + 2 * q2 + 3 * q3 + 4) >> 3, q2 - tc, q2 + tc);
        } else {
            const int delta = port_clip((((q0 - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
            P_(l, 0) = port_clip_pixel(p0 + delta, bd);
            Q_(l, 0) = port_clip_pixel(q0 - delta, bd);
        }
    }
}

/* host pointers in the same struct layout as VVCB200DeblockTables */
PORT_EXPORT int port_deblock_picture(const PortPlanes *pic, const VVCB200DeblockTables *t, const VVCB200DeblockParams *p)
{
    const int bd = p->bit_depth, W = pic->width[0], H = pic->height[0];
    const int ctb_width = (W + (1 << p->ctb_log2_size) - 1) >> p->ctb_log2_size;
    for (int dir = 0; dir < 2; dir++) {
        const int hor = dir;
        for (int c = 0; c < 3; c++) {
            const int hs = c ? 1 : 0;
            const int grid = c ? 16 : 4;
            const ptrdiff_t stride = pic->stride[c] / 2;
            for (int y = hor ? grid : 0; y < H; y += hor ? grid : 4) {
                for (int x = hor ? 0 : grid; x < W; x += hor ? 4 : grid) {
                    const int u = (y >> 2) * t->stride4 + (x >> 2);
                    const int pu = hor ? u - t->stride4 : u - 1;
                    const int e = t->d_edge[dir][c][u];
                    const int bs = e & 3, lp = (e >> 2) & 7, lq = (e >> 5) & 7;
                    Seg s;
                    int qp;
                    if (!bs)
                        continue;
                    s.pix = (uint16_t *)(pic->plane[c] + (size_t)(y >> hs) * pic->stride[c]) + (x >> hs);
                    s.xs = hor ? stride : 1;
                    s.ys = hor ? 1 : stride;
                    if (!c) {
                        qp = (t->d_qp[0][pu] + t->d_qp[0][u] + 1) >> 1;
                        if (p->ladf_enabled) {
                            const Seg *sp = &s; const Seg *s = sp;
                            const int level = (P_(0, 0) + P_(3, 0) + Q_(0, 0) + Q_(3, 0)) >> 2;
                            int off = p->ladf_lowest_interval_qp_offset;
                            for (int i = 0; i < p->num_ladf_intervals - 1 && level > p->ladf_interval_lower_bound[i + 1]; i++)
                                off = p->ladf_qp_offset[i];
                            qp += off;
                        }
                    } else {
                        qp = (t->d_qp[c][pu] + t->d_qp[c][u] - 2 * 6 * (bd - 8) + 1) >> 1;
                    }
                    const int ctb = (y >> p->ctb_log2_size) * ctb_width + (x >> p->ctb_log2_size);
                    const int beta = beta_tab[port_clip(qp + t->d_ctb_offsets[ctb * 6 + c], 0, 63)];
                    const int tc = tc_tab[port_clip(qp + 2 * (bs - 1) + (t->d_ctb_offsets[ctb * 6 + 3 + c] & -2), 0, 65)];
                    if (!c)
                        luma_segment(&s, beta, tc, lp, lq, hor && !(y & ((1 << p->ctb_log2_size) - 1)), bd);
                    else
                        chroma_segment(&s, beta, tc, lp, lq, 1, bd);
                }
            }
        }
    }
    return 0;
}
