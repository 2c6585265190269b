// deblock.cu — stages DEBLOCK_V and DEBLOCK_H as two picture-level passes.
// Reference semantics: drivers ff_vvc_deblock_vertical/horizontal (libavcodec/vvc/vvc_filter.c:861-1003: tc / beta
// from tctable/betatable, QP averaging, LADF, chroma filter lengths), kernels lf.filter_luma / filter_chroma /
// ladf_level (vvc_filter_template.c:466-804, h26x/h2656_deblock_template.c:25-99).
//
// Mapping: one thread per 4-sample edge segment (the reference's decision granularity).  Picture-wide
// "all vertical edges, then all horizontal edges" equals the reference's CTU-ordered filtering because the
// samples one edge modifies are never read or modified by a neighbouring edge of the same direction
// (lengths are limited by the transform/sub-block sizes on both sides, vvc_filter.c:374-397,418-427).
// Lanes run along x in both passes so every row access of a warp is contiguous.
#include "common.cuh"

namespace vvcb200 {

__constant__ uint16_t c_tctable[66] = {
      0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,
      0,   0,   3,   4,   4,   4,   4,   5,   5,   5,   5,   7,   7,   8,   9,  10,
     10,  11,  13,  14,  15,  17,  19,  21,  24,  25,  29,  33,  36,  41,  45,  51,
     57,  64,  71,  80,  89, 100, 112, 125, 141, 157, 177, 198, 222, 250, 280, 314,
    352, 395,
};
__constant__ uint8_t c_betatable[64] = {
      0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,   0,
      6,   7,   8,   9,  10,  11,  12,  13,  14,  15,  16,  17,  18,  20,  22,  24,
     26,  28,  30,  32,  34,  36,  38,  40,  42,  44,  46,  48,  50,  52,  54,  56,
     58,  60,  62,  64,  66,  68,  70,  72,  74,  76,  78,  80,  82,  84,  86,  88,
};

struct DeblockArgs {
    PlanesView pic;
    const uint8_t *edge[3];
    const int8_t *qp[3];
    const int8_t *ctb_offsets;
    int stride4, rows4, ctb_log2, ctb_width;
    int bd, horizontal;                 // horizontal = 1: horizontal edges (second pass)
    VVCB200DeblockParams p;
};

// Sample accessor of one segment: line l (along the edge), tap i across it (i >= 0: Q side, i < 0: P side).
struct SegRef {
    uint16_t *pix;
    int xs, ys;
    __device__ __forceinline__ int P(int l, int i) const { return pix[-(i + 1) * xs + l * ys]; }
    __device__ __forceinline__ int Q(int l, int i) const { return pix[i * xs + l * ys]; }
    __device__ __forceinline__ void setP(int l, int i, int v) const { pix[-(i + 1) * xs + l * ys] = (uint16_t)v; }
    __device__ __forceinline__ void setQ(int l, int i, int v) const { pix[i * xs + l * ys] = (uint16_t)v; }
};

__device__ __forceinline__ int scale_tc(int tc, int bd)
{
    return bd < 10 ? (tc + (1 << (9 - bd))) >> (10 - bd) : tc << (bd - 10);
}

// loop_filter_luma_large (vvc_filter_template.c:466-552), one line
template <class Seg>
__device__ __forceinline__ void luma_large_line(const Seg &s, int l, int tc, int max_len_p, int max_len_q)
{
    // how far the mean `m` below reaches on each side (the reference reads 7 either way; only these matter)
    int np, nq;
    if (max_len_p == max_len_q)            { np = nq = max_len_p == 5 ? 5 : 7; }
    else if (max_len_p + max_len_q == 12)  { np = nq = 6; }
    else if (max_len_p + max_len_q == 8)   { np = nq = 4; }
    else if (max_len_q == 7)               { np = 3; nq = 7; }
    else                                   { np = 7; nq = 3; }
    np = max(np, max_len_p);
    nq = max(nq, max_len_q);
    int p[7], q[7];
#pragma unroll
    for (int i = 0; i < 7; i++) {
        p[i] = i < np ? s.P(l, i) : 0;
        q[i] = i < nq ? s.Q(l, i) : 0;
    }
    int m;
    if (max_len_p == 5 && max_len_q == 5)
        m = (p[4] + p[3] + 2 * (p[2] + p[1] + p[0] + q[0] + q[1] + q[2]) + q[3] + q[4] + 8) >> 4;
    else if (max_len_p == max_len_q)
        m = (p[6] + p[5] + p[4] + p[3] + p[2] + p[1] + 2 * (p[0] + q[0]) + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + 8) >> 4;
    else if (max_len_p + max_len_q == 12)
        m = (p[5] + p[4] + p[3] + p[2] + 2 * (p[1] + p[0] + q[0] + q[1]) + q[2] + q[3] + q[4] + q[5] + 8) >> 4;
    else if (max_len_p + max_len_q == 8)
        m = (p[3] + p[2] + p[1] + p[0] + q[0] + q[1] + q[2] + q[3] + 4) >> 3;
    else if (max_len_q == 7)
        m = (2 * (p[2] + p[1] + p[0] + q[0]) + p[0] + p[1] + q[1] + q[2] + q[3] + q[4] + q[5] + q[6] + 8) >> 4;
    else
        m = (p[6] + p[5] + p[4] + p[3] + p[2] + p[1] + 2 * (q[2] + q[1] + q[0] + p[0]) + q[0] + q[1] + 8) >> 4;

    // interpolation weights of m (the rest, 64 - w, goes to the far-side reference) and clip factors
    const int w3[3] = { 53, 32, 11 },               c3[3] = { 6, 4, 2 };
    const int w5[5] = { 58, 45, 32, 19, 6 },         c5[5] = { 6, 5, 4, 3, 2 };
    const int w7[7] = { 59, 50, 41, 32, 23, 14, 5 }, c7[7] = { 6, 5, 4, 3, 2, 1, 1 };
    const int refp = (s.P(l, max_len_p) + s.P(l, max_len_p - 1) + 1) >> 1;
    const int refq = (s.Q(l, max_len_q) + s.Q(l, max_len_q - 1) + 1) >> 1;
#pragma unroll
    for (int i = 0; i < 7; i++) {
        if (i < max_len_p) {
            const int w = max_len_p == 3 ? w3[i % 3] : max_len_p == 5 ? w5[i % 5] : w7[i];
            const int c = max_len_p == 3 ? c3[i % 3] : max_len_p == 5 ? c5[i % 5] : c7[i];
            const int lim = (tc * c) >> 1;
            s.setP(l, i, p[i] + clip3(((m * w + refp * (64 - w) + 32) >> 6) - p[i], -lim, lim));
        }
        if (i < max_len_q) {
            const int w = max_len_q == 3 ? w3[i % 3] : max_len_q == 5 ? w5[i % 5] : w7[i];
            const int c = max_len_q == 3 ? c3[i % 3] : max_len_q == 5 ? c5[i % 5] : c7[i];
            const int lim = (tc * c) >> 1;
            s.setQ(l, i, q[i] + clip3(((m * w + refq * (64 - w) + 32) >> 6) - q[i], -lim, lim));
        }
    }
}

// vvc_loop_filter_luma (vvc_filter_template.c:554-652) for one 4-line segment
template <class Seg>
__device__ __forceinline__ void deblock_luma_segment(const Seg &s, int beta_in, int tc_in, int max_len_p, int max_len_q,
                                     int hor_ctu_edge, int bd)
{
    const int tc = scale_tc(tc_in, bd);
    if (!tc)
        return;
    const int pixel_max = (1 << bd) - 1;
    const int p00 = s.P(0, 0), p01 = s.P(0, 1), p02 = s.P(0, 2), q00 = s.Q(0, 0), q01 = s.Q(0, 1), q02 = s.Q(0, 2);
    const int p30 = s.P(3, 0), p31 = s.P(3, 1), p32 = s.P(3, 2), q30 = s.Q(3, 0), q31 = s.Q(3, 1), q32 = s.Q(3, 2);
    const int dp0 = abs(p02 - 2 * p01 + p00), dq0 = abs(q02 - 2 * q01 + q00);
    const int dp3 = abs(p32 - 2 * p31 + p30), dq3 = abs(q32 - 2 * q31 + q30);
    const int d0 = dp0 + dq0, d3 = dp3 + dq3;
    const int tc25 = (tc * 5 + 1) >> 1;
    const int large_p = max_len_p > 3 && !hor_ctu_edge;
    const int large_q = max_len_q > 3;
    const int beta = beta_in << (bd - 8);
    const int beta_3 = beta >> 3, beta_2 = beta >> 2;

    if (large_p || large_q) {
        const int dp0l = large_p ? ((dp0 + abs(s.P(0, 5) - 2 * s.P(0, 4) + s.P(0, 3)) + 1) >> 1) : dp0;
        const int dq0l = large_q ? ((dq0 + abs(s.Q(0, 5) - 2 * s.Q(0, 4) + s.Q(0, 3)) + 1) >> 1) : dq0;
        const int dp3l = large_p ? ((dp3 + abs(s.P(3, 5) - 2 * s.P(3, 4) + s.P(3, 3)) + 1) >> 1) : dp3;
        const int dq3l = large_q ? ((dq3 + abs(s.Q(3, 5) - 2 * s.Q(3, 4) + s.Q(3, 3)) + 1) >> 1) : dq3;
        const int d0l = dp0l + dq0l, d3l = dp3l + dq3l;
        const int beta53 = beta * 3 >> 5, beta_4 = beta >> 4;
        max_len_p = large_p ? max_len_p : 3;        // kept for the fall-through below, as in the reference
        max_len_q = large_q ? max_len_q : 3;
        if (d0l + d3l < beta) {
            const int sp0l = abs(s.P(0, 3) - p00) + (max_len_p == 7 ? abs(s.P(0, 7) - s.P(0, 6) - s.P(0, 5) + s.P(0, 4)) : 0);
            const int sq0l = abs(q00 - s.Q(0, 3)) + (max_len_q == 7 ? abs(s.Q(0, 4) - s.Q(0, 5) - s.Q(0, 6) + s.Q(0, 7)) : 0);
            const int sp3l = abs(s.P(3, 3) - p30) + (max_len_p == 7 ? abs(s.P(3, 7) - s.P(3, 6) - s.P(3, 5) + s.P(3, 4)) : 0);
            const int sq3l = abs(q30 - s.Q(3, 3)) + (max_len_q == 7 ? abs(s.Q(3, 4) - s.Q(3, 5) - s.Q(3, 6) + s.Q(3, 7)) : 0);
            const int sp0 = large_p ? ((sp0l + abs(s.P(0, 3) - s.P(0, max_len_p)) + 1) >> 1) : sp0l;
            const int sp3 = large_p ? ((sp3l + abs(s.P(3, 3) - s.P(3, max_len_p)) + 1) >> 1) : sp3l;
            const int sq0 = large_q ? ((sq0l + abs(s.Q(0, 3) - s.Q(0, max_len_q)) + 1) >> 1) : sq0l;
            const int sq3 = large_q ? ((sq3l + abs(s.Q(3, 3) - s.Q(3, max_len_q)) + 1) >> 1) : sq3l;
            if (sp0 + sq0 < beta53 && abs(p00 - q00) < tc25 && sp3 + sq3 < beta53 && abs(p30 - q30) < tc25 &&
                (d0l << 1) < beta_4 && (d3l << 1) < beta_4) {
                for (int l = 0; l < 4; l++)
                    luma_large_line(s, l, tc, max_len_p, max_len_q);
                return;
            }
        }
    }
    if (d0 + d3 >= beta)
        return;
    if (max_len_p > 2 && max_len_q > 2 &&
        abs(s.P(0, 3) - p00) + abs(s.Q(0, 3) - q00) < beta_3 && abs(p00 - q00) < tc25 &&
        abs(s.P(3, 3) - p30) + abs(s.Q(3, 3) - q30) < beta_3 && abs(p30 - q30) < tc25 &&
        (d0 << 1) < beta_2 && (d3 << 1) < beta_2) {
        // loop_filter_luma_strong (h2656_deblock_template.c:25-51)
        const int tc2 = tc << 1, tc3 = tc * 3;
        for (int l = 0; l < 4; l++) {
            const int p3 = s.P(l, 3), p2 = s.P(l, 2), p1 = s.P(l, 1), p0 = s.P(l, 0);
            const int q0 = s.Q(l, 0), q1 = s.Q(l, 1), q2 = s.Q(l, 2), q3 = s.Q(l, 3);
            s.setP(l, 0, p0 + clip3(((p2 + 2 * p1 + 2 * p0 + 2 * q0 + q1 + 4) >> 3) - p0, -tc3, tc3));
            s.setP(l, 1, p1 + clip3(((p2 + p1 + p0 + q0 + 2) >> 2) - p1, -tc2, tc2));
            s.setP(l, 2, p2 + clip3(((2 * p3 + 3 * p2 + p1 + p0 + q0 + 4) >> 3) - p2, -tc, tc));
            s.setQ(l, 0, q0 + clip3(((p1 + 2 * p0 + 2 * q0 + 2 * q1 + q2 + 4) >> 3) - q0, -tc3, tc3));
            s.setQ(l, 1, q1 + clip3(((p0 + q0 + q1 + q2 + 2) >> 2) - q1, -tc2, tc2));
            s.setQ(l, 2, q2 + clip3(((2 * q3 + 3 * q2 + q1 + q0 + p0 + 4) >> 3) - q2, -tc, tc));
        }
    } else {
        // loop_filter_luma_weak (h2656_deblock_template.c:53-83)
        int nd_p = 1, nd_q = 1;
        if (max_len_p > 1 && max_len_q > 1) {
            if (dp0 + dp3 < ((beta + (beta >> 1)) >> 3))
                nd_p = 2;
            if (dq0 + dq3 < ((beta + (beta >> 1)) >> 3))
                nd_q = 2;
        }
        const int tc_2 = tc >> 1;
        for (int l = 0; l < 4; l++) {
            const int p2 = s.P(l, 2), p1 = s.P(l, 1), p0 = s.P(l, 0), q0 = s.Q(l, 0), q1 = s.Q(l, 1), q2 = s.Q(l, 2);
            int delta0 = (9 * (q0 - p0) - 3 * (q1 - p1) + 8) >> 4;
            if (abs(delta0) < 10 * tc) {
                delta0 = clip3(delta0, -tc, tc);
                s.setP(l, 0, clip3(p0 + delta0, 0, pixel_max));
                s.setQ(l, 0, clip3(q0 - delta0, 0, pixel_max));
                if (nd_p > 1)
                    s.setP(l, 1, clip3(p1 + clip3((((p2 + p0 + 1) >> 1) - p1 + delta0) >> 1, -tc_2, tc_2), 0, pixel_max));
                if (nd_q > 1)
                    s.setQ(l, 1, clip3(q1 + clip3((((q2 + q0 + 1) >> 1) - q1 - delta0) >> 1, -tc_2, tc_2), 0, pixel_max));
            }
        }
    }
}

// vvc_loop_filter_chroma (vvc_filter_template.c:700-771) for one segment of `size` lines
template <class Seg>
__device__ __forceinline__ void deblock_chroma_segment(const Seg &s, int beta_in, int tc_in, int max_len_p, int max_len_q,
                                       int shift, int bd)
{
    const int tc = scale_tc(tc_in, bd);
    if (!tc || !max_len_p || !max_len_q)
        return;
    const int size = shift ? 2 : 4;
    const int n = shift ? 1 : 3;                    // the second decision line
    const int pixel_max = (1 << bd) - 1;
    const int beta = beta_in << (bd - 8);
    const int beta_3 = beta >> 3, beta_2 = beta >> 2;
    const int tc25 = (tc * 5 + 1) >> 1;

    if (max_len_q == 3) {
        const int p1n = s.P(n, 1);
        const int p2n = max_len_p == 1 ? p1n : s.P(n, 2);
        const int p0n = s.P(n, 0), q0n = s.Q(n, 0), q1n = s.Q(n, 1), q2n = s.Q(n, 2);
        const int p3 = max_len_p == 1 ? s.P(0, 1) : s.P(0, 3);
        const int p2 = max_len_p == 1 ? s.P(0, 1) : s.P(0, 2);
        const int p1 = s.P(0, 1), p0 = s.P(0, 0);
        const int dp0 = abs(p2 - 2 * p1 + p0);
        const int dq0 = abs(s.Q(0, 2) - 2 * s.Q(0, 1) + s.Q(0, 0));
        const int dp1 = abs(p2n - 2 * p1n + p0n);
        const int dq1 = abs(q2n - 2 * q1n + q0n);
        const int d0 = dp0 + dq0, d1 = dp1 + dq1;
        if (d0 + d1 < beta) {
            const int p3n = max_len_p == 1 ? p1n : s.P(n, 3);
            const int q3n = s.Q(n, 3);
            const int dsam0 = (d0 << 1) < beta_2 && (abs(p3 - p0) + abs(s.Q(0, 0) - s.Q(0, 3)) < beta_3) && abs(p0 - s.Q(0, 0)) < tc25;
            const int dsam1 = (d1 << 1) < beta_2 && (abs(p3n - p0n) + abs(q0n - q3n) < beta_3) && abs(p0n - q0n) < tc25;
            if (!dsam0 || !dsam1)
                max_len_p = max_len_q = 1;
        } else {
            max_len_p = max_len_q = 1;
        }
    }
    for (int l = 0; l < size; l++) {
        if (max_len_p == 3 && max_len_q == 3) {
            const int p3 = s.P(l, 3), p2 = s.P(l, 2), p1 = s.P(l, 1), p0 = s.P(l, 0);
            const int q0 = s.Q(l, 0), q1 = s.Q(l, 1), q2 = s.Q(l, 2), q3 = s.Q(l, 3);
            s.setP(l, 0, clip3((p3 + p2 + p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3, p0 - tc, p0 + tc));
            s.setP(l, 1, clip3((2 * p3 + p2 + 2 * p1 + p0 + q0 + q1 + 4) >> 3, p1 - tc, p1 + tc));
            s.setP(l, 2, clip3((3 * p3 + 2 * p2 + p1 + p0 + q0 + 4) >> 3, p2 - tc, p2 + tc));
            s.setQ(l, 0, clip3((p2 + p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3, q0 - tc, q0 + tc));
            s.setQ(l, 1, clip3((p1 + p0 + q0 + 2 * q1 + q2 + 2 * q3 + 4) >> 3, q1 - tc, q1 + tc));
            s.setQ(l, 2, clip3((p0 + q0 + q1 + 2 * q2 + 3 * q3 + 4) >> 3, q2 - tc, q2 + tc));
        } else if (max_len_q == 3) {
            const int p1 = s.P(l, 1), p0 = s.P(l, 0), q0 = s.Q(l, 0), q1 = s.Q(l, 1), q2 = s.Q(l, 2), q3 = s.Q(l, 3);
            s.setP(l, 0, clip3((3 * p1 + 2 * p0 + q0 + q1 + q2 + 4) >> 3, p0 - tc, p0 + tc));
            s.setQ(l, 0, clip3((2 * p1 + p0 + 2 * q0 + q1 + q2 + q3 + 4) >> 3, q0 - tc, q0 + tc));
            s.setQ(l, 1, clip3((p1 + p0 + q0 + 2 * q1 + q2 + 2 * q3 + 4) >> 3, q1 - tc, q1 + tc));
            s.setQ(l, 2, clip3((p0 + q0 + q1 + 2 * q2 + 3 * q3 + 4) >> 3, q2 - tc, q2 + tc));
        } else {
            const int p1 = s.P(l, 1), p0 = s.P(l, 0), q0 = s.Q(l, 0), q1 = s.Q(l, 1);
            const int delta0 = clip3((((q0 - p0) * 4) + p1 - q1 + 4) >> 3, -tc, tc);
            s.setP(l, 0, clip3(p0 + delta0, 0, pixel_max));
            s.setQ(l, 0, clip3(q0 - delta0, 0, pixel_max));
        }
    }
}

// grid.x covers the units of one row band, grid.y the unit rows, grid.z the component
#ifndef VVC_DEBLOCK_MIN_BLOCKS
#define VVC_DEBLOCK_MIN_BLOCKS 12     // 40 registers: the kernel waits on memory (59 % long-scoreboard stalls), so warps in
#endif                                // flight beat registers — 12 measured best of 1 / 8 / 10 / 12 / 16 (60 -> 44 us per picture)
__global__ void __launch_bounds__(128, VVC_DEBLOCK_MIN_BLOCKS) deblock_kernel(const DeblockArgs a)
{
    const int c = blockIdx.z;
    const int hor = a.horizontal;
    // candidate edges: luma every unit; chroma every 4th unit across the edge (8 chroma samples)
    int ux = blockIdx.x * blockDim.x + threadIdx.x;
    int uy = blockIdx.y;
    if (c) {
        if (hor) uy *= 4; else ux *= 4;
    }
    if (ux >= a.stride4 || uy >= a.rows4 || (hor ? uy == 0 : ux == 0))
        return;
    const int x = ux << 2, y = uy << 2;                       // luma coordinates of the segment start
    if (x >= a.pic.w[0] || y >= a.pic.h[0])
        return;
    const int e = a.edge[c][uy * a.stride4 + ux];
    const int bs = e & 3;
    if (!bs)
        return;
    int max_len_p = (e >> 2) & 7, max_len_q = (e >> 5) & 7;
    const int hs = c ? 1 : 0;
    SegRef s;
    s.pix = a.pic.p[c] + (size_t)(y >> hs) * a.pic.stride[c] + (x >> hs);
    s.xs = hor ? a.pic.stride[c] : 1;
    s.ys = hor ? 1 : a.pic.stride[c];

    const int pu = hor ? (uy - 1) * a.stride4 + ux : uy * a.stride4 + ux - 1;
    const int qu = uy * a.stride4 + ux;
    int qp;
    if (!c) {
        qp = (a.qp[0][pu] + a.qp[0][qu] + 1) >> 1;
        if (a.p.ladf_enabled) {
            const int level = (s.P(0, 0) + s.P(3, 0) + s.Q(0, 0) + s.Q(3, 0)) >> 2;
            int off = a.p.ladf_lowest_interval_qp_offset;
            for (int i = 0; i < a.p.num_ladf_intervals - 1 && level > a.p.ladf_interval_lower_bound[i + 1]; i++)
                off = a.p.ladf_qp_offset[i];
            qp += off;
        }
    } else {
        qp = (a.qp[c][pu] + a.qp[c][qu] - 2 * 6 * (a.bd - 8) + 1) >> 1;
    }
    const int ctb = (y >> a.ctb_log2) * a.ctb_width + (x >> a.ctb_log2);
    const int beta_offset = a.ctb_offsets[ctb * 6 + c];
    const int tc_offset = a.ctb_offsets[ctb * 6 + 3 + c];
    const int beta = c_betatable[clip3(qp + beta_offset, 0, 63)];
    const int tc = c_tctable[clip3(qp + 2 * (bs - 1) + (tc_offset & -2), 0, 65)];
    if (!c) {
        const int hor_ctu_edge = hor && !(y & ((1 << a.ctb_log2) - 1));
        deblock_luma_segment(s, beta, tc, max_len_p, max_len_q, hor_ctu_edge, a.bd);
    } else {
        deblock_chroma_segment(s, beta, tc, max_len_p, max_len_q, 1, a.bd);
    }
}


// ------------------------------------------------------------------------------------------------------------------
// Tile-resident variant: DEBLOCK_V and DEBLOCK_H in one out-of-place kernel (what Level B runs).
//
// A CTA owns an aligned 64x64 tile of one plane.  It stages the tile plus an apron (8 columns either side: one 128-bit
// chunk; 8 luma / 4 chroma rows) with 128-bit row loads, optionally mapping luma through the inverse LMCS LUT on the way
// in (lmcs.filter is pointwise, so ff_vvc_lmcs_filter folds into this load), filters the vertical edges x0 .. x0 + 64 over
// all staged rows and then the horizontal edges y0 .. y0 + 64 over the tile's columns, both in shared memory, and writes
// the tile once.  Why a 64-aligned tile is closed under this: no transform block crosses a multiple of 64 luma samples
// (max TB 64, ternary splits only below 64) and chroma edges sit on an 8-sample grid with lengths <= 3, so the only edge
// whose filter reaches across a tile side is the edge ON that side — which both neighbours evaluate (the apron gives
// each the samples the other owns).  Inside a pass the reference's own invariant applies: what one edge modifies no
// other edge of that direction reads or modifies (vvc_filter.c:374-397, 418-427), so the edges of a pass are
// independent threads here exactly as they are in deblock_kernel.
//
// Thread mapping: one thread per decision segment (4 luma / 2 chroma lines), as in deblock_kernel — the same segment
// routines run on the staged window.  (One thread per line with the segment's decisions shared by shuffles was measured
// first: 8 segments per warp instead of 32 make every warp walk almost the same union of filter paths, 3x the warp
// instructions.)
constexpr int kDT = 64;            // tile side, samples of the plane
constexpr int kDCA = 8;            // column apron: one 128-bit chunk
constexpr int kDThreads = 352;      // one round per pass for a luma tile: 17 edges x 20 segments = 340 threads
constexpr int kDStride = 88;       // staged row stride in samples: 176 B = 11 x 16 B, so that the 128-bit rows of 8
                                   // consecutive lines fall into different bank groups

struct DeblockTileArgs {
    PlanesView dst, src;
    const uint8_t *edge[2][3];
    const int8_t *qp[3];
    const int8_t *ctb_offsets;
    const uint16_t *lut;           // inverse LMCS LUT (1 << bd entries) applied to luma while staging, or nullptr
    int stride4, rows4, ctb_log2, ctb_width, bd;
    VVCB200DeblockParams p;
};

// Accessor of one segment in the staged window: compile-time steps across the edge (XS) and along it (YS).
template <int XS, int YS>
struct SegWin {
    uint16_t *pix;
    __device__ __forceinline__ int P(int l, int i) const { return pix[-(i + 1) * XS + l * YS]; }
    __device__ __forceinline__ int Q(int l, int i) const { return pix[i * XS + l * YS]; }
    __device__ __forceinline__ void setP(int l, int i, int v) const { pix[-(i + 1) * XS + l * YS] = (uint16_t)v; }
    __device__ __forceinline__ void setQ(int l, int i, int v) const { pix[i * XS + l * YS] = (uint16_t)v; }
};

// One pass over the staged window: one thread per decision segment (4 luma / 2 chroma lines), exactly the routines of the
// picture-level kernel on shared memory.  Consecutive lanes take consecutive edges of one segment row (V) or
// consecutive segments of one edge (H), which spreads a warp's 16-bit accesses over the banks.
template <bool CHROMA, bool HOR>
__device__ __forceinline__ void tile_edge_pass(uint16_t *win, const DeblockTileArgs &a, int c, int x0, int y0)
{
    constexpr int ES = CHROMA ? 8 : 4;          // edge spacing in samples of this plane
    constexpr int SEG = CHROMA ? 2 : 4;         // lines per segment
    constexpr int AR = CHROMA ? 4 : 8;          // row apron
    constexpr int NE = kDT / ES + 1;            // edges per pass, both tile sides included
    constexpr int NSEG = (HOR ? kDT : kDT + 2 * AR) / SEG;
    constexpr int N = NE * NSEG;
    const int W = a.src.w[c], H = a.src.h[c];
    const int hs = CHROMA ? 1 : 0;
    for (int item = threadIdx.x; item < N; item += blockDim.x) {
        int e, sg;
        if (HOR) { e = item / NSEG; sg = item - e * NSEG; }
        else     { sg = item / NE; e = item - sg * NE; }
        const int wx = HOR ? kDCA + sg * SEG : kDCA + e * ES;   // window coordinates of line 0's q0
        const int wy = HOR ? AR + e * ES : sg * SEG;
        const int gx = x0 - kDCA + wx, gy = y0 - AR + wy;       // plane coordinates
        if (gx >= W || gy >= H || gy < 0 || (HOR ? gy == 0 : gx == 0))
            continue;
        const int ux = (gx << hs) >> 2, uy = (gy << hs) >> 2;
        const int eb = a.edge[HOR][c][uy * a.stride4 + ux];
        const int bs = eb & 3;
        if (!bs)
            continue;
        const int max_len_p = (eb >> 2) & 7, max_len_q = (eb >> 5) & 7;
        SegWin<HOR ? kDStride : 1, HOR ? 1 : kDStride> s;
        s.pix = win + wy * kDStride + wx;
        const int qu = uy * a.stride4 + ux;
        const int pu = HOR ? qu - a.stride4 : qu - 1;
        const int ctb = ((gy << hs) >> a.ctb_log2) * a.ctb_width + ((gx << hs) >> a.ctb_log2);
        int qp;
        if (!CHROMA) {
            qp = (a.qp[0][pu] + a.qp[0][qu] + 1) >> 1;
            if (a.p.ladf_enabled) {
                const int level = (s.P(0, 0) + s.P(3, 0) + s.Q(0, 0) + s.Q(3, 0)) >> 2;
                int off = a.p.ladf_lowest_interval_qp_offset;
                for (int i = 0; i < a.p.num_ladf_intervals - 1 && level > a.p.ladf_interval_lower_bound[i + 1]; i++)
                    off = a.p.ladf_qp_offset[i];
                qp += off;
            }
        } else {
            qp = (a.qp[c][pu] + a.qp[c][qu] - 2 * 6 * (a.bd - 8) + 1) >> 1;
        }
        const int beta_offset = a.ctb_offsets[ctb * 6 + c];
        const int tc_offset = a.ctb_offsets[ctb * 6 + 3 + c];
        const int beta = c_betatable[clip3(qp + beta_offset, 0, 63)];
        const int tc = c_tctable[clip3(qp + 2 * (bs - 1) + (tc_offset & -2), 0, 65)];
        if (!CHROMA) {
            const int hor_ctu_edge = HOR && !(gy & ((1 << a.ctb_log2) - 1));
            deblock_luma_segment(s, beta, tc, max_len_p, max_len_q, hor_ctu_edge, a.bd);
        } else {
            deblock_chroma_segment(s, beta, tc, max_len_p, max_len_q, 1, a.bd);
        }
    }
}

template <bool CHROMA, bool LUT>
__global__ void __launch_bounds__(kDThreads, 3) deblock_tile_kernel(const DeblockTileArgs a)
{
    constexpr int AR = CHROMA ? 4 : 8;
    constexpr int WR = kDT + 2 * AR;
    __shared__ __align__(16) uint16_t win[WR * kDStride];
    __shared__ uint16_t s_lut[LUT ? 4096 : 1];
    const int c = CHROMA ? 1 + blockIdx.z : 0;
    const int x0 = blockIdx.x * kDT, y0 = blockIdx.y * kDT;
    const int W = a.src.w[c], H = a.src.h[c];
    const int t = threadIdx.x;
    if (LUT) {
        for (int i = t; i < (1 << a.bd); i += kDThreads)
            s_lut[i] = a.lut[i];
        __syncthreads();
    }
    // ---- stage: 128-bit chunks; anything outside the plane's rows / pitch reads as zero and is never used
    {
        const uint16_t *src = a.src.p[c];
        const int stride = a.src.stride[c];
        constexpr int CH = (kDT + 2 * kDCA) / 8;                 // 10 chunks per row
        for (int i = t; i < WR * CH; i += kDThreads) {
            const int r = i / CH, k = i - r * CH;
            const int gx = x0 - kDCA + 8 * k, gy = y0 - AR + r;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (gy >= 0 && gy < H && gx >= 0 && gx + 8 <= stride)
                v = *reinterpret_cast<const uint4 *>(src + (size_t)gy * stride + gx);
            if (LUT) {
                uint32_t *q = reinterpret_cast<uint32_t *>(&v);
                const uint32_t mask = (1u << a.bd) - 1;
#pragma unroll
                for (int j = 0; j < 4; j++)
                    q[j] = (uint32_t)s_lut[q[j] & mask] | ((uint32_t)s_lut[(q[j] >> 16) & mask] << 16);
            }
            *reinterpret_cast<uint4 *>(win + r * kDStride + 8 * k) = v;
        }
    }
    __syncthreads();
    tile_edge_pass<CHROMA, false>(win, a, c, x0, y0);
    __syncthreads();
    tile_edge_pass<CHROMA, true>(win, a, c, x0, y0);
    __syncthreads();
    {
        uint16_t *dst = a.dst.p[c];
        const int stride = a.dst.stride[c];
        for (int i = t; i < kDT * (kDT / 8); i += kDThreads) {
            const int r = i >> 3, k = i & 7;
            const int gx = x0 + 8 * k, gy = y0 + r;
            if (gy >= H || gx >= W)
                continue;
            const uint16_t *s = win + (AR + r) * kDStride + kDCA + 8 * k;
            if (gx + 8 <= W)
                *reinterpret_cast<uint4 *>(dst + (size_t)gy * stride + gx) = *reinterpret_cast<const uint4 *>(s);
            else
                for (int j = 0; gx + j < W; j++)
                    dst[(size_t)gy * stride + gx + j] = s[j];
        }
    }
}


// ------------------------------------------------------------------------------------------------------------------
// Boundary strengths / filter lengths on the device (vvcb200_deblock_derive_edges): one thread per luma 4x4 unit derives
// the unit's vertical-edge and horizontal-edge bytes of all three planes.  Every rule cites the reference it restates.
struct BsArgs {
    uint8_t *out[2][3];
    const VVCB200MotionUnit *motion;
    const VVCB200PartitionUnit *part;
    const uint8_t *ctb_edges;
    int width, height, stride4, rows4, ctb_log2, ctb_width;
};

// boundary_strength(), vvc_filter.c:308-372; reference pictures compared through ref_id
__device__ __forceinline__ int motion_bs(const VVCB200MotionUnit &c, const VVCB200MotionUnit &n)
{
    auto far = [](const int32_t (&a)[2], const int32_t (&b)[2]) { return abs(a[0] - b[0]) >= 8 || abs(a[1] - b[1]) >= 8; };
    if (c.pred_flag == 3 && n.pred_flag == 3) {
        if (c.ref_id[0] == n.ref_id[0] && c.ref_id[0] == c.ref_id[1] && n.ref_id[0] == n.ref_id[1])
            return (far(n.mv[0], c.mv[0]) || far(n.mv[1], c.mv[1])) && (far(n.mv[1], c.mv[0]) || far(n.mv[0], c.mv[1]));
        if (n.ref_id[0] == c.ref_id[0] && n.ref_id[1] == c.ref_id[1])
            return far(n.mv[0], c.mv[0]) || far(n.mv[1], c.mv[1]);
        if (n.ref_id[1] == c.ref_id[0] && n.ref_id[0] == c.ref_id[1])
            return far(n.mv[1], c.mv[0]) || far(n.mv[0], c.mv[1]);
        return 1;
    }
    if (c.pred_flag != 3 && n.pred_flag != 3) {
        const int lc = (c.pred_flag & 1) ? 0 : 1, ln = (n.pred_flag & 1) ? 0 : 1;
        if (c.ref_id[lc] != n.ref_id[ln])
            return 1;
        return far(c.mv[lc], n.mv[ln]);
    }
    return 1;
}

__global__ void __launch_bounds__(256) deblock_bs_kernel(const BsArgs a)
{
    const int ux = blockIdx.x * blockDim.x + threadIdx.x, uy = blockIdx.y;
    if (ux >= a.stride4 || uy >= a.rows4)
        return;
    const int x = ux << 2, y = uy << 2, u = uy * a.stride4 + ux;
    uint8_t res[2][3] = { { 0, 0, 0 }, { 0, 0, 0 } };
    if (x < a.width && y < a.height) {
        const VVCB200PartitionUnit q = a.part[u];
        const VVCB200MotionUnit mq = a.motion[u];
        const int ctb_mask = (1 << a.ctb_log2) - 1;
        const int ce = a.ctb_edges ? a.ctb_edges[(y >> a.ctb_log2) * a.ctb_width + (x >> a.ctb_log2)] : 0;
        const bool sb_cu = mq.pred_flag != 0 && (q.cu_off_x & 32);            // !is_intra && (msf || iaf), vvc_filter.c:497-500
#pragma unroll
        for (int dir = 0; dir < 2; dir++) {                                   // 0: vertical edge at x, 1: horizontal edge at y
            const int pos = dir ? y : x;
            if (pos == 0)
                continue;
            const int pu = dir ? u - a.stride4 : u - 1;
            const VVCB200PartitionUnit p = a.part[pu];
            const VVCB200MotionUnit mp = a.motion[pu];
            // a slice / tile limit on this CTB side that must not be filtered across (vvc_filter.c:505-512, 566-573)
            const bool barred = !(pos & ctb_mask) && (ce & (dir ? 2 : 1));
            const bool intra_side = mq.pred_flag == 0 || mp.pred_flag == 0 || mq.ciip_flag || mp.ciip_flag;
            // ---- luma
            const int tb_off = dir ? q.tb_off >> 4 : q.tb_off & 15;           // offset of this unit in its TB, units of 4
            const int cu_off = dir ? q.cu_off_y : q.cu_off_x & 31;            // offset in its coding unit
            const int log2_q = dir ? (q.tb_size >> 3) & 7 : q.tb_size & 7, log2_p = dir ? (p.tb_size >> 3) & 7 : p.tb_size & 7;
            const bool has_sb = sb_cu && (q.cu_off_x & (dir ? 128 : 64));     // cb_width / cb_height > 8
            if (tb_off == 0) {
                if (!barred) {                                                // transform edge, vvc_filter.c:514-556
                    int bs;
                    if ((q.tb_size & 128) && (p.tb_size & 128))
                        bs = 0;
                    else if (intra_side)
                        bs = 2;
                    else if ((q.tb_size & 64) || (p.tb_size & 64))
                        bs = 1;
                    else if (cu_off && ((cu_off & 1) || !has_sb))             // inside a CU, not on the 8-sample grid or no sub-blocks
                        bs = 0;
                    else
                        bs = motion_bs(mq, mp);
                    // derive_max_filter_length_luma, vvc_filter.c:374-397
                    int lp, lq;
                    if (log2_p <= 2 || log2_q <= 2) {
                        lp = lq = 1;
                    } else {
                        lp = log2_p >= 5 ? 7 : 3;
                        lq = log2_q >= 5 ? 7 : 3;
                    }
                    if (has_sb)
                        lq = min(5, lq);
                    if (p.cu_off_x & 32)
                        lp = min(5, lp);
                    res[dir][0] = (uint8_t)(bs | (lp << 2) | (lq << 5));
                }
            } else if (sb_cu && !(cu_off & 1)) {
                // edge of the 8x8 sub-block grid inside a TB (vvc_deblock_subblock_bs_*, vvc_filter.c:399-475)
                const int i = tb_off << 2, size = 1 << log2_q;
                const int len = (i == 4 || i == size - 4) ? 1 : (i == 8 || i == size - 8) ? 2 : 3;
                res[dir][0] = (uint8_t)(motion_bs(mq, mp) | (len << 2) | (len << 5));
            }
            // ---- chroma: edges of the chroma transform blocks on the 8-chroma-sample grid (vvc_filter.c:646-749)
            const int ctb_off = dir ? q.ctb_off >> 4 : q.ctb_off & 15;
            if (ctb_off == 0 && !(pos & 15) && !barred) {
                const bool pcm = (q.cflags & 2) && (p.cflags & 2);
                const bool joint = (q.cflags & 1) || (p.cflags & 1);
                const int sz_q = dir ? (q.ctb_size >> 3) & 7 : q.ctb_size & 7, sz_p = dir ? (p.ctb_size >> 3) & 7 : p.ctb_size & 7;
                const bool big = sz_p >= 3 && sz_q >= 3;                      // max_filter_length_chroma, vvc_filter.c:796-817
                const bool ctu_edge = dir && !(y & ctb_mask);
#pragma unroll
                for (int c = 1; c < 3; c++) {
                    const bool cbf = ((q.ctb_size | p.ctb_size) & (c == 1 ? 64 : 128)) || joint;
                    const int bs = pcm ? 0 : intra_side ? 2 : cbf ? 1 : 0;
                    const int lq = big ? 3 : bs == 2, lp = big ? (ctu_edge ? 1 : 3) : bs == 2;
                    res[dir][c] = (uint8_t)(bs | (lp << 2) | (lq << 5));
                }
            }
        }
    }
    for (int dir = 0; dir < 2; dir++)
        for (int c = 0; c < 3; c++)
            a.out[dir][c][u] = res[dir][c];
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_deblock_picture(const VVCB200Planes *pic, const VVCB200DeblockTables *t,
                                       const VVCB200DeblockParams *p, void *stream)
{
    if (!pic || !t || !p || t->stride4 <= 0 || !t->d_ctb_offsets) {
        set_error("vvcb200_deblock_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (p->bit_depth != 10 && p->bit_depth != 12) {
        set_error("unsupported bit depth %d (10 or 12: 8-bit pictures are uint8 planes in the reference, which are not built)", p->bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (p->ctb_log2_size < 5 || p->ctb_log2_size > 7 || p->num_ladf_intervals > 5) {
        set_error("vvcb200_deblock_picture: ctb_log2_size %d / ladf intervals %d out of range", p->ctb_log2_size, p->num_ladf_intervals);
        return VVCB200_ERR_ARG;
    }
    DeblockArgs a = {};
    a.pic = make_view(pic);
    a.stride4 = t->stride4;
    a.rows4 = (pic->height[0] + 3) >> 2;
    a.ctb_log2 = p->ctb_log2_size;
    a.ctb_width = (pic->width[0] + (1 << p->ctb_log2_size) - 1) >> p->ctb_log2_size;
    a.bd = p->bit_depth;
    a.ctb_offsets = t->d_ctb_offsets;
    a.p = *p;
    for (int c = 0; c < 3; c++)
        a.qp[c] = t->d_qp[c];
    const int units_x = (pic->width[0] + 3) >> 2;
    for (int dir = 0; dir < 2; dir++) {
        for (int c = 0; c < 3; c++)
            a.edge[c] = t->d_edge[dir][c];
        a.horizontal = dir;
        dim3 block(128);
        dim3 grid((units_x + 127) / 128, a.rows4, 3);
        deblock_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(a);
        VVC_LAUNCH_CHECK();
    }
    return 0;
}


extern "C" int vvcb200_deblock_lmcs_picture(const VVCB200Planes *dst, const VVCB200Planes *src,
                                            const VVCB200DeblockTables *t, const VVCB200DeblockParams *p,
                                            const uint16_t *d_inv_lut, void *stream)
{
    return vvcb200_deblock_lmcs_planes(dst, src, t, p, d_inv_lut, 3, stream);
}

extern "C" int vvcb200_deblock_lmcs_planes(const VVCB200Planes *dst, const VVCB200Planes *src,
                                           const VVCB200DeblockTables *t, const VVCB200DeblockParams *p,
                                           const uint16_t *d_inv_lut, int planes, void *stream)
{
    if (!dst || !src || !t || !p || t->stride4 <= 0 || !t->d_ctb_offsets || planes < 1 || planes > 3) {
        set_error("vvcb200_deblock_lmcs_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (p->bit_depth != 10 && p->bit_depth != 12) {
        set_error("unsupported bit depth %d (10 or 12)", p->bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (p->ctb_log2_size < 5 || p->ctb_log2_size > 7 || p->num_ladf_intervals > 5) {
        set_error("vvcb200_deblock_lmcs_picture: ctb_log2_size %d / ladf intervals %d out of range", p->ctb_log2_size, p->num_ladf_intervals);
        return VVCB200_ERR_ARG;
    }
    DeblockTileArgs a = {};
    a.dst = make_view(dst);
    a.src = make_view(src);
    for (int c = 0; c < 3; c++) {
        if (((uintptr_t)a.dst.p[c] | (uintptr_t)a.src.p[c]) & 15 || ((a.dst.stride[c] | a.src.stride[c]) & 7) ||
            a.dst.w[c] != a.src.w[c] || a.dst.h[c] != a.src.h[c] || a.dst.p[c] == a.src.p[c]) {
            set_error("vvcb200_deblock_lmcs_picture: planes must be distinct, equally sized, 16-byte aligned with 16-byte multiple strides");
            return VVCB200_ERR_ARG;
        }
    }
    a.stride4 = t->stride4;
    a.rows4 = (src->height[0] + 3) >> 2;
    a.ctb_log2 = p->ctb_log2_size;
    a.ctb_width = (src->width[0] + (1 << p->ctb_log2_size) - 1) >> p->ctb_log2_size;
    a.bd = p->bit_depth;
    a.ctb_offsets = t->d_ctb_offsets;
    a.lut = d_inv_lut;
    a.p = *p;
    for (int c = 0; c < 3; c++) {
        a.qp[c] = t->d_qp[c];
        a.edge[0][c] = t->d_edge[0][c];
        a.edge[1][c] = t->d_edge[1][c];
    }
    // luma and the two chroma planes are independent grids: side by side
    const cudaStream_t main_stream = (cudaStream_t)stream;
    ForkStreams *f = planes == 3 ? fork_streams(main_stream) : nullptr;
    cudaStream_t side = main_stream;
    if (f) {
        VVC_CUDA(cudaEventRecord(f->fork, main_stream));
        side = f->side[0];
        VVC_CUDA(cudaStreamWaitEvent(side, f->fork, 0));
    }
    if (planes & 1) {
        dim3 gl((src->width[0] + kDT - 1) / kDT, (src->height[0] + kDT - 1) / kDT, 1);
        if (d_inv_lut)
            deblock_tile_kernel<false, true><<<gl, kDThreads, 0, main_stream>>>(a);
        else
            deblock_tile_kernel<false, false><<<gl, kDThreads, 0, main_stream>>>(a);
        VVC_LAUNCH_CHECK();
    }
    if (planes & 2) {
        dim3 gc((src->width[1] + kDT - 1) / kDT, (src->height[1] + kDT - 1) / kDT, 2);
        deblock_tile_kernel<true, false><<<gc, kDThreads, 0, side>>>(a);
        VVC_LAUNCH_CHECK();
    }
    if (f) {
        VVC_CUDA(cudaEventRecord(f->join[0], side));
        VVC_CUDA(cudaStreamWaitEvent(main_stream, f->join[0], 0));
    }
    return 0;
}


extern "C" int vvcb200_deblock_derive_edges(uint8_t *const d_edge_out[2][3], const VVCB200MotionUnit *d_motion,
                                            const VVCB200PartitionUnit *d_part, const uint8_t *d_ctb_edges, int width,
                                            int height, int stride4, int ctb_log2_size, void *stream)
{
    if (!d_edge_out || !d_motion || !d_part || width <= 0 || height <= 0 || stride4 < (width + 3) / 4 || ctb_log2_size < 5 ||
        ctb_log2_size > 7) {
        set_error("vvcb200_deblock_derive_edges: bad argument");
        return VVCB200_ERR_ARG;
    }
    BsArgs a;
    for (int d = 0; d < 2; d++)
        for (int c = 0; c < 3; c++) {
            if (!d_edge_out[d][c]) {
                set_error("vvcb200_deblock_derive_edges: missing output table");
                return VVCB200_ERR_ARG;
            }
            a.out[d][c] = d_edge_out[d][c];
        }
    a.motion = d_motion;
    a.part = d_part;
    a.ctb_edges = d_ctb_edges;
    a.width = width;
    a.height = height;
    a.stride4 = stride4;
    a.rows4 = (height + 3) >> 2;
    a.ctb_log2 = ctb_log2_size;
    a.ctb_width = (width + (1 << ctb_log2_size) - 1) >> ctb_log2_size;
    dim3 grid((stride4 + 255) / 256, a.rows4);
    deblock_bs_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    VVC_LAUNCH_CHECK();
    return 0;
}
