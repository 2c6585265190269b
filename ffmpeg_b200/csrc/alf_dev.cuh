// alf_dev.cuh — device code of the ALF stage shared by the stand-alone ALF kernels (alf.cu) and the fused SAO + ALF
// kernel (saoalf.cu): CTB geometry / padding rules, virtual-boundary row rules, and the luma tile pipeline after staging
// (gradients -> class -> coefficients -> 7x7 filter).  Reference: vvc_filter.c:1005-1319, vvc_filter_template.c:38-408.
#pragma once
#include "common.cuh"

#ifndef VVC_TABLE_QUAL
#define VVC_TABLE_QUAL static __device__ const
#endif
#include "vvc_tables.h"

namespace vvcb200 {

constexpr int kTile   = 32;
constexpr int kHalo   = 3;
constexpr int kColOrg = 4;                       // staged column of the tile's first sample (even: sample pairs are 32-bit aligned)
constexpr int kPad    = kTile + 2 * kHalo;       // 38
constexpr int kPadStr = kPad + 2;                // 40 samples per staged row
constexpr int kCells  = kTile / 2 + 2;           // 18 gradient cells per tile side

struct AlfArgs {
    PlanesView dst, src;
    const VVCB200AlfCtb *ctbs;
    const VVCB200AlfParams *params;
    int ctb_log2, ctb_width, bd;
    const uint32_t *table;       // alf_coef_table_kernel's output (kTable instantiations), or nullptr
};

// Luma coefficients and clip values for every (filter set, class, transposition), derived once per picture instead of once
// per 4x4 block (alf_recon_coeff_and_clip, vvc_filter_template.c:365-408): word [32 + ((set * 25 + cls) * 4 + tr) * 12 + j] =
// coefficient (int16, low half) | clip value (high half); word [set] != 0: some coefficient of the set does not fit int8.
constexpr int kAlfSets = 23;                              // 16 fixed sets + up to 7 APS
constexpr int kAlfTableHead = 32;
constexpr int kAlfTableWords = kAlfTableHead + kAlfSets * 25 * 4 * 12;

__device__ __forceinline__ void alf_coef_entry(const VVCB200AlfParams *params, int bd, int set, int cls, int tr, int j, int &cf, int &clip)
{
    const int16_t *coeff;
    const uint8_t *clip_idx = nullptr;
    if (set < 16) {
        coeff = VVC_ALF_FIX_FILT_COEFF + VVC_ALF_CLASS_TO_FILT_MAP[set * 25 + cls] * 12;
    } else {
        const int k = min(set - 16, 6);
        coeff = params->luma_coeff[k][VVC_ALF_APS_CLASS_TO_FILT_MAP[cls]];
        clip_idx = params->luma_clip_idx[k][cls];
    }
    // transposition tables of alf_recon_coeff_and_clip (vvc_filter_template.c:386-391), 4 bits per entry
    const unsigned long long i1 = 0x62037b518a49ull, i2 = 0xba9456781230ull, i3 = 0x62015b734a89ull;
    const int idx = tr == 0 ? j : (int)(((tr == 1 ? i1 : tr == 2 ? i2 : i3) >> (4 * j)) & 15);
    const int ci = clip_idx ? clip_idx[idx] : 0;
    cf = coeff[idx];
    clip = 1 << (bd - (ci == 0 ? 0 : ci == 1 ? 3 : ci == 2 ? 5 : 7));
}

static __global__ void alf_coef_table_kernel(const VVCB200AlfParams *params, int bd, uint32_t *table)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= kAlfSets * 25 * 4 * 12)
        return;
    const int j = i % 12, tr = (i / 12) & 3, cls = (i / 48) % 25, set = i / (48 * 25);
    int cf, clip;
    alf_coef_entry(params, bd, set, cls, tr, j, cf, clip);
    if (cf < -128 || cf > 127)
        table[set] = 1;                                   // (the head was cleared by the caller)
    table[kAlfTableHead + i] = (uint32_t)(cf & 0xffff) | ((uint32_t)clip << 16);
}

struct CtbGeom {
    int x0, y0, w, h;      // CTB rectangle in the component's plane
    int edges;             // bit0 L, bit1 T, bit2 R, bit3 B incl. picture borders
};

__device__ __forceinline__ CtbGeom ctb_geom(const AlfArgs &a, int c, int x, int y, int edges)
{
    CtbGeom g;
    const int l2 = a.ctb_log2 - (c ? 1 : 0);
    g.x0 = (x >> l2) << l2;
    g.y0 = (y >> l2) << l2;
    g.w = min(1 << l2, a.src.w[c] - g.x0);
    g.h = min(1 << l2, a.src.h[c] - g.y0);
    g.edges = edges | (g.x0 == 0 ? 1 : 0) | (g.y0 == 0 ? 2 : 0) |
              (g.x0 + g.w >= a.src.w[c] ? 4 : 0) | (g.y0 + g.h >= a.src.h[c] ? 8 : 0);
    return g;
}

// alf_prepare_buffer as a coordinate clamp: where CTB g finds the sample it reads at plane position (x, y)
__device__ __forceinline__ void pad_coords(const CtbGeom &g, int &x, int &y)
{
    if (x < g.x0 && (g.edges & 1)) x = g.x0;
    if (x >= g.x0 + g.w && (g.edges & 4)) x = g.x0 + g.w - 1;
    if (y < g.y0 && (g.edges & 2)) y = g.y0;
    if (y >= g.y0 + g.h && (g.edges & 8)) y = g.y0 + g.h - 1;
}

__device__ __forceinline__ int padded_sample(const uint16_t *plane, int stride, const CtbGeom &g, int x, int y)
{
    pad_coords(g, x, y);
    return plane[(size_t)y * stride + x];
}

// row offsets of the taps one / two / three rows away for a sample at CTB row yy (vvc_filter_template.c:83-101)
__device__ __forceinline__ void vb_rows(int yy, int vb, int reach, int &o1, int &o2, int &o3, bool &near_vb)
{
    o1 = 1; o2 = 2; o3 = 3;
    if (yy < vb && yy >= vb - reach - 1) {
        o1 = yy == vb - 1 ? 0 : 1;
        o2 = yy >= vb - 2 ? o1 : 2;
        o3 = yy >= vb - 3 ? o2 : 3;
    } else if (yy >= vb && yy <= vb + reach) {
        o1 = yy == vb ? 0 : 1;
        o2 = yy <= vb + 1 ? o1 : 2;
        o3 = yy <= vb + 2 ? o2 : 3;
    }
    near_vb = yy == vb - 1 || yy == vb;
}

__device__ __forceinline__ int alf_clip2(int curr, int v0, int v1, int c)
{
    return clip3(v0 - curr, -c, c) + clip3(v1 - curr, -c, c);
}

struct AlfLumaShared {
    __align__(16) uint16_t pix[kPad * kPadStr];          // the tile + halo as its CTB sees it (alf_prepare_buffer)
    __align__(8) uint16_t grad[kCells * kCells][4];
    int16_t  coef[64][12];
    int16_t  clip[64][12];
    uint8_t  cls[64];
    int wide;                                            // some coefficient of the tile does not fit int8 (never in a conforming stream)
};

// Everything after staging for one 32x32 luma tile at (tx0, ty0) of CTB `g` (record `cp`): s_pix (row stride kStr, 32-bit
// aligned) holds the padded window — sample (tx0 + c - kColOrg, ty0 + r - kHalo) at s_pix[r * kStr + c] — which is S.pix
// after alf_prepare_buffer's clamping, or any window of the pre-ALF samples when the CTB has no edge to clamp at;
// S.wide was cleared, and a __syncthreads() lies between those writes and this call.  256 threads.
template <bool kTable, int kStr = kPadStr>
__device__ __forceinline__ void alf_luma_filter_tile(const AlfArgs &a, const VVCB200AlfCtb &cp, const CtbGeom &g, int tx0,
                                                     int ty0, AlfLumaShared &S, const uint16_t *s_pix)
{
    uint16_t (*const s_grad)[4] = S.grad;
    int16_t (*const s_coef)[12] = S.coef;
    int16_t (*const s_clip)[12] = S.clip;
    uint8_t *const s_cls = S.cls;
    int &s_wide = S.wide;
    uint16_t *dst = a.dst.p[0];
    const int dstride = a.dst.stride[0];
    const int t = threadIdx.x;
    const int vb = (1 << a.ctb_log2) - 4;                // ALF_VB_POS_ABOVE_LUMA, CTB-relative (vvc_filter.c:1304)
    const int ty_ctb = ty0 - g.y0;                       // tile origin inside the CTB
    const int tw = min(kTile, g.x0 + g.w - tx0), th = min(kTile, g.y0 + g.h - ty0);

    // ---- gradients per 2x2 cell (alf_classify first loop); cell (gy, gx) has p0 at tile (2gy-2, 2gx-2)
    for (int i = t; i < kCells * kCells; i += 256) {
        const int gy = i / kCells, gx = i - gy * kCells;
        const int r0 = 2 * gy - 2 + kHalo, c0 = 2 * gx - 2 + kColOrg;   // staged coordinates of p0
        const int ygrad = ty_ctb + 2 * gy;                                // the reference's loop variable y
        int ra = r0 - 1, rb = r0, rc = r0 + 1, rd = r0 + 2;               // s0..s3
        if (ygrad == vb)
            rd = rc;
        else if (ygrad == vb + 2)
            ra = rb;
        const uint16_t *s0 = s_pix + ra * kStr, *s1 = s_pix + rb * kStr;
        const uint16_t *s2 = s_pix + rc * kStr, *s3 = s_pix + rd * kStr;
        const int val0 = s1[c0] << 1, val1 = s2[c0 + 1] << 1;
        s_grad[i][0] = (uint16_t)(abs(val0 - s0[c0] - s2[c0]) + abs(val1 - s1[c0 + 1] - s3[c0 + 1]));
        s_grad[i][1] = (uint16_t)(abs(val0 - s1[c0 - 1] - s1[c0 + 1]) + abs(val1 - s2[c0] - s2[c0 + 2]));
        s_grad[i][2] = (uint16_t)(abs(val0 - s0[c0 - 1] - s2[c0 + 1]) + abs(val1 - s1[c0] - s3[c0 + 2]));
        s_grad[i][3] = (uint16_t)(abs(val0 - s0[c0 + 1] - s2[c0 - 1]) + abs(val1 - s1[c0 + 2] - s3[c0]));
    }
    __syncthreads();

    // ---- class / transpose per 4x4 block, then its 12 coefficients and clips (alf_get_idx, recon_coeff_and_clip)
    {                                                    // four threads per block, one per column of gradient cells
        const int blk = t >> 2, j = t & 3;
        const int bx = blk & 7, by = blk >> 3;
        const int yb = ty_ctb + by * 4;
        int start = 0, end = 4, ac = 2;
        if (yb + 4 == vb) { end = 3; ac = 3; }
        else if (yb == vb) { start = 1; ac = 3; }
        int sum[4] = { 0, 0, 0, 0 };
        for (int i = start; i < end; i++) {
            const uint2 c = *reinterpret_cast<const uint2 *>(s_grad[(by * 2 + i) * kCells + bx * 2 + j]);
            sum[0] += c.x & 0xffff; sum[1] += c.x >> 16; sum[2] += c.y & 0xffff; sum[3] += c.y >> 16;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            sum[k] += __shfl_xor_sync(0xffffffffu, sum[k], 1);
            sum[k] += __shfl_xor_sync(0xffffffffu, sum[k], 2);
        }
        const int dir_hv = sum[0] <= sum[1];
        const int hv1 = max(sum[0], sum[1]), hv0 = min(sum[0], sum[1]);
        const int dir_d = sum[2] <= sum[3];
        const int d1 = max(sum[2], sum[3]), d0 = min(sum[2], sum[3]);
        const int dir1 = (unsigned long long)d1 * hv0 <= (unsigned long long)hv1 * d0;
        const int hvd1 = dir1 ? hv1 : d1, hvd0 = dir1 ? hv0 : d0;
        const int act = min(max(((sum[0] + sum[1]) * ac) >> (a.bd - 1), 0), 15);
        int cls = act == 0 ? 0 : act == 1 ? 1 : act < 7 ? 2 : act < 15 ? 3 : 4;
        if (hvd1 * 2 > 9 * hvd0)
            cls += ((dir1 << 1) + 2) * 5;
        else if (hvd1 > 2 * hvd0)
            cls += ((dir1 << 1) + 1) * 5;
        if (j == 0)
            s_cls[blk] = (uint8_t)(cls | ((dir_d * 2 + dir_hv) << 5));
    }
    __syncthreads();
    if (!kTable) {
        // the 64 x 12 coefficient / clip fetches (global tables) spread over the whole CTA instead of 12 in a row on 64 threads
        for (int i = t; i < 64 * 12; i += 256) {
            const int blk = i / 12, j = i - blk * 12;
            int cf, clip;
            alf_coef_entry(a.params, a.bd, cp.filt_set_idx_y, s_cls[blk] & 31, s_cls[blk] >> 5, j, cf, clip);
            if (cf < -128 || cf > 127)
                s_wide = 1;
            s_coef[blk][j] = (int16_t)cf;
            s_clip[blk][j] = (int16_t)clip;
        }
        __syncthreads();
    }

    // ---- filter: each thread owns 4 horizontally adjacent samples of one 4x4 block
    {
        const int row = t >> 3, xg = t & 7;
        if (row < th && xg * 4 < tw) {
            const int blk = (row >> 2) * 8 + xg;
            int f[12], cl[12];
            bool wide;
            if (kTable) {                                 // the block's row of the per-picture table: three 128-bit loads
                const int set = min((int)cp.filt_set_idx_y, kAlfSets - 1);
                const uint4 *e = reinterpret_cast<const uint4 *>(a.table + kAlfTableHead +
                                                                ((set * 25 + (s_cls[blk] & 31)) * 4 + (s_cls[blk] >> 5)) * 12);
                const uint4 e0 = __ldg(e), e1 = __ldg(e + 1), e2 = __ldg(e + 2);
                const uint32_t ew[12] = { e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w, e2.x, e2.y, e2.z, e2.w };
#pragma unroll
                for (int j = 0; j < 12; j++) {
                    f[j] = (int16_t)(ew[j] & 0xffff);
                    cl[j] = (int)(ew[j] >> 16);
                }
                wide = __ldg(a.table + set) != 0;
            } else {
#pragma unroll
                for (int j = 0; j < 12; j++) {
                    f[j] = s_coef[blk][j];
                    cl[j] = s_clip[blk][j];
                }
                wide = s_wide != 0;
            }
            int o1, o2, o3;
            bool near_vb;
            vb_rows(ty_ctb + row, vb, 3, o1, o2, o3, near_vb);
            const uint16_t *p0 = s_pix + (row + kHalo) * kStr + xg * 4 + kColOrg;
            const uint16_t *p1 = p0 + o1 * kStr, *p2 = p0 - o1 * kStr;
            const uint16_t *p3 = p0 + o2 * kStr, *p4 = p0 - o2 * kStr;
            const uint16_t *p5 = p0 + o3 * kStr, *p6 = p0 - o3 * kStr;
            uint16_t out[4];
            if (!wide) {
                // Two samples per 32-bit word through the 16x2 integer pipe (VIADD.16x2 / VIADDMNMX.S16x2 on sm_100a), the
                // tap products through IDP.2A.  Differences, clips and their sums stay inside int16 for bit depths <= 12.
                // w(p, k) = staged word k of a row relative to this thread's first sample: samples (2k, 2k + 1).
                auto w = [](const uint16_t *p, int k) { return reinterpret_cast<const uint32_t *>(p)[k]; };
                // pair starting at sample s of the row: aligned word for even s, funnel shift of two words for odd s
                auto pr = [&](const uint16_t *p, int s0) {
                    return (s0 & 1) ? __funnelshift_r(w(p, (s0 - 1) >> 1), w(p, (s0 + 1) >> 1), 16) : w(p, s0 >> 1);
                };
                const uint32_t cA = w(p0, 0), cB = w(p0, 1);               // the four current samples
                int sum[4] = { 0, 0, 0, 0 };
                auto tap = [&](int k, const uint16_t *pa, int da, const uint16_t *pb, int db) {
                    const uint32_t c2 = (uint32_t)cl[k] * 0x10001u, n2 = __vneg2(c2);
                    const uint32_t tA = __vadd2(__vmaxs2(__vmins2(__vsub2(pr(pa, da), cA), c2), n2),
                                                __vmaxs2(__vmins2(__vsub2(pr(pb, db), cA), c2), n2));
                    const uint32_t tB = __vadd2(__vmaxs2(__vmins2(__vsub2(pr(pa, da + 2), cB), c2), n2),
                                                __vmaxs2(__vmins2(__vsub2(pr(pb, db + 2), cB), c2), n2));
                    const int flo = f[k] & 0xff, fhi = flo << 8;
                    sum[0] = __dp2a_lo((int)tA, flo, sum[0]);
                    sum[1] = __dp2a_lo((int)tA, fhi, sum[1]);
                    sum[2] = __dp2a_lo((int)tB, flo, sum[2]);
                    sum[3] = __dp2a_lo((int)tB, fhi, sum[3]);
                };
                tap(0, p5, 0, p6, 0);
                tap(1, p3, 1, p4, -1);
                tap(2, p3, 0, p4, 0);
                tap(3, p3, -1, p4, 1);
                tap(4, p1, 2, p2, -2);
                tap(5, p1, 1, p2, -1);
                tap(6, p1, 0, p2, 0);
                tap(7, p1, -1, p2, 1);
                tap(8, p1, -2, p2, 2);
                tap(9, p0, 3, p0, -3);
                tap(10, p0, 2, p0, -2);
                tap(11, p0, 1, p0, -1);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int curr = p0[j];
                    const int sv = near_vb ? (sum[j] + 512) >> 10 : (sum[j] + 64) >> 7;
                    out[j] = (uint16_t)clip_pixel(sv + curr, a.bd);
                }
            } else
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int curr = p0[j];
                int sum = 0;
                sum += f[0]  * alf_clip2(curr, p5[j],     p6[j],     cl[0]);
                sum += f[1]  * alf_clip2(curr, p3[j + 1], p4[j - 1], cl[1]);
                sum += f[2]  * alf_clip2(curr, p3[j],     p4[j],     cl[2]);
                sum += f[3]  * alf_clip2(curr, p3[j - 1], p4[j + 1], cl[3]);
                sum += f[4]  * alf_clip2(curr, p1[j + 2], p2[j - 2], cl[4]);
                sum += f[5]  * alf_clip2(curr, p1[j + 1], p2[j - 1], cl[5]);
                sum += f[6]  * alf_clip2(curr, p1[j],     p2[j],     cl[6]);
                sum += f[7]  * alf_clip2(curr, p1[j - 1], p2[j + 1], cl[7]);
                sum += f[8]  * alf_clip2(curr, p1[j - 2], p2[j + 2], cl[8]);
                sum += f[9]  * alf_clip2(curr, p0[j + 3], p0[j - 3], cl[9]);
                sum += f[10] * alf_clip2(curr, p0[j + 2], p0[j - 2], cl[10]);
                sum += f[11] * alf_clip2(curr, p0[j + 1], p0[j - 1], cl[11]);
                sum = near_vb ? (sum + 512) >> 10 : (sum + 64) >> 7;
                out[j] = (uint16_t)clip_pixel(sum + curr, a.bd);
            }
            uint16_t *d = dst + (size_t)(ty0 + row) * dstride + tx0 + xg * 4;
            *reinterpret_cast<uint2 *>(d) = make_uint2(out[0] | ((uint32_t)out[1] << 16), out[2] | ((uint32_t)out[3] << 16));
        }
    }
}

// One chroma tile of side T (<= 32, never straddling a chroma CTB) at (tx0, ty0) of component c: 5x5 ALF from the padded
// (T + 4)^2 window in s_pix, row stride S (staged by the caller when cp.ctb_flag[c] is set, __syncthreads() in between;
// or a window of the pre-ALF samples when the CTB has no edge to clamp at) and CC-ALF
// (alf_filter_cc, vvc_filter_template.c:223-263).  cur(x, y): the pre-ALF sample of this plane; luma(x, y): the pre-ALF
// luma sample at a plane position (already padded for the luma CTB by this function).  256 threads.
template <class CurAt, class LumaAt>
__device__ __forceinline__ void alf_chroma_filter_tile(const AlfArgs &a, const VVCB200AlfCtb &cp, const CtbGeom &g, int c,
                                                       int T, int tx0, int ty0, const uint16_t *s_pix, int S, CurAt cur,
                                                       LumaAt luma)
{
    const int W = a.src.w[c], H = a.src.h[c];
    const int l2 = a.ctb_log2 - 1;
    uint16_t *dst = a.dst.p[c];
    const int dstride = a.dst.stride[c];
    const int t = threadIdx.x;
    const bool do_alf = cp.ctb_flag[c];
    const int cc = cp.cc_idc[c - 1];
    const bool do_cc = cc && a.params->cc_enabled[c - 1];
    const int ty_ctb = ty0 - g.y0;
    int f[6], cl[6];
    bool wide = false;                                   // a coefficient outside int8: never in a conforming stream
    if (do_alf) {
        const int alt = min((int)cp.alt_idx[c - 1], 7);
#pragma unroll
        for (int j = 0; j < 6; j++) {
            const int ci = a.params->chroma_clip_idx[alt][j];
            f[j] = a.params->chroma_coeff[alt][j];
            cl[j] = 1 << (a.bd - (ci == 0 ? 0 : ci == 1 ? 3 : ci == 2 ? 5 : 7));
            wide = wide || f[j] < -128 || f[j] > 127;
        }
    }
    int fc[7];
    CtbGeom gl = g;
    if (do_cc) {
#pragma unroll
        for (int j = 0; j < 7; j++)
            fc[j] = a.params->cc_coeff[c - 1][min(cc - 1, 3)][j];
        gl = ctb_geom(a, 0, tx0 * 2, ty0 * 2, cp.edges);
    }
    const int vb = (1 << l2) - 2;                        // ALF_VB_POS_ABOVE_CHROMA
    const int vb_luma = (1 << a.ctb_log2) - 4;
    const int half = T >> 1, lh = 31 - __clz(half);      // T is 16 or 32; a thread owns two horizontally adjacent samples
    const int pmax = (1 << a.bd) - 1;
    for (int i = t; i < T * half; i += 256) {
        const int row = i >> lh, col = (i & (half - 1)) * 2;
        const int x = tx0 + col, y = ty0 + row;
        if (x >= W || y >= H)                            // (plane widths are even: a pair is inside or outside as one)
            continue;
        int v0, v1;
        if (do_alf) {
            int o1, o2, o3;
            bool near_vb;
            vb_rows(ty_ctb + row, vb, 1, o1, o2, o3, near_vb);
            const uint16_t *p0 = s_pix + (row + 2) * S + col + 2;
            const uint16_t *p1 = p0 + o1 * S, *p2 = p0 - o1 * S, *p3 = p0 + o2 * S, *p4 = p0 - o2 * S;
            int sum0 = 0, sum1 = 0;
            const int c0 = p0[0], c1 = p0[1];
            if (!wide) {
                // the pair through the 16x2 integer pipe, as in the luma filter: differences, clips and their sums stay in int16
                auto w = [](const uint16_t *p, int k) { return reinterpret_cast<const uint32_t *>(p)[k]; };
                auto pr = [&](const uint16_t *p, int s0) {
                    return (s0 & 1) ? __funnelshift_r(w(p, (s0 - 1) >> 1), w(p, (s0 + 1) >> 1), 16) : w(p, s0 >> 1);
                };
                const uint32_t cw = w(p0, 0);
                auto tap = [&](int k, const uint16_t *pa, int da, const uint16_t *pb, int db) {
                    const uint32_t c2 = (uint32_t)cl[k] * 0x10001u, n2 = __vneg2(c2);
                    const uint32_t tA = __vadd2(__vmaxs2(__vmins2(__vsub2(pr(pa, da), cw), c2), n2),
                                                __vmaxs2(__vmins2(__vsub2(pr(pb, db), cw), c2), n2));
                    const int flo = f[k] & 0xff;
                    sum0 = __dp2a_lo((int)tA, flo, sum0);
                    sum1 = __dp2a_lo((int)tA, flo << 8, sum1);
                };
                tap(0, p3, 0, p4, 0);
                tap(1, p1, 1, p2, -1);
                tap(2, p1, 0, p2, 0);
                tap(3, p1, -1, p2, 1);
                tap(4, p0, 2, p0, -2);
                tap(5, p0, 1, p0, -1);
            } else {
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const int curr = p0[j];
                    int sum = 0;
                    sum += f[0] * alf_clip2(curr, p3[j],     p4[j],     cl[0]);
                    sum += f[1] * alf_clip2(curr, p1[j + 1], p2[j - 1], cl[1]);
                    sum += f[2] * alf_clip2(curr, p1[j],     p2[j],     cl[2]);
                    sum += f[3] * alf_clip2(curr, p1[j - 1], p2[j + 1], cl[3]);
                    sum += f[4] * alf_clip2(curr, p0[j + 2], p0[j - 2], cl[4]);
                    sum += f[5] * alf_clip2(curr, p0[j + 1], p0[j - 1], cl[5]);
                    if (j == 0) sum0 = sum; else sum1 = sum;
                }
            }
            v0 = min(max((near_vb ? (sum0 + 512) >> 10 : (sum0 + 64) >> 7) + c0, 0), pmax);
            v1 = min(max((near_vb ? (sum1 + 512) >> 10 : (sum1 + 64) >> 7) + c1, 0), pmax);
        } else {
            v0 = cur(x, y);
            v1 = cur(x + 1, y);
        }
        if (do_cc) {
            // 4:2:0: chroma (x, y) sits at luma (2x, 2y); rows by the virtual-boundary rule, then alf_prepare_buffer's clamp —
            // separable, and the centre column / row is always inside the CTB
            const int ly = y * 2;
            const int pos = (ty_ctb + row) * 2;
            int r0 = ly - 1, r2 = ly + 1, r3 = ly + 2;
            if (pos == vb_luma - 2 || pos == vb_luma + 1)
                r3 = r2;
            else if (pos == vb_luma - 1 || pos == vb_luma)
                r3 = r2 = r0 = ly;
            int xs[5] = { 2 * x - 1, 2 * x, 2 * x + 1, 2 * x + 2, 2 * x + 3 };
            if (gl.edges) {
                pad_coords(gl, xs[0], r0);                // (xs[1..3] lie between two samples of the CTB: never clamped)
                pad_coords(gl, xs[2], r2);
                pad_coords(gl, xs[4], r3);
            }
            const int half_range = 1 << (a.bd - 1);
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int xm = xs[2 * j], xc = xs[2 * j + 1], xp = xs[2 * j + 2];
                const int cu = luma(xc, ly);
                int sum = fc[0] * (luma(xc, r0) - cu) + fc[1] * (luma(xm, ly) - cu) + fc[2] * (luma(xp, ly) - cu) +
                          fc[3] * (luma(xm, r2) - cu) + fc[4] * (luma(xc, r2) - cu) + fc[5] * (luma(xp, r2) - cu) +
                          fc[6] * (luma(xc, r3) - cu);
                sum = clip3((sum + 64) >> 7, -half_range, half_range - 1);
                if (j == 0) v0 = min(max(v0 + sum, 0), pmax); else v1 = min(max(v1 + sum, 0), pmax);
            }
        }
        *reinterpret_cast<uint32_t *>(dst + (size_t)y * dstride + x) = (uint32_t)v0 | ((uint32_t)v1 << 16);
    }
}

}  // namespace vvcb200
