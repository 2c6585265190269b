// inter.cu — stage INTER: motion-compensated prediction of a picture's sub-block batch.
// Reference semantics: 8-tap luma / 4-tap chroma separable interpolation (inter.put / put_uni / put_uni_w,
// libavcodec/h26x/h2656_inter_template.c:29-577), avg / w_avg incl. BCW (vvc_inter_template.c:25-58,
// vvc_inter.c:31,145-174), MV -> offset/fraction arithmetic of vvc_inter.c:176-368, replicate-edge fetch of
// ff_emulated_edge_mc (videodsp_template.c:27-105, here a coordinate clamp), forward LMCS (vvc_inter.c:888-891).
//
// Mapping.  The batch is ordered by block width (16, 8, 4).  A warp takes 16 luma columns' worth of blocks of
// one width class: one 16-wide, two 8-wide or four 4-wide blocks, each owned by a lane group of 2*w lanes;
// the two chroma planes of a block are then handled by the two halves of that group.  Per list the
// (h+7)x(w+7) reference window is staged once in shared memory with clamped coordinates.  Both filter passes
// use __dp2a on sample pairs: two adjacent 16-bit samples sit in one 32-bit word and meet a pair of int8 taps,
// so an 8-tap output costs 4 (even phase) or 5 (odd phase) instructions instead of 8 multiply-adds.  The
// horizontal pass writes its 16-bit rows transposed so that the vertical pass finds vertical neighbours in one
// word as well.  The two lists meet in registers; every reference sample is fetched from L2/HBM once per
// block and the prediction is written once.
//
// The four reference variants (pixels / h / v / hv) are folded into one separable form, which is
// bit-identical: a zero fraction selects the identity tap {..,64,..}, and (64*s) >> (bd-8) == s << (14-bd),
// (64*t) >> 6 == t exactly.  All sums are exact in int32 (|taps| sum < 128, samples < 2^12, rows < 2^15).
#include "common.cuh"

#define VVC_TABLE_QUAL static __device__ const
#define VVC_TABLES_WITH_GPM_WEIGHTS
#include "vvc_tables.h"

namespace vvcb200 {

constexpr int kInterWarps = 8;
constexpr int kMaxRefs    = 16;
constexpr int kWinElems   = 1288 + 24;          // staged windows of one warp (u16), largest class + slack
constexpr int kTmpElems   = 16 * 26 + 16;       // transposed horizontal-pass rows of one warp (int16)
constexpr int kBufElems   = kWinElems + kTmpElems;   // 1744: also exactly what BDOF needs (2 ringed blocks + 4 gradients)

struct InterArgs {
    PlanesView dst;
    PlanesView tmp14;                            // optional int16 output of the list-0 intermediate
    PlanesView refs[kMaxRefs];
    const VVCB200PredBlock *blocks;
    const uint16_t *lmcs_fwd;
    int32_t *refined_mv;                         // optional: the vectors actually used (DMVR feedback), n x 2 x 2
    const VVCB200ProfTable *prof;                // optional: PROF offset tables of the affine coding units
    int n, n16, n8;                              // blocks of width 16 come first, then width 8, then width 4
    int first, last;                             // the blocks of this launch's width class
    int bd, to_tmp14;
};

struct Window { int x0, x1, y0, y1; };           // valid sample coordinates of one fetch (picture, or DMVR window)

__device__ __forceinline__ Window fetch_window(int pw, int ph, bool dmvr, int x_sb, int y_sb, int bw, int bh, int before, int after)
{
    Window w = { 0, pw - 1, 0, ph - 1 };
    if (dmvr) {
        // emulated_edge_dmvr (vvc_inter.c:60-89): fetches are confined to what the un-refined vector would have read
        w.x0 = clip3(x_sb - before, 0, pw - 1);
        w.y0 = clip3(y_sb - before, 0, ph - 1);
        w.x1 = w.x0 + max(min(pw, x_sb + bw + after) - w.x0, 1) - 1;
        w.y1 = w.y0 + max(min(ph, y_sb + bh + after) - w.y0, 1) - 1;
    }
    return w;
}

__device__ __forceinline__ int bcw_weight(int idx)
{
    return idx == 1 ? 5 : idx == 2 ? 3 : idx == 3 ? 10 : idx == 4 ? -2 : 4;
}

__device__ __forceinline__ int pack_taps(int lo, int hi) { return (lo & 0xff) | ((hi & 0xff) << 8); }

// tap words of one filter for the two phases: even[k] = (f[2k], f[2k+1]); odd[k] = (f[2k-1], f[2k]) with f[-1] = f[TAPS] = 0
template <int TAPS>
struct TapWords {
    int even[TAPS / 2], odd[TAPS / 2 + 1];
    __device__ __forceinline__ void load(const int8_t *f)
    {
#pragma unroll
        for (int k = 0; k < TAPS / 2; k++)
            even[k] = pack_taps(f[2 * k], f[2 * k + 1]);
        odd[0] = pack_taps(0, f[0]);
#pragma unroll
        for (int k = 1; k < TAPS / 2; k++)
            odd[k] = pack_taps(f[2 * k - 1], f[2 * k]);
        odd[TAPS / 2] = pack_taps(f[TAPS - 1], 0);
    }
};

// out[i], i < 8: TAPS-tap filter at offset i of the sample row held as words wd[0..7] (two samples per word)
template <int TAPS>
__device__ __forceinline__ void filter8(int (&out)[8], const int (&wd)[8], const TapWords<TAPS> &t)
{
#pragma unroll
    for (int i = 0; i < 8; i++) {
        int acc = 0;
        if (i & 1) {
#pragma unroll
            for (int k = 0; k <= TAPS / 2; k++)
                if ((i >> 1) + k < 8)
                    acc = __dp2a_lo(wd[(i >> 1) + k], t.odd[k], acc);
        } else {
#pragma unroll
            for (int k = 0; k < TAPS / 2; k++)
                if ((i >> 1) + k < 8)
                    acc = __dp2a_lo(wd[(i >> 1) + k], t.even[k], acc);
        }
        out[i] = acc;
    }
}

// One component, one list, one lane group.  val[k] receives the 14-bit intermediate of the lane's output
// sample (x = li % w, y = (li / w) * outs + k).
//   li / gs  : lane index inside its group / lanes per group (gs = 2 * w, or w for the chroma halves)
//   win / tmp: this group's staging areas; ws / ts their row strides (even, odd number of words)
template <int TAPS>
__device__ __forceinline__ void interpolate(int (&val)[8], bool active, int li, int gs, int w, int h, int log2w,
                                            const uint16_t *src, int stride, const Window &fw, int xi, int yi,
                                            const int8_t *hf, const int8_t *vf, int bd,
                                            uint16_t *win, int ws, int16_t *tmp, int ts, int trim)
{
    // trim = 1: the filter set's outer taps are zero (the 6-tap affine luma set): the window's outer ring and the first /
    // last intermediate row are multiplied by zero whatever they hold, so they are neither fetched nor computed
    constexpr int BEFORE = TAPS / 2 - 1;
    const int rows = h + TAPS - 1, cols = w + TAPS - 1;

    __syncwarp();
    if (active) {
        // stage the window: a lane owns columns li, li + gs, ... and walks down the rows
        // (eight loads are issued before the first store so that they are in flight together: the dependent
        // load -> store walk down the rows was the kernel's main memory stall)
        for (int c = li + trim; c < cols - trim; c += gs) {
            const int sx = clip3(xi - BEFORE + c, fw.x0, fw.x1);
            for (int r0 = trim; r0 < rows - trim; r0 += 8) {
                uint16_t v[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int sy = clip3(yi - BEFORE + r0 + k, fw.y0, fw.y1);
                    v[k] = __ldg(src + (size_t)sy * stride + sx);
                }
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (r0 + k < rows - trim)
                        win[(r0 + k) * ws + c] = v[k];
            }
        }
    }
    __syncwarp();
    if (active) {
        // horizontal pass: an item is one row of one strip of up to 8 outputs
        TapWords<TAPS> t;
        t.load(hf);
        const int strips = w > 8 ? 2 : 1, sw = min(w, 8);
        for (int i = li; i < (rows - 2 * trim) * strips; i += gs) {
            const int r = trim + (strips == 2 ? i >> 1 : i), s = strips == 2 ? i & 1 : 0;
            const uint32_t *p = reinterpret_cast<const uint32_t *>(win + r * ws + s * 8);
            int wd[8], out[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                wd[k] = (int)p[k];
            filter8<TAPS>(out, wd, t);
#pragma unroll
            for (int x = 0; x < 8; x++)
                if (x < sw)
                    tmp[(s * 8 + x) * ts + r] = (int16_t)(out[x] >> (bd - 8));
        }
    }
    __syncwarp();
    if (active) {
        // vertical pass: lane (x, chunk) produces `outs` vertically adjacent samples of column x
        const int nchunk = h >= 4 ? 2 : 1, outs = h / nchunk;
        const int x = li & (w - 1), chunk = li >> log2w;
        if (chunk < nchunk) {
            TapWords<TAPS> t;
            t.load(vf);
            const uint32_t *p = reinterpret_cast<const uint32_t *>(tmp + x * ts + chunk * outs);
            int wd[8];
#pragma unroll
            for (int k = 0; k < 8; k++)
                wd[k] = (int)p[k];
            filter8<TAPS>(val, wd, t);
#pragma unroll
            for (int k = 0; k < 8; k++)
                val[k] >>= 6;
        }
    }
}

// ---- DMVR (8.5.3): dmvr_mv_refine / parametric_mv_refine (vvc_inter.c:642-748), bilinear dmvr[][] (vvc_inter_template.c:
// 324-413), sad (vvcdsp.c:49-65).  One lane group refines one block; `buf` holds the two (w+4)x(h+4) bilinear
// predictions and the 25 SADs.  mv[][] is updated in place (all lanes of the group get the result); returns false if
// BDOF must be switched off for the block.  Called by whole warps; `go` selects the groups that really refine.
// Reductions over aligned lane groups of `gs` lanes.  The group size may differ between the lanes of a warp (blocks of
// different height share a warp), so every lane executes every shuffle and only the add is conditional.
__device__ __forceinline__ int group_sum(int v, int gs)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const int t = __shfl_xor_sync(0xffffffffu, v, o);
        if (o < gs)
            v += t;
    }
    return v;
}

__device__ __forceinline__ int group_min(int v, int gs)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        const int t = __shfl_xor_sync(0xffffffffu, v, o);
        if (o < gs)
            v = min(v, t);
    }
    return v;
}

__device__ __forceinline__ int dmvr_parabola(int sm, int sc, int sp)
{
    int denom = ((sm + sp) - (sc << 1)) << 3;
    if (!denom)
        return 0;
    if (sm == sc)
        return -8;
    if (sp == sc)
        return 8;
    int num = (sm - sp) * 16, q = 0;
    const bool neg = num < 0;
    if (neg)
        num = -num;
#pragma unroll
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

__device__ __forceinline__ bool dmvr_refine(const InterArgs &a, const VVCB200InterParams &p, const VVCB200PredBlock &b,
                                            bool go, int li, int gs, int16_t *buf, int (&mv)[2][2])
{
    const int w = b.w, h = b.h, bd = a.bd;
    const int pw4 = w + 4, ph4 = h + 4;
    int16_t *pred[2] = { buf, buf + 20 * pw4 };
    int *sads = reinterpret_cast<int *>(buf + 40 * pw4 + ((40 * pw4) & 1));
    __syncwarp();
    if (go) {
#pragma unroll
        for (int l = 0; l < 2; l++) {
            const PlanesView &ref = a.refs[p.ref_slot[l][b.ref_idx[l] & 15] & (kMaxRefs - 1)];
            const uint16_t *src = ref.p[0];
            const int stride = ref.stride[0], pw = ref.w[0], ph = ref.h[0];
            const int mx = mv[l][0] & 15, my = mv[l][1] & 15;
            const int xi = b.x0 + (mv[l][0] >> 4) - 2, yi = b.y0 + (mv[l][1] >> 4) - 2;
            const int fx0 = VVC_DMVR_FILTERS[mx * 2], fx1 = VVC_DMVR_FILTERS[mx * 2 + 1];
            const int fy0 = VVC_DMVR_FILTERS[my * 2], fy1 = VVC_DMVR_FILTERS[my * 2 + 1];
            const int sh1 = bd - 6, of1 = 1 << (sh1 - 1);
            if (li < pw4) {
                // a lane owns one column of the (w + 4) x (h + 4) bilinear block; its two source columns are read four
                // rows at a time so that eight loads are in flight together (row by row, each row waited for the one
                // before: 40 L2 round trips in a row made this search the longest thing a warp of the grid does)
                const int sx0 = clip3(xi + li, 0, pw - 1), sx1 = clip3(xi + li + 1, 0, pw - 1);
                const uint16_t *first = src + (size_t)clip3(yi, 0, ph - 1) * stride;
                int top0 = __ldg(first + sx0), top1 = __ldg(first + sx1);
                for (int rb = 0; rb < ph4; rb += 4) {
                    int c0[4], c1[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint16_t *rr = src + (size_t)clip3(yi + rb + k + 1, 0, ph - 1) * stride;
                        c0[k] = __ldg(rr + sx0);
                        c1[k] = __ldg(rr + sx1);
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        if (rb + k < ph4) {
                            const int a0 = k ? c0[k - 1] : top0, a1 = k ? c1[k - 1] : top1;     // row r: (sx0, sx1); row r + 1: c0, c1
                            int v;
                            if (!mx && !my)
                                v = bd > 10 ? (a0 + (1 << (bd - 11))) >> (bd - 10) : a0 << (10 - bd);
                            else if (!my)
                                v = (fx0 * a0 + fx1 * a1 + of1) >> sh1;
                            else if (!mx)
                                v = (fy0 * a0 + fy1 * c0[k] + of1) >> sh1;
                            else {
                                const int t0 = (fx0 * a0 + fx1 * a1 + of1) >> sh1;
                                const int t1 = (fx0 * c0[k] + fx1 * c1[k] + of1) >> sh1;
                                v = (fy0 * t0 + fy1 * t1 + 8) >> 4;
                            }
                            pred[l][(rb + k) * pw4 + li] = (int16_t)v;
                        }
                    }
                    top0 = c0[3];
                    top1 = c1[3];
                }
            }
        }
    }
    __syncwarp();
    // centre SAD, every other row (vvcdsp.c:49-65), all lanes of the group share the samples
    int part = 0;
    if (go) {
        const int log2w = 31 - __clz(w);
        for (int i = li; i < (w * h) >> 1; i += gs) {
            const int x = i & (w - 1), y = (i >> log2w) << 1;
            part += abs(pred[0][(2 + y) * pw4 + 2 + x] - pred[1][(2 + y) * pw4 + 2 + x]);
        }
    }
    int min_sad = group_sum(part, gs);
    min_sad -= min_sad >> 2;
    const bool search = go && min_sad >= w * h;
    if (search) {
        // one search position per lane (24 besides the centre)
        for (int pos = li; pos < 25; pos += gs) {
            int sad = min_sad;
            if (pos != 12) {
                const int dx = pos % 5 - 2, dy = pos / 5 - 2;
                const int16_t *p0 = pred[0] + (2 + dy) * pw4 + 2 + dx, *p1 = pred[1] + (2 - dy) * pw4 + 2 - dx;
                sad = 0;
                for (int y = 0; y < h; y += 2)
                    for (int x = 0; x < w; x++)
                        sad += abs(p0[y * pw4 + x] - p1[y * pw4 + x]);
            }
            sads[pos] = sad;
        }
    }
    __syncwarp();
    int best = 0x7fffffff;
    if (search)
        for (int pos = li; pos < 25; pos += gs)
            best = min(best, sads[pos] * 32 + (pos == 12 ? 0 : pos + 1));   // centre wins ties, then raster order
    best = group_min(best, gs);
    if (search) {
        const int pos = (best & 31) ? (best & 31) - 1 : 12;
        const int min_dx = pos % 5, min_dy = pos / 5;
        min_sad = best >> 5;
        int dmx = (min_dx - 2) * 16, dmy = (min_dy - 2) * 16;
        if (min_dx != 0 && min_dx != 4 && min_dy != 0 && min_dy != 4) {
            dmx += dmvr_parabola(sads[pos - 1], sads[pos], sads[pos + 1]);
            dmy += dmvr_parabola(sads[pos - 5], sads[pos], sads[pos + 5]);
        }
        mv[0][0] = clip3(mv[0][0] + dmx, -(1 << 17), (1 << 17) - 1);
        mv[0][1] = clip3(mv[0][1] + dmy, -(1 << 17), (1 << 17) - 1);
        mv[1][0] = clip3(mv[1][0] - dmx, -(1 << 17), (1 << 17) - 1);
        mv[1][1] = clip3(mv[1][1] - dmy, -(1 << 17), (1 << 17) - 1);
    }
    __syncwarp();
    return !go || min_sad >= 2 * w * h;
}

// ---- BDOF: bdof_fetch_samples / prof_grad_filter / derive_bdof_vx_vy / apply_bdof (vvc_inter_template.c:101-128, 135-158,
// 237-317).  `buf`: the two 14-bit predictions with a one-sample ring of integer reference samples, then the four
// gradient blocks.  The reference pads ring and gradients with replicated edge values after the gradients are formed;
// here that is an index clamp on read.  v0 / v1: the lane's own prediction samples (x = li % w, rows chunk * outs + k).
__device__ __forceinline__ void bdof_block(const InterArgs &a, const VVCB200InterParams &p, const VVCB200PredBlock &b,
                                           bool go, int li, int gs, int16_t *buf, int8_t *vxy, const int (&v0)[8],
                                           const int (&v1)[8], const int (&mv)[2][2], const Window (&fw)[2],
                                           const uint16_t *lut)
{
    const int w = b.w, h = b.h, bd = a.bd;
    const int log2w = 31 - __clz(w);
    const int ps = w + 2;                                   // ringed row stride
    int16_t *pa[2] = { buf, buf + (h + 2) * ps };
    int16_t *gh[2] = { buf + 2 * (h + 2) * ps, buf + 2 * (h + 2) * ps + w * h };
    int16_t *gv[2] = { gh[1] + w * h, gh[1] + 2 * w * h };
    const int outs = h >> 1;
    __syncwarp();
    if (go) {
        const int x = li & (w - 1), chunk = li >> log2w;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (k < outs) {
                pa[0][(chunk * outs + k + 1) * ps + x + 1] = (int16_t)v0[k];
                pa[1][(chunk * outs + k + 1) * ps + x + 1] = (int16_t)v1[k];
            }
        // ring: integer samples at the rounded position, << (14 - bd)
#pragma unroll
        for (int l = 0; l < 2; l++) {
            const PlanesView &ref = a.refs[p.ref_slot[l][b.ref_idx[l] & 15] & (kMaxRefs - 1)];
            const int ox = b.x0 + (mv[l][0] >> 4) + ((mv[l][0] & 15) >> 3), oy = b.y0 + (mv[l][1] >> 4) + ((mv[l][1] & 15) >> 3);
            const int n_ring = 2 * (w + 2) + 2 * h;
            for (int i = li; i < n_ring; i += gs) {
                int rx, ry;
                if (i < w + 2) { rx = i - 1; ry = -1; }
                else if (i < 2 * (w + 2)) { rx = i - (w + 2) - 1; ry = h; }
                else { const int j = i - 2 * (w + 2); rx = (j & 1) ? w : -1; ry = j >> 1; }
                const int sx = clip3(ox + rx, fw[l].x0, fw[l].x1), sy = clip3(oy + ry, fw[l].y0, fw[l].y1);
                pa[l][(ry + 1) * ps + rx + 1] = (int16_t)(ref.p[0][(size_t)sy * ref.stride[0] + sx] << (14 - bd));
            }
        }
    }
    __syncwarp();
    if (go) {
        for (int i = li; i < w * h; i += gs) {
            const int x = i & (w - 1), y = i >> log2w;
#pragma unroll
            for (int l = 0; l < 2; l++) {
                const int16_t *c = pa[l] + (y + 1) * ps + x + 1;
                gh[l][i] = (int16_t)((c[1] >> 6) - (c[-1] >> 6));
                gv[l][i] = (int16_t)((c[ps] >> 6) - (c[-ps] >> 6));
            }
        }
    }
    __syncwarp();
    {
        // per 4x4: sums over its 6x6 neighbourhood, `lps` lanes per 4x4
        const int nsb = (w * h) >> 4, lps = gs / nsb;
        const int sb = li / lps, sub = li - sb * lps;
        const int bx = (sb & ((w >> 2) - 1)) << 2, by = (sb / (w >> 2)) << 2;
        int sgx2 = 0, sgy2 = 0, sgxgy = 0, sgxdi = 0, sgydi = 0;
        if (go) {
            for (int j = sub; j < 36; j += lps) {
                const int cy = clip3(by - 1 + j / 6, 0, h - 1), cx = clip3(bx - 1 + j % 6, 0, w - 1);
                const int i = (cy << log2w) + cx;
                const int diff = (pa[0][(cy + 1) * ps + cx + 1] >> 4) - (pa[1][(cy + 1) * ps + cx + 1] >> 4);
                const int th = (gh[0][i] + gh[1][i]) >> 1, tv = (gv[0][i] + gv[1][i]) >> 1;
                const int sh_ = (th > 0) - (th < 0), sv = (tv > 0) - (tv < 0);
                sgx2 += abs(th);
                sgy2 += abs(tv);
                sgxgy += sv * th;
                sgxdi -= sh_ * diff;
                sgydi -= sv * diff;
            }
        }
        sgx2 = group_sum(sgx2, lps);
        sgy2 = group_sum(sgy2, lps);
        sgxgy = group_sum(sgxgy, lps);
        sgxdi = group_sum(sgxdi, lps);
        sgydi = group_sum(sgydi, lps);
        if (go && sub == 0) {
            int vx = 0, vy = 0;
            if (sgx2 > 0)
                vx = clip3((sgxdi * 4) >> (31 - __clz(sgx2)), -15, 15);
            if (sgy2 > 0)
                vy = clip3(((sgydi * 4) - ((vx * sgxgy) >> 1)) >> (31 - __clz(sgy2)), -15, 15);
            vxy[2 * sb] = (int8_t)vx;
            vxy[2 * sb + 1] = (int8_t)vy;
        }
    }
    __syncwarp();
    if (go) {
        const int sh4 = 15 - bd, of4 = 1 << (sh4 - 1);
        for (int i = li; i < w * h; i += gs) {
            const int x = i & (w - 1), y = i >> log2w;
            const int sb = (y >> 2) * (w >> 2) + (x >> 2);
            const int vx = vxy[2 * sb], vy = vxy[2 * sb + 1];
            const int off = vx * (gh[0][i] - gh[1][i]) + vy * (gv[0][i] - gv[1][i]);
            int r = clip_pixel((pa[0][(y + 1) * ps + x + 1] + of4 + pa[1][(y + 1) * ps + x + 1] + off) >> sh4, bd);
            if ((b.flags & VVCB200_PB_LMCS) && lut)
                r = lut[r];
            a.dst.p[0][(size_t)(b.y0 + y) * a.dst.stride[0] + b.x0 + x] = (uint16_t)r;
        }
    }
    __syncwarp();
}

// ---- PROF (fetch_samples + apply_prof / apply_prof_uni[_w], vvc_inter_template.c:130-235) on a 4x4 affine sub-block:
// the 14-bit prediction gets a ring of integer reference samples, gradients are taken on the ringed block and every
// sample moves by gh * diff_mv_x + gv * diff_mv_y (clipped); the usual uni / bi combination follows.
__device__ __forceinline__ void prof_block(const InterArgs &a, const VVCB200PredBlock &b, bool go, int li, int gs, int l,
                                           int16_t *r /* 6 x 6 */, int (&v)[8], const PlanesView &ref, const int (&mv)[2][2],
                                           const Window &fw)
{
    const int x = li & 3, chunk = li >> 2;               // 8 lanes per 4x4 block: two rows each
    __syncwarp();
    if (go) {
        r[(chunk * 2 + 1) * 6 + x + 1] = (int16_t)v[0];
        r[(chunk * 2 + 2) * 6 + x + 1] = (int16_t)v[1];
        const int ox = b.x0 + (mv[l][0] >> 4) + ((mv[l][0] & 15) >> 3), oy = b.y0 + (mv[l][1] >> 4) + ((mv[l][1] & 15) >> 3);
        for (int i = li; i < 20; i += gs) {
            int rx, ry;
            if (i < 6) { rx = i - 1; ry = -1; }
            else if (i < 12) { rx = i - 7; ry = 4; }
            else { rx = (i & 1) ? 4 : -1; ry = (i - 12) >> 1; }
            const int sx = clip3(ox + rx, fw.x0, fw.x1), sy = clip3(oy + ry, fw.y0, fw.y1);
            r[(ry + 1) * 6 + rx + 1] = (int16_t)(ref.p[0][(size_t)sy * ref.stride[0] + sx] << (14 - a.bd));
        }
    }
    __syncwarp();
    if (go) {
        const VVCB200ProfTable *pt = a.prof + b.aux;
        const int limit = 1 << max(13, a.bd + 1);
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int y = chunk * 2 + k;
            const int16_t *q = r + (y + 1) * 6 + x + 1;
            const int gh = (q[1] >> 6) - (q[-1] >> 6), gv = (q[6] >> 6) - (q[-6] >> 6);
            const int di = gh * pt->diff_mv_x[l][y * 4 + x] + gv * pt->diff_mv_y[l][y * 4 + x];
            v[k] = q[0] + clip3(di, -limit, limit - 1);
        }
    }
    __syncwarp();
}

// One component of one block by one lane group (both lists + weighting + store).  WC: the block's width in this
// component, fixed by the width class of the launch.
// kFancyX false: the instantiation for records that use none of DMVR / BDOF / PROF / GPM — two thirds of the kernel's code
// is those tools, and the plain records (most of a picture) run a quarter faster without it in the instruction cache.
template <int TAPS, int WC, bool kFancyX>
__device__ __forceinline__ void predict_component(const InterArgs &a, const VVCB200InterParams &p, const VVCB200PredBlock &b,
                                                  bool active, int c, int li, int gs, uint16_t *win, int ws,
                                                  int16_t *tmp, int ts, const int (&mv)[2][2], bool dmvr, bool bdof,
                                                  int16_t *bdof_buf, int8_t *bdof_vxy, const uint16_t *lut)
{
    const int bd = a.bd;
    const int hs = c ? 1 : 0;
    const int w = WC, h = max(b.h >> hs, 1);
    const int x0 = b.x0 >> hs, y0 = b.y0 >> hs;
    const int log2w = 31 - __clz(w);
    const bool gpm = kFancyX && (b.flags & VVCB200_PB_GPM);
    const int set = (c || gpm) ? 0 : (b.flags & VVCB200_PB_AFFINE) ? 2 : (b.flags & VVCB200_PB_HPEL_IF) ? 1 : 0;
    int v0[8], v1[8];
    Window fw[2];

#pragma unroll
    for (int l = 0; l < 2; l++) {
        const bool use = active && (b.pred_flag & (1 << l));
        const int mvx = mv[l][0], mvy = mv[l][1];
        const int mx = c ? (mvx & 31) : (mvx & 15);
        const int my = c ? (mvy & 31) : (mvy & 15);
        const int xi = x0 + (mvx >> (4 + hs)), yi = y0 + (mvy >> (4 + hs));
        const int8_t *hf = c ? VVC_CHROMA_FILTERS + (set * 32 + mx) * 4 : VVC_LUMA_FILTERS + (set * 16 + mx) * 8;
        const int8_t *vf = c ? VVC_CHROMA_FILTERS + (set * 32 + my) * 4 : VVC_LUMA_FILTERS + (set * 16 + my) * 8;
        // GPM: prediction l is gpm_mv[l], which names its own reference list (vvc_inter.c:507-509)
        const int list = gpm ? (b.aux >> (18 + l)) & 1 : l;
        const PlanesView &ref = a.refs[p.ref_slot[list][b.ref_idx[l] & 15] & (kMaxRefs - 1)];
        fw[l] = fetch_window(ref.w[c], ref.h[c], dmvr, x0 + (b.mv[l][0] >> (4 + hs)), y0 + (b.mv[l][1] >> (4 + hs)), w, h,
                             TAPS / 2 - 1, TAPS / 2);
        if (l == 0)
            interpolate<TAPS>(v0, use, li, gs, w, h, log2w, ref.p[c], ref.stride[c], fw[l], xi, yi, hf, vf, bd,
                              win, ws, tmp, ts, TAPS == 8 && set == 2);
        else
            interpolate<TAPS>(v1, use, li, gs, w, h, log2w, ref.p[c], ref.stride[c], fw[l], xi, yi, hf, vf, bd,
                              win, ws, tmp, ts, TAPS == 8 && set == 2);
        if (TAPS == 8 && kFancyX && a.prof) {
            const bool go = use && w == 4 && h == 4 && (b.flags & (l ? VVCB200_PB_PROF1 : VVCB200_PB_PROF0));
            if (__any_sync(0xffffffffu, go)) {
                if (l == 0)
                    prof_block(a, b, go, li, gs, l, tmp, v0, ref, mv, fw[l]);
                else
                    prof_block(a, b, go, li, gs, l, tmp, v1, ref, mv, fw[l]);
            }
        }
    }
    if (TAPS == 8 && kFancyX && !a.to_tmp14 && __any_sync(0xffffffffu, active && bdof))
        bdof_block(a, p, b, active && bdof, li, gs, bdof_buf, bdof_vxy, v0, v1, mv, fw, lut);
    if (!active || (TAPS == 8 && kFancyX && bdof && !a.to_tmp14))
        return;

    // ---- combine (weights are uniform inside a lane group)
    int mode;                       // 0 uni, 1 uni_w, 2 avg, 3 w_avg, 4 GPM blend
    int wgt0 = 0, wgt1 = 0, off = 0, sh = 0, ox = 0;
    const int l_uni = (b.pred_flag & 1) ? 0 : 1;
    // GPM weight addressing of pred_gpm_blk (vvc_inter.c:466-505): weight of luma position (X, Y) inside the coding unit
    const uint8_t *gpm_w = nullptr;
    int gpm_x = 0, gpm_y = 0, gpm_sx = 1, gpm_sy = 1;
    if (gpm) {
        const int pidx = b.aux & 63, lw = (b.aux >> 6) & 3, lh = (b.aux >> 8) & 3;
        const int angle = VVC_GPM_ANGLE_IDX[pidx], mirror = VVC_GPM_ANGLE_TO_MIRROR[angle];
        const int off_x = VVC_GPM_WEIGHTS_OFFSET_X[(pidx * 4 + lh) * 4 + lw], off_y = VVC_GPM_WEIGHTS_OFFSET_Y[(pidx * 4 + lh) * 4 + lw];
        gpm_w = VVC_GPM_WEIGHTS + VVC_GPM_ANGLE_TO_WEIGHTS_IDX[angle] * 112 * 112;
        gpm_sx = mirror == 1 ? -1 : 1;
        gpm_sy = mirror == 2 ? -1 : 1;
        gpm_x = (mirror == 1 ? 111 - off_x : off_x) + gpm_sx * (int)(((b.aux >> 10) & 15) << 2);
        gpm_y = (mirror == 2 ? 111 - off_y : off_y) + gpm_sy * (int)(((b.aux >> 14) & 15) << 2);
        mode = 4;
        sh = max(5, 17 - bd);
        off = 1 << (sh - 1);
    } else if (b.pred_flag != 3) {
        sh = 14 - bd;
        if (b.flags & VVCB200_PB_WP) {
            mode = 1;
            wgt0 = p.weight[l_uni][c][b.ref_idx[l_uni] & 15];
            ox   = p.offset[l_uni][c][b.ref_idx[l_uni] & 15] * (1 << (bd - 8));
            sh  += p.log2_denom[c > 0];
            off  = 1 << (sh - 1);
        } else {
            mode = 0;
            off  = 1 << (sh - 1);
        }
    } else {
        const bool wp = (b.flags & VVCB200_PB_WP) && !(b.flags & VVCB200_PB_DMVR);
        const int bcw = b.bcw_idx;
        sh = max(3, 15 - bd);
        if ((!wp && !bcw) || (bcw && (b.flags & VVCB200_PB_CIIP))) {
            mode = 2;
            off  = 1 << (sh - 1);
        } else {
            mode = 3;
            int o0 = 0, o1 = 0;
            if (bcw) {
                wgt1 = bcw_weight(bcw);
                wgt0 = 8 - wgt1;
                sh  += 2;
            } else {
                wgt0 = p.weight[0][c][b.ref_idx[0] & 15];
                wgt1 = p.weight[1][c][b.ref_idx[1] & 15];
                o0   = p.offset[0][c][b.ref_idx[0] & 15];
                o1   = p.offset[1][c][b.ref_idx[1] & 15];
                sh  += p.log2_denom[c > 0];
            }
            off = (((o0 + o1) << (bd - 8)) + 1) << (sh - 1);
        }
    }
    const bool lmcs = c == 0 && (b.flags & VVCB200_PB_LMCS) && lut;
    const int nchunk = h >= 4 ? 2 : 1, outs = h / nchunk;
    const int x = li & (w - 1), chunk = li >> log2w;
    if (chunk >= nchunk)
        return;
    if (a.to_tmp14) {
        int16_t *o = reinterpret_cast<int16_t *>(a.tmp14.p[c]) + (size_t)(y0 + chunk * outs) * a.tmp14.stride[c] + x0 + x;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (k < outs)
                o[(size_t)k * a.tmp14.stride[c]] = (int16_t)(l_uni == 0 ? v0[k] : v1[k]);
        return;
    }
    // every mode as one multiply-add form r = (v0 * wgt0 + v1 * wgt1 + off) >> sh (+ ox): the mode is decided once per block,
    // not once per sample, and the row pointer walks down the block
    const int pmax = (1 << bd) - 1;
    const int dstride = a.dst.stride[c];
    uint16_t *o = a.dst.p[c] + (size_t)(y0 + chunk * outs) * dstride + x0 + x;
    if (mode == 4) {
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (k < outs) {
                const int y = chunk * outs + k;
                const int wg = gpm_w[(gpm_y + gpm_sy * (y << hs)) * 112 + gpm_x + gpm_sx * (x << hs)];
                int r = min(max((v0[k] * wg + v1[k] * (8 - wg) + off) >> sh, 0), pmax);
                if (lmcs)
                    r = lut[r];
                o[(size_t)k * dstride] = (uint16_t)r;
            }
        return;
    }
    if (mode < 2) {                                      // one list: ((v * w + off) >> sh) + ox, w = 1 without weights
        if (mode == 0)
            wgt0 = 1;
#pragma unroll
        for (int k = 0; k < 8; k++)
            if (k < outs) {
                int r = min(max((((l_uni == 0 ? v0[k] : v1[k]) * wgt0 + off) >> sh) + ox, 0), pmax);
                if (lmcs)
                    r = lut[r];
                o[(size_t)k * dstride] = (uint16_t)r;
            }
        return;
    }
    if (mode == 2)
        wgt0 = wgt1 = 1;
#pragma unroll
    for (int k = 0; k < 8; k++)
        if (k < outs) {
            int r = min(max((v0[k] * wgt0 + v1[k] * wgt1 + off) >> sh, 0), pmax);
            if (lmcs)
                r = lut[r];
            o[(size_t)k * dstride] = (uint16_t)r;
        }
}

#ifndef VVC_INTER_MIN_BLOCKS
#define VVC_INTER_MIN_BLOCKS 4        // 64 registers.  Measured: 5 / 6 blocks (48 / 40 registers) spill and lose 8 / 12 % of the
#endif                                // picture; 3 blocks (80 registers) gain under 1 %
// One launch per width class (kCls 0 / 1 / 2: blocks of width 16 / 8 / 4, a.first .. a.last): widths, lane groups and
// staging strides are compile-time constants, the width-4 kernel carries no DMVR / BDOF code (they need 8x8) and the
// others no PROF code (4x4 sub-blocks only) — the kernel is bound by instruction issue and fetch.
// kPart: 0 every record; 1 only the warps none of whose records uses DMVR / BDOF / PROF / GPM (the instantiation carries
// no code for them); 2 only the other warps.  Parts 1 and 2 are launched over the same range and between them do what
// part 0 does: no contract on the record order is needed, a sorted batch just makes whole CTAs leave at once.
constexpr int kFancyFlags = VVCB200_PB_DMVR | VVCB200_PB_BDOF | VVCB200_PB_GPM | VVCB200_PB_PROF0 | VVCB200_PB_PROF1;
template <int kCls, int kPart>
__global__ void __launch_bounds__(kInterWarps * 32, VVC_INTER_MIN_BLOCKS)
inter_kernel(const InterArgs a, const VVCB200InterParams p)
{
    constexpr bool kFancyX = kPart != 1;
    __shared__ __align__(16) uint16_t s_buf[kInterWarps][kBufElems];
    __shared__ int s_mv[kInterWarps][4][4];             // refined vectors of the warp's blocks, luma phase -> chroma phase
    __shared__ int8_t s_vxy[kInterWarps][2][32];        // BDOF (vx, vy) per 4x4 of the warp's (up to two) blocks
    // the forward LMCS LUT of the slice in shared memory: every luma sample of a mapped block goes through it
    __shared__ uint16_t s_lut[4096];
    const uint16_t *lut = nullptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wi = blockIdx.x * kInterWarps + warp;

    // the blocks this warp owns: one of width 16, two of width 8 or four of width 4
    constexpr int cls = kCls;
    const int first = a.first + (wi << kCls), last = a.last;
    bool mine = first < last;
    if (kPart != 0) {
        const int bi = first + (lane >> (5 - cls));
        const bool fancy = bi < last && (a.blocks[bi].flags & kFancyFlags);
        mine = mine && (__any_sync(0xffffffffu, fancy) == (kPart == 2));
        if (!__syncthreads_or(mine))
            return;
    }
    if (a.lmcs_fwd) {
        for (int i = threadIdx.x; i < (1 << a.bd); i += kInterWarps * 32)
            s_lut[i] = __ldg(a.lmcs_fwd + i);
        __syncthreads();
        lut = s_lut;
    }
    if (!mine)
        return;
    uint16_t *win = s_buf[warp];
    int16_t *tmp = reinterpret_cast<int16_t *>(s_buf[warp]) + kWinElems;

    {   // luma: groups of 2 * w lanes
        const int gs = 32 >> cls, g = lane >> (5 - cls), li = lane & (gs - 1);
        const int bi = first + g;
        const bool active = bi < last;
        const VVCB200PredBlock b = a.blocks[active ? bi : first];
        const bool luma = active && !(b.flags & VVCB200_PB_CHROMA_ONLY);
        int mv[2][2] = { { b.mv[0][0], b.mv[0][1] }, { b.mv[1][0], b.mv[1][1] } };
        const bool dmvr = kFancyX && kCls < 2 && (b.flags & VVCB200_PB_DMVR) && b.pred_flag == 3 && b.h >= 8;
        bool bdof = kFancyX && kCls < 2 && luma && (b.flags & VVCB200_PB_BDOF) && b.pred_flag == 3 && b.h >= 8;
        if (kFancyX && kCls < 2 && __any_sync(0xffffffffu, active && dmvr)) {
            // each group's scratch: 2 predictions of (w+4) x 20 samples + 25 SADs
            const int region = cls == 0 ? kBufElems : kBufElems / 2;
            if (!dmvr_refine(a, p, b, active && dmvr, li, gs, reinterpret_cast<int16_t *>(s_buf[warp]) + g * region, mv))
                bdof = false;
        }
        if (active && li == 0) {
#pragma unroll
            for (int k = 0; k < 4; k++)
                s_mv[warp][g][k] = mv[k >> 1][k & 1];
            if (a.refined_mv)
#pragma unroll
                for (int k = 0; k < 4; k++)
                    a.refined_mv[(size_t)bi * 4 + k] = mv[k >> 1][k & 1];
        }
        const int ws = (16 >> cls) + 10;                // 26 / 18 / 14 samples: an odd number of words
        const int win_sz = 23 * ws, tmp_sz = (16 >> cls) * 26;
        const int bdof_region = cls == 0 ? 0 : kBufElems / 2;
        predict_component<8, (16 >> kCls), kFancyX>(a, p, b, luma, 0, li, gs, win + g * win_sz, ws, tmp + g * tmp_sz, 26, mv, active && dmvr, bdof,
                             reinterpret_cast<int16_t *>(s_buf[warp]) + g * bdof_region, s_vxy[warp][g & 1], lut);
    }
    __syncwarp();
    {   // chroma: each luma group splits into a Cb half and a Cr half
        const int gs = 16 >> cls, g = lane / gs, li = lane & (gs - 1);
        const int bi = first + (g >> 1);
        const bool active = bi < last;
        const VVCB200PredBlock b = a.blocks[active ? bi : first];
        int mv[2][2];
#pragma unroll
        for (int k = 0; k < 4; k++)
            mv[k >> 1][k & 1] = s_mv[warp][(g >> 1) & 3][k];
        const bool dmvr = kFancyX && kCls < 2 && active && (b.flags & VVCB200_PB_DMVR) && b.pred_flag == 3 && b.h >= 8;
        const int wc = 8 >> cls;
        const int ws = wc + 6;                          // 14 / 10 / 8 samples
        const int win_sz = 11 * ws + 8, tmp_sz = wc * 14;
        predict_component<4, (8 >> kCls), kFancyX>(a, p, b, active && !(b.flags & VVCB200_PB_LUMA_ONLY), 1 + (g & 1), li, gs,
                             win + g * win_sz, ws, tmp + g * tmp_sz, 14, mv, dmvr, false, nullptr, nullptr, nullptr);
    }
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_inter_batch(const VVCB200Planes *dst, const VVCB200Planes *refs, int n_refs,
                                   const VVCB200PredBlock *d_blocks, int n, const VVCB200InterParams *params,
                                   const uint16_t *d_lmcs_fwd_lut, const VVCB200Planes *d_tmp14, int32_t *d_refined_mv,
                                   const VVCB200ProfTable *d_prof, void *stream)
{
    if (!params || (!dst && !d_tmp14) || !refs || n_refs <= 0 || n_refs > kMaxRefs || n < 0 || (n && !d_blocks)) {
        set_error("vvcb200_inter_batch: bad argument (n_refs must be 1..%d)", kMaxRefs);
        return VVCB200_ERR_ARG;
    }
    if (params->bit_depth != 10 && params->bit_depth != 12) {
        set_error("unsupported bit depth %d (10 or 12: 8-bit pictures are uint8 planes in the reference, which are not built)", params->bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (params->n_w16 < 0 || params->n_w8 < 0 || params->n_w16 + params->n_w8 > n) {
        set_error("vvcb200_inter_batch: n_w16 %d + n_w8 %d exceed the batch of %d blocks", params->n_w16, params->n_w8, n);
        return VVCB200_ERR_ARG;
    }
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < 16; i++)
            if (params->ref_slot[l][i] >= kMaxRefs) {
                set_error("vvcb200_inter_batch: ref_slot[%d][%d] = %d out of range", l, i, params->ref_slot[l][i]);
                return VVCB200_ERR_ARG;
            }
    if (n == 0)
        return 0;
    InterArgs a = {};
    if (dst)
        a.dst = make_view(dst);
    if (d_tmp14) {
        a.tmp14 = make_view(d_tmp14);
        a.to_tmp14 = 1;
    }
    for (int i = 0; i < kMaxRefs; i++)
        a.refs[i] = make_view(&refs[i < n_refs ? i : 0]);
    a.blocks = d_blocks;
    a.lmcs_fwd = d_lmcs_fwd_lut;
    a.refined_mv = d_refined_mv;
    a.prof = d_prof;
    a.n = n;
    a.n16 = params->n_w16;
    a.n8 = params->n_w8;
    a.bd = params->bit_depth;
    const int bound[4] = { 0, a.n16, a.n16 + a.n8, n };
    const cudaStream_t main_stream = (cudaStream_t)stream;
    // The launches: classes 0 / 1 as two grids over the same records (warps with a DMVR / BDOF / GPM record, the slow ones,
    // first; then the warps without, on a kernel a third the size), class 2 as one.  They write disjoint blocks: side by side.
    struct Launch { int cls, part; } launches[5];
    int n_launch = 0;
    for (int cls = 0; cls < 3; cls++) {
        if (bound[cls + 1] == bound[cls])
            continue;
        if (cls < 2) {
            launches[n_launch++] = { cls, 2 };
            launches[n_launch++] = { cls, 1 };
        } else {
            launches[n_launch++] = { cls, 0 };
        }
    }
    ForkStreams *f = n_launch > 1 ? fork_streams(main_stream) : nullptr;
    if (f)
        VVC_CUDA(cudaEventRecord(f->fork, main_stream));
    int used = 0;                                                         // side streams in use
    for (int i = 0; i < n_launch; i++) {
        const int cls = launches[i].cls, part = launches[i].part;
        a.first = bound[cls];
        a.last = bound[cls + 1];
        const int count = a.last - a.first;
        cudaStream_t st = main_stream;                                    // the first launch stays on the caller's stream
        if (f && i > 0 && used < 4) {
            st = f->side[used];
            VVC_CUDA(cudaStreamWaitEvent(st, f->fork, 0));
        }
        const int warps = (count + (1 << cls) - 1) >> cls;
        const int grid = (warps + kInterWarps - 1) / kInterWarps;
        if (cls == 0 && part == 2)
            inter_kernel<0, 2><<<grid, kInterWarps * 32, 0, st>>>(a, *params);
        else if (cls == 0)
            inter_kernel<0, 1><<<grid, kInterWarps * 32, 0, st>>>(a, *params);
        else if (cls == 1 && part == 2)
            inter_kernel<1, 2><<<grid, kInterWarps * 32, 0, st>>>(a, *params);
        else if (cls == 1)
            inter_kernel<1, 1><<<grid, kInterWarps * 32, 0, st>>>(a, *params);
        else
            inter_kernel<2, 0><<<grid, kInterWarps * 32, 0, st>>>(a, *params);
        VVC_LAUNCH_CHECK();
        if (st != main_stream) {
            VVC_CUDA(cudaEventRecord(f->join[used], st));
            VVC_CUDA(cudaStreamWaitEvent(main_stream, f->join[used], 0));
            used++;
        }
    }
    return 0;
}
