// residual.cu — dequant + BDPCM + inverse LFNST + inverse DCT-II/DST-VII/DCT-VIII + add_residual for
// a whole picture's batch of transform blocks (the body of itransform(), reference
// libavcodec/vvc/vvc_intra.c:432-478, with the DSP entries it calls: vvcdsp.c:94-138 (2-D/1-D drivers),
// vvc_itx_1d.c (1-D cores), vvcdsp_template.c:32-95 (add_residual, joint, bdpcm)).
//
// Mapping: one warp owns one transform block at a time; CTAs of 4 warps walk the (size-sorted)
// descriptor list with a grid stride so that all CTAs see the same mix of block sizes.  The block lives in shared memory between the two 1-D passes
// (levels -> dequantized 16-bit -> column pass -> 16-bit clip -> row pass -> epilogue), so HBM sees
// the int16 levels once and the destination samples once (read-modify-write).  All arithmetic is
// exact int32 on CUDA cores: accumulators wrap mod 2^32 exactly like the reference's `int`.
#include "common.cuh"

#define VVC_TABLE_QUAL static __device__ const
#include "vvc_tables.h"

namespace vvcb200 {

constexpr int kWarpsPerCta = 4;
constexpr int kChunk       = 32;            // transform blocks per CTA (Level-A kernels)
constexpr int kPackedWords = 1536;          // VVC_TX_PACKED: every transform matrix, dp2a-packed (6 KB)
constexpr int kMatPad      = kPackedWords * 4;
constexpr int kInElems     = 32 * 32;       // generic path: non-zero input box (zero-out keeps <= 32 x 32)
constexpr int kMidElems    = 64 * 32;       // generic path: after the column pass, h rows x nzw columns
constexpr int kPairStride  = 34;            // fast path: int16 row stride of the transposed input / the middle
                                            // block; 17 words -> lanes on different rows hit different banks
constexpr int kFastIn      = 32 * kPairStride;     // inT[x][k]
constexpr int kFastMid     = 64 * kPairStride;     // mid[y][x]
constexpr int kWarpElems   = kFastIn + kFastMid;   // 3264 int16 per warp (>= kInElems + kMidElems = 3072)
static_assert(kWarpElems >= kInElems + kMidElems, "generic path must fit");

struct ScalingLists {                        // passed by value (kernel parameter space, uniform reads)
    uint8_t rec[28][64];
    uint8_t dc[16];
};

struct ResidualArgs {
    PlanesView dst;                          // pixel planes (read-modify-write) when !to_residual
    PlanesView res;                          // int16 residual planes when to_residual
    const int16_t *levels;
    const VVCB200TransformBlock *tbs;
    int n;
    int n_class[4];                          // blocks per lane-group class (32/16/8/4 lanes per block); the rest is generic
    int first, last;                         // the blocks of this launch's class
    int bd, range, c_sign;
    int to_residual;
    int sparse;                              // levels come as bitmap + non-zero values (VVCB200ResidualParams.levels_sparse)
};

// The levels of one TB's scan box (nzw x nzh, row-major) into `dst`, by the `lpt` lanes of a group (lane li).
// Dense: the box as it is.  Sparse: ceil(n / 32) bitmap words (bit i % 32 of word i / 32 set = level i is non-zero) followed
// by the non-zero levels in box order; a lane walks its positions i = li, li + lpt, ... with a running count of the bits
// before the word it is in, so a level's place is that count plus the bits below its own in the word.
__device__ __forceinline__ void fetch_levels(const ResidualArgs &a, const VVCB200TransformBlock &d, int16_t *dst, int li, int lpt)
{
    const int n = d.nzw * d.nzh;
    const int16_t *lv = a.levels + d.coeff_off;
    const bool narrow = d.flags & VVCB200_TB_LEVELS8;    // int8 levels, two per slot
    if (!a.sparse) {
        if (narrow) {
            const int8_t *l8 = reinterpret_cast<const int8_t *>(lv);
            for (int i = li; i < n; i += lpt)
                dst[i] = l8[i];
        } else {
            for (int i = li; i < n; i += lpt)
                dst[i] = lv[i];
        }
        return;
    }
    const uint32_t *bitmap = reinterpret_cast<const uint32_t *>(lv);
    const int16_t *nz = lv + 2 * ((n + 31) >> 5);
    const int8_t *nz8 = reinterpret_cast<const int8_t *>(nz);
    int cur = -1, before = 0;
    uint32_t word = 0;
    for (int i = li; i < n; i += lpt) {
        const int wi = i >> 5;
        while (cur < wi) {                               // (at most once per word: lpt <= 32)
            before += __popc(word);
            word = __ldg(bitmap + ++cur);
        }
        const uint32_t bit = 1u << (i & 31);
        const int k = before + __popc(word & (bit - 1));
        dst[i] = (word & bit) ? (narrow ? (int16_t)nz8[k] : nz[k]) : (int16_t)0;
    }
}

// Packed matrix of (type, log2N) in shared memory: word [kp * (N/2) + jp], see tools/gen_tables.py
__device__ __forceinline__ const uint32_t *tx_packed(const uint32_t *smat, int type, int log2n)
{
    return smat + VVC_TX_PACKED_OFF[type * 8 + log2n];
}

// Element M[k][j] of a packed matrix (generic path only)
struct TxMatrix {
    const int8_t *bytes;
    int half_n;
    __device__ __forceinline__ int at(int k, int j) const
    {
        return bytes[(((k >> 1) * half_n + (j >> 1)) << 2) + (k & 1) + ((j & 1) << 1)];
    }
};

__device__ __forceinline__ TxMatrix tx_matrix(const uint32_t *smat, int type, int log2n)
{
    TxMatrix m;
    m.bytes = reinterpret_cast<const int8_t *>(tx_packed(smat, type, log2n));
    m.half_n = 1 << (log2n - 1);
    return m;
}

__device__ __forceinline__ void load_matrices(uint32_t *smat)
{
    const uint4 *src = reinterpret_cast<const uint4 *>(VVC_TX_PACKED);
    uint4 *dst = reinterpret_cast<uint4 *>(smat);
    for (int i = threadIdx.x; i < kPackedWords / 4; i += blockDim.x)
        dst[i] = src[i];
}

// The 2-D / 1-D inverse transform of one block by one warp (vvcdsp.c:94-138).
//   in : non-zero box, element (y, x) at in[y * in_stride + x], x < nzw, y < nzh
//   mid: scratch for the column-pass result (h x nzw)
//   emit(y, x, r) receives every residual sample of the w x h block exactly once.
template <typename T, typename Emit>
__device__ __forceinline__ void inverse_transform(const T *in, int in_stride, T *mid, const uint32_t *smat,
                                                  int trh, int trv, int log2w, int log2h, int nzw, int nzh,
                                                  int range, int bd, int lane, Emit emit)
{
    const int w = 1 << log2w, h = 1 << log2h;
    const bool dconly = trh == VVCB200_DCT2 && trv == VVCB200_DCT2 && nzw == 1 && nzh == 1;

    if (w > 1 && h > 1) {
        const int sh1 = 5 + range - bd;
        if (dconly && w == h) {
            const int t  = (int)((unsigned)in[0] * 64u + 64u) >> 7;
            const int dc = (int)((unsigned)t * 64u + (1u << (sh1 - 1))) >> sh1;
            for (int i = lane; i < w * h; i += 32)
                emit(i >> log2w, i & (w - 1), dc);
            return;
        }
        const TxMatrix mv = tx_matrix(smat, trv, log2h);
        for (int i = lane; i < h * nzw; i += 32) {
            const int j = i / nzw, x = i - j * nzw;
            unsigned acc = 0;
            for (int k = 0; k < nzh; k++)
                acc += (unsigned)(int)in[k * in_stride + x] * (unsigned)mv.at(k, j);
            mid[i] = (T)clip_intp2(((int)acc + 64) >> 7, range);
        }
        __syncwarp();
        const TxMatrix mh = tx_matrix(smat, trh, log2w);
        for (int i = lane; i < w * h; i += 32) {
            const int y = i >> log2w, x = i & (w - 1);
            unsigned acc = 0;
            for (int k = 0; k < nzw; k++)
                acc += (unsigned)(int)mid[y * nzw + k] * (unsigned)mh.at(k, x);
            emit(y, x, (int)(acc + (1u << (sh1 - 1))) >> sh1);
        }
    } else {
        const int sh = 6 + range - bd;
        const int n = w * h;                       // the 1-D length
        if (dconly) {
            const int dc = (int)((unsigned)in[0] * 64u + (1u << (sh - 1))) >> sh;
            for (int i = lane; i < n; i += 32)
                emit(w > 1 ? 0 : i, w > 1 ? i : 0, dc);
            return;
        }
        const int type = w > 1 ? trh : trv;
        const int nz   = w > 1 ? nzw : nzh;
        const int step = w > 1 ? 1 : in_stride;
        const TxMatrix m = tx_matrix(smat, type, w > 1 ? log2w : log2h);
        for (int i = lane; i < n; i += 32) {
            unsigned acc = 0;
            for (int k = 0; k < nz; k++)
                acc += (unsigned)(int)in[k * step] * (unsigned)m.at(k, i);
            emit(w > 1 ? 0 : i, w > 1 ? i : 0, (int)(acc + (1u << (sh - 1))) >> sh);
        }
    }
}

// ---- fast path: dp2a register-tiled passes ---------------------------------------------------------------------
// One 1-D pass for one lane: R consecutive outputs (first output j0, even) from n_kp input pairs.
//   src : this lane's input pairs, word kp = (x[2kp], x[2kp+1]) as int16x2
//   mat : packed matrix of the pass, row kp has half_n words, word jp serves outputs 2jp (lo bytes) and 2jp+1 (hi bytes)
// 16-bit inputs x 8-bit matrix entries accumulate exactly in int32 (32 terms * 2^15 * 90 < 2^31).
template <int R>
__device__ __forceinline__ void dp2a_pass(int (&acc)[R], const uint32_t *src, int n_kp, const uint32_t *mat, int half_n, int j0)
{
#pragma unroll
    for (int r = 0; r < R; r++)
        acc[r] = 0;
    const uint32_t *m = mat + (j0 >> 1);
    for (int kp = 0; kp < n_kp; kp++, m += half_n) {
        const int a = (int)src[kp];
        if constexpr (R >= 8) {
#pragma unroll
            for (int q = 0; q < R / 8; q++) {
                const uint4 w = *reinterpret_cast<const uint4 *>(m + 4 * q);
                acc[8 * q + 0] = __dp2a_lo(a, (int)w.x, acc[8 * q + 0]);
                acc[8 * q + 1] = __dp2a_hi(a, (int)w.x, acc[8 * q + 1]);
                acc[8 * q + 2] = __dp2a_lo(a, (int)w.y, acc[8 * q + 2]);
                acc[8 * q + 3] = __dp2a_hi(a, (int)w.y, acc[8 * q + 3]);
                acc[8 * q + 4] = __dp2a_lo(a, (int)w.z, acc[8 * q + 4]);
                acc[8 * q + 5] = __dp2a_hi(a, (int)w.z, acc[8 * q + 5]);
                acc[8 * q + 6] = __dp2a_lo(a, (int)w.w, acc[8 * q + 6]);
                acc[8 * q + 7] = __dp2a_hi(a, (int)w.w, acc[8 * q + 7]);
            }
        } else if constexpr (R == 4) {
            const uint2 w = *reinterpret_cast<const uint2 *>(m);
            acc[0] = __dp2a_lo(a, (int)w.x, acc[0]);
            acc[1] = __dp2a_hi(a, (int)w.x, acc[1]);
            acc[2] = __dp2a_lo(a, (int)w.y, acc[2]);
            acc[3] = __dp2a_hi(a, (int)w.y, acc[3]);
        } else {
            const uint32_t w = m[0];
            acc[0] = __dp2a_lo(a, (int)w, acc[0]);
            acc[1] = __dp2a_hi(a, (int)w, acc[1]);
        }
    }
}

// Column pass of one block: the lane computes outputs j0 .. j0 + R - 1 of column x and stores them,
// scaled and clipped to 16 bits, into mid[j][x] (scale_clip, vvcdsp.c:67-80).
template <int R>
__device__ __forceinline__ void fast_column_pass(const int16_t *inT, int16_t *mid, const uint32_t *mat, int half_n,
                                                 int x, int j0, int nzw, int n_kp, int range)
{
    int acc[R];
    dp2a_pass<R>(acc, reinterpret_cast<const uint32_t *>(inT + x * kPairStride), n_kp, mat, half_n, j0);
#pragma unroll
    for (int r = 0; r < R; r++) {
        const int j = j0 + r;
        mid[j * kPairStride + x] = (int16_t)clip_intp2((acc[r] + 64) >> 7, range);
        if ((nzw & 1) && x == nzw - 1)
            mid[j * kPairStride + nzw] = 0;          // pad the last input pair of the row pass
    }
}

// Row pass: the lane computes outputs x0 .. x0 + R - 1 of row y and hands them to the epilogue in
// groups of two adjacent samples.
template <int R, typename Epi>
__device__ __forceinline__ void fast_row_pass(const int16_t *mid, const uint32_t *mat, int half_n, int y, int x0,
                                              int n_kp, int sh1, const Epi &ep)
{
    // The R destination samples of this lane (adjacent, in row y) are fetched BEFORE the arithmetic, in the widest
    // pieces their address allows, so that the load latency hides behind the row transform instead of stalling the
    // read-modify-write at its end (the kernel's main stall, see profiles/).
    constexpr int W = R / 2;                                               // 32-bit words
    uint32_t old[W];
    const bool rmw = ep.plain();
    uint32_t *p = reinterpret_cast<uint32_t *>(ep.dst + y * ep.stride + x0);
    const unsigned al = (unsigned)(uintptr_t)p;
    const int vec = (W >= 4 && !(al & 15)) ? 4 : (W >= 2 && !(al & 7)) ? 2 : 1;
    if (rmw) {
        if (vec == 4) {
#pragma unroll
            for (int i = 0; i < W / 4; i++) {
                const uint4 v = reinterpret_cast<const uint4 *>(p)[i];
                old[4 * i] = v.x; old[4 * i + 1] = v.y; old[4 * i + 2] = v.z; old[4 * i + 3] = v.w;
            }
        } else if (vec == 2) {
#pragma unroll
            for (int i = 0; i < W / 2; i++) {
                const uint2 v = reinterpret_cast<const uint2 *>(p)[i];
                old[2 * i] = v.x; old[2 * i + 1] = v.y;
            }
        } else {
#pragma unroll
            for (int i = 0; i < W; i++)
                old[i] = p[i];
        }
    }
    int acc[R];
    dp2a_pass<R>(acc, reinterpret_cast<const uint32_t *>(mid + y * kPairStride), n_kp, mat, half_n, x0);
    const int add = 1 << (sh1 - 1);
    if (!rmw) {
#pragma unroll
        for (int r = 0; r < R; r += 2)
            ep.two(y, x0 + r, (acc[r] + add) >> sh1, (acc[r + 1] + add) >> sh1);
        return;
    }
#pragma unroll
    for (int i = 0; i < W; i++) {
        const int r0 = (acc[2 * i] + add) >> sh1, r1 = (acc[2 * i + 1] + add) >> sh1;
        old[i] = (uint32_t)clip_pixel((int)(old[i] & 0xffff) + r0, ep.bd) | ((uint32_t)clip_pixel((int)(old[i] >> 16) + r1, ep.bd) << 16);
    }
    if (vec == 4) {
#pragma unroll
        for (int i = 0; i < W / 4; i++)
            reinterpret_cast<uint4 *>(p)[i] = make_uint4(old[4 * i], old[4 * i + 1], old[4 * i + 2], old[4 * i + 3]);
    } else if (vec == 2) {
#pragma unroll
        for (int i = 0; i < W / 2; i++)
            reinterpret_cast<uint2 *>(p)[i] = make_uint2(old[2 * i], old[2 * i + 1]);
    } else {
#pragma unroll
        for (int i = 0; i < W; i++)
            p[i] = old[i];
    }
}

// Table 38 (vvc_intra.c:342-362)
__device__ __forceinline__ int scaling_id(int inter, int c_idx, int log2max)
{
    const int base = log2max == 1 ? 0 : 2 + (log2max - 2) * 6;      // 0, 2, 8, 14, 20, 26
    if (log2max == 1)
        return inter && c_idx == 2 ? 1 : 0;
    if (log2max == 6)
        return c_idx ? (inter ? 24 + (c_idx == 2) : 20 + c_idx) : 26 + inter;
    return base + (inter ? 3 : 0) + c_idx;
}

__device__ __forceinline__ int lmcs_scale_residual(int r, int scale, int bd)
{
    const int c = clip_intp2(r, bd);
    const int a = abs(c);
    const int s = (a * scale + (1 << 10)) >> 11;
    return c < 0 ? -s : s;
}

// dequantisation of one level (scale_coeff, vvc_intra.c:389-395); products wrap mod 2^32 like the reference's int
struct Dequant {
    unsigned scale, bd_off;
    int bd_shift, id, l2m, log2_w, log2_h, range;
    bool use_sl;
    __device__ __forceinline__ int apply(int c, int x, int y, const ScalingLists &sl) const
    {
        unsigned m = 16;
        if (use_sl) {
            const int off = ((y << l2m) >> log2_h) << l2m;
            m = sl.rec[id][off + ((x << l2m) >> log2_w)];
            if (id >= 14 && !(x | y))
                m = sl.dc[id - 14];
        }
        const int v = (int)((unsigned)c * scale * m + bd_off) >> bd_shift;
        return clip_intp2(v, range);
    }
};

__device__ __forceinline__ Dequant make_dequant(const VVCB200TransformBlock &d, bool ts, int bd, int range)
{
    Dequant q;
    const int log_sum = d.log2_w + d.log2_h;
    const int rect = ts ? 0 : (log_sum & 1);
    const int dq = (d.flags & VVCB200_TB_DEP_QUANT) ? 1 : 0;
    q.bd_shift = ts ? 10 : bd + rect + (log_sum >> 1) + 10 - range + dq;
    q.bd_off = (1u << q.bd_shift) >> 1;
    const int qp = d.qp + (ts ? 0 : dq);
    const int rem = qp % 6;
    const int ls0 = rem == 0 ? 40 : rem == 1 ? 45 : rem == 2 ? 51 : rem == 3 ? 57 : rem == 4 ? 64 : 72;
    const int ls1 = rem == 0 ? 57 : rem == 1 ? 64 : rem == 2 ? 72 : rem == 3 ? 80 : rem == 4 ? 90 : 102;
    q.scale = (unsigned)(rect ? ls1 : ls0) << (qp / 6);
    q.use_sl = (d.flags & VVCB200_TB_SCALING) && !ts;
    q.id = scaling_id((d.flags & VVCB200_TB_INTER) != 0, d.c_idx, max(d.log2_w, d.log2_h));
    q.l2m = q.id < 2 ? 1 : q.id < 8 ? 2 : 3;
    q.log2_w = d.log2_w;
    q.log2_h = d.log2_h;
    q.range = range;
    return q;
}

// Everything the epilogue of one block needs (add_residual / joint CbCr / LMCS chroma scale, vvc_intra.c:166-186,451-477)
struct Epilogue {
    uint16_t *dst, *dst2;       // first sample of the block in its plane / in the joint plane (pixels or int16 residual)
    int stride, stride2;
    int bd, c_sign, jshift, lscale;
    bool joint, lmcs, to_residual;

    __device__ __forceinline__ void one(int y, int x, int r) const
    {
        int r2 = joint ? (r * c_sign) >> jshift : 0;
        if (lmcs) {
            r = lmcs_scale_residual(r, lscale, bd);
            r2 = lmcs_scale_residual(r2, lscale, bd);
        }
        if (to_residual) {
            reinterpret_cast<int16_t *>(dst)[y * stride + x] = (int16_t)clip_intp2(r, 15);
            if (joint)
                reinterpret_cast<int16_t *>(dst2)[y * stride2 + x] = (int16_t)clip_intp2(r2, 15);
        } else {
            dst[y * stride + x] = (uint16_t)clip_pixel(dst[y * stride + x] + r, bd);
            if (joint)
                dst2[y * stride2 + x] = (uint16_t)clip_pixel(dst2[y * stride2 + x] + r2, bd);
        }
    }
    // plain add into the picture (no joint plane, no chroma scaling, not into the residual planes)
    __device__ __forceinline__ bool plain() const { return !(joint || lmcs || to_residual); }
    // two horizontally adjacent samples: one 32-bit read-modify-write (x even, planes 4-byte aligned)
    __device__ __forceinline__ void two(int y, int x, int r0, int r1) const
    {
        if (joint || lmcs || to_residual) {
            one(y, x, r0);
            one(y, x + 1, r1);
            return;
        }
        uint32_t *p = reinterpret_cast<uint32_t *>(dst + y * stride + x);
        const uint32_t v = *p;
        *p = (uint32_t)clip_pixel((int)(v & 0xffff) + r0, bd) | ((uint32_t)clip_pixel((int)(v >> 16) + r1, bd) << 16);
    }
};

__device__ __forceinline__ Epilogue make_epilogue(const ResidualArgs &a, const VVCB200TransformBlock &d)
{
    Epilogue e;
    const bool to_res = a.to_residual || (d.flags & VVCB200_TB_TO_RESIDUAL);
    const PlanesView &pv = to_res ? a.res : a.dst;
    const int other = d.c_idx ? 3 - d.c_idx : 0;
    e.stride = pv.stride[d.c_idx];
    e.stride2 = pv.stride[other];
    e.dst = pv.p[d.c_idx] + (size_t)d.y0 * e.stride + d.x0;
    e.dst2 = pv.p[other] + (size_t)d.y0 * e.stride2 + d.x0;
    e.bd = a.bd;
    e.c_sign = a.c_sign;
    e.jshift = (d.flags & VVCB200_TB_JOINT_SHIFT) ? 1 : 0;
    e.lscale = d.lmcs_scale;
    e.joint = (d.flags & VVCB200_TB_JOINT) && d.c_idx;
    e.lmcs = d.flags & VVCB200_TB_LMCS_SCALE;
    e.to_residual = to_res;
    return e;
}

// ---- fast classes: `lpt` lanes per block (32, 16, 8 or 4), all blocks of the warp transformed side by side ---------
// li = lane inside the block's lane group; inT / fmid = this block's staging areas.  Called by all lanes of the warp.
__device__ __forceinline__ void fast_block(const ResidualArgs &a, const ScalingLists &sl, const VVCB200TransformBlock &d,
                                           bool active, int li, int lpt, int log2lpt, int16_t *inT, int16_t *fmid,
                                           const uint32_t *smat)
{
    const int w = 1 << d.log2_w, h = 1 << d.log2_h;
    const int nzw = d.nzw, nzh = d.nzh;
    const Epilogue ep = make_epilogue(a, d);
    const int sh1 = 5 + a.range - a.bd;
    const bool dconly = d.trh == VVCB200_DCT2 && d.trv == VVCB200_DCT2 && nzw == 1 && nzh == 1 && w == h;

    if (active && !ep.to_residual && li < h)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(ep.dst + li * ep.stride));
    // levels: coalesced fetch into the (still free) middle buffer
    if (active)
        fetch_levels(a, d, fmid, li, lpt);
    __syncwarp();
    if (active && dconly) {
        // DC-only square DCT-II block (vvcdsp.c:101-108): one value for the whole block
        const Dequant q = make_dequant(d, false, a.bd, a.range);
        const int c = fmid[0] ? q.apply(fmid[0], 0, 0, sl) : 0;
        const int t = (c * 64 + 64) >> 7;
        const int dc = (t * 64 + (1 << (sh1 - 1))) >> sh1;
        for (int i = li; i < (w * h) >> 1; i += lpt) {
            const int y = i >> (d.log2_w - 1), x = (i & ((w >> 1) - 1)) << 1;
            ep.two(y, x, dc, dc);
        }
    }
    // dequant + transpose: lane x owns column x of the non-zero box
    if (active && !dconly && li < nzw) {
        const Dequant q = make_dequant(d, false, a.bd, a.range);
        for (int y = 0; y < nzh; y++) {
            const int c = fmid[y * nzw + li];
            inT[li * kPairStride + y] = (int16_t)(c ? q.apply(c, li, y, sl) : 0);
        }
        if (nzh & 1)
            inT[li * kPairStride + nzh] = 0;
    }
    __syncwarp();
    if (active && !dconly) {
        // column pass: lane = g * cw + x owns column x and outputs [g * per_lane, (g + 1) * per_lane)
        const int log2cw = nzw <= 1 ? 0 : 32 - __clz(nzw - 1);             // nzw rounded up to a power of two
        const int lg = min(log2lpt - log2cw, d.log2_h - 1);                // log2(lane groups per column)
        const int per_lane = h >> lg;                                      // even
        const int x = li & ((1 << log2cw) - 1), g = li >> log2cw;
        const uint32_t *mat = tx_packed(smat, d.trv, d.log2_h);
        const int n_kp = (nzh + 1) >> 1, half = h >> 1;
        if (x < nzw && g < (1 << lg)) {
            switch (per_lane) {                    // at most 16 accumulators live at a time (registers = occupancy)
            case 2:  fast_column_pass<2>(inT, fmid, mat, half, x, g * 2, nzw, n_kp, a.range); break;
            case 4:  fast_column_pass<4>(inT, fmid, mat, half, x, g * 4, nzw, n_kp, a.range); break;
            case 8:  fast_column_pass<8>(inT, fmid, mat, half, x, g * 8, nzw, n_kp, a.range); break;
            default:
                for (int j0 = g * per_lane; j0 < (g + 1) * per_lane; j0 += 16)
                    fast_column_pass<16>(inT, fmid, mat, half, x, j0, nzw, n_kp, a.range);
                break;
            }
        }
    }
    __syncwarp();
    if (active && !dconly) {
        // row pass + epilogue: lane = g * rows + y owns row y and outputs [g * per_lane, (g + 1) * per_lane)
        const int log2rows = min((int)d.log2_h, log2lpt);
        const int lg = min(log2lpt - log2rows, d.log2_w - 1);
        const int per_lane = w >> lg;
        const int yl = li & ((1 << log2rows) - 1), g = li >> log2rows;
        const uint32_t *mat = tx_packed(smat, d.trh, d.log2_w);
        const int n_kp = (nzw + 1) >> 1, half = w >> 1;
        if (g < (1 << lg)) {
            for (int y = yl; y < h; y += lpt) {
                switch (per_lane) {
                case 2:  fast_row_pass<2>(fmid, mat, half, y, g * 2, n_kp, sh1, ep); break;
                case 4:  fast_row_pass<4>(fmid, mat, half, y, g * 4, n_kp, sh1, ep); break;
                case 8:  fast_row_pass<8>(fmid, mat, half, y, g * 8, n_kp, sh1, ep); break;
                default:
                    for (int x0 = g * per_lane; x0 < (g + 1) * per_lane; x0 += 16)
                        fast_row_pass<16>(fmid, mat, half, y, x0, n_kp, sh1, ep);
                    break;
                }
            }
        }
    }
}

// ---- generic class: TS, BDPCM, LFNST, 1-D blocks — one warp per block ---------------------------------------------
__device__ __forceinline__ void generic_block(const ResidualArgs &a, const ScalingLists &sl, const VVCB200TransformBlock &d,
                                              int lane, int16_t *in, int16_t *mid, const uint32_t *smat)
{
    const int w = 1 << d.log2_w, h = 1 << d.log2_h;
    const bool ts = d.flags & VVCB200_TB_TS;
    int nzw = d.nzw, nzh = d.nzh, in_stride = d.nzw;
    const Dequant q = make_dequant(d, ts, a.bd, a.range);
    const Epilogue ep = make_epilogue(a, d);
    auto emit = [&](int y, int x, int r) { ep.one(y, x, r); };

    // ---- levels into shared memory (+ BDPCM prefix sums on the raw levels)
    if (d.flags & VVCB200_TB_BDPCM) {
        in_stride = w;
        fetch_levels(a, d, mid, lane, 32);                // the scan box first (mid is free until the column pass)
        __syncwarp();
        for (int i = lane; i < w * h; i += 32) {
            const int y = i >> d.log2_w, x = i & (w - 1);
            in[i] = (x < nzw && y < nzh) ? mid[y * nzw + x] : (int16_t)0;
        }
        __syncwarp();
        if (d.flags & VVCB200_TB_BDPCM_VERT) {
            for (int x = lane; x < w; x += 32) {
                int run = in[x];
                for (int y = 1; y < h; y++) {
                    run = clip_intp2(run + in[y * w + x], a.range);
                    in[y * w + x] = (int16_t)run;
                }
            }
            nzh = h;
        } else {
            for (int y = lane; y < h; y += 32) {
                int run = in[y * w];
                for (int x = 1; x < w; x++) {
                    run = clip_intp2(run + in[y * w + x], a.range);
                    in[y * w + x] = (int16_t)run;
                }
            }
            nzw = w;
        }
        __syncwarp();
    } else {
        fetch_levels(a, d, in, lane, 32);
        __syncwarp();
    }
    // ---- dequant (vvc_intra.c:277-417)
    for (int i = lane; i < nzw * nzh; i += 32) {
        const int y = i / nzw, x = i - y * nzw;
        const int c = in[y * in_stride + x];
        if (c)
            in[y * in_stride + x] = (int16_t)q.apply(c, x, y, sl);
    }
    __syncwarp();
    // ---- inverse LFNST (vvc_intra.c:65-127, vvc_itx_1d.c:708-721)
    if (d.lfnst && !ts) {
        const int idx = (d.lfnst & 3) - 1, set = (d.lfnst >> 2) & 3;
        const bool transpose = (d.lfnst >> 4) & 1;
        const bool big = w >= 8 && h >= 8;
        const int n_out = big ? 48 : 16, size = big ? 8 : 4;
        const int nz = ((w == 8 && h == 8) || (w == 4 && h == 4)) ? 8 : 16;
        const int8_t *m = big ? VVC_LFNST_8X8 + (set * 2 + idx) * 16 * 48 : VVC_LFNST_4X4 + (set * 2 + idx) * 16 * 16;
        if (lane < 16) {
            const int p = VVC_DIAG_SCAN_4X4[lane];
            const int x = p & 3, y = p >> 2;
            mid[lane] = (lane < nz && x < nzw && y < nzh) ? in[y * in_stride + x] : (int16_t)0;
        }
        __syncwarp();
        int out[2];
        for (int r = 0; r < 2; r++) {
            const int i = lane + 32 * r;             // raster position in the size x size region
            int y = i / size, x = i - y * size;
            out[r] = 0;
            if (i < size * size) {
                if (transpose) { const int sw = x; x = y; y = sw; }
                const int j = size == 4 ? y * 4 + x : (y < 4 ? y * 8 + x : (x < 4 ? 32 + (y - 4) * 4 + x : -1));
                if (j >= 0) {
                    int acc = 0;
                    for (int k = 0; k < nz; k++)
                        acc += mid[k] * m[k * n_out + j];
                    out[r] = clip_intp2((acc + 64) >> 7, a.range);
                }
            }
        }
        __syncwarp();
        for (int r = 0; r < 2; r++)
            if (lane + 32 * r < size * size)
                in[lane + 32 * r] = (int16_t)out[r];
        nzw = nzh = in_stride = size;
        __syncwarp();
    }
    // ---- inverse transform + epilogue
    if (ts) {
        for (int i = lane; i < w * h; i += 32) {
            const int y = i >> d.log2_w, x = i & (w - 1);
            emit(y, x, (x < nzw && y < nzh) ? (int)in[y * in_stride + x] : 0);
        }
    } else {
        inverse_transform<int16_t>(in, in_stride, mid, smat, d.trh, d.trv, d.log2_w, d.log2_h, nzw, nzh,
                                   a.range, a.bd, lane, emit);
    }
}

// Does block d take the dp2a path, and with how many lanes?  Returns the class (0: 32 lanes ... 3: 4 lanes, 4: generic).
// The host orders the batch by this class (vvcb200_residual_class, exported for emitters).
__host__ __device__ inline int residual_class(const VVCB200TransformBlock &d)
{
    const bool fast = !(d.flags & (VVCB200_TB_TS | VVCB200_TB_BDPCM)) && !d.lfnst && d.log2_w >= 1 && d.log2_h >= 1;
    if (!fast)
        return 4;
    const int m = d.log2_w > d.log2_h ? d.log2_w : d.log2_h;
    return m >= 5 ? 0 : m == 4 ? 1 : m == 3 ? 2 : 3;
}

// One launch per class (kCls 0 .. 3: fast blocks of 32 / 16 / 8 / 4 lanes, 4: the generic ones), blocks a.first .. a.last,
// side by side on forked streams: each kernel carries only its own path — the fast kernels no transform-skip / BDPCM /
// LFNST code and no passes of sizes their class never meets, the generic one no packed fast passes — which matters for
// a kernel that spends a quarter of its stall samples on instruction fetch.
template <int kCls>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 7)      // 7 CTAs/SM is also the shared-memory limit (32 KB each)
residual_kernel(const ResidualArgs a, const ScalingLists sl)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint32_t *smat = reinterpret_cast<uint32_t *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int16_t *buf = reinterpret_cast<int16_t *>(smem_raw + kMatPad) + warp * kWarpElems;

    load_matrices(smat);
    __syncthreads();

    const int wi = blockIdx.x * kWarpsPerCta + warp;
    if (kCls < 4) {
        constexpr int cls = kCls < 4 ? kCls : 0;
        const int first = a.first + (wi << cls), last = a.last;
        if (first >= last)
            return;
        constexpr int log2lpt = 5 - cls, lpt = 32 >> cls;
        const int g = lane >> log2lpt, li = lane & (lpt - 1);
        const bool active = first + g < last;
        VVCB200TransformBlock d = a.tbs[active ? first + g : first];
        // the class bounds the block's larger side (vvcb200_residual_class): say so, and the passes of larger sizes go
        constexpr int max_log2 = cls == 0 ? 6 : 5 - cls;
        d.log2_w = min((int)d.log2_w, max_log2);
        d.log2_h = min((int)d.log2_h, max_log2);
        // per-block staging: transposed input (lpt columns) + middle block (lpt rows; 64 for the 32-lane class)
        constexpr int in_sz = lpt * kPairStride, region = cls == 0 ? kWarpElems : 2 * in_sz;
        fast_block(a, sl, d, active, li, lpt, log2lpt, buf + g * region, buf + g * region + in_sz, smat);
    } else {
        const int t = a.first + wi;
        if (t < a.last)
            generic_block(a, sl, a.tbs[t], lane, buf, buf + kInElems, smat);
    }
}

// ---- Level A: itx.itx[...] on int32 blocks in place, add_residual on int32 blocks --------------------
struct ItxArgs {
    int32_t *coeffs;
    const uint32_t *block_off;
    const VVCB200TransformBlock *tbs;
    int n, range, bd;
};

__global__ void __launch_bounds__(kWarpsPerCta * 32)
itx_kernel(const ItxArgs a)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint32_t *smat = reinterpret_cast<uint32_t *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int32_t *in  = reinterpret_cast<int32_t *>(smem_raw + kMatPad) + warp * (kInElems + kMidElems);
    int32_t *mid = in + kInElems;

    load_matrices(smat);
    __syncthreads();
    for (int t = blockIdx.x + warp * gridDim.x; t < a.n; t += kWarpsPerCta * gridDim.x) {
        const VVCB200TransformBlock d = a.tbs[t];
        const int w = 1 << d.log2_w;
        int32_t *blk = a.coeffs + a.block_off[t];
        __syncwarp();
        for (int i = lane; i < d.nzw * d.nzh; i += 32) {
            const int y = i / d.nzw, x = i - y * d.nzw;
            in[i] = blk[y * w + x];
        }
        __syncwarp();
        inverse_transform<int32_t>(in, d.nzw, mid, smat, d.trh, d.trv, d.log2_w, d.log2_h, d.nzw, d.nzh,
                                   a.range, a.bd, lane, [&](int y, int x, int r) { blk[y * w + x] = r; });
    }
}

__global__ void add_residual_kernel(PlanesView dst, const int32_t *res, const uint32_t *block_off,
                                    const VVCB200TransformBlock *tbs, int n, int bd)
{
    const int t = blockIdx.x;
    if (t >= n)
        return;
    const VVCB200TransformBlock d = tbs[t];
    const int w = 1 << d.log2_w, h = 1 << d.log2_h;
    const int32_t *r = res + block_off[t];
    for (int i = threadIdx.x; i < w * h; i += blockDim.x) {
        const int y = i >> d.log2_w, x = i & (w - 1);
        uint16_t *p = dst.p[d.c_idx] + (size_t)(d.y0 + y) * dst.stride[d.c_idx] + d.x0 + x;
        *p = (uint16_t)clip_pixel(*p + r[i], bd);
    }
}

static int check_bd(int bd)
{
    if (bd != 8 && bd != 10 && bd != 12) {
        set_error("unsupported bit depth %d (8, 10 or 12)", bd);
        return VVCB200_ERR_ARG;
    }
    return 0;
}

// dynamic shared memory above 48 KB needs a per-device opt-in; done once per device per kernel
template <typename K>
static int opt_in_smem(K kernel, size_t bytes, bool *ready /* [64] */)
{
    int dev = 0;
    VVC_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && ready[dev])
        return 0;
    VVC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    if (dev >= 0 && dev < 64)
        ready[dev] = true;
    return 0;
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_residual_batch(const VVCB200Planes *dst, const VVCB200Planes *d_residual,
                                      const int16_t *d_levels, const VVCB200TransformBlock *d_tbs, int n,
                                      const VVCB200ResidualParams *params, void *stream)
{
    if (!params || (!dst && !d_residual) || n < 0 || (n && (!d_levels || !d_tbs))) {
        set_error("vvcb200_residual_batch: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (int r = check_bd(params->bit_depth))
        return r;
    if (params->log2_transform_range != 15) {
        set_error("log2_transform_range %d not supported (15 only)", params->log2_transform_range);
        return VVCB200_ERR_UNSUP;
    }
    if (n == 0)
        return 0;
    ResidualArgs a = {};
    if (dst)
        a.dst = make_view(dst);
    if (d_residual)
        a.res = make_view(d_residual);
    a.to_residual = dst == nullptr;          // otherwise only the TBs flagged VVCB200_TB_TO_RESIDUAL go to the residual planes
    a.levels = d_levels;
    a.tbs = d_tbs;
    a.n = n;
    a.bd = params->bit_depth;
    a.range = params->log2_transform_range;
    a.c_sign = params->joint_c_sign;
    a.sparse = params->levels_sparse != 0;
    if (a.sparse && ((uintptr_t)d_levels & 3)) {
        set_error("vvcb200_residual_batch: a sparse level buffer must be 4-byte aligned");
        return VVCB200_ERR_ARG;
    }
    ScalingLists sl;
    memcpy(sl.rec, params->scaling_matrix_rec, sizeof(sl.rec));
    memset(sl.dc, 16, sizeof(sl.dc));
    memcpy(sl.dc, params->scaling_matrix_dc_rec, 14);

    int rest = n, warps = 0;
    for (int c = 0; c < 4; c++) {
        a.n_class[c] = params->n_class[c];
        if (a.n_class[c] < 0 || a.n_class[c] > rest) {
            set_error("vvcb200_residual_batch: n_class[%d] = %d does not fit the batch of %d blocks", c, a.n_class[c], n);
            return VVCB200_ERR_ARG;
        }
        rest -= a.n_class[c];
        warps += (a.n_class[c] + (1 << c) - 1) >> c;
    }
    const size_t smem = kMatPad + kWarpsPerCta * kWarpElems * sizeof(int16_t);
    static bool ready[5][64];
    if (int r = opt_in_smem(residual_kernel<0>, smem, ready[0])) return r;
    if (int r = opt_in_smem(residual_kernel<1>, smem, ready[1])) return r;
    if (int r = opt_in_smem(residual_kernel<2>, smem, ready[2])) return r;
    if (int r = opt_in_smem(residual_kernel<3>, smem, ready[3])) return r;
    if (int r = opt_in_smem(residual_kernel<4>, smem, ready[4])) return r;
    const cudaStream_t main_stream = (cudaStream_t)stream;
    int classes = 0;
    for (int c = 0; c < 5; c++)
        classes += (c < 4 ? a.n_class[c] : rest) > 0;
    ForkStreams *f = classes > 1 ? fork_streams(main_stream) : nullptr;        // the classes are independent grids: side by side
    if (f)
        VVC_CUDA(cudaEventRecord(f->fork, main_stream));
    int base = 0, used = 0, launched = 0;
    for (int c = 0; c < 5; c++) {
        const int count = c < 4 ? a.n_class[c] : rest;
        a.first = base;
        a.last = base + count;
        base += count;
        if (!count)
            continue;
        cudaStream_t st = main_stream;                                    // the first class present stays on the caller's stream
        if (f && launched > 0 && used < 4) {
            st = f->side[used];
            VVC_CUDA(cudaStreamWaitEvent(st, f->fork, 0));
        }
        const int cw = c < 4 ? (count + (1 << c) - 1) >> c : count;     // generic blocks: one warp each
        const int grid = (cw + kWarpsPerCta - 1) / kWarpsPerCta;
        switch (c) {
        case 0:  residual_kernel<0><<<grid, kWarpsPerCta * 32, smem, st>>>(a, sl); break;
        case 1:  residual_kernel<1><<<grid, kWarpsPerCta * 32, smem, st>>>(a, sl); break;
        case 2:  residual_kernel<2><<<grid, kWarpsPerCta * 32, smem, st>>>(a, sl); break;
        case 3:  residual_kernel<3><<<grid, kWarpsPerCta * 32, smem, st>>>(a, sl); break;
        default: residual_kernel<4><<<grid, kWarpsPerCta * 32, smem, st>>>(a, sl); break;
        }
        VVC_LAUNCH_CHECK();
        launched++;
        if (st != main_stream) {
            VVC_CUDA(cudaEventRecord(f->join[used], st));
            VVC_CUDA(cudaStreamWaitEvent(main_stream, f->join[used], 0));
            used++;
        }
    }
    return 0;
}

extern "C" int vvcb200_residual_class(const VVCB200TransformBlock *tb)
{
    return tb ? residual_class(*tb) : VVCB200_ERR_ARG;
}

extern "C" int vvcb200_dsp_itx_batch(int32_t *d_coeffs, const uint32_t *d_block_off,
                                     const VVCB200TransformBlock *d_tbs, int n,
                                     int log2_transform_range, int bit_depth, void *stream)
{
    if (n < 0 || (n && (!d_coeffs || !d_block_off || !d_tbs))) {
        set_error("vvcb200_dsp_itx_batch: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (int r = check_bd(bit_depth))
        return r;
    if (n == 0)
        return 0;
    ItxArgs a = { d_coeffs, d_block_off, d_tbs, n, log2_transform_range, bit_depth };
    const size_t smem = kMatPad + kWarpsPerCta * (kInElems + kMidElems) * sizeof(int32_t);
    static bool ready[64];
    if (int r = opt_in_smem(itx_kernel, smem, ready))
        return r;
    itx_kernel<<<(n + kChunk - 1) / kChunk, kWarpsPerCta * 32, smem, (cudaStream_t)stream>>>(a);
    VVC_LAUNCH_CHECK();
    return 0;
}

extern "C" int vvcb200_dsp_add_residual_batch(const VVCB200Planes *dst, const int32_t *d_res, const uint32_t *d_block_off,
                                              const VVCB200TransformBlock *d_tbs, int n, int bit_depth, void *stream)
{
    if (!dst || n < 0 || (n && (!d_res || !d_block_off || !d_tbs))) {
        set_error("vvcb200_dsp_add_residual_batch: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (int r = check_bd(bit_depth))
        return r;
    if (n == 0)
        return 0;
    add_residual_kernel<<<n, 128, 0, (cudaStream_t)stream>>>(make_view(dst), d_res, d_block_off, d_tbs, n, bit_depth);
    VVC_LAUNCH_CHECK();
    return 0;
}
