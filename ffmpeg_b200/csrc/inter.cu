// inter.cu — stage INTER: motion-compensated prediction of a picture's sub-block batch.
// Reference semantics: 8-tap luma / 4-tap chroma separable interpolation (inter.put / put_uni / put_uni_w,
// libavcodec/h26x/h2656_inter_template.c:29-577), avg / w_avg incl. BCW (vvc_inter_template.c:25-58,
// vvc_inter.c:31,145-174), MV -> offset/fraction arithmetic of vvc_inter.c:176-368, replicate-edge fetch of
// ff_emulated_edge_mc (videodsp_template.c:27-105, here a coordinate clamp), forward LMCS (vvc_inter.c:888-891).
//
// Mapping: one warp per sub-block (<= 16x16 luma + its two chroma blocks).  Per list the (h+7)x(w+7)
// reference window is staged once in shared memory with clamped coordinates, filtered horizontally into a
// 16-bit shared tile and vertically into registers; the two lists meet in registers, so every reference
// sample is read from L2/HBM once per block and the prediction is written once.
//
// The four reference variants (pixels / h / v / hv) are folded into one separable form, which is
// bit-identical: a zero fraction selects the identity tap {..,64,..}, and (64*s) >> (bd-8) == s << (14-bd),
// (64*t) >> 6 == t exactly.
#include "common.cuh"

#define VVC_TABLE_QUAL static __device__ const
#include "vvc_tables.h"

namespace vvcb200 {

constexpr int kInterWarps = 8;
constexpr int kWinStride  = 24;                 // 16 + 7 columns, padded
constexpr int kWinRows    = 23;
constexpr int kMaxRefs    = 16;

struct InterArgs {
    PlanesView dst;
    PlanesView tmp14;                            // optional int16 output of the list-0 intermediate
    PlanesView refs[kMaxRefs];
    const VVCB200PredBlock *blocks;
    const uint16_t *lmcs_fwd;
    int n, bd, to_tmp14;
};

__device__ __forceinline__ int bcw_weight(int idx)
{
    return idx == 1 ? 5 : idx == 2 ? 3 : idx == 3 ? 10 : idx == 4 ? -2 : 4;
}

// One component of one block, one warp.  val[] (registers) receives the 14-bit intermediate of this
// list for the lane's output samples i = lane + 32 * k  ->  (y, x) = (i / w, i % w).
template <int TAPS>
__device__ __forceinline__ void interpolate(int (&val)[8], const PlanesView &ref, int c, int xi, int yi, int w, int h,
                                            int log2w, const int8_t *hf, const int8_t *vf, int mx, int my, int bd,
                                            uint16_t *win, int16_t *tmp, int lane)
{
    constexpr int BEFORE = TAPS / 2 - 1;
    const int rows = my ? h + TAPS - 1 : h;
    const int cols = mx ? w + TAPS - 1 : w;
    const int ry0  = yi - (my ? BEFORE : 0);
    const int rx0  = xi - (mx ? BEFORE : 0);
    const uint16_t *src = ref.p[c];
    const int stride = ref.stride[c], pw = ref.w[c], ph = ref.h[c];

    __syncwarp();
    for (int i = lane; i < rows * cols; i += 32) {
        const int r = i / cols, x = i - r * cols;
        const int sy = clip3(ry0 + r, 0, ph - 1), sx = clip3(rx0 + x, 0, pw - 1);
        win[r * kWinStride + x] = __ldg(src + (size_t)sy * stride + sx);
    }
    __syncwarp();
    int f[TAPS];
    if (mx) {
#pragma unroll
        for (int t = 0; t < TAPS; t++)
            f[t] = hf[t];
    }
    for (int i = lane; i < rows * w; i += 32) {
        const int r = i >> log2w, x = i & (w - 1);
        int s;
        if (mx) {
            s = 0;
#pragma unroll
            for (int t = 0; t < TAPS; t++)
                s += f[t] * win[r * kWinStride + x + t];
        } else {
            s = 64 * win[r * kWinStride + x];
        }
        tmp[r * 16 + x] = (int16_t)(s >> (bd - 8));
    }
    __syncwarp();
    if (my) {
#pragma unroll
        for (int t = 0; t < TAPS; t++)
            f[t] = vf[t];
    }
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int i = lane + 32 * k;
        if (i < w * h) {
            const int y = i >> log2w, x = i & (w - 1);
            if (my) {
                int s = 0;
#pragma unroll
                for (int t = 0; t < TAPS; t++)
                    s += f[t] * tmp[(y + t) * 16 + x];
                val[k] = s >> 6;
            } else {
                val[k] = tmp[y * 16 + x];
            }
        }
    }
}

template <int TAPS>
__device__ __forceinline__ void predict_component(const InterArgs &a, const VVCB200InterParams &p, const VVCB200PredBlock &b,
                                                  int c, uint16_t *win, int16_t *tmp, int lane)
{
    const int bd = a.bd;
    const int hs = c ? 1 : 0;
    const int w = b.w >> hs, h = b.h >> hs;
    const int x0 = b.x0 >> hs, y0 = b.y0 >> hs;
    const int log2w = 31 - __clz(w);
    const int set = c ? 0 : (b.flags & VVCB200_PB_AFFINE) ? 2 : (b.flags & VVCB200_PB_HPEL_IF) ? 1 : 0;
    int v0[8], v1[8];

#pragma unroll
    for (int l = 0; l < 2; l++) {
        if (!(b.pred_flag & (1 << l)))
            continue;
        const int mvx = b.mv[l][0], mvy = b.mv[l][1];
        const int mx = c ? (mvx & 31) : (mvx & 15);
        const int my = c ? (mvy & 31) : (mvy & 15);
        const int xi = x0 + (mvx >> (4 + hs)), yi = y0 + (mvy >> (4 + hs));
        const int8_t *hf = c ? VVC_CHROMA_FILTERS + (set * 32 + mx) * 4 : VVC_LUMA_FILTERS + (set * 16 + mx) * 8;
        const int8_t *vf = c ? VVC_CHROMA_FILTERS + (set * 32 + my) * 4 : VVC_LUMA_FILTERS + (set * 16 + my) * 8;
        const PlanesView &ref = a.refs[p.ref_slot[l][b.ref_idx[l]] & (kMaxRefs - 1)];
        if (l == 0)
            interpolate<TAPS>(v0, ref, c, xi, yi, w, h, log2w, hf, vf, mx, my, bd, win, tmp, lane);
        else
            interpolate<TAPS>(v1, ref, c, xi, yi, w, h, log2w, hf, vf, mx, my, bd, win, tmp, lane);
    }

    // ---- combine (weights are warp-uniform)
    int mode;                       // 0 uni, 1 uni_w, 2 avg, 3 w_avg
    int wgt0 = 0, wgt1 = 0, off = 0, sh = 0, ox = 0;
    const int l_uni = (b.pred_flag & 1) ? 0 : 1;
    if (b.pred_flag != 3) {
        sh = 14 - bd;
        if (b.flags & VVCB200_PB_WP) {
            mode = 1;
            wgt0 = p.weight[l_uni][c][b.ref_idx[l_uni]];
            ox   = p.offset[l_uni][c][b.ref_idx[l_uni]] * (1 << (bd - 8));
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
                wgt0 = p.weight[0][c][b.ref_idx[0]];
                wgt1 = p.weight[1][c][b.ref_idx[1]];
                o0   = p.offset[0][c][b.ref_idx[0]];
                o1   = p.offset[1][c][b.ref_idx[1]];
                sh  += p.log2_denom[c > 0];
            }
            off = (((o0 + o1) << (bd - 8)) + 1) << (sh - 1);
        }
    }
    const bool lmcs = c == 0 && (b.flags & VVCB200_PB_LMCS) && a.lmcs_fwd;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int i = lane + 32 * k;
        if (i < w * h) {
            const int y = i >> log2w, x = i & (w - 1);
            if (a.to_tmp14) {
                int16_t *o = reinterpret_cast<int16_t *>(a.tmp14.p[c]) + (size_t)(y0 + y) * a.tmp14.stride[c] + x0 + x;
                *o = (int16_t)(l_uni == 0 ? v0[k] : v1[k]);
                continue;
            }
            int r;
            if (mode == 0)
                r = ((l_uni == 0 ? v0[k] : v1[k]) + off) >> sh;
            else if (mode == 1)
                r = (((l_uni == 0 ? v0[k] : v1[k]) * wgt0 + off) >> sh) + ox;
            else if (mode == 2)
                r = (v0[k] + v1[k] + off) >> sh;
            else
                r = (v0[k] * wgt0 + v1[k] * wgt1 + off) >> sh;
            r = clip_pixel(r, bd);
            if (lmcs)
                r = __ldg(a.lmcs_fwd + r);
            a.dst.p[c][(size_t)(y0 + y) * a.dst.stride[c] + x0 + x] = (uint16_t)r;
        }
    }
}

__global__ void __launch_bounds__(kInterWarps * 32)
inter_kernel(const InterArgs a, const VVCB200InterParams p)
{
    __shared__ __align__(16) uint16_t s_win[kInterWarps][kWinRows * kWinStride];
    __shared__ __align__(16) int16_t  s_tmp[kInterWarps][kWinRows * 16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = blockIdx.x * kInterWarps + warp;
    if (t >= a.n)
        return;
    const VVCB200PredBlock b = a.blocks[t];
    if (!(b.flags & VVCB200_PB_CHROMA_ONLY))
        predict_component<8>(a, p, b, 0, s_win[warp], s_tmp[warp], lane);
    if (!(b.flags & VVCB200_PB_LUMA_ONLY)) {
        predict_component<4>(a, p, b, 1, s_win[warp], s_tmp[warp], lane);
        predict_component<4>(a, p, b, 2, s_win[warp], s_tmp[warp], lane);
    }
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_inter_batch(const VVCB200Planes *dst, const VVCB200Planes *refs, int n_refs,
                                   const VVCB200PredBlock *d_blocks, int n, const VVCB200InterParams *params,
                                   const uint16_t *d_lmcs_fwd_lut, const VVCB200Planes *d_tmp14, void *stream)
{
    if (!params || (!dst && !d_tmp14) || !refs || n_refs <= 0 || n_refs > kMaxRefs || n < 0 || (n && !d_blocks)) {
        set_error("vvcb200_inter_batch: bad argument (n_refs must be 1..%d)", kMaxRefs);
        return VVCB200_ERR_ARG;
    }
    if (params->bit_depth != 8 && params->bit_depth != 10 && params->bit_depth != 12) {
        set_error("unsupported bit depth %d (8, 10 or 12)", params->bit_depth);
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
    a.n = n;
    a.bd = params->bit_depth;
    inter_kernel<<<(n + kInterWarps - 1) / kInterWarps, kInterWarps * 32, 0, (cudaStream_t)stream>>>(a, *params);
    VVC_LAUNCH_CHECK();
    return 0;
}
