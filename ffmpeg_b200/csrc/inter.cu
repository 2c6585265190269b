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
#include "vvc_tables.h"

namespace vvcb200 {

constexpr int kInterWarps = 8;
constexpr int kMaxRefs    = 16;
constexpr int kWinElems   = 1288 + 24;          // staged windows of one warp (u16), largest class + slack
constexpr int kTmpElems   = 16 * 26 + 16;       // transposed horizontal-pass rows of one warp (int16)

struct InterArgs {
    PlanesView dst;
    PlanesView tmp14;                            // optional int16 output of the list-0 intermediate
    PlanesView refs[kMaxRefs];
    const VVCB200PredBlock *blocks;
    const uint16_t *lmcs_fwd;
    int n, n16, n8;                              // blocks of width 16 come first, then width 8, then width 4
    int bd, to_tmp14;
};

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
                                            const uint16_t *src, int stride, int pw, int ph, int xi, int yi,
                                            const int8_t *hf, const int8_t *vf, int bd,
                                            uint16_t *win, int ws, int16_t *tmp, int ts)
{
    constexpr int BEFORE = TAPS / 2 - 1;
    const int rows = h + TAPS - 1, cols = w + TAPS - 1;

    __syncwarp();
    if (active) {
        // stage the window: a lane owns columns li, li + gs, ... and walks down the rows
        for (int c = li; c < cols; c += gs) {
            const int sx = clip3(xi - BEFORE + c, 0, pw - 1);
            for (int r = 0; r < rows; r++) {
                const int sy = clip3(yi - BEFORE + r, 0, ph - 1);
                win[r * ws + c] = __ldg(src + (size_t)sy * stride + sx);
            }
        }
    }
    __syncwarp();
    if (active) {
        // horizontal pass: an item is one row of one strip of up to 8 outputs
        TapWords<TAPS> t;
        t.load(hf);
        const int strips = w > 8 ? 2 : 1, sw = min(w, 8);
        for (int i = li; i < rows * strips; i += gs) {
            const int r = strips == 2 ? i >> 1 : i, s = strips == 2 ? i & 1 : 0;
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

// One component of one block by one lane group (both lists + weighting + store).
template <int TAPS>
__device__ __forceinline__ void predict_component(const InterArgs &a, const VVCB200InterParams &p, const VVCB200PredBlock &b,
                                                  bool active, int c, int li, int gs, uint16_t *win, int ws,
                                                  int16_t *tmp, int ts)
{
    const int bd = a.bd;
    const int hs = c ? 1 : 0;
    const int w = max(b.w >> hs, 1), h = max(b.h >> hs, 1);
    const int x0 = b.x0 >> hs, y0 = b.y0 >> hs;
    const int log2w = 31 - __clz(w);
    const int set = c ? 0 : (b.flags & VVCB200_PB_AFFINE) ? 2 : (b.flags & VVCB200_PB_HPEL_IF) ? 1 : 0;
    int v0[8], v1[8];

#pragma unroll
    for (int l = 0; l < 2; l++) {
        const bool use = active && (b.pred_flag & (1 << l));
        const int mvx = b.mv[l][0], mvy = b.mv[l][1];
        const int mx = c ? (mvx & 31) : (mvx & 15);
        const int my = c ? (mvy & 31) : (mvy & 15);
        const int xi = x0 + (mvx >> (4 + hs)), yi = y0 + (mvy >> (4 + hs));
        const int8_t *hf = c ? VVC_CHROMA_FILTERS + (set * 32 + mx) * 4 : VVC_LUMA_FILTERS + (set * 16 + mx) * 8;
        const int8_t *vf = c ? VVC_CHROMA_FILTERS + (set * 32 + my) * 4 : VVC_LUMA_FILTERS + (set * 16 + my) * 8;
        const PlanesView &ref = a.refs[p.ref_slot[l][b.ref_idx[l] & 15] & (kMaxRefs - 1)];
        if (l == 0)
            interpolate<TAPS>(v0, use, li, gs, w, h, log2w, ref.p[c], ref.stride[c], ref.w[c], ref.h[c], xi, yi, hf, vf, bd,
                              win, ws, tmp, ts);
        else
            interpolate<TAPS>(v1, use, li, gs, w, h, log2w, ref.p[c], ref.stride[c], ref.w[c], ref.h[c], xi, yi, hf, vf, bd,
                              win, ws, tmp, ts);
    }
    if (!active)
        return;

    // ---- combine (weights are uniform inside a lane group)
    int mode;                       // 0 uni, 1 uni_w, 2 avg, 3 w_avg
    int wgt0 = 0, wgt1 = 0, off = 0, sh = 0, ox = 0;
    const int l_uni = (b.pred_flag & 1) ? 0 : 1;
    if (b.pred_flag != 3) {
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
    const bool lmcs = c == 0 && (b.flags & VVCB200_PB_LMCS) && a.lmcs_fwd;
    const int nchunk = h >= 4 ? 2 : 1, outs = h / nchunk;
    const int x = li & (w - 1), chunk = li >> log2w;
    if (chunk >= nchunk)
        return;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        if (k < outs) {
            const int y = chunk * outs + k;
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
    __shared__ __align__(16) uint16_t s_win[kInterWarps][kWinElems];
    __shared__ __align__(16) int16_t  s_tmp[kInterWarps][kTmpElems];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wi = blockIdx.x * kInterWarps + warp;

    // which width class and which blocks does this warp own?
    const int n4 = a.n - a.n16 - a.n8;
    const int warps8 = (a.n8 + 1) >> 1, warps4 = (n4 + 3) >> 2;
    int cls, first, last;                               // class: 0 -> w 16, 1 -> w 8, 2 -> w 4
    if (wi < a.n16) {
        cls = 0; first = wi; last = a.n16;
    } else if (wi < a.n16 + warps8) {
        cls = 1; first = a.n16 + (wi - a.n16) * 2; last = a.n16 + a.n8;
    } else if (wi < a.n16 + warps8 + warps4) {
        cls = 2; first = a.n16 + a.n8 + (wi - a.n16 - warps8) * 4; last = a.n;
    } else {
        return;
    }
    uint16_t *win = s_win[warp];
    int16_t *tmp = s_tmp[warp];

    {   // luma: groups of 2 * w lanes
        const int gs = 32 >> cls, g = lane >> (5 - cls), li = lane & (gs - 1);
        const int bi = first + g;
        const bool active = bi < last;
        const VVCB200PredBlock b = a.blocks[active ? bi : first];
        const int ws = (16 >> cls) + 10;                // 26 / 18 / 14 samples: an odd number of words
        const int win_sz = 23 * ws, tmp_sz = (16 >> cls) * 26;
        predict_component<8>(a, p, b, active && !(b.flags & VVCB200_PB_CHROMA_ONLY), 0, li, gs,
                             win + g * win_sz, ws, tmp + g * tmp_sz, 26);
    }
    {   // chroma: each luma group splits into a Cb half and a Cr half
        const int gs = 16 >> cls, g = lane / gs, li = lane & (gs - 1);
        const int bi = first + (g >> 1);
        const bool active = bi < last;
        const VVCB200PredBlock b = a.blocks[active ? bi : first];
        const int wc = 8 >> cls;
        const int ws = wc + 6;                          // 14 / 10 / 8 samples
        const int win_sz = 11 * ws + 8, tmp_sz = wc * 14;
        predict_component<4>(a, p, b, active && !(b.flags & VVCB200_PB_LUMA_ONLY), 1 + (g & 1), li, gs,
                             win + g * win_sz, ws, tmp + g * tmp_sz, 14);
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
    a.n = n;
    a.n16 = params->n_w16;
    a.n8 = params->n_w8;
    a.bd = params->bit_depth;
    const int n4 = n - a.n16 - a.n8;
    const int warps = a.n16 + ((a.n8 + 1) >> 1) + ((n4 + 3) >> 2);
    inter_kernel<<<(warps + kInterWarps - 1) / kInterWarps, kInterWarps * 32, 0, (cudaStream_t)stream>>>(a, *params);
    VVC_LAUNCH_CHECK();
    return 0;
}
