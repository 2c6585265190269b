// alf.cu — stage ALF as out-of-place picture passes: luma (4x4 classification + 7x7 diamond), chroma (5x5
// diamond) + CC-ALF.
// Reference semantics: ff_vvc_alf_filter and its buffer preparation (libavcodec/vvc/vvc_filter.c:1005-1319),
// alf.classify / recon_coeff_and_clip / filter[luma|chroma] / filter_cc (vvc_filter_template.c:38-408).
//
// The reference pads each CTB with its neighbours' pre-ALF samples, or — at a picture / slice / tile limit —
// with replicated CTB samples (alf_prepare_buffer).  Here that is a per-CTB coordinate clamp applied while a
// 32x32 tile (+3 halo) is staged into shared memory from the SAO output plane.  Virtual-boundary handling
// (row substitution around ctb_size-4, the >>10 rounding next to it, the reduced classification window)
// follows the template exactly.
#include "alf_dev.cuh"

namespace vvcb200 {

#ifndef VVC_ALF_LUMA_MIN_BLOCKS
#define VVC_ALF_LUMA_MIN_BLOCKS 4
#endif
#ifndef VVC_ALF_CHROMA_MIN_BLOCKS
#define VVC_ALF_CHROMA_MIN_BLOCKS 5   // 48 registers (6 or 8 blocks, i.e. 40 / 32 registers, measured 2 % slower on the stage; 1 block, 80 registers, 4 %)
#endif
__global__ void __launch_bounds__(256, VVC_ALF_LUMA_MIN_BLOCKS) alf_luma_kernel(const AlfArgs a)
{
    __shared__ AlfLumaShared S;

    const int tx0 = blockIdx.x * kTile, ty0 = blockIdx.y * kTile;
    const int W = a.src.w[0], H = a.src.h[0];
    const int ctb = (ty0 >> a.ctb_log2) * a.ctb_width + (tx0 >> a.ctb_log2);
    const VVCB200AlfCtb cp = a.ctbs[ctb];
    const uint16_t *src = a.src.p[0];
    const int sstride = a.src.stride[0];
    uint16_t *dst = a.dst.p[0];
    const int dstride = a.dst.stride[0];
    const int t = threadIdx.x;

    if (!cp.ctb_flag[0]) {                               // ALF off for this CTB: pass the SAO samples through
        for (int i = t; i < kTile * kTile / 2; i += 256) {
            const int y = ty0 + i / (kTile / 2), x = tx0 + (i % (kTile / 2)) * 2;
            if (x < W && y < H)
                *reinterpret_cast<uint32_t *>(dst + (size_t)y * dstride + x) =
                    *reinterpret_cast<const uint32_t *>(src + (size_t)y * sstride + x);
        }
        return;
    }
    if (t == 0)
        S.wide = 0;
    const CtbGeom g = ctb_geom(a, 0, tx0, ty0, cp.edges);

    // ---- stage the tile + halo
    {
        // halo positions not clamped by an edge flag are real neighbours inside the picture (picture borders are
        // always flagged by ctb_geom), so every read is in bounds.  All of a thread's loads are issued before the first
        // store so that they are in flight together (this loop was the kernel's main stall).
        constexpr int kIter = (kPad * kPad + 255) / 256;
        uint16_t v[kIter];
#pragma unroll
        for (int k = 0; k < kIter; k++) {
            const int i = t + k * 256;
            const int r = i / kPad, col = i - r * kPad;
            if (g.edges == 0)                            // interior CTB: nothing to clamp
                v[k] = i < kPad * kPad ? src[(ptrdiff_t)(ty0 + r - kHalo) * sstride + tx0 + col - kHalo] : (uint16_t)0;
            else
                v[k] = i < kPad * kPad ? (uint16_t)padded_sample(src, sstride, g, tx0 + col - kHalo, ty0 + r - kHalo) : (uint16_t)0;
        }
#pragma unroll
        for (int k = 0; k < kIter; k++) {
            const int i = t + k * 256;
            const int r = i / kPad, col = i - r * kPad;
            if (i < kPad * kPad)
                S.pix[r * kPadStr + col + kColOrg - kHalo] = v[k];
        }
    }
    __syncthreads();
    alf_luma_filter_tile<false>(a, cp, g, tx0, ty0, S, S.pix);
}

// chroma ALF (5x5) + CC-ALF; grid.z selects Cb / Cr
__global__ void __launch_bounds__(256, VVC_ALF_CHROMA_MIN_BLOCKS) alf_chroma_kernel(const AlfArgs a)
{
    __shared__ __align__(16) uint16_t s_pix[(kTile + 4) * (kTile + 4)];
    const int T = min(kTile, 1 << (a.ctb_log2 - 1));      // a tile never straddles a chroma CTB
    const int S = T + 4;
    const int c = 1 + blockIdx.z;
    const int tx0 = blockIdx.x * T, ty0 = blockIdx.y * T;
    const int l2 = a.ctb_log2 - 1;
    const int ctb = (ty0 >> l2) * a.ctb_width + (tx0 >> l2);
    const VVCB200AlfCtb cp = a.ctbs[ctb];
    const uint16_t *src = a.src.p[c];
    const int sstride = a.src.stride[c];
    const int t = threadIdx.x;
    const CtbGeom g = ctb_geom(a, c, tx0, ty0, cp.edges);

    if (cp.ctb_flag[c]) {
        if (g.edges == 0) {                              // interior CTB: the halo is simply the neighbours' samples
            const uint16_t *org = src + (ptrdiff_t)(ty0 - 2) * sstride + tx0 - 2;
            for (int i = t; i < S * S; i += 256) {
                const int r = i / S, col = i - r * S;
                s_pix[i] = org[(ptrdiff_t)r * sstride + col];
            }
        } else {
            for (int i = t; i < S * S; i += 256) {
                const int r = i / S, col = i - r * S;
                s_pix[i] = (uint16_t)padded_sample(src, sstride, g, tx0 + col - 2, ty0 + r - 2);
            }
        }
        __syncthreads();
    }
    const uint16_t *lp = a.src.p[0];
    const int ls = a.src.stride[0];
    alf_chroma_filter_tile(a, cp, g, c, T, tx0, ty0, s_pix, S,
                           [&](int x, int y) { return src[(size_t)y * sstride + x]; },
                           [&](int x, int y) { return lp[(size_t)y * ls + x]; });
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_alf_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200AlfCtb *d_ctbs,
                                   const VVCB200AlfParams *d_params, int ctb_log2_size, int bit_depth, void *stream)
{
    if (!dst || !src || !d_ctbs || !d_params || ctb_log2_size < 5 || ctb_log2_size > 7 ||
        (bit_depth != 10 && bit_depth != 12)) {
        set_error("vvcb200_alf_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    AlfArgs a;
    a.dst = make_view(dst);
    a.src = make_view(src);
    a.ctbs = d_ctbs;
    a.params = d_params;
    a.ctb_log2 = ctb_log2_size;
    a.ctb_width = (src->width[0] + (1 << ctb_log2_size) - 1) >> ctb_log2_size;
    a.bd = bit_depth;
    a.table = nullptr;
    if ((a.dst.stride[0] & 3) || ((uintptr_t)a.dst.p[0] & 7) || (a.src.stride[0] & 1) || ((uintptr_t)a.src.p[0] & 3)) {
        set_error("vvcb200_alf_picture: luma planes must be 8-byte aligned with strides that are multiples of 8 bytes");
        return VVCB200_ERR_ARG;
    }
    for (int c = 1; c < 3; c++)
        if ((a.dst.stride[c] & 1) || ((uintptr_t)a.dst.p[c] & 3) || (a.dst.w[c] & 1)) {
            set_error("vvcb200_alf_picture: chroma planes must be 4-byte aligned with even strides and widths (sample pairs)");
            return VVCB200_ERR_ARG;
        }
    // luma and chroma write different planes of dst and only read src: two independent grids, side by side
    const cudaStream_t main_stream = (cudaStream_t)stream;
    ForkStreams *f = fork_streams(main_stream);
    cudaStream_t side = main_stream;
    if (f) {
        VVC_CUDA(cudaEventRecord(f->fork, main_stream));
        side = f->side[0];
        VVC_CUDA(cudaStreamWaitEvent(side, f->fork, 0));
    }
    dim3 gl((src->width[0] + kTile - 1) / kTile, (src->height[0] + kTile - 1) / kTile, 1);
    alf_luma_kernel<<<gl, 256, 0, main_stream>>>(a);
    VVC_LAUNCH_CHECK();
    const int ct = kTile < (1 << (ctb_log2_size - 1)) ? kTile : (1 << (ctb_log2_size - 1));
    dim3 gc((src->width[1] + ct - 1) / ct, (src->height[1] + ct - 1) / ct, 2);
    alf_chroma_kernel<<<gc, 256, 0, side>>>(a);
    VVC_LAUNCH_CHECK();
    if (f) {
        VVC_CUDA(cudaEventRecord(f->join[0], side));
        VVC_CUDA(cudaStreamWaitEvent(main_stream, f->join[0], 0));
    }
    return 0;
}
