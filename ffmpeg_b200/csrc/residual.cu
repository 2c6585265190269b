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
constexpr int kChunk       = 32;            // transform blocks per CTA
constexpr int kDct2Bytes   = 3412;
constexpr int kDstBytes    = 1360;
constexpr int kMatBytes    = kDct2Bytes + 2 * kDstBytes;   // DCT2 | DST7 | DCT8
constexpr int kMatPad      = (kMatBytes + 15) & ~15;
constexpr int kInElems     = 32 * 32;       // non-zero input box (zero-out keeps <= 32 x 32)
constexpr int kMidElems    = 64 * 32;       // after the column pass: h rows x nzw columns

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
    int bd, range, c_sign;
    int to_residual;
};

__device__ __forceinline__ const int8_t *tx_matrix(const int8_t *smat, int type, int log2n)
{
    if (type == VVCB200_DCT2)
        return smat + VVC_DCT2_OFF[log2n - 1];
    return smat + kDct2Bytes + (type == VVCB200_DCT8 ? kDstBytes : 0) + VVC_DST7_OFF[log2n - 2];
}

__device__ __forceinline__ void load_matrices(int8_t *smat)
{
    // 16-byte vector copy of the three matrix families into shared memory
    static_assert(kDct2Bytes % 4 == 0 && kDstBytes % 4 == 0, "word copy");
    uint32_t *d = reinterpret_cast<uint32_t *>(smat);
    const uint32_t *s0 = reinterpret_cast<const uint32_t *>(VVC_DCT2_MAT);
    const uint32_t *s1 = reinterpret_cast<const uint32_t *>(VVC_DST7_MAT);
    const uint32_t *s2 = reinterpret_cast<const uint32_t *>(VVC_DCT8_MAT);
    for (int i = threadIdx.x; i < kDct2Bytes / 4; i += blockDim.x)
        d[i] = s0[i];
    for (int i = threadIdx.x; i < kDstBytes / 4; i += blockDim.x) {
        d[kDct2Bytes / 4 + i] = s1[i];
        d[(kDct2Bytes + kDstBytes) / 4 + i] = s2[i];
    }
}

// The 2-D / 1-D inverse transform of one block by one warp (vvcdsp.c:94-138).
//   in : non-zero box, element (y, x) at in[y * in_stride + x], x < nzw, y < nzh
//   mid: scratch for the column-pass result (h x nzw)
//   emit(y, x, r) receives every residual sample of the w x h block exactly once.
template <typename T, typename Emit>
__device__ __forceinline__ void inverse_transform(const T *in, int in_stride, T *mid, const int8_t *smat,
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
        const int8_t *mv = tx_matrix(smat, trv, log2h);
        for (int i = lane; i < h * nzw; i += 32) {
            const int j = i / nzw, x = i - j * nzw;
            unsigned acc = 0;
            for (int k = 0; k < nzh; k++)
                acc += (unsigned)(int)in[k * in_stride + x] * (unsigned)(int)mv[k * h + j];
            mid[i] = (T)clip_intp2(((int)acc + 64) >> 7, range);
        }
        __syncwarp();
        const int8_t *mh = tx_matrix(smat, trh, log2w);
        for (int i = lane; i < w * h; i += 32) {
            const int y = i >> log2w, x = i & (w - 1);
            unsigned acc = 0;
            for (int k = 0; k < nzw; k++)
                acc += (unsigned)(int)mid[y * nzw + k] * (unsigned)(int)mh[k * w + x];
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
        const int8_t *m = tx_matrix(smat, type, w > 1 ? log2w : log2h);
        for (int i = lane; i < n; i += 32) {
            unsigned acc = 0;
            for (int k = 0; k < nz; k++)
                acc += (unsigned)(int)in[k * step] * (unsigned)(int)m[k * n + i];
            emit(w > 1 ? 0 : i, w > 1 ? i : 0, (int)(acc + (1u << (sh - 1))) >> sh);
        }
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

__global__ void __launch_bounds__(kWarpsPerCta * 32)
residual_kernel(const ResidualArgs a, const ScalingLists sl)
{
    extern __shared__ __align__(16) uint8_t smem_raw[];
    int8_t *smat = reinterpret_cast<int8_t *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int16_t *in  = reinterpret_cast<int16_t *>(smem_raw + kMatPad) + warp * (kInElems + kMidElems);
    int16_t *mid = in + kInElems;

    load_matrices(smat);
    __syncthreads();

    // The list is sorted by block area (largest first).  CTA b takes blocks b, b + grid, b + 2*grid, ... so
    // every CTA gets the same size profile (no tail of CTAs stuck with the 64x64 blocks).
    for (int t = blockIdx.x + warp * gridDim.x; t < a.n; t += kWarpsPerCta * gridDim.x) {
        const VVCB200TransformBlock d = a.tbs[t];
        const int w = 1 << d.log2_w, h = 1 << d.log2_h;
        const bool ts = d.flags & VVCB200_TB_TS;
        const int16_t *lv = a.levels + d.coeff_off;
        int nzw = d.nzw, nzh = d.nzh, in_stride = d.nzw;

        __syncwarp();
        // ---- levels into shared memory (+ BDPCM prefix sums on the raw levels)
        if (d.flags & VVCB200_TB_BDPCM) {
            in_stride = w;
            for (int i = lane; i < w * h; i += 32) {
                const int y = i >> d.log2_w, x = i & (w - 1);
                in[i] = (x < nzw && y < nzh) ? lv[y * nzw + x] : (int16_t)0;
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
            for (int i = lane; i < nzw * nzh; i += 32)
                in[i] = lv[i];
            __syncwarp();
        }
        // ---- dequant (vvc_intra.c:277-417)
        {
            const int log_sum = d.log2_w + d.log2_h;
            const int rect = ts ? 0 : (log_sum & 1);
            const int dq = (d.flags & VVCB200_TB_DEP_QUANT) ? 1 : 0;
            const int bd_shift = ts ? 10 : a.bd + rect + (log_sum >> 1) + 10 - a.range + dq;
            const unsigned bd_off = (1u << bd_shift) >> 1;
            const int qp = d.qp + (ts ? 0 : dq);
            const int rem = qp % 6;
            const int ls0 = rem == 0 ? 40 : rem == 1 ? 45 : rem == 2 ? 51 : rem == 3 ? 57 : rem == 4 ? 64 : 72;
            const int ls1 = rem == 0 ? 57 : rem == 1 ? 64 : rem == 2 ? 72 : rem == 3 ? 80 : rem == 4 ? 90 : 102;
            const unsigned scale = (unsigned)(rect ? ls1 : ls0) << (qp / 6);
            const bool use_sl = (d.flags & VVCB200_TB_SCALING) && !ts;
            const int id = scaling_id((d.flags & VVCB200_TB_INTER) != 0, d.c_idx, max(d.log2_w, d.log2_h));
            const int l2m = id < 2 ? 1 : id < 8 ? 2 : 3;
            for (int i = lane; i < nzw * nzh; i += 32) {
                const int y = i / nzw, x = i - y * nzw;
                const int c = in[y * in_stride + x];
                if (c) {
                    unsigned m = 16;
                    if (use_sl) {
                        const int off = ((y << l2m) >> d.log2_h) << l2m;
                        m = sl.rec[id][off + ((x << l2m) >> d.log2_w)];
                        if (id >= 14 && i == 0)
                            m = sl.dc[id - 14];
                    }
                    const int v = (int)((unsigned)c * scale * m + bd_off) >> bd_shift;
                    in[y * in_stride + x] = (int16_t)clip_intp2(v, a.range);
                }
            }
            __syncwarp();
        }
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
                    if (transpose) { const int s = x; x = y; y = s; }
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
        const int other = 3 - d.c_idx;
        const bool joint = (d.flags & VVCB200_TB_JOINT) && d.c_idx;
        const int jshift = (d.flags & VVCB200_TB_JOINT_SHIFT) ? 1 : 0;
        const bool lmcs = d.flags & VVCB200_TB_LMCS_SCALE;
        const int lscale = d.lmcs_scale;
        auto emit = [&](int y, int x, int r) {
            int r2 = joint ? (r * a.c_sign) >> jshift : 0;
            if (lmcs) {
                r = lmcs_scale_residual(r, lscale, a.bd);
                r2 = lmcs_scale_residual(r2, lscale, a.bd);
            }
            if (a.to_residual) {
                int16_t *p = reinterpret_cast<int16_t *>(a.res.p[d.c_idx]) + (size_t)(d.y0 + y) * a.res.stride[d.c_idx] + d.x0 + x;
                *p = (int16_t)clip_intp2(r, 15);
                if (joint) {
                    int16_t *q = reinterpret_cast<int16_t *>(a.res.p[other]) + (size_t)(d.y0 + y) * a.res.stride[other] + d.x0 + x;
                    *q = (int16_t)clip_intp2(r2, 15);
                }
            } else {
                uint16_t *p = a.dst.p[d.c_idx] + (size_t)(d.y0 + y) * a.dst.stride[d.c_idx] + d.x0 + x;
                *p = (uint16_t)clip_pixel(*p + r, a.bd);
                if (joint) {
                    uint16_t *q = a.dst.p[other] + (size_t)(d.y0 + y) * a.dst.stride[other] + d.x0 + x;
                    *q = (uint16_t)clip_pixel(*q + r2, a.bd);
                }
            }
        };
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
    int8_t *smat = reinterpret_cast<int8_t *>(smem_raw);
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
    a.to_residual = d_residual != nullptr;
    a.levels = d_levels;
    a.tbs = d_tbs;
    a.n = n;
    a.bd = params->bit_depth;
    a.range = params->log2_transform_range;
    a.c_sign = params->joint_c_sign;
    ScalingLists sl;
    memcpy(sl.rec, params->scaling_matrix_rec, sizeof(sl.rec));
    memset(sl.dc, 16, sizeof(sl.dc));
    memcpy(sl.dc, params->scaling_matrix_dc_rec, 14);

    const size_t smem = kMatPad + kWarpsPerCta * (kInElems + kMidElems) * sizeof(int16_t);
    static bool ready[64];
    if (int r = opt_in_smem(residual_kernel, smem, ready))
        return r;
    const int grid = (n + kChunk - 1) / kChunk;
    residual_kernel<<<grid, kWarpsPerCta * 32, smem, (cudaStream_t)stream>>>(a, sl);
    VVC_LAUNCH_CHECK();
    return 0;
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
