// engine.cu — Level B of the C ABI: the per-GPU picture engine (context, device planes, staging, stage order).
// Host-side C++ only; the kernels live in the stage files and are reached through their Level-A entries,
// so both levels run exactly the same device code.
#include <new>
#include <string.h>
#include "common.cuh"

using namespace vvcb200;

struct DevBuffer {
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap)
            return 0;
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 4 + 4096;
        VVC_CUDA(cudaMalloc(&p, want));
        cap = want;
        return 0;
    }
    void release()
    {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct VVCB200Ctx {
    int device = 0;
    int width = 0, height = 0, bit_depth = 10, hshift = 1, vshift = 1;
    cudaStream_t stream = nullptr;
    VVCB200Planes cur = {};        // picture under reconstruction
    void *cur_base = nullptr;
    DevBuffer levels, tbs;
};

namespace {

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess)
            prev = -1;
        if (prev != dev)
            ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard()
    {
        if (prev >= 0)
            cudaSetDevice(prev);
    }
};

int alloc_planes(VVCB200Ctx *c, VVCB200Planes *pl, void **base)
{
    size_t off[3], total = 0;
    for (int i = 0; i < 3; i++) {
        const int w = i ? c->width >> c->hshift : c->width;
        const int h = i ? c->height >> c->vshift : c->height;
        const int pitch = (w * 2 + 255) & ~255;          // 256-byte rows: every row start is 128-bit/TMA aligned
        pl->width[i] = w;
        pl->height[i] = h;
        pl->stride[i] = pitch;
        off[i] = total;
        total += (size_t)pitch * h;
    }
    VVC_CUDA(cudaMalloc(base, total));
    VVC_CUDA(cudaMemsetAsync(*base, 0, total, c->stream));
    for (int i = 0; i < 3; i++)
        pl->d_plane[i] = (uint8_t *)*base + off[i];
    return 0;
}

}  // namespace

extern "C" {

int vvcb200_ctx_create(VVCB200Ctx **out, int device, int width, int height, int chroma_format_idc, int bit_depth)
{
    if (!out || width <= 0 || height <= 0 || width > 16384 || height > 16384 || (width & 7) || (height & 7)) {
        set_error("vvcb200_ctx_create: bad dimensions %dx%d (multiples of 8 up to 16384)", width, height);
        return VVCB200_ERR_ARG;
    }
    if (bit_depth != 8 && bit_depth != 10 && bit_depth != 12) {
        set_error("vvcb200_ctx_create: unsupported bit depth %d", bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (chroma_format_idc != 1) {
        set_error("vvcb200_ctx_create: chroma_format_idc %d not built (4:2:0 only)", chroma_format_idc);
        return VVCB200_ERR_UNSUP;
    }
    int ndev = 0;
    VVC_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) {
        set_error("vvcb200_ctx_create: no CUDA device %d (%d visible); there is no CPU fallback", device, ndev);
        return VVCB200_ERR_CUDA;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cudaSetDevice(%d) failed", device);
        return VVCB200_ERR_CUDA;
    }
    VVCB200Ctx *c = new (std::nothrow) VVCB200Ctx;
    if (!c)
        return VVCB200_ERR_NOMEM;
    c->device = device;
    c->width = width;
    c->height = height;
    c->bit_depth = bit_depth;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete c;
        return cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__);
    }
    if (int r = alloc_planes(c, &c->cur, &c->cur_base)) {
        vvcb200_ctx_destroy(c);
        return r;
    }
    *out = c;
    return 0;
}

void vvcb200_ctx_destroy(VVCB200Ctx *c)
{
    if (!c)
        return;
    DeviceGuard g(c->device);
    if (c->stream)
        cudaStreamSynchronize(c->stream);
    c->levels.release();
    c->tbs.release();
    if (c->cur_base)
        cudaFree(c->cur_base);
    if (c->stream)
        cudaStreamDestroy(c->stream);
    delete c;
}

int vvcb200_ctx_sync(VVCB200Ctx *c)
{
    if (!c)
        return VVCB200_ERR_ARG;
    VVC_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

void *vvcb200_ctx_stream(VVCB200Ctx *c) { return c ? (void *)c->stream : nullptr; }

int vvcb200_ctx_current(VVCB200Ctx *c, VVCB200Planes *out)
{
    if (!c || !out)
        return VVCB200_ERR_ARG;
    *out = c->cur;
    return 0;
}

void *vvcb200_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        set_error("cudaHostAlloc(%zu) failed", bytes);
        return nullptr;
    }
    return p;
}

void vvcb200_host_free(void *p)
{
    if (p)
        cudaFreeHost(p);
}

int vvcb200_pic_begin(VVCB200Ctx *c)
{
    if (!c)
        return VVCB200_ERR_ARG;
    return 0;
}

int vvcb200_pic_upload(VVCB200Ctx *c, const uint16_t *const planes[3], const int32_t strides[3])
{
    if (!c || !planes || !strides)
        return VVCB200_ERR_ARG;
    DeviceGuard g(c->device);
    for (int i = 0; i < 3; i++)
        VVC_CUDA(cudaMemcpy2DAsync(c->cur.d_plane[i], c->cur.stride[i], planes[i], strides[i],
                                   (size_t)c->cur.width[i] * 2, c->cur.height[i], cudaMemcpyHostToDevice, c->stream));
    return 0;
}

int vvcb200_pic_download(VVCB200Ctx *c, uint16_t *const planes[3], const int32_t strides[3])
{
    if (!c || !planes || !strides)
        return VVCB200_ERR_ARG;
    DeviceGuard g(c->device);
    for (int i = 0; i < 3; i++)
        VVC_CUDA(cudaMemcpy2DAsync(planes[i], strides[i], c->cur.d_plane[i], c->cur.stride[i],
                                   (size_t)c->cur.width[i] * 2, c->cur.height[i], cudaMemcpyDeviceToHost, c->stream));
    return 0;
}

int vvcb200_pic_residual(VVCB200Ctx *c, const int16_t *levels, size_t n_levels,
                         const VVCB200TransformBlock *tbs, int n, const VVCB200ResidualParams *params)
{
    if (!c || !params || n < 0 || (n && (!levels || !tbs))) {
        set_error("vvcb200_pic_residual: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (params->bit_depth != c->bit_depth) {
        set_error("vvcb200_pic_residual: params bit depth %d != context %d", params->bit_depth, c->bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (n == 0)
        return 0;
    DeviceGuard g(c->device);
    if (int r = c->levels.reserve(n_levels * sizeof(int16_t)))
        return r;
    if (int r = c->tbs.reserve((size_t)n * sizeof(VVCB200TransformBlock)))
        return r;
    VVC_CUDA(cudaMemcpyAsync(c->levels.p, levels, n_levels * sizeof(int16_t), cudaMemcpyHostToDevice, c->stream));
    VVC_CUDA(cudaMemcpyAsync(c->tbs.p, tbs, (size_t)n * sizeof(VVCB200TransformBlock), cudaMemcpyHostToDevice, c->stream));
    return vvcb200_residual_batch(&c->cur, nullptr, (const int16_t *)c->levels.p,
                                  (const VVCB200TransformBlock *)c->tbs.p, n, params, c->stream);
}

}  // extern "C"

// ---- host-side descriptor producers ----------------------------------------------------------------------
extern "C" int vvcb200_pack_deblock_edges(uint8_t *const edge_out[2][3], const uint8_t *const bs[2][3],
                                          const uint8_t *const len_p[2], const uint8_t *const len_q[2],
                                          const uint8_t *tb_w_c, const uint8_t *tb_h_c,
                                          int stride4, int rows4, int ctb_log2_size)
{
    if (!edge_out || !bs || !len_p || !len_q || !tb_w_c || !tb_h_c || stride4 <= 0 || rows4 <= 0) {
        set_error("vvcb200_pack_deblock_edges: bad argument");
        return VVCB200_ERR_ARG;
    }
    const int ctb_units = 1 << (ctb_log2_size - 2);
    for (int dir = 0; dir < 2; dir++) {
        const int vertical = dir == 0;
        for (int uy = 0; uy < rows4; uy++) {
            for (int ux = 0; ux < stride4; ux++) {
                const int u = uy * stride4 + ux;
                const int b0 = bs[dir][0][u] & 3;
                edge_out[dir][0][u] = (uint8_t)(b0 | ((len_p[dir][u] & 7) << 2) | ((len_q[dir][u] & 7) << 5));
                // chroma lengths: max_filter_length_chroma(), vvc_filter.c:796-817
                const int pu = vertical ? (ux ? u - 1 : u) : (uy ? u - stride4 : u);
                const uint8_t *size = vertical ? tb_w_c : tb_h_c;
                const int big = size[pu] >= 8 && size[u] >= 8;
                const int ctu_edge = !vertical && !(uy % ctb_units);
                for (int c = 1; c < 3; c++) {
                    const int b = bs[dir][c][u] & 3;
                    int lp, lq;
                    if (big) {
                        lq = 3;
                        lp = ctu_edge ? 1 : 3;
                    } else {
                        lp = lq = b == 2;
                    }
                    edge_out[dir][c][u] = (uint8_t)(b | (lp << 2) | (lq << 5));
                }
            }
        }
    }
    return 0;
}
