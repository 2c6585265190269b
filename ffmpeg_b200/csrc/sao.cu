// sao.cu — stage SAO as one out-of-place picture pass (+ the LMCS LUT pass, which is the same shape).
// Reference semantics: ff_vvc_sao_filter (libavcodec/vvc/vvc_filter.c:154-298) and sao.band_filter /
// edge_filter / edge_restore[0|1] (libavcodec/h26x/h2656_sao_template.c:24-215); lmcs.filter
// (vvc_filter_template.c:25-36) driven by ff_vvc_lmcs_filter (vvc_filter.c:1322-1332).
//
// Mapping: one thread per 2 horizontally adjacent samples (32-bit stores), lanes along x.  Neighbours are
// read from the source plane directly: the 3x3 footprint of a warp is 3 contiguous row segments, served by L1.
#include "common.cuh"

namespace vvcb200 {

struct SaoArgs {
    PlanesView dst, src;
    const VVCB200SaoCtb *ctbs;
    int ctb_log2, ctb_width, bd;
};

// value of one sample; (x, y) plane coordinates of component c, (w, h) plane size
__device__ __forceinline__ int sao_sample(const SaoArgs &a, const VVCB200SaoCtb &p, int c, int x, int y,
                                          int cx0, int cy0, int cw, int ch)
{
    const uint16_t *src = a.src.p[c];
    const int stride = a.src.stride[c];
    const int s = src[(size_t)y * stride + x];
    const int type = p.type_idx[c];
    if (type == VVCB200_SAO_OFF)
        return s;
    const int16_t *ov = p.offset_val[c];
    if (type == VVCB200_SAO_BAND) {
        const int k = (((s >> (a.bd - 5)) & 31) - p.band_position[c]) & 31;
        return clip_pixel(s + (k < 4 ? ov[k + 1] : 0), a.bd);
    }
    const int eo = p.eo_class[c];
    const int w = a.src.w[c], h = a.src.h[c];
    const int lx = x - cx0, ly = y - cy0;                 // position inside the CTB
    const bool b_left = cx0 == 0, b_top = cy0 == 0, b_right = cx0 + cw == w, b_bottom = cy0 + ch == h;
    // picture borders (edge_restore_0/1 first half): the class needs a neighbour that does not exist
    if (eo != 1 && ((b_left && lx == 0) || (b_right && lx == cw - 1)))
        return clip_pixel(s + ov[0], a.bd);
    if (eo != 0 && ((b_top && ly == 0) || (b_bottom && ly == ch - 1)))
        return clip_pixel(s + ov[0], a.bd);
    // slice / tile boundaries SAO may not look across (edge_restore_1 second half)
    const int nf = p.no_filter;
    if (nf) {
        const int init_x = (eo != 1 && b_left) ? 1 : 0, init_y = (eo != 0 && b_top) ? 1 : 0;
        const int width = cw - ((eo != 1 && b_right) ? 1 : 0), height = ch - ((eo != 0 && b_bottom) ? 1 : 0);
        const int save_ul = !(nf & 16) && eo == 2 && !b_left && !b_top;
        const int save_ur = !(nf & 32) && eo == 3 && !b_top && !b_right;
        const int save_lr = !(nf & 64) && eo == 2 && !b_right && !b_bottom;
        const int save_ll = !(nf & 128) && eo == 3 && !b_left && !b_bottom;
        bool keep = false;
        if ((nf & 1) && eo != 1 && lx == 0 && ly >= init_y + save_ul && ly < height - save_ll) keep = true;
        if ((nf & 2) && eo != 1 && lx == width - 1 && ly >= init_y + save_ur && ly < height - save_lr) keep = true;
        if ((nf & 4) && eo != 0 && ly == 0 && lx >= init_x + save_ul && lx < width - save_ur) keep = true;
        if ((nf & 8) && eo != 0 && ly == height - 1 && lx >= init_x + save_ll && lx < width - save_lr) keep = true;
        if ((nf & 16) && eo == 2 && lx == 0 && ly == 0) keep = true;
        if ((nf & 32) && eo == 3 && lx == width - 1 && ly == 0) keep = true;
        if ((nf & 64) && eo == 2 && lx == width - 1 && ly == height - 1) keep = true;
        if ((nf & 128) && eo == 3 && lx == 0 && ly == height - 1) keep = true;
        if (keep)
            return s;
    }
    // neighbour offsets per class (h2656_sao_template.c:54-59); coordinates clamped only to stay in the allocation,
    // every case where a clamped read would matter was handled above
    const int ax = eo == 1 ? 0 : (eo == 3 ? 1 : -1), ay = eo == 0 ? 0 : -1;
    const int na = src[(size_t)clip3(y + ay, 0, h - 1) * stride + clip3(x + ax, 0, w - 1)];
    const int nb = src[(size_t)clip3(y - ay, 0, h - 1) * stride + clip3(x - ax, 0, w - 1)];
    const int d = (s > na) - (s < na) + (s > nb) - (s < nb);      // -2..2
    const int idx = d == -2 ? 1 : d == -1 ? 2 : d == 0 ? 0 : d == 1 ? 3 : 4;
    return clip_pixel(s + ov[idx], a.bd);
}

__global__ void __launch_bounds__(256) sao_kernel(const SaoArgs a)
{
    const int c = blockIdx.z;
    const int w = a.src.w[c], h = a.src.h[c];
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    const int y = blockIdx.y;
    if (x >= w || y >= h)
        return;
    const int l2 = a.ctb_log2 - (c ? 1 : 0);
    const int cx = x >> l2, cy = y >> l2;
    const VVCB200SaoCtb p = a.ctbs[cy * a.ctb_width + cx];
    const int cx0 = cx << l2, cy0 = cy << l2;
    const int cw = min(1 << l2, w - cx0), ch = min(1 << l2, h - cy0);
    const int v0 = sao_sample(a, p, c, x, y, cx0, cy0, cw, ch);
    const int v1 = x + 1 < w ? sao_sample(a, p, c, x + 1, y, cx0, cy0, cw, ch) : 0;
    uint16_t *d = a.dst.p[c] + (size_t)y * a.dst.stride[c] + x;
    if (x + 1 < w)
        *reinterpret_cast<uint32_t *>(d) = (uint32_t)v0 | ((uint32_t)v1 << 16);
    else
        *d = (uint16_t)v0;
}

// Vectorised variant: one thread per 8 horizontally adjacent samples (128-bit accesses).  Threads whose samples
// touch a picture border or a no-filter CTB side take the per-sample routine above; everything else runs the
// straight-line code below.  Needs 16-byte aligned planes and strides.
template <int EO>
__device__ __forceinline__ void sao_edge8(uint32_t (&out)[4], const uint16_t *src, int stride, int x, int y,
                                          const int16_t *ov, int bd, const uint4 &vc, const uint4 &vu, const uint4 &vd)
{
    // neighbour a = (x + AX, y + AY), b = (x - AX, y - AY).  The three rows (vc: y, vu: y - 1, vd: y + 1) were fetched
    // before the CTB's parameters were known, so that both round trips overlap; only the two samples beside the
    // vectors are read here, and those sit in lines the neighbouring threads have just asked for.
    constexpr int AX = EO == 1 ? 0 : (EO == 3 ? 1 : -1), AY = EO == 0 ? 0 : -1;
    const uint16_t *rc = src + (size_t)y * stride + x;
    const uint16_t *ra = rc + AY * stride, *rb = rc - AY * stride;
    uint16_t c[8], a[8], b[8];
    *reinterpret_cast<uint4 *>(c) = vc;
    if (AX == 0) {
        *reinterpret_cast<uint4 *>(a) = vu;
        *reinterpret_cast<uint4 *>(b) = vd;
    } else {
        uint16_t ta[8], tb[8];
        *reinterpret_cast<uint4 *>(ta) = AY == 0 ? vc : vu;
        *reinterpret_cast<uint4 *>(tb) = AY == 0 ? vc : vd;
        // shift by one sample: a[j] = rowA[j + AX], b[j] = rowB[j - AX]
        const uint16_t ea = AX < 0 ? ra[-1] : ra[8], eb = AX < 0 ? rb[8] : rb[-1];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int ja = j + AX, jb = j - AX;
            a[j] = ja < 0 || ja > 7 ? ea : ta[ja];
            b[j] = jb < 0 || jb > 7 ? eb : tb[jb];
        }
    }
    // The five offsets (|offset| < 128 at any bit depth up to 12) packed as bytes in the order of d + 2 = 0..4, so that one
    // PRMT both selects and sign-extends: selector nibble i picks byte i, bit 3 of a nibble replicates that byte's sign.
    const uint32_t lo = (uint32_t)(ov[1] & 0xff) | ((uint32_t)(ov[2] & 0xff) << 8) | ((uint32_t)(ov[0] & 0xff) << 16) |
                        ((uint32_t)(ov[3] & 0xff) << 24);
    const uint32_t hi = (uint32_t)(ov[4] & 0xff);
    const int pmax = (1 << bd) - 1;
    int r[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int s = c[j];
        // sign(s - a) + sign(s - b) + 2 with clamps instead of four compares
        const int d2 = min(max(s - (int)a[j], -1), 1) + min(max(s - (int)b[j], -1), 1) + 2;
        int off;                                                          // (the __byte_perm intrinsic masks the sign-replicate bits)
        asm("prmt.b32 %0, %1, %2, %3;" : "=r"(off) : "r"(lo), "r"(hi), "r"(d2 * 0x1111 + 0x8880));
        r[j] = min(max(s + off, 0), pmax);
    }
#pragma unroll
    for (int j = 0; j < 4; j++)
        out[j] = (uint32_t)r[2 * j] | ((uint32_t)r[2 * j + 1] << 16);
}

#ifndef VVC_SAO_MIN_BLOCKS
#define VVC_SAO_MIN_BLOCKS 10         // 48 registers; 12 / 16 blocks (40 / 32 registers) measured no faster
#endif
__global__ void __launch_bounds__(128, VVC_SAO_MIN_BLOCKS) sao_kernel_v8(const SaoArgs a)
{
    const int c = blockIdx.z;
    const int w = a.src.w[c], h = a.src.h[c];
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
    const int y = blockIdx.y;
    if (x >= w || y >= h)
        return;
    const uint16_t *src = a.src.p[c];
    const int stride = a.src.stride[c];
    // the sample rows first: they do not depend on the CTB's parameters, whose fetch then overlaps them (rows y - 1 and
    // y + 1 are only used by the edge classes; clamped so that the speculative reads stay inside the plane)
    uint4 vc, vu, vd;
    if (x + 8 <= w) {
        vc = *reinterpret_cast<const uint4 *>(src + (size_t)y * stride + x);
        vu = *reinterpret_cast<const uint4 *>(src + (size_t)max(y - 1, 0) * stride + x);
        vd = *reinterpret_cast<const uint4 *>(src + (size_t)min(y + 1, h - 1) * stride + x);
    }
    const int l2 = a.ctb_log2 - (c ? 1 : 0);
    const int cx = x >> l2, cy = y >> l2;
    const VVCB200SaoCtb *pp = a.ctbs + cy * a.ctb_width + cx;
    const int type = pp->type_idx[c];
    uint16_t *d = a.dst.p[c] + (size_t)y * a.dst.stride[c] + x;
    const int cx0 = cx << l2, cy0 = cy << l2;
    const int cw = min(1 << l2, w - cx0), ch = min(1 << l2, h - cy0);
    bool slow = x + 8 > w;
    if (type == VVCB200_SAO_EDGE) {
        // samples on a CTB side that is a picture border or must not be filtered across need the per-sample rules
        const bool side = x == cx0 || x + 8 >= cx0 + cw || y == cy0 || y == cy0 + ch - 1;
        const bool border = cx0 == 0 || cy0 == 0 || cx0 + cw == w || cy0 + ch == h;
        slow = slow || (side && (border || pp->no_filter));
    }
    if (slow) {
        const VVCB200SaoCtb p = *pp;
        for (int j = 0; j < 8 && x + j < w; j++)
            d[j] = (uint16_t)sao_sample(a, p, c, x + j, y, cx0, cy0, cw, ch);
        return;
    }
    uint32_t out[4];
    if (type == VVCB200_SAO_OFF) {
        *reinterpret_cast<uint4 *>(d) = vc;
        return;
    }
    const int16_t *ov = pp->offset_val[c];
    if (type == VVCB200_SAO_BAND) {
        uint16_t s[8];
        *reinterpret_cast<uint4 *>(s) = vc;
        const int pos = pp->band_position[c], sh = a.bd - 5, pmax = (1 << a.bd) - 1;
        // offsets of the four bands as bytes 0..3, byte 4 = 0 for every other band; one PRMT selects and sign-extends
        const uint32_t lo = (uint32_t)(ov[1] & 0xff) | ((uint32_t)(ov[2] & 0xff) << 8) | ((uint32_t)(ov[3] & 0xff) << 16) |
                            ((uint32_t)(ov[4] & 0xff) << 24);
        int r[8];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int k = min(((s[j] >> sh) - pos) & 31, 4);
            int off;
            asm("prmt.b32 %0, %1, %2, %3;" : "=r"(off) : "r"(lo), "r"(0u), "r"(k * 0x1111 + 0x8880));
            r[j] = min(max(s[j] + off, 0), pmax);
        }
#pragma unroll
        for (int j = 0; j < 4; j++)
            out[j] = (uint32_t)r[2 * j] | ((uint32_t)r[2 * j + 1] << 16);
    } else {
        switch (pp->eo_class[c]) {
        case 0:  sao_edge8<0>(out, src, stride, x, y, ov, a.bd, vc, vu, vd); break;
        case 1:  sao_edge8<1>(out, src, stride, x, y, ov, a.bd, vc, vu, vd); break;
        case 2:  sao_edge8<2>(out, src, stride, x, y, ov, a.bd, vc, vu, vd); break;
        default: sao_edge8<3>(out, src, stride, x, y, ov, a.bd, vc, vu, vd); break;
        }
    }
    *reinterpret_cast<uint4 *>(d) = make_uint4(out[0], out[1], out[2], out[3]);
}

__global__ void __launch_bounds__(256) lmcs_kernel(PlanesView pic, const uint16_t *lut, int mask)
{
    __shared__ uint16_t s_lut[4096];
    for (int i = threadIdx.x; i <= mask; i += blockDim.x)
        s_lut[i] = lut[i];
    __syncthreads();
    const int w = pic.w[0], h = pic.h[0];
    const int groups = (w + 7) >> 3;                         // 8 samples (128 bits) per thread
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < groups * h; i += gridDim.x * blockDim.x) {
        const int y = i / groups, x = (i - y * groups) << 3;
        uint16_t *row = pic.p[0] + (size_t)y * pic.stride[0] + x;
        if (x + 8 <= w) {
            uint4 v = *reinterpret_cast<uint4 *>(row);
            uint32_t *q = reinterpret_cast<uint32_t *>(&v);
#pragma unroll
            for (int k = 0; k < 4; k++)
                q[k] = (uint32_t)s_lut[q[k] & mask] | ((uint32_t)s_lut[(q[k] >> 16) & mask] << 16);
            *reinterpret_cast<uint4 *>(row) = v;
        } else {
            for (int k = 0; x + k < w; k++)
                row[k] = s_lut[row[k] & mask];
        }
    }
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" int vvcb200_sao_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200SaoCtb *d_ctbs,
                                   int ctb_log2_size, int bit_depth, void *stream)
{
    if (!dst || !src || !d_ctbs || ctb_log2_size < 5 || ctb_log2_size > 7 ||
        (bit_depth != 10 && bit_depth != 12)) {
        set_error("vvcb200_sao_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    SaoArgs a;
    a.dst = make_view(dst);
    a.src = make_view(src);
    a.ctbs = d_ctbs;
    a.ctb_log2 = ctb_log2_size;
    a.ctb_width = (src->width[0] + (1 << ctb_log2_size) - 1) >> ctb_log2_size;
    a.bd = bit_depth;
    bool aligned = true;
    for (int c = 0; c < 3; c++)
        aligned = aligned && !(((uintptr_t)a.src.p[c] | (uintptr_t)a.dst.p[c]) & 15) && !((a.src.stride[c] | a.dst.stride[c]) & 7);
    if (aligned) {
        dim3 grid((src->width[0] / 8 + 127) / 128, src->height[0], 3);
        sao_kernel_v8<<<grid, 128, 0, (cudaStream_t)stream>>>(a);
    } else {
        // same arithmetic, scalar accesses: planes that do not meet the 16-byte alignment of the engine's own pictures
        dim3 grid((src->width[0] / 2 + 255) / 256, src->height[0], 3);
        sao_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    }
    VVC_LAUNCH_CHECK();
    return 0;
}

extern "C" int vvcb200_lmcs_picture(const VVCB200Planes *pic, const uint16_t *d_lut, int bit_depth, void *stream)
{
    if (!pic || !d_lut || (bit_depth != 10 && bit_depth != 12)) {
        set_error("vvcb200_lmcs_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    const PlanesView v = make_view(pic);
    if (v.stride[0] % 8 || ((uintptr_t)v.p[0] & 15)) {
        set_error("vvcb200_lmcs_picture: luma plane must be 16-byte aligned with a 16-byte multiple stride");
        return VVCB200_ERR_ARG;
    }
    lmcs_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(v, d_lut, (1 << bit_depth) - 1);
    VVC_LAUNCH_CHECK();
    return 0;
}
