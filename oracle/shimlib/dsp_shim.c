/*
 * oracle/shimlib/dsp_shim.c — TEST INFRASTRUCTURE (SURVEY.md §8 b: the checkasm-style view of the boundary).
 *
 * vvcb200_dsp_init(VVCDSPContext *c, int bd) mirrors ff_vvc_dsp_init (libavcodec/vvc/vvcdsp.h:170): it fills the
 * reference's own function table, then replaces every member that has a block-level view on the device by a function
 * with the member's exact signature that pushes the one block through libvvcb200.so synchronously (copy in, one Level-A
 * call, copy out).  The members that stay the reference's are listed, with the reason, in vvcshim_report().
 * vvcshim_check() is the driver: modelled on tests/checkasm/vvc_mc.c:68-324 and vvc_itx.c, it calls the reference table
 * and the shimmed table with the same random inputs over the reference tests' own shapes and compares the outputs.
 * Compiled against the reference's headers where they lie; links libvvcref.so (the reference's code), libvvcb200.so and
 * the CUDA runtime.  Nothing here is product code.
 */
#include "libavcodec/vvc/vvcdsp.h"
#include "libavcodec/vvc/vvc_data.h"
#include "libavcodec/vvc/vvc_ctu.h"
#include "vvc_b200.h"

#include <cuda_runtime_api.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define SHIM_EXPORT __attribute__((visibility("default")))

static int g_bd = 10;                     /* the table has no user pointer: the last vvcb200_dsp_init() decides */
static int g_device_errors;               /* a device call that failed (the output then keeps the reference's bytes) */
static char *g_scratch;                   /* one device arena, carved up per call */
static size_t g_scratch_size, g_scratch_used;

static void *dev_alloc(size_t bytes)
{
    if (!g_scratch) {
        g_scratch_size = 64u << 20;
        if (cudaMalloc((void **)&g_scratch, g_scratch_size) != cudaSuccess) {
            g_scratch = NULL;
            g_device_errors++;
            return NULL;
        }
    }
    bytes = (bytes + 255) & ~(size_t)255;
    if (g_scratch_used + bytes > g_scratch_size) {
        g_device_errors++;
        return NULL;
    }
    g_scratch_used += bytes;
    return g_scratch + g_scratch_used - bytes;
}

static void dev_reset(void) { g_scratch_used = 0; }

static int ok(int r)
{
    if (r != 0 || cudaDeviceSynchronize() != cudaSuccess) {
        g_device_errors++;
        return 0;
    }
    return 1;
}

static int ilog2(int v) { int l = 0; while ((1 << l) < v) l++; return l; }

/* ---- itx.itx[trh][trv][log2 w][log2 h] (vvcdsp.h:117-118) -> vvcb200_dsp_itx_batch ------------------------------ */
static void run_itx(int trh, int trv, int lw, int lh, int *coeffs, size_t nzw, size_t nzh, intptr_t range, intptr_t bd)
{
    const size_t n = (size_t)1 << (lw + lh);
    VVCB200TransformBlock tb = { 0 };
    uint32_t off = 0;
    int32_t *d_c;
    uint32_t *d_off;
    VVCB200TransformBlock *d_tb;

    dev_reset();
    d_c = dev_alloc(n * 4);
    d_off = dev_alloc(4);
    d_tb = dev_alloc(sizeof(tb));
    if (!d_c || !d_off || !d_tb)
        return;
    tb.log2_w = lw; tb.log2_h = lh; tb.nzw = nzw; tb.nzh = nzh; tb.trh = trh; tb.trv = trv;
    cudaMemcpy(d_c, coeffs, n * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_off, &off, 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_tb, &tb, sizeof(tb), cudaMemcpyHostToDevice);
    if (ok(vvcb200_dsp_itx_batch(d_c, d_off, d_tb, 1, (int)range, (int)bd, NULL)))
        cudaMemcpy(coeffs, d_c, n * 4, cudaMemcpyDeviceToHost);
}

#define ITX0(a, b, w, h) \
    static void shim_itx_##a##b##w##h(int *c, size_t nzw, size_t nzh, intptr_t range, intptr_t bd) \
    { run_itx(a, b, w, h, c, nzw, nzh, range, bd); }
#define ITX1(a, b, w) ITX0(a, b, w, 0) ITX0(a, b, w, 1) ITX0(a, b, w, 2) ITX0(a, b, w, 3) ITX0(a, b, w, 4) ITX0(a, b, w, 5) ITX0(a, b, w, 6)
#define ITX2(a, b) ITX1(a, b, 0) ITX1(a, b, 1) ITX1(a, b, 2) ITX1(a, b, 3) ITX1(a, b, 4) ITX1(a, b, 5) ITX1(a, b, 6)
#define ITX3(a) ITX2(a, 0) ITX2(a, 1) ITX2(a, 2)
ITX3(0) ITX3(1) ITX3(2)

#define SET0(a, b, w, h) if (c->itx.itx[a][b][w][h]) { c->itx.itx[a][b][w][h] = shim_itx_##a##b##w##h; n++; }
#define SET1(a, b, w) SET0(a, b, w, 0) SET0(a, b, w, 1) SET0(a, b, w, 2) SET0(a, b, w, 3) SET0(a, b, w, 4) SET0(a, b, w, 5) SET0(a, b, w, 6)
#define SET2(a, b) SET1(a, b, 0) SET1(a, b, 1) SET1(a, b, 2) SET1(a, b, 3) SET1(a, b, 4) SET1(a, b, 5) SET1(a, b, 6)
#define SET3(a) SET2(a, 0) SET2(a, 1) SET2(a, 2)

/* ---- a device picture of one plane's worth of samples ------------------------------------------------------- */
/* 4:2:0 picture of luma size W x H (even, multiples of 8 keep every row 16-byte aligned): three planes in the arena */
static int make_planes(VVCB200Planes *p, int W, int H, int bytes_per_sample)
{
    memset(p, 0, sizeof(*p));
    for (int c = 0; c < 3; c++) {
        const int w = c ? W / 2 : W, h = c ? H / 2 : H;
        const int stride = ((w * bytes_per_sample + 63) & ~63);
        p->d_plane[c] = dev_alloc((size_t)stride * h);
        if (!p->d_plane[c])
            return 0;
        cudaMemset(p->d_plane[c], 0, (size_t)stride * h);
        p->stride[c] = stride;
        p->width[c] = w;
        p->height[c] = h;
    }
    return 1;
}

/* ---- itx.add_residual (vvcdsp.h:113) -> vvcb200_dsp_add_residual_batch ------------------------------------- */
static void shim_add_residual(uint8_t *dst, const int *res, int w, int h, ptrdiff_t stride)
{
    const int W = (w + 7) & ~7, H = (h + 1) & ~1;
    VVCB200Planes pl;
    VVCB200TransformBlock tb = { 0 };
    uint32_t off = 0;

    dev_reset();
    if (!make_planes(&pl, W < 16 ? 16 : W, H < 2 ? 2 : H, 2))
        return;
    int32_t *d_res = dev_alloc((size_t)w * h * 4);
    uint32_t *d_off = dev_alloc(4);
    VVCB200TransformBlock *d_tb = dev_alloc(sizeof(tb));
    if (!d_res || !d_off || !d_tb)
        return;
    tb.log2_w = ilog2(w); tb.log2_h = ilog2(h);
    cudaMemcpy2D(pl.d_plane[0], pl.stride[0], dst, stride, (size_t)w * 2, h, cudaMemcpyHostToDevice);
    cudaMemcpy(d_res, res, (size_t)w * h * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_off, &off, 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_tb, &tb, sizeof(tb), cudaMemcpyHostToDevice);
    if (ok(vvcb200_dsp_add_residual_batch(&pl, d_res, d_off, d_tb, 1, g_bd, NULL)))
        cudaMemcpy2D(dst, stride, pl.d_plane[0], pl.stride[0], (size_t)w * 2, h, cudaMemcpyDeviceToHost);
}

/* ---- lmcs.filter (vvcdsp.h:124) -> vvcb200_lmcs_picture ---------------------------------------------------- */
static void shim_lmcs_filter(uint8_t *dst, ptrdiff_t dst_stride, int width, int height, const uint8_t *lut)
{
    const int W = (width + 15) & ~15, H = (height + 1) & ~1;
    VVCB200Planes pl;

    dev_reset();
    if (!make_planes(&pl, W, H, 2))
        return;
    uint16_t *d_lut = dev_alloc(sizeof(uint16_t) << g_bd);
    if (!d_lut)
        return;
    cudaMemcpy2D(pl.d_plane[0], pl.stride[0], dst, dst_stride, (size_t)width * 2, height, cudaMemcpyHostToDevice);
    cudaMemcpy(d_lut, lut, sizeof(uint16_t) << g_bd, cudaMemcpyHostToDevice);
    if (ok(vvcb200_lmcs_picture(&pl, d_lut, g_bd, NULL)))
        cudaMemcpy2D(dst, dst_stride, pl.d_plane[0], pl.stride[0], (size_t)width * 2, height, cudaMemcpyDeviceToHost);
}

/* ---- inter.put / put_uni / put_uni_w [luma, chroma][log2 w - 1][my != 0][mx != 0] (vvcdsp.h:49-59) -> vvcb200_inter_batch
 * The table members take tap pointers; the device takes (filter set, fraction), as the decoder's call sites do
 * (vvc_inter.c:180-330): the pointers are looked up in ff_vvc_inter_luma_filters / ff_vvc_inter_chroma_filters.  Calls no
 * decoder makes — taps of two different sets, chroma sets other than 0 — go to the reference and are counted. */
static VVCDSPContext g_ref;                               /* the reference's table of g_bd */
static int g_unrepresentable;

typedef struct TapId { int set, frac; } TapId;

static int find_taps(TapId *t, const int8_t *f, int chroma)
{
    const int8_t *base = chroma ? &ff_vvc_inter_chroma_filters[0][0][0] : &ff_vvc_inter_luma_filters[0][0][0];
    const int taps = chroma ? 4 : 8, fracs = chroma ? 32 : 16;
    const ptrdiff_t d = f - base;
    if (d < 0 || d >= 3 * fracs * taps || d % taps)
        return 0;
    t->set = (int)(d / (fracs * taps));
    t->frac = (int)(d / taps) % fracs;
    return 1;
}

/* mode 0: put (int16 intermediate, stride MAX_PB_SIZE); 1: put_uni; 2: put_uni_w */
static int run_put(int mode, int chroma, int j, int i, void *dst, ptrdiff_t dst_stride, const uint8_t *src, ptrdiff_t src_stride,
                   int height, const int8_t *hf, const int8_t *vf, int width, int denom, int wx, int ox)
{
    TapId th, tv;
    if (!find_taps(&th, hf, chroma) || !find_taps(&tv, vf, chroma))
        return 0;
    /* a slot that does not filter in a direction ignores that direction's taps */
    const int mx = i ? th.frac : 0, my = j ? tv.frac : 0;
    const int set = i ? th.set : tv.set;
    if ((i && j && th.set != tv.set) || (chroma && set != 0))
        return 0;
    /* the block and the ring its filters reach, as a picture of its own: luma 3 before / 4 after, chroma 1 / 2 */
    const int before = chroma ? 1 : 3, after = chroma ? 2 : 4;
    const int org = chroma ? 2 : 4;                          /* block origin inside the plane (luma origin even) */
    const int pw = org + width + after, ph = org + height + after;
    const int W = ((chroma ? 2 * pw : pw) + 15) & ~15, H = ((chroma ? 2 * ph : ph) + 1) & ~1;
    const int c = chroma ? 1 : 0;
    VVCB200Planes ref, out;
    VVCB200InterParams p;
    VVCB200PredBlock *blocks, *d_blocks;
    int n = 0;

    dev_reset();
    if (!make_planes(&ref, W, H, 2) || !make_planes(&out, W, H, 2))
        return 0;
    cudaMemcpy2D((uint8_t *)ref.d_plane[c] + (size_t)(org - before) * ref.stride[c] + (org - before) * 2, ref.stride[c],
                 src - before * src_stride - before * 2, src_stride, (size_t)(width + before + after) * 2, height + before + after,
                 cudaMemcpyHostToDevice);
    memset(&p, 0, sizeof(p));
    p.bit_depth = g_bd;
    p.num_ref_idx_active[0] = 1;
    if (mode == 2) {
        p.log2_denom[0] = p.log2_denom[1] = denom;
        p.weight[0][c][0] = wx;
        p.offset[0][c][0] = ox;
    }
    /* sub-blocks of at most 16x16 luma samples (8x8 chroma), all with the one vector */
    const int lw = chroma ? 2 * width : width, lh = chroma ? 2 * height : height;   /* the block in luma units */
    const int sbw = lw < 16 ? lw : 16, sbh = lh < 16 ? lh : 16;
    blocks = calloc((size_t)(lw / sbw) * (lh / sbh), sizeof(*blocks));
    for (int y = 0; y < lh; y += sbh)
        for (int x = 0; x < lw; x += sbw) {
            VVCB200PredBlock *b = &blocks[n++];
            b->x0 = (chroma ? 2 * org : org) + x;
            b->y0 = (chroma ? 2 * org : org) + y;
            b->w = sbw;
            b->h = sbh;
            b->pred_flag = 1;
            b->mv[0][0] = mx;                                 /* 1/16 luma = 1/32 chroma sample */
            b->mv[0][1] = my;
            b->flags = (chroma ? VVCB200_PB_CHROMA_ONLY : VVCB200_PB_LUMA_ONLY) | (mode == 2 ? VVCB200_PB_WP : 0) |
                       (set == 1 ? VVCB200_PB_HPEL_IF : set == 2 ? VVCB200_PB_AFFINE : 0);
        }
    p.n_w16 = sbw == 16 ? n : 0;
    p.n_w8 = sbw == 8 ? n : 0;
    d_blocks = dev_alloc((size_t)n * sizeof(*blocks));
    if (!d_blocks) {
        free(blocks);
        return 0;
    }
    cudaMemcpy(d_blocks, blocks, (size_t)n * sizeof(*blocks), cudaMemcpyHostToDevice);
    free(blocks);
    if (!ok(vvcb200_inter_batch(mode ? &out : NULL, &ref, 1, d_blocks, n, &p, NULL, mode ? NULL : &out, NULL, NULL, NULL)))
        return 0;
    cudaMemcpy2D(dst, mode ? dst_stride : MAX_PB_SIZE * 2, (uint8_t *)out.d_plane[c] + (size_t)org * out.stride[c] + org * 2,
                 out.stride[c], (size_t)width * 2, height, cudaMemcpyDeviceToHost);
    return 1;
}

#define PUT0(C, K, J, I) \
    static void shim_put_##C##K##J##I(int16_t *dst, const uint8_t *src, ptrdiff_t ss, int h, const int8_t *hf, const int8_t *vf, int w) \
    { if (!run_put(0, C, J, I, dst, 0, src, ss, h, hf, vf, w, 0, 0, 0)) { g_unrepresentable++; g_ref.inter.put[C][K][J][I](dst, src, ss, h, hf, vf, w); } } \
    static void shim_uni_##C##K##J##I(uint8_t *dst, ptrdiff_t ds, const uint8_t *src, ptrdiff_t ss, int h, const int8_t *hf, const int8_t *vf, int w) \
    { if (!run_put(1, C, J, I, dst, ds, src, ss, h, hf, vf, w, 0, 0, 0)) { g_unrepresentable++; g_ref.inter.put_uni[C][K][J][I](dst, ds, src, ss, h, hf, vf, w); } } \
    static void shim_uniw_##C##K##J##I(uint8_t *dst, ptrdiff_t ds, const uint8_t *src, ptrdiff_t ss, int h, int denom, int wx, int ox, \
                                       const int8_t *hf, const int8_t *vf, int w) \
    { if (!run_put(2, C, J, I, dst, ds, src, ss, h, hf, vf, w, denom, wx, ox)) { g_unrepresentable++; g_ref.inter.put_uni_w[C][K][J][I](dst, ds, src, ss, h, denom, wx, ox, hf, vf, w); } }
#define PUT1(C, K) PUT0(C, K, 0, 0) PUT0(C, K, 0, 1) PUT0(C, K, 1, 0) PUT0(C, K, 1, 1)
#define PUT2(C) PUT1(C, 0) PUT1(C, 1) PUT1(C, 2) PUT1(C, 3) PUT1(C, 4) PUT1(C, 5) PUT1(C, 6)
PUT2(0) PUT2(1)

#define PSET0(C, K, J, I) \
    if (c->inter.put[C][K][J][I])       { c->inter.put[C][K][J][I] = shim_put_##C##K##J##I; n++; } \
    if (c->inter.put_uni[C][K][J][I])   { c->inter.put_uni[C][K][J][I] = shim_uni_##C##K##J##I; n++; } \
    if (c->inter.put_uni_w[C][K][J][I]) { c->inter.put_uni_w[C][K][J][I] = shim_uniw_##C##K##J##I; n++; }
#define PSET1(C, K) PSET0(C, K, 0, 0) PSET0(C, K, 0, 1) PSET0(C, K, 1, 0) PSET0(C, K, 1, 1)
#define PSET2(C) PSET1(C, 0) PSET1(C, 1) PSET1(C, 2) PSET1(C, 3) PSET1(C, 4) PSET1(C, 5) PSET1(C, 6)

/* Returns the number of table members now served by the device. */
SHIM_EXPORT int vvcb200_dsp_init(VVCDSPContext *c, int bd)
{
    int n = 0;
    memset(c, 0, sizeof(*c));               /* (ff_vvc_dsp_init writes the populated members only) */
    memset(&g_ref, 0, sizeof(g_ref));
    ff_vvc_dsp_init(c, bd);
    ff_vvc_dsp_init(&g_ref, bd);
    g_bd = bd;
    if (bd != 10 && bd != 12)               /* 8-bit pictures are uint8 planes in the reference: not built */
        return 0;
    SET3(0) SET3(1) SET3(2)
    c->itx.add_residual = shim_add_residual; n++;
    c->lmcs.filter = shim_lmcs_filter; n++;
    PSET2(0) PSET2(1)
    return n;
}

/* ---- what stays the reference's, and why -------------------------------------------------------------------- */
SHIM_EXPORT const char *vvcshim_report(void)
{
    return
        "served by the device: itx.itx[][][][] (every populated member), itx.add_residual, lmcs.filter, inter.put / put_uni / "
        "put_uni_w [luma, chroma][7][2][2]\n"
        "left to the reference (no block-level device view; each is covered at stage level against the reference's stage code):\n"
        "  inter.avg, w_avg, put_gpm, apply_bdof, apply_prof*, prof_grad_filter: they take the 14-bit intermediates of put(), which "
        "on the device never leave the kernel that made them (tests/test_gpu_inter.py: bi, weighted, BCW, GPM, BDOF, PROF blocks "
        "against predict_inter)\n"
        "  inter.sad, dmvr[2][2], fetch_samples, bdof_fetch_samples: steps of the DMVR / BDOF / PROF searches inside inter_kernel; "
        "sad returns a value to host control flow (test_dmvr_bdof_match_oracle compares pixels and refined vectors)\n"
        "  inter.put_ciip, intra.*: prediction from reconstructed neighbours, driven by VVCLocalContext state, not by a block "
        "(tests/test_gpu_intra.py against ff_vvc_reconstruct)\n"
        "  itx.add_residual_joint, pred_residual_joint, transform_bdpcm: fused into residual_kernel (tests/test_gpu_residual.py "
        "tools joint, bdpcm against itransform)\n"
        "  lf.filter_luma / filter_chroma / ladf_level: per-edge calls with host-computed tc / beta; the device filters a whole "
        "picture from the edge tables (tests/test_gpu_deblock.py against ff_vvc_deblock_vertical / horizontal); ladf_level returns "
        "a value\n"
        "  sao.band_filter / edge_filter / edge_restore, alf.filter / filter_cc / classify / recon_coeff_and_clip: the device "
        "pass takes CTB records and classifies, derives coefficients and filters in one kernel; the members' per-block "
        "coefficient arrays and pre-padded source buffers have no device counterpart (tests/test_gpu_loopfilter.py against "
        "ff_vvc_sao_filter / ff_vvc_alf_filter, every CTB size, slices, tiles)\n"
        "  8-bit tables (uint8 planes) and log2_transform_range > 15: not built\n";
}

/* ---- the driver ------------------------------------------------------------------------------------------- */
static uint32_t g_rng;
static uint32_t rnd(void) { g_rng = g_rng * 1664525u + 1013904223u; return g_rng >> 8; }

enum { F_ITX, F_ADD, F_LMCS, F_PUT, F_UNI, F_UNIW, F_COUNT };

#define SRC_PAD   8
#define SRC_STRIDE ((MAX_CTU_SIZE + 2 * SRC_PAD) * 2)
#define SRC_BYTES (SRC_STRIDE * (MAX_CTU_SIZE + 2 * SRC_PAD))

/* out[f * 3 + 0] members exercised, + 1 calls compared, + 2 calls that differed; returns the device errors seen */
SHIM_EXPORT int vvcshim_check(int bd, unsigned seed, int *out)
{
    VVCDSPContext ref, dev;
    const int pmax = (1 << bd) - 1;
    g_rng = seed * 2654435761u + 12345u;
    g_device_errors = g_unrepresentable = 0;
    memset(out, 0, sizeof(int) * 3 * F_COUNT);
    memset(&ref, 0, sizeof(ref));
    ff_vvc_dsp_init(&ref, bd);
    vvcb200_dsp_init(&dev, bd);

    /* tests/checkasm/vvc_itx.c:58-92: every populated member, a random scan box, full-range coefficients */
    {
        int *c0 = malloc(sizeof(int) * 64 * 64), *c1 = malloc(sizeof(int) * 64 * 64);
        for (int lw = 0; lw < 7; lw++)
            for (int lh = 0; lh < 7; lh++)
                for (int trh = 0; trh < 3; trh++)
                    for (int trv = 0; trv < 3; trv++) {
                        if (!ref.itx.itx[trh][trv][lw][lh])
                            continue;
                        out[F_ITX * 3]++;
                        for (int rep = 0; rep < 3; rep++) {
                            const int w = 1 << lw, h = 1 << lh;
                            const int zw = FFMIN(trh == DCT2 ? 32 : 16, w), zh = FFMIN(trv == DCT2 ? 32 : 16, h);
                            const int nzw = rep == 0 ? zw : rnd() % zw + 1, nzh = rep == 0 ? zh : rnd() % zh + 1;
                            for (int y = 0; y < h; y++)
                                for (int x = 0; x < w; x++)
                                    c0[y * w + x] = c1[y * w + x] = (x < nzw && y < nzh) ? av_clip_intp2((int)(rnd() << 8) >> 8, 15) : 0;
                            ref.itx.itx[trh][trv][lw][lh](c0, nzw, nzh, 15, bd);
                            dev.itx.itx[trh][trv][lw][lh](c1, nzw, nzh, 15, bd);
                            out[F_ITX * 3 + 1]++;
                            out[F_ITX * 3 + 2] += memcmp(c0, c1, sizeof(int) * w * h) != 0;
                        }
                    }
        /* itx.add_residual: every transform-block shape */
        uint16_t *d0 = malloc(64 * 80 * 2), *d1 = malloc(64 * 80 * 2);
        out[F_ADD * 3] = 1;
        for (int w = 2; w <= 64; w <<= 1)
            for (int h = 2; h <= 64; h <<= 1) {
                for (int i = 0; i < 64 * 80; i++)
                    d0[i] = d1[i] = rnd() & pmax;
                for (int i = 0; i < w * h; i++)
                    c0[i] = (int)(rnd() % (4u << bd)) - (2 << bd);
                ref.itx.add_residual((uint8_t *)d0, c0, w, h, 80 * 2);
                dev.itx.add_residual((uint8_t *)d1, c0, w, h, 80 * 2);
                out[F_ADD * 3 + 1]++;
                out[F_ADD * 3 + 2] += memcmp(d0, d1, 64 * 80 * 2) != 0;
            }
        free(c0); free(c1); free(d0); free(d1);
    }
    /* lmcs.filter: a random LUT over blocks of every CTU-row shape */
    {
        uint16_t *lut = malloc(sizeof(uint16_t) << bd), *d0 = malloc(128 * 136 * 2), *d1 = malloc(128 * 136 * 2);
        out[F_LMCS * 3] = 1;
        for (int i = 0; i <= pmax; i++)
            lut[i] = rnd() & pmax;
        for (int w = 8; w <= 128; w <<= 1)
            for (int h = 8; h <= 128; h <<= 1) {
                for (int i = 0; i < 128 * 136; i++)
                    d0[i] = d1[i] = rnd() & pmax;
                ref.lmcs.filter((uint8_t *)d0, 136 * 2, w, h, (const uint8_t *)lut);
                dev.lmcs.filter((uint8_t *)d1, 136 * 2, w, h, (const uint8_t *)lut);
                out[F_LMCS * 3 + 1]++;
                out[F_LMCS * 3 + 2] += memcmp(d0, d1, 128 * 136 * 2) != 0;
            }
        free(lut); free(d0); free(d1);
    }
    /* tests/checkasm/vvc_mc.c:68-324: put / put_uni / put_uni_w, luma and chroma, pixels / h / v / hv, every w x h.
     * (taps as the decoder passes them: both from one filter set, fractions that match the member) */
    {
        uint8_t *src = malloc(SRC_BYTES);
        int16_t *t0 = malloc(MAX_PB_SIZE * MAX_PB_SIZE * 2), *t1 = malloc(MAX_PB_SIZE * MAX_PB_SIZE * 2);
        uint16_t *p0 = malloc(MAX_PB_SIZE * MAX_PB_SIZE * 2), *p1 = malloc(MAX_PB_SIZE * MAX_PB_SIZE * 2);
        for (int chroma = 0; chroma < 2; chroma++)
            for (int j = 0; j < 2; j++)
                for (int i = 0; i < 2; i++)
                    for (int w = chroma ? 2 : 4; w <= (chroma ? 64 : MAX_CTU_SIZE); w *= 2) {
                        const int k = av_log2(w) - 1;
                        if (ref.inter.put[chroma][k][j][i]) out[F_PUT * 3]++;
                        if (ref.inter.put_uni[chroma][k][j][i]) out[F_UNI * 3]++;
                        if (ref.inter.put_uni_w[chroma][k][j][i]) out[F_UNIW * 3]++;
                        for (int h = chroma ? 2 : 4; h <= (chroma ? 64 : MAX_CTU_SIZE); h *= 2) {
                            const int fr = chroma ? 32 : 16, set = chroma ? 0 : rnd() % 3;
                            const int mx = i ? rnd() % (fr - 1) + 1 : 0, my = j ? rnd() % (fr - 1) + 1 : 0;
                            const int8_t *hf = chroma ? ff_vvc_inter_chroma_filters[set][mx] : ff_vvc_inter_luma_filters[set][mx];
                            const int8_t *vf = chroma ? ff_vvc_inter_chroma_filters[set][my] : ff_vvc_inter_luma_filters[set][my];
                            const uint8_t *s = src + SRC_PAD * SRC_STRIDE + SRC_PAD * 2;
                            const int denom = rnd() % 8, wx = (int)(rnd() % 256) - 128, ox = (int)(rnd() % 256) - 128;
                            for (int n = 0; n < SRC_BYTES / 2; n++)
                                ((uint16_t *)src)[n] = rnd() & pmax;
                            if (ref.inter.put[chroma][k][j][i]) {
                                memset(t0, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                memset(t1, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                ref.inter.put[chroma][k][j][i](t0, s, SRC_STRIDE, h, hf, vf, w);
                                dev.inter.put[chroma][k][j][i](t1, s, SRC_STRIDE, h, hf, vf, w);
                                out[F_PUT * 3 + 1]++;
                                out[F_PUT * 3 + 2] += memcmp(t0, t1, MAX_PB_SIZE * MAX_PB_SIZE * 2) != 0;
                            }
                            if (ref.inter.put_uni[chroma][k][j][i]) {
                                memset(p0, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                memset(p1, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                ref.inter.put_uni[chroma][k][j][i]((uint8_t *)p0, MAX_PB_SIZE * 2, s, SRC_STRIDE, h, hf, vf, w);
                                dev.inter.put_uni[chroma][k][j][i]((uint8_t *)p1, MAX_PB_SIZE * 2, s, SRC_STRIDE, h, hf, vf, w);
                                out[F_UNI * 3 + 1]++;
                                out[F_UNI * 3 + 2] += memcmp(p0, p1, MAX_PB_SIZE * MAX_PB_SIZE * 2) != 0;
                            }
                            if (ref.inter.put_uni_w[chroma][k][j][i]) {
                                memset(p0, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                memset(p1, 0, MAX_PB_SIZE * MAX_PB_SIZE * 2);
                                ref.inter.put_uni_w[chroma][k][j][i]((uint8_t *)p0, MAX_PB_SIZE * 2, s, SRC_STRIDE, h, denom, wx, ox, hf, vf, w);
                                dev.inter.put_uni_w[chroma][k][j][i]((uint8_t *)p1, MAX_PB_SIZE * 2, s, SRC_STRIDE, h, denom, wx, ox, hf, vf, w);
                                out[F_UNIW * 3 + 1]++;
                                out[F_UNIW * 3 + 2] += memcmp(p0, p1, MAX_PB_SIZE * MAX_PB_SIZE * 2) != 0;
                            }
                        }
                    }
        free(src); free(t0); free(t1); free(p0); free(p1);
    }
    return g_device_errors + (g_unrepresentable << 16);
}
