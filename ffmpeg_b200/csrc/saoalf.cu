// saoalf.cu — stages SAO and ALF (+ CC-ALF) as ONE out-of-place picture pass (what Level B runs).
// Reference semantics: ff_vvc_sao_filter (libavcodec/vvc/vvc_filter.c:154-298) with sao.band_filter / edge_filter /
// edge_restore (libavcodec/h26x/h2656_sao_template.c:24-215), then ff_vvc_alf_filter (vvc_filter.c:1005-1319) with
// alf.classify / recon_coeff_and_clip / filter[luma|chroma] / filter_cc (vvc_filter_template.c:38-408).
//
// ALF and CC-ALF read SAO output only (vvc_filter.c:1297-1316), and SAO reads deblocked samples only, so a CTA that
// holds the deblocked samples of a region plus a ring of 4 can produce the SAO output of the region plus a ring of 3
// — everything ALF needs — without the SAO picture ever existing in memory.  One CTA owns a 64x64 luma region and the
// two co-located 32x32 chroma regions:
//   1. stage the deblocked windows with 128-bit row loads (luma 80 x 72, chroma 48 x 38 per plane);
//   2. SAO on the windows in shared memory, 8 samples per thread in packed 16x2 arithmetic (band / edge classes; CTB
//      sides that are picture borders or carry a no-filter flag take the per-sample rules);
//   3. ALF luma per 32x32 sub-tile (a sub-tile never straddles a CTB): the padded window of alf_prepare_buffer is a
//      clamped copy out of the SAO window, then the shared tile pipeline of alf_dev.cuh;
//   4. ALF chroma + CC-ALF per plane, CC-ALF reading the SAO'd luma window in shared memory.
// Samples outside the picture are never consumed: picture borders are always "edges" for ALF (ctb_geom) and SAO's
// border rules never read across them.
#include "alf_dev.cuh"

namespace vvcb200 {

constexpr int kFT   = 64;                 // luma region side
constexpr int kFYS  = 88;                 // staged luma row stride (samples): 80 columns X0-8 .. X0+71, 176 B = 11 x 16 B
constexpr int kFYR  = kFT + 8;            // deblocked luma rows Y0-4 .. Y0+67
constexpr int kFYSR = kFT + 6;            // SAO'd luma rows Y0-3 .. Y0+66
constexpr int kFCS  = 56;                 // staged chroma row stride: 48 columns x0c-8 .. x0c+39, 112 B = 7 x 16 B
constexpr int kFCR  = kFT / 2 + 6;        // deblocked chroma rows y0c-3 .. y0c+34
constexpr int kFCSR = kFT / 2 + 4;        // SAO'd chroma rows y0c-2 .. y0c+33

struct SaoAlfArgs {
    AlfArgs alf;                          // alf.ctbs == nullptr: ALF off everywhere
    const VVCB200SaoCtb *sao;             // nullptr: SAO off everywhere
};

// A window of one plane in shared memory: sample (x, y) of the plane at p[(y - y0) * stride + (x - x0)].
struct PlaneWin {
    const uint16_t *p;
    int stride, x0, y0, w, h;             // origin in plane coordinates, size in samples
    __device__ __forceinline__ int at(int x, int y) const
    {
        return p[clip3(y - y0, 0, h - 1) * stride + clip3(x - x0, 0, w - 1)];
    }
};

// One sample by the reference's rules, including everything edge_restore does (picture borders, no-filter sides).
// (x, y): plane coordinates; the CTB is (cx0, cy0, cw, ch); (W, H): plane size.
__device__ __forceinline__ int sao_sample_win(const PlaneWin &src, const VVCB200SaoCtb &p, int c, int x, int y, int cx0,
                                              int cy0, int cw, int ch, int W, int H, int bd)
{
    const int s = src.at(x, y);
    const int type = p.type_idx[c];
    if (type == VVCB200_SAO_OFF)
        return s;
    const int16_t *ov = p.offset_val[c];
    if (type == VVCB200_SAO_BAND) {
        const int k = (((s >> (bd - 5)) & 31) - p.band_position[c]) & 31;
        return clip_pixel(s + (k < 4 ? ov[k + 1] : 0), bd);
    }
    const int eo = p.eo_class[c];
    const int lx = x - cx0, ly = y - cy0;
    const bool b_left = cx0 == 0, b_top = cy0 == 0, b_right = cx0 + cw == W, b_bottom = cy0 + ch == H;
    if (eo != 1 && ((b_left && lx == 0) || (b_right && lx == cw - 1)))
        return clip_pixel(s + ov[0], bd);
    if (eo != 0 && ((b_top && ly == 0) || (b_bottom && ly == ch - 1)))
        return clip_pixel(s + ov[0], bd);
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
    const int ax = eo == 1 ? 0 : (eo == 3 ? 1 : -1), ay = eo == 0 ? 0 : -1;
    const int na = src.at(x + ax, y + ay), nb = src.at(x - ax, y - ay);
    const int d = (s > na) - (s < na) + (s > nb) - (s < nb);
    const int idx = d == -2 ? 1 : d == -1 ? 2 : d == 0 ? 0 : d == 1 ? 3 : 4;
    return clip_pixel(s + ov[idx], bd);
}

// The 8 samples of a chunk one by one, for the few chunks that touch a picture border or a no-filter side.  Deliberately a
// real call with a rolled loop: inlined eight-fold into both call sites this rule set was two thirds of the kernel's code,
// and the instruction cache paid for it on every other chunk.
__device__ __noinline__ uint4 sao_chunk_slow(const uint16_t *wp, int wstride, int wx0, int wy0, int ww, int wh,
                                             const VVCB200SaoCtb *pp, int c, int x, int y, int cx0, int cy0, int cw, int ch,
                                             int W, int H, int bd)
{
    const PlaneWin src = { wp, wstride, wx0, wy0, ww, wh };
    const VVCB200SaoCtb p = *pp;
    unsigned long long lo = 0, hi = 0;
#pragma unroll 1
    for (int j = 0; j < 8; j++) {
        const unsigned long long v = x + j < W ? (unsigned long long)sao_sample_win(src, p, c, x + j, y, cx0, cy0, cw, ch, W, H, bd) : 0;
        if (j < 4)
            lo |= v << (16 * j);
        else
            hi |= v << (16 * (j - 4));
    }
    return make_uint4((uint32_t)lo, (uint32_t)(lo >> 32), (uint32_t)hi, (uint32_t)(hi >> 32));
}

// Offsets looked up for both halves of a packed index word (each half 0..4) with two PRMTs: byte i of (lo, hi) holds entry
// i, selector nibbles (k, k | 8) pick the byte and replicate its sign, which sign-extends it into the 16-bit half.
__device__ __forceinline__ uint32_t lut5_s16x2(uint32_t idx2, uint32_t lo, uint32_t hi)
{
    const uint32_t t = idx2 * 0x11u;                              // per half: k | k << 4
    const uint32_t sel = __byte_perm(t, 0, 0x4420) | 0x8080u;     // bytes: t.b0, t.b2  ->  nibbles (klo, klo|8, khi, khi|8)
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(sel));
    return r;
}

// SAO of the 8 samples at plane position (x, y), x a multiple of 8, from the deblocked window `src`; result as 4 words.
__device__ __forceinline__ uint4 sao_chunk(const SaoAlfArgs &a, const PlaneWin &src, int c, int x, int y)
{
    const int W = a.alf.src.w[c], H = a.alf.src.h[c], bd = a.alf.bd;
    const uint16_t *rc = src.p + (y - src.y0) * src.stride + (x - src.x0);
    const uint4 vc = *reinterpret_cast<const uint4 *>(rc);
    if (!a.sao || x < 0 || y < 0 || x >= W || y >= H)
        return vc;
    const int l2 = a.alf.ctb_log2 - (c ? 1 : 0);
    const int cx = x >> l2, cy = y >> l2;
    const VVCB200SaoCtb *pp = a.sao + cy * a.alf.ctb_width + cx;
    const int type = pp->type_idx[c];
    if (type == VVCB200_SAO_OFF)
        return vc;
    const int cx0 = cx << l2, cy0 = cy << l2;
    const int cw = min(1 << l2, W - cx0), ch = min(1 << l2, H - cy0);
    bool slow = x + 8 > W;
    if (type == VVCB200_SAO_EDGE) {
        // only the chunks on a CTB side that is a picture border or carries a no-filter flag (its own or a corner's)
        const int nf = pp->no_filter;
        const bool left = cx0 == 0 || (nf & (1 | 16 | 128)), right = cx0 + cw == W || (nf & (2 | 32 | 64));
        const bool top = cy0 == 0 || (nf & (4 | 16 | 32)), bottom = cy0 + ch == H || (nf & (8 | 64 | 128));
        slow = slow || (x == cx0 && left) || (x + 8 >= cx0 + cw && right) || (y == cy0 && top) || (y == cy0 + ch - 1 && bottom);
    }
    const uint32_t cw4[4] = { vc.x, vc.y, vc.z, vc.w };
    uint32_t out[4];
    if (slow)
        return sao_chunk_slow(src.p, src.stride, src.x0, src.y0, src.w, src.h, pp, c, x, y, cx0, cy0, cw, ch, W, H, bd);
    const int16_t *ov = pp->offset_val[c];
    const uint32_t pmax2 = (uint32_t)((1 << bd) - 1) * 0x10001u;
    if (type == VVCB200_SAO_BAND) {
        // bands pos .. pos + 3 carry offsets 1 .. 4; every other band entry 4 of the table (zero)
        const uint32_t lo = (uint32_t)(ov[1] & 0xff) | ((uint32_t)(ov[2] & 0xff) << 8) | ((uint32_t)(ov[3] & 0xff) << 16) |
                            ((uint32_t)(ov[4] & 0xff) << 24);
        const uint32_t pos2 = (uint32_t)pp->band_position[c] * 0x10001u;
        const int sh = bd - 5;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t k = __vminu2(__vsub2((cw4[j] >> sh) & 0x001f001fu, pos2) & 0x001f001fu, 0x00040004u);
            out[j] = __vmins2(__vmaxs2(__vadd2(cw4[j], lut5_s16x2(k, lo, 0u)), 0u), pmax2);
        }
        return make_uint4(out[0], out[1], out[2], out[3]);
    }
    // edge classes: neighbour a = (x + AX, y + AY), b = (x - AX, y - AY)   (h2656_sao_template.c:54-59)
    const int eo = pp->eo_class[c];
    const int ax = eo == 1 ? 0 : (eo == 3 ? 1 : -1), ay = eo == 0 ? 0 : -1;
    const uint16_t *ra = rc + ay * src.stride, *rb = rc - ay * src.stride;
    const uint4 va = *reinterpret_cast<const uint4 *>(ra), vb = *reinterpret_cast<const uint4 *>(rb);
    uint32_t aw[4] = { va.x, va.y, va.z, va.w }, bw[4] = { vb.x, vb.y, vb.z, vb.w };
    if (ax) {
        // shift the rows by one sample: the sample just outside the chunk comes from the window (clamped at the window's
        // own ends, where the result is never consumed)
        const int col = x - src.x0;
        const uint32_t left_a = ra[col > 0 ? -1 : 0], right_a = ra[col + 8 < src.w ? 8 : 7];
        const uint32_t left_b = rb[col > 0 ? -1 : 0], right_b = rb[col + 8 < src.w ? 8 : 7];
        uint32_t ta[4], tb[4];
        if (ax < 0) {                     // a[j] = rowA[j - 1], b[j] = rowB[j + 1]
            ta[0] = __funnelshift_r(left_a << 16, aw[0], 16);
            ta[1] = __funnelshift_r(aw[0], aw[1], 16);
            ta[2] = __funnelshift_r(aw[1], aw[2], 16);
            ta[3] = __funnelshift_r(aw[2], aw[3], 16);
            tb[0] = __funnelshift_r(bw[0], bw[1], 16);
            tb[1] = __funnelshift_r(bw[1], bw[2], 16);
            tb[2] = __funnelshift_r(bw[2], bw[3], 16);
            tb[3] = __funnelshift_r(bw[3], right_b, 16);
        } else {                          // a[j] = rowA[j + 1], b[j] = rowB[j - 1]
            ta[0] = __funnelshift_r(aw[0], aw[1], 16);
            ta[1] = __funnelshift_r(aw[1], aw[2], 16);
            ta[2] = __funnelshift_r(aw[2], aw[3], 16);
            ta[3] = __funnelshift_r(aw[3], right_a, 16);
            tb[0] = __funnelshift_r(left_b << 16, bw[0], 16);
            tb[1] = __funnelshift_r(bw[0], bw[1], 16);
            tb[2] = __funnelshift_r(bw[1], bw[2], 16);
            tb[3] = __funnelshift_r(bw[2], bw[3], 16);
        }
#pragma unroll
        for (int j = 0; j < 4; j++) {
            aw[j] = ta[j];
            bw[j] = tb[j];
        }
    }
    // table in the order of sign(s - a) + sign(s - b) + 2 = 0..4: offsets 1, 2, 0, 3, 4 (h2656_sao_template.c:50)
    const uint32_t lo = (uint32_t)(ov[1] & 0xff) | ((uint32_t)(ov[2] & 0xff) << 8) | ((uint32_t)(ov[0] & 0xff) << 16) |
                        ((uint32_t)(ov[3] & 0xff) << 24);
    const uint32_t hi = (uint32_t)(ov[4] & 0xff);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint32_t da = __vmins2(__vmaxs2(__vsub2(cw4[j], aw[j]), 0xffffffffu), 0x00010001u);
        const uint32_t db = __vmins2(__vmaxs2(__vsub2(cw4[j], bw[j]), 0xffffffffu), 0x00010001u);
        const uint32_t k = __vadd2(__vadd2(da, db), 0x00020002u);
        out[j] = __vmins2(__vmaxs2(__vadd2(cw4[j], lut5_s16x2(k, lo, hi)), 0u), pmax2);
    }
    return make_uint4(out[0], out[1], out[2], out[3]);
}

#ifndef VVC_SAOALF_MINB
#define VVC_SAOALF_MINB 4
#endif
template <bool kTable>
__global__ void __launch_bounds__(256, VVC_SAOALF_MINB) sao_alf_kernel(const SaoAlfArgs A)
{
    __shared__ __align__(16) uint16_t s_sao_y[kFYSR * kFYS];              // SAO'd luma, rows Y0-3 .., columns X0-8 ..
    __shared__ __align__(16) uint16_t s_sao_c[2][kFCSR * kFCS];           // SAO'd chroma, rows y0c-2 .., columns x0c-8 ..
    // deblocked windows during steps 1-2, the ALF tile state afterwards
    constexpr int kDbkBytes = (kFYR * kFYS + 2 * kFCR * kFCS) * 2;
    static_assert(sizeof(AlfLumaShared) <= kDbkBytes, "ALF tile state must fit the dead deblocked windows");
    __shared__ __align__(16) unsigned char s_work[kDbkBytes];
    uint16_t *const s_dbk_y = reinterpret_cast<uint16_t *>(s_work);
    uint16_t *const s_dbk_c = s_dbk_y + kFYR * kFYS;
    AlfLumaShared &S = *reinterpret_cast<AlfLumaShared *>(s_work);

    const AlfArgs &a = A.alf;
    const int t = threadIdx.x;
    const int X0 = blockIdx.x * kFT, Y0 = blockIdx.y * kFT;
    const int xc0 = X0 >> 1, yc0 = Y0 >> 1;

    // ---- 1. stage the deblocked windows (rows / chunks outside the plane's allocation read as zero, never consumed)
    {
        constexpr int YCH = 10, CCH = 6;
        constexpr int NY = kFYR * YCH, NC = kFCR * CCH;
        for (int i = t; i < NY + 2 * NC; i += 256) {
            int c, r, k, gx, gy;
            uint16_t *d;
            if (i < NY) {
                c = 0; r = i / YCH; k = i - r * YCH;
                gx = X0 - 8 + 8 * k; gy = Y0 - 4 + r;
                d = s_dbk_y + r * kFYS + 8 * k;
            } else {
                const int j = i - NY;
                c = 1 + (j >= NC);
                const int jj = j - (c - 1) * NC;
                r = jj / CCH; k = jj - r * CCH;
                gx = xc0 - 8 + 8 * k; gy = yc0 - 3 + r;
                d = s_dbk_c + (c - 1) * kFCR * kFCS + r * kFCS + 8 * k;
            }
            uint4 v = make_uint4(0, 0, 0, 0);
            if (gy >= 0 && gy < a.src.h[c] && gx >= 0 && gx + 8 <= a.src.stride[c])
                v = *reinterpret_cast<const uint4 *>(a.src.p[c] + (size_t)gy * a.src.stride[c] + gx);
            *reinterpret_cast<uint4 *>(d) = v;
        }
    }
    __syncthreads();

    // ---- 2. SAO into the SAO windows
    {
        constexpr int YCH = 10, CCH = 6;
        constexpr int NY = kFYSR * YCH, NC = kFCSR * CCH;
        const PlaneWin wy = { s_dbk_y, kFYS, X0 - 8, Y0 - 4, 80, kFYR };
        for (int i = t; i < NY + 2 * NC; i += 256) {
            if (i < NY) {
                const int r = i / YCH, k = i - r * YCH;
                *reinterpret_cast<uint4 *>(s_sao_y + r * kFYS + 8 * k) = sao_chunk(A, wy, 0, X0 - 8 + 8 * k, Y0 - 3 + r);
            } else {
                const int j = i - NY;
                const int c = 1 + (j >= NC);
                const int jj = j - (c - 1) * NC;
                const int r = jj / CCH, k = jj - r * CCH;
                const PlaneWin wc = { s_dbk_c + (c - 1) * kFCR * kFCS, kFCS, xc0 - 8, yc0 - 3, 48, kFCR };
                *reinterpret_cast<uint4 *>(s_sao_c[c - 1] + r * kFCS + 8 * k) = sao_chunk(A, wc, c, xc0 - 8 + 8 * k, yc0 - 2 + r);
            }
        }
    }
    __syncthreads();

    VVCB200AlfCtb off = {};
    auto luma_at = [&](int x, int y) { return s_sao_y[(y - (Y0 - 3)) * kFYS + (x - (X0 - 8))]; };

    // ---- 3. ALF luma, four 32x32 sub-tiles
    for (int st = 0; st < 4; st++) {
        const int tx0 = X0 + (st & 1) * kTile, ty0 = Y0 + (st >> 1) * kTile;
        if (tx0 >= a.src.w[0] || ty0 >= a.src.h[0])
            continue;
        const VVCB200AlfCtb cp = a.ctbs ? a.ctbs[(ty0 >> a.ctb_log2) * a.ctb_width + (tx0 >> a.ctb_log2)] : off;
        if (!cp.ctb_flag[0]) {                            // ALF off for this CTB: the SAO samples are the output
            uint16_t *dst = a.dst.p[0];
            const int dstride = a.dst.stride[0];
            for (int i = t; i < kTile * kTile / 4; i += 256) {
                const int r = i >> 3, x = tx0 + (i & 7) * 4, y = ty0 + r;
                if (x < a.src.w[0] && y < a.src.h[0])
                    *reinterpret_cast<uint2 *>(dst + (size_t)y * dstride + x) =
                        *reinterpret_cast<const uint2 *>(&s_sao_y[(y - (Y0 - 3)) * kFYS + (x - (X0 - 8))]);
            }
            continue;
        }
        const CtbGeom g = ctb_geom(a, 0, tx0, ty0, cp.edges);
        if (t == 0)
            S.wide = 0;
        if (!g.edges) {
            // a CTB with no side to clamp at (four in five of a 4K picture's): alf_prepare_buffer's window IS the SAO
            // window, the filter reads it where it lies
            alf_luma_filter_tile<kTable, kFYS>(a, cp, g, tx0, ty0, S,
                                               s_sao_y + (ty0 - kHalo - (Y0 - 3)) * kFYS + (tx0 - kColOrg - (X0 - 8)));
            __syncthreads();
            continue;
        }
        for (int r = t >> 5; r < kPad; r += 8)            // a warp per staged row, lanes along it
            for (int col = t & 31; col < kPad; col += 32) {
                int x = tx0 + col - kHalo, y = ty0 + r - kHalo;
                pad_coords(g, x, y);
                S.pix[r * kPadStr + col + kColOrg - kHalo] = luma_at(x, y);
            }
        __syncthreads();
        alf_luma_filter_tile<kTable>(a, cp, g, tx0, ty0, S, S.pix);
        __syncthreads();
    }

    // ---- 4. ALF chroma + CC-ALF, per plane and per sub-tile that stays inside one chroma CTB
    const int T = min(kTile, 1 << (a.ctb_log2 - 1));
    const int per_side = (kFT / 2) / T;
    for (int c = 1; c < 3; c++) {
        const uint16_t *sc = s_sao_c[c - 1];
        auto cur_at = [&](int x, int y) { return sc[(y - (yc0 - 2)) * kFCS + (x - (xc0 - 8))]; };
        for (int st = 0; st < per_side * per_side; st++) {
            const int tx0 = xc0 + (st % per_side) * T, ty0 = yc0 + (st / per_side) * T;
            if (tx0 >= a.src.w[c] || ty0 >= a.src.h[c])
                continue;
            const int l2 = a.ctb_log2 - 1;
            const VVCB200AlfCtb cp = a.ctbs ? a.ctbs[(ty0 >> l2) * a.ctb_width + (tx0 >> l2)] : off;
            const CtbGeom g = ctb_geom(a, c, tx0, ty0, cp.edges);
            const int Sd = T + 4;
            if (!g.edges) {                               // as for luma: nothing to clamp, the SAO window is the padded window
                alf_chroma_filter_tile(a, cp, g, c, T, tx0, ty0, sc + (ty0 - 2 - (yc0 - 2)) * kFCS + (tx0 - 2 - (xc0 - 8)), kFCS,
                                       cur_at, luma_at);
                continue;
            }
            if (cp.ctb_flag[c]) {
                for (int r = t >> 5; r < Sd; r += 8)
                    for (int col = t & 31; col < Sd; col += 32) {
                        int x = tx0 + col - 2, y = ty0 + r - 2;
                        pad_coords(g, x, y);
                        S.pix[r * Sd + col] = cur_at(x, y);
                    }
            }
            __syncthreads();
            alf_chroma_filter_tile(a, cp, g, c, T, tx0, ty0, S.pix, Sd, cur_at, luma_at);
            __syncthreads();
        }
    }
}

}  // namespace vvcb200

using namespace vvcb200;

extern "C" size_t vvcb200_sao_alf_scratch_bytes(void) { return (size_t)kAlfTableWords * sizeof(uint32_t); }

extern "C" int vvcb200_sao_alf_picture(const VVCB200Planes *dst, const VVCB200Planes *src, const VVCB200SaoCtb *d_sao,
                                       const VVCB200AlfCtb *d_alf, const VVCB200AlfParams *d_params, int ctb_log2_size,
                                       int bit_depth, void *d_scratch, void *stream)
{
    if (!dst || !src || (d_alf && !d_params) || ctb_log2_size < 5 || ctb_log2_size > 7 || (bit_depth != 10 && bit_depth != 12)) {
        set_error("vvcb200_sao_alf_picture: bad argument");
        return VVCB200_ERR_ARG;
    }
    SaoAlfArgs A;
    A.alf.dst = make_view(dst);
    A.alf.src = make_view(src);
    A.alf.ctbs = d_alf;
    A.alf.params = d_params;
    A.alf.ctb_log2 = ctb_log2_size;
    A.alf.ctb_width = (src->width[0] + (1 << ctb_log2_size) - 1) >> ctb_log2_size;
    A.alf.bd = bit_depth;
    A.sao = d_sao;
    for (int c = 0; c < 3; c++) {
        if (((uintptr_t)A.alf.src.p[c] & 15) || (A.alf.src.stride[c] & 7) || ((uintptr_t)A.alf.dst.p[c] & 7) ||
            (A.alf.dst.stride[c] & 3) || A.alf.dst.w[c] != A.alf.src.w[c] || A.alf.dst.h[c] != A.alf.src.h[c] ||
            A.alf.dst.p[c] == A.alf.src.p[c]) {
            set_error("vvcb200_sao_alf_picture: planes must be distinct and equally sized; source planes 16-byte aligned with "
                      "16-byte multiple strides, destination planes 8-byte aligned with 8-byte multiple strides");
            return VVCB200_ERR_ARG;
        }
    }
    if (src->width[1] * 2 != src->width[0] || src->height[1] * 2 != src->height[0]) {
        set_error("vvcb200_sao_alf_picture: 4:2:0 pictures only");
        return VVCB200_ERR_UNSUP;
    }
    dim3 grid((src->width[0] + kFT - 1) / kFT, (src->height[0] + kFT - 1) / kFT, 1);
    A.alf.table = nullptr;
    if (d_alf && d_scratch) {
        if ((uintptr_t)d_scratch & 15) {
            set_error("vvcb200_sao_alf_picture: d_scratch must be 16-byte aligned");
            return VVCB200_ERR_ARG;
        }
        // the picture's coefficient / clip table first: a few microseconds that every 4x4 block then does not spend
        VVC_CUDA(cudaMemsetAsync(d_scratch, 0, kAlfTableHead * sizeof(uint32_t), (cudaStream_t)stream));
        alf_coef_table_kernel<<<(kAlfSets * 25 * 4 * 12 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_params, bit_depth, (uint32_t *)d_scratch);
        VVC_LAUNCH_CHECK();
        A.alf.table = (const uint32_t *)d_scratch;
        sao_alf_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
    } else {
        sao_alf_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(A);
    }
    VVC_LAUNCH_CHECK();
    return 0;
}
