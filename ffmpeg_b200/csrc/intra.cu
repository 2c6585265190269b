// intra.cu — stage RECON, intra part: intra prediction (+ CIIP blend, + residual) of a picture's intra blocks.
//
// Replaces, for a whole picture, the per-CTU loop of ff_vvc_reconstruct (vvc_intra.c:498-536) together with the
// intra DSP of vvc_intra_template.c (edge preparation 450-598, planar/DC/angular/PDPC 600-705 + 838-1001,
// MIP 707-836, CCLM 29-375).
//
// Intra prediction reads reconstructed neighbours, so the work is a dependency graph, not a batch.  The reference
// walks it as a CTU wavefront (vvc_thread.c:156-184); here the graph is resolved per block:
//   * every record marks the 4x4 (chroma: 4x2) units it will write as "pending" (mark kernel);
//   * warps draw records from a ticket counter in array order — a topological order, see vvcb200_intra_schedule —
//     and spin until the units holding the neighbour samples they read are no longer pending;
//   * a warp that finished a record publishes its samples (__threadfence) and clears the record's marks.
// A ticket is only ever held by a resident warp and waits only on lower tickets, so the scheme cannot deadlock
// whatever the grid size; parallelism is what the picture's block graph offers (blocks of one anti-diagonal of
// CTUs and independent blocks inside a CTU run concurrently), which is far more than one warp per CTU row.
#include "common.cuh"

#define VVC_TABLE_QUAL static __device__ const
#include "vvc_tables.h"

namespace vvcb200 {
namespace {

constexpr int kWarps   = 4;       // warps per CTA
constexpr int kEdgeLen = 256;     // samples per edge array: indices -72 .. 183 are addressable
constexpr int kEdgeOrg = 72;
constexpr unsigned kFull = 0xffffffffu;

struct WarpScratch {
    uint16_t edge[4][kEdgeLen];   // left, top, filtered left, filtered top
    int      red[16];             // MIP: reduced boundary
    uint16_t mip[64];             // MIP: reduced prediction, row-major at the reduced size
};

struct IntraArgs {
    PlanesView pic;
    const int16_t *res[3];
    int res_stride[3];            // in samples
    const VVCB200IntraBlock *blocks;
    int n;
    int bd, ctb_log2, collocated;
    uint8_t *marks[3];            // pending marks per plane
    int mpitch[3];
    int *ticket;
    uint16_t lmcs_pivot[17], lmcs_coeff[16];
    int lmcs_min_bin, lmcs_max_bin;
};

__device__ __forceinline__ int ilog2(int v) { return 31 - __clz(v); }
__device__ __forceinline__ int unit_shift(int c) { return c ? 1 : 2; }     // rows per unit: 4 (luma) / 2 (chroma)

__device__ __forceinline__ VVCB200IntraBlock load_record(const VVCB200IntraBlock *p)
{
    union { VVCB200IntraBlock b; uint32_t w[7]; } u;
    const uint32_t *q = reinterpret_cast<const uint32_t *>(p);
#pragma unroll
    for (int i = 0; i < 7; i++)
        u.w[i] = __ldg(q + i);
    return u.b;
}

// ---- pending marks -------------------------------------------------------------------------------------------
__device__ __forceinline__ bool row_pending(const IntraArgs &a, int c, int uy, int ux0, int n, int lane)
{
    bool p = false;
    for (int i = lane; i < n; i += 32)
        p |= *((const volatile uint8_t *)a.marks[c] + (size_t)uy * a.mpitch[c] + ux0 + i) != 0;
    return p;
}

__device__ __forceinline__ bool col_pending(const IntraArgs &a, int c, int ux, int uy0, int n, int lane)
{
    bool p = false;
    for (int i = lane; i < n; i += 32)
        p |= *((const volatile uint8_t *)a.marks[c] + (size_t)(uy0 + i) * a.mpitch[c] + ux) != 0;
    return p;
}

// units touched by samples [p0, p0 + n) along an axis with 1 << s samples per unit
__device__ __forceinline__ int unit_count(int p0, int n, int s) { return n > 0 ? ((p0 + n - 1) >> s) - (p0 >> s) + 1 : 0; }

__device__ void set_marks(const IntraArgs &a, int c, int x0, int y0, int w, int h, uint8_t v, int lane)
{
    const int s = unit_shift(c);
    const int ux = x0 >> 2, uy = y0 >> s, nx = unit_count(x0, w, 2), ny = unit_count(y0, h, s);
    for (int i = lane; i < nx * ny; i += 32)
        *((volatile uint8_t *)a.marks[c] + (size_t)(uy + i / nx) * a.mpitch[c] + ux + i % nx) = v;
}

// the samples a record writes: an ISP luma record stands for its whole coding block
__device__ __forceinline__ void record_extent(const VVCB200IntraBlock &b, int &w, int &h)
{
    const bool isp = (b.flags & VVCB200_IB_ISP) && !b.c_idx;
    w = isp ? b.cb_w : b.w;
    h = isp ? b.cb_h : b.h;
}

__global__ void intra_mark_kernel(IntraArgs a)
{
    const int lane = threadIdx.x & 31;
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (t == 0 && lane == 0)
        *a.ticket = 0;
    if (t >= a.n)
        return;
    const VVCB200IntraBlock b = load_record(a.blocks + t);
    int w, h;
    record_extent(b, w, h);
    set_marks(a, b.c_idx, b.x0, b.y0, w, h, 1, lane);
    if (b.mode >= 81)
        set_marks(a, 2, b.x0, b.y0, w, h, 1, lane);
}

// ---- mode helpers (vvc_intra.c:538-660) ---------------------------------------------------------------------------
__device__ __forceinline__ bool ref_filter_mode(int mode)
{
    // modes whose angle hits integer sample positions: the reference smooths the edge instead of interpolating
    switch (mode) {
    case -14: case -12: case -10: case -6: case 0: case 2: case 34: case 66: case 72: case 76: case 78: case 80:
        return true;
    }
    return false;
}

__device__ __forceinline__ bool is_plain_mode(int mode) { return mode == 0 || mode == 1 || mode == 18 || mode == 50; }

__device__ __forceinline__ int nscale_of(int lw, int lh, int mode)
{
    if (is_plain_mode(mode))
        return (lw + lh - 2) >> 2;
    const int ns = (mode >= 50 ? lh : lw) - ilog2(3 * VVC_INTRA_INV_ANGLE[mode + 14] - 2) + 8;
    return min(ns, 2);
}

__device__ __forceinline__ bool need_pdpc(int w, int h, int lw, int lh, bool bdpcm, int mode, int ref_idx)
{
    if (w >= 4 && h >= 4 && !ref_idx && !bdpcm) {
        if (is_plain_mode(mode))
            return true;
        if (mode > 18 && mode < 50)
            return false;
        return nscale_of(lw, lh, mode) >= 0;
    }
    return false;
}

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1)
        v += __shfl_xor_sync(kFull, v, o);
    return v;
}

// ---- luma-dependent chroma residual scaling (vvc_intra_template.c:376-449) ------------------------------------------
// The factor of a VPDU comes from the mean of the reconstructed luma samples left of and above it.
__device__ int vpdu_chroma_scale(const IntraArgs &a, const VVCB200IntraBlock &b, int lane)
{
    const int size = min(1 << a.ctb_log2, 64), ls = ilog2(size);
    const int x = b.vpdu_x << ls, y = b.vpdu_y << ls;
    const bool al = b.vpdu_avail & 1, at = b.vpdu_avail & 2;
    const int rows = min(size, a.pic.h[0] - y), cols = min(size, a.pic.w[0] - x);
    for (;;) {
        bool p = false;
        if (al) p |= col_pending(a, 0, (x >> 2) - 1, y >> 2, unit_count(y, rows, 2), lane);
        if (at) p |= row_pending(a, 0, (y >> 2) - 1, x >> 2, unit_count(x, cols, 2), lane);
        if (!__any_sync(kFull, p))
            break;
        __nanosleep(64);
    }
    __threadfence();
    const uint16_t *luma = a.pic.p[0] + (size_t)y * a.pic.stride[0] + x;
    int sum = 0, cnt = 0;
    if (al) {                                                             // past the picture edge the last sample repeats
        for (int i = lane; i < size; i += 32)
            sum += __ldcg(luma - 1 + (ptrdiff_t)a.pic.stride[0] * min(i, rows - 1));
        cnt = size;
    }
    if (at) {
        for (int i = lane; i < size; i += 32)
            sum += __ldcg(luma - a.pic.stride[0] + min(i, cols - 1));
        cnt += size;
    }
    sum = warp_sum(sum);
    const int mean = cnt ? (sum + (cnt >> 1)) >> ilog2(cnt) : 1 << (a.bd - 1);
    int i = a.lmcs_min_bin;
    while (i <= a.lmcs_max_bin && mean >= a.lmcs_pivot[i + 1])
        i++;
    return a.lmcs_coeff[min(i, 15)];
}

__device__ __forceinline__ int scale_chroma_residual(int r, int scale, int bd)
{
    const int c = clip_intp2(r, bd);
    return c > 0 ? (c * scale + (1 << 10)) >> 11 : -((-c * scale + (1 << 10)) >> 11);
}

// ---- one angular / planar / DC / MIP block --------------------------------------------------------------------------
struct Store {                    // what happens to a predicted sample on its way to the picture
    uint16_t *dst;
    const int16_t *res;
    int dstride, rstride, bd, ciip_w, lscale;
    __device__ __forceinline__ void put(int x, int y, int pred) const
    {
        uint16_t *d = dst + (size_t)y * dstride + x;
        if (ciip_w)                                                       // put_ciip, vvc_inter_template.c:60-76
            pred = (pred * ciip_w + (int)__ldcg(d) * (4 - ciip_w) + 2) >> 2;
        if (res) {                                                        // add_residual, vvc_itx_1d / vvcdsp.h:154
            int r = res[(size_t)y * rstride + x];
            if (lscale)
                r = scale_chroma_residual(r, lscale, bd);
            pred = clip_pixel(pred + r, bd);
        }
        *d = (uint16_t)pred;
    }
};

// skip_wait: bit 0 = the row above (and the corner), bit 1 = the column to the left (and the corner) were written by
// this warp itself (earlier ISP sub-partitions) and carry this record's own marks
__device__ void predict_record(const IntraArgs &a, const VVCB200IntraBlock &b, WarpScratch &ws, int lane, int skip_wait = 0)
{
    const int c = b.c_idx, w = b.w, h = b.h, mode = b.mode, ref_idx = b.ref_idx, bd = a.bd;
    const int lw = ilog2(w), lh = ilog2(h);
    const bool is_mip = b.flags & VVCB200_IB_MIP, isp = b.flags & VVCB200_IB_ISP;
    const bool pdpc = need_pdpc(w, h, lw, lh, b.flags & VVCB200_IB_BDPCM, mode, ref_idx);
    const bool rff = !is_mip && ref_filter_mode(mode);
    const bool smooth = !ref_idx && w * h > 32 && !c && !isp && rff;      // [1 2 1] edge smoothing
    const int ref_line = ref_idx == 3 ? -4 : -1 - ref_idx;
    const int stride = a.pic.stride[c];
    uint16_t *src = a.pic.p[c] + (size_t)b.y0 * stride + b.x0;
    const int lscale = (b.flags & VVCB200_IB_LMCS_SCALE) ? vpdu_chroma_scale(a, b, lane) : 0;

    if (b.flags & VVCB200_IB_ADD_ONLY) {                                  // the prediction is there already
        if (b.flags & VVCB200_IB_RESIDUAL) {
            const int16_t *res = a.res[c] + (size_t)b.y0 * a.res_stride[c] + b.x0;
            for (int idx = lane; idx < w * h; idx += 32) {
                const int x = idx & (w - 1), y = idx >> lw;
                int r = res[(size_t)y * a.res_stride[c] + x];
                if (lscale)
                    r = scale_chroma_residual(r, lscale, bd);
                uint16_t *d = src + (size_t)y * stride + x;
                *d = (uint16_t)clip_pixel((int)__ldcg(d) + r, bd);
            }
        }
        return;
    }

    // how much of each edge the mode reads (vvc_intra_template.c:470-520)
    int left_size, ul, ut, refw = 0, refh = 0;
    if (is_mip || mode == 0) {
        left_size = h + 1;
        ul = left_size + smooth;
        ut = w + 1 + smooth;
    } else if (mode == 1) {
        ul = left_size = h;
        ut = w;
    } else if (mode == 50) {
        ul = left_size = pdpc ? h : 1;
        ut = w;
    } else if (mode == 18) {
        ul = left_size = h;
        ut = pdpc ? w : 1;
    } else {
        if (!isp || c) { refw = w * 2; refh = h * 2; }
        else           { refw = b.cb_w + w; refh = b.cb_h + h; }
        ut = refw;
        ul = left_size = refh;
    }
    const int la = min(ul, (int)b.avail_left), ta = min(ut, (int)b.avail_top);

    // ---- wait for the neighbours this block reads
    {
        const int s = unit_shift(c), ux = b.x0 >> 2, uy = b.y0 >> s;
        const int nl = unit_count(b.y0, la, s), nt = unit_count(b.x0, ta, 2);
        for (;;) {
            bool p = false;
            if (nl && !(skip_wait & 2)) p |= col_pending(a, c, ux - 1, uy, nl, lane);
            if (nt && !(skip_wait & 1)) p |= row_pending(a, c, uy - 1, ux, nt, lane);
            if (b.up_left && !skip_wait) p |= row_pending(a, c, uy - 1, ux - 1, 1, lane);
            if (!__any_sync(kFull, p))
                break;
            __nanosleep(64);
        }
        __threadfence();
    }

    // ---- gather the edges into shared memory
    uint16_t *left = ws.edge[0] + kEdgeOrg, *top = ws.edge[1] + kEdgeOrg;
    for (int i = lane; i < la; i += 32)
        left[i] = __ldcg(src + ref_line + (ptrdiff_t)stride * i);
    for (int i = lane; i < ta; i += 32)
        top[i] = __ldcg(src + i + (ptrdiff_t)stride * ref_line);
    if (lane < -ref_line) {                                               // the corner run, indices -1 .. ref_line
        const int i = -1 - lane;
        int vl, vt;
        if (b.up_left) {
            vl = __ldcg(src + ref_line + (ptrdiff_t)stride * i);
            vt = __ldcg(src + i + (ptrdiff_t)stride * ref_line);
        } else if (la) {
            vl = vt = __ldcg(src + ref_line);
        } else if (ta) {
            vl = vt = __ldcg(src + (ptrdiff_t)stride * ref_line);
        } else {
            vl = vt = 1 << (bd - 1);
        }
        left[i] = vl;
        top[i] = vt;
    }
    __syncwarp();
    {                                                                     // substitution of what is not there
        const int vt = top[ta - 1], vl = left[la - 1];
        __syncwarp();
        for (int i = ta + lane; i < ut; i += 32)
            top[i] = vt;
        for (int i = la + lane; i < ul; i += 32)
            left[i] = vl;
    }
    __syncwarp();
    if (smooth) {
        uint16_t *fl = ws.edge[2] + kEdgeOrg, *ft = ws.edge[3] + kEdgeOrg;
        const int keep_last = left_size == ul;
        if (lane == 0)
            fl[-1] = ft[-1] = (left[0] + 2 * left[-1] + top[0] + 2) >> 2;
        for (int i = lane; i < ul - keep_last; i += 32)
            fl[i] = (left[i - 1] + 2 * left[i] + left[i + 1] + 2) >> 2;
        for (int i = lane; i < ut - keep_last; i += 32)
            ft[i] = (top[i - 1] + 2 * top[i] + top[i + 1] + 2) >> 2;
        if (keep_last && lane == 0) {
            ft[ut - 1] = top[ut - 1];
            fl[ul - 1] = left[ul - 1];
        }
        left = fl;
        top = ft;
        __syncwarp();
    }
    int interp = 0;                                                       // 1: the smoothing interpolation filter
    const bool angular = !is_mip && mode != 0 && mode != 1;
    int angle = 0;
    if (angular) {
        if (!(rff || ref_idx || isp)) {
            const int d50 = abs(mode - 50), d18 = abs(mode - 18);
            const int thres = (int)((0x020e18ull >> (8 * (((lw + lh) >> 1) - 2))) & 0xff);   // {24, 14, 2, 0, 0}
            interp = min(d50, d18) > thres;
        }
        if (mode != 50 && mode != 18) {
            angle = VVC_INTRA_ANGLE[mode + 14];
            const int inv = VVC_INTRA_INV_ANGLE[mode + 14];
            uint16_t *mainr = mode >= 34 ? top : left, *side = mode >= 34 ? left : top;
            const int nm = mode >= 34 ? w : h, ns = mode >= 34 ? h : w, refn = mode >= 34 ? refw : refh;
            if (angle < 0) {                                              // project the side edge onto the main one
                uint16_t *p = mainr - (ref_idx + 1);
                for (int x = -ns + lane; x < 0; x += 32)
                    p[x] = side[-1 - ref_idx + min((x * inv + 256) >> 9, ns)];
            } else {
                const int last = refn + max(nm / ns, 1) * ref_idx + 1, v = mainr[refn - 1];
                for (int i = refn + lane; i <= last; i += 32)
                    mainr[i] = v;
            }
            __syncwarp();
        }
    }

    Store st;
    st.dst = src; st.dstride = stride; st.bd = bd;
    st.ciip_w = (b.flags & VVCB200_IB_CIIP) ? b.ciip_weight : 0;
    st.res = (b.flags & VVCB200_IB_RESIDUAL) ? a.res[c] + (size_t)b.y0 * a.res_stride[c] + b.x0 : nullptr;
    st.rstride = a.res_stride[c];
    st.lscale = lscale;
    const int n = w * h;

    if (is_mip) {                                                         // vvc_intra_template.c:707-836
        const int size_id = (w == 4 && h == 4) ? 0 : ((w == 4 || h == 4) || (w == 8 && h == 8)) ? 1 : 2;
        const int bsz = size_id ? 4 : 2, psz = size_id == 2 ? 8 : 4, lp = size_id == 2 ? 3 : 2;
        const int in_size = 2 * bsz - (size_id == 2);
        const int nout = psz * psz;
        const bool tr = b.flags & VVCB200_IB_MIP_TRANSPOSED;
        const uint8_t *m = VVC_MIP_MAT + VVC_MIP_OFF[size_id] + b.mip_mode * nout * in_size;
        if (lane < 2 * bsz) {                                             // boundary down-sampling
            const int pass = lane >= bsz, i = lane - pass * bsz;
            const uint16_t *ref = pass ? left : top;
            const int dwn = (pass ? h : w) / bsz;
            int r = 0;
            for (int j = 0; j < dwn; j++)
                r += ref[i * dwn + j];
            r = dwn == 1 ? r : (r + (dwn >> 1)) >> ilog2(dwn);
            ws.red[(pass != tr ? bsz : 0) + i] = r;                       // transposed: left first
        }
        __syncwarp();
        const int temp0 = ws.red[0];
        int in[8], ow;
        {
            int off = 1;
            if (size_id != 2) { off = 0; ow = (1 << (bd - 1)) - temp0; }
            else ow = ws.red[1] - temp0;
            in[0] = ow;
#pragma unroll
            for (int i = 1; i < 8; i++) {
                in[i] = i < in_size ? ws.red[i + off] - temp0 : 0;
                ow += in[i];
            }
            ow = 32 - 32 * ow;
        }
        for (int o = lane; o < nout; o += 32) {
            const uint8_t *mo = m + o * in_size;
            int p = 0;
#pragma unroll
            for (int i = 0; i < 8; i++)
                if (i < in_size)
                    p += in[i] * mo[i];
            p = clip_pixel(((p + ow) >> 6) + temp0, bd);
            const int oy = o >> lp, ox = o & (psz - 1);
            ws.mip[tr ? ox * psz + oy : o] = p;                           // stored at (row, column) of the reduced block
        }
        __syncwarp();
        const int uh_ = w >> lp, uv = h >> lp, luh = ilog2(uh_), luv = ilog2(uv);
        // H(j, x): row j of the reduced block stretched horizontally to the full width
        auto hval = [&](int j, int x) {
            const int jx = x >> luh, k = (x & (uh_ - 1)) + 1;
            const int after = ws.mip[j * psz + jx];
            if (k == uh_)
                return after;
            const int before = jx ? ws.mip[j * psz + jx - 1] : left[j * uv + uv - 1];
            return ((uh_ - k) * before + k * after + (uh_ >> 1)) >> luh;
        };
        for (int idx = lane; idx < n; idx += 32) {
            const int x = idx & (w - 1), y = idx >> lw;
            const int jy = y >> luv, k = (y & (uv - 1)) + 1;
            const int after = hval(jy, x);
            int v = after;
            if (k != uv) {
                const int before = jy ? hval(jy - 1, x) : top[x];
                v = ((uv - k) * before + k * after + (uv >> 1)) >> luv;
            }
            st.put(x, y, v);
        }
        return;
    }

    const int scale = (lw + lh - 2) >> 2;
    if (mode == 0 || mode == 1 || mode == 18 || mode == 50) {
        int dc = 0;
        if (mode == 1) {
            int sum = 0;
            if (w >= h) for (int i = lane; i < w; i += 32) sum += top[i];
            if (w <= h) for (int i = lane; i < h; i += 32) sum += left[i];
            sum = warp_sum(sum);
            const int offset = w == h ? w << 1 : max(w, h);
            dc = (sum + (offset >> 1)) >> ilog2(offset);
        }
        const int tl = top[-1], topw = top[w], lefth = left[h];           // top[w] / left[h] only meaningful for planar
        for (int idx = lane; idx < n; idx += 32) {
            const int x = idx & (w - 1), y = idx >> lw;
            int val;
            if (mode == 0) {
                const int pv = ((h - 1 - y) * top[x] + (y + 1) * lefth) << lw;
                const int ph = ((w - 1 - x) * left[y] + (x + 1) * topw) << lh;
                val = (pv + ph + n) >> (lw + lh + 1);
            } else if (mode == 1) {
                val = dc;
            } else if (mode == 50) {
                val = top[x];
            } else {
                val = left[y];
            }
            if (pdpc) {                                                   // vvc_intra_template.c:630-660
                const int sx = min((x << 1) >> scale, 31), sy = min((y << 1) >> scale, 31);
                int l, t, wl, wt;
                if (mode <= 1) {
                    l = left[y]; t = top[x];
                    wl = 32 >> sx; wt = 32 >> sy;
                } else {
                    l = left[y] - tl + val; t = top[x] - tl + val;
                    wl = mode == 50 ? 32 >> sx : 0;
                    wt = mode == 18 ? 32 >> sy : 0;
                }
                val = clip_pixel(val + ((wl * (l - val) + wt * (t - val) + 32) >> 6), bd);
            }
            st.put(x, y, val);
        }
        return;
    }

    // angular (vvc_intra_template.c:838-1001): `along` runs with the prediction direction's axis, `cross` across it
    const bool vertical = mode >= 34;
    const int inv = pdpc ? VVC_INTRA_INV_ANGLE[mode + 14] : 0, nscale = pdpc ? nscale_of(lw, lh, mode) : 0;
    const uint16_t *main_ref = (vertical ? top : left) - (1 + ref_idx), *side = vertical ? left : top;
    const int8_t *ftab = VVC_INTRA_LUMA_FILTER + interp * 128;
    for (int idx = lane; idx < n; idx += 32) {
        const int x = idx & (w - 1), y = idx >> lw;
        const int along = vertical ? y : x, cross = vertical ? x : y;
        const int pos = (1 + ref_idx + along) * angle;
        const int fact = pos & 31;
        const uint16_t *p = main_ref + cross + (pos >> 5) + ref_idx;
        int pred;
        if (!fact && (c || !interp)) {
            pred = p[1];
        } else if (!c) {
            const int8_t *f = ftab + fact * 4;
            pred = clip_pixel((p[0] * f[0] + p[1] * f[1] + p[2] * f[2] + p[3] * f[3] + 32) >> 6, bd);
        } else {
            pred = ((32 - fact) * p[1] + fact * p[2] + 16) >> 5;
        }
        if (pdpc && cross < (3 << nscale)) {
            const int s = side[along + ((256 + (cross + 1) * inv) >> 9)];
            const int wgt = 32 >> ((cross << 1) >> nscale);
            pred = clip_pixel(pred + (((s - pred) * wgt + 32) >> 6), bd);
        }
        st.put(x, y, pred);
    }
}

// ---- CCLM (vvc_intra_template.c:29-375): one record predicts Cb and Cr --------------------------------------------
__device__ void cclm_record(const IntraArgs &a, const VVCB200IntraBlock &b, int lane)
{
    const int bd = a.bd, w = b.w, h = b.h, x = b.x0, y = b.y0, x0 = x * 2, y0 = y * 2, mode = b.mode;
    const bool avail_t = b.avail_top > 0, avail_l = b.avail_left > 0;
    const int ls = a.pic.stride[0];
    const uint16_t *luma = a.pic.p[0] + (size_t)y0 * ls + x0;
    const int cs[2] = { a.pic.stride[1], a.pic.stride[2] };
    uint16_t *dst[2] = { a.pic.p[1] + (size_t)y * cs[0] + x, a.pic.p[2] + (size_t)y * cs[1] + x };
    const int mn = min(w, h);
    int num[2];
    if (mode == 81) {
        num[0] = avail_t ? w : 0;
        num[1] = avail_l ? h : 0;
    } else {
        num[0] = (avail_t && mode == 83) ? min(w + mn, (int)b.avail_top) : 0;
        num[1] = (avail_l && mode == 82) ? min(h + mn, (int)b.avail_left) : 0;
    }

    // ---- wait: own luma block, luma neighbours under the down-sampling footprint, chroma neighbours of both planes
    {
        const int lux = x0 >> 2, luy = y0 >> 2, cux = x >> 2, cuy = y >> 1;
        const int lnt = avail_t ? unit_count(x0, 2 * max(w, num[0]), 2) : 0;
        const int lnl = avail_l ? unit_count(y0, 2 * max(h, num[1]), 2) : 0;
        const int cnt_ = unit_count(x, num[0], 2), cnl = unit_count(y, num[1], 1);
        const int lnx = unit_count(x0, 2 * w, 2), lny = unit_count(y0, 2 * h, 2);
        for (;;) {
            bool p = row_pending(a, 0, luy, lux, lnx, lane) | row_pending(a, 0, luy + lny - 1, lux, lnx, lane);
            if (lnt) p |= row_pending(a, 0, luy - 1, lux, lnt, lane);
            if (lnl) p |= col_pending(a, 0, lux - 1, luy, lnl, lane);
            if (avail_t && avail_l) p |= row_pending(a, 0, luy - 1, lux - 1, 1, lane);
            for (int c = 1; c <= 2; c++) {
                if (cnt_) p |= row_pending(a, c, cuy - 1, cux, cnt_, lane);
                if (cnl) p |= col_pending(a, c, cux - 1, cuy, cnl, lane);
            }
            if (!__any_sync(kFull, p))
                break;
            __nanosleep(64);
        }
        __threadfence();
    }

    const int lscale = (b.flags & VVCB200_IB_LMCS_SCALE) ? vpdu_chroma_scale(a, b, lane) : 0;
    const int16_t *res[2] = { nullptr, nullptr };
    if (b.flags & VVCB200_IB_RESIDUAL)
        for (int c = 0; c < 2; c++)
            if (b.res_planes & (1 << c))
                res[c] = a.res[c + 1] + (size_t)y * a.res_stride[c + 1] + x;
    const int n = w * h, lw = ilog2(w);
    auto emit = [&](int c, int i, int j, int v) {
        if (res[c]) {
            int r = res[c][(size_t)j * a.res_stride[c + 1] + i];
            if (lscale)
                r = scale_chroma_residual(r, lscale, bd);
            v = clip_pixel(v + r, bd);
        }
        dst[c][(size_t)j * cs[c] + i] = (uint16_t)v;
    };
    if (!avail_t && !avail_l) {
        for (int idx = lane; idx < n; idx += 32)
            for (int c = 0; c < 2; c++)
                emit(c, idx & (w - 1), idx >> lw, 1 << (bd - 1));
        return;
    }
#define L(dx, dy) ((int)__ldcg(luma + (dx) + (ptrdiff_t)ls * (dy)))
    int ka[2] = { 0, 0 }, kb[2] = { 1 << (bd - 1), 1 << (bd - 1) }, kk[2] = { 0, 0 };
    if (num[0] || num[1]) {
        const int is4 = !avail_t || !avail_l || mode != 81;
        int cnt[2], start[2], step[2];
        for (int i = 0; i < 2; i++) {
            start[i] = num[i] >> (2 + is4);
            step[i] = max(num[i] >> (1 + is4), 1);
            cnt[i] = min(num[i], (1 + is4) << 1);
        }
        int v0 = 0, v1 = 0, v2 = 0;                                       // lane i gathers neighbour sample i
        if (lane < cnt[0]) {
            const int pos = start[0] + lane * step[0], lx = pos * 2;
            const int xl = (lx || avail_l) ? lx - 1 : lx;
            if (y0 & ((1 << a.ctb_log2) - 1)) {
                if (a.collocated)
                    v0 = (L(lx, -3) + L(xl, -2) + 4 * L(lx, -2) + L(lx + 1, -2) + L(lx, -1) + 4) >> 3;
                else
                    v0 = (L(xl, -2) + L(xl, -1) + 2 * (L(lx, -2) + L(lx, -1)) + L(lx + 1, -2) + L(lx + 1, -1) + 4) >> 3;
            } else {
                v0 = (L(xl, -1) + 2 * L(lx, -1) + L(lx + 1, -1) + 2) >> 2; // one luma line above a CTU boundary
            }
            v1 = __ldcg(dst[0] + pos - cs[0]);
            v2 = __ldcg(dst[1] + pos - cs[1]);
        } else if (lane < cnt[0] + cnt[1]) {
            const int pos = start[1] + (lane - cnt[0]) * step[1], ly = pos * 2;
            if (a.collocated) {
                const int yt = (ly || avail_t) ? ly - 1 : ly;
                v0 = (L(-3, ly) + L(-2, yt) + 4 * L(-2, ly) + L(-1, ly) + L(-2, ly + 1) + 4) >> 3;
            } else {
                v0 = (L(-3, ly) + L(-3, ly + 1) + 2 * L(-2, ly) + 2 * L(-2, ly + 1) + L(-1, ly) + L(-1, ly + 1) + 4) >> 3;
            }
            v1 = __ldcg(dst[0] - 1 + (ptrdiff_t)pos * cs[0]);
            v2 = __ldcg(dst[1] - 1 + (ptrdiff_t)pos * cs[1]);
        }
        int sel[3][4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            sel[0][i] = __shfl_sync(kFull, v0, i);
            sel[1][i] = __shfl_sync(kFull, v1, i);
            sel[2][i] = __shfl_sync(kFull, v2, i);
        }
        if (cnt[0] + cnt[1] == 2)
#pragma unroll
            for (int c = 0; c < 3; c++) {
                sel[c][3] = sel[c][0]; sel[c][2] = sel[c][1]; sel[c][0] = sel[c][1]; sel[c][1] = sel[c][3];
            }
        // the two smallest and the two largest luma samples, as index pairs
        int mi0 = 0, mi1 = 2, ma0 = 1, ma1 = 3, t;
#define SEL0(i) ((i) == 0 ? sel[0][0] : (i) == 1 ? sel[0][1] : (i) == 2 ? sel[0][2] : sel[0][3])
#define SELC(c, i) ((i) == 0 ? sel[c][0] : (i) == 1 ? sel[c][1] : (i) == 2 ? sel[c][2] : sel[c][3])
#define SWAPI(p, q) (t = p, p = q, q = t)
        if (SEL0(mi0) > SEL0(mi1)) SWAPI(mi0, mi1);
        if (SEL0(ma0) > SEL0(ma1)) SWAPI(ma0, ma1);
        if (SEL0(mi0) > SEL0(ma1)) { SWAPI(mi0, ma0); SWAPI(mi1, ma1); }
        if (SEL0(mi1) > SEL0(ma0)) SWAPI(mi1, ma0);
        int mx[3], mnv[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            mx[c] = (SELC(c, ma0) + SELC(c, ma1) + 1) >> 1;
            mnv[c] = (SELC(c, mi0) + SELC(c, mi1) + 1) >> 1;
        }
#undef SEL0
#undef SELC
#undef SWAPI
        const int diff = mx[0] - mnv[0];
        if (!diff) {
            for (int i = 0; i < 2; i++) { ka[i] = kk[i] = 0; kb[i] = mnv[i + 1]; }
        } else {
            // div_sig_table of the reference packed 4 bits per entry: {0,7,6,5,5,4,4,3,3,2,2,1,1,1,1,0}
            const unsigned long long div_sig = 0x0111122334455670ull;
            for (int i = 0; i < 2; i++) {
                const int diffc = mx[i + 1] - mnv[i + 1];
                int xx = ilog2(diff);
                const int norm = ((diff << 4) >> xx) & 15;
                xx += norm ? 1 : 0;
                const int yy = diffc ? ilog2(abs(diffc)) + 1 : 0;
                const int v = (int)((div_sig >> (4 * norm)) & 15) | 8;
                int av = (diffc * v + ((1 << yy) >> 1)) >> yy;
                const int sh = 3 + xx - yy;
                if (sh < 1)
                    av = (av < 0 ? -1 : av > 0) * 15;
                ka[i] = av;
                kk[i] = max(sh, 1);
                kb[i] = mnv[i + 1] - ((av * mnv[0]) >> kk[i]);
            }
        }
    }
    for (int idx = lane; idx < n; idx += 32) {                            // down-sampled luma -> both chroma planes
        const int i = idx & (w - 1), j = idx >> lw;
        const int lx = 2 * i, ly = 2 * j;
        const int xl = (i || avail_l) ? lx - 1 : lx;
        int ds;
        if (a.collocated) {
            const int yt = (j || avail_t) ? ly - 1 : ly;
            ds = (L(xl, ly) + L(lx, yt) + 4 * L(lx, ly) + L(lx + 1, ly) + L(lx, ly + 1) + 4) >> 3;
        } else {
            ds = (L(xl, ly) + L(xl, ly + 1) + 2 * L(lx, ly) + 2 * L(lx, ly + 1) + L(lx + 1, ly) + L(lx + 1, ly + 1) + 4) >> 3;
        }
        for (int c = 0; c < 2; c++)
            emit(c, i, j, clip_pixel(((ds * ka[c]) >> kk[c]) + kb[c], bd));
    }
#undef L
}

__global__ void __launch_bounds__(kWarps * 32) intra_kernel(IntraArgs a)
{
    __shared__ WarpScratch s_ws[kWarps];
    const int lane = threadIdx.x & 31;
    WarpScratch &ws = s_ws[threadIdx.x >> 5];
    for (;;) {
        int t = 0;
        if (lane == 0)
            t = atomicAdd(a.ticket, 1);
        t = __shfl_sync(kFull, t, 0);
        if (t >= a.n)
            break;
        const VVCB200IntraBlock b = load_record(a.blocks + t);
        if (b.mode >= 81) {
            cclm_record(a, b, lane);
        } else if ((b.flags & VVCB200_IB_ISP) && !b.c_idx) {
            // ISP: the sub-partitions of the coding block one after the other, each predicted from its reconstructed
            // predecessor (vvc_intra.c:217-226, 244-252); w x h is the prediction unit
            const bool hor = b.w == b.cb_w;
            const int n_parts = hor ? b.cb_h / b.h : b.cb_w / b.w;
            for (int k = 0; k < n_parts; k++) {
                VVCB200IntraBlock s = b;
                if (hor) {
                    s.y0 += k * b.h;
                    s.avail_left = max((int)b.avail_left - k * b.h, 0);
                    if (k) { s.avail_top = b.w; s.up_left = b.avail_left > 0; }
                } else {
                    s.x0 += k * b.w;
                    s.avail_top = max((int)b.avail_top - k * b.w, 0);
                    if (k) { s.avail_left = b.h; s.up_left = b.avail_top > 0; }
                }
                predict_record(a, s, ws, lane, k ? (hor ? 1 : 2) : 0);
                __threadfence();                                          // the next sub-partition reads these samples
                __syncwarp();
            }
        } else {
            predict_record(a, b, ws, lane);
        }
        __threadfence();                                                  // samples before marks
        __syncwarp();
        int ew, eh;
        record_extent(b, ew, eh);
        set_marks(a, b.c_idx, b.x0, b.y0, ew, eh, 0, lane);
        if (b.mode >= 81)
            set_marks(a, 2, b.x0, b.y0, ew, eh, 0, lane);
        __syncwarp();
    }
}

struct ScratchLayout {
    size_t off[3], total;
    int pitch[3];
};

ScratchLayout scratch_layout(int width, int height)
{
    ScratchLayout l;
    size_t o = 256;                                                       // the ticket counter's own sector
    for (int c = 0; c < 3; c++) {
        const int w = c ? (width + 1) >> 1 : width, h = c ? (height + 1) >> 1 : height;
        l.pitch[c] = (((w + 3) >> 2) + 15) & ~15;
        const int rows = c ? (h + 1) >> 1 : (h + 3) >> 2;
        l.off[c] = o;
        o += (size_t)l.pitch[c] * rows;
        o = (o + 255) & ~(size_t)255;
    }
    l.total = o;
    return l;
}

}  // namespace
}  // namespace vvcb200

using namespace vvcb200;

extern "C" size_t vvcb200_intra_scratch_bytes(int width, int height)
{
    if (width <= 0 || height <= 0)
        return 0;
    return scratch_layout(width, height).total;
}

extern "C" int vvcb200_intra_schedule(const VVCB200IntraBlock *in, const uint32_t *ctu_first, int ctus_w, int ctus_h,
                                      VVCB200IntraBlock *out)
{
    if (!in || !ctu_first || !out || ctus_w <= 0 || ctus_h <= 0) {
        set_error("vvcb200_intra_schedule: bad arguments");
        return VVCB200_ERR_ARG;
    }
    size_t o = 0;
    for (int d = 0; d <= ctus_w - 1 + 2 * (ctus_h - 1); d++) {
        // CTUs of this anti-diagonal: (d - 2y, y); round-robin over them so that a window of consecutive tickets
        // spans every CTU that can run now
        const int y_lo = d - (ctus_w - 1) > 0 ? (d - (ctus_w - 1) + 1) / 2 : 0;
        const int y_hi = d / 2 < ctus_h - 1 ? d / 2 : ctus_h - 1;
        uint32_t longest = 0;
        for (int y = y_lo; y <= y_hi; y++) {
            const int rs = y * ctus_w + d - 2 * y;
            const uint32_t len = ctu_first[rs + 1] - ctu_first[rs];
            longest = len > longest ? len : longest;
        }
        for (uint32_t k = 0; k < longest; k++)
            for (int y = y_lo; y <= y_hi; y++) {
                const int rs = y * ctus_w + d - 2 * y;
                if (ctu_first[rs] + k < ctu_first[rs + 1])
                    out[o++] = in[ctu_first[rs] + k];
            }
    }
    return 0;
}

extern "C" int vvcb200_intra_picture(const VVCB200Planes *pic, const VVCB200Planes *d_residual,
                                     const VVCB200IntraBlock *d_blocks, int n, const VVCB200IntraParams *params,
                                     void *d_scratch, void *stream)
{
    if (!pic || !params || n < 0 || (n && (!d_blocks || !d_scratch))) {
        set_error("vvcb200_intra_picture: bad arguments");
        return VVCB200_ERR_ARG;
    }
    if (params->bit_depth < 8 || params->bit_depth > 12 || params->ctb_log2_size < 5 || params->ctb_log2_size > 7) {
        set_error("vvcb200_intra_picture: unsupported bit depth %d / CTB size 2^%d", params->bit_depth, params->ctb_log2_size);
        return VVCB200_ERR_ARG;
    }
    if (!n)
        return 0;
    IntraArgs a;
    a.pic = make_view(pic);
    for (int c = 0; c < 3; c++) {
        a.res[c] = d_residual ? (const int16_t *)d_residual->d_plane[c] : nullptr;
        a.res_stride[c] = d_residual ? d_residual->stride[c] / 2 : 0;
    }
    a.blocks = d_blocks;
    a.n = n;
    a.bd = params->bit_depth;
    a.ctb_log2 = params->ctb_log2_size;
    a.collocated = params->chroma_vertical_collocated;
    for (int i = 0; i < 17; i++)
        a.lmcs_pivot[i] = params->lmcs_pivot[i];
    for (int i = 0; i < 16; i++)
        a.lmcs_coeff[i] = params->lmcs_chroma_scale_coeff[i];
    a.lmcs_min_bin = params->lmcs_min_bin_idx;
    a.lmcs_max_bin = params->lmcs_max_bin_idx < 15 ? params->lmcs_max_bin_idx : 15;
    const ScratchLayout l = scratch_layout(pic->width[0], pic->height[0]);
    a.ticket = (int *)d_scratch;
    for (int c = 0; c < 3; c++) {
        a.marks[c] = (uint8_t *)d_scratch + l.off[c];
        a.mpitch[c] = l.pitch[c];
    }
    cudaStream_t s = (cudaStream_t)stream;
    intra_mark_kernel<<<(n + 3) / 4, 128, 0, s>>>(a);
    VVC_LAUNCH_CHECK();
    const int ctas = min((n + kWarps - 1) / kWarps, sm_count() * 4);
    intra_kernel<<<ctas, kWarps * 32, 0, s>>>(a);
    VVC_LAUNCH_CHECK();
    return 0;
}

// ---- the same reconstruction with the scratch living in the module image instead of caller memory ---------------
// Some integrations cannot hand us scratch memory or let us create any: a decoder thread sitting between two pictures
// must not call the CUDA allocator, because every allocation and release serialises against all work in flight on
// the device and would stall the other frame contexts' streams; and a player that has given the whole card to its
// own pool at start-up can see a late driver allocation fail outright.  For those callers the marks and the ticket
// counter live in a fixed array that is part of the loaded module, sized for pictures up to 1920x1088.
// Nothing here allocates.  The address of the array is looked up once per device (with that device current) and
// remembered; that first call should happen before the application starts capturing a stream.  The array is shared
// by every caller on the device: two of these calls overlapping in time (different streams or threads) race on it.
namespace vvcb200 {
namespace {
constexpr size_t kStaticScratchBytes = 256 * 1024;                        // enough for 1920x1088 (261376 bytes)
constexpr int kMaxDevices = 64;
static __device__ __align__(256) unsigned char g_intra_scratch[kStaticScratchBytes];
void *g_intra_scratch_ptr[kMaxDevices];                                   // per device, filled on first use
}  // namespace
}  // namespace vvcb200

#include <mutex>

extern "C" int vvcb200_intra_picture_static(const VVCB200Planes *pic, const VVCB200Planes *d_residual,
                                            const VVCB200IntraBlock *d_blocks, int n,
                                            const VVCB200IntraParams *params, void *stream)
{
    static std::mutex lock;
    if (!pic || !params || n < 0 || (n && !d_blocks)) {
        set_error("vvcb200_intra_picture_static: bad arguments");
        return VVCB200_ERR_ARG;
    }
    if (!n)
        return 0;
    const size_t need = vvcb200_intra_scratch_bytes(pic->width[0], pic->height[0]);
    if (!need || need > kStaticScratchBytes) {
        set_error("vvcb200_intra_picture_static: a %dx%d picture needs %zu bytes of marks, the built-in scratch holds %zu "
                  "(pictures up to 1920x1088); use vvcb200_intra_picture with caller memory",
                  pic->width[0], pic->height[0], need, kStaticScratchBytes);
        return VVCB200_ERR_ARG;
    }
    int dev = -1;
    VVC_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("vvcb200_intra_picture_static: device index %d out of range", dev);
        return VVCB200_ERR_ARG;
    }
    void *ptr;
    {
        std::lock_guard<std::mutex> g(lock);
        if (!g_intra_scratch_ptr[dev]) {
            void *p = nullptr;
            VVC_CUDA(cudaGetSymbolAddress((void **)&p, g_intra_scratch));
            g_intra_scratch_ptr[dev] = p;
        }
        ptr = g_intra_scratch_ptr[dev];
    }
    // the marks have to read zero when the mark kernel starts; an earlier call that failed half-way may have left some
    VVC_CUDA(cudaMemsetAsync(ptr, 0, need, (cudaStream_t)stream));
    return vvcb200_intra_picture(pic, d_residual, d_blocks, n, params, ptr, stream);
}
