// intra.cu — stage RECON, intra part: intra prediction (+ CIIP blend, + residual) of a picture's intra blocks.
//
// Replaces, for a whole picture, the per-CTU loop of ff_vvc_reconstruct (vvc_intra.c:498-536) together with the
// intra DSP of vvc_intra_template.c (edge preparation 450-598, planar/DC/angular/PDPC 600-705 + 838-1001,
// MIP 707-836, CCLM 29-375).
//
// Intra prediction reads reconstructed neighbours, so the work is a dependency graph, not a batch.  The reference
// walks it as a CTU wavefront (vvc_thread.c:156-184); here the graph is resolved per block:
//   * a pre-pass fills the samples every intra record will write with a sentinel (0xffff, not a sample value at any
//     supported bit depth); warps draw records from a ticket counter in array order — a topological order, see
//     vvcb200_intra_schedule — and re-read the neighbour samples they need until none of them is the sentinel: the
//     read that detects readiness is the read that fetches the data, and the writer needs no fence and no flag;
//   * blocks that are overwritten in place (CIIP blend, add-only chroma of inter units) cannot carry a sentinel: they
//     mark their 4x4 (chroma: 4x2) units "pending" in the scratch maps instead, publish with a fence and clear the marks.
// A ticket is only ever held by a resident warp and waits only on lower tickets, so the scheme cannot deadlock
// whatever the grid size; parallelism is what the picture's block graph offers (blocks of one anti-diagonal of
// CTUs and independent blocks inside a CTU run concurrently), which is far more than one warp per CTU row.
#include "common.cuh"

#define VVC_TABLE_QUAL static __device__ const
#include "vvc_tables.h"

namespace vvcb200 {
namespace {

constexpr int kWarps   = 4;       // warps per CTA
#ifndef VVC_INTRA_CTAS_PER_SM
#define VVC_INTRA_CTAS_PER_SM 4   // 16 warps per SM: what 128 registers per thread and ~50 KB of shared memory per CTA allow
#endif
constexpr int kCtasPerSm = VVC_INTRA_CTAS_PER_SM;
constexpr int kSlots   = 1;       // records a warp holds at a time (1: it simply waits for its record)
constexpr int kEdgeLen = 256;     // samples per edge array: indices -72 .. 183 are addressable
constexpr int kEdgeOrg = 72;
constexpr unsigned kFull = 0xffffffffu;
#ifndef VVC_INTRA_BACKOFF_MAX
#define VVC_INTRA_BACKOFF_MAX 256
#endif
constexpr unsigned kBackoffMax = VVC_INTRA_BACKOFF_MAX;                   // ns between two looks at neighbours that are not there yet
constexpr unsigned kSentinel = 0xffffu;                                   // "not reconstructed yet"
#ifdef VVC_INTRA_PROFILE
__device__ unsigned long long g_prof[8];      // cycles: 0 ticket+record, 1 record total, 2 publish; 3 records; 4 wait (inside 1)
#define PROF_T() clock64()
#define PROF_ADD(i, v) do { if (lane == 0) atomicAdd(&g_prof[i], (unsigned long long)(v)); } while (0)
#else
#define PROF_T() 0ll
#define PROF_ADD(i, v) do { } while (0)
#endif

// Staging area of a warp: everything a record reads that does not depend on its neighbours (its residual, the inter
// prediction of a CIIP block) is fetched with cp.async BEFORE the warp starts waiting, so that the latency of those
// reads never sits on the picture's dependency chain; CCLM stages the co-located luma window here after the wait.
// The per-warp buffers hold records of up to kSmallSamples samples, which is nearly all of them: the kernel lives on
// the number of resident warps (each spends most of its time in L2 round trips), and buffers sized for the largest
// block would allow 8 warps per SM instead of 12 .. 16.  A larger record borrows the CTA's one arena, and only once
// everything it waits for is there, so that an arena holder never blocks on another record.
#ifndef VVC_INTRA_SMALL
#define VVC_INTRA_SMALL 1024
#define VVC_INTRA_SMALL_EXTRA 2624
#endif
constexpr int kSmallSamples = VVC_INTRA_SMALL;
constexpr int kSmallRes   = kSmallSamples * 2;                            // residual block, or Cb + Cr residual of a CCLM record
constexpr int kSmallExtra = VVC_INTRA_SMALL_EXTRA;                        // CIIP inter prediction; CCLM luma window (16x16, 8x32 chroma)
constexpr int kArenaRes   = 64 * 64 * 2;
constexpr int kArenaExtra = 65 * 68 * 2 + 16;
struct __align__(16) WarpScratch {
    uint8_t  stage_res[kSmallRes];
    uint8_t  stage_extra[kSmallExtra];
    VVCB200IntraBlock slot[kSlots]; // records this warp holds and looks at in turn
    uint16_t edge[4][kEdgeLen];   // left, top, filtered left, filtered top
    int      red[16];             // MIP: reduced boundary
    uint16_t mip[64];             // MIP: reduced prediction, row-major at the reduced size
};
struct __align__(16) CtaArena {
    uint8_t stage_res[kArenaRes];
    uint8_t stage_extra[kArenaExtra];
    int     lock;
};

// Small tables every record consults, copied once per CTA: after each acquire fence the L1 is invalidated, so a table
// left in global memory would cost an L2 round trip per lookup on the dependency chain.
struct CtaTables {
    int16_t angle[95], inv_angle[95];
    int8_t  luma_filter[256];
    uint16_t mip_off[4];
    uint8_t mip_mat[4736];
};

struct IntraArgs {
    PlanesView pic;
    const int16_t *res[3];
    int res_stride[3];            // in samples
    uint16_t *shadow[3];          // where the prediction of in-place blocks is kept while the picture holds the sentinel, or null
    int shadow_stride[3];         // in samples
    const VVCB200IntraBlock *blocks;
    int n;
    int bd, ctb_log2, collocated;
    uint8_t *marks[3];            // pending marks per plane (in-place records only)
    int mpitch[3];
    int *ticket;                  // next record of this launch
    int *inplace_flag;            // set by the pre-pass when the picture has in-place records that go by marks
    int skip_first, skip_n;       // pre-pass: records left alone (the add-only phase of a phased picture)
    const int *phase_info;        // device: n_luma, n_add_only, n_chroma as vvcb200_intra_schedule_device left them, or null
    int phase;                    // with phase_info: which records this launch is for (kPhase*)
    unsigned *vpdu_scale;         // phased pictures: chroma residual scale per VPDU, 0 = not derived yet, else 0x10000 | scale
    int vpdu_pitch;               // VPDUs per row of that table
    uint16_t lmcs_pivot[17], lmcs_coeff[16];
    int lmcs_min_bin, lmcs_max_bin;
};

// Launches whose record range is only known on the device (the schedule was made there) read it here.
enum { kPhaseLuma = 0, kPhaseAddOnly = 1, kPhaseChroma = 2, kPhaseAllChroma = 3, kPhaseMark = 4 };
struct RecordRange {
    const VVCB200IntraBlock *blocks;
    int n, skip_first, skip_n;
};
// (returns a copy of the few fields instead of patching the argument struct: a kernel that writes to its parameter makes
// the compiler move all of it into local memory)
__device__ __forceinline__ RecordRange resolve_phase(const IntraArgs &a)
{
    RecordRange r = { a.blocks, a.n, a.skip_first, a.skip_n };
    if (!a.phase_info)
        return r;
    const int n0 = a.phase_info[0], n1 = a.phase_info[1], n2 = a.phase_info[2];
    switch (a.phase) {
    case kPhaseLuma:      r.n = n0; break;
    case kPhaseAddOnly:   r.blocks += n0; r.n = n1; break;
    case kPhaseChroma:    r.blocks += n0 + n1; r.n = n2; break;
    case kPhaseAllChroma: r.blocks += n0; r.n = n1 + n2; break;
    default:              r.n = n0 + n1 + n2; r.skip_first = n0; r.skip_n = n1; break;
    }
    return r;
}

__device__ __forceinline__ int ilog2(int v) { return 31 - __clz(v); }
__host__ __device__ __forceinline__ int unit_shift(int c) { return c ? 1 : 2; }     // rows per unit: 4 (luma) / 2 (chroma)

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
// Marks and the samples they guard are exchanged through L2: producers publish with fence.acq_rel.gpu before clearing
// a mark, consumers poll with gpu-scope relaxed loads, fence, then read samples with ld.cg.  (__threadfence() would be
// a sequentially consistent fence, MEMBAR.SC.GPU, several times more expensive on the critical path.)
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// w samples per row (multiple of 4), h rows, global -> shared (compact rows of w), asynchronously.  8-byte pieces: block
// origins are only guaranteed to be multiples of 4 samples (ternary splits).
// One copy of this in the kernel, not one per call site (they made up 30 % of the code, and the kernel is bound by
// instruction fetch): a call costs a few cycles, and the copies are asynchronous anyway.
__device__ __noinline__ void stage_block_async(void *sdst, const void *g, int gstride /* samples */, int w, int h, int lane)
{
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sdst);
    const uint16_t *gp = static_cast<const uint16_t *>(g);
    const int per_row = w >> 2, n = per_row * h, lpr = 31 - __clz(per_row);   // w is a power of two
#pragma unroll 1
    for (int i = lane; i < n; i += 32) {
        const int y = i >> lpr, x = (i & (per_row - 1)) << 2;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sbase + (y * w + x) * 2), "l"(gp + (size_t)y * gstride + x) : "memory");
    }
}
__device__ __forceinline__ void stage_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// The CTA's arena for records too large for a warp's own staging buffers.  Taken only by a warp that has nothing left
// to wait for, so every holder runs to its release.
__device__ __forceinline__ void arena_acquire(CtaArena &ar, int lane)
{
    if (lane == 0)
        while (atomicCAS(&ar.lock, 0, 1))
            __nanosleep(64);
    __syncwarp();
}
__device__ __forceinline__ void arena_release(CtaArena &ar, int lane)
{
    __syncwarp();
    if (lane == 0) {
        __threadfence_block();
        atomicExch(&ar.lock, 0);
    }
}

__device__ __forceinline__ unsigned load_mark(const uint8_t *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u8 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void store_mark(uint8_t *p, unsigned v)
{
    asm volatile("st.relaxed.gpu.global.u8 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// marks: 0 = never marked, 1 = pending, 2 = was pending, now written.  Returns the OR of the marks of the run.
__device__ __forceinline__ unsigned row_marks(const IntraArgs &a, int c, int uy, int ux0, int n, int lane)
{
    unsigned m = 0;
    for (int i = lane; i < n; i += 32)
        m |= load_mark(a.marks[c] + (size_t)uy * a.mpitch[c] + ux0 + i);
    return m;
}

__device__ __forceinline__ unsigned col_marks(const IntraArgs &a, int c, int ux, int uy0, int n, int lane)
{
    unsigned m = 0;
    for (int i = lane; i < n; i += 32)
        m |= load_mark(a.marks[c] + (size_t)(uy0 + i) * a.mpitch[c] + ux);
    return m;
}

// units touched by samples [p0, p0 + n) along an axis with 1 << s samples per unit
__host__ __device__ __forceinline__ int unit_count(int p0, int n, int s) { return n > 0 ? ((p0 + n - 1) >> s) - (p0 >> s) + 1 : 0; }

// lane / lanes: this thread's index in the group of threads that shares the block, and the group's size
__device__ void set_marks(const IntraArgs &a, int c, int x0, int y0, int w, int h, uint8_t v, int lane, int lanes = 32)
{
    const int s = unit_shift(c);
    const int ux = x0 >> 2, uy = y0 >> s, nx = unit_count(x0, w, 2), ny = unit_count(y0, h, s);
    for (int i = lane; i < nx * ny; i += lanes)
        store_mark(a.marks[c] + (size_t)(uy + i / nx) * a.mpitch[c] + ux + i % nx, v);
}

// the samples a record writes: an ISP luma record stands for its whole coding block
__device__ __forceinline__ void record_extent(const VVCB200IntraBlock &b, int &w, int &h)
{
    const bool isp = (b.flags & VVCB200_IB_ISP) && !b.c_idx;
    w = isp ? b.cb_w : b.w;
    h = isp ? b.cb_h : b.h;
}

__device__ __forceinline__ bool in_place(const VVCB200IntraBlock &b) { return b.flags & (VVCB200_IB_CIIP | VVCB200_IB_ADD_ONLY); }

// Threads per record.  (Most records are 4 .. 16 pieces of 8 bytes, but 8 threads per record measured slower than a
// warp per record: the records sharing a warp differ in size and their loops serialise.)
constexpr int kMarkLanes = 32;
__global__ void intra_mark_kernel(const IntraArgs a)
{
    const RecordRange rr = resolve_phase(a);
    const int lane = threadIdx.x & (kMarkLanes - 1);
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) / kMarkLanes;
    if (t >= rr.n || (t >= rr.skip_first && t < rr.skip_first + rr.skip_n))
        return;
    const VVCB200IntraBlock b = load_record(rr.blocks + t);
    int w, h;
    record_extent(b, w, h);
    if ((b.flags & VVCB200_IB_ADD_ONLY) && !(b.flags & VVCB200_IB_RESIDUAL))
        return;                                                           // nothing to add: the samples are final already
    const bool keep = in_place(b);                                        // the block's current samples are an input
    if (keep && !a.shadow[b.c_idx]) {
        set_marks(a, b.c_idx, b.x0, b.y0, w, h, 1, lane, kMarkLanes);
        if (lane == 0)
            *a.inplace_flag = 1;
        return;
    }
    for (int c = b.c_idx; c <= (b.mode >= 81 ? 2 : b.c_idx); c++) {       // sentinel over everything the record will write
        uint16_t *dst = a.pic.p[c] + (size_t)b.y0 * a.pic.stride[c] + b.x0;
        uint16_t *sh = keep ? a.shadow[c] + (size_t)b.y0 * a.shadow_stride[c] + b.x0 : nullptr;
        const int per_row = w >> 2, lpr = 31 - __clz(per_row), total = per_row * h;
        // four pieces per thread and round: the loads of a copied block are in flight together (a 64x64 CIIP block walked
        // piece by piece was the tail of the kernel)
        for (int i0 = lane; i0 < total; i0 += 4 * kMarkLanes) {
            uint2 v[4];
            uint2 *d[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int i = min(i0 + k * kMarkLanes, total - 1);
                const int y = i >> lpr, x = (i & (per_row - 1)) << 2;
                d[k] = reinterpret_cast<uint2 *>(dst + (size_t)y * a.pic.stride[c] + x);
                if (keep)
                    v[k] = *d[k];
            }
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int i = i0 + k * kMarkLanes;
                if (i >= total)
                    break;
                if (keep) {
                    const int y = i >> lpr, x = (i & (per_row - 1)) << 2;
                    *reinterpret_cast<uint2 *>(sh + (size_t)y * a.shadow_stride[c] + x) = v[k];
                }
                *d[k] = make_uint2(0xffffffffu, 0xffffffffu);
            }
        }
    }
}

// ---- mode helpers (vvc_intra.c:538-660) ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ bool ref_filter_mode(int mode)
{
    // modes whose angle hits integer sample positions: the reference smooths the edge instead of interpolating
    switch (mode) {
    case -14: case -12: case -10: case -6: case 0: case 2: case 34: case 66: case 72: case 76: case 78: case 80:
        return true;
    }
    return false;
}

__device__ __forceinline__ bool is_plain_mode(int mode) { return mode == 0 || mode == 1 || mode == 18 || mode == 50; }

__device__ __forceinline__ int nscale_of(const CtaTables &tb, int lw, int lh, int mode)
{
    if (is_plain_mode(mode))
        return (lw + lh - 2) >> 2;
    const int ns = (mode >= 50 ? lh : lw) - ilog2(3 * tb.inv_angle[mode + 14] - 2) + 8;
    return min(ns, 2);
}

__device__ __forceinline__ bool need_pdpc(const CtaTables &tb, int w, int h, int lw, int lh, bool bdpcm, int mode, int ref_idx)
{
    if (w >= 4 && h >= 4 && !ref_idx && !bdpcm) {
        if (is_plain_mode(mode))
            return true;
        if (mode > 18 && mode < 50)
            return false;
        return nscale_of(tb, lw, lh, mode) >= 0;
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

// ---- what a record reads, as runs of mark units: shared by the device (what to wait for) and by the host scheduler
// (which records a record depends on) so that the two can never disagree ------------------------------------------
struct Span {
    int c, horizontal;            // plane; run goes right (1) or down (0)
    int ux, uy, n;                // first unit and length
};
constexpr int kMaxSpans = 12;

__host__ __device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__host__ __device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

struct EdgeExtent {
    int ul, ut;                   // samples of the left / top edge the mode reads (vvc_intra_template.c:470-520)
    int left_size, refw, refh;
    bool smooth;                  // [1 2 1] smoothing of the edges
};

__host__ __device__ inline EdgeExtent edge_extent(const VVCB200IntraBlock &b)
{
    EdgeExtent e;
    const int c = b.c_idx, w = b.w, h = b.h, mode = b.mode;
    const bool is_mip = b.flags & VVCB200_IB_MIP, isp = b.flags & VVCB200_IB_ISP;
    const bool pdpc_plain = w >= 4 && h >= 4 && !b.ref_idx && !(b.flags & VVCB200_IB_BDPCM);
    e.smooth = !b.ref_idx && w * h > 32 && !c && !isp && !is_mip && ref_filter_mode(mode);
    e.refw = e.refh = 0;
    if (is_mip || mode == 0) {
        e.left_size = h + 1;
        e.ul = e.left_size + e.smooth;
        e.ut = w + 1 + e.smooth;
    } else if (mode == 1) {
        e.ul = e.left_size = h;
        e.ut = w;
    } else if (mode == 50) {
        e.ul = e.left_size = pdpc_plain ? h : 1;
        e.ut = w;
    } else if (mode == 18) {
        e.ul = e.left_size = h;
        e.ut = pdpc_plain ? w : 1;
    } else {
        if (!isp || c) { e.refw = w * 2; e.refh = h * 2; }
        else           { e.refw = b.cb_w + w; e.refh = b.cb_h + h; }
        e.ut = e.refw;
        e.ul = e.left_size = e.refh;
    }
    return e;
}

// the luma column / row next to the VPDU that a luma-dependent chroma scale is derived from
__host__ __device__ inline int vpdu_deps(const VVCB200IntraBlock &b, int ctb_log2, int pic_w, int pic_h, Span *out)
{
    int n = 0;
    if (!(b.flags & VVCB200_IB_LMCS_SCALE))
        return 0;
    const int size = imin(1 << ctb_log2, 64);
    const int x = b.vpdu_x * size, y = b.vpdu_y * size;
    const int rows = imin(size, pic_h - y), cols = imin(size, pic_w - x);
    if (b.vpdu_avail & 1) out[n++] = Span{ 0, 0, (x >> 2) - 1, y >> 2, unit_count(y, rows, 2) };
    if (b.vpdu_avail & 2) out[n++] = Span{ 0, 1, x >> 2, (y >> 2) - 1, unit_count(x, cols, 2) };
    return n;
}

// skip_wait: bit 0 = the row above (and the corner), bit 1 = the column to the left (and the corner) were written by
// the same warp (earlier ISP sub-partitions) and carry this record's own marks
__host__ __device__ inline int predict_deps(const VVCB200IntraBlock &b, int skip_wait, Span *out)
{
    if (b.flags & VVCB200_IB_ADD_ONLY)
        return 0;
    const EdgeExtent e = edge_extent(b);
    const int la = imin(e.ul, b.avail_left), ta = imin(e.ut, b.avail_top);
    const int c = b.c_idx, s = unit_shift(c), ux = b.x0 >> 2, uy = b.y0 >> s;
    int n = 0;
    if (la && !(skip_wait & 2)) out[n++] = Span{ c, 0, ux - 1, uy, unit_count(b.y0, la, s) };
    if (ta && !(skip_wait & 1)) out[n++] = Span{ c, 1, ux, uy - 1, unit_count(b.x0, ta, 2) };
    if (b.up_left && !skip_wait) out[n++] = Span{ c, 1, ux - 1, uy - 1, 1 };
    return n;
}

struct CclmCounts { int num[2]; };
__host__ __device__ inline CclmCounts cclm_counts(const VVCB200IntraBlock &b)
{
    CclmCounts k;
    const bool avail_t = b.avail_top > 0, avail_l = b.avail_left > 0;
    const int mn = imin(b.w, b.h);
    if (b.mode == 81) {
        k.num[0] = avail_t ? b.w : 0;
        k.num[1] = avail_l ? b.h : 0;
    } else {
        k.num[0] = (avail_t && b.mode == 83) ? imin(b.w + mn, b.avail_top) : 0;
        k.num[1] = (avail_l && b.mode == 82) ? imin(b.h + mn, b.avail_left) : 0;
    }
    return k;
}

// CCLM: own luma block, luma neighbours under the down-sampling footprint, chroma neighbours of both planes
__host__ __device__ inline int cclm_deps(const VVCB200IntraBlock &b, Span *out)
{
    const CclmCounts k = cclm_counts(b);
    const bool avail_t = b.avail_top > 0, avail_l = b.avail_left > 0;
    const int w = b.w, h = b.h, x = b.x0, y = b.y0, x0 = x * 2, y0 = y * 2;
    const int lux = x0 >> 2, luy = y0 >> 2, cux = x >> 2, cuy = y >> 1;
    const int lnx = unit_count(x0, 2 * w, 2), lny = unit_count(y0, 2 * h, 2);
    int n = 0;
    out[n++] = Span{ 0, 1, lux, luy, lnx };
    out[n++] = Span{ 0, 1, lux, luy + lny - 1, lnx };
    if (avail_t) out[n++] = Span{ 0, 1, lux, luy - 1, unit_count(x0, 2 * imax(w, k.num[0]), 2) };
    if (avail_l) out[n++] = Span{ 0, 0, lux - 1, luy, unit_count(y0, 2 * imax(h, k.num[1]), 2) };
    if (avail_t && avail_l) out[n++] = Span{ 0, 1, lux - 1, luy - 1, 1 };
    for (int c = 1; c <= 2; c++) {
        if (k.num[0]) out[n++] = Span{ c, 1, cux, cuy - 1, unit_count(x, k.num[0], 2) };
        if (k.num[1]) out[n++] = Span{ c, 0, cux - 1, cuy, unit_count(y, k.num[1], 1) };
    }
    return n;
}

// sub-partition k of an ISP record (vvc_intra.c:217-226): geometry and the availability that follows from the
// coding block's; returns the skip_wait bits
// Transform partitions of an ISP coding block (vvc_ctu.c: 4x8 / 8x4 blocks split in two, all others in four) and the
// width of one when the split is vertical (a horizontal split always has h < cb_h, a vertical one h == cb_h; with
// partitions narrower than 4 the prediction unit is 4 wide and spans several of them, vvc_intra.c:217-226).
__host__ __device__ inline int isp_tb_count(const VVCB200IntraBlock &b) { return b.cb_w * b.cb_h == 32 ? 2 : 4; }
__host__ __device__ inline bool isp_vertical(const VVCB200IntraBlock &b) { return b.h == b.cb_h; }

__host__ __device__ inline int isp_part(const VVCB200IntraBlock &b, int k, VVCB200IntraBlock &s)
{
    const bool hor = b.w == b.cb_w;
    s = b;
    // which of the transform partitions inside prediction unit k carry a residual (isp_coded): the others add nothing,
    // whatever the residual planes hold there
    if (b.flags & VVCB200_IB_RESIDUAL) {
        int bits;
        if (isp_vertical(b)) {
            const int ratio = b.w / (b.cb_w / isp_tb_count(b));
            bits = (b.isp_coded >> (k * ratio)) & ((1 << ratio) - 1);
        } else {
            bits = (b.isp_coded >> k) & 1;
        }
        s.isp_coded = (uint8_t)bits;
        if (!bits)
            s.flags &= ~VVCB200_IB_RESIDUAL;
    }
    if (hor) {
        s.y0 += k * b.h;
        s.avail_left = imax((int)b.avail_left - k * b.h, 0);
        if (k) { s.avail_top = b.w; s.up_left = b.avail_left > 0; }
    } else {
        s.x0 += k * b.w;
        s.avail_top = imax((int)b.avail_top - k * b.w, 0);
        if (k) { s.avail_left = b.h; s.up_left = b.avail_top > 0; }
    }
    return k ? (hor ? 1 : 2) : 0;
}
__host__ __device__ inline int isp_parts(const VVCB200IntraBlock &b) { return b.w == b.cb_w ? b.cb_h / b.h : b.cb_w / b.w; }

// Are the units in the runs free of pending marks?  blocking: poll until they are.  The acquire fence is only needed
// when some unit was written by an in-place record (its samples were published before its mark went from 1 to 2).
__device__ bool spans_ready(const IntraArgs &a, const Span *sp, int n, int lane, bool blocking)
{
    if (!n)
        return true;
    unsigned m;
    for (unsigned backoff = 32;; backoff = min(backoff * 2, kBackoffMax)) {
        m = 0;
        for (int i = 0; i < n; i++)
            m |= sp[i].horizontal ? row_marks(a, sp[i].c, sp[i].uy, sp[i].ux, sp[i].n, lane)
                                  : col_marks(a, sp[i].c, sp[i].ux, sp[i].uy, sp[i].n, lane);
        m = __reduce_or_sync(kFull, m);
        if (!(m & 1))
            break;
        if (!blocking)
            return false;
        __nanosleep(backoff);
    }
    if (m & 2)
        fence_gpu();
    return true;
}

// ---- luma-dependent chroma residual scaling (vvc_intra_template.c:376-449) ------------------------------------------
// The factor of a VPDU comes from the mean of the reconstructed luma samples left of and above it.
// Returns false (try mode only) when the luma it depends on is not there yet; the factor comes back in `scale`.
// Phased pictures never get here with luma outstanding: vpdu_scale_kernel has filled the per-VPDU table.
__device__ bool derive_vpdu_scale(const IntraArgs &a, const VVCB200IntraBlock &b, bool has_marks, int lane, bool blocking,
                                  int &scale)
{
    const int size = min(1 << a.ctb_log2, 64), ls = ilog2(size);
    const int x = b.vpdu_x << ls, y = b.vpdu_y << ls;
    const bool al = b.vpdu_avail & 1, at = b.vpdu_avail & 2;
    const int rows = min(size, a.pic.h[0] - y), cols = min(size, a.pic.w[0] - x);
    if (has_marks) {
        Span sp[2];
        if (!spans_ready(a, sp, vpdu_deps(b, a.ctb_log2, a.pic.w[0], a.pic.h[0], sp), lane, blocking))
            return false;
    }
    const uint16_t *luma = a.pic.p[0] + (size_t)y * a.pic.stride[0] + x;
    int sum, cnt;
    for (unsigned backoff = 32;; backoff = min(backoff * 2, kBackoffMax)) {
        bool bad = false;
        sum = cnt = 0;
        if (al) {                                                         // past the picture edge the last sample repeats
            for (int i = lane; i < size; i += 32) {
                const unsigned v = __ldcg(luma - 1 + (ptrdiff_t)a.pic.stride[0] * min(i, rows - 1));
                bad |= v == kSentinel;
                sum += v;
            }
            cnt = size;
        }
        if (at) {
            for (int i = lane; i < size; i += 32) {
                const unsigned v = __ldcg(luma - a.pic.stride[0] + min(i, cols - 1));
                bad |= v == kSentinel;
                sum += v;
            }
            cnt += size;
        }
        if (!__any_sync(kFull, bad))
            break;
        if (!blocking)
            return false;
        __nanosleep(backoff);
    }
    sum = warp_sum(sum);
    const int mean = cnt ? (sum + (cnt >> 1)) >> ilog2(cnt) : 1 << (a.bd - 1);
    int i = a.lmcs_min_bin;
    while (i <= a.lmcs_max_bin && mean >= a.lmcs_pivot[i + 1])
        i++;
    scale = a.lmcs_coeff[min(i, 15)];
    return true;
}

__device__ __forceinline__ bool vpdu_chroma_scale(const IntraArgs &a, const VVCB200IntraBlock &b, bool has_marks, int lane,
                                                  bool blocking, int &scale)
{
    scale = 0;
    if (!(b.flags & VVCB200_IB_LMCS_SCALE))
        return true;
    if (a.vpdu_scale) {
        scale = (int)(__ldcg(a.vpdu_scale + b.vpdu_y * a.vpdu_pitch + b.vpdu_x) & 0xffff);
        return true;
    }
    return derive_vpdu_scale(a, b, has_marks, lane, blocking, scale);
}

// Phased pictures, between the luma wavefront and the chroma phases: the factor of every VPDU some record asks for,
// derived once.  One warp per chroma record; the first to claim its VPDU's table entry (0 -> 1) derives the factor and
// stores it as 0x10000 | factor, the others leave.  (Deriving on demand in the add-only batch does not work: the
// thirty-odd records of a VPDU are neighbours in the array, in flight together, and all miss together.)
__global__ void __launch_bounds__(256) vpdu_scale_kernel(const IntraArgs a)
{
    const RecordRange rr = resolve_phase(a);
    const int lane = threadIdx.x & 31;
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (t >= rr.n)
        return;
    const VVCB200IntraBlock b = load_record(rr.blocks + t);
    if (!(b.flags & VVCB200_IB_LMCS_SCALE))
        return;
    unsigned *entry = a.vpdu_scale + b.vpdu_y * a.vpdu_pitch + b.vpdu_x;
    unsigned old = 1;
    if (lane == 0)
        old = atomicCAS(entry, 0u, 1u);
    if (__shfl_sync(kFull, old, 0) != 0)
        return;
    int scale = 0;
    derive_vpdu_scale(a, b, false, lane, true, scale);
    if (lane == 0)
        *entry = 0x10000u | (unsigned)scale;
}

__device__ __forceinline__ int scale_chroma_residual(int r, int scale, int bd)
{
    const int c = clip_intp2(r, bd);
    return c > 0 ? (c * scale + (1 << 10)) >> 11 : -((-c * scale + (1 << 10)) >> 11);
}

// ---- the add-only phase of a phased picture ---------------------------------------------------------------------------
// Chroma blocks of inter units whose residual waits for the luma-dependent scale (VVCB200_IB_ADD_ONLY): once every luma
// sample of the picture is final they depend on nothing else, so they run as a plain batch — one warp per record —
// between the luma wavefront and the chroma wavefront instead of as tickets of the latter.
__global__ void __launch_bounds__(256) intra_add_only_kernel(const IntraArgs a)
{
    const RecordRange rr = resolve_phase(a);
    const int lane = threadIdx.x & 31;
    const int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (t >= rr.n)
        return;
    const VVCB200IntraBlock b = load_record(rr.blocks + t);
    if (!(b.flags & VVCB200_IB_RESIDUAL))
        return;
    int lscale = 0;
    vpdu_chroma_scale(a, b, false, lane, true, lscale);
    const int c = b.c_idx, w = b.w, h = b.h, lw = ilog2(w);
    uint16_t *dst = a.pic.p[c] + (size_t)b.y0 * a.pic.stride[c] + b.x0;
    const int16_t *res = a.res[c] + (size_t)b.y0 * a.res_stride[c] + b.x0;
    for (int idx = lane; idx < w * h; idx += 32) {
        const int x = idx & (w - 1), y = idx >> lw;
        int r = res[(size_t)y * a.res_stride[c] + x];
        if (lscale)
            r = scale_chroma_residual(r, lscale, a.bd);
        uint16_t *d = dst + (size_t)y * a.pic.stride[c] + x;
        *d = (uint16_t)clip_pixel((int)*d + r, a.bd);
    }
}

// ---- one angular / planar / DC / MIP block --------------------------------------------------------------------------
struct Store {                    // what happens to a predicted sample on its way to the picture
    uint16_t *dst;
    const int16_t *res;           // staged residual block (shared memory, rows of w), or null
    const uint16_t *inter;        // CIIP: staged inter prediction (shared memory, rows of w)
    int dstride, w, bd, ciip_w, lscale;
    int res_cols, res_shift;      // ISP: bit (x >> res_shift) of res_cols says whether column x lies in a coded partition
    __device__ __forceinline__ void put(int x, int y, int pred) const
    {
        uint16_t *d = dst + (size_t)y * dstride + x;
        if (ciip_w)                                                       // put_ciip, vvc_inter_template.c:60-76
            pred = (pred * ciip_w + (int)inter[y * w + x] * (4 - ciip_w) + 2) >> 2;
        if (res && ((res_cols >> (x >> res_shift)) & 1)) {                // add_residual, vvc_itx_1d / vvcdsp.h:154
            int r = res[y * w + x];
            if (lscale)
                r = scale_chroma_residual(r, lscale, bd);
            pred = clip_pixel(pred + r, bd);
        }
        *d = (uint16_t)pred;
    }
};

// blocking = false: one look at the neighbours; returns false, having written nothing, when they are not all there.
// arena_held: set when the record took the CTA's arena (the caller releases it once the record has been written).
__device__ bool predict_record(const IntraArgs &a, const CtaTables &tb, const VVCB200IntraBlock &b, WarpScratch &ws, CtaArena &ar,
                               bool &arena_held, int lane, bool has_marks, bool scaling, bool blocking, int skip_wait = 0)
{
    const long long p0 = PROF_T();
    const int c = b.c_idx, w = b.w, h = b.h, mode = b.mode, ref_idx = b.ref_idx, bd = a.bd;
    const int lw = ilog2(w), lh = ilog2(h);
    const bool is_mip = b.flags & VVCB200_IB_MIP, isp = b.flags & VVCB200_IB_ISP;
    const bool pdpc = need_pdpc(tb, w, h, lw, lh, b.flags & VVCB200_IB_BDPCM, mode, ref_idx);
    const bool rff = !is_mip && ref_filter_mode(mode);
    const EdgeExtent ee = edge_extent(b);
    const bool smooth = ee.smooth;
    const int ref_line = ref_idx == 3 ? -4 : -1 - ref_idx;
    const int stride = a.pic.stride[c];
    uint16_t *src = a.pic.p[c] + (size_t)b.y0 * stride + b.x0;
    const bool has_res = b.flags & VVCB200_IB_RESIDUAL;
    const bool ciip = b.flags & VVCB200_IB_CIIP, add_only = b.flags & VVCB200_IB_ADD_ONLY;
    // Reads that do not depend on the neighbours (the residual, the inter prediction of an in-place block): a warp that
    // is going to wait for this record starts them now so that they land meanwhile; one that only takes a look starts
    // them once it knows the record will run.
    const bool stages = has_res || ciip;
    const bool big = stages && w * h > kSmallSamples;                     // staged in the arena, after the wait
    uint8_t *const res_buf = big ? ar.stage_res : ws.stage_res, *const extra_buf = big ? ar.stage_extra : ws.stage_extra;
    auto stage_inputs = [&]() {
        if (big) {
            arena_acquire(ar, lane);
            arena_held = true;
        }
        if (has_res)
            stage_block_async(res_buf, a.res[c] + (size_t)b.y0 * a.res_stride[c] + b.x0, a.res_stride[c], w, h, lane);
        if (ciip || (add_only && has_res)) {                              // the block's own samples before this record
            if (a.shadow[c])
                stage_block_async(extra_buf, a.shadow[c] + (size_t)b.y0 * a.shadow_stride[c] + b.x0, a.shadow_stride[c], w, h, lane);
            else
                stage_block_async(extra_buf, src, stride, w, h, lane);
        }
    };
    const bool early = blocking && !big;
    if (early)
        stage_inputs();
    int lscale = 0;
    if (scaling && !vpdu_chroma_scale(a, b, has_marks, lane, blocking, lscale))
        return false;

    if (add_only) {                                                       // the prediction is there already
        if (has_res) {
            if (!early)
                stage_inputs();
            const int16_t *res = reinterpret_cast<const int16_t *>(res_buf);
            const uint16_t *cur = reinterpret_cast<const uint16_t *>(extra_buf);
            stage_wait();
            __syncwarp();
            for (int idx = lane; idx < w * h; idx += 32) {
                const int x = idx & (w - 1), y = idx >> lw;
                int r = res[idx];
                if (lscale)
                    r = scale_chroma_residual(r, lscale, bd);
                src[(size_t)y * stride + x] = (uint16_t)clip_pixel((int)cur[idx] + r, bd);
            }
        }
        return true;
    }

    const int left_size = ee.left_size, ul = ee.ul, ut = ee.ut, refw = ee.refw, refh = ee.refh;
    const int la = min(ul, (int)b.avail_left), ta = min(ut, (int)b.avail_top);

    // ---- neighbours written in place (CIIP, add-only) announce themselves through the marks
    const long long p1 = PROF_T();
    PROF_ADD(7, p1 - p0);
    if (has_marks) {
        Span sp[3];
        if (!spans_ready(a, sp, predict_deps(b, skip_wait, sp), lane, blocking))
            return false;
    }
    PROF_ADD(4, PROF_T() - p1);

    // ---- which of the edge samples the prediction really consumes.  The reference fetches `ul` / `ut` samples whatever
    // the mode; samples that never reach the output need not be waited for (their value, sentinel or not, is
    // irrelevant), which removes the above-right / below-left dependency for most modes and shortens the chains.
    int need_l = ul, need_t = ut;
    if (is_mip) {
        need_l = h; need_t = w;
    } else if (mode == 1) {
        if (!pdpc) { need_t = w >= h ? w : 0; need_l = h >= w ? h : 0; }
    } else if (mode == 50) {
        need_l = pdpc ? h : 0;
    } else if (mode == 18) {
        need_t = pdpc ? w : 0;
    } else if (mode != 0) {
        const bool vertical = mode >= 34;
        const int ang = tb.angle[mode + 14];
        const int n_main = vertical ? w : h, n_al = vertical ? h : w;
        int need_main, need_side;
        if (ang > 0) {
            need_main = n_main + (((n_al + ref_idx) * ang) >> 5) + 2 + smooth;
            need_side = 0;
            if (pdpc) {
                const int ns = nscale_of(tb, lw, lh, mode);
                need_side = n_al + ((256 + min(n_main, 3 << ns) * tb.inv_angle[mode + 14]) >> 9) + 1 + smooth;
            }
        } else {
            need_main = n_main + 2 + smooth;
            need_side = n_al + 2 + smooth;
        }
        need_t = min(ut, vertical ? need_main : need_side);
        need_l = min(ul, vertical ? need_side : need_main);
    }

    // ---- gather the edges into shared memory; a sentinel among them means "not reconstructed yet": read again
    const long long g0 = PROF_T();
    uint16_t *left = ws.edge[0] + kEdgeOrg, *top = ws.edge[1] + kEdgeOrg;
    for (unsigned backoff = 32;; backoff = min(backoff * 2, kBackoffMax)) {
        unsigned lv[4], tv[4], vl = 0, vt = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = lane + 32 * k;
            lv[k] = i < la ? __ldcg(src + ref_line + (ptrdiff_t)stride * i) : 0;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = lane + 32 * k;
            tv[k] = i < ta ? __ldcg(src + i + (ptrdiff_t)stride * ref_line) : 0;
        }
        if (lane < -ref_line) {                                           // the corner run, indices -1 .. ref_line
            const int i = -1 - lane;
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
        }
        bool bad = vl == kSentinel || vt == kSentinel;
#pragma unroll
        for (int k = 0; k < 4; k++)
            bad |= (lv[k] == kSentinel && lane + 32 * k < need_l) || (tv[k] == kSentinel && lane + 32 * k < need_t);
        if (__any_sync(kFull, bad)) {
            if (!blocking)
                return false;
            __nanosleep(backoff);
            continue;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int i = lane + 32 * k;
            if (i < la) left[i] = lv[k];
            if (i < ta) top[i] = tv[k];
        }
        if (lane < -ref_line) {
            left[-1 - lane] = vl;
            top[-1 - lane] = vt;
        }
        break;
    }
    if (!early)
        stage_inputs();
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
            angle = tb.angle[mode + 14];
            const int inv = tb.inv_angle[mode + 14];
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
    st.ciip_w = ciip ? b.ciip_weight : 0;
    st.res = has_res ? reinterpret_cast<const int16_t *>(res_buf) : nullptr;
    st.inter = reinterpret_cast<const uint16_t *>(extra_buf);
    st.w = w;
    st.lscale = lscale;
    st.res_cols = 1;
    st.res_shift = 8;                                                     // every column maps to bit 0
    if (isp && !c && has_res) {
        st.res_cols = b.isp_coded;
        if (isp_vertical(b))
            st.res_shift = ilog2(b.cb_w / isp_tb_count(b));
    }
    const long long g1 = PROF_T();
    stage_wait();                                                         // residual / inter prediction are in shared memory
    __syncwarp();
    PROF_ADD(5, g1 - g0);
    PROF_ADD(6, PROF_T() - g1);
    const int n = w * h;

    if (is_mip) {                                                         // vvc_intra_template.c:707-836
        const int size_id = (w == 4 && h == 4) ? 0 : ((w == 4 || h == 4) || (w == 8 && h == 8)) ? 1 : 2;
        const int bsz = size_id ? 4 : 2, psz = size_id == 2 ? 8 : 4, lp = size_id == 2 ? 3 : 2;
        const int in_size = 2 * bsz - (size_id == 2);
        const int nout = psz * psz;
        const bool tr = b.flags & VVCB200_IB_MIP_TRANSPOSED;
        const uint8_t *m = tb.mip_mat + tb.mip_off[size_id] + b.mip_mode * nout * in_size;
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
        return true;
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
        return true;
    }

    // angular (vvc_intra_template.c:838-1001): `along` runs with the prediction direction's axis, `cross` across it
    const bool vertical = mode >= 34;
    const int inv = pdpc ? tb.inv_angle[mode + 14] : 0, nscale = pdpc ? nscale_of(tb, lw, lh, mode) : 0;
    const uint16_t *main_ref = (vertical ? top : left) - (1 + ref_idx), *side = vertical ? left : top;
    const int8_t *ftab = tb.luma_filter + interp * 128;
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
    return true;
}

__device__ __forceinline__ int checked_load(const uint16_t *p, bool &bad)
{
    const unsigned v = __ldcg(p);
    bad |= v == kSentinel;
    return (int)v;
}

// ---- CCLM (vvc_intra_template.c:29-375): one record predicts Cb and Cr --------------------------------------------
__device__ bool cclm_record(const IntraArgs &a, const VVCB200IntraBlock &b, WarpScratch &ws, CtaArena &ar, bool &arena_held,
                            int lane, bool has_marks, bool scaling, bool blocking)
{
    const int bd = a.bd, w = b.w, h = b.h, x = b.x0, y = b.y0, x0 = x * 2, y0 = y * 2, mode = b.mode;
    const bool avail_t = b.avail_top > 0, avail_l = b.avail_left > 0;
    const int ls = a.pic.stride[0];
    const uint16_t *luma = a.pic.p[0] + (size_t)y0 * ls + x0;
    const int cs[2] = { a.pic.stride[1], a.pic.stride[2] };
    uint16_t *dst[2] = { a.pic.p[1] + (size_t)y * cs[0] + x, a.pic.p[2] + (size_t)y * cs[1] + x };
    const CclmCounts kc = cclm_counts(b);
    const int num[2] = { kc.num[0], kc.num[1] };
    const int n = w * h, lw = ilog2(w);
    const int16_t *res[2] = { nullptr, nullptr };
    const int wx0 = avail_l ? -4 : 0, wy0 = avail_t ? -1 : 0, wpitch = 2 * w + 4;
    // too large for the warp's own buffers: residuals and luma window go to the CTA's arena, taken after the last wait
    const bool big = 2 * n * 2 > kSmallRes || wpitch * (2 * h + 1) * 2 > kSmallExtra;
    bool staged = false;
    auto stage_residuals = [&]() {                                        // before waiting, or once the record is known to run
        if (staged)
            return;
        staged = true;
        if (big && !arena_held) {
            arena_acquire(ar, lane);
            arena_held = true;
        }
        if (b.flags & VVCB200_IB_RESIDUAL)
            for (int c = 0; c < 2; c++)
                if (b.res_planes & (1 << c)) {
                    int16_t *sres = reinterpret_cast<int16_t *>(big ? ar.stage_res : ws.stage_res) + c * n;
                    stage_block_async(sres, a.res[c + 1] + (size_t)y * a.res_stride[c + 1] + x, a.res_stride[c + 1], w, h, lane);
                    res[c] = sres;
                }
    };
    if (blocking && !big)
        stage_residuals();
    if (has_marks) {
        Span sp[kMaxSpans];
        if (!spans_ready(a, sp, cclm_deps(b, sp), lane, blocking))
            return false;
    }
    int lscale = 0;
    if (scaling && !vpdu_chroma_scale(a, b, has_marks, lane, blocking, lscale))
        return false;
    __syncwarp();
    auto emit = [&](int c, int i, int j, int v) {
        if (res[c]) {
            int r = res[c][j * w + i];
            if (lscale)
                r = scale_chroma_residual(r, lscale, bd);
            v = clip_pixel(v + r, bd);
        }
        dst[c][(size_t)j * cs[c] + i] = (uint16_t)v;
    };
    if (!avail_t && !avail_l) {
        stage_residuals();
        stage_wait();
        __syncwarp();
        for (int idx = lane; idx < n; idx += 32)
            for (int c = 0; c < 2; c++)
                emit(c, idx & (w - 1), idx >> lw, 1 << (bd - 1));
        return true;
    }
    // The co-located luma window, columns -4 (or 0) .. 2w - 1, rows -1 (or 0) .. 2h - 1, into shared memory in 8-byte
    // pieces.  It comes after the model: the block's own luma is usually the last thing to arrive, and by then only the
    // fetch, the down-sampling and the stores are left.  Poll its last row cheaply, then fetch and verify everything.
    uint16_t *win = reinterpret_cast<uint16_t *>(big ? ar.stage_extra : ws.stage_extra);
    auto stage_window = [&]() -> bool {
        const int per_row = (2 * w - wx0) >> 2, rows = 2 * h - wy0;
        for (unsigned backoff = 32;; backoff = min(backoff * 2, kBackoffMax)) {
            bool bad = false;
            for (int i = lane; i < 2 * w; i += 32)
                bad |= __ldcg(luma + (ptrdiff_t)ls * (2 * h - 1) + i) == kSentinel;
            if (!__any_sync(kFull, bad)) {
                if (big && !arena_held) {                                 // look only: the arena is not ours while we may still wait
                    for (int i = lane; i < per_row * rows; i += 32) {
                        const int r = i / per_row, q = i - r * per_row;
                        const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(luma + (ptrdiff_t)ls * (wy0 + r) + wx0 + 4 * q));
                        bad |= (v.x & 0xffffu) == kSentinel || (v.x >> 16) == kSentinel || (v.y & 0xffffu) == kSentinel || (v.y >> 16) == kSentinel;
                    }
                    if (__any_sync(kFull, bad)) {
                        if (!blocking)
                            return false;
                        __nanosleep(backoff);
                        continue;
                    }
                    arena_acquire(ar, lane);
                    arena_held = true;
                }
                for (int i = lane; i < per_row * rows; i += 32) {
                    const int r = i / per_row, q = i - r * per_row;
                    const uint2 v = __ldcg(reinterpret_cast<const uint2 *>(luma + (ptrdiff_t)ls * (wy0 + r) + wx0 + 4 * q));
                    bad |= (v.x & 0xffffu) == kSentinel || (v.x >> 16) == kSentinel || (v.y & 0xffffu) == kSentinel || (v.y >> 16) == kSentinel;
                    *reinterpret_cast<uint2 *>(win + (r + 1 + wy0) * wpitch + 4 + wx0 + 4 * q) = v;
                }
                if (!__any_sync(kFull, bad))
                    return true;
            }
            if (!blocking)
                return false;
            __nanosleep(backoff);
        }
    };
    // A record with buffers of its own fetches the window first: it then travels while the warp waits for the chroma
    // neighbours, which is what the chains of an intra picture run through.  A record that borrows the arena may only
    // take it once nothing is left to wait for, so its window comes after the neighbours.
    if (!big && !stage_window())
        return false;
#define L(dx, dy) checked_load(luma + (dx) + (ptrdiff_t)ls * (dy), bad)
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
        for (unsigned backoff = 32;; backoff = min(backoff * 2, kBackoffMax)) {  // until none of the neighbours is a sentinel
        bool bad = false;
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
            v1 = checked_load(dst[0] + pos - cs[0], bad);
            v2 = checked_load(dst[1] + pos - cs[1], bad);
        } else if (lane < cnt[0] + cnt[1]) {
            const int pos = start[1] + (lane - cnt[0]) * step[1], ly = pos * 2;
            if (a.collocated) {
                const int yt = (ly || avail_t) ? ly - 1 : ly;
                v0 = (L(-3, ly) + L(-2, yt) + 4 * L(-2, ly) + L(-1, ly) + L(-2, ly + 1) + 4) >> 3;
            } else {
                v0 = (L(-3, ly) + L(-3, ly + 1) + 2 * L(-2, ly) + 2 * L(-2, ly + 1) + L(-1, ly) + L(-1, ly + 1) + 4) >> 3;
            }
            v1 = checked_load(dst[0] - 1 + (ptrdiff_t)pos * cs[0], bad);
            v2 = checked_load(dst[1] - 1 + (ptrdiff_t)pos * cs[1], bad);
        }
        if (!__any_sync(kFull, bad))
            break;
        if (!blocking)
            return false;
        __nanosleep(backoff);
        }
        if (!big)
            stage_residuals();                                            // travel while the model is derived
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
    if (big && !stage_window())
        return false;
    stage_residuals();
    stage_wait();
    __syncwarp();
#define W(dx, dy) ((int)win[((dy) + 1) * wpitch + 4 + (dx)])
    for (int idx = lane; idx < n; idx += 32) {                            // down-sampled luma -> both chroma planes
        const int i = idx & (w - 1), j = idx >> lw;
        const int lx = 2 * i, ly = 2 * j;
        const int xl = (i || avail_l) ? lx - 1 : lx;
        int ds;
        if (a.collocated) {
            const int yt = (j || avail_t) ? ly - 1 : ly;
            ds = (W(xl, ly) + W(lx, yt) + 4 * W(lx, ly) + W(lx + 1, ly) + W(lx, ly + 1) + 4) >> 3;
        } else {
            ds = (W(xl, ly) + W(xl, ly + 1) + 2 * W(lx, ly) + 2 * W(lx, ly + 1) + W(lx + 1, ly) + W(lx + 1, ly + 1) + 4) >> 3;
        }
        for (int c = 0; c < 2; c++)
            emit(c, i, j, clip_pixel(((ds * ka[c]) >> kk[c]) + kb[c], bd));
    }
#undef W
#undef L
    return true;
}

// Everything an ISP coding block reads from outside itself: the column left of it and the row above it as far as any
// sub-partition reaches (twice the block's size, limited by what the host reported available), and the corner.
__device__ bool isp_external_ready(const IntraArgs &a, const VVCB200IntraBlock &b, bool has_marks, int lane)
{
    if (has_marks) {
        Span sp[3];
        const int parts = isp_parts(b);
        for (int q = 0; q < parts; q++) {
            VVCB200IntraBlock sub;
            const int skip = isp_part(b, q, sub);
            if (!spans_ready(a, sp, predict_deps(sub, skip, sp), lane, false))
                return false;
        }
    }
    const int stride = a.pic.stride[0];
    const uint16_t *src = a.pic.p[0] + (size_t)b.y0 * stride + b.x0;
    const int nl = min((int)b.avail_left, 2 * b.cb_h), nt = min((int)b.avail_top, 2 * b.cb_w);
    bool bad = false;
    for (int i = lane; i < nl; i += 32)
        bad |= __ldcg(src - 1 + (ptrdiff_t)stride * i) == kSentinel;
    for (int i = lane; i < nt; i += 32)
        bad |= __ldcg(src - stride + i) == kSentinel;
    if (lane == 0 && b.up_left)
        bad |= __ldcg(src - stride - 1) == kSentinel;
    return !__any_sync(kFull, bad);
}

// kMarks: some blocks are rewritten in place and go by marks (no shadow picture).  The other instantiation carries none
// of that code, which matters for a kernel whose steps wait on instruction fetch.
// kScale: the picture uses luma-dependent chroma residual scaling (VVCB200IntraParams.lmcs_chroma_scale).
template <bool kMarks, bool kScale>
__global__ void __launch_bounds__(kWarps * 32, kCtasPerSm) intra_kernel(const IntraArgs a)
{
    extern __shared__ __align__(16) uint8_t s_dyn[];
    __shared__ CtaTables tb;
    const RecordRange rr = resolve_phase(a);
    const int lane = threadIdx.x & 31;
    WarpScratch &ws = reinterpret_cast<WarpScratch *>(s_dyn)[threadIdx.x >> 5];
    CtaArena &ar = *reinterpret_cast<CtaArena *>(s_dyn + sizeof(WarpScratch) * kWarps);
    if (threadIdx.x == 0)
        ar.lock = 0;
    for (int i = threadIdx.x; i < 95; i += blockDim.x) {
        tb.angle[i] = VVC_INTRA_ANGLE[i];
        tb.inv_angle[i] = VVC_INTRA_INV_ANGLE[i];
    }
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
        tb.luma_filter[i] = VVC_INTRA_LUMA_FILTER[i];
    for (int i = threadIdx.x; i < 4736; i += blockDim.x)
        tb.mip_mat[i] = VVC_MIP_MAT[i];
    if (threadIdx.x < 4)
        tb.mip_off[threadIdx.x] = VVC_MIP_OFF[threadIdx.x];
    __syncthreads();
    const bool has_marks = kMarks && *a.inplace_flag != 0;                 // written by the pre-pass

    // Each warp holds up to kSlots records drawn in ticket order.  With one slot it waits for that record's neighbours
    // (tight polling: the lowest latency along a dependency chain, which is what bounds the kernel).  With more slots it
    // looks at its records in turn and runs whichever is complete; measured on B200 that widens the window of records
    // that can start early but roughly doubles the time a finished neighbour goes unnoticed, and loses.
    // Neither can deadlock: the lowest unfinished ticket depends only on finished records and its holder keeps looking.
    constexpr bool kWait = kSlots == 1;
    VVCB200IntraBlock *slot = ws.slot;
    unsigned held = 0;                                                    // bit k: slot k is occupied (warp-uniform)
    bool tickets_left = true;
    unsigned backoff = 32;
    for (;;) {
        const long long pt0 = PROF_T();
        while (tickets_left && held != (1u << kSlots) - 1) {              // refill in ticket order
            int t = 0;
            if (lane == 0)
                t = atomicAdd(a.ticket, 1);
            t = __shfl_sync(kFull, t, 0);
            if (t >= rr.n) {
                tickets_left = false;
                break;
            }
            const int k = __ffs(~held) - 1;
            if (lane < 7)
                reinterpret_cast<uint32_t *>(&slot[k])[lane] = __ldg(reinterpret_cast<const uint32_t *>(rr.blocks + t) + lane);
            held |= 1u << k;
        }
        __syncwarp();
        if (!held)
            break;
        PROF_ADD(0, PROF_T() - pt0);
        bool progress = false;
        for (int k = 0; k < kSlots && !progress; k++) {
            if (!(held & (1u << k)))
                continue;
            const VVCB200IntraBlock b = slot[k];
            const long long pt1 = PROF_T();
            bool done, arena_held = false;
            if (b.mode >= 81) {
                done = cclm_record(a, b, ws, ar, arena_held, lane, has_marks, kScale, kWait);
            } else {
                // ISP: the sub-partitions of the coding block one after the other, each predicted from its
                // reconstructed predecessor (vvc_intra.c:217-226, 244-252); w x h is the prediction unit.  Everything the
                // partitions read from outside the block must be there before the first one starts; after that the warp
                // only ever waits for its own stores.  Every other record is a single "partition" (one call site keeps
                // the kernel's code, and with it the pressure on the instruction cache, small).
                const bool isp = (b.flags & VVCB200_IB_ISP) && !b.c_idx;
                const int n_parts = isp ? isp_parts(b) : 1;
                done = !isp || kWait || isp_external_ready(a, b, has_marks, lane);
                for (int q = 0; done && q < n_parts; q++) {
                    VVCB200IntraBlock sub = b;
                    const int skip = isp ? isp_part(b, q, sub) : 0;
                    done = predict_record(a, tb, sub, ws, ar, arena_held, lane, has_marks, kScale, kWait || isp, skip);
                    __syncwarp();
                    if (arena_held && q + 1 < n_parts) {                  // the next partition may wait for this one's stores
                        arena_release(ar, lane);
                        arena_held = false;
                    }
                }
            }
            __syncwarp();
            if (arena_held)
                arena_release(ar, lane);
            if (!done)
                continue;
            const long long pt2 = PROF_T();
            PROF_ADD(1, pt2 - pt1);
            PROF_ADD(3, 1);
            if (has_marks && in_place(b)) {                               // samples before marks
                fence_gpu();
                __syncwarp();
                set_marks(a, b.c_idx, b.x0, b.y0, b.w, b.h, 2, lane);
                __syncwarp();
            }
            PROF_ADD(2, PROF_T() - pt2);
            held &= ~(1u << k);
            progress = true;
        }
        if (progress) {
            backoff = 32;
        } else {
            __nanosleep(backoff);
            backoff = min(backoff * 2, 512u);
        }
    }
}

// ---- the record order made on the device (vvcb200_intra_schedule_device) ---------------------------------------------
// Same order as the host helper's diagonal form, per phase: CTU anti-diagonals x + 2y, the CTUs of a diagonal dealt
// record by record, each CTU's records of a phase in decoding order.
// 1. one warp per CTU: every record's phase and its index k among the CTU's records of that phase; the CTU's three counts
__global__ void __launch_bounds__(256) intra_rank_kernel(const VVCB200IntraBlock *in, const uint32_t *ctu_first, int nctu,
                                                         uint32_t *rank, uint16_t *ctu_of, int *len)
{
    const int lane = threadIdx.x & 31;
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= nctu)
        return;
    const uint32_t first = ctu_first[c], last = ctu_first[c + 1];
    int cnt[3] = { 0, 0, 0 };
    for (uint32_t i0 = first; i0 < last; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < last;
        int p = -1;
        if (valid) {
            const uint32_t *w = reinterpret_cast<const uint32_t *>(in + i);
            const uint32_t w1 = __ldg(w + 1), w2 = __ldg(w + 2);                  // (w, h, c_idx, mode), (ref_idx, flags, ...)
            const int c_idx = (w1 >> 16) & 0xff, flags = (w2 >> 8) & 0xff;
            p = !c_idx ? 0 : (flags & VVCB200_IB_ADD_ONLY) ? 1 : 2;
        }
        int k = 0;
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const unsigned m = __ballot_sync(kFull, p == q);
            if (p == q)
                k = cnt[q] + __popc(m & ((1u << lane) - 1));
            cnt[q] += __popc(m);
        }
        if (valid) {
            rank[i] = ((uint32_t)p << 30) | (uint32_t)k;
            ctu_of[i] = (uint16_t)c;
        }
    }
    if (lane < 3)
        len[c * 3 + lane] = lane == 0 ? cnt[0] : lane == 1 ? cnt[1] : cnt[2];
}

__device__ __forceinline__ void diagonal_rows(int d, int ctus_w, int ctus_h, int &y_lo, int &y_hi)
{
    y_lo = d - (ctus_w - 1) > 0 ? (d - (ctus_w - 1) + 1) / 2 : 0;
    y_hi = d / 2 < ctus_h - 1 ? d / 2 : ctus_h - 1;
}

// 2. one CTA: records per (phase, diagonal), their exclusive prefix, the three phase totals
__global__ void __launch_bounds__(256) intra_base_kernel(const int *len, int ctus_w, int ctus_h, int *diag_base, int *info)
{
    const int nd = ctus_w + 2 * (ctus_h - 1);
    for (int d = threadIdx.x; d < nd; d += blockDim.x) {
        int y_lo, y_hi, tot[3] = { 0, 0, 0 };
        diagonal_rows(d, ctus_w, ctus_h, y_lo, y_hi);
        for (int y = y_lo; y <= y_hi; y++) {
            const int *l = len + (y * ctus_w + d - 2 * y) * 3;
            tot[0] += l[0]; tot[1] += l[1]; tot[2] += l[2];
        }
        for (int p = 0; p < 3; p++)
            diag_base[p * nd + d] = tot[p];
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int p = threadIdx.x;
        int run = 0;
        for (int d = 0; d < nd; d++) {
            const int v = diag_base[p * nd + d];
            diag_base[p * nd + d] = run;
            run += v;
        }
        info[p] = run;
    }
}

// 3. one thread per record: its place, then the copy
__global__ void __launch_bounds__(256) intra_scatter_kernel(const VVCB200IntraBlock *in, int n, const uint32_t *rank,
                                                            const uint16_t *ctu_of, const int *len, const int *diag_base,
                                                            const int *info, int ctus_w, int ctus_h, VVCB200IntraBlock *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    const uint32_t r = rank[i];
    const int p = r >> 30, k = r & 0x3fffffff, c = ctu_of[i];
    const int cy = c / ctus_w, cx = c - cy * ctus_w, d = cx + 2 * cy, nd = ctus_w + 2 * (ctus_h - 1);
    int y_lo, y_hi;
    diagonal_rows(d, ctus_w, ctus_h, y_lo, y_hi);
    int pos = diag_base[p * nd + d] + (p > 0 ? info[0] : 0) + (p > 1 ? info[1] : 0);
    for (int y = y_lo; y <= y_hi; y++) {
        const int L = len[(y * ctus_w + d - 2 * y) * 3 + p];
        pos += min(L, k) + (y < cy && L > k ? 1 : 0);
    }
    const uint32_t *src = reinterpret_cast<const uint32_t *>(in + i);
    uint32_t *dst = reinterpret_cast<uint32_t *>(out + pos);
#pragma unroll
    for (int j = 0; j < 7; j++)
        dst[j] = __ldg(src + j);
}

struct SchedLayout {
    size_t rank, ctu_of, len, diag, total;
};

SchedLayout sched_layout(int n, int ctus_w, int ctus_h)
{
    SchedLayout l;
    size_t o = 0;
    auto take = [&](size_t bytes) { const size_t at = o; o = (o + bytes + 255) & ~(size_t)255; return at; };
    l.rank = take((size_t)n * 4);
    l.ctu_of = take((size_t)n * 2);
    l.len = take((size_t)ctus_w * ctus_h * 3 * 4);
    l.diag = take((size_t)(ctus_w + 2 * (ctus_h - 1)) * 3 * 4);
    l.total = o;
    return l;
}

struct ScratchLayout {
    size_t off[3], marks_total;   // ticket sector + marks
    size_t vpdu_off, total;       // + the per-VPDU scale table of phased pictures
    int pitch[3], vpdu_pitch;
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
    l.marks_total = o;
    l.vpdu_off = o;
    l.vpdu_pitch = (width + 31) >> 5;                                     // VPDUs are 64x64, or the CTB when that is 32x32
    o += (size_t)l.vpdu_pitch * ((height + 31) >> 5) * sizeof(unsigned);
    l.total = (o + 255) & ~(size_t)255;
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

#include <stdlib.h>
#include <vector>

// Host scheduler.  Depth of every record in the dependency graph: 1 + the largest depth among the records that wrote
// the units it waits for (the same runs the kernel would poll).  Records come CTU by CTU in decoding order, so every
// writer a record depends on has been seen.  Sorting by depth is a topological order in which a window of
// consecutive tickets holds the records that are ready now anywhere in the picture.
// Map entries are (epoch << 16 | depth): a new picture bumps the epoch instead of clearing the maps; a record only
// stamps the units other records can read (its top row, bottom row and right column).
namespace vvcb200 {
struct IntraScheduler {
    std::vector<uint32_t> map[3];
    std::vector<uint16_t> depth;
    std::vector<uint32_t> start;
    int uw = 0, uh = 0;
    uint32_t epoch = 0;

    int run(const VVCB200IntraBlock *in, const uint32_t *ctu_first, int ctus_w, int ctus_h, int ctb_log2, int width,
            int height, VVCB200IntraBlock *out, int32_t *n_levels)
    {
        const uint32_t n = ctu_first[ctus_w * ctus_h];
        const int w_units = ((ctus_w << ctb_log2) >> 2) + 2, h_units = ((ctus_h << ctb_log2) >> 2) + 2;   // one guard column / row
        if (w_units != uw || h_units != uh || ++epoch >= 0xffff) {
            uw = w_units;
            uh = h_units;
            epoch = 1;
            for (int c = 0; c < 3; c++)
                map[c].assign((size_t)uw * uh, 0);                        // chroma: half the samples, 2-row units: same count
        }
        depth.resize(n);
        const uint32_t tag = epoch << 16;
        uint32_t max_depth = 0;
        Span sp[kMaxSpans + 3];
        for (uint32_t i = 0; i < n; i++) {
            const VVCB200IntraBlock &b = in[i];
            if (b.c_idx > 2 || (b.x0 >> 2) >= uw - 2 || (b.y0 >> unit_shift(b.c_idx)) >= uh - 2)
                return -(int)i - 1;
            uint32_t d = 0;
            auto take = [&](int cnt) {
                for (int k = 0; k < cnt; k++) {
                    const uint32_t *m = map[sp[k].c].data() + (size_t)(sp[k].uy + 1) * uw + sp[k].ux + 1;
                    const size_t step = sp[k].horizontal ? 1 : (size_t)uw;
                    uint32_t best = 0;
                    for (int j = 0; j < sp[k].n; j++) {
                        const uint32_t v = m[j * step] ^ tag;                 // < 65536 iff stamped in this picture
                        best = v < 65536u && v > best ? v : best;
                    }
                    d = best > d ? best : d;
                }
            };
            take(vpdu_deps(b, ctb_log2, width, height, sp));
            const bool isp = (b.flags & VVCB200_IB_ISP) && !b.c_idx;
            if (b.mode >= 81) {
                take(cclm_deps(b, sp));
            } else if (isp) {
                const int parts = isp_parts(b);
                for (int k = 0; k < parts; k++) {
                    VVCB200IntraBlock s;
                    const int skip = isp_part(b, k, s);
                    take(predict_deps(s, skip, sp));
                }
            } else {
                take(predict_deps(b, 0, sp));
            }
            depth[i] = (uint16_t)++d;
            max_depth = d > max_depth ? d : max_depth;
            if (d >= 0xffff)
                return -(int)i - 1;
            const int w = isp ? b.cb_w : b.w, h = isp ? b.cb_h : b.h;
            const uint32_t stamp = tag | d;
            for (int c = b.c_idx; c <= (b.mode >= 81 ? 2 : b.c_idx); c++) {
                const int s = unit_shift(c);
                const int nx = unit_count(b.x0, w, 2), ny = unit_count(b.y0, h, s);
                uint32_t *m = map[c].data() + (size_t)((b.y0 >> s) + 1) * uw + (b.x0 >> 2) + 1;
                for (int x = 0; x < nx; x++)
                    m[x] = m[(size_t)(ny - 1) * uw + x] = stamp;
                for (int y = 1; y + 1 < ny; y++)
                    m[(size_t)y * uw + nx - 1] = stamp;
            }
        }
        start.assign(max_depth + 2, 0);                                   // counting sort by depth, stable
        for (uint32_t i = 0; i < n; i++)
            start[depth[i] + 1]++;
        for (uint32_t d = 1; d <= max_depth + 1; d++)
            start[d] += start[d - 1];
        for (uint32_t i = 0; i < n; i++)
            out[start[depth[i]]++] = in[i];
        if (n_levels)
            *n_levels = (int32_t)max_depth;
        return 0;
    }
};

IntraScheduler *intra_scheduler_new() { return new IntraScheduler; }
void intra_scheduler_free(IntraScheduler *s) { delete s; }
int intra_scheduler_run(IntraScheduler *s, const VVCB200IntraBlock *in, const uint32_t *ctu_first, int ctus_w, int ctus_h,
                        int ctb_log2, int width, int height, VVCB200IntraBlock *out, int32_t *n_levels)
{
    return s->run(in, ctu_first, ctus_w, ctus_h, ctb_log2, width, height, out, n_levels);
}
}  // namespace vvcb200

namespace vvcb200 {
// phase of a record: 0 luma (and anything that is not a chroma add-only record), 1 chroma add-only, 2 other chroma
static inline int intra_phase(const VVCB200IntraBlock &b)
{
    return !b.c_idx ? 0 : (b.flags & VVCB200_IB_ADD_ONLY) ? 1 : 2;
}
}  // namespace vvcb200

extern "C" int vvcb200_intra_schedule(const VVCB200IntraBlock *in, const uint32_t *ctu_first, int ctus_w, int ctus_h,
                                      int ctb_log2_size, int width, int height, int order, VVCB200IntraBlock *out,
                                      int32_t *n_levels, int32_t *phase_counts)
{
    if (!in || !ctu_first || !out || ctus_w <= 0 || ctus_h <= 0 || ctb_log2_size < 5 || ctb_log2_size > 7 ||
        width <= 0 || height <= 0 || width > (ctus_w << ctb_log2_size) || height > (ctus_h << ctb_log2_size) ||
        (order != VVCB200_INTRA_ORDER_DIAGONAL && order != VVCB200_INTRA_ORDER_DEPTH)) {
        set_error("vvcb200_intra_schedule: bad arguments");
        return VVCB200_ERR_ARG;
    }
    if (order == VVCB200_INTRA_ORDER_DIAGONAL) {
        // CTU anti-diagonals x + 2y (the above-right CTU comes first, vvc_thread.c:161), the CTUs of a diagonal dealt
        // round-robin so that a window of consecutive tickets spans every CTU that can run now.  O(n).
        size_t cursor[3] = { 0, 0, 0 };                                   // where the next record of each phase goes
        if (phase_counts) {
            const uint32_t n = ctu_first[ctus_w * ctus_h];
            size_t cnt[3] = { 0, 0, 0 };
            for (uint32_t i = 0; i < n; i++)
                cnt[intra_phase(in[i])]++;
            cursor[1] = cnt[0];
            cursor[2] = cnt[0] + cnt[1];
            phase_counts[0] = (int32_t)cnt[0];
            phase_counts[1] = (int32_t)cnt[1];
        }
        for (int d = 0; d <= ctus_w - 1 + 2 * (ctus_h - 1); d++) {
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
                    if (ctu_first[rs] + k < ctu_first[rs + 1]) {
                        const VVCB200IntraBlock &b = in[ctu_first[rs] + k];
                        out[cursor[phase_counts ? intra_phase(b) : 0]++] = b;
                    }
                }
        }
        if (n_levels)
            *n_levels = 0;
        return 0;
    }
    static thread_local IntraScheduler sched;                             // keeps its maps from call to call
    const int r = sched.run(in, ctu_first, ctus_w, ctus_h, ctb_log2_size, width, height, out, n_levels);
    if (r) {
        set_error("vvcb200_intra_schedule: record %d lies outside the picture or the dependency chain is too long", -r - 1);
        return VVCB200_ERR_ARG;
    }
    if (phase_counts) {                                                   // stable partition of the depth order
        static thread_local std::vector<VVCB200IntraBlock> tmp;
        const uint32_t n = ctu_first[ctus_w * ctus_h];
        tmp.assign(out, out + n);
        size_t cnt[3] = { 0, 0, 0 };
        for (uint32_t i = 0; i < n; i++)
            cnt[intra_phase(tmp[i])]++;
        size_t cursor[3] = { 0, cnt[0], cnt[0] + cnt[1] };
        for (uint32_t i = 0; i < n; i++)
            out[cursor[intra_phase(tmp[i])]++] = tmp[i];
        phase_counts[0] = (int32_t)cnt[0];
        phase_counts[1] = (int32_t)cnt[1];
    }
    return 0;
}

// what: bit 0 = the pre-pass (scratch cleared, sentinels, in-place predictions moved to the shadow picture), bit 1 = the
// wavefronts.  The pre-pass needs the inter prediction only, not the residual stage: a caller that issues it on a stream of
// its own (vvcb200_intra_prepare) lets it run beside the residual kernels.
static int intra_impl(const VVCB200Planes *pic, const VVCB200Planes *d_residual, const VVCB200Planes *d_shadow,
                      const VVCB200IntraBlock *d_blocks, int n, const VVCB200IntraParams *params, void *d_scratch, void *stream,
                      int what)
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
    // the pre-pass and the staging copies move 8 bytes at a time
    for (int c = 0; c < 3; c++) {
        if (((uintptr_t)pic->d_plane[c] & 7) || (pic->stride[c] & 7) ||
            (d_residual && (((uintptr_t)d_residual->d_plane[c] & 7) || (d_residual->stride[c] & 7)))) {
            set_error("vvcb200_intra_picture: planes of pic and d_residual must be 8-byte aligned with strides that are multiples "
                      "of 8 bytes (plane %d)", c);
            return VVCB200_ERR_ARG;
        }
    }
    IntraArgs a;
    a.pic = make_view(pic);
    for (int c = 0; c < 3; c++) {
        a.res[c] = d_residual ? (const int16_t *)d_residual->d_plane[c] : nullptr;
        a.res_stride[c] = d_residual ? d_residual->stride[c] / 2 : 0;
    }
    for (int c = 0; c < 3; c++) {
        a.shadow[c] = d_shadow ? (uint16_t *)d_shadow->d_plane[c] : nullptr;
        a.shadow_stride[c] = d_shadow ? d_shadow->stride[c] / 2 : 0;
    }
    if (d_shadow)
        for (int c = 0; c < 3; c++)
            if (!d_shadow->d_plane[c] || d_shadow->width[c] < pic->width[c] || d_shadow->height[c] < pic->height[c] ||
                d_shadow->d_plane[c] == pic->d_plane[c] || (d_shadow->stride[c] & 7) || ((uintptr_t)d_shadow->d_plane[c] & 7)) {
                set_error("vvcb200_intra_picture: d_shadow must be a second picture of at least the size of pic, planes and "
                          "strides multiples of 8 bytes");
                return VVCB200_ERR_ARG;
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
    a.inplace_flag = (int *)d_scratch + 1;
    for (int c = 0; c < 3; c++) {
        a.marks[c] = (uint8_t *)d_scratch + l.off[c];
        a.mpitch[c] = l.pitch[c];
    }
    // phases: luma wavefront, chroma add-only batch, chroma wavefront (see VVCB200IntraParams)
    const int *d_counts = params->d_phase_counts;                          // the order was made on the device
    const int n_luma = d_counts ? 0 : params->n_luma, n_add = d_counts ? 0 : params->n_add_only;
    const bool phased = d_counts || n_luma > 0 || n_add > 0;
    if (n_luma < 0 || n_add < 0 || (long long)n_luma + n_add > n) {
        set_error("vvcb200_intra_picture: n_luma %d + n_add_only %d exceed the %d records", n_luma, n_add, n);
        return VVCB200_ERR_ARG;
    }
    if ((d_counts || (phased && n_add)) && !d_residual) {
        set_error("vvcb200_intra_picture: add-only records need the residual planes");
        return VVCB200_ERR_ARG;
    }
    a.skip_first = n_luma;
    a.skip_n = phased ? n_add : 0;
    a.phase_info = d_counts;
    a.phase = kPhaseMark;
    a.vpdu_scale = nullptr;
    a.vpdu_pitch = l.vpdu_pitch;
    cudaStream_t s = (cudaStream_t)stream;
    if (phased && params->lmcs_chroma_scale)
        a.vpdu_scale = (unsigned *)((uint8_t *)d_scratch + l.vpdu_off);
    if (what & 1) {
        // ticket counters and in-place flag; the marks too unless a shadow picture makes them unnecessary
        VVC_CUDA(cudaMemsetAsync(d_scratch, 0, d_shadow ? 256 : l.marks_total, s));
        if (a.vpdu_scale)
            VVC_CUDA(cudaMemsetAsync(a.vpdu_scale, 0, l.total - l.vpdu_off, s));
        intra_mark_kernel<<<(n + 128 / kMarkLanes - 1) / (128 / kMarkLanes), 128, 0, s>>>(a);
        VVC_LAUNCH_CHECK();
    }
    if (!(what & 2))
        return 0;
    // luma is final once the luma wavefront (or the only wavefront) has run: the event a caller may have asked for
    cudaEvent_t luma_done = (cudaEvent_t)params->luma_done_event;
    // Pictures with scattered intra units are bound by the number of records in flight and take every warp the SM holds;
    // pictures that are mostly intra are bound by their dependency chains, where more pollers only lengthen a step
    // (4K all-intra: 4.4 ms with 12 warps per SM, 4.8 ms with 16; random access: 2 % the other way).
    const long long areas = ((long long)pic->width[0] * pic->height[0] + 4095) / 4096;       // 64x64 luma areas
    const int per_sm = n > 30 * areas ? min(kCtasPerSm, 3) : kCtasPerSm;
    int ctas = min((n + kWarps - 1) / kWarps, sm_count() * per_sm);
#ifdef VVC_INTRA_PROFILE
    if (const char *e = getenv("VVC_INTRA_CTAS"))                         // development: occupancy experiments
        ctas = min(ctas, atoi(e));
#endif
    const size_t smem = sizeof(WarpScratch) * kWarps + sizeof(CtaArena);
    static bool attr_set[64];                                             // per device: allow > 48 KB of shared memory
    int dev = 0;
    VVC_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        VVC_CUDA(cudaFuncSetAttribute(intra_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VVC_CUDA(cudaFuncSetAttribute(intra_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VVC_CUDA(cudaFuncSetAttribute(intra_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        VVC_CUDA(cudaFuncSetAttribute(intra_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[dev] = true;
    }
    const bool marks = !d_shadow;
    const VVCB200IntraBlock *all = d_blocks;
    // one wavefront launch over records [first, first + count) with its own ticket counter
    auto wavefront = [&](int first, int count, int ticket_idx, bool scale, int phase) -> int {
        if (count <= 0)
            return 0;
        IntraArgs w = a;
        w.blocks = all + first;
        w.n = count;
        w.phase = phase;                                                  // (with d_counts the kernel finds its own range)
        w.ticket = (int *)d_scratch + ticket_idx;
        const int grid = min(ctas, (count + kWarps - 1) / kWarps);
        if (marks && scale)
            intra_kernel<true, true><<<grid, kWarps * 32, smem, s>>>(w);
        else if (marks)
            intra_kernel<true, false><<<grid, kWarps * 32, smem, s>>>(w);
        else if (scale)
            intra_kernel<false, true><<<grid, kWarps * 32, smem, s>>>(w);
        else
            intra_kernel<false, false><<<grid, kWarps * 32, smem, s>>>(w);
        VVC_LAUNCH_CHECK();
        return 0;
    };
    const bool scale = params->lmcs_chroma_scale != 0;
    if (!phased) {
        if (int r = wavefront(0, n, 0, scale, 0))
            return r;
        if (luma_done)
            VVC_CUDA(cudaEventRecord(luma_done, s));
        return 0;
    }
    if (d_counts) {
        // the ranges live on the device: every launch is sized for all n records and takes its own share
        if (int r = wavefront(0, n, 0, false, kPhaseLuma))
            return r;
        if (luma_done)
            VVC_CUDA(cudaEventRecord(luma_done, s));
        IntraArgs w = a;
        w.n = n;
        if (a.vpdu_scale) {
            w.phase = kPhaseAllChroma;
            vpdu_scale_kernel<<<(n + 7) / 8, 256, 0, s>>>(w);
            VVC_LAUNCH_CHECK();
        }
        w.phase = kPhaseAddOnly;
        intra_add_only_kernel<<<(n + 7) / 8, 256, 0, s>>>(w);
        VVC_LAUNCH_CHECK();
        return wavefront(0, n, 2, scale, kPhaseChroma);
    }
    if (int r = wavefront(0, n_luma, 0, false, 0))                        // luma records never carry the chroma scale
        return r;
    if (luma_done)
        VVC_CUDA(cudaEventRecord(luma_done, s));
    if (a.vpdu_scale && n > n_luma) {                                     // every chroma record's VPDU factor, once
        IntraArgs w = a;
        w.blocks = all + n_luma;
        w.n = n - n_luma;
        vpdu_scale_kernel<<<(w.n + 7) / 8, 256, 0, s>>>(w);
        VVC_LAUNCH_CHECK();
    }
    if (n_add) {
        IntraArgs w = a;
        w.blocks = all + n_luma;
        w.n = n_add;
        intra_add_only_kernel<<<(n_add + 7) / 8, 256, 0, s>>>(w);
        VVC_LAUNCH_CHECK();
    }
    return wavefront(n_luma + n_add, n - n_luma - n_add, 2, scale, 0);
}

extern "C" int vvcb200_intra_picture(const VVCB200Planes *pic, const VVCB200Planes *d_residual, const VVCB200Planes *d_shadow,
                                     const VVCB200IntraBlock *d_blocks, int n, const VVCB200IntraParams *params,
                                     void *d_scratch, void *stream)
{
    if (params && !n && params->luma_done_event)                          // nothing to do, but the caller waits for the event
        VVC_CUDA(cudaEventRecord((cudaEvent_t)params->luma_done_event, (cudaStream_t)stream));
    return intra_impl(pic, d_residual, d_shadow, d_blocks, n, params, d_scratch, stream, params && params->prepared ? 2 : 3);
}

extern "C" int vvcb200_intra_prepare(const VVCB200Planes *pic, const VVCB200Planes *d_residual, const VVCB200Planes *d_shadow,
                                     const VVCB200IntraBlock *d_blocks, int n, const VVCB200IntraParams *params,
                                     void *d_scratch, void *stream)
{
    return intra_impl(pic, d_residual, d_shadow, d_blocks, n, params, d_scratch, stream, 1);
}

extern "C" size_t vvcb200_intra_schedule_scratch_bytes(int n, int ctus_w, int ctus_h)
{
    if (n < 0 || ctus_w <= 0 || ctus_h <= 0)
        return 0;
    return sched_layout(n, ctus_w, ctus_h).total;
}

extern "C" int vvcb200_intra_schedule_device(const VVCB200IntraBlock *d_in, const uint32_t *d_ctu_first, int n, int ctus_w,
                                             int ctus_h, VVCB200IntraBlock *d_out, int32_t *d_phase_counts, void *d_scratch,
                                             void *stream)
{
    if (n < 0 || ctus_w <= 0 || ctus_h <= 0 || ctus_w * ctus_h > 65535 || !d_phase_counts ||
        (n && (!d_in || !d_ctu_first || !d_out || !d_scratch || d_in == d_out))) {
        set_error("vvcb200_intra_schedule_device: bad arguments");
        return VVCB200_ERR_ARG;
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (!n) {
        VVC_CUDA(cudaMemsetAsync(d_phase_counts, 0, 3 * sizeof(int32_t), s));
        return 0;
    }
    const SchedLayout l = sched_layout(n, ctus_w, ctus_h);
    uint8_t *base = (uint8_t *)d_scratch;
    uint32_t *rank = (uint32_t *)(base + l.rank);
    uint16_t *ctu_of = (uint16_t *)(base + l.ctu_of);
    int *len = (int *)(base + l.len), *diag = (int *)(base + l.diag);
    const int nctu = ctus_w * ctus_h;
    intra_rank_kernel<<<(nctu + 7) / 8, 256, 0, s>>>(d_in, d_ctu_first, nctu, rank, ctu_of, len);
    VVC_LAUNCH_CHECK();
    intra_base_kernel<<<1, 256, 0, s>>>(len, ctus_w, ctus_h, diag, d_phase_counts);
    VVC_LAUNCH_CHECK();
    intra_scatter_kernel<<<(n + 255) / 256, 256, 0, s>>>(d_in, n, rank, ctu_of, len, diag, d_phase_counts, ctus_w, ctus_h, d_out);
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
    const size_t need = pic->width[0] > 0 && pic->height[0] > 0 ? scratch_layout(pic->width[0], pic->height[0]).marks_total : 0;
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
    // One wavefront over all records (the per-VPDU table of the phased form would not fit the array); the callee clears the
    // ticket sector and the marks, which an earlier call may have left set.
    VVCB200IntraParams one = *params;
    one.n_luma = one.n_add_only = 0;
    one.d_phase_counts = nullptr;
    return vvcb200_intra_picture(pic, d_residual, nullptr, d_blocks, n, &one, ptr, stream);
}

#ifdef VVC_INTRA_PROFILE
extern "C" VVCB200_API int vvcb200_intra_profile(unsigned long long *out8, int reset)
{
    VVC_CUDA(cudaMemcpyFromSymbol(out8, vvcb200::g_prof, sizeof(unsigned long long) * 8));
    if (reset) {
        unsigned long long z[8] = { 0 };
        VVC_CUDA(cudaMemcpyToSymbol(vvcb200::g_prof, z, sizeof(z)));
    }
    return 0;
}
#endif
