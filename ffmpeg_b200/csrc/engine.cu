// engine.cu — Level B of the C ABI: the per-GPU picture engine (context, DPB planes, staging, stage order).
// Host-side C++ only; the kernels live in the stage files and are reached through their Level-A entries,
// so both levels run exactly the same device code.  Everything is enqueued on the context's stream; the
// caller's host buffers are read by asynchronous copies (pinned memory gives real overlap).
#include <algorithm>
#include <new>
#include <string.h>
#include <vector>
#include "common.cuh"

using namespace vvcb200;

namespace {

struct DevBuffer {
    void *p = nullptr;             // what the kernels read: the owned allocation, or a place in an arena's device mirror
    void *own = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes)
    {
        if (bytes <= cap) {
            p = own;
            return 0;
        }
        if (own)
            cudaFree(own);
        own = p = nullptr;
        cap = 0;
        const size_t want = bytes + bytes / 4 + 4096;
        VVC_CUDA(cudaMalloc(&own, want));
        cap = want;
        p = own;
        return 0;
    }
    void release()
    {
        if (own)
            cudaFree(own);
        own = p = nullptr;
        cap = 0;
    }
};

struct DevPicture {
    VVCB200Planes planes = {};
    void *base = nullptr;
};

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

}  // namespace

// Pictures in flight.  A lane is a compute stream with its working pictures; consecutive pictures alternate between the
// two lanes, so that the kernels of picture n + 1 overlap those of picture n wherever n + 1 does not read n: the intra
// stage is bound by dependency latency and leaves most issue slots of the GPU free for the neighbour's INTER / residual
// kernels.  Pictures that do depend on each other (through the DPB) are ordered with per-slot events.  A stage set is
// the device copy of one picture's host descriptor batches; four sets rotate (two per lane), so that the uploads of a
// lane's next picture (copy stream) overlap the kernels of its current one.
struct Lane {
    cudaStream_t stream = nullptr; // this lane's compute stream
    DevPicture scratch;            // SAO output / ALF input; before that, the intra stage's shadow picture
    DevPicture residual;           // int16 residual of intra / CIIP blocks, between stage RESIDUAL and stage INTRA (lazy)
    DevBuffer intra_marks;
    cudaEvent_t tail = nullptr;    // "everything this lane has queued so far", recorded when the other lane must get behind it
    // Two things inside a picture need not wait for the stage before them and go to a helper stream: the intra pre-pass
    // (needs the inter prediction, runs beside the residual kernels) and luma deblocking (needs final luma, runs beside the
    // chroma phases of the intra stage).  Both are joined back into the lane's stream before anything reads their output.
    cudaStream_t helper = nullptr;
    cudaEvent_t inter_done = nullptr, luma_done = nullptr, helper_done = nullptr;
    bool have_inter = false, luma_pending = false;   // of the lane's current picture
};
constexpr int kLanes = 2, kSets = 4;   // two descriptor sets per lane: the next picture of a lane uploads while this one runs
struct StageSet {
    Lane *lane = nullptr;          // sets[i] belongs to lanes[i % kLanes]
    DevBuffer levels, tbs, blocks, refined, prof, lut_fwd, lut_inv, db_tables, sao, alf, alf_params, alf_table, iblocks;
    DevBuffer iblocks_in, ctu_first, sched_scratch, phase_counts;   // intra records as emitted, their CTU ranges, the device scheduler's state
    cudaEvent_t uploaded = nullptr;    // last upload into this set (recorded on the upload stream)
    cudaEvent_t consumed = nullptr;    // last kernel that read this set (recorded on the compute stream)
    cudaEvent_t mv_ready = nullptr;        // INTER has written `refined` (compute stream)
    cudaEvent_t mv_copied = nullptr;       // `refined` has reached the host (download stream)
    bool in_use = false, mv_pending = false;
};

// A pinned host block handed out by vvcb200_ctx_arena_alloc with one device mirror per stage set: a batch array that lies
// in it is uploaded to the same offset of the current set's mirror, and the uploads of one stage call that lie next to
// each other travel as ONE copy.  (Measured: with a picture download running, every additional upload of a picture costs
// about 10 us of link time whatever its size — fifteen uploads per picture made the end-to-end loop 0.72 ms, one 0.52.)
struct Arena {
    uint8_t *host = nullptr;
    size_t bytes = 0;
    uint8_t *mirror[4] = { nullptr, nullptr, nullptr, nullptr };   // kSets
};
struct PendingCopy { int arena; size_t lo, hi; };

struct VVCB200Ctx {
    std::vector<Arena> arenas;
    std::vector<PendingCopy> pending;  // arena ranges the current stage call wants on the device
    int device = 0;
    int width = 0, height = 0, bit_depth = 10, hshift = 1, vshift = 1;
    cudaStream_t h2d = nullptr;        // descriptor uploads
    cudaStream_t d2h = nullptr;        // picture downloads
    std::vector<DevPicture> dpb;   // reference pictures + the picture under reconstruction
    std::vector<cudaEvent_t> downloaded;   // per DPB slot: last download of that slot (d2h stream)
    std::vector<char> download_pending;
    std::vector<cudaEvent_t> written;      // per DPB slot: last kernel / copy that wrote the slot (its writer's stream)
    std::vector<signed char> writer;       // per DPB slot: lane of that write, -1 if none yet
    std::vector<unsigned char> readers;    // per DPB slot: lanes that have used it as a reference since
    std::vector<cudaEvent_t> released;     // per DPB slot: a consumer's "done reading" event (vvcb200_pic_release), or null
    int cur = 0;
    Lane lanes[kLanes];
    StageSet sets[kSets];
    int set = 0;
    int validate = 1;              // 0: trust the emitter, 1: spot checks (class boundaries + every 64th record), 2: every record
    // The loop-filter passes are out of place and ping-pong between the DPB buffer and the lane's scratch picture:
    bool in_scratch = false;       // the current picture's newest samples are in its lane's scratch picture, not in dpb[cur]
    bool lmcs_pending = false;     // vvcb200_pic_lmcs was called; the mapping rides on the deblocking pass's load if one follows
};

static inline int lane_of(int set) { return set % kLanes; }
static inline cudaStream_t lane_stream(VVCB200Ctx *c) { return c->lanes[lane_of(c->set)].stream; }

namespace {

int alloc_picture(VVCB200Ctx *c, DevPicture *pic)
{
    size_t off[3], total = 0;
    for (int i = 0; i < 3; i++) {
        const int w = i ? c->width >> c->hshift : c->width;
        const int h = i ? c->height >> c->vshift : c->height;
        const int pitch = (w * 2 + 255) & ~255;          // 256-byte rows: every row start is 128-bit/TMA aligned
        pic->planes.width[i] = w;
        pic->planes.height[i] = h;
        pic->planes.stride[i] = pitch;
        off[i] = total;
        total += (size_t)pitch * h;
    }
    VVC_CUDA(cudaMalloc(&pic->base, total));
    VVC_CUDA(cudaMemsetAsync(pic->base, 0, total, lane_stream(c)));
    for (int i = 0; i < 3; i++)
        pic->planes.d_plane[i] = (uint8_t *)pic->base + off[i];
    return 0;
}

DevPicture *slot_of(VVCB200Ctx *c, int slot)
{
    if (slot < 0)
        return &c->dpb[c->cur];
    return slot < (int)c->dpb.size() ? &c->dpb[slot] : nullptr;
}

// Host -> device copy of a descriptor batch on the upload stream.  A buffer that has to grow is only reallocated after
// the kernels still reading its set have finished (cudaFree would wait for them anyway, but on the whole device).
// host range inside an arena: remember it for the merged copy, hand out its place in the current set's mirror
static void *arena_place(VVCB200Ctx *c, const void *host, size_t bytes)
{
    const uint8_t *h = (const uint8_t *)host;
    for (size_t i = 0; i < c->arenas.size(); i++) {
        const Arena &a = c->arenas[i];
        if (h >= a.host && h + bytes <= a.host + a.bytes) {
            const size_t lo = (size_t)(h - a.host);
            if (bytes)
                c->pending.push_back({ (int)i, lo, lo + bytes });
            return a.mirror[c->set] + lo;
        }
    }
    return nullptr;
}

int upload(VVCB200Ctx *c, DevBuffer &b, const void *host, size_t bytes)
{
    if (void *d = arena_place(c, host, bytes)) {
        b.p = d;
        return 0;
    }
    if (bytes > b.cap) {
        VVC_CUDA(cudaStreamSynchronize(lane_stream(c)));
        VVC_CUDA(cudaStreamSynchronize(c->h2d));
    }
    if (int r = b.reserve(bytes))
        return r;
    VVC_CUDA(cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, c->h2d));
    return 0;
}

// the compute stream may read the current set once everything uploaded so far has arrived
int uploads_visible(VVCB200Ctx *c)
{
    StageSet &s = c->sets[c->set];
    if (!c->pending.empty()) {
        // neighbouring ranges as one copy; a gap of up to 64 KiB goes along (its bytes land in the mirror and mean nothing)
        std::sort(c->pending.begin(), c->pending.end(), [](const PendingCopy &x, const PendingCopy &y) {
            return x.arena != y.arena ? x.arena < y.arena : x.lo < y.lo;
        });
        size_t i = 0;
        while (i < c->pending.size()) {
            PendingCopy m = c->pending[i++];
            while (i < c->pending.size() && c->pending[i].arena == m.arena && c->pending[i].lo <= m.hi + 65536) {
                m.hi = std::max(m.hi, c->pending[i].hi);
                i++;
            }
            const Arena &a = c->arenas[m.arena];
            const cudaError_t e = cudaMemcpyAsync(a.mirror[c->set] + m.lo, a.host + m.lo, m.hi - m.lo, cudaMemcpyHostToDevice, c->h2d);
            if (e != cudaSuccess) {
                c->pending.clear();
                VVC_CUDA(e);
            }
        }
        c->pending.clear();
    }
    VVC_CUDA(cudaEventRecord(s.uploaded, c->h2d));
    VVC_CUDA(cudaStreamWaitEvent(s.lane->stream, s.uploaded, 0));
    return 0;
}

// the current set is being read by what its compute stream holds now, and the current picture is written up to here
int set_consumed(VVCB200Ctx *c)
{
    StageSet &s = c->sets[c->set];
    VVC_CUDA(cudaEventRecord(s.consumed, s.lane->stream));
    VVC_CUDA(cudaEventRecord(c->written[c->cur], s.lane->stream));
    c->writer[c->cur] = (signed char)lane_of(c->set);
    s.in_use = true;
    return 0;
}

// The current lane is about to overwrite DPB slot `slot`: wait for the other lane's write of it and for the other lane's
// pictures that use it as a reference — everything that lane has queued so far, marked with an event recorded now.
int order_before_write(VVCB200Ctx *c, int slot)
{
    const int other = lane_of(c->set) ^ 1;
    if ((c->readers[slot] >> other) & 1) {
        VVC_CUDA(cudaEventRecord(c->lanes[other].tail, c->lanes[other].stream));
        VVC_CUDA(cudaStreamWaitEvent(lane_stream(c), c->lanes[other].tail, 0));
    } else if (c->writer[slot] == other) {
        VVC_CUDA(cudaStreamWaitEvent(lane_stream(c), c->written[slot], 0));
    }
    if (c->download_pending[slot]) {      // the slot's previous content is still on its way to the host
        VVC_CUDA(cudaStreamWaitEvent(lane_stream(c), c->downloaded[slot], 0));
        c->download_pending[slot] = 0;
    }
    if (c->released[slot]) {              // an exported picture whose consumer may still be reading
        VVC_CUDA(cudaStreamWaitEvent(lane_stream(c), c->released[slot], 0));
        c->released[slot] = nullptr;
    }
    c->readers[slot] = 0;
    return 0;
}

// The current lane is about to read DPB slot `slot` as a reference picture.
int order_before_read(VVCB200Ctx *c, int slot)
{
    if (c->writer[slot] == (lane_of(c->set) ^ 1))
        VVC_CUDA(cudaStreamWaitEvent(lane_stream(c), c->written[slot], 0));
    c->readers[slot] |= (unsigned char)(1u << lane_of(c->set));
    return 0;
}

int ensure_residual(VVCB200Ctx *c, Lane *lane)
{
    if (lane->residual.base)
        return 0;
    return alloc_picture(c, &lane->residual);
}

// Bring the current picture to rest in its DPB slot: run a mapping that no deblocking pass picked up, and if the newest
// samples sit in the lane's scratch picture, exchange the two buffers (both belong to this lane's stream order).
int settle_picture(VVCB200Ctx *c)
{
    StageSet &S = c->sets[c->set];
    if (c->lmcs_pending) {
        c->lmcs_pending = false;
        VVCB200Planes *pl = c->in_scratch ? &S.lane->scratch.planes : &c->dpb[c->cur].planes;
        if (int r = vvcb200_lmcs_picture(pl, (const uint16_t *)S.lut_inv.p, c->bit_depth, S.lane->stream))
            return r;
        if (int r = set_consumed(c))
            return r;
    }
    if (c->in_scratch) {
        c->in_scratch = false;
        DevPicture t = c->dpb[c->cur];
        c->dpb[c->cur] = S.lane->scratch;
        S.lane->scratch = t;
    }
    return 0;
}

}  // namespace

extern "C" {

int vvcb200_ctx_create(VVCB200Ctx **out, int device, int width, int height, int chroma_format_idc, int bit_depth,
                       int dpb_size)
{
    if (!out || width <= 0 || height <= 0 || width > 16384 || height > 16384 || (width & 7) || (height & 7)) {
        set_error("vvcb200_ctx_create: bad dimensions %dx%d (multiples of 8 up to 16384)", width, height);
        return VVCB200_ERR_ARG;
    }
    if (bit_depth != 10 && bit_depth != 12) {
        set_error("vvcb200_ctx_create: unsupported bit depth %d (10 or 12; 8-bit pictures are uint8 planes in the reference, not built)", bit_depth);
        return VVCB200_ERR_ARG;
    }
    if (dpb_size < 1 || dpb_size > 17) {
        set_error("vvcb200_ctx_create: dpb_size %d out of range (1..17)", dpb_size);
        return VVCB200_ERR_TOO_BIG;
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
    cudaError_t e = cudaStreamCreateWithFlags(&c->h2d, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->d2h, cudaStreamNonBlocking);
    for (int i = 0; i < kLanes && e == cudaSuccess; i++) {
        e = cudaStreamCreateWithFlags(&c->lanes[i].stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->lanes[i].tail, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->lanes[i].helper, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->lanes[i].inter_done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->lanes[i].luma_done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->lanes[i].helper_done, cudaEventDisableTiming);
    }
    for (int i = 0; i < kSets && e == cudaSuccess; i++) {
        c->sets[i].lane = &c->lanes[lane_of(i)];
        e = cudaEventCreateWithFlags(&c->sets[i].uploaded, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->sets[i].consumed, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->sets[i].mv_ready, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->sets[i].mv_copied, cudaEventDisableTiming);
    }
    c->downloaded.assign(dpb_size, nullptr);
    c->written.assign(dpb_size, nullptr);
    c->download_pending.assign(dpb_size, 0);
    c->writer.assign(dpb_size, -1);
    c->readers.assign(dpb_size, 0);
    c->released.assign(dpb_size, nullptr);
    for (int i = 0; i < dpb_size && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&c->downloaded[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->written[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        vvcb200_ctx_destroy(c);
        return cuda_fail(e, "stream / event creation", __FILE__, __LINE__);
    }
    c->dpb.resize(dpb_size);
    int r = alloc_picture(c, &c->lanes[0].scratch);
    if (!r)
        r = alloc_picture(c, &c->lanes[1].scratch);
    for (int i = 0; !r && i < dpb_size; i++)
        r = alloc_picture(c, &c->dpb[i]);
    // Working memory for the descriptor batches of a typical picture of this size, so that the steady state never meets the
    // allocator (an allocation serialises against everything in flight on the device): dense levels, TBs of 32 samples,
    // prediction and intra records of 32 luma samples on average; a picture that needs more grows its buffer once.
    const size_t wh = (size_t)width * height;
    const int ctb32 = ((width + 31) >> 5) * ((height + 31) >> 5);
    for (int i = 0; !r && i < kSets; i++) {
        StageSet &ss = c->sets[i];
        struct { DevBuffer *b; size_t bytes; } want[] = {
            { &ss.levels, 3 * wh }, { &ss.tbs, wh * 3 / 64 * sizeof(VVCB200TransformBlock) },
            { &ss.blocks, wh / 32 * sizeof(VVCB200PredBlock) }, { &ss.refined, wh / 32 * 16 },
            { &ss.prof, wh / 256 * sizeof(VVCB200ProfTable) / 4 }, { &ss.lut_fwd, sizeof(uint16_t) << bit_depth },
            { &ss.lut_inv, sizeof(uint16_t) << bit_depth }, { &ss.db_tables, (wh / 16 + 64) * 9 + (size_t)ctb32 * 6 + 64 },
            { &ss.sao, (size_t)ctb32 * sizeof(VVCB200SaoCtb) }, { &ss.alf, (size_t)ctb32 * sizeof(VVCB200AlfCtb) },
            { &ss.alf_params, sizeof(VVCB200AlfParams) }, { &ss.alf_table, vvcb200_sao_alf_scratch_bytes() },
            { &ss.iblocks, wh / 48 * sizeof(VVCB200IntraBlock) }, { &ss.iblocks_in, wh / 48 * sizeof(VVCB200IntraBlock) },
            { &ss.ctu_first, ((size_t)ctb32 + 1) * 4 },
            { &ss.sched_scratch, vvcb200_intra_schedule_scratch_bytes((int)(wh / 48), (width + 31) >> 5, (height + 31) >> 5) },
            { &ss.phase_counts, 256 },
        };
        for (auto &w : want)
            if (!r)
                r = w.b->reserve(w.bytes);
    }
    for (int i = 0; !r && i < kLanes; i++) {
        r = c->lanes[i].intra_marks.reserve(vvcb200_intra_scratch_bytes(width, height));
        if (!r)
            r = alloc_picture(c, &c->lanes[i].residual);
    }
    if (!r && cudaDeviceSynchronize() != cudaSuccess)                   // the pictures are cleared before either lane runs
        r = VVCB200_ERR_CUDA;
    if (r) {
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
    for (cudaStream_t st : { c->lanes[0].stream, c->lanes[1].stream, c->h2d, c->d2h })
        if (st)
            cudaStreamSynchronize(st);
    for (StageSet &ss : c->sets) {
        DevBuffer *bufs[] = { &ss.levels, &ss.tbs, &ss.blocks, &ss.refined, &ss.prof, &ss.lut_fwd, &ss.lut_inv, &ss.db_tables,
                              &ss.sao, &ss.alf, &ss.alf_params, &ss.alf_table, &ss.iblocks, &ss.iblocks_in, &ss.ctu_first,
                              &ss.sched_scratch, &ss.phase_counts };
        for (DevBuffer *b : bufs)
            b->release();
        if (ss.uploaded) cudaEventDestroy(ss.uploaded);
        if (ss.consumed) cudaEventDestroy(ss.consumed);
        if (ss.mv_ready) cudaEventDestroy(ss.mv_ready);
        if (ss.mv_copied) cudaEventDestroy(ss.mv_copied);
    }
    for (Lane &l : c->lanes) {
        if (l.stream) fork_streams_release(l.stream);
        if (l.tail) cudaEventDestroy(l.tail);
        if (l.helper) { cudaStreamSynchronize(l.helper); cudaStreamDestroy(l.helper); }
        for (cudaEvent_t ev : { l.inter_done, l.luma_done, l.helper_done })
            if (ev) cudaEventDestroy(ev);
        l.intra_marks.release();
        if (l.scratch.base) cudaFree(l.scratch.base);
        if (l.residual.base) cudaFree(l.residual.base);
    }
    for (Arena &a : c->arenas) {
        for (uint8_t *m : a.mirror)
            if (m)
                cudaFree(m);
        if (a.host)
            cudaFreeHost(a.host);
    }
    for (cudaEvent_t ev : c->downloaded)
        if (ev)
            cudaEventDestroy(ev);
    for (cudaEvent_t ev : c->written)
        if (ev)
            cudaEventDestroy(ev);
    for (DevPicture &p : c->dpb)
        if (p.base)
            cudaFree(p.base);
    for (cudaStream_t st : { c->lanes[0].stream, c->lanes[1].stream, c->h2d, c->d2h })
        if (st)
            cudaStreamDestroy(st);
    delete c;
}

int vvcb200_ctx_sync(VVCB200Ctx *c)
{
    if (!c)
        return VVCB200_ERR_ARG;
    DeviceGuard g(c->device);
    if (int r = settle_picture(c))
        return r;
    VVC_CUDA(cudaStreamSynchronize(c->h2d));
    VVC_CUDA(cudaStreamSynchronize(c->lanes[0].stream));
    VVC_CUDA(cudaStreamSynchronize(c->lanes[1].stream));
    VVC_CUDA(cudaStreamSynchronize(c->d2h));
    for (char &p : c->download_pending)
        p = 0;
    for (StageSet &ss : c->sets)
        ss.in_use = ss.mv_pending = false;
    for (size_t i = 0; i < c->writer.size(); i++) {                       // nothing is in flight: no ordering left to do
        c->writer[i] = -1;
        c->readers[i] = 0;
    }
    // (c->released stays: the consumer's event belongs to another stream)
    return 0;
}

int vvcb200_ctx_set_validation(VVCB200Ctx *c, int level)
{
    if (!c || level < 0 || level > 2) {
        set_error("vvcb200_ctx_set_validation: bad argument");
        return VVCB200_ERR_ARG;
    }
    c->validate = level;
    return 0;
}

void *vvcb200_ctx_stream(VVCB200Ctx *c) { return c ? (void *)lane_stream(c) : nullptr; }

int vvcb200_ctx_planes(VVCB200Ctx *c, int slot, VVCB200Planes *out)
{
    if (c) {
        DeviceGuard g(c->device);
        if (int r = settle_picture(c))
            return r;
    }
    DevPicture *p = c ? slot_of(c, slot) : nullptr;
    if (!p || !out) {
        set_error("vvcb200_ctx_planes: bad context or slot %d", slot);
        return VVCB200_ERR_ARG;
    }
    *out = p->planes;
    return 0;
}

void *vvcb200_ctx_arena_alloc(VVCB200Ctx *c, size_t bytes)
{
    if (!c || !bytes) {
        set_error("vvcb200_ctx_arena_alloc: bad argument");
        return nullptr;
    }
    DeviceGuard g(c->device);
    Arena a;
    a.bytes = (bytes + 255) & ~(size_t)255;
    bool ok = cudaHostAlloc((void **)&a.host, a.bytes, cudaHostAllocDefault) == cudaSuccess;
    for (int i = 0; ok && i < kSets; i++)
        ok = cudaMalloc((void **)&a.mirror[i], a.bytes) == cudaSuccess;
    if (!ok) {
        set_error("vvcb200_ctx_arena_alloc: %zu bytes of pinned host memory + %d device mirrors: %s", a.bytes, kSets,
                  cudaGetErrorString(cudaGetLastError()));
        for (uint8_t *m : a.mirror)
            if (m)
                cudaFree(m);
        if (a.host)
            cudaFreeHost(a.host);
        return nullptr;
    }
    c->arenas.push_back(a);
    return a.host;
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

int vvcb200_pic_begin(VVCB200Ctx *c, int slot)
{
    if (!c || slot < 0 || slot >= (int)c->dpb.size()) {
        set_error("vvcb200_pic_begin: bad context or slot %d", slot);
        return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    if (int r = settle_picture(c))        // the picture before this one
        return r;
    c->cur = slot;
    c->set = (c->set + 1) % kSets;
    StageSet &s = c->sets[c->set];
    s.lane->have_inter = s.lane->luma_pending = false;
    if (s.in_use)                         // uploads of this picture must not overtake the kernels of the picture before last
        VVC_CUDA(cudaStreamWaitEvent(c->h2d, s.consumed, 0));
    return order_before_write(c, slot);
}

int vvcb200_pic_upload(VVCB200Ctx *c, int slot, const uint16_t *const planes[3], const int32_t strides[3])
{
    DevPicture *p = c ? slot_of(c, slot) : nullptr;
    if (!p || !planes || !strides) {
        set_error("vvcb200_pic_upload: bad argument");
        return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    const int idx = slot < 0 ? c->cur : slot;
    if (idx == c->cur) {
        if (int r = settle_picture(c))
            return r;
        p = &c->dpb[idx];
    } else if (int r = order_before_write(c, idx)) {
        return r;
    }
    for (int i = 0; i < 3; i++)
        VVC_CUDA(cudaMemcpy2DAsync(p->planes.d_plane[i], p->planes.stride[i], planes[i], strides[i],
                                   (size_t)p->planes.width[i] * 2, p->planes.height[i], cudaMemcpyHostToDevice, lane_stream(c)));
    VVC_CUDA(cudaEventRecord(c->written[idx], lane_stream(c)));
    c->writer[idx] = (signed char)lane_of(c->set);
    return 0;
}

int vvcb200_pic_download(VVCB200Ctx *c, int slot, uint16_t *const planes[3], const int32_t strides[3])
{
    DevPicture *p = c ? slot_of(c, slot) : nullptr;
    if (!p || !planes || !strides) {
        set_error("vvcb200_pic_download: bad argument");
        return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    const int idx = slot < 0 ? c->cur : slot;
    if (idx == c->cur) {
        if (int r = settle_picture(c))
            return r;
        p = &c->dpb[idx];
    }
    if (c->writer[idx] >= 0)              // whatever wrote the slot last, in whichever lane
        VVC_CUDA(cudaStreamWaitEvent(c->d2h, c->written[idx], 0));
    // a frame whose three planes lie back to back in host memory exactly as they do on the device leaves as one copy
    bool one = true;
    for (int i = 0; i < 3; i++) {
        const size_t row = (size_t)p->planes.width[i] * 2, plane = row * p->planes.height[i];
        one = one && (size_t)p->planes.stride[i] == row && (size_t)strides[i] == row;
        if (i < 2)
            one = one && (uint8_t *)p->planes.d_plane[i + 1] == (uint8_t *)p->planes.d_plane[i] + plane &&
                  (uint8_t *)planes[i + 1] == (uint8_t *)planes[i] + plane;
    }
    if (one) {
        size_t total = 0;
        for (int i = 0; i < 3; i++)
            total += (size_t)p->planes.width[i] * 2 * p->planes.height[i];
        VVC_CUDA(cudaMemcpyAsync(planes[0], p->planes.d_plane[0], total, cudaMemcpyDeviceToHost, c->d2h));
    } else {
        for (int i = 0; i < 3; i++)
            VVC_CUDA(cudaMemcpy2DAsync(planes[i], strides[i], p->planes.d_plane[i], p->planes.stride[i],
                                       (size_t)p->planes.width[i] * 2, p->planes.height[i], cudaMemcpyDeviceToHost, c->d2h));
    }
    VVC_CUDA(cudaEventRecord(c->downloaded[idx], c->d2h));
    c->download_pending[idx] = 1;
    return 0;
}

static int window_of(VVCB200Ctx *c, const VVCB200Window *w, VVCB200Window &out, const char *who)
{
    out = w ? *w : VVCB200Window{ 0, 0, 0, 0 };
    if (out.crop_left < 0 || out.crop_right < 0 || out.crop_top < 0 || out.crop_bottom < 0 ||
        ((out.crop_left | out.crop_right | out.crop_top | out.crop_bottom) & 1) ||
        out.crop_left + out.crop_right >= c->width || out.crop_top + out.crop_bottom >= c->height) {
        set_error("%s: bad window (crop %d %d %d %d of %dx%d; even luma values)", who, out.crop_left, out.crop_right,
                  out.crop_top, out.crop_bottom, c->width, c->height);
        return VVCB200_ERR_ARG;
    }
    return 0;
}

static void window_planes(VVCB200Ctx *c, const VVCB200Planes &pl, const VVCB200Window &w, VVCB200Planes &out)
{
    for (int i = 0; i < 3; i++) {
        const int hs = i ? c->hshift : 0, vs = i ? c->vshift : 0;
        out.d_plane[i] = (uint8_t *)pl.d_plane[i] + (size_t)(w.crop_top >> vs) * pl.stride[i] + (size_t)(w.crop_left >> hs) * 2;
        out.stride[i] = pl.stride[i];
        out.width[i] = pl.width[i] - ((w.crop_left + w.crop_right) >> hs);
        out.height[i] = pl.height[i] - ((w.crop_top + w.crop_bottom) >> vs);
    }
}

int vvcb200_pic_export(VVCB200Ctx *c, int slot, const VVCB200Window *window, VVCB200Export *out)
{
    if (!c || !out || slot >= (int)c->dpb.size()) {
        set_error("vvcb200_pic_export: bad argument");
        return VVCB200_ERR_ARG;
    }
    VVCB200Window w;
    if (int r = window_of(c, window, w, "vvcb200_pic_export"))
        return r;
    DeviceGuard g(c->device);
    const int idx = slot < 0 ? c->cur : slot;
    if (idx == c->cur)
        if (int r = settle_picture(c))
            return r;
    window_planes(c, c->dpb[idx].planes, w, out->planes);
    if (c->writer[idx] < 0)               // nothing in flight wrote it: an event that is complete already
        VVC_CUDA(cudaEventRecord(c->written[idx], c->lanes[0].stream));
    out->ready = (void *)c->written[idx];
    return 0;
}

int vvcb200_pic_release(VVCB200Ctx *c, int slot, void *consumer_done)
{
    if (!c || slot >= (int)c->dpb.size()) {
        set_error("vvcb200_pic_release: bad argument");
        return VVCB200_ERR_ARG;
    }
    c->released[slot < 0 ? c->cur : slot] = (cudaEvent_t)consumer_done;
    return 0;
}

int vvcb200_pic_download_window(VVCB200Ctx *c, int slot, const VVCB200Window *window, uint16_t *const planes[3],
                                const int32_t strides[3])
{
    if (!c || !planes || !strides || slot >= (int)c->dpb.size()) {
        set_error("vvcb200_pic_download_window: bad argument");
        return VVCB200_ERR_ARG;
    }
    VVCB200Window w;
    if (int r = window_of(c, window, w, "vvcb200_pic_download_window"))
        return r;
    DeviceGuard g(c->device);
    const int idx = slot < 0 ? c->cur : slot;
    if (idx == c->cur)
        if (int r = settle_picture(c))
            return r;
    VVCB200Planes pl;
    window_planes(c, c->dpb[idx].planes, w, pl);
    if (c->writer[idx] >= 0)
        VVC_CUDA(cudaStreamWaitEvent(c->d2h, c->written[idx], 0));
    for (int i = 0; i < 3; i++)
        VVC_CUDA(cudaMemcpy2DAsync(planes[i], strides[i], pl.d_plane[i], pl.stride[i], (size_t)pl.width[i] * 2, pl.height[i],
                                   cudaMemcpyDeviceToHost, c->d2h));
    VVC_CUDA(cudaEventRecord(c->downloaded[idx], c->d2h));
    c->download_pending[idx] = 1;
    return 0;
}

int vvcb200_pic_inter(VVCB200Ctx *c, const VVCB200PredBlock *blocks, int n, const VVCB200InterParams *params,
                      const uint16_t *lmcs_fwd_lut, int32_t *refined_mv, const VVCB200ProfTable *prof, int n_prof)
{
    if (!c || !params || n < 0 || (n && !blocks)) {
        set_error("vvcb200_pic_inter: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (n == 0)
        return 0;
    // the batch must come in width order (fatal to get wrong); how thoroughly that is checked is the caller's choice
    const int b16 = params->n_w16, b8 = params->n_w16 + params->n_w8;
    auto check_block = [&](int i) {
        if (i < 0 || i >= n)
            return true;
        const int want = i < b16 ? 16 : i < b8 ? 8 : 4;
        if (blocks[i].w == want)
            return true;
        set_error("vvcb200_pic_inter: block %d has width %d where the width order (n_w16 %d, n_w8 %d) wants %d",
                  i, blocks[i].w, params->n_w16, params->n_w8, want);
        return false;
    };
    if (c->validate) {
        for (int i = 0; i < n; i += c->validate == 1 ? 64 : 1)
            if (!check_block(i))
                return VVCB200_ERR_ARG;
        for (int i : { b16 - 1, b16, b8 - 1, b8, n - 1 })               // both sides of every class boundary
            if (!check_block(i))
                return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    if (int r = settle_picture(c))
        return r;
    StageSet &S = c->sets[c->set];
    if (int r = upload(c, S.blocks, blocks, (size_t)n * sizeof(*blocks)))
        return r;
    const uint16_t *d_lut = nullptr;
    if (lmcs_fwd_lut) {
        if (int r = upload(c, S.lut_fwd, lmcs_fwd_lut, sizeof(uint16_t) << c->bit_depth))
            return r;
        d_lut = (const uint16_t *)S.lut_fwd.p;
    }
    VVCB200Planes refs[16];
    const int n_refs = (int)c->dpb.size() < 16 ? (int)c->dpb.size() : 16;
    for (int i = 0; i < n_refs; i++)
        refs[i] = c->dpb[i].planes;
    for (int l = 0; l < 2; l++)
        for (int i = 0; i < 16; i++)
            if (params->ref_slot[l][i] >= n_refs) {
                set_error("vvcb200_pic_inter: ref_slot[%d][%d] = %d is not a DPB slot of this context", l, i, params->ref_slot[l][i]);
                return VVCB200_ERR_ARG;
            }
    for (int l = 0; l < 2; l++) {                 // reference pictures the other lane is still writing
        const int active = params->num_ref_idx_active[l] > 0 && params->num_ref_idx_active[l] <= 16 ? params->num_ref_idx_active[l] : 16;
        for (int i = 0; i < active; i++)
            if (params->ref_slot[l][i] != c->cur)
                if (int r = order_before_read(c, params->ref_slot[l][i]))
                    return r;
    }
    int32_t *d_mv = nullptr;
    if (refined_mv) {
        if ((size_t)n * 4 * sizeof(int32_t) > S.refined.cap)
            VVC_CUDA(cudaStreamSynchronize(S.lane->stream));
        if (int r = S.refined.reserve((size_t)n * 4 * sizeof(int32_t)))
            return r;
        d_mv = (int32_t *)S.refined.p;
        if (S.mv_pending) {               // this set's previous vectors must have left before the kernels overwrite them
            VVC_CUDA(cudaStreamWaitEvent(S.lane->stream, S.mv_copied, 0));
            S.mv_pending = false;
        }
    }
    const VVCB200ProfTable *d_prof = nullptr;
    if (prof && n_prof > 0) {
        if (int r = upload(c, S.prof, prof, (size_t)n_prof * sizeof(*prof)))
            return r;
        d_prof = (const VVCB200ProfTable *)S.prof.p;
    }
    if (int r = uploads_visible(c))
        return r;
    if (int r = vvcb200_inter_batch(&c->dpb[c->cur].planes, refs, n_refs, (const VVCB200PredBlock *)S.blocks.p, n, params,
                                    d_lut, nullptr, d_mv, d_prof, S.lane->stream))
        return r;
    VVC_CUDA(cudaEventRecord(S.lane->inter_done, S.lane->stream));
    S.lane->have_inter = true;
    if (int r = set_consumed(c))
        return r;
    if (refined_mv) {   // DMVR feedback for the host's MV derivation of later pictures (vvc_inter.c:750-762), on the
                        // download stream so that the lane goes straight on with the residual stage
        VVC_CUDA(cudaEventRecord(S.mv_ready, S.lane->stream));
        VVC_CUDA(cudaStreamWaitEvent(c->d2h, S.mv_ready, 0));
        if (params->dmvr_first) {
            // the batch is in the recommended order: the records DMVR may move are a prefix of the width-16 and of the width-8
            // class; every other entry would only tell the host the vectors it sent
            const int start[2] = { 0, b16 }, end[2] = { b16, b8 };
            for (int k = 0; k < 2; k++) {
                int m = start[k];
                while (m < end[k] && (blocks[m].flags & VVCB200_PB_DMVR))
                    m++;
                if (m > start[k])
                    VVC_CUDA(cudaMemcpyAsync(refined_mv + (size_t)start[k] * 4, d_mv + (size_t)start[k] * 4,
                                             (size_t)(m - start[k]) * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, c->d2h));
            }
        } else {
            VVC_CUDA(cudaMemcpyAsync(refined_mv, d_mv, (size_t)n * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, c->d2h));
        }
        VVC_CUDA(cudaEventRecord(S.mv_copied, c->d2h));
        S.mv_pending = true;
    }
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
    {   // the batch must come in class order (see vvcb200_residual_class)
        int bound[5], acc = 0;
        for (int k = 0; k < 4; k++) {
            acc += params->n_class[k];
            bound[k] = acc;
        }
        bound[4] = n;
        auto check_tb = [&](int i) {
            if (i < 0 || i >= n)
                return true;
            int k = 0;
            while (k < 4 && i >= bound[k])
                k++;
            if (vvcb200_residual_class(&tbs[i]) == k)
                return true;
            set_error("vvcb200_pic_residual: block %d is of class %d where the class order (n_class %d %d %d %d) wants %d",
                      i, vvcb200_residual_class(&tbs[i]), params->n_class[0], params->n_class[1], params->n_class[2],
                      params->n_class[3], k);
            return false;
        };
        if (c->validate) {
            for (int i = 0; i < n; i += c->validate == 1 ? 64 : 1)
                if (!check_tb(i))
                    return VVCB200_ERR_ARG;
            for (int k = 0; k < 4; k++)                                   // both sides of every class boundary
                if (!check_tb(bound[k] - 1) || !check_tb(bound[k]))
                    return VVCB200_ERR_ARG;
            if (!check_tb(n - 1))
                return VVCB200_ERR_ARG;
        }
    }
    DeviceGuard g(c->device);
    if (int r = settle_picture(c))
        return r;
    StageSet &S = c->sets[c->set];
    if (int r = upload(c, S.levels, levels, n_levels * sizeof(int16_t)))
        return r;
    if (int r = upload(c, S.tbs, tbs, (size_t)n * sizeof(VVCB200TransformBlock)))
        return r;
    const VVCB200Planes *res = nullptr;
    for (int i = 0; i < n && !res; i++)
        if (tbs[i].flags & VVCB200_TB_TO_RESIDUAL) {
            if (!S.lane->residual.base)
                if (int r = alloc_picture(c, &S.lane->residual))
                    return r;
            res = &S.lane->residual.planes;
        }
    if (int r = uploads_visible(c))
        return r;
    if (int r = vvcb200_residual_batch(&c->dpb[c->cur].planes, res, (const int16_t *)S.levels.p,
                                       (const VVCB200TransformBlock *)S.tbs.p, n, params, S.lane->stream))
        return r;
    return set_consumed(c);
}

int vvcb200_pic_intra(VVCB200Ctx *c, const VVCB200IntraBlock *blocks, const uint32_t *ctu_first, int ctus_w, int ctus_h,
                      const VVCB200IntraParams *params)
{
    if (!c || !params || !ctu_first || ctus_w <= 0 || ctus_h <= 0) {
        set_error("vvcb200_pic_intra: bad argument");
        return VVCB200_ERR_ARG;
    }
    const int shift = params->ctb_log2_size;
    if (shift < 5 || shift > 7 || ctus_w != (c->width + (1 << shift) - 1) >> shift ||
        ctus_h != (c->height + (1 << shift) - 1) >> shift) {
        set_error("vvcb200_pic_intra: %dx%d CTUs of 2^%d do not tile the context's %dx%d picture", ctus_w, ctus_h, shift,
                  c->width, c->height);
        return VVCB200_ERR_ARG;
    }
    const int n = (int)ctu_first[ctus_w * ctus_h];
    if (n == 0)
        return 0;
    if (!blocks || ctu_first[0] != 0) {
        set_error("vvcb200_pic_intra: bad record ranges");
        return VVCB200_ERR_ARG;
    }
    StageSet &S = c->sets[c->set];
    const bool have_res = S.lane->residual.base != nullptr;                     // residual planes of this lane exist
    for (int i = 0; i < n && c->validate; i = (i == n - 1 || c->validate == 2) ? i + 1 : (i + 64 < n ? i + 64 : n - 1)) {
        const VVCB200IntraBlock &b = blocks[i];
        if ((b.flags & VVCB200_IB_RESIDUAL) && !have_res) {
            set_error("vvcb200_pic_intra: records add a residual but vvcb200_pic_residual stored none for this picture");
            return VVCB200_ERR_ARG;
        }
        const int pw = b.c_idx ? c->width >> c->hshift : c->width, ph = b.c_idx ? c->height >> c->vshift : c->height;
        if (b.c_idx > 2 || !b.w || !b.h || b.w > 64 || b.h > 64 || (b.w & (b.w - 1)) || (b.h & (b.h - 1)) ||
            b.x0 + b.w > pw || b.y0 + b.h > ph || (b.x0 & 3) || (b.y0 & (b.c_idx ? 1 : 3)) || b.mode > 83 || b.mode < -14 ||
            (b.mode >= 81 && b.c_idx != 1)) {
            set_error("vvcb200_pic_intra: record %d (plane %d, %dx%d at %d,%d, mode %d) is malformed or outside the picture",
                      i, b.c_idx, b.w, b.h, b.x0, b.y0, b.mode);
            return VVCB200_ERR_ARG;
        }
    }
    DeviceGuard g(c->device);
    if (int r = settle_picture(c))
        return r;
    // The records go up as the parser emitted them; the wavefront order is made on the device (three small kernels), so the
    // calling thread spends no time per record.
    const int nctu = ctus_w * ctus_h;
    if (int r = upload(c, S.iblocks_in, blocks, (size_t)n * sizeof(VVCB200IntraBlock)))
        return r;
    if (int r = upload(c, S.ctu_first, ctu_first, (size_t)(nctu + 1) * sizeof(uint32_t)))
        return r;
    const size_t need[3] = { (size_t)n * sizeof(VVCB200IntraBlock), vvcb200_intra_schedule_scratch_bytes(n, ctus_w, ctus_h),
                             vvcb200_intra_scratch_bytes(c->width, c->height) };
    DevBuffer *bufs[3] = { &S.iblocks, &S.sched_scratch, &S.lane->intra_marks };
    for (int i = 0; i < 3; i++)
        if (need[i] > bufs[i]->cap) {                 // (only before the buffers have reached their working size)
            VVC_CUDA(cudaStreamSynchronize(S.lane->stream));
            if (int r = bufs[i]->reserve(need[i]))
                return r;
        }
    if (int r = S.phase_counts.reserve(256))
        return r;
    if (int r = uploads_visible(c))
        return r;
    // The record order and the pre-pass need the records and the inter prediction, not the residual stage: on a picture
    // that had an INTER call they go to the lane's helper stream, behind that call only, and run beside the residual kernels
    // (everything older on the lane — the previous picture's use of the marks and of the scratch picture — lies before it).
    Lane &L = *S.lane;
    const bool beside = L.have_inter;
    cudaStream_t ps = beside ? L.helper : L.stream;
    if (beside) {
        VVC_CUDA(cudaStreamWaitEvent(L.helper, S.uploaded, 0));
        VVC_CUDA(cudaStreamWaitEvent(L.helper, L.inter_done, 0));
    }
    if (int r = vvcb200_intra_schedule_device((const VVCB200IntraBlock *)S.iblocks_in.p, (const uint32_t *)S.ctu_first.p, n, ctus_w,
                                              ctus_h, (VVCB200IntraBlock *)S.iblocks.p, (int32_t *)S.phase_counts.p,
                                              S.sched_scratch.p, ps))
        return r;
    VVCB200IntraParams phased = *params;
    phased.n_luma = phased.n_add_only = 0;
    phased.d_phase_counts = (const int32_t *)S.phase_counts.p;
    phased.prepared = 0;
    phased.luma_done_event = L.luma_done;             // luma deblocking starts there (vvcb200_pic_deblock)
    if (int r = ensure_residual(c, S.lane))           // (a picture without stored residuals still hands the planes over)
        return r;
    // the loop filters' scratch picture is idle until SAO: it keeps the prediction of the blocks rewritten in place
    if (beside) {
        if (int r = vvcb200_intra_prepare(&c->dpb[c->cur].planes, &L.residual.planes, &L.scratch.planes,
                                          (const VVCB200IntraBlock *)S.iblocks.p, n, &phased, L.intra_marks.p, ps))
            return r;
        VVC_CUDA(cudaEventRecord(L.helper_done, L.helper));
        VVC_CUDA(cudaStreamWaitEvent(L.stream, L.helper_done, 0));
        phased.prepared = 1;
    }
    if (int r = vvcb200_intra_picture(&c->dpb[c->cur].planes, &L.residual.planes, &L.scratch.planes,
                                      (const VVCB200IntraBlock *)S.iblocks.p, n, &phased, L.intra_marks.p, L.stream))
        return r;
    L.luma_pending = true;
    return set_consumed(c);
}

int vvcb200_pic_lmcs(VVCB200Ctx *c, const uint16_t *inv_lut)
{
    if (!c || !inv_lut) {
        set_error("vvcb200_pic_lmcs: bad argument");
        return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    StageSet &S = c->sets[c->set];
    if (c->lmcs_pending)                  // a second mapping of the same picture: run the first one now
        if (int r = settle_picture(c))
            return r;
    if (int r = upload(c, S.lut_inv, inv_lut, sizeof(uint16_t) << c->bit_depth))
        return r;
    // The mapping is pointwise on luma: it is applied by the deblocking pass while it loads its tiles
    // (vvcb200_deblock_lmcs_picture), or by settle_picture() if no deblocking call follows for this picture.
    c->lmcs_pending = true;
    return 0;
}

int vvcb200_pic_deblock(VVCB200Ctx *c, const VVCB200DeblockTables *t, const VVCB200DeblockParams *params)
{
    if (!c || !t || !params || t->stride4 <= 0 || !t->d_ctb_offsets) {
        set_error("vvcb200_pic_deblock: bad argument");
        return VVCB200_ERR_ARG;
    }
    DeviceGuard g(c->device);
    StageSet &S = c->sets[c->set];
    const size_t rows4 = (size_t)(c->height + 3) >> 2;
    const size_t units = (size_t)t->stride4 * rows4;
    const int shift = params->ctb_log2_size;
    const size_t n_ctb = (size_t)((c->width + (1 << shift) - 1) >> shift) * ((c->height + (1 << shift) - 1) >> shift);
    // one device arena: 6 edge tables, 3 qp tables, the per-CTB offsets
    if (units * 9 + n_ctb * 6 + 64 > S.db_tables.cap) {
        VVC_CUDA(cudaStreamSynchronize(S.lane->stream));
        VVC_CUDA(cudaStreamSynchronize(c->h2d));
    }
    if (int r = S.db_tables.reserve(units * 9 + n_ctb * 6 + 64))
        return r;
    uint8_t *d = (uint8_t *)S.db_tables.p;
    VVCB200DeblockTables dt = *t;
    // ten copies: the host tables are separate allocations (a copy may not span two pinned allocations even when they
    // happen to be adjacent)
    for (int dir = 0; dir < 2; dir++)
        for (int k = 0; k < 3; k++) {
            if (!t->d_edge[dir][k]) {
                set_error("vvcb200_pic_deblock: missing edge table");
                return VVCB200_ERR_ARG;
            }
            if (void *m = arena_place(c, t->d_edge[dir][k], units)) {
                dt.d_edge[dir][k] = (const uint8_t *)m;
            } else {
                VVC_CUDA(cudaMemcpyAsync(d, t->d_edge[dir][k], units, cudaMemcpyHostToDevice, c->h2d));
                dt.d_edge[dir][k] = d;
            }
            d += units;
        }
    for (int k = 0; k < 3; k++) {
        if (!t->d_qp[k]) {
            set_error("vvcb200_pic_deblock: missing qp table");
            return VVCB200_ERR_ARG;
        }
        if (void *m = arena_place(c, t->d_qp[k], units)) {
            dt.d_qp[k] = (const int8_t *)m;
        } else {
            VVC_CUDA(cudaMemcpyAsync(d, t->d_qp[k], units, cudaMemcpyHostToDevice, c->h2d));
            dt.d_qp[k] = (const int8_t *)d;
        }
        d += units;
    }
    if (void *m = arena_place(c, t->d_ctb_offsets, n_ctb * 6)) {
        dt.d_ctb_offsets = (const int8_t *)m;
    } else {
        VVC_CUDA(cudaMemcpyAsync(d, t->d_ctb_offsets, n_ctb * 6, cudaMemcpyHostToDevice, c->h2d));
        dt.d_ctb_offsets = (const int8_t *)d;
    }
    if (int r = uploads_visible(c))
        return r;
    if (c->in_scratch)                    // a second deblocking call for this picture: start from its DPB buffer again
        if (int r = settle_picture(c))
            return r;
    const uint16_t *lut = c->lmcs_pending ? (const uint16_t *)S.lut_inv.p : nullptr;
    Lane &L = *S.lane;
    if (L.luma_pending && L.have_inter) {
        // the intra stage left an event where luma became final: the luma half starts there, on the helper stream, beside
        // the stage's chroma phases; the chroma half follows the stage on the lane.  (Pictures with inter units only: on an
        // all-intra picture the chroma wavefront is a dependency chain, and a grid that takes SMs from its pollers costs
        // more than it hides — 11.1 instead of 10.0 ms at 8K.)
        L.luma_pending = false;
        VVC_CUDA(cudaStreamWaitEvent(L.helper, S.uploaded, 0));
        VVC_CUDA(cudaStreamWaitEvent(L.helper, L.luma_done, 0));
        if (int r = vvcb200_deblock_lmcs_planes(&L.scratch.planes, &c->dpb[c->cur].planes, &dt, params, lut, 1, L.helper))
            return r;
        if (int r = vvcb200_deblock_lmcs_planes(&L.scratch.planes, &c->dpb[c->cur].planes, &dt, params, lut, 2, L.stream))
            return r;
        VVC_CUDA(cudaEventRecord(L.helper_done, L.helper));
        VVC_CUDA(cudaStreamWaitEvent(L.stream, L.helper_done, 0));
    } else if (int r = vvcb200_deblock_lmcs_picture(&L.scratch.planes, &c->dpb[c->cur].planes, &dt, params, lut, L.stream))
        return r;
    c->lmcs_pending = false;
    c->in_scratch = true;
    return set_consumed(c);
}

int vvcb200_pic_sao_alf(VVCB200Ctx *c, const VVCB200SaoCtb *sao, const VVCB200AlfCtb *alf,
                        const VVCB200AlfParams *alf_params, int ctb_log2_size)
{
    if (!c || ctb_log2_size < 5 || ctb_log2_size > 7 || (alf && !alf_params)) {
        set_error("vvcb200_pic_sao_alf: bad argument");
        return VVCB200_ERR_ARG;
    }
    if (!sao && !alf)
        return 0;
    DeviceGuard g(c->device);
    StageSet &S = c->sets[c->set];
    const int sh = ctb_log2_size;
    const size_t n_ctb = (size_t)((c->width + (1 << sh) - 1) >> sh) * ((c->height + (1 << sh) - 1) >> sh);
    if (sao)
        if (int r = upload(c, S.sao, sao, n_ctb * sizeof(*sao)))
            return r;
    if (alf) {
        if (int r = upload(c, S.alf, alf, n_ctb * sizeof(*alf)))
            return r;
        if (int r = upload(c, S.alf_params, alf_params, sizeof(*alf_params)))
            return r;
        if (int r = S.alf_table.reserve(vvcb200_sao_alf_scratch_bytes()))
            return r;
    }
    if (int r = uploads_visible(c))
        return r;
    if (c->lmcs_pending) {                // no deblocking pass took the mapping along: run it where the picture is now
        const bool keep = c->in_scratch;
        c->in_scratch = false;            // (settle_picture must not move the picture)
        VVCB200Planes *pl = keep ? &S.lane->scratch.planes : &c->dpb[c->cur].planes;
        c->lmcs_pending = false;
        if (int r = vvcb200_lmcs_picture(pl, (const uint16_t *)S.lut_inv.p, c->bit_depth, S.lane->stream))
            return r;
        c->in_scratch = keep;
    }
    // one fused pass from wherever the picture is into the other buffer
    VVCB200Planes *cur = &c->dpb[c->cur].planes, *tmp = &S.lane->scratch.planes;
    VVCB200Planes *src = c->in_scratch ? tmp : cur, *dst = c->in_scratch ? cur : tmp;
    if (int r = vvcb200_sao_alf_picture(dst, src, sao ? (const VVCB200SaoCtb *)S.sao.p : nullptr,
                                        alf ? (const VVCB200AlfCtb *)S.alf.p : nullptr,
                                        alf ? (const VVCB200AlfParams *)S.alf_params.p : nullptr, ctb_log2_size, c->bit_depth,
                                        alf ? S.alf_table.p : nullptr, S.lane->stream))
        return r;
    c->in_scratch = !c->in_scratch;
    return set_consumed(c);
}

}  // extern "C"

// ---- host-side descriptor producers ----------------------------------------------------------------------
extern "C" int vvcb200_pack_levels(const int16_t *dense, VVCB200TransformBlock *tbs, int n, int sparse, int narrow, int16_t *out,
                                   size_t out_cap, size_t *out_len)
{
    if (!dense || !tbs || !out || !out_len || n < 0) {
        set_error("vvcb200_pack_levels: bad argument");
        return VVCB200_ERR_ARG;
    }
    size_t o = 0;
    for (int t = 0; t < n; t++) {
        const int box = tbs[t].nzw * tbs[t].nzh, words = sparse ? (box + 31) >> 5 : 0;
        const int16_t *lv = dense + tbs[t].coeff_off;
        if (o + 2 * (size_t)words + box + 2 > out_cap) {
            set_error("vvcb200_pack_levels: output buffer too small");
            return VVCB200_ERR_TOO_BIG;
        }
        bool fits8 = narrow != 0;
        for (int i = 0; fits8 && i < box; i++)
            fits8 = lv[i] >= -128 && lv[i] <= 127;
        uint32_t *bitmap = reinterpret_cast<uint32_t *>(out + o);     // sparse: o stays even
        int16_t *body = out + o + 2 * words;
        int8_t *body8 = reinterpret_cast<int8_t *>(body);
        int count = 0;
        if (sparse) {
            for (int w = 0; w < words; w++) {
                uint32_t bits = 0;
                const int lim = box - 32 * w < 32 ? box - 32 * w : 32;
                for (int b = 0; b < lim; b++)
                    if (lv[32 * w + b]) {
                        bits |= 1u << b;
                        if (fits8)
                            body8[count++] = (int8_t)lv[32 * w + b];
                        else
                            body[count++] = lv[32 * w + b];
                    }
                bitmap[w] = bits;
            }
        } else {
            for (int i = 0; i < box; i++)
                if (fits8)
                    body8[count++] = (int8_t)lv[i];
                else
                    body[count++] = lv[i];
        }
        if (fits8) {
            while (count & (sparse ? 3 : 1))                          // whole int16 slots; sparse: the next bitmap stays 4-byte aligned
                body8[count++] = 0;
            count >>= 1;
        } else if (sparse && (count & 1)) {
            body[count++] = 0;
        }
        tbs[t].flags = (uint16_t)((tbs[t].flags & ~VVCB200_TB_LEVELS8) | (fits8 ? VVCB200_TB_LEVELS8 : 0));
        tbs[t].coeff_off = (uint32_t)o;
        o += 2 * (size_t)words + count;
    }
    *out_len = o;
    return 0;
}

extern "C" int vvcb200_pack_levels_sparse(const int16_t *dense, VVCB200TransformBlock *tbs, int n, int16_t *out, size_t out_cap,
                                          size_t *out_len)
{
    return vvcb200_pack_levels(dense, tbs, n, 1, 0, out, out_cap, out_len);
}

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
