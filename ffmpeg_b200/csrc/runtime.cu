// runtime.cu — error reporting, launch accounting and device queries behind the C ABI.
#include <atomic>
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace vvcb200 {

static thread_local char g_err[512] = "";
static std::atomic<uint64_t> g_launches{0};

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line)
{
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return e == cudaErrorMemoryAllocation ? VVCB200_ERR_NOMEM : VVCB200_ERR_CUDA;
}

void count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

int sm_count()
{
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
        return 148;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

ForkStreams *fork_streams(cudaStream_t main_stream)
{
    // one set per calling stream: helper streams shared between two pictures' streams would order them behind each other
    struct Entry { int dev; cudaStream_t main; ForkStreams f; };
    static thread_local Entry pool[16];
    static thread_local int used = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess)
        return nullptr;
    for (int i = 0; i < used; i++)
        if (pool[i].dev == dev && pool[i].main == main_stream)
            return &pool[i].f;
    if (used == 16) {                               // all sets taken: hand an old one over (its streams serve any caller)
        static thread_local int next = 0;
        Entry &e = pool[next++ & 15];
        e.dev = dev;
        e.main = main_stream;
        return &e.f;
    }
    Entry &e = pool[used];
    ForkStreams f;
    bool good = cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < 4 && good; i++)
        good = cudaStreamCreateWithFlags(&f.side[i], cudaStreamNonBlocking) == cudaSuccess &&
               cudaEventCreateWithFlags(&f.join[i], cudaEventDisableTiming) == cudaSuccess;
    if (!good)
        return nullptr;
    f.ok = true;
    e.dev = dev;
    e.main = main_stream;
    e.f = f;
    used++;
    return &e.f;
}

}  // namespace vvcb200

extern "C" {

const char *vvcb200_last_error(void) { return vvcb200::g_err; }
int vvcb200_abi_version(void) { return VVCB200_ABI_VERSION; }
uint64_t vvcb200_launch_count(void) { return vvcb200::g_launches.load(std::memory_order_relaxed); }

}
