// runtime.cu — error reporting, launch accounting and device queries behind the C ABI.
#include <atomic>
#include <mutex>
#include <new>
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

namespace {
struct ForkEntry { int dev; cudaStream_t main; ForkStreams f; };
std::mutex g_fork_mutex;
ForkEntry *g_fork_pool[64];
int g_fork_used = 0;

void destroy_fork(ForkStreams &f)
{
    for (int i = 0; i < 4; i++) {
        if (f.side[i]) cudaStreamDestroy(f.side[i]);
        if (f.join[i]) cudaEventDestroy(f.join[i]);
    }
    if (f.fork) cudaEventDestroy(f.fork);
    f = ForkStreams();
}
}  // namespace

ForkStreams *fork_streams(cudaStream_t main_stream)
{
    // one set per device and calling stream: helper streams shared between two pictures' streams would order them behind
    // each other, and a stream or event of one device cannot serve a launch on another
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess)
        return nullptr;
    std::lock_guard<std::mutex> lock(g_fork_mutex);
    for (int i = 0; i < g_fork_used; i++)
        if (g_fork_pool[i]->dev == dev && g_fork_pool[i]->main == main_stream)
            return &g_fork_pool[i]->f;
    if (g_fork_used == 64)
        return nullptr;
    ForkEntry *e = new (std::nothrow) ForkEntry;
    if (!e)
        return nullptr;
    ForkStreams &f = e->f;
    bool good = cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < 4 && good; i++)
        good = cudaStreamCreateWithFlags(&f.side[i], cudaStreamNonBlocking) == cudaSuccess &&
               cudaEventCreateWithFlags(&f.join[i], cudaEventDisableTiming) == cudaSuccess;
    if (!good) {
        destroy_fork(f);
        delete e;
        return nullptr;
    }
    f.ok = true;
    e->dev = dev;
    e->main = main_stream;
    g_fork_pool[g_fork_used++] = e;
    return &e->f;
}

void fork_streams_release(cudaStream_t main_stream)
{
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess)
        return;
    std::lock_guard<std::mutex> lock(g_fork_mutex);
    for (int i = 0; i < g_fork_used; i++)
        if (g_fork_pool[i]->dev == dev && g_fork_pool[i]->main == main_stream) {
            destroy_fork(g_fork_pool[i]->f);
            delete g_fork_pool[i];
            g_fork_pool[i] = g_fork_pool[--g_fork_used];
            return;
        }
}

}  // namespace vvcb200

extern "C" {

const char *vvcb200_last_error(void) { return vvcb200::g_err; }
int vvcb200_abi_version(void) { return VVCB200_ABI_VERSION; }
uint64_t vvcb200_launch_count(void) { return vvcb200::g_launches.load(std::memory_order_relaxed); }

}
