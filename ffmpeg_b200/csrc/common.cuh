// common.cuh — shared helpers of the sm_100a VVC engine (device inlines + host-side error plumbing).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "vvc_b200.h"

#define VVC_DEVICE_TABLES
namespace vvcb200 {

// ---- host side ---------------------------------------------------------------------------------
void set_error(const char *fmt, ...);
int  cuda_fail(cudaError_t e, const char *what, const char *file, int line);
void count_launch(int n = 1);

#define VVC_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess)                                                     \
            return vvcb200::cuda_fail(e__, #call, __FILE__, __LINE__);              \
    } while (0)

#define VVC_LAUNCH_CHECK()                                                          \
    do {                                                                            \
        cudaError_t e__ = cudaGetLastError();                                       \
        if (e__ != cudaSuccess)                                                     \
            return vvcb200::cuda_fail(e__, "kernel launch", __FILE__, __LINE__);    \
        vvcb200::count_launch();                                                    \
    } while (0)

int sm_count();   // SMs of the current device (cached per device)

// Helper streams for entry points that launch several independent grids: on one stream each grid would wait for the
// tail of the one before.  The grids are forked from the caller's stream and joined back into it with events, so the
// call keeps its stream semantics.  One set per host thread, device and calling stream (up to 16; streams are never
// destroyed behind the library's back in the engine); nullptr when none can be had (the caller stays on its one stream).
struct ForkStreams {
    cudaStream_t side[4] = { nullptr, nullptr, nullptr, nullptr };
    cudaEvent_t fork = nullptr, join[4] = { nullptr, nullptr, nullptr, nullptr };
    bool ok = false;
};
ForkStreams *fork_streams(cudaStream_t main_stream);
void fork_streams_release(cudaStream_t main_stream);   // the owner of `main_stream` is about to destroy it (current device)

// ---- device side -------------------------------------------------------------------------------
__device__ __forceinline__ int clip3(int v, int lo, int hi) { return min(max(v, lo), hi); }
__device__ __forceinline__ int clip_intp2(int v, int p) { return min(max(v, -(1 << p)), (1 << p) - 1); }
__device__ __forceinline__ int clip_pixel(int v, int bd) { return min(max(v, 0), (1 << bd) - 1); }

struct PlanesView {           // VVCB200Planes flattened for kernel parameters
    uint16_t *p[3];
    int stride[3];            // in SAMPLES
    int w[3], h[3];
};

inline PlanesView make_view(const VVCB200Planes *pl)
{
    PlanesView v;
    for (int c = 0; c < 3; c++) {
        v.p[c] = (uint16_t *)pl->d_plane[c];
        v.stride[c] = pl->stride[c] / 2;
        v.w[c] = pl->width[c];
        v.h[c] = pl->height[c];
    }
    return v;
}

}  // namespace vvcb200
