// Shared helpers for the sm_100a kernels of the association hot path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math_constants.h>

#define BM_WARP 32
#define BM_FULL 0xffffffffu

// Keep fp64 arithmetic un-contracted where the operation order mirrors the numpy
// reference (a*b+c must round twice, as numpy does).  nvcc is invoked with -fmad=false
// for the whole library; this macro documents the sites that rely on it.
#define BM_NO_FMA

namespace bm {

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

__device__ __forceinline__ double shfl_d(double v, int src) {
    return __shfl_sync(BM_FULL, v, src);
}
__device__ __forceinline__ double shfl_xor_d(double v, int m) {
    return __shfl_xor_sync(BM_FULL, v, m);
}

// Stable compaction by ONE warp: out[k0 + rank] = val(i) for every i in [0, n) with
// pred(i), in ascending i.  Returns k0 + count on all lanes.
template <typename Pred, typename Val>
__device__ __forceinline__ int warp_compact(int n, int* out, int k0, Pred pred, Val val) {
    const int lane = lane_id();
    for (int base = 0; base < n; base += 32) {
        const int i = base + lane;
        const bool ok = i < n && pred(i);
        const unsigned b = __ballot_sync(BM_FULL, ok);
        if (ok) out[k0 + __popc(b & ((1u << lane) - 1))] = val(i);
        k0 += __popc(b);
    }
    __syncwarp();
    return k0;
}

// Error bits recorded per stream in the engine header (never silently ignored: the
// host wrapper raises when any bit is set).
enum : int {
    BM_ERR_TRACK_OVERFLOW = 1,   // more live tracks than the engine capacity
    BM_ERR_DET_OVERFLOW = 2,     // ndets > capacity
    BM_ERR_SCRATCH = 4,          // cost matrix larger than the scratch region
};

// Opt a kernel into the device's whole dynamic shared-memory range.  The attribute belongs to the kernel
// AND the current device and is process-global, so an engine must never set it to its OWN need: a second
// engine with smaller capacities would lower it under the first one's launches (and an engine on another
// device would never get the opt-in at all).  Called at every engine create with that engine's device
// current; the value only depends on the device, so repeating it is harmless.
template <typename Kernel>
inline cudaError_t smem_optin(Kernel kernel) {
    int dev = 0, optin = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (e != cudaSuccess) return e;
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, kernel);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes);
}

}  // namespace bm
