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

// Error bits recorded per stream in the engine header (never silently ignored: the
// host wrapper raises when any bit is set).
enum : int {
    BM_ERR_TRACK_OVERFLOW = 1,   // more live tracks than the engine capacity
    BM_ERR_DET_OVERFLOW = 2,     // ndets > capacity
    BM_ERR_SCRATCH = 4,          // cost matrix larger than the scratch region
};

}  // namespace bm
