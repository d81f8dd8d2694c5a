// common.cuh -- shared host/device helpers for libdvsof (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/dvsof.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libdvsof is written for Blackwell (sm_100a); build with -gencode arch=compute_100a,code=sm_100a"
#endif

namespace dvsof {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// host: record a thread-local error string, return the code
int fail(int code, const char* fmt, ...);

inline int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
    return DVSOF_OK;
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- device helpers ----
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// streaming (evict-first) loads for data touched exactly once
template <typename T>
__device__ __forceinline__ T ld_stream(const T* p) { return __ldcs(p); }

}  // namespace dvsof
