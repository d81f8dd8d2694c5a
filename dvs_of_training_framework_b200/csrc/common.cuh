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

// Launch through cudaLaunchKernelEx; `pdl` = programmatic dependent launch: the kernel may start
// while its predecessor on the stream still runs and must call griddepcontrol.wait before it
// touches anything the predecessor writes.
template <typename... KArgs, typename... Args>
inline int launch_kernel(const char* what, void (*kern)(KArgs...), int grid, int block, size_t smem,
                         cudaStream_t stream, bool pdl, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
    return DVSOF_OK;
}

// Function attributes (dynamic shared memory opt-in) are per DEVICE: `mask` holds one bit per device
// ordinal on which the caller has configured its kernels.  Returns true when the current device
// still needs it (a race between threads only repeats the cheap call); devices >= 64 always do.
inline bool needs_device_setup(const unsigned long long& mask, int* dev_out) {
    int dev = 0;
    cudaGetDevice(&dev);
    *dev_out = dev;
    return dev >= 64 || !((__atomic_load_n(&mask, __ATOMIC_RELAXED) >> dev) & 1ull);
}
inline void device_setup_done(unsigned long long& mask, int dev) {
    if (dev < 64) __atomic_fetch_or(&mask, 1ull << dev, __ATOMIC_RELAXED);
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

// programmatic dependent launch (PDL): release the next kernel on the stream early / block until
// every kernel this one depends on has completed and its writes are visible.  Both are no-ops
// when the launch did not carry the PDL attribute (see launch_kernel above).
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// A pointer to data the PREDECESSOR writes must be passed through here after the wait: loads
// through `const T* __restrict__` are invariant loads for the compiler, which moves them above the
// wait's asm statement (seen in SASS: the match_idx loads of the forward loss sweep sat 260
// instructions before ACQBULK and read the previous contents of the buffer).  The empty asm makes
// the pointer's value unknown until the wait has executed, so nothing can be loaded through it
// earlier.
template <typename P>                               // P: a (possibly __restrict__) pointer type
__device__ __forceinline__ void pdl_acquire(P& p) {
    static_assert(sizeof(P) == 8, "pointers only");
    asm volatile("" : "+l"(p) : : "memory");
}
template <typename... Ps>
__device__ __forceinline__ void pdl_wait_for(Ps&... ptrs) {
    pdl_wait();
    (pdl_acquire(ptrs), ...);
}

// streaming (evict-first) loads for data touched exactly once
template <typename T>
__device__ __forceinline__ T ld_stream(const T* p) { return __ldcs(p); }

}  // namespace dvsof
