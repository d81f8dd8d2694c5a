// ubench_atomics.cu -- B200 micro-benchmark behind the encoder design:
// lane rate of (a) global no-return reductions RED.ADD.F32 + RED.MAX.S32 on spread addresses,
// (b) shared-memory atomics on spread addresses, (c) cluster-remote shared-memory atomics.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_atomics ubench_atomics.cu
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ uint32_t rng(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

// (a) global REDs: each thread issues `iters` pairs (add.f32 + max.s32) at random cells
__global__ void k_redg(float* out, uint32_t cells_mask, int iters, int pair) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    for (int i = 0; i < iters; ++i) {
        const uint32_t c = rng(s) & cells_mask;
        atomicAdd(out + c, 1.0f);
        if (pair) atomicMax(reinterpret_cast<int*>(out) + cells_mask + 1 + c, (int)(s >> 3));
    }
}

// (b) shared atomics on a per-CTA slab of `cells` u32 counters + `cells` s32 maxima
__global__ void k_atoms(unsigned* sink, int cells, int iters, int pair) {
    extern __shared__ unsigned slab[];
    for (int i = threadIdx.x; i < 2 * cells; i += blockDim.x) slab[i] = 0;
    __syncthreads();
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 777u;
    for (int i = 0; i < iters; ++i) {
        const uint32_t c = rng(s) % cells;
        atomicAdd(slab + c, 1u);
        if (pair) atomicMax(reinterpret_cast<int*>(slab) + cells + c, (int)(s >> 3));
    }
    __syncthreads();
    unsigned acc = 0;
    for (int i = threadIdx.x; i < 2 * cells; i += blockDim.x) acc += slab[i];
    if (acc == 0xdeadbeef) sink[0] = acc;
}

// (c) cluster of CL CTAs, slab distributed over the cluster; target CTA = cell / cells_per_cta
template <int CL>
__global__ void __cluster_dims__(CL, 1, 1) k_dsmem(unsigned* sink, int cells, int iters, int pair) {
    extern __shared__ unsigned slab[];
    cg::cluster_group cluster = cg::this_cluster();
    for (int i = threadIdx.x; i < 2 * cells; i += blockDim.x) slab[i] = 0;
    cluster.sync();
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 99u;
    for (int i = 0; i < iters; ++i) {
        const uint32_t r = rng(s);
        const uint32_t c = r % cells;
        const unsigned dst = (r >> 20) % CL;
        unsigned* remote = cluster.map_shared_rank(slab, dst);
        atomicAdd(remote + c, 1u);
        if (pair) atomicMax(reinterpret_cast<int*>(remote) + cells + c, (int)(s >> 3));
    }
    cluster.sync();
    unsigned acc = 0;
    for (int i = threadIdx.x; i < 2 * cells; i += blockDim.x) acc += slab[i];
    if (acc == 0xdeadbeef) sink[0] = acc;
}

template <typename F>
static float time_ms(F launch) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    launch();
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) printf("  CUDA error: %s\n", cudaGetErrorString(e));
    return ms;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount;
    const double clk = 1.965e9;   // nominal max; printed rates are absolute, cyc/lane is indicative
    printf("{\"device\": \"%s\", \"sms\": %d,\n", p.name, sms);
    float* out; unsigned* sink;
    cudaMalloc(&sink, 64);
    for (int log2cells = 21; log2cells <= 24; log2cells += 3) {      // 2M cells (8 MB) / 16M (64 MB) per plane
        const uint32_t cells = 1u << log2cells;
        cudaMalloc(&out, (size_t)cells * 8);
        cudaMemset(out, 0, (size_t)cells * 8);
        for (int pair = 0; pair <= 1; ++pair) {
            const int iters = 512, grid = sms * 8, thr = 256;
            const float ms = time_ms([&] { k_redg<<<grid, thr>>>(out, cells - 1, iters, pair); });
            const double ops = (double)grid * thr * iters * (1 + pair);
            printf(" \"redg_cells2^%d_pair%d\": {\"Gops\": %.1f, \"cyc_per_lane_per_sm\": %.3f},\n",
                   log2cells, pair, ops / ms / 1e6, ms * 1e-3 * clk * sms / ops);
        }
        cudaFree(out);
    }
    for (int cells : {8192, 24576}) {                                   // 64 KB / 192 KB slab
        const int smem = cells * 8;
        cudaFuncSetAttribute(k_atoms, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        for (int thr : {256, 1024})
            for (int pair = 0; pair <= 1; ++pair) {
                const int iters = 2048, grid = sms;
                const float ms = time_ms([&] { k_atoms<<<grid, thr, smem>>>(sink, cells, iters, pair); });
                const double ops = (double)grid * thr * iters * (1 + pair);
                printf(" \"atoms_cells%d_thr%d_pair%d\": {\"Gops\": %.1f, \"cyc_per_lane_per_sm\": %.4f},\n",
                       cells, thr, pair, ops / ms / 1e6, ms * 1e-3 * clk * sms / ops);
            }
    }
    {
        const int cells = 16384, smem = cells * 8, thr = 512, iters = 1024;   // 128 KB per CTA
        auto run = [&](auto kern, int cl) {
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            int max_clusters = 0;
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cl * 64); cfg.blockDim = dim3(thr); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at; at.id = cudaLaunchAttributeClusterDimension;
            at.val.clusterDim.x = cl; at.val.clusterDim.y = 1; at.val.clusterDim.z = 1;
            cfg.attrs = &at; cfg.numAttrs = 1;
            cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg);
            const int grid = (max_clusters > 0 ? max_clusters : 1) * cl;
            for (int pair = 0; pair <= 1; ++pair) {
                const float ms = time_ms([&] { kern<<<grid, thr, smem>>>(sink, cells, iters, pair); });
                const double ops = (double)grid * thr * iters * (1 + pair);
                printf(" \"dsmem_cl%d_pair%d\": {\"clusters\": %d, \"Gops\": %.1f, \"cyc_per_lane_per_cta\": %.4f},\n",
                       cl, pair, max_clusters, ops / ms / 1e6, ms * 1e-3 * clk * grid / ops);
            }
        };
        run(k_dsmem<2>, 2);
        run(k_dsmem<4>, 4);
        run(k_dsmem<8>, 8);
    }
    printf(" \"done\": true}\n");
    return 0;
}
