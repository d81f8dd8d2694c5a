// voxel.cuh -- event stream -> temporal-bilinear voxel grid (B, C, H, W), an EXTENSION: the reference
// has no voxel grid (its network input is the 4-channel event image of EV_FlowNet/net.py:88-150);
// BASELINE.json's north_star and configs name "bins", so the published definition (Zhu et al.,
// "Unsupervised Event-based Learning of Optical Flow, Depth and Egomotion", CVPR 2019, eq. 1-3; the
// events_to_voxel_grid of the E2VID code base) is built on the encoder's own machinery.  Parity is
// UNPINNED by construction (no reference implementation, no golden vectors from it).
//
//   t*  = (C - 1) * (t - start[b]) / dt[b]          start, dt exactly as the event image (net.py:136-141)
//   out[b, k, y, x] = sum over the pixel's events of  p * max(0, 1 - |t* - k|)
// evaluated per event as k0 = floor(t*), f = t* - k0:  p (1 - f) into bin k0, p f into bin k0 + 1,
// each only when the bin exists (events before `start` or after the last event of the sample keep the
// part of their tent that falls inside [0, C - 1]).
//
// Two device paths, as for the event image: direct (two RED.ADD.F32 per event into the zeroed grid)
// and binned -- pass 1 is encode_bin_kernel unchanged (the 6-byte records do not depend on C), pass 2
// keeps a [C][rows x W] fp32 slab of one (sample, row band) in shared memory, applies two shared
// atomic adds per record and writes the C plane bands once.  fp32 sums: the order of the additions
// is not fixed, results agree with an exact sum to ~1e-7 of the absolute sum of the weights.
#pragma once
#include "common.cuh"
#include "encode_binned.cuh"

namespace dvsof {

constexpr int kVoxelMaxBins = 64;
constexpr size_t kVoxelSlabPref = 72 * 1024;     // three slabs per SM
constexpr size_t kVoxelSlabMax = 200 * 1024;     // one slab per SM
constexpr int kVoxelRedThreads = 512;

// the tile geometry of the event-image plan with a slab of C floats per pixel
inline BinPlan make_voxel_plan(int64_t n, int64_t B, int H, int W, int C) {
    BinPlan p = make_bin_plan(0, B, H, W);           // sd | ends
    if (n <= 0 || B <= 0 || B > kBinMaxSamples || C < 1 || C > kVoxelMaxBins) return p;
    const size_t px_bytes = (size_t)C * sizeof(float);
    int r = 0;
    while (((size_t)2 << r) * W * px_bytes <= kVoxelSlabPref) ++r;
    while (r > 0 && (1 << (r - 1)) >= H) --r;
    while (((H + (1 << r) - 1) >> r) > kBinMaxTiles - 1 && ((size_t)2 << r) * W * px_bytes <= kVoxelSlabMax) ++r;
    p.r_log2 = r; p.R = 1 << r;
    p.T = (H + p.R - 1) >> r;
    p.cells = p.R * W * 2;                            // pass-1 code space: pixel in the band x polarity
    if (p.T > kBinMaxTiles - 1 || p.cells > 8192 || (size_t)p.R * W * px_bytes > kVoxelSlabMax) return p;
    p.max_chunks = n / kBinChunk + B + 1;
    p.off_codes = p.total;
    p.off_ts = p.off_codes + round_up((size_t)p.max_chunks * kBinChunk * sizeof(uint16_t), 256);
    p.off_table = p.off_ts + round_up((size_t)p.max_chunks * kBinChunk * sizeof(float), 256);
    p.total = p.off_table + round_up((size_t)p.max_chunks * (p.T + 1) * sizeof(uint32_t), 256);
    p.ok = 1;
    return p;
}

// the two bins of one event and their weights (fp32, one rounding per operation)
__device__ __forceinline__ bool aligned16_dev(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

struct VoxelTaps { int k0; float w0, w1; };
__device__ __forceinline__ VoxelTaps voxel_taps(float t, float2 sd, float bins_m1, float sign) {
    const float tn = __fdiv_rn(__fsub_rn(t, sd.x), sd.y);        // net.py:141
    const float ts = __fmul_rn(tn, bins_m1);
    const float fl = floorf(ts);
    const float f = __fsub_rn(ts, fl);                           // exact
    VoxelTaps v;
    // far outside the grid (or not finite): no bin exists; the clamp keeps the conversion defined
    v.k0 = (fl >= -2.f && fl <= 1.0e6f) ? (int)fl : -2;
    v.w1 = __fmul_rn(sign, f);
    v.w0 = __fmul_rn(sign, __fsub_rn(1.f, f));
    return v;
}

// ---------------------------------------------------------------------------------------
// direct path
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned voxel_scatter_event(int64_t xi, int64_t yi, float ti, int64_t pi, int64_t bi,
                                                        const float2* __restrict__ sd, int n_samples, int C,
                                                        int H, int W, float* __restrict__ out) {
    if (pi != 1 && pi != -1) return DVSOF_ST_BAD_POLARITY;
    if ((uint64_t)bi >= (uint64_t)n_samples) return DVSOF_ST_BAD_SAMPLE;
    if ((uint64_t)xi >= (uint64_t)W || (uint64_t)yi >= (uint64_t)H) return DVSOF_ST_BAD_PIXEL;
    const VoxelTaps v = voxel_taps(ti, __ldg(sd + (int)bi), (float)(C - 1), (float)(int)pi);
    const size_t plane = (size_t)H * W;
    float* px = out + (size_t)bi * C * plane + ((unsigned)yi * (unsigned)W + (unsigned)xi);
    if ((unsigned)v.k0 < (unsigned)C) atomicAdd(px + (size_t)v.k0 * plane, v.w0);
    if ((unsigned)(v.k0 + 1) < (unsigned)C && v.w1 != 0.f) atomicAdd(px + (size_t)(v.k0 + 1) * plane, v.w1);
    return 0u;
}

template <bool VEC>
__global__ void __launch_bounds__(256)
voxel_direct_kernel(const int64_t* __restrict__ x, const int64_t* __restrict__ y,
                    const float* __restrict__ t, const int64_t* __restrict__ p,
                    const int64_t* __restrict__ b, int64_t n, const float2* __restrict__ sd,
                    int n_samples, int C, int H, int W, float* __restrict__ out,
                    uint32_t* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned flags = 0;
    int64_t done = 0;
    bool waited = false;
    if (VEC) {
        const int64_t nvec = n >> 2;
        done = nvec << 2;
        for (int64_t i = tid; i < nvec; i += stride) {
            const longlong2* x2 = reinterpret_cast<const longlong2*>(x) + 2 * i;
            const longlong2* y2 = reinterpret_cast<const longlong2*>(y) + 2 * i;
            const longlong2* p2 = reinterpret_cast<const longlong2*>(p) + 2 * i;
            const longlong2* b2 = reinterpret_cast<const longlong2*>(b) + 2 * i;
            const longlong2 xa = __ldcs(x2), xb = __ldcs(x2 + 1), ya = __ldcs(y2), yb = __ldcs(y2 + 1);
            const longlong2 pa = __ldcs(p2), pb = __ldcs(p2 + 1), ba = __ldcs(b2), bb = __ldcs(b2 + 1);
            const float4 tt = __ldcs(reinterpret_cast<const float4*>(t) + i);
            const int64_t nxt = 4 * (i + 1);
            const long long bnext = nxt < n ? b[nxt] : bb.y;
            if (!waited) { pdl_wait_for(sd); waited = true; }     // grid zeroed + `sd` ready from here on
            if (ba.x > ba.y || ba.y > bb.x || bb.x > bb.y || bb.y > bnext) flags |= DVSOF_ST_NOT_GROUPED;
            flags |= voxel_scatter_event(xa.x, ya.x, tt.x, pa.x, ba.x, sd, n_samples, C, H, W, out);
            flags |= voxel_scatter_event(xa.y, ya.y, tt.y, pa.y, ba.y, sd, n_samples, C, H, W, out);
            flags |= voxel_scatter_event(xb.x, yb.x, tt.z, pb.x, bb.x, sd, n_samples, C, H, W, out);
            flags |= voxel_scatter_event(xb.y, yb.y, tt.w, pb.y, bb.y, sd, n_samples, C, H, W, out);
        }
    }
    if (!waited) pdl_wait_for(sd);
    for (int64_t i = done + tid; i < n; i += stride) {
        const int64_t bi = __ldcs(b + i);
        if (i + 1 < n && bi > b[i + 1]) flags |= DVSOF_ST_NOT_GROUPED;
        flags |= voxel_scatter_event(__ldcs(x + i), __ldcs(y + i), __ldcs(t + i), __ldcs(p + i), bi, sd,
                                     n_samples, C, H, W, out);
    }
    if (flags) atomicOr(status, flags);
}

// ---------------------------------------------------------------------------------------
// binned path, pass 2: one CTA per (sample, row band), slab [C][R * W] fp32 in shared memory
// ---------------------------------------------------------------------------------------
template <int kLP>                                    // records per lane and run in flight (2 runs at a time)
__global__ void __launch_bounds__(kVoxelRedThreads)
voxel_reduce_kernel(const uint16_t* __restrict__ codes_g, const float* __restrict__ ts_g,
                    const uint32_t* __restrict__ table, const int64_t* __restrict__ ends,
                    const float2* __restrict__ sd, int n_samples, int C, int H, int W, int r_log2, int T,
                    float* __restrict__ out) {
    extern __shared__ __align__(16) float vslab[];      // [C][npx]
    __shared__ int prefix[kBinMaxSamples + 1];
    const int s = blockIdx.x / T, tile = blockIdx.x - s * T;
    const int npx = (1 << r_log2) * W;
    const int n_slab = npx * C;
    if ((n_slab & 3) == 0) {                             // the slab is private: zeroed while pass 1 still runs
        float4* v4 = reinterpret_cast<float4*>(vslab);
        for (int i = threadIdx.x; i < n_slab / 4; i += kVoxelRedThreads) v4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
        for (int i = threadIdx.x; i < n_slab; i += kVoxelRedThreads) vslab[i] = 0.f;
    }
    pdl_wait_for(codes_g, ts_g, table, ends, sd);
    chunk_prefix<kVoxelRedThreads>(ends, n_samples, prefix);   // ends with __syncthreads()
    const int cb = prefix[s], nc = prefix[s + 1] - prefix[s];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kVoxelRedThreads / 32;
    const int my_n = nc > warp ? (nc - warp + kWarps - 1) / kWarps : 0;
    const float2 nd = sd[s];
    const float bins_m1 = (float)(C - 1);
    auto apply = [&](unsigned c, float t) {
        if (c == 0xffffffffu) return;
        // the records carry the order-preserving key of t (encode_binned.cuh): back to the raw t
        const VoxelTaps v = voxel_taps(key_to_float(__float_as_int(t)), nd, bins_m1, (c & 1u) ? -1.f : 1.f);     // code bit 0: 1 = negative
        float* px = vslab + (c >> 1);
        if ((unsigned)v.k0 < (unsigned)C) atomicAdd(px + v.k0 * npx, v.w0);
        if ((unsigned)(v.k0 + 1) < (unsigned)C && v.w1 != 0.f) atomicAdd(px + (v.k0 + 1) * npx, v.w1);
    };
    // the chunks of the sample are dealt to the warps round-robin; a lane holds the bounds of one of
    // the warp's next 32 runs; two runs x kLP records per lane are in flight (encode_reduce_kernel's lean loop)
    constexpr int kLR = 2;
    constexpr size_t kStride = (size_t)kWarps * kBinChunk;
    for (int k0 = 0; k0 < my_n; k0 += 32) {
        unsigned lo = 0, hi = 0;
        if (k0 + lane < my_n) {
            const uint32_t* row = table + (size_t)(cb + (k0 + lane) * kWarps + warp) * (T + 1) + tile;
            lo = __ldg(row); hi = __ldg(row + 1);
        }
        const int m = min(32, my_n - k0);
        const uint16_t* cg = codes_g + (size_t)(cb + k0 * kWarps + warp) * kBinChunk;
        const float* tg = ts_g + (size_t)(cb + k0 * kWarps + warp) * kBinChunk;
        for (int j = 0; j < m; j += kLR) {
            unsigned ra[kLR], rb[kLR];
#pragma unroll
            for (int q = 0; q < kLR; ++q) {
                ra[q] = __shfl_sync(0xffffffffu, lo, (j + q) & 31);
                rb[q] = __shfl_sync(0xffffffffu, hi, (j + q) & 31);
                if (j + q >= m) rb[q] = ra[q];
            }
            const uint16_t* c0 = cg + j * kStride;
            const float* t0 = tg + j * kStride;
            unsigned code[kLR * kLP];
            float tv[kLR * kLP];
#pragma unroll
            for (int q = 0; q < kLR; ++q)
#pragma unroll
                for (int u = 0; u < kLP; ++u) {
                    const unsigned i = ra[q] + u * 32 + lane;
                    code[q * kLP + u] = 0xffffffffu;
                    if (i < rb[q]) {
                        code[q * kLP + u] = __ldcs(c0 + q * kStride + i);
                        tv[q * kLP + u] = __ldcs(t0 + q * kStride + i);
                    }
                }
#pragma unroll
            for (int u = 0; u < kLR * kLP; ++u) apply(code[u], tv[u]);
#pragma unroll
            for (int q = 0; q < kLR; ++q)
                for (unsigned i = ra[q] + 32 * kLP + lane; __any_sync(0xffffffffu, i < rb[q]); i += 32)
                    if (i < rb[q]) apply(__ldcs(c0 + q * kStride + i), __ldcs(t0 + q * kStride + i));
        }
    }
    __syncthreads();
    const int y0 = tile << r_log2;
    const size_t plane = (size_t)H * W;
    const int n_px = min(1 << r_log2, H - y0) * W;       // whole image rows: contiguous in every plane
    float* o0 = out + (size_t)s * C * plane + (size_t)y0 * W;
    if (((npx | n_px) & 3) == 0 && (plane & 3) == 0 && aligned16_dev(o0)) {
        for (int k = 0; k < C; ++k) {
            float4* o4 = reinterpret_cast<float4*>(o0 + (size_t)k * plane);
            const float4* s4 = reinterpret_cast<const float4*>(vslab + k * npx);
            for (int i = threadIdx.x; i < n_px / 4; i += kVoxelRedThreads) __stcs(o4 + i, s4[i]);
        }
    } else {
        for (int k = 0; k < C; ++k)
            for (int i = threadIdx.x; i < n_px; i += kVoxelRedThreads) o0[(size_t)k * plane + i] = vslab[k * npx + i];
    }
}

}  // namespace dvsof
