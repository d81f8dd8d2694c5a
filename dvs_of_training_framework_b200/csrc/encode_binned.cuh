// encode_binned.cuh -- two-pass, shared-memory-privatised event-image encoder for LARGE event
// streams (included by encode.cu; same results as the direct RED kernel, bit for bit).
//
// Why: the direct kernel issues two L2 reductions per event.  On B200 that path saturates at
// ~190 G scattered REDs/s (scripts/ubench_atomics.cu; ncu: lts__t_tag_requests 76 % at 164 G/s),
// i.e. <= 95 Gevents/s = 52 % of the HBM roofline of the 36 B/event stream -- while shared-memory
// atomics run 17x faster (3.2 T/s).  A whole sample's slab (H*W*2 cells * 8 B = 1 MB at 256^2)
// does not fit in one SM and cluster-remote shared atomics are as slow as global REDs (82-184 G/s),
// so the stream is first PARTITIONED by output row band:
//
//   pass 1 (bin):    a CTA takes a chunk of kBinChunk consecutive events of ONE sample, counting-
//                    sorts them by tile (band of 2^r rows) in shared memory (one shared atomic per
//                    event: histogram and rank inside the tile's run at once) and writes them back
//                    as 6-byte records {u16 cell-in-tile, f32 t} -- contiguous per (chunk, tile)
//                    run -- plus the chunk's run table.  HBM: 36 B read + 6 B written per event.
//   pass 2 (reduce): a CTA owns one (sample, tile): its slab of {u32 count, s32 max-key} lives in
//                    shared memory; it streams its runs of every chunk of the sample (6 B/event),
//                    applies ATOMS.ADD / ATOMS.MAX, then writes its 4 output plane bands ONCE with
//                    coalesced stores (no memset of the image, no global atomics).
//
// Exactness: counts are integers; the timestamp plane is max_i f(t_i) with f(t) = (t - start) / dt
// monotone non-decreasing (IEEE sub and div by dt > 0 are monotone), so it equals f(max_i t_i):
// pass 2 keeps the max of the RAW t under an order-preserving integer key and normalises once per
// cell at the flush.  max(0, .) against the zero-initialised plane = scatter_max(out=zeros).
#pragma once
#include "common.cuh"

namespace dvsof {

#ifndef DVSOF_BIN_CHUNK
#define DVSOF_BIN_CHUNK 4096
#endif
#ifndef DVSOF_BIN_THREADS
#define DVSOF_BIN_THREADS 256
#endif
#ifndef DVSOF_BIN_INFLIGHT
#define DVSOF_BIN_INFLIGHT 4
#endif
#ifndef DVSOF_BIN_MIN_CTAS
#define DVSOF_BIN_MIN_CTAS 3
#endif
constexpr int kBinChunk = DVSOF_BIN_CHUNK;       // events per pass-1 chunk
constexpr int kBinThreads = DVSOF_BIN_THREADS;   // pass-1 CTA: 16 events per thread, DVSOF_BIN_INFLIGHT at a time
#ifndef DVSOF_RED_BATCH
#define DVSOF_RED_BATCH 5
#endif
constexpr int kRedBatchLong = DVSOF_RED_BATCH;          // pass-2 loads in flight per lane
#ifndef DVSOF_RED_BATCH_SHORT
#define DVSOF_RED_BATCH_SHORT -1
#endif
constexpr int kRedBatchShort = DVSOF_RED_BATCH_SHORT;   // the same for short runs (wide images); 0 = four runs per warp at a time, -1 = the lean loop (default); the long-run batch: -k = the lean loop with k records per lane and run
constexpr int kBinMaxTiles = 128;     // tiles (row bands) per sample
constexpr int kBinMaxSamples = 512;   // samples per batch on this path (per-CTA chunk lookup)
#ifndef DVSOF_RED_THREADS
#define DVSOF_RED_THREADS 512
#endif
constexpr int kRedThreads = DVSOF_RED_THREADS;      // pass-2 CTA: 16 warps stream the runs of one slab, 3 CTAs per SM
constexpr size_t kBinSlabBytes = 32 * 1024;      // preferred slab (3 CTAs/SM and 32 row bands at 256^2)
constexpr size_t kBinSlabBytesMax = 64 * 1024;   // wide images
static_assert(kBinChunk <= 4096 && kBinMaxTiles <= 128 && kBinSlabBytesMax / 8 <= 8192, "packed record of pass 1");
constexpr int kBinStageBytes = kBinChunk * (sizeof(float) + sizeof(uint16_t));   // pass-1 staging

struct BinPlan {
    int ok;                 // geometry feasible
    int r_log2, R, T;       // rows per tile (2^r_log2) and tiles per sample
    int cells;              // R * W * 2 slab cells
    int64_t max_chunks;     // upper bound of sum_s ceil(n_s / kBinChunk)
    size_t off_ends, off_codes, off_ts, off_table, total;   // workspace layout (after the sd array)
};

inline BinPlan make_bin_plan(int64_t n, int64_t B, int H, int W) {
    BinPlan p = {};
    p.off_ends = round_up((size_t)(B > 0 ? B : 1) * sizeof(float2), 256);
    p.total = p.off_ends + round_up((size_t)(B > 0 ? B : 1) * sizeof(int64_t), 256);
    if (n <= 0 || B <= 0 || B > kBinMaxSamples) return p;
    int r = 0;
    while (((size_t)2 << r) * W * 16 <= kBinSlabBytes) ++r;        // largest 2^r rows within the preferred slab
    if ((size_t)(1 << r) * W * 16 > kBinSlabBytesMax) return p;       // one row does not fit
    while (r > 0 && (1 << (r - 1)) >= H) --r;                          // no taller than the image
    // wide images: more rows per slab (up to kBinSlabBytesMax) until the tile index fits its 7 bits
    while (((H + (1 << r) - 1) >> r) > kBinMaxTiles - 1 && ((size_t)2 << r) * W * 16 <= kBinSlabBytesMax) ++r;
    p.r_log2 = r; p.R = 1 << r;
    p.T = (H + p.R - 1) >> r;
    p.cells = p.R * W * 2;
    // packed pass-1 record: code 13 bits, tile 7 (127 = the dropped marker's), rank in a chunk 12
    if (p.T > kBinMaxTiles - 1 || p.cells > 8192 || (size_t)p.cells * 8 > kBinSlabBytesMax) return p;
    p.max_chunks = n / kBinChunk + B + 1;
    p.off_codes = p.total;
    p.off_ts = p.off_codes + round_up((size_t)p.max_chunks * kBinChunk * sizeof(uint16_t), 256);
    p.off_table = p.off_ts + round_up((size_t)p.max_chunks * kBinChunk * sizeof(float), 256);
    p.total = p.off_table + round_up((size_t)p.max_chunks * (p.T + 1) * sizeof(uint32_t), 256);
    p.ok = 1;
    return p;
}

// order-preserving map float -> signed int (all finite values and infinities); self-inverse
__device__ __forceinline__ int ordered_key(float t) {
    const int b = __float_as_int(t);
    return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float key_to_float(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

// Every CTA derives the chunk layout from the per-sample ends (B <= kBinMaxSamples):
// prefix[s] = number of chunks of samples < s.  Returns through shared memory.
template <int THREADS>
__device__ __forceinline__ void chunk_prefix(const int64_t* __restrict__ ends, int n_samples,
                                             int* prefix /* [kBinMaxSamples + 1] shared */) {
    for (int s = threadIdx.x; s < n_samples; s += THREADS) {
        const int64_t lo = s ? ends[s - 1] : 0, hi = ends[s];
        const int64_t cnt = hi > lo ? hi - lo : 0;
        prefix[s + 1] = (int)((cnt + kBinChunk - 1) / kBinChunk);
    }
    if (threadIdx.x == 0) prefix[0] = 0;
    __syncthreads();
    if (threadIdx.x < 32) {                      // one warp: serial over 32-wide blocks, scan inside
        int carry = 0;
        for (int base = 0; base < n_samples; base += 32) {
            const int i = base + threadIdx.x;
            int v = i < n_samples ? prefix[i + 1] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, o);
                if ((int)threadIdx.x >= o) v += u;
            }
            if (i < n_samples) prefix[i + 1] = v + carry;
            carry += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
}

}  // namespace dvsof

namespace dvsof {

// ---------------------------------------------------------------------------------------
// pass 1: partition one chunk of one sample by tile (counting sort in shared memory)
// ---------------------------------------------------------------------------------------
// kCompact: the storage form of utils/dataset.py:286-289 -- x, y int16, polarity uint8 (non-zero =
// positive), no sample_index array (a chunk's sample is known from `ends` = the per-sample prefix).
template <bool kCompact>
__global__ void __launch_bounds__(kBinThreads, DVSOF_BIN_MIN_CTAS)
encode_bin_kernel(const void* __restrict__ x_, const void* __restrict__ y_,
                  const float* __restrict__ t, const void* __restrict__ p_,
                  const int64_t* __restrict__ b, int64_t n, const int64_t* __restrict__ ends,
                  int n_samples, int H, int W, int r_log2, int T,
                  uint16_t* __restrict__ codes_g, float* __restrict__ ts_g,
                  uint32_t* __restrict__ table, uint32_t* __restrict__ status) {
    __shared__ int prefix[kBinMaxSamples + 1];
    __shared__ unsigned hist[kBinMaxTiles + 1];      // counts, then exclusive bases (T + 1 entries)
    __shared__ unsigned cursor[kBinMaxTiles];
    extern __shared__ __align__(16) unsigned char stage_raw[];   // kBinStageBytes (dynamic: > 48 KB total)
    float* ts_s = reinterpret_cast<float*>(stage_raw);
    uint16_t* codes_s = reinterpret_cast<uint16_t*>(stage_raw + kBinChunk * sizeof(float));
    __shared__ int s_sample;
    pdl_launch_dependents();
    pdl_wait_for(ends);                                     // ends[] come from the prepare kernel
    chunk_prefix<kBinThreads>(ends, n_samples, prefix);
    const int chunk = blockIdx.x;
    unsigned flags = 0;
    if (chunk == 0 && threadIdx.x == 0) {             // the layout itself: monotone, covers [0, n)
        int64_t prev = 0;
        for (int s = 0; s < n_samples; ++s) { if (ends[s] < prev) flags |= DVSOF_ST_NOT_GROUPED; prev = ends[s]; }
        if (prev != n) flags |= DVSOF_ST_BAD_SAMPLE;  // events with sample_index >= B (or < 0 at the front)
    }
    if (chunk >= prefix[n_samples]) { if (flags) atomicOr(status, flags); return; }
    if (threadIdx.x < 32) {                           // which sample owns this chunk
        for (int base = 0; base < n_samples; base += 32) {
            const int s = base + threadIdx.x;
            if (s < n_samples && prefix[s] <= chunk && chunk < prefix[s + 1]) s_sample = s;
        }
    }
    for (int k = threadIdx.x; k <= T; k += kBinThreads) hist[k] = 0;
    __syncthreads();
    const int s = s_sample;
    const int64_t s_begin = s ? ends[s - 1] : 0;
    const int64_t e0 = s_begin + (int64_t)(chunk - prefix[s]) * kBinChunk;
    const int64_t e1 = min(e0 + (int64_t)kBinChunk, ends[s]);
    constexpr int kPer = kBinChunk / kBinThreads;     // 16 events per thread
    constexpr int kFly = DVSOF_BIN_INFLIGHT;
    static_assert(kPer % kFly == 0, "whole load steps");
    unsigned packed[kPer];                            // rank | tile | code, 0xffffffff = dropped
    float tv[kPer];
    // Storage form, full chunk on a 16-event boundary of 16-byte aligned arrays: a thread takes two
    // groups of 8 CONSECUTIVE events with 16-byte loads (x, y: 8 x int16; t: 2 x float4; polarity:
    // 8 bytes).  The scalar loop below keeps one 2-byte load per lane in flight -- 64 B per warp
    // instruction, far too little outstanding traffic to cover the DRAM latency at 9 B/event.  The
    // order inside a run is free, so the other event-to-thread mapping changes nothing downstream.
    bool vec8 = false;
    if constexpr (kCompact) {
        static_assert(kPer == 16, "two groups of 8 events per thread");
        vec8 = e1 - e0 == kBinChunk && ((e0 & 15) == 0) &&
               ((reinterpret_cast<uintptr_t>(x_) | reinterpret_cast<uintptr_t>(y_) |
                 reinterpret_cast<uintptr_t>(t) | reinterpret_cast<uintptr_t>(p_)) & 15u) == 0;
        if (vec8) {                                   // CTA-uniform
            uint4 xv[2], yv[2];
            uint2 pv[2];
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                const int64_t eb = e0 + (int64_t)g * (kBinChunk / 2) + 8 * threadIdx.x;
                xv[g] = __ldcs(reinterpret_cast<const uint4*>(static_cast<const int16_t*>(x_) + eb));
                yv[g] = __ldcs(reinterpret_cast<const uint4*>(static_cast<const int16_t*>(y_) + eb));
                pv[g] = __ldcs(reinterpret_cast<const uint2*>(static_cast<const uint8_t*>(p_) + eb));
                const float4 ta = __ldcs(reinterpret_cast<const float4*>(t + eb));
                const float4 tb = __ldcs(reinterpret_cast<const float4*>(t + eb) + 1);
                tv[8 * g + 0] = ta.x; tv[8 * g + 1] = ta.y; tv[8 * g + 2] = ta.z; tv[8 * g + 3] = ta.w;
                tv[8 * g + 4] = tb.x; tv[8 * g + 5] = tb.y; tv[8 * g + 6] = tb.z; tv[8 * g + 7] = tb.w;
            }
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                const unsigned xw[4] = {xv[g].x, xv[g].y, xv[g].z, xv[g].w};
                const unsigned yw[4] = {yv[g].x, yv[g].y, yv[g].z, yv[g].w};
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int xi = (int)(short)((j & 1) ? (xw[j >> 1] >> 16) : (xw[j >> 1] & 0xffffu));
                    const int yi = (int)(short)((j & 1) ? (yw[j >> 1] >> 16) : (yw[j >> 1] & 0xffffu));
                    const unsigned pb = (((j < 4) ? pv[g].x : pv[g].y) >> (8 * (j & 3))) & 0xffu;
                    packed[8 * g + j] = 0xffffffffu;
                    if ((unsigned)xi >= (unsigned)W || (unsigned)yi >= (unsigned)H) { flags |= DVSOF_ST_BAD_PIXEL; continue; }
                    const unsigned yy = (unsigned)yi, tile = yy >> r_log2;
                    const unsigned code = (((yy & ((1u << r_log2) - 1u)) * (unsigned)W + (unsigned)xi) << 1) | (pb ? 0u : 1u);
                    const unsigned rank = atomicAdd(&hist[tile], 1u);
                    packed[8 * g + j] = (rank << 20) | (tile << 13) | code;
                }
            }
        }
    }
    if (!vec8)
#pragma unroll
    for (int k = 0; k < kPer; k += kFly) {            // kFly events in flight per thread and step
        int64_t xi[kFly], yi[kFly], pi[kFly], bi[kFly];
        bool ok[kFly];
#pragma unroll
        for (int j = 0; j < kFly; ++j) {
            const int64_t i = e0 + (int64_t)(k + j) * kBinThreads + threadIdx.x;
            ok[j] = i < e1;
            if (ok[j]) {
                if (kCompact) {
                    xi[j] = __ldcs(static_cast<const int16_t*>(x_) + i);
                    yi[j] = __ldcs(static_cast<const int16_t*>(y_) + i);
                    pi[j] = __ldcs(static_cast<const uint8_t*>(p_) + i) ? 1 : -1;
                    bi[j] = s;
                } else {
                    xi[j] = __ldcs(static_cast<const int64_t*>(x_) + i);
                    yi[j] = __ldcs(static_cast<const int64_t*>(y_) + i);
                    pi[j] = __ldcs(static_cast<const int64_t*>(p_) + i);
                    bi[j] = __ldcs(b + i);
                }
                tv[k + j] = __ldcs(t + i);
            }
        }
#pragma unroll
        for (int j = 0; j < kFly; ++j) {
            packed[k + j] = 0xffffffffu;
            if (!ok[j]) continue;
            if (pi[j] != 1 && pi[j] != -1) { flags |= DVSOF_ST_BAD_POLARITY; continue; }
            if (bi[j] != (int64_t)s) { flags |= DVSOF_ST_NOT_GROUPED; continue; }
            if ((uint64_t)xi[j] >= (uint64_t)W || (uint64_t)yi[j] >= (uint64_t)H) { flags |= DVSOF_ST_BAD_PIXEL; continue; }
            const unsigned yy = (unsigned)yi[j], tile = yy >> r_log2;
            const unsigned c = (unsigned)(1 - (int)pi[j]) >> 1;                   // net.py:142
            const unsigned code = (((yy & ((1u << r_log2) - 1u)) * (unsigned)W + (unsigned)xi[j]) << 1) | c;
            // ONE contended shared atomic per event: its return value is the event's rank inside
            // the tile's run (the order inside a run is free: sum and max commute)
            const unsigned rank = atomicAdd(&hist[tile], 1u);
            packed[k + j] = (rank << 20) | (tile << 13) | code;    // 12 + 7 + 13 bits (make_bin_plan)
        }
    }
    __syncthreads();
    if (threadIdx.x < 32) {                            // exclusive scan of the T tile counts
        unsigned carry = 0;
        for (int base = 0; base <= T; base += 32) {
            const int k = base + threadIdx.x;
            const unsigned cnt = k < T ? hist[k] : 0u;
            unsigned v = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned u = __shfl_up_sync(0xffffffffu, v, o);
                if ((int)threadIdx.x >= o) v += u;
            }
            const unsigned excl = carry + v - cnt;
            if (k <= T) {
                hist[k] = excl;                                      // hist[T] = total kept events
                if (k < T) cursor[k] = excl;
                table[(size_t)chunk * (T + 1) + k] = excl;
            }
            carry += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        if (packed[k] == 0xffffffffu) continue;
        const unsigned pos = cursor[(packed[k] >> 13) & 127u] + (packed[k] >> 20);
        codes_s[pos] = (uint16_t)(packed[k] & 0x1fffu);
        ts_s[pos] = tv[k];
    }
    __syncthreads();
    const unsigned total = hist[T];
    uint32_t* cg = reinterpret_cast<uint32_t*>(codes_g + (size_t)chunk * kBinChunk);
    const uint32_t* cs = reinterpret_cast<const uint32_t*>(codes_s);
    float* tg = ts_g + (size_t)chunk * kBinChunk;
    for (unsigned i = threadIdx.x; i < (total + 1) / 2; i += kBinThreads) cg[i] = cs[i];
    for (unsigned i = threadIdx.x; i < total; i += kBinThreads) tg[i] = ts_s[i];
    if (flags) atomicOr(status, flags);
}

// ---------------------------------------------------------------------------------------
// pass 2: one CTA per (sample, tile): shared-memory slab, then one coalesced write of the bands
// ---------------------------------------------------------------------------------------
#ifndef DVSOF_RED_MIN_CTAS
#define DVSOF_RED_MIN_CTAS 3
#endif
template <int kRedBatch>                       // records per lane and batch: 4 for runs of ~128, 1 for short runs (wide images)
__global__ void __launch_bounds__(kRedThreads, DVSOF_RED_MIN_CTAS)
encode_reduce_kernel(const uint16_t* __restrict__ codes_g, const float* __restrict__ ts_g,
                     const uint32_t* __restrict__ table, const int64_t* __restrict__ ends,
                     const float2* __restrict__ sd, int n_samples, int H, int W, int r_log2, int T,
                     float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned slab[];   // [cells] counts, then [cells] max keys
    __shared__ int prefix[kBinMaxSamples + 1];
    pdl_wait_for(codes_g, ts_g, table, ends, sd);                                       // records + tables come from pass 1
    const int s = blockIdx.x / T, tile = blockIdx.x - s * T;
    const int R = 1 << r_log2, cells = R * W * 2;
    unsigned* cnt = slab;
    int* key = reinterpret_cast<int*>(slab + cells);
    {   // cells is a multiple of 4 (W * 2 * R with W % 2 == 0 ... checked: falls back to scalar stores otherwise)
        if ((cells & 3) == 0) {
            uint4* c4 = reinterpret_cast<uint4*>(cnt);
            int4* k4 = reinterpret_cast<int4*>(key);
            for (int i = threadIdx.x; i < cells / 4; i += kRedThreads) {
                c4[i] = make_uint4(0u, 0u, 0u, 0u);
                k4[i] = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
            }
        } else {
            for (int i = threadIdx.x; i < cells; i += kRedThreads) { cnt[i] = 0u; key[i] = INT_MIN; }
        }
    }
    chunk_prefix<kRedThreads>(ends, n_samples, prefix);   // ends with __syncthreads()
    const int cb = prefix[s], nc = prefix[s + 1] - prefix[s];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // Runs are short (kBinChunk / T records, one batch of 32 x kRedBatch), so a warp's time is the
    // latency of one dependent trip to DRAM per run unless the next run is already in flight:
    // chunks are dealt to the warps round-robin (k-th chunk of warp w = k * kWarps + w: every warp
    // works even when a sample has few chunks), a lane holds the bounds of one of the warp's next
    // 32 chunks, and the batch after the current one is requested before the current one is
    // applied.  Everything that steers the loop is warp-uniform.
    constexpr int kWarps = kRedThreads / 32;
    constexpr unsigned kStep = 32 * (kRedBatch > 0 ? kRedBatch : 1);
    const int my_n = nc > warp ? (nc - warp + kWarps - 1) / kWarps : 0;
    if constexpr (kRedBatch < 0) {
        // Short runs, the lean way: no software pipelining at all -- 16 warps x 4 CTAs per SM are
        // enough loads in flight, and a run of ~34 records costs ~30 instructions instead of ~100.
        for (int k0 = 0; k0 < my_n; k0 += 32) {
            unsigned lo = 0, hi = 0;
            if (k0 + lane < my_n) {
                const uint32_t* row = table + (size_t)(cb + (k0 + lane) * kWarps + warp) * (T + 1) + tile;
                lo = __ldg(row); hi = __ldg(row + 1);
            }
            const int m = min(32, my_n - k0);
            const uint16_t* cg = codes_g + (size_t)(cb + k0 * kWarps + warp) * kBinChunk;
            const float* tg = ts_g + (size_t)(cb + k0 * kWarps + warp) * kBinChunk;
            constexpr size_t kStride = (size_t)kWarps * kBinChunk;       // records between two chunks of a warp
            auto apply = [&](unsigned c, float t) {
                if (c != 0xffffffffu) {
                    atomicAdd(&cnt[c], 1u);                      // scatter_sum, net.py:147-148
                    atomicMax(&key[c], ordered_key(t));          // scatter_max of the raw t, net.py:149
                }
            };
            // kLR runs x kLP records per lane in flight (short runs: 2 x 2 = 64 records cover 34 +- 6;
            // long runs of ~128: 2 x 5)
            constexpr int kLR = 2, kLP = kRedBatch == -1 ? 2 : -kRedBatch;
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
                for (int q = 0; q < kLR; ++q)                      // rare: a run longer than 32 kLP records
                    for (unsigned i = ra[q] + 32 * kLP + lane; __any_sync(0xffffffffu, i < rb[q]); i += 32)
                        if (i < rb[q]) apply(__ldcs(c0 + q * kStride + i), __ldcs(t0 + q * kStride + i));
            }
        }
    } else if constexpr (kRedBatch == 0) {
        // Short runs (wide images: kBinChunk / T ~ 34 records): FOUR runs per warp at a time, 8 lanes
        // and up to kGT records per lane each, the next four requested before the current ones are
        // applied.  A slab's CTA is a chain of DRAM round trips (table entry -> records -> atomics);
        // with one run in flight per warp it was 16-32 of them per warp (82 us for cfg4's 1920 slabs).
        constexpr int kGT = 6, kG = 4;                        // 48 records cover a run of 34 +- 6
        const int grp = lane >> 3, sub = lane & 7;
        for (int k0 = 0; k0 < my_n; k0 += 32) {
            unsigned lo = 0, hi = 0;
            if (k0 + lane < my_n) {
                const uint32_t* row = table + (size_t)(cb + (k0 + lane) * kWarps + warp) * (T + 1) + tile;
                lo = __ldg(row); hi = __ldg(row + 1);
            }
            const int m = min(32, my_n - k0);
            unsigned a = 0, b = 0, na = 0, nb = 0;             // record range of this group's run (current / next)
            size_t base = 0, nbase = 0;
            auto fetch = [&](int j0, unsigned& ra, unsigned& rb, size_t& rbase, unsigned* code, float* tv) {
                const int j = j0 + grp;
                ra = __shfl_sync(0xffffffffu, lo, j & 31);
                rb = __shfl_sync(0xffffffffu, hi, j & 31);
                if (j >= m) rb = ra;
                rbase = (size_t)(cb + (k0 + j) * kWarps + warp) * kBinChunk;
#pragma unroll
                for (int u = 0; u < kGT; ++u) {
                    const unsigned i = ra + u * 8 + sub;
                    code[u] = 0xffffffffu;
                    if (i < rb) { code[u] = __ldcs(codes_g + rbase + i); tv[u] = __ldcs(ts_g + rbase + i); }
                }
            };
            auto apply = [&](unsigned c, float t) {
                if (c != 0xffffffffu) {
                    atomicAdd(&cnt[c], 1u);                      // scatter_sum, net.py:147-148
                    atomicMax(&key[c], ordered_key(t));          // scatter_max of the raw t, net.py:149
                }
            };
            unsigned code[kGT], ncode[kGT];
            float tv[kGT], ntv[kGT];
            fetch(0, a, b, base, code, tv);
            for (int j0 = 0; j0 < m; j0 += kG) {
                if (j0 + kG < m) fetch(j0 + kG, na, nb, nbase, ncode, ntv);
#pragma unroll
                for (int u = 0; u < kGT; ++u) apply(code[u], tv[u]);
                for (unsigned i = a + kGT * 8 + sub; __any_sync(0xffffffffu, i < b); i += 8)   // rare: a run above 48 records
                    if (i < b) apply(__ldcs(codes_g + base + i), __ldcs(ts_g + base + i));
#pragma unroll
                for (int u = 0; u < kGT; ++u) { code[u] = ncode[u]; tv[u] = ntv[u]; }
                a = na; b = nb; base = nbase;
            }
        }
    } else {
    for (int k0 = 0; k0 < my_n; k0 += 32) {
        unsigned lo = 0, hi = 0;
        if (k0 + lane < my_n) {
            const uint32_t* row = table + (size_t)(cb + (k0 + lane) * kWarps + warp) * (T + 1) + tile;
            lo = __ldg(row); hi = __ldg(row + 1);
        }
        const int m = min(32, my_n - k0);
        int j = -1;
        unsigned i0 = 0, h = 0;
        auto advance = [&]() {                            // next batch: same run, or the next non-empty one
            i0 += kStep;
            while (i0 >= h && ++j < m) { i0 = __shfl_sync(0xffffffffu, lo, j); h = __shfl_sync(0xffffffffu, hi, j); }
        };
        auto load = [&](unsigned* code, float* tv) {
            const size_t base = (size_t)(cb + (k0 + j) * kWarps + warp) * kBinChunk;
#pragma unroll
            for (int u = 0; u < kRedBatch; ++u) {
                const unsigned i = i0 + u * 32 + lane;
                code[u] = 0xffffffffu;
                if (i < h) { code[u] = __ldcs(codes_g + base + i); tv[u] = __ldcs(ts_g + base + i); }
            }
        };
        unsigned code[kRedBatch], ncode[kRedBatch];
        float tv[kRedBatch], ntv[kRedBatch];
        advance();
        if (j < m) load(code, tv);
        while (j < m) {
            advance();
            if (j < m) load(ncode, ntv);
#pragma unroll
            for (int u = 0; u < kRedBatch; ++u) {
                if (code[u] != 0xffffffffu) {
                    atomicAdd(&cnt[code[u]], 1u);                // scatter_sum, net.py:147-148
                    atomicMax(&key[code[u]], ordered_key(tv[u]));   // scatter_max of the raw t, net.py:149
                }
                code[u] = ncode[u]; tv[u] = ntv[u];
            }
        }
    }
    }
    __syncthreads();
    const float2 nd = sd[s];
    const int y0 = tile << r_log2;
    const size_t plane = (size_t)H * W;
    float* o = out + (size_t)s * 4 * plane;
    // the slab's rows are whole image rows: its pixels are contiguous in every output plane
    const int n_px = min(R, H - y0) * W;
    float* o0 = o + (size_t)y0 * W;
    for (int i = threadIdx.x; i < n_px; i += kRedThreads) {
        const size_t px = (size_t)i;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int k = key[2 * i + c];
            float tn = 0.f;
            if (k != INT_MIN) {
                tn = __fdiv_rn(__fsub_rn(key_to_float(k), nd.x), nd.y);          // net.py:141
                if (__float_as_int(tn) <= 0) tn = 0.f;                           // max against the zero plane
            }
            o0[c * plane + px] = (float)cnt[2 * i + c];
            o0[(c + 2) * plane + px] = tn;
        }
    }
}

}  // namespace dvsof
