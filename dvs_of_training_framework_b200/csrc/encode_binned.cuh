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
//                    as 6-byte records {u16 cell-in-tile, s32 order-preserving key of t} -- contiguous per (chunk, tile)
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
#define DVSOF_RED_BATCH 4
#endif
#ifndef DVSOF_RED_BATCH_SHORT
#define DVSOF_RED_BATCH_SHORT 2
#endif
#ifndef DVSOF_RED_RUNS
#define DVSOF_RED_RUNS 2
#endif
constexpr int kRedBatchLong = DVSOF_RED_BATCH;          // pass 2: records per lane and run in flight (runs of ~128 records)
constexpr int kRedBatchShort = DVSOF_RED_BATCH_SHORT;   // the same for short runs (wide images: ~34 records)
constexpr int kRedRuns = DVSOF_RED_RUNS;                // runs in flight per warp
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
        ts_s[pos] = __int_as_float(ordered_key(tv[k]));      // pass 2 takes the maximum of the keys as they are
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
// pass 2: one CTA per (sample, tile): shared-memory slab of {u32 count} and {s32 max key} per cell,
// then one coalesced write of the four plane bands.
// The chunks of the sample are dealt to the 16 warps round-robin (k-th chunk of warp w = k * 16 + w:
// every warp works even when a sample has few chunks); a lane holds the run bounds of one of the warp's
// next 32 chunks; kLR runs x kLP records per lane are requested, then applied -- no software pipeline,
// 16 warps x 3 CTAs per SM keep enough loads in flight.  Written for the instruction count (the first
// version of this loop compiled to 53 instructions per record under its 42-register budget: the
// shared-memory base and the key array's offset re-derived per record, a branch around each pair of
// atomics, two 64-bit address computations per load, ten moves per batch of the software pipeline):
//   * the key array sits at a FIXED distance behind the counts (kKeyOffsetBytes): one address, one
//     immediate offset;
//   * both atomics are unconditional (slab_apply), no divergence regions;
//   * a run is addressed once; its records are 32 apart at immediate offsets; pass 1 delivers the key.
// What bounds it (profiles/r02zj_*): the shared-memory atomics themselves.  With the record loads
// replaced by hashed cell numbers the kernel still takes 206 of its 286 us at cfg3 (5.3 lane-atomics
// per clock and SM; scripts/ubench_atomics.cu reaches 11 in isolation with 64 warps per SM); forcing
// every lane of a warp onto its own bank gains 10 %; 40 % fewer instructions, splitting each slab over
// a cluster of 2-4 CTAs against the 2.3-round tail, more or fewer records in flight: no change.
// ---------------------------------------------------------------------------------------
#ifndef DVSOF_RED_MIN_CTAS
#define DVSOF_RED_MIN_CTAS 3
#endif
constexpr unsigned kKeyOffsetBytes = kBinSlabBytesMax / 2;       // cells <= 8192 counts of 4 bytes
// at least 32 cells: lanes without a record touch cell `lane` (slab_apply)
inline size_t reduce_smem_bytes(int cells) { return (size_t)kKeyOffsetBytes + (size_t)(cells < 32 ? 32 : cells) * 4; }

// One record into the slab.  EVERY lane executes both atomics: ptxas wraps a predicated shared atomic
// in a divergence region (BSSY / BRA / BSYNC: four instructions per atomic), so a lane without a
// record adds 0 to, and maximises INT_MIN into, cell `lane` -- no-ops on 32 distinct banks.
__device__ __forceinline__ void slab_apply(unsigned* cnt, unsigned code, int key, unsigned inc) {
    unsigned* a = cnt + code;
    atomicAdd(a, inc);                                                       // scatter_sum, net.py:147-148
    atomicMax(reinterpret_cast<int*>(a + kKeyOffsetBytes / 4), key);         // scatter_max (key of the raw t), net.py:149
}

template <int kLP>                                   // records per lane and run in flight (kLR runs at a time)
__global__ void __launch_bounds__(kRedThreads, DVSOF_RED_MIN_CTAS)
encode_reduce_kernel(const uint16_t* __restrict__ codes_g, const float* __restrict__ ts_g,
                     const uint32_t* __restrict__ table, const int64_t* __restrict__ ends,
                     const float2* __restrict__ sd, int n_samples, int H, int W, int r_log2, int T,
                     float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned slab[];   // [cells] counts ... [kKeyOffsetBytes:] [cells] max keys
    __shared__ int prefix[kBinMaxSamples + 1];
    const int s = blockIdx.x / T, tile = blockIdx.x - s * T;
    const int R = 1 << r_log2, cells = R * W * 2;
    unsigned* cnt = slab;
    int* key = reinterpret_cast<int*>(slab + kKeyOffsetBytes / 4);
    if ((cells & 3) == 0) {                            // private: zeroed while pass 1 still runs
        uint4* c4 = reinterpret_cast<uint4*>(cnt);
        int4* k4 = reinterpret_cast<int4*>(key);
        for (int i = threadIdx.x; i < cells / 4; i += kRedThreads) {
            c4[i] = make_uint4(0u, 0u, 0u, 0u);
            k4[i] = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
        }
    } else {
        for (int i = threadIdx.x; i < cells; i += kRedThreads) { cnt[i] = 0u; key[i] = INT_MIN; }
    }
    pdl_wait_for(codes_g, ts_g, table, ends, sd);      // records + tables come from pass 1
    chunk_prefix<kRedThreads>(ends, n_samples, prefix);   // ends with __syncthreads()
    const int cb = prefix[s], nc = prefix[s + 1] - prefix[s];
    const int lane = threadIdx.x & 31;
    constexpr int kWarps = kRedThreads / 32;
    const int warp = threadIdx.x >> 5;
    constexpr int kLR = kRedRuns;
    constexpr size_t kStride = (size_t)kWarps * kBinChunk;        // records between two chunks of a warp
    const int my_n = nc > warp ? (nc - warp + kWarps - 1) / kWarps : 0;
    const int* keys_g = reinterpret_cast<const int*>(ts_g);
    for (int k0 = 0; k0 < my_n; k0 += 32) {
        unsigned lo = 0, hi = 0;
        if (k0 + lane < my_n) {
            const uint32_t* row = table + (size_t)(cb + (k0 + lane) * kWarps + warp) * (T + 1) + tile;
            lo = __ldg(row); hi = __ldg(row + 1);
        }
        const int m = min(32, my_n - k0);
        const size_t base0 = (size_t)(cb + k0 * kWarps + warp) * kBinChunk + lane;
        for (int j = 0; j < m; j += kLR) {
            const uint16_t* cp[kLR];
            const int* kp[kLR];
            int left[kLR];                               // records of the run from this lane's first one on
#pragma unroll
            for (int q = 0; q < kLR; ++q) {
                const unsigned a = __shfl_sync(0xffffffffu, lo, (j + q) & 31);
                const unsigned b = __shfl_sync(0xffffffffu, hi, (j + q) & 31);
                const size_t first = base0 + (size_t)(j + q) * kStride + a;
                cp[q] = codes_g + first;
                kp[q] = keys_g + first;
                left[q] = (j + q < m) ? (int)(b - a) - lane : 0;
            }
            unsigned code[kLR * kLP];
            int kv[kLR * kLP];
#pragma unroll
            for (int q = 0; q < kLR; ++q)
#pragma unroll
                for (int u = 0; u < kLP; ++u) {
                    code[q * kLP + u] = (unsigned)lane; kv[q * kLP + u] = INT_MIN;
                    if (u * 32 < left[q]) {
                        code[q * kLP + u] = __ldcs(cp[q] + u * 32);
                        kv[q * kLP + u] = __ldcs(kp[q] + u * 32);
                    }
                }
#pragma unroll
            for (int q = 0; q < kLR; ++q)
#pragma unroll
                for (int u = 0; u < kLP; ++u) slab_apply(cnt, code[q * kLP + u], kv[q * kLP + u], u * 32 < left[q] ? 1u : 0u);
#pragma unroll
            for (int q = 0; q < kLR; ++q)                    // a run longer than 32 kLP records
                for (int off = 32 * kLP; __any_sync(0xffffffffu, off < left[q]); off += 32) {
                    const bool v = off < left[q];
                    unsigned c = (unsigned)lane; int k = INT_MIN;
                    if (v) { c = __ldcs(cp[q] + off); k = __ldcs(kp[q] + off); }
                    slab_apply(cnt, c, k, v ? 1u : 0u);
                }
        }
    }
    __syncthreads();
    const float2 nd = sd[s];
    const int y0 = tile << r_log2;
    const size_t plane = (size_t)H * W;
    const int n_px = min(R, H - y0) * W;               // whole image rows: contiguous in every output plane
    float* o0 = out + (size_t)s * 4 * plane + (size_t)y0 * W;
    for (int i = threadIdx.x; i < n_px; i += kRedThreads) {
        const uint2 c2 = *reinterpret_cast<const uint2*>(cnt + 2 * i);
        const int2 k2 = *reinterpret_cast<const int2*>(key + 2 * i);
        const unsigned cc[2] = {c2.x, c2.y};
        const int kk[2] = {k2.x, k2.y};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            float tn = 0.f;
            if (kk[c] != INT_MIN) {
                tn = __fdiv_rn(__fsub_rn(key_to_float(kk[c]), nd.x), nd.y);          // net.py:141
                if (__float_as_int(tn) <= 0) tn = 0.f;                               // max against the zero plane
            }
            o0[c * plane + i] = (float)cc[c];
            o0[(c + 2) * plane + i] = tn;
        }
    }
}

}  // namespace dvsof
