// encode.cu -- event stream -> 4-channel event image (counts + last normalised timestamp per
// polarity).  Replaces EV_FlowNet/net.py:88-150 of the reference (see include/dvsof.h).
//
// Data layout in HBM: events are structure-of-arrays exactly as the reference's collate
// delivers them (x,y,polarity,sample_index int64; timestamp fp32 => 36 B/event), streamed once
// with 16-byte evict-first loads.  The output (B,4,H,W) fp32 is zeroed, then updated with
// no-return L2 reductions: RED.ADD.F32 for the counts (exact: integer-valued fp32 sums) and
// RED.MAX.S32 on the bit pattern for the timestamps (signed-int order == float order for
// t^ >= 0, and any negative t^ loses against the initial +0.0, which is exactly
// scatter_max(..., out=zeros)).  Both are order independent => bit-exact vs the oracle.
#include <stdlib.h>
#include "common.cuh"
#include "encode_binned.cuh"
#include "voxel.cuh"

namespace dvsof {

// ---------------------------------------------------------------------------------------
// prepare: zero the event image AND find per-sample (start, dt) in ONE launch.
// CTA s (mod grid) runs a 256-ary upper-bound search in sample_index for sample s (3 dependent
// DRAM round trips at 1 M events, 4 at 160 M -- the old warp-wide version needed 2 x 4..6), the
// rest of the grid streams zeros with 16 B stores meanwhile.  Dependent kernels are released
// right away (programmatic dependent launch) and block in griddepcontrol.wait before they touch
// `out` / `sd`, so their own event loads overlap this kernel.
// ---------------------------------------------------------------------------------------

constexpr int kPrepThreads = 256;

// first index i in [0,n) with b[i] > s (n if none); all kPrepThreads threads of the CTA cooperate
__device__ __forceinline__ int64_t cta_upper_bound(const int64_t* __restrict__ b, int64_t n,
                                                   int64_t s, int* warp_first) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t step = (hi - lo + kPrepThreads - 1) / kPrepThreads;
        const int64_t pos = lo + (int64_t)(tid + 1) * step - 1;
        const bool gt = (pos < hi) ? (__ldg(b + pos) > s) : true;
        const unsigned m = __ballot_sync(0xffffffffu, gt);
        if (lane == 0) warp_first[warp] = m ? warp * 32 + (__ffs(m) - 1) : kPrepThreads;
        __syncthreads();
        int first = kPrepThreads;
#pragma unroll
        for (int w = 0; w < kPrepThreads / 32; ++w) first = min(first, warp_first[w]);
        __syncthreads();
        const int64_t cand = lo + (int64_t)(first + 1) * step - 1;   // first probe that is > s
        lo = lo + (int64_t)first * step;
        hi = cand < hi ? cand : hi;
    }
    return lo;
}

__global__ void __launch_bounds__(kPrepThreads)
encode_prepare_kernel(const int64_t* __restrict__ b, const float* __restrict__ t, int64_t n,
                      const float* __restrict__ timestamps, int n_samples,
                      float2* __restrict__ sd, int64_t* __restrict__ ends,
                      float* __restrict__ out, size_t n_floats) {
    __shared__ int warp_first[kPrepThreads / 32];
    pdl_launch_dependents();
    if (n > 0) {
        for (int s = blockIdx.x; s < n_samples; s += gridDim.x) {
            const int64_t end = cta_upper_bound(b, n, (int64_t)s, warp_first);
            if (threadIdx.x == 0) {
                const float start = timestamps[2 * s];
                float dt = 0.f;
                if (end > 0 && b[end - 1] == (int64_t)s) dt = __fsub_rn(t[end - 1], start);  // net.py:136-137
                sd[s] = make_float2(start, fmaxf(dt, 1e-9f));                                  // net.py:138
                ends[s] = end;                      // exclusive end of sample s in the event stream
            }
        }
    }
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    if ((reinterpret_cast<uintptr_t>(out) & 15u) == 0) {
        float4* o4 = reinterpret_cast<float4*>(out);
        const size_t n4 = n_floats >> 2;
        for (size_t i = tid; i < n4; i += stride) o4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (size_t i = (n4 << 2) + tid; i < n_floats; i += stride) out[i] = 0.f;
    } else {
        for (size_t i = tid; i < n_floats; i += stride) out[i] = 0.f;
    }
}

__global__ void sample_bounds_compact_kernel(const int64_t* __restrict__ sample_begin,
                                             const float* __restrict__ t, int64_t n,
                                             const float* __restrict__ timestamps, int n_samples,
                                             float2* __restrict__ sd,
                                             uint32_t* __restrict__ status) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_samples) return;
    int64_t begin = sample_begin[s], end = sample_begin[s + 1];
    if (begin < 0 || end < begin || end > n || (s == 0 && begin != 0) ||
        (s == n_samples - 1 && end != n)) {
        atomicOr(status, DVSOF_ST_NOT_GROUPED);
        begin = end = 0;
    }
    const float start = timestamps[2 * s];
    float dt = 0.f;
    if (end > begin) dt = __fsub_rn(t[end - 1], start);
    sd[s] = make_float2(start, fmaxf(dt, 1e-9f));
}

// ---------------------------------------------------------------------------------------
// scatter of one event
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned scatter_event(int64_t xi, int64_t yi, float ti, int64_t pi,
                                                  int64_t bi, const float2* __restrict__ sd,
                                                  int n_samples, int H, int W,
                                                  float* __restrict__ out) {
    // validate in 64 bit, then continue in 32 bit (fewer live registers per event in flight)
    if (pi != 1 && pi != -1) return DVSOF_ST_BAD_POLARITY;
    if ((uint64_t)bi >= (uint64_t)n_samples) return DVSOF_ST_BAD_SAMPLE;
    if ((uint64_t)xi >= (uint64_t)W || (uint64_t)yi >= (uint64_t)H) return DVSOF_ST_BAD_PIXEL;
    const unsigned x32 = (unsigned)xi, y32 = (unsigned)yi, b32 = (unsigned)bi;
    const unsigned c = (unsigned)(1 - (int)pi) >> 1;                 // net.py:142
    const float2 s = __ldg(sd + b32);
    const float tn = __fdiv_rn(__fsub_rn(ti, s.x), s.y);           // net.py:141
    const size_t plane = (size_t)H * W;
    float* cell = out + (size_t)(b32 * 4u + c) * plane + (y32 * (unsigned)W + x32);  // net.py:144
    atomicAdd(cell, 1.0f);                                          // scatter_sum, net.py:147
    atomicMax(reinterpret_cast<int*>(cell + 2 * plane), __float_as_int(tn));  // scatter_max, net.py:149
    return 0u;
}

// ---------------------------------------------------------------------------------------
// int64 structure-of-arrays stream, 4 events per thread per iteration (16 B loads)
// ---------------------------------------------------------------------------------------
template <int MINB>
__global__ void __launch_bounds__(256, MINB)
encode_vec4_kernel(const int64_t* __restrict__ x, const int64_t* __restrict__ y,
                   const float* __restrict__ t, const int64_t* __restrict__ p,
                   const int64_t* __restrict__ b, int64_t n, const float2* __restrict__ sd,
                   int n_samples, int H, int W, float* __restrict__ out,
                   uint32_t* __restrict__ status) {
    const int lane = threadIdx.x & 31;
    const int64_t nvec = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned flags = 0;
    bool waited = false;
    // warp-uniform trip count so the shuffle below is always convergent
    for (int64_t base = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < nvec;
         base += stride) {
        const int64_t i = base + lane;
        const bool valid = i < nvec;
        longlong2 xa, xb, ya, yb, pa, pb, ba, bb;
        float4 tt;
        if (valid) {
            const longlong2* x2 = reinterpret_cast<const longlong2*>(x) + 2 * i;
            const longlong2* y2 = reinterpret_cast<const longlong2*>(y) + 2 * i;
            const longlong2* p2 = reinterpret_cast<const longlong2*>(p) + 2 * i;
            const longlong2* b2 = reinterpret_cast<const longlong2*>(b) + 2 * i;
            xa = __ldcs(x2); xb = __ldcs(x2 + 1);
            ya = __ldcs(y2); yb = __ldcs(y2 + 1);
            pa = __ldcs(p2); pb = __ldcs(p2 + 1);
            ba = __ldcs(b2); bb = __ldcs(b2 + 1);
            tt = __ldcs(reinterpret_cast<const float4*>(t) + i);
        } else {
            ba = bb = make_longlong2(0, 0);
        }
        // grouping check (net.py:131-134): sample_index must be non-decreasing
        long long bnext = __shfl_down_sync(0xffffffffu, ba.x, 1);
        if (!waited) { pdl_wait_for(sd); waited = true; }   // `out` zeroed + `sd` ready from here on
        if (valid) {
            const int64_t nxt = 4 * (i + 1);
            if (lane == 31 || i + 1 >= nvec) bnext = (nxt < n) ? b[nxt] : bb.y;
            if (ba.x > ba.y || ba.y > bb.x || bb.x > bb.y || bb.y > bnext)
                flags |= DVSOF_ST_NOT_GROUPED;
            flags |= scatter_event(xa.x, ya.x, tt.x, pa.x, ba.x, sd, n_samples, H, W, out);
            flags |= scatter_event(xa.y, ya.y, tt.y, pa.y, ba.y, sd, n_samples, H, W, out);
            flags |= scatter_event(xb.x, yb.x, tt.z, pb.x, bb.x, sd, n_samples, H, W, out);
            flags |= scatter_event(xb.y, yb.y, tt.w, pb.y, bb.y, sd, n_samples, H, W, out);
        }
    }
    if (!waited) pdl_wait_for(sd);
    // tail (n % 4 events): first threads of block 0
    if (blockIdx.x == 0) {
        const int64_t i = (nvec << 2) + threadIdx.x;
        if (i < n) {
            if (i + 1 < n && b[i] > b[i + 1]) flags |= DVSOF_ST_NOT_GROUPED;
            flags |= scatter_event(x[i], y[i], t[i], p[i], b[i], sd, n_samples, H, W, out);
        }
    }
    if (flags) atomicOr(status, flags);
}

// unaligned views (e.g. a DDP shard that starts at an odd event index): scalar loads
__global__ void __launch_bounds__(256)
encode_scalar_kernel(const int64_t* __restrict__ x, const int64_t* __restrict__ y,
                     const float* __restrict__ t, const int64_t* __restrict__ p,
                     const int64_t* __restrict__ b, int64_t n, const float2* __restrict__ sd,
                     int n_samples, int H, int W, float* __restrict__ out,
                     uint32_t* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned flags = 0;
    pdl_wait_for(sd);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int64_t bi = __ldcs(b + i);
        if (i + 1 < n && bi > b[i + 1]) flags |= DVSOF_ST_NOT_GROUPED;
        flags |= scatter_event(__ldcs(x + i), __ldcs(y + i), __ldcs(t + i), __ldcs(p + i), bi, sd,
                               n_samples, H, W, out);
    }
    if (flags) atomicOr(status, flags);
}

// ---------------------------------------------------------------------------------------
// compact storage form (utils/dataset.py:286-289): x,y int16, polarity bool, t fp32 = 9 B/event;
// sample membership from the per-sample prefix `sample_begin` (B+1).  8 events per thread.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned scatter_compact(int xi, int yi, float ti, int c, int s,
                                                    const float2* __restrict__ sd, int H, int W,
                                                    float* __restrict__ out) {
    if ((unsigned)xi >= (unsigned)W || (unsigned)yi >= (unsigned)H) return DVSOF_ST_BAD_PIXEL;
    const float2 v = __ldg(sd + s);
    const float tn = __fdiv_rn(__fsub_rn(ti, v.x), v.y);
    const size_t plane = (size_t)H * W;
    const size_t cell = ((size_t)s * 4 + c) * plane + (size_t)yi * W + xi;
    atomicAdd(out + cell, 1.0f);
    atomicMax(reinterpret_cast<int*>(out + cell + 2 * plane), __float_as_int(tn));
    return 0u;
}

__device__ __forceinline__ int find_sample(const int64_t* __restrict__ sample_begin, int n_samples,
                                           int64_t e) {
    // largest s with sample_begin[s] <= e
    int lo = 0, hi = n_samples;  // answer in [lo, hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(sample_begin + mid) <= e) lo = mid; else hi = mid;
    }
    return lo;
}

template <bool VEC>
__global__ void __launch_bounds__(256)
encode_compact_kernel(const int16_t* __restrict__ x, const int16_t* __restrict__ y,
                      const float* __restrict__ t, const uint8_t* __restrict__ p,
                      const int64_t* __restrict__ sample_begin, int64_t n,
                      const float2* __restrict__ sd, int n_samples, int H, int W,
                      float* __restrict__ out, uint32_t* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned flags = 0;
    int64_t done = 0;
    if (VEC) {
        const int64_t nvec = n >> 3;
        done = nvec << 3;
        for (int64_t i = tid; i < nvec; i += stride) {
            const int4 xv = __ldcs(reinterpret_cast<const int4*>(x) + i);
            const int4 yv = __ldcs(reinterpret_cast<const int4*>(y) + i);
            const uint2 pv = __ldcs(reinterpret_cast<const uint2*>(p) + i);
            const float4 ta = __ldcs(reinterpret_cast<const float4*>(t) + 2 * i);
            const float4 tb = __ldcs(reinterpret_cast<const float4*>(t) + 2 * i + 1);
            const int xw[4] = {xv.x, xv.y, xv.z, xv.w};
            const int yw[4] = {yv.x, yv.y, yv.z, yv.w};
            const float tv[8] = {ta.x, ta.y, ta.z, ta.w, tb.x, tb.y, tb.z, tb.w};
            int64_t e = i << 3;
            int s = find_sample(sample_begin, n_samples, e);
            int64_t next = __ldg(sample_begin + s + 1);
#pragma unroll
            for (int k = 0; k < 8; ++k, ++e) {
                while (e >= next && s + 1 < n_samples) { ++s; next = __ldg(sample_begin + s + 1); }
                const int xi = (int)(short)((k & 1) ? (xw[k >> 1] >> 16) : (xw[k >> 1] & 0xffff));
                const int yi = (int)(short)((k & 1) ? (yw[k >> 1] >> 16) : (yw[k >> 1] & 0xffff));
                const unsigned pw = (k < 4) ? pv.x : pv.y;
                const int pos = (pw >> (8 * (k & 3))) & 0xff;
                flags |= scatter_compact(xi, yi, tv[k], pos ? 0 : 1, s, sd, H, W, out);
            }
        }
    }
    for (int64_t e = done + tid; e < n; e += stride) {
        const int s = find_sample(sample_begin, n_samples, e);
        flags |= scatter_compact(x[e], y[e], t[e], p[e] ? 0 : 1, s, sd, H, W, out);
    }
    if (flags) atomicOr(status, flags);
}

}  // namespace dvsof

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
using namespace dvsof;

static int encode_grid(int64_t work_items, int threads, int ctas_per_sm) {
    int64_t blocks = (work_items + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * ctas_per_sm;  // persistent: multiple of the SM count
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

struct EncodeArgs {
    const int64_t *x, *y; const float* t; const int64_t *p, *b; int64_t n;
    const float2* sd; int B, H, W; float* out; uint32_t* status;
    // the storage form (binned path only): x, y int16, polarity uint8; ends = sample_begin + 1
    const void *cx = nullptr, *cy = nullptr, *cp = nullptr; const int64_t* cends = nullptr;
};

template <typename K>
static int launch_stream(K kern, const char* name, int grid, int threads, size_t smem,
                         cudaStream_t stream, bool pdl, const EncodeArgs& a) {
    return launch_kernel(name, kern, grid, threads, smem, stream, pdl, a.x, a.y, a.t, a.p, a.b, a.n,
                         a.sd, a.B, a.H, a.W, a.out, a.status);
}

static int check_encode_args(int64_t n, int64_t B, int H, int W, const void* out, const void* ws,
                             const void* status) {
    if (n < 0 || B <= 0 || H <= 0 || W <= 0)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: bad sizes n=%lld B=%lld H=%d W=%d",
                    (long long)n, (long long)B, H, W);
    if (B > INT32_MAX / 8) return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: too many samples");
    if (!out || !ws || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: out/workspace/status must be non-null");
    return DVSOF_OK;
}

// which path: DVSOF_ENCODE_PATH=direct|binned forces one (tests); default: binned for streams
// large and dense enough to pay for the second pass (>= 4 M events, >= 8 events per output cell)
static int encode_path_override() {   // read per call: tests switch it at run time
    const char* e = getenv("DVSOF_ENCODE_PATH");
    if (!e) return 0;
    return e[0] == 'd' ? 1 : (e[0] == 'b' ? 2 : 0);
}

static bool use_binned(const BinPlan& plan, int64_t n, int64_t B, int H, int W, size_t workspace_bytes) {
    if (!plan.ok || workspace_bytes < plan.total) return false;
    const int ov = encode_path_override();
    if (ov == 1) return false;
    if (ov == 2) return true;
    // measured (profiles/r01bz_binned_launches_cfg3.csv, cfg3: 160 M events on 32 x 256^2): bin 1.02 ms +
    // reduce 0.31 ms vs 1.96 ms direct.  The records cost 12 B/event of extra traffic and pass 2 runs
    // B * T CTAs, so it pays for dense streams: from ~11 events per pixel with 8 samples of 256^2
    // (256 pass-2 CTAs; profiles/r01by_encoder_size_sweep_direct_vs_binned.jsonl), and already at 3.4
    // events per pixel on cfg4 (16 x 480 x 640: 1920 pass-2 CTAs, 217 us vs 234 us direct;
    // profiles/r01cb_binned_launches_cfg4.csv).
    const int64_t pixels = B * (int64_t)H * W;
    const bool many_ctas = B * (int64_t)plan.T >= 2 * 3 * kNumSMs;     // two full waves of pass 2
    return n >= (4 << 20) && n >= (many_ctas ? 3 : 8) * pixels;
}

extern "C" size_t dvsof_encode_workspace_bytes(int64_t n_events, int64_t n_samples, int32_t H, int32_t W) {
    return make_bin_plan(n_events, n_samples, H, W).total;
}

static int launch_prepare(const float* t, const int64_t* b, int64_t n, const float* timestamps,
                          int64_t B, int32_t H, int32_t W, float* out, void* workspace, bool zero,
                          cudaStream_t stream) {
    const BinPlan plan = make_bin_plan(0, B, H, W);      // sd | ends offsets only
    const size_t n_floats = zero ? (size_t)B * 4 * H * W : 0;
    // enough CTAs to stream the zeros at full rate, at least one per sample for the search
    int64_t grid = (int64_t)((n_floats / 4 + kPrepThreads * 8 - 1) / (kPrepThreads * 8));
    if (grid > kNumSMs * 8) grid = kNumSMs * 8;
    if (grid < B) grid = B < kNumSMs * 8 ? B : kNumSMs * 8;
    if (grid < 1) grid = 1;
    return launch_kernel("encode_prepare_kernel", encode_prepare_kernel, (int)grid, kPrepThreads, 0,
                         stream, false, b, t, n, timestamps, (int)B, reinterpret_cast<float2*>(workspace),
                         reinterpret_cast<int64_t*>(static_cast<char*>(workspace) + plan.off_ends),
                         out, n_floats);                 // n == 0: zeros only, net.py:115-116
}

extern "C" int dvsof_encode_prepare(const float* t, const int64_t* b, int64_t n,
                                   const float* timestamps, int64_t B, int32_t H, int32_t W,
                                   float* out, void* workspace, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, workspace)) return rc;
    if (n > 0 && (!t || !b || !timestamps))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: null event array");
    return launch_prepare(t, b, n, timestamps, B, H, W, out, workspace, true, (cudaStream_t)stream_);
}

static int encode_scatter_impl(const EncodeArgs& a, cudaStream_t stream, bool pdl) {
    const bool vec = aligned16(a.x) && aligned16(a.y) && aligned16(a.t) && aligned16(a.p) && aligned16(a.b);
    if (!vec)
        return launch_stream(encode_scalar_kernel, "encode_scalar_kernel", encode_grid(a.n, 256, 8),
                             256, 0, stream, pdl, a);
    // up to 128 CTAs per SM in the grid (3 resident at a time), i.e. one pass of 4 events per thread
    // for anything below ~78 M events: fresh CTAs keep the load phase of one and the RED phase of
    // another interleaved.  Measured (profiles/r01h_variants.jsonl): cfg4 54.6 -> 76.4 Gev/s and
    // cfg3 78.0 -> 81.8 Gev/s over an 8-per-SM persistent grid.  The other schedules tried in round 1
    // (2 events per thread, a register software pipeline, cp.async.bulk chunk rings: same RED-bound
    // throughput, profiles/r01b_variants.jsonl) left the product library with commit "r02: experiment
    // kernels out of libdvsof"; they are in the history of this file.
    return launch_stream(encode_vec4_kernel<3>, "encode_vec4_kernel<3>",
                         encode_grid((a.n >> 2) + 1, 256, 128), 256, 0, stream, pdl, a);
}

extern "C" int dvsof_encode_scatter(const int64_t* x, const int64_t* y, const float* t,
                                   const int64_t* p, const int64_t* b, int64_t n, int64_t B,
                                   int32_t H, int32_t W, float* out, const void* workspace,
                                   uint32_t* status, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, status)) return rc;
    if (n == 0) return DVSOF_OK;
    if (!x || !y || !t || !p || !b) return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: null event array");
    const EncodeArgs a{x, y, t, p, b, n, reinterpret_cast<const float2*>(workspace), (int)B, H, W, out, status};
    return encode_scatter_impl(a, (cudaStream_t)stream_, false);
}

static int encode_binned_impl(const EncodeArgs& a, const BinPlan& plan, void* workspace, cudaStream_t stream,
                              int passes);

extern "C" int dvsof_encode_event_image(const int64_t* x, const int64_t* y, const float* t,
                                        const int64_t* p, const int64_t* b, int64_t n,
                                        const float* timestamps, int64_t B, int32_t H, int32_t W,
                                        float* out, void* workspace, size_t workspace_bytes,
                                        uint32_t* status, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, status)) return rc;
    if (n > 0 && (!x || !y || !p || !t || !b || !timestamps))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: null event array");
    if (workspace_bytes < make_bin_plan(0, B, H, W).total)
        return fail(DVSOF_ERR_WORKSPACE, "encode: workspace of %zu bytes is too small", workspace_bytes);
    cudaStream_t stream = (cudaStream_t)stream_;
    const BinPlan plan = make_bin_plan(n, B, H, W);
    const bool binned = n > 0 && use_binned(plan, n, B, H, W, workspace_bytes);
    // binned: pass 2 writes every output cell, so the image is not zeroed first
    if (int rc = launch_prepare(t, b, n, timestamps, B, H, W, out, workspace, !binned, stream)) return rc;
    if (n == 0) return DVSOF_OK;
    const EncodeArgs a{x, y, t, p, b, n, reinterpret_cast<const float2*>(workspace), (int)B, H, W, out, status};
    if (binned) return encode_binned_impl(a, plan, workspace, stream, 3);
    static const bool no_pdl = getenv("DVSOF_NO_PDL") != nullptr;
    // scatter is released while prepare still runs; it blocks in griddepcontrol.wait
    return encode_scatter_impl(a, stream, !no_pdl);
}

static int encode_binned_impl(const EncodeArgs& a, const BinPlan& plan, void* workspace, cudaStream_t stream,
                              int passes) {
    char* ws = static_cast<char*>(workspace);
    const bool compact = a.cx != nullptr;
    const int64_t* ends = compact ? a.cends : reinterpret_cast<const int64_t*>(ws + plan.off_ends);
    uint16_t* codes = reinterpret_cast<uint16_t*>(ws + plan.off_codes);
    float* ts = reinterpret_cast<float*>(ws + plan.off_ts);
    uint32_t* table = reinterpret_cast<uint32_t*>(ws + plan.off_table);
    static unsigned long long configured = 0;            // one bit per device (attributes are per device)
    int dev = 0;
    if (needs_device_setup(configured, &dev)) {
        cudaError_t e = cudaFuncSetAttribute(encode_bin_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBinStageBytes);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(encode_bin_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBinStageBytes);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(encode_reduce_kernel<kRedBatchLong>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBinSlabBytesMax);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(encode_reduce_kernel<kRedBatchShort>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBinSlabBytesMax);
        if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "encode_binned smem attr: %s", cudaGetErrorString(e));
        device_setup_done(configured, dev);
    }
    if (plan.max_chunks > INT32_MAX || (int64_t)a.B * plan.T > INT32_MAX)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_binned: grid too large");
    if (passes & 1) {
        if (int rc = compact
                ? launch_kernel("encode_bin_kernel<compact>", encode_bin_kernel<true>, (int)plan.max_chunks, kBinThreads,
                                kBinStageBytes, stream, passes == 3, a.cx, a.cy, a.t, a.cp, (const int64_t*)nullptr, a.n,
                                ends, a.B, a.H, a.W, plan.r_log2, plan.T, codes, ts, table, a.status)
                : launch_kernel("encode_bin_kernel", encode_bin_kernel<false>, (int)plan.max_chunks, kBinThreads,
                                kBinStageBytes, stream, passes == 3, (const void*)a.x, (const void*)a.y, a.t,
                                (const void*)a.p, a.b, a.n, ends, a.B, a.H, a.W, plan.r_log2, plan.T, codes, ts, table,
                                a.status))
            return rc;
    }
    if (!(passes & 2)) return DVSOF_OK;
    const bool chained = passes == 3;
    // runs hold ~kBinChunk / T records: two records per lane and run in flight when that is at most two warps' worth
    if (kBinChunk / plan.T <= 64)
        return launch_kernel("encode_reduce_kernel<short>", encode_reduce_kernel<kRedBatchShort>, a.B * plan.T, kRedThreads,
                             reduce_smem_bytes(plan.cells), stream, chained, codes, ts, table, ends, a.sd, a.B, a.H, a.W,
                             plan.r_log2, plan.T, a.out);
    return launch_kernel("encode_reduce_kernel", encode_reduce_kernel<kRedBatchLong>, a.B * plan.T, kRedThreads,
                         reduce_smem_bytes(plan.cells), stream, chained, codes, ts, table, ends, a.sd, a.B, a.H, a.W,
                         plan.r_log2, plan.T, a.out);
}

// 1 when dvsof_encode_event_image takes the binned path for these sizes and this workspace, else 0
extern "C" int dvsof_encode_path(int64_t n, int64_t B, int32_t H, int32_t W, size_t workspace_bytes) {
    if (n <= 0 || B <= 0 || H <= 0 || W <= 0) return 0;
    const BinPlan plan = make_bin_plan(n, B, H, W);
    return use_binned(plan, n, B, H, W, workspace_bytes) ? 1 : 0;
}

// One pass of the binned path alone (kernel-level timing, bench.py): which = 1 the counting sort
// of the event chunks into records, 2 the per-slab reduction of those records into the image.  The
// per-sample bounds (dvsof_encode_prepare) and, for pass 2, the records of an earlier pass 1 must be
// in `workspace`.  Fails when the sizes do not admit the binned path.
extern "C" int dvsof_encode_binned_pass(const int64_t* x, const int64_t* y, const float* t,
                                        const int64_t* p, const int64_t* b, int64_t n, int64_t B,
                                        int32_t H, int32_t W, float* out, void* workspace,
                                        size_t workspace_bytes, uint32_t* status, int32_t which,
                                        void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, status)) return rc;
    if (n <= 0 || !x || !y || !p || !t || !b || (which != 1 && which != 2))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_binned_pass: bad arguments");
    const BinPlan plan = make_bin_plan(n, B, H, W);
    if (!plan.ok || workspace_bytes < plan.total)
        return fail(DVSOF_ERR_WORKSPACE, "encode_binned_pass: sizes / workspace do not admit the binned path");
    const EncodeArgs a{x, y, t, p, b, n, reinterpret_cast<const float2*>(workspace), (int)B, H, W, out, status};
    return encode_binned_impl(a, plan, workspace, (cudaStream_t)stream_, which);
}

extern "C" int dvsof_encode_event_image_compact(const int16_t* x, const int16_t* y, const float* t,
                                                const uint8_t* polarity01,
                                                const int64_t* sample_begin, int64_t n,
                                                const float* timestamps, int64_t B, int32_t H,
                                                int32_t W, float* out, void* workspace,
                                                size_t workspace_bytes, uint32_t* status, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, status)) return rc;
    if (workspace_bytes < make_bin_plan(0, B, H, W).total)
        return fail(DVSOF_ERR_WORKSPACE, "encode_compact: workspace of %zu bytes is too small", workspace_bytes);
    cudaStream_t stream = (cudaStream_t)stream_;
    const BinPlan plan = make_bin_plan(n, B, H, W);
    const bool binned = n > 0 && use_binned(plan, n, B, H, W, workspace_bytes);
    if (!binned) {                                   // binned: pass 2 writes every output cell
        const size_t out_bytes = (size_t)B * 4 * H * W * sizeof(float);
        cudaError_t e = cudaMemsetAsync(out, 0, out_bytes, stream);
        if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "encode memset: %s", cudaGetErrorString(e));
    }
    if (n == 0) return DVSOF_OK;
    if (!x || !y || !t || !polarity01 || !sample_begin || !timestamps)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_compact: null event array");
    float2* sd = reinterpret_cast<float2*>(workspace);
    sample_bounds_compact_kernel<<<(int)((B + 127) / 128), 128, 0, stream>>>(
        sample_begin, t, n, timestamps, (int)B, sd, status);
    if (int rc = check_launch("sample_bounds_compact_kernel")) return rc;
    if (binned) {
        // the same two passes as the int64 form; pass 1 reads 9 B/event instead of 36 and takes the
        // sample of a chunk from the prefix (ends[s] = sample_begin[s + 1])
        EncodeArgs a{nullptr, nullptr, t, nullptr, nullptr, n, sd, (int)B, H, W, out, status};
        a.cx = x; a.cy = y; a.cp = polarity01; a.cends = sample_begin + 1;
        return encode_binned_impl(a, plan, workspace, stream, 3);
    }
    const bool vec = aligned16(x) && aligned16(y) && aligned16(t) &&
                     ((reinterpret_cast<uintptr_t>(polarity01) & 7u) == 0);
    if (vec) {
        const int grid = encode_grid((n >> 3) + 1, 256, 8);
        encode_compact_kernel<true><<<grid, 256, 0, stream>>>(x, y, t, polarity01, sample_begin, n,
                                                              sd, (int)B, H, W, out, status);
    } else {
        const int grid = encode_grid(n, 256, 8);
        encode_compact_kernel<false><<<grid, 256, 0, stream>>>(x, y, t, polarity01, sample_begin,
                                                               n, sd, (int)B, H, W, out, status);
    }
    return check_launch("encode_compact_kernel");
}

// ---------------------------------------------------------------------------------------
// dvsof_encode_event_image_static: the direct path of dvsof_encode_event_image for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Its only
// scratch, the per-sample (start, dt) pairs and exclusive ends that encode_prepare_kernel writes,
// lives in a file-scope device array instead of a caller workspace: 64 KiB held by every GPU that
// loads the library, until exit.  No counterpart in the reference; see include/dvsof.h.
//
// The array's name is a host-side stand-in in host code, not a device address: the address comes
// from cudaGetSymbolAddress, once per device, with that device current.  Every word the scatter
// reads (sd[0..B)) is rewritten by the prepare kernel of the same call, so nothing has to start at
// zero and the previous call's contents are never seen.  All calls on one device share the array:
// two calls that overlap on the GPU (different streams) race on it.
// ---------------------------------------------------------------------------------------
constexpr int64_t kStaticMaxSamples = DVSOF_STATIC_MAX_SAMPLES;
constexpr size_t kStaticScratchBytes = (size_t)kStaticMaxSamples * (sizeof(float2) + sizeof(int64_t));
static __device__ __align__(256) unsigned char g_encode_static_scratch[kStaticScratchBytes];

static int static_scratch_address(void** out) {
    constexpr int kMaxDevices = 64;
    static void* address[kMaxDevices] = {};              // per device ordinal; a race repeats the lookup
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "encode_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_static: device ordinal %d not supported (max %d)", dev, kMaxDevices - 1);
    void* p = __atomic_load_n(&address[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        e = cudaGetSymbolAddress(&p, g_encode_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(DVSOF_ERR_CUDA, "encode_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        __atomic_store_n(&address[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return DVSOF_OK;
}

extern "C" int dvsof_encode_event_image_static(const int64_t* x, const int64_t* y, const float* t,
                                               const int64_t* p, const int64_t* b, int64_t n,
                                               const float* timestamps, int64_t B, int32_t H, int32_t W,
                                               float* out, uint32_t* status, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, out, status)) return rc;
    if (B > kStaticMaxSamples)
        return fail(DVSOF_ERR_WORKSPACE, "encode_static: %lld samples exceed the static scratch (at most %lld); "
                    "use dvsof_encode_event_image with a caller workspace", (long long)B, (long long)kStaticMaxSamples);
    if (n > 0 && (!x || !y || !p || !t || !b || !timestamps))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode: null event array");
    static_assert(kStaticScratchBytes % 256 == 0, "sd | ends layout of make_bin_plan");
    if (make_bin_plan(0, B, H, W).total > kStaticScratchBytes)
        return fail(DVSOF_ERR_WORKSPACE, "encode_static: scratch layout exceeds the static array");
    void* scratch = nullptr;
    if (int rc = static_scratch_address(&scratch)) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (int rc = launch_prepare(t, b, n, timestamps, B, H, W, out, scratch, true, stream)) return rc;
    if (n == 0) return DVSOF_OK;
    const EncodeArgs a{x, y, t, p, b, n, reinterpret_cast<const float2*>(scratch), (int)B, H, W, out, status};
    static const bool no_pdl = getenv("DVSOF_NO_PDL") != nullptr;
    return encode_scatter_impl(a, stream, !no_pdl);
}

extern "C" int64_t dvsof_encode_static_max_samples(void) { return kStaticMaxSamples; }

// ---------------------------------------------------------------------------------------
// Voxel grid (extension, csrc/voxel.cuh): same stream, same per-sample bounds, C temporal bins.
// ---------------------------------------------------------------------------------------
static bool voxel_use_binned(const BinPlan& plan, int64_t n, int64_t B, int H, int W, size_t workspace_bytes) {
    if (!plan.ok || workspace_bytes < plan.total) return false;
    const int ov = encode_path_override();
    if (ov == 1) return false;
    if (ov == 2) return true;
    const int64_t pixels = B * (int64_t)H * W;
    const bool many_ctas = B * (int64_t)plan.T >= 2 * 3 * kNumSMs;
    return n >= (4 << 20) && n >= (many_ctas ? 3 : 8) * pixels;
}

extern "C" size_t dvsof_voxel_workspace_bytes(int64_t n_events, int64_t n_samples, int32_t H, int32_t W,
                                              int32_t n_bins) {
    return make_voxel_plan(n_events, n_samples, H, W, n_bins).total;
}

extern "C" int dvsof_voxel_path(int64_t n, int64_t B, int32_t H, int32_t W, int32_t C, size_t workspace_bytes) {
    if (n <= 0 || B <= 0 || H <= 0 || W <= 0) return 0;
    return voxel_use_binned(make_voxel_plan(n, B, H, W, C), n, B, H, W, workspace_bytes) ? 1 : 0;
}

extern "C" int dvsof_encode_voxel_grid(const int64_t* x, const int64_t* y, const float* t,
                                       const int64_t* p, const int64_t* b, int64_t n,
                                       const float* timestamps, int64_t B, int32_t C, int32_t H, int32_t W,
                                       float* out, void* workspace, size_t workspace_bytes,
                                       uint32_t* status, void* stream_) {
    if (int rc = check_encode_args(n, B, H, W, out, workspace, status)) return rc;
    if (C < 1 || C > kVoxelMaxBins)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "voxel: n_bins = %d outside [1, %d]", C, kVoxelMaxBins);
    if (n > 0 && (!x || !y || !p || !t || !b || !timestamps))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "voxel: null event array");
    if (workspace_bytes < make_bin_plan(0, B, H, W).total)
        return fail(DVSOF_ERR_WORKSPACE, "voxel: workspace of %zu bytes is too small", workspace_bytes);
    cudaStream_t stream = (cudaStream_t)stream_;
    const BinPlan plan = make_voxel_plan(n, B, H, W, C);
    const bool binned = n > 0 && voxel_use_binned(plan, n, B, H, W, workspace_bytes);
    // the prepare kernel zeroes "4 planes" of H * W' cells: hand it the C planes as one wide image
    // (it only uses the product); binned: pass 2 writes every cell
    {
        const size_t n_floats = binned ? 0 : (size_t)B * C * H * W;
        int64_t grid = (int64_t)((n_floats / 4 + kPrepThreads * 8 - 1) / (kPrepThreads * 8));
        if (grid > kNumSMs * 8) grid = kNumSMs * 8;
        if (grid < B) grid = B < kNumSMs * 8 ? B : kNumSMs * 8;
        if (grid < 1) grid = 1;
        const BinPlan base = make_bin_plan(0, B, H, W);
        if (int rc = launch_kernel("encode_prepare_kernel", encode_prepare_kernel, (int)grid, kPrepThreads, 0, stream,
                                   false, b, t, n, timestamps, (int)B, reinterpret_cast<float2*>(workspace),
                                   reinterpret_cast<int64_t*>(static_cast<char*>(workspace) + base.off_ends),
                                   out, n_floats))
            return rc;
    }
    if (n == 0) return DVSOF_OK;
    const float2* sd = reinterpret_cast<const float2*>(workspace);
    if (!binned) {
        const bool vec = aligned16(x) && aligned16(y) && aligned16(t) && aligned16(p) && aligned16(b);
        if (vec)
            return launch_kernel("voxel_direct_kernel<vec>", voxel_direct_kernel<true>, encode_grid((n >> 2) + 1, 256, 64),
                                 256, 0, stream, true, x, y, t, p, b, n, sd, (int)B, (int)C, H, W, out, status);
        return launch_kernel("voxel_direct_kernel", voxel_direct_kernel<false>, encode_grid(n, 256, 8), 256, 0,
                             stream, true, x, y, t, p, b, n, sd, (int)B, (int)C, H, W, out, status);
    }
    // pass 1 of the event-image path as it is (the records do not depend on C), then the voxel pass 2
    const EncodeArgs a{x, y, t, p, b, n, sd, (int)B, H, W, out, status};
    if (int rc = encode_binned_impl(a, plan, workspace, stream, 1)) return rc;
    char* ws = static_cast<char*>(workspace);
    const size_t slab = (size_t)plan.R * W * C * sizeof(float);
    static unsigned long long configured = 0;
    int dev = 0;
    if (needs_device_setup(configured, &dev)) {
        cudaError_t e = cudaFuncSetAttribute(voxel_reduce_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVoxelSlabMax);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(voxel_reduce_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVoxelSlabMax);
        if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "voxel smem attr: %s", cudaGetErrorString(e));
        device_setup_done(configured, dev);
    }
    const uint16_t* codes = reinterpret_cast<const uint16_t*>(ws + plan.off_codes);
    const float* ts = reinterpret_cast<const float*>(ws + plan.off_ts);
    const uint32_t* table = reinterpret_cast<const uint32_t*>(ws + plan.off_table);
    const int64_t* ends = reinterpret_cast<const int64_t*>(ws + plan.off_ends);
    if (kBinChunk / plan.T <= 64)
        return launch_kernel("voxel_reduce_kernel<2>", voxel_reduce_kernel<2>, (int)(B * plan.T), kVoxelRedThreads, slab,
                             stream, true, codes, ts, table, ends, sd, (int)B, (int)C, H, W, plan.r_log2, plan.T, out);
    return launch_kernel("voxel_reduce_kernel<5>", voxel_reduce_kernel<5>, (int)(B * plan.T), kVoxelRedThreads, slab,
                         stream, true, codes, ts, table, ends, sd, (int)B, (int)C, H, W, plan.r_log2, plan.T, out);
}
