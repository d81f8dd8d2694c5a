// transforms.cu -- event transforms around the encoder: count image (utils/data.py:120-136),
// EventCrop (utils/data.py:24-42) and the native LUT remap `map_events`
// (utils/modules_to_build/transformations/src/transformation.cpp:35-85).
//
// All three are integer / data-movement work, bit-exact by construction.  The two with a
// data-dependent output size use the same ordered-compaction skeleton:
//   pass 1  per-1024-row chunk fan-out sums            -> chunk_sums[nb]
//   pass 2  single-CTA exclusive scan of chunk sums    -> chunk offsets, total (*n_out)
//   pass 3  recompute fan-out, CTA-local scan, write rows at chunk offset + local offset
// so the input is read twice and nothing per-row is materialised in HBM.
#include "common.cuh"

namespace dvsof {

constexpr int kCompactThreads = 256;
constexpr int kRowsPerThread = 4;
constexpr int kChunk = kCompactThreads * kRowsPerThread;  // rows per CTA

// exclusive scan of one int per thread across the CTA; returns the exclusive prefix, *total = sum
__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
    __shared__ int64_t warp_tot[kCompactThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int64_t up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    int64_t warp_prefix = 0, sum = 0;
#pragma unroll
    for (int w = 0; w < kCompactThreads / 32; ++w) {
        if (w == warp) warp_prefix = sum;
        sum += warp_tot[w];
    }
    __syncthreads();
    *total = sum;
    return warp_prefix + inc - v;
}

template <class Count>
__global__ void __launch_bounds__(kCompactThreads)
chunk_sums_kernel(Count count, int64_t n, int64_t* __restrict__ chunk_sums) {
    const int64_t base = (int64_t)blockIdx.x * kChunk + (int64_t)threadIdx.x * kRowsPerThread;
    int64_t c = 0;
#pragma unroll
    for (int k = 0; k < kRowsPerThread; ++k)
        if (base + k < n) c += count(base + k);
    int64_t total;
    block_exclusive_scan(c, &total);
    if (threadIdx.x == 0) chunk_sums[blockIdx.x] = total;
}

// in-place exclusive scan of chunk_sums[0..nb), total -> *n_out.  One CTA, sequential carry.
__global__ void __launch_bounds__(kCompactThreads)
scan_chunk_sums_kernel(int64_t* __restrict__ chunk_sums, int64_t nb, int64_t* __restrict__ n_out) {
    int64_t carry = 0;
    for (int64_t base = 0; base < nb; base += kCompactThreads) {
        const int64_t i = base + threadIdx.x;
        const int64_t v = (i < nb) ? chunk_sums[i] : 0;
        int64_t total;
        const int64_t ex = block_exclusive_scan(v, &total);
        if (i < nb) chunk_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) *n_out = carry;
}

template <class Count, class Write>
__global__ void __launch_bounds__(kCompactThreads)
compact_write_kernel(Count count, Write write, int64_t n, const int64_t* __restrict__ chunk_offsets) {
    const int64_t base = (int64_t)blockIdx.x * kChunk + (int64_t)threadIdx.x * kRowsPerThread;
    int64_t cnt[kRowsPerThread];
    int64_t c = 0;
#pragma unroll
    for (int k = 0; k < kRowsPerThread; ++k) {
        cnt[k] = (base + k < n) ? count(base + k) : 0;
        c += cnt[k];
    }
    int64_t total;
    int64_t pos = chunk_offsets[blockIdx.x] + block_exclusive_scan(c, &total);
#pragma unroll
    for (int k = 0; k < kRowsPerThread; ++k) {
        if (cnt[k]) write(base + k, pos, cnt[k]);
        pos += cnt[k];
    }
}

// ------------------------------ count image -------------------------------------------
__global__ void __launch_bounds__(256)
count_image_kernel(const int64_t* __restrict__ x, const int64_t* __restrict__ y, int64_t n, int H,
                   int W, unsigned long long* __restrict__ out, uint32_t* __restrict__ status) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    unsigned flags = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int64_t xi = __ldcs(x + i), yi = __ldcs(y + i);
        if ((uint64_t)xi >= (uint64_t)W || (uint64_t)yi >= (uint64_t)H) {
            flags |= DVSOF_ST_BAD_PIXEL;
            continue;
        }
        atomicAdd(out + (size_t)yi * W + xi, 1ull);
    }
    if (flags) atomicOr(status, flags);
}

// ------------------------------ event crop --------------------------------------------
template <typename T>
struct CropCount {
    const T* ev; int E; double y0, x0, y1, x1;
    __device__ int64_t operator()(int64_t r) const {
        const double x = (double)ev[r * E], y = (double)ev[r * E + 1];
        return (x >= x0 && x < x1 && y >= y0 && y < y1) ? 1 : 0;   // utils/data.py:35-38
    }
};
template <typename T>
struct CropWrite {
    const T* ev; T* out; int E; double y0, x0;
    __device__ void operator()(int64_t r, int64_t pos, int64_t) const {
        const T* src = ev + r * E;
        T* dst = out + pos * E;
        dst[0] = (T)((double)src[0] - x0);                          // utils/data.py:41
        dst[1] = (T)((double)src[1] - y0);
        for (int c = 2; c < E; ++c) dst[c] = src[c];
    }
};

// ------------------------------ map_events --------------------------------------------
// lut_ws layout: off[HW+1] | cnt[HW] | cursor[HW] | (int64) chunk sums | total | lut_dst[M] | tmp[M]
struct LutView {
    int32_t* off; int32_t* cnt; int32_t* cursor; int32_t* tmp; int32_t* lut_dst; int64_t* chunks; int64_t* total; char* end;
};
__host__ __device__ inline LutView lut_view(void* ws, int64_t HW, int64_t M) {
    LutView v;
    char* p = reinterpret_cast<char*>(ws);
    auto take = [&](size_t bytes) { char* q = p; p += (bytes + 255) / 256 * 256; return q; };
    v.off = reinterpret_cast<int32_t*>(take((HW + 1) * 4));
    v.cnt = reinterpret_cast<int32_t*>(take(HW * 4));
    v.cursor = reinterpret_cast<int32_t*>(take(HW * 4));
    v.chunks = reinterpret_cast<int64_t*>(take(((HW + kChunk - 1) / kChunk + 1) * 8));
    v.total = reinterpret_cast<int64_t*>(take(8));
    v.lut_dst = reinterpret_cast<int32_t*>(take((M > 0 ? M : 1) * 4));  // position independent of M
    v.tmp = reinterpret_cast<int32_t*>(take((M > 0 ? M : 1) * 4));
    v.end = p;
    return v;
}

__global__ void lut_hist_kernel(const uint64_t* __restrict__ src, const uint64_t* __restrict__ dst,
                                int64_t M, int64_t HW, int32_t* __restrict__ cnt,
                                uint32_t* __restrict__ status) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    if (src[k] >= (uint64_t)HW || dst[k] >= (uint64_t)HW) { atomicOr(status, DVSOF_ST_BAD_LUT); return; }
    atomicAdd(cnt + src[k], 1);
}
struct LutCount {
    const int32_t* cnt;
    __device__ int64_t operator()(int64_t r) const { return cnt[r]; }
};
struct LutOffWrite {   // exclusive offsets of the per-pixel lists
    int32_t* off;
    __device__ void operator()(int64_t r, int64_t pos, int64_t) const { off[r] = (int32_t)pos; }
};
__global__ void lut_fill_kernel(const uint64_t* __restrict__ src, const uint64_t* __restrict__ dst,
                                int64_t M, int64_t HW, const int32_t* __restrict__ off,
                                int32_t* __restrict__ cursor, int32_t* __restrict__ tmp) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    const uint64_t s = src[k];
    if (s >= (uint64_t)HW || dst[k] >= (uint64_t)HW) return;
    const int slot = atomicAdd(cursor + s, 1);
    tmp[off[s] + slot] = (int32_t)k;
}
// restore LUT (insertion) order inside every pixel list: rank of k among its list-mates
__global__ void lut_rank_kernel(const uint64_t* __restrict__ src, const uint64_t* __restrict__ dst,
                                int64_t M, int64_t HW, const int32_t* __restrict__ off,
                                const int32_t* __restrict__ cnt, const int32_t* __restrict__ tmp,
                                int32_t* __restrict__ lut_dst) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= M) return;
    const uint64_t s = src[k];
    if (s >= (uint64_t)HW || dst[k] >= (uint64_t)HW) return;
    const int begin = off[s], len = cnt[s];
    int rank = 0;
    for (int j = 0; j < len; ++j) rank += (tmp[begin + j] < (int32_t)k);
    lut_dst[begin + rank] = (int32_t)dst[k];
}

__device__ __forceinline__ int event_pixel_key(const float* ev, int W) {
    // transformation.cpp:62,74: ravel(int(y), int(x)) with float->int truncation
    return __float2int_rz(ev[1]) * W + __float2int_rz(ev[0]);
}
struct MapCount {
    const float* ev; int E; int W; int64_t HW; const int32_t* cnt;
    __device__ int64_t operator()(int64_t r) const {
        const int key = event_pixel_key(ev + r * E, W);
        return (key >= 0 && key < HW) ? cnt[key] : 0;
    }
};
struct MapWrite {
    const float* ev; float* out; int E; int W; const int32_t* off; const int32_t* lut_dst;
    __device__ void operator()(int64_t r, int64_t pos, int64_t c) const {
        const float* src = ev + r * E;
        const int key = event_pixel_key(src, W);
        const int32_t* d = lut_dst + off[key];
        for (int64_t j = 0; j < c; ++j) {
            float* dst = out + (pos + j) * E;
            dst[0] = (float)(d[j] % W);                             // transformation.cpp:78-80
            dst[1] = (float)(d[j] / W);
            for (int col = 2; col < E; ++col) dst[col] = src[col];  // :81 memcpy
        }
    }
};

// ------------------------------ fused augmentation ------------------------------------
// utils/dataset.py:753-769 applied to the (N, E) fp32 event rows in ONE count + scan + write:
//   flip      x <- W - x - 1                       (dataset.py:757, fp32 like numpy: two roundings)
//   rotation  (x, y) <- every LUT target of pixel (int(y), int(x)), in LUT order; unmapped pixels
//             drop the event                       (data.py:213-216 -> transformation.cpp:35-85)
//   crop      keep x in [x0, x0+w), y in [y0, y0+h), shift by (x0, y0)     (data.py:24-42)
// Equal to running the three reference steps one after the other (order preserved), without the
// two intermediate event arrays.
struct AugParams {
    const float* ev; int E; int W; int64_t HW;
    int flip; float Wf;
    const int32_t* cnt; const int32_t* off; const int32_t* lut_dst;   // cnt == nullptr: no rotation
    int crop; double y0, x0, y1, x1;
    __device__ __forceinline__ float flipped_x(const float* row) const {
        return flip ? __fsub_rn(__fsub_rn(Wf, row[0]), 1.f) : row[0];
    }
    __device__ __forceinline__ bool inside(float x, float y) const {
        return !crop || ((double)x >= x0 && (double)x < x1 && (double)y >= y0 && (double)y < y1);
    }
};
struct AugCount {
    AugParams p;
    __device__ int64_t operator()(int64_t r) const {
        const float* row = p.ev + r * p.E;
        const float x = p.flipped_x(row), y = row[1];
        if (!p.cnt) return p.inside(x, y) ? 1 : 0;
        const int key = __float2int_rz(y) * p.W + __float2int_rz(x);
        if (key < 0 || key >= p.HW) return 0;
        const int32_t* d = p.lut_dst + p.off[key];
        int64_t c = 0;
        for (int j = 0, m = p.cnt[key]; j < m; ++j) c += p.inside((float)(d[j] % p.W), (float)(d[j] / p.W)) ? 1 : 0;
        return c;
    }
};
struct AugWrite {
    AugParams p; float* out;
    __device__ __forceinline__ void emit(float* dst, const float* row, float x, float y) const {
        dst[0] = p.crop ? (float)((double)x - p.x0) : x;               // data.py:41
        dst[1] = p.crop ? (float)((double)y - p.y0) : y;
        for (int col = 2; col < p.E; ++col) dst[col] = row[col];
    }
    __device__ void operator()(int64_t r, int64_t pos, int64_t) const {
        const float* row = p.ev + r * p.E;
        const float x = p.flipped_x(row), y = row[1];
        if (!p.cnt) { emit(out + pos * p.E, row, x, y); return; }
        const int key = __float2int_rz(y) * p.W + __float2int_rz(x);
        const int32_t* d = p.lut_dst + p.off[key];
        for (int j = 0, m = p.cnt[key]; j < m; ++j) {
            const float cx = (float)(d[j] % p.W), cy = (float)(d[j] / p.W);
            if (p.inside(cx, cy)) { emit(out + pos * p.E, row, cx, cy); ++pos; }
        }
    }
};

template <class Count>
static int run_scan(Count count, int64_t n, int64_t* chunks, int64_t* n_out, cudaStream_t stream) {
    const int64_t nb = (n + kChunk - 1) / kChunk;
    if (nb > 0) chunk_sums_kernel<<<(unsigned)nb, kCompactThreads, 0, stream>>>(count, n, chunks);
    scan_chunk_sums_kernel<<<1, kCompactThreads, 0, stream>>>(chunks, nb, n_out);
    return check_launch("compaction scan");
}

}  // namespace dvsof

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
using namespace dvsof;

extern "C" int dvsof_count_image(const int64_t* x, const int64_t* y, int64_t n, int32_t H,
                                 int32_t W, unsigned long long* out, uint32_t* status,
                                 void* stream_) {
    if (n < 0 || H <= 0 || W <= 0 || !out || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "count_image: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    cudaError_t e = cudaMemsetAsync(out, 0, (size_t)H * W * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "count_image memset: %s", cudaGetErrorString(e));
    if (n == 0) return DVSOF_OK;
    int64_t blocks = (n + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    count_image_kernel<<<(unsigned)blocks, 256, 0, stream>>>(x, y, n, H, W, out, status);
    return check_launch("count_image_kernel");
}

extern "C" size_t dvsof_compact_workspace_bytes(int64_t n_rows) {
    const int64_t nb = (n_rows + kChunk - 1) / kChunk;
    return round_up((size_t)(nb + 1) * sizeof(int64_t), 256);
}

template <typename T>
static int event_crop_t(const void* events, int64_t n, int E, double y0, double x0, double h,
                        double w, void* out, int64_t* n_out, void* ws, cudaStream_t stream) {
    CropCount<T> count{(const T*)events, E, y0, x0, y0 + h, x0 + w};
    CropWrite<T> write{(const T*)events, (T*)out, E, y0, x0};
    int64_t* chunks = reinterpret_cast<int64_t*>(ws);
    if (int rc = run_scan(count, n, chunks, n_out, stream)) return rc;
    const int64_t nb = (n + kChunk - 1) / kChunk;
    if (nb > 0)
        compact_write_kernel<<<(unsigned)nb, kCompactThreads, 0, stream>>>(count, write, n, chunks);
    return check_launch("event_crop write");
}

extern "C" int dvsof_event_crop(const void* events, int64_t n, int32_t n_cols, int32_t elem_size,
                                double y0, double x0, double h, double w, void* out,
                                int64_t* n_out, void* workspace, void* stream_) {
    if (n < 0 || n_cols < 2 || !n_out || !workspace || (n > 0 && (!events || !out)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "event_crop: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    if (elem_size == 4) return event_crop_t<float>(events, n, n_cols, y0, x0, h, w, out, n_out, workspace, stream);
    if (elem_size == 8) return event_crop_t<double>(events, n, n_cols, y0, x0, h, w, out, n_out, workspace, stream);
    return fail(DVSOF_ERR_INVALID_ARGUMENT, "event_crop: elem_size must be 4 or 8, got %d", elem_size);
}

extern "C" size_t dvsof_map_lut_workspace_bytes(int32_t H, int32_t W, int64_t M) {
    const int64_t HW = (int64_t)H * W;
    LutView v = lut_view(nullptr, HW, M);
    return (size_t)(v.end - (char*)nullptr);
}

extern "C" int dvsof_map_events_build(const uint64_t* src, const uint64_t* dst, int64_t M,
                                      int32_t H, int32_t W, void* lut_ws, uint32_t* status,
                                      void* stream_) {
    if (M < 0 || H <= 0 || W <= 0 || !lut_ws || !status || (M > 0 && (!src || !dst)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "map_events_build: bad arguments");
    const int64_t HW = (int64_t)H * W;
    if (HW >= INT32_MAX || M >= INT32_MAX)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "map_events_build: image or LUT too large");
    cudaStream_t stream = (cudaStream_t)stream_;
    LutView v = lut_view(lut_ws, HW, M);
    // cnt and cursor are adjacent in the workspace: one memset clears both
    cudaError_t e = cudaMemsetAsync(v.cnt, 0, (size_t)((char*)v.chunks - (char*)v.cnt), stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "map build memset: %s", cudaGetErrorString(e));
    const unsigned mb = (unsigned)((M + 255) / 256);
    if (M > 0) lut_hist_kernel<<<mb, 256, 0, stream>>>(src, dst, M, HW, v.cnt, status);
    LutCount count{v.cnt};
    if (int rc = run_scan(count, HW, v.chunks, v.total, stream)) return rc;
    const int64_t nb = (HW + kChunk - 1) / kChunk;
    compact_write_kernel<<<(unsigned)nb, kCompactThreads, 0, stream>>>(count, LutOffWrite{v.off}, HW, v.chunks);
    if (M > 0) {
        lut_fill_kernel<<<mb, 256, 0, stream>>>(src, dst, M, HW, v.off, v.cursor, v.tmp);
        lut_rank_kernel<<<mb, 256, 0, stream>>>(src, dst, M, HW, v.off, v.cnt, v.tmp, v.lut_dst);
    }
    return check_launch("map_events_build");
}

extern "C" int dvsof_map_events_count(const float* events, int64_t n, int32_t n_cols, int32_t H,
                                      int32_t W, const void* lut_ws, void* ev_ws, int64_t* n_out,
                                      void* stream_) {
    if (n < 0 || n_cols < 2 || !lut_ws || !ev_ws || !n_out || (n > 0 && !events))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "map_events_count: bad arguments");
    const int64_t HW = (int64_t)H * W;
    LutView v = lut_view(const_cast<void*>(lut_ws), HW, 0);
    MapCount count{events, n_cols, W, HW, v.cnt};
    return run_scan(count, n, reinterpret_cast<int64_t*>(ev_ws), n_out, (cudaStream_t)stream_);
}

extern "C" int dvsof_map_events_write(const float* events, int64_t n, int32_t n_cols, int32_t H,
                                      int32_t W, const void* lut_ws, const void* ev_ws, float* out,
                                      void* stream_) {
    if (n < 0 || n_cols < 2 || !lut_ws || !ev_ws)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "map_events_write: bad arguments");
    if (n == 0) return DVSOF_OK;
    const int64_t HW = (int64_t)H * W;
    LutView v = lut_view(const_cast<void*>(lut_ws), HW, 0);
    MapCount count{events, n_cols, W, HW, v.cnt};
    MapWrite write{events, out, n_cols, W, v.off, v.lut_dst};
    const int64_t nb = (n + kChunk - 1) / kChunk;
    compact_write_kernel<<<(unsigned)nb, kCompactThreads, 0, (cudaStream_t)stream_>>>(
        count, write, n, reinterpret_cast<const int64_t*>(ev_ws));
    return check_launch("map_events_write");
}


static int aug_params(AugParams* p, const float* events, int64_t n, int32_t n_cols, int32_t H, int32_t W,
                      int32_t flip, const void* lut_ws, const double* box) {
    if (n < 0 || n_cols < 2 || H <= 0 || W <= 0 || (n > 0 && !events))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "augment_events: bad arguments");
    const int64_t HW = (int64_t)H * W;
    p->ev = events; p->E = n_cols; p->W = W; p->HW = HW;
    p->flip = flip != 0; p->Wf = (float)W;
    p->cnt = p->off = p->lut_dst = nullptr;
    if (lut_ws) {
        LutView v = lut_view(const_cast<void*>(lut_ws), HW, 0);
        p->cnt = v.cnt; p->off = v.off; p->lut_dst = v.lut_dst;
    }
    p->crop = box != nullptr;
    if (box) { p->y0 = box[0]; p->x0 = box[1]; p->y1 = box[0] + box[2]; p->x1 = box[1] + box[3]; }
    else p->y0 = p->x0 = p->y1 = p->x1 = 0;
    return DVSOF_OK;
}

extern "C" int dvsof_augment_events_count(const float* events, int64_t n, int32_t n_cols, int32_t H,
                                          int32_t W, int32_t flip, const void* lut_ws,
                                          const double* box, void* ev_ws, int64_t* n_out,
                                          void* stream_) {
    AugParams p;
    if (int rc = aug_params(&p, events, n, n_cols, H, W, flip, lut_ws, box)) return rc;
    if (!ev_ws || !n_out) return fail(DVSOF_ERR_INVALID_ARGUMENT, "augment_events_count: null workspace");
    return run_scan(AugCount{p}, n, reinterpret_cast<int64_t*>(ev_ws), n_out, (cudaStream_t)stream_);
}

extern "C" int dvsof_augment_events_write(const float* events, int64_t n, int32_t n_cols, int32_t H,
                                          int32_t W, int32_t flip, const void* lut_ws,
                                          const double* box, const void* ev_ws, float* out,
                                          void* stream_) {
    AugParams p;
    if (int rc = aug_params(&p, events, n, n_cols, H, W, flip, lut_ws, box)) return rc;
    if (!ev_ws) return fail(DVSOF_ERR_INVALID_ARGUMENT, "augment_events_write: null workspace");
    if (n == 0) return DVSOF_OK;
    const int64_t nb = (n + kChunk - 1) / kChunk;
    compact_write_kernel<<<(unsigned)nb, kCompactThreads, 0, (cudaStream_t)stream_>>>(
        AugCount{p}, AugWrite{p, out}, n, reinterpret_cast<const int64_t*>(ev_ws));
    return check_launch("augment_events_write");
}

// ---------------------------------------------------------------------------------------
// utils/data.py:139-152 frame_generator: np.searchsorted(t, frames.ravel(), side='right') --
// for every window boundary the number of events with t <= boundary (events sorted by t).
// One thread per boundary, binary search; comparisons in fp64 as numpy promotes them.
// ---------------------------------------------------------------------------------------
namespace dvsof {
template <typename T>
__global__ void window_offsets_kernel(const T* __restrict__ t, int64_t n,
                                      const double* __restrict__ bounds, int64_t n_bounds,
                                      int64_t* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_bounds) return;
    const double v = bounds[k];
    int64_t lo = 0, hi = n;                        // first i with t[i] > v
    while (lo < hi) {
        const int64_t mid = lo + ((hi - lo) >> 1);
        if ((double)__ldg(t + mid) <= v) lo = mid + 1; else hi = mid;
    }
    out[k] = lo;
}
}  // namespace dvsof

extern "C" int dvsof_window_offsets(const void* t, int32_t elem_size, int64_t n_events,
                                    const double* bounds, int64_t n_bounds, int64_t* out,
                                    void* stream_) {
    using namespace dvsof;
    if (n_events < 0 || n_bounds < 0 || (elem_size != 4 && elem_size != 8) || (n_bounds && (!bounds || !out)) ||
        (n_events && !t))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "window_offsets: bad arguments");
    if (n_bounds == 0) return DVSOF_OK;
    const int threads = 128, blocks = (int)((n_bounds + threads - 1) / threads);
    cudaStream_t stream = (cudaStream_t)stream_;
    if (elem_size == 4)
        return launch_kernel("window_offsets_kernel", window_offsets_kernel<float>, blocks, threads, 0, stream,
                             false, static_cast<const float*>(t), n_events, bounds, n_bounds, out);
    return launch_kernel("window_offsets_kernel", window_offsets_kernel<double>, blocks, threads, 0, stream,
                         false, static_cast<const double*>(t), n_events, bounds, n_bounds, out);
}

// ---------------------------------------------------------------------------------------
// Batch codec on device: the event part of utils/dataset.py:240-305
// `encode_batch` (training form -> 9 B/event storage form + run lengths) and :335-373
// `decode_batch` (storage form -> int64 training form: widening plus the run-length expansion of
// events_per_element / elements_per_sample into element_index / sample_index).
// Both are one streaming pass: decode reads 5 B and writes 40 B per event, encode reads 40 B and
// writes 5 B; the timestamps are shared between the two forms and never copied.
// ---------------------------------------------------------------------------------------
namespace dvsof {

constexpr int kCodecPer = 4;                              // consecutive events per thread
constexpr int kCodecChunk = kCompactThreads * kCodecPer;

// workspace: elem_begin[E+1] | sample_shift[B+1] (int64) | elem_sample[E] | elem_local[E] (int32)
struct CodecView {
    int64_t* elem_begin; int64_t* sample_shift; int32_t* elem_sample; int32_t* elem_local;
};
__host__ __device__ inline CodecView codec_view(void* ws, int64_t E, int64_t B) {
    CodecView v;
    char* p = reinterpret_cast<char*>(ws);
    auto take = [&](size_t bytes) { char* q = p; p += (bytes + 255) / 256 * 256; return q; };
    v.elem_begin = reinterpret_cast<int64_t*>(take((E + 1) * 8));
    v.sample_shift = reinterpret_cast<int64_t*>(take((B + 1) * 8));
    v.elem_sample = reinterpret_cast<int32_t*>(take((E > 0 ? E : 1) * 4));
    v.elem_local = reinterpret_cast<int32_t*>(take((E > 0 ? E : 1) * 4));
    return v;
}

// exclusive prefix of src[0..n) into dst[0..n], dst[n] = total; one CTA
__device__ void cta_prefix(const int64_t* __restrict__ src, int64_t n, int64_t* __restrict__ dst) {
    int64_t carry = 0;
    for (int64_t base = 0; base < n; base += kCompactThreads) {
        const int64_t i = base + threadIdx.x;
        const int64_t v = (i < n) ? src[i] : 0;
        int64_t total;
        const int64_t ex = block_exclusive_scan(v, &total);
        if (i < n) dst[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) dst[n] = carry;
}

// last index k in [0, n] with a[k] <= v (a ascending, a[0] = 0 <= v)
__device__ __forceinline__ int64_t upper_slot(const int64_t* __restrict__ a, int64_t n, int64_t v) {
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi + 1) >> 1;
        if (a[mid] <= v) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Plan of a decode: which sample an element belongs to and its index inside the sample
// (utils/dataset.py:357-368: events_per_element is sliced by the prefix of elements_per_sample;
// elements past the end of a truncated events_per_element simply do not exist).
__global__ void __launch_bounds__(kCompactThreads)
codec_plan_kernel(const int64_t* __restrict__ events_per_element, int64_t E,
                  const int64_t* __restrict__ elements_per_sample, int64_t B, int64_t n_events,
                  CodecView v, uint32_t* status) {
    pdl_launch_dependents();                       // the per-event kernel may start its prologue
    cta_prefix(elements_per_sample, B, v.sample_shift);
    if (events_per_element) cta_prefix(events_per_element, E, v.elem_begin);
    __syncthreads();
    if (events_per_element) {
        for (int64_t j = threadIdx.x; j < E; j += kCompactThreads) {
            const int64_t s = upper_slot(v.sample_shift, B, j);      // empty samples are skipped
            v.elem_sample[j] = (int32_t)(s < B ? s : B - 1);
            v.elem_local[j] = (int32_t)(j - v.sample_shift[s < B ? s : B - 1]);
            if (events_per_element[j] < 0) atomicOr(status, DVSOF_ST_BAD_COUNTS);
        }
        if (threadIdx.x == 0 && (v.elem_begin[E] != n_events || v.sample_shift[B] < E))
            atomicOr(status, DVSOF_ST_BAD_COUNTS);
    }
}

template <bool kVec>
__global__ void __launch_bounds__(kCompactThreads)
decode_events_kernel(const int16_t* __restrict__ x, const int16_t* __restrict__ y,
                     const uint8_t* __restrict__ pol, int64_t n, CodecView v, int64_t E,
                     int64_t* __restrict__ ox, int64_t* __restrict__ oy, int64_t* __restrict__ op,
                     int64_t* __restrict__ oe, int64_t* __restrict__ os) {
    const int64_t i0 = (int64_t)blockIdx.x * kCodecChunk + (int64_t)threadIdx.x * kCodecPer;
    if (i0 >= n) return;
    const int cnt = (int)(n - i0 < kCodecPer ? n - i0 : kCodecPer);
    int64_t xs[kCodecPer], ys[kCodecPer], ps[kCodecPer], es[kCodecPer], ss[kCodecPer];
    // the event arrays are inputs of the call; only the plan (read below) comes from codec_plan_kernel
    if (kVec && cnt == kCodecPer) {
        const short4 xv = *reinterpret_cast<const short4*>(x + i0), yv = *reinterpret_cast<const short4*>(y + i0);
        const uchar4 pv = *reinterpret_cast<const uchar4*>(pol + i0);
        xs[0] = xv.x; xs[1] = xv.y; xs[2] = xv.z; xs[3] = xv.w;
        ys[0] = yv.x; ys[1] = yv.y; ys[2] = yv.z; ys[3] = yv.w;
        ps[0] = pv.x; ps[1] = pv.y; ps[2] = pv.z; ps[3] = pv.w;
    } else {
#pragma unroll
        for (int k = 0; k < kCodecPer; ++k) {
            xs[k] = k < cnt ? x[i0 + k] : 0; ys[k] = k < cnt ? y[i0 + k] : 0; ps[k] = k < cnt ? pol[i0 + k] : 0;
        }
    }
    pdl_wait_for(v.elem_begin, v.elem_local, v.elem_sample);
    // element of the first event by search, of the following ones by walking (empty elements skipped)
    int64_t e = E > 0 ? upper_slot(v.elem_begin, E, i0) : 0;
    if (e >= E) e = E > 0 ? E - 1 : 0;                       // counts do not cover the events: flagged by the plan
    int64_t next = E > 0 ? v.elem_begin[e + 1] : n;
#pragma unroll
    for (int k = 0; k < kCodecPer; ++k) {
        while (i0 + k >= next && e + 1 < E) { ++e; next = v.elem_begin[e + 1]; }
        es[k] = E > 0 ? v.elem_local[e] : 0;
        ss[k] = E > 0 ? v.elem_sample[e] : 0;
        ps[k] = (ps[k] != 0 ? 1 : 0) * 2 - 1;               // dataset.py:351: bool -> long * 2 - 1
    }
    if (kVec && cnt == kCodecPer) {
        int64_t* const outs[5] = {ox, oy, op, oe, os};
        const int64_t* const vals[5] = {xs, ys, ps, es, ss};
#pragma unroll
        for (int a = 0; a < 5; ++a) {
            longlong2* d = reinterpret_cast<longlong2*>(outs[a] + i0);
            __stcs(d, make_longlong2(vals[a][0], vals[a][1]));
            __stcs(d + 1, make_longlong2(vals[a][2], vals[a][3]));
        }
    } else {
        for (int k = 0; k < cnt; ++k) {
            ox[i0 + k] = xs[k]; oy[i0 + k] = ys[k]; op[i0 + k] = ps[k]; oe[i0 + k] = es[k]; os[i0 + k] = ss[k];
        }
    }
}

// encode: narrowing casts exactly as torch's (`.to(torch.short)` wraps, dataset.py:286-291) and the
// histogram of dataset.py:293-302 (`np.add.at(events_per_element, e + element_shift[s], 1)`).
// Events arrive grouped by (sample, element), so a warp usually holds one or two distinct
// elements: one atomic per distinct element per warp.
template <bool kVec>
__global__ void __launch_bounds__(kCompactThreads)
encode_events_kernel(const int64_t* __restrict__ x, const int64_t* __restrict__ y,
                     const int64_t* __restrict__ pol, const int64_t* __restrict__ elem,
                     const int64_t* __restrict__ smp, int64_t n, CodecView v, int64_t B, int64_t E,
                     int16_t* __restrict__ ox, int16_t* __restrict__ oy, uint8_t* __restrict__ op,
                     unsigned long long* __restrict__ events_per_element, uint32_t* status) {
    const int64_t i0 = (int64_t)blockIdx.x * kCodecChunk + (int64_t)threadIdx.x * kCodecPer;
    const int cnt = i0 >= n ? 0 : (int)(n - i0 < kCodecPer ? n - i0 : kCodecPer);
    int64_t xs[kCodecPer], ys[kCodecPer], ps[kCodecPer], es[kCodecPer], ss[kCodecPer], key[kCodecPer];
    uint32_t flags = 0;
    if (kVec && cnt == kCodecPer) {                       // 16-byte streaming loads
        const int64_t* const ins[5] = {x, y, pol, elem, smp};
        int64_t* const regs[5] = {xs, ys, ps, es, ss};
#pragma unroll
        for (int a = 0; a < 5; ++a) {
            const longlong2* q = reinterpret_cast<const longlong2*>(ins[a] + i0);
            const longlong2 lo = __ldcs(q), hi = __ldcs(q + 1);
            regs[a][0] = lo.x; regs[a][1] = lo.y; regs[a][2] = hi.x; regs[a][3] = hi.y;
        }
    } else {
#pragma unroll
        for (int k = 0; k < kCodecPer; ++k) {
            const bool on = k < cnt;
            xs[k] = on ? __ldcs(x + i0 + k) : 0;
            ys[k] = on ? __ldcs(y + i0 + k) : 0;
            ps[k] = on ? __ldcs(pol + i0 + k) : 0;
            es[k] = on ? __ldcs(elem + i0 + k) : 0;
            ss[k] = on ? __ldcs(smp + i0 + k) : 0;
        }
    }
    pdl_wait_for(v.sample_shift);                                  // sample_shift comes from codec_plan_kernel
#pragma unroll
    for (int k = 0; k < kCodecPer; ++k) {
        const bool on = k < cnt;
        key[k] = -1;
        if (on) {
            const int64_t s = (int16_t)ss[k];                         // dataset.py:291 `.to(torch.short)`
            const int64_t e = es[k];
            if (s < 0 || s >= B) flags |= DVSOF_ST_BAD_SAMPLE;
            else {
                const int64_t g = e + v.sample_shift[s];               // dataset.py:297
                if (g < 0 || g >= E) flags |= DVSOF_ST_BAD_COUNTS;     // the reference: IndexError
                else key[k] = g;
            }
        }
    }
    if (cnt > 0) {
        if (kVec && cnt == kCodecPer) {
            *reinterpret_cast<short4*>(ox + i0) = make_short4((short)xs[0], (short)xs[1], (short)xs[2], (short)xs[3]);
            *reinterpret_cast<short4*>(oy + i0) = make_short4((short)ys[0], (short)ys[1], (short)ys[2], (short)ys[3]);
            // ((p + 1) / 2).to(bool): true unless p == -1 (dataset.py:289)
            *reinterpret_cast<uchar4*>(op + i0) = make_uchar4(ps[0] != -1, ps[1] != -1, ps[2] != -1, ps[3] != -1);
        } else {
            for (int k = 0; k < cnt; ++k) {
                ox[i0 + k] = (int16_t)xs[k]; oy[i0 + k] = (int16_t)ys[k]; op[i0 + k] = ps[k] != -1;
            }
        }
    }
    // histogram: one atomic per distinct element per warp and slot.  When every thread's events
    // share one element (the usual case: elements hold thousands of events) a single slot carries
    // the whole chunk.
    // Counts of the CTA's first element are gathered in shared memory and leave with one atomic.
    __shared__ int64_t cta_key;
    __shared__ unsigned cta_count;
    if (threadIdx.x == 0) { cta_key = key[0]; cta_count = 0u; }
    __syncthreads();
    const int64_t k0 = cta_key;
    const int lane = threadIdx.x & 31;
    const bool same = key[0] == key[1] && key[1] == key[2] && key[2] == key[3] && (cnt == kCodecPer || cnt == 0);
    const bool merged = __all_sync(0xffffffffu, same);
#pragma unroll
    for (int k = 0; k < kCodecPer; ++k) {
        const int64_t fk = merged ? (k == 0 ? key[0] : -1) : key[k];
        if (__all_sync(0xffffffffu, fk < 0)) continue;                   // warp-uniform
        const unsigned peers = __match_any_sync(0xffffffffu, fk);
        const unsigned total = __popc(peers) * (merged ? (unsigned)kCodecPer : 1u);
        if (fk >= 0 && lane == __ffs(peers) - 1) {
            if (fk == k0) atomicAdd(&cta_count, total);
            else atomicAdd(events_per_element + fk, (unsigned long long)total);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0 && k0 >= 0 && cta_count) atomicAdd(events_per_element + k0, (unsigned long long)cta_count);
    if (flags) atomicOr(status, flags);
}

}  // namespace dvsof

extern "C" size_t dvsof_codec_workspace_bytes(int64_t n_elements, int64_t n_samples) {
    if (n_elements < 0 || n_samples < 0) return 0;
    using dvsof::round_up;
    const size_t e = (size_t)(n_elements > 0 ? n_elements : 1);
    return round_up((size_t)(n_elements + 1) * 8, 256) + round_up((size_t)(n_samples + 1) * 8, 256) +
           2 * round_up(e * 4, 256);
}

static bool codec_vec_ok(const void* a, const void* b, const void* c) {
    return dvsof::aligned16(a) && dvsof::aligned16(b) && dvsof::aligned16(c);
}

extern "C" int dvsof_decode_events(const int16_t* x, const int16_t* y, const uint8_t* polarity01,
                                   int64_t n_events, const int64_t* events_per_element,
                                   int64_t n_elements, const int64_t* elements_per_sample,
                                   int64_t n_samples, int64_t* out_x, int64_t* out_y,
                                   int64_t* out_polarity, int64_t* out_element_index,
                                   int64_t* out_sample_index, void* workspace, uint32_t* status,
                                   void* stream_) {
    using namespace dvsof;
    if (n_events < 0 || n_elements < 0 || n_samples < 1 || !elements_per_sample || !workspace || !status ||
        (n_elements > 0 && !events_per_element) || n_elements > INT32_MAX ||
        (n_events > 0 && (!x || !y || !polarity01 || !out_x || !out_y || !out_polarity ||
                          !out_element_index || !out_sample_index)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "decode_events: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    const CodecView v = codec_view(workspace, n_elements, n_samples);
    codec_plan_kernel<<<1, kCompactThreads, 0, stream>>>(events_per_element ? events_per_element : elements_per_sample,
                                                         n_elements, elements_per_sample, n_samples, n_events, v, status);
    if (int rc = check_launch("codec_plan_kernel")) return rc;
    if (n_events == 0) return DVSOF_OK;
    const int64_t grid = (n_events + kCodecChunk - 1) / kCodecChunk;
    if (grid > INT32_MAX) return fail(DVSOF_ERR_INVALID_ARGUMENT, "decode_events: too many events");
    const bool vec = codec_vec_ok(x, y, out_x) && codec_vec_ok(out_y, out_polarity, out_element_index) &&
                     aligned16(out_sample_index) && (reinterpret_cast<uintptr_t>(polarity01) & 3u) == 0;
    return launch_kernel("decode_events_kernel", vec ? decode_events_kernel<true> : decode_events_kernel<false>,
                         (int)grid, kCompactThreads, 0, stream, true, x, y, polarity01, n_events, v, n_elements,
                         out_x, out_y, out_polarity, out_element_index, out_sample_index);
}

extern "C" int dvsof_encode_events(const int64_t* x, const int64_t* y, const int64_t* polarity,
                                   const int64_t* element_index, const int64_t* sample_index,
                                   int64_t n_events, const int64_t* elements_per_sample,
                                   int64_t n_samples, int64_t n_elements, int16_t* out_x,
                                   int16_t* out_y, uint8_t* out_polarity01, int64_t* events_per_element,
                                   void* workspace, uint32_t* status, void* stream_) {
    using namespace dvsof;
    if (n_events < 0 || n_elements < 0 || n_samples < 1 || !elements_per_sample || !workspace || !status ||
        (n_elements > 0 && !events_per_element) ||
        (n_events > 0 && (!x || !y || !polarity || !element_index || !sample_index || !out_x || !out_y ||
                          !out_polarity01)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_events: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    const CodecView v = codec_view(workspace, n_elements, n_samples);
    if (n_elements > 0) {
        cudaError_t e = cudaMemsetAsync(events_per_element, 0, sizeof(int64_t) * n_elements, stream);
        if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "encode_events memset: %s", cudaGetErrorString(e));
    }
    codec_plan_kernel<<<1, kCompactThreads, 0, stream>>>(nullptr, 0, elements_per_sample, n_samples, 0, v, status);
    if (int rc = check_launch("codec_plan_kernel")) return rc;
    if (n_events == 0) return DVSOF_OK;
    const int64_t grid = (n_events + kCodecChunk - 1) / kCodecChunk;
    if (grid > INT32_MAX) return fail(DVSOF_ERR_INVALID_ARGUMENT, "encode_events: too many events");
    const bool vec = (reinterpret_cast<uintptr_t>(out_x) & 7u) == 0 && (reinterpret_cast<uintptr_t>(out_y) & 7u) == 0 &&
                     (reinterpret_cast<uintptr_t>(out_polarity01) & 3u) == 0 && codec_vec_ok(x, y, polarity) &&
                     aligned16(element_index) && aligned16(sample_index);
    auto* epe = reinterpret_cast<unsigned long long*>(events_per_element);
    return launch_kernel("encode_events_kernel", vec ? encode_events_kernel<true> : encode_events_kernel<false>,
                         (int)grid, kCompactThreads, 0, stream, true, x, y, polarity, element_index, sample_index,
                         n_events, v, n_samples, n_elements, out_x, out_y, out_polarity01, epe, status);
}
