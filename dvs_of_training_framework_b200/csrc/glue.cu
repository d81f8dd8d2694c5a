// glue.cu -- data-movement glue of the train step's network (EV_FlowNet/net.py:72-85: every
// decoder block starts with `interpolate(x, scale_factor=2)` (nearest) followed by
// `ReflectionPad2d(1)` before its cuDNN convolution).  As two ATen ops this writes the 4x
// upsampled tensor, reads it back and writes the padded one (and the same again in backward):
// 7.5 ms of a 34 ms step at 32 x 256x256 (profiles/r01at_train_profile_nchw.txt).  Here it is one
// gather per direction: x (N,C,H,W) -> y (N,C,2H+2,2W+2),
//     y[i, j] = x[r(i) / 2, c(j) / 2],   r(i) = reflect(i - 1) over [0, 2H),  c likewise,
// and the backward sums, for every x[h, w], the y positions that read it -- a fixed set, so no
// atomics and a fixed summation order.  The convolutions stay on cuDNN.
#include "common.cuh"

namespace dvsof {

constexpr int kGlueWarps = 8;

__device__ __forceinline__ int reflect_idx(int i, int n) {          // ReflectionPad2d(1) over [0, n)
    i = i < 0 ? -i : i;
    return i >= n ? 2 * n - 2 - i : i;
}

__global__ void __launch_bounds__(kGlueWarps * 32)
upsample2x_reflect_pad_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t planes,
                                  int H, int W) {
    const int Ho = 2 * H + 2, Wo = 2 * W + 2;
    const int lane = threadIdx.x & 31;
    const int64_t rows = planes * Ho;
    for (int64_t row = (int64_t)blockIdx.x * kGlueWarps + (threadIdx.x >> 5); row < rows;
         row += (int64_t)gridDim.x * kGlueWarps) {
        const int64_t p = row / Ho;
        const int i = (int)(row - p * Ho);
        const int a = reflect_idx(i - 1, 2 * H) >> 1;
        const float* src = x + (p * H + a) * W;
        float* dst = y + row * Wo;
        for (int j = lane; j < Wo; j += 32) dst[j] = __ldg(src + (reflect_idx(j - 1, 2 * W) >> 1));
    }
}

// gx[h, w] = sum over the upsampled pixels (a, b) in the 2x2 block of (h, w) of the padded-gradient
// positions that read (a, b): (a+1, b+1), plus row 0 for a == 1, row 2H+1 for a == 2H-2 (same for
// columns).  Rows and columns are added in ascending order.
__global__ void __launch_bounds__(kGlueWarps * 32)
upsample2x_reflect_pad_bwd_kernel(const float* __restrict__ gy, float* __restrict__ gx, int64_t planes,
                                  int H, int W) {
    const int Ho = 2 * H + 2, Wo = 2 * W + 2;
    const int lane = threadIdx.x & 31;
    const int64_t rows = planes * H;
    for (int64_t row = (int64_t)blockIdx.x * kGlueWarps + (threadIdx.x >> 5); row < rows;
         row += (int64_t)gridDim.x * kGlueWarps) {
        const int64_t p = row / H;
        const int h = (int)(row - p * H);
        int ri[4], nr = 0;                                  // padded rows feeding x row h, ascending
        if (2 * h + 1 == 1 || 2 * h == 1) ri[nr++] = 0;     // a == 1 (only h == 0 has it)
        ri[nr++] = 2 * h + 1;
        ri[nr++] = 2 * h + 2;
        if (2 * h == 2 * H - 2 || 2 * h + 1 == 2 * H - 2) ri[nr++] = 2 * H + 1;   // a == 2H-2 (h == H-1)
        const float* base = gy + p * Ho * Wo;
        for (int w = lane; w < W; w += 32) {
            const bool left = w == 0, right = w == W - 1;
            float acc = 0.f;
            for (int k = 0; k < nr; ++k) {
                const float* r = base + (int64_t)ri[k] * Wo;
                float s = left ? __ldg(r) : 0.f;
                s += __ldg(r + 2 * w + 1);
                s += __ldg(r + 2 * w + 2);
                if (right) s += __ldg(r + 2 * W + 1);
                acc += s;
            }
            gx[row * W + w] = acc;
        }
    }
}

// channels_last variants: a row of the tensor is W x C floats, a pixel is C contiguous floats.
// Work item = V consecutive channels of one pixel (V = 4, 2 or 1 by divisibility of C); a CTA walks
// output rows, its threads the flattened (pixel, channel) index of the row.  o / C through a float
// reciprocal: exact below 2^21 items per row (checked on the host).
template <int V> struct VecT;
template <> struct VecT<4> { using T = float4; };
template <> struct VecT<2> { using T = float2; };
template <> struct VecT<1> { using T = float; };
__device__ __forceinline__ float4 vadd(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float2 vadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float vadd(float a, float b) { return a + b; }

template <int V>
__global__ void __launch_bounds__(256)
upsample2x_reflect_pad_fwd_nhwc_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t N, int C,
                                       int H, int W) {
    using T = typename VecT<V>::T;
    const int Ho = 2 * H + 2, Wo = 2 * W + 2, CV = C / V;
    const float inv_cv = 1.f / (float)CV;
    const int row_items = Wo * CV;
    for (int64_t row = blockIdx.x; row < N * Ho; row += gridDim.x) {
        const int64_t n = row / Ho;
        const int i = (int)(row - n * Ho);
        const int a = reflect_idx(i - 1, 2 * H) >> 1;
        const T* src = reinterpret_cast<const T*>(x + (n * H + a) * (int64_t)W * C);
        T* dst = reinterpret_cast<T*>(y + row * (int64_t)Wo * C);
        for (int o = threadIdx.x; o < row_items; o += 256) {
            const int j = (int)(((float)o + 0.5f) * inv_cv), cv = o - j * CV;
            const int b = reflect_idx(j - 1, 2 * W) >> 1;
            dst[o] = __ldg(src + b * CV + cv);
        }
    }
}

template <int V>
__global__ void __launch_bounds__(256)
upsample2x_reflect_pad_bwd_nhwc_kernel(const float* __restrict__ gy, float* __restrict__ gx, int64_t N, int C,
                                       int H, int W) {
    using T = typename VecT<V>::T;
    const int Ho = 2 * H + 2, Wo = 2 * W + 2, CV = C / V;
    const float inv_cv = 1.f / (float)CV;
    const int row_items = W * CV;
    for (int64_t row = blockIdx.x; row < N * H; row += gridDim.x) {
        const int64_t n = row / H;
        const int h = (int)(row - n * H);
        int ri[4], nr = 0;
        if (h == 0) ri[nr++] = 0;
        ri[nr++] = 2 * h + 1;
        ri[nr++] = 2 * h + 2;
        if (h == H - 1) ri[nr++] = 2 * H + 1;
        const T* base = reinterpret_cast<const T*>(gy + n * (int64_t)Ho * Wo * C);
        T* dst = reinterpret_cast<T*>(gx + row * (int64_t)W * C);
        for (int o = threadIdx.x; o < row_items; o += 256) {
            const int w = (int)(((float)o + 0.5f) * inv_cv), cv = o - w * CV;
            T acc = T();
            for (int k = 0; k < nr; ++k) {
                const T* r = base + (int64_t)ri[k] * Wo * CV + cv;
                T s = w == 0 ? __ldg(r) : T();
                s = vadd(s, __ldg(r + (2 * w + 1) * CV));
                s = vadd(s, __ldg(r + (2 * w + 2) * CV));
                if (w == W - 1) s = vadd(s, __ldg(r + (2 * W + 1) * CV));
                acc = vadd(acc, s);
            }
            dst[o] = acc;
        }
    }
}

// The same gather with the channel concatenation of the decoder input folded in (net.py:246-262:
// `cat((h, skip), 1)` feeds the upsampling): up to kMaxParts channels_last sources, optional zero
// channels at the end; the backward writes one contiguous gradient per source, so neither the
// concatenated tensor nor the strided views of its gradient ever exist.
constexpr int kMaxParts = 4;
struct CatParts {
    float* p[kMaxParts];          // sources (forward) / gradient destinations (backward; may be null)
    int c_begin[kMaxParts + 1];   // channel offsets; c_begin[n_parts] = channels that come from sources
    int n_parts;
};
__device__ __forceinline__ int cat_part(const CatParts& cp, int c) {
    int k = 0;
#pragma unroll
    for (int i = 1; i <= kMaxParts; ++i) k += (i <= cp.n_parts && c >= cp.c_begin[i]) ? 1 : 0;
    return k;                                                // == n_parts: the zero channels
}

template <int V>
__global__ void __launch_bounds__(256)
upsample2x_reflect_pad_cat_fwd_kernel(const __grid_constant__ CatParts cp, float* __restrict__ y, int64_t N,
                                      int C, int H, int W) {
    using T = typename VecT<V>::T;
    const int Ho = 2 * H + 2, Wo = 2 * W + 2, CV = C / V;
    const float inv_cv = 1.f / (float)CV;
    const int row_items = Wo * CV;
    for (int64_t row = blockIdx.x; row < N * Ho; row += gridDim.x) {
        const int64_t n = row / Ho;
        const int i = (int)(row - n * Ho);
        const int a = reflect_idx(i - 1, 2 * H) >> 1;
        const int64_t src_row = (n * H + a) * (int64_t)W;
        T* dst = reinterpret_cast<T*>(y + row * (int64_t)Wo * C);
        for (int o = threadIdx.x; o < row_items; o += 256) {
            const int j = (int)(((float)o + 0.5f) * inv_cv), c = (o - j * CV) * V;
            const int b = reflect_idx(j - 1, 2 * W) >> 1;
            const int k = cat_part(cp, c);
            T v = T();
            if (k < cp.n_parts) {
                const int ck = cp.c_begin[k + 1] - cp.c_begin[k];
                v = __ldg(reinterpret_cast<const T*>(cp.p[k] + (src_row + b) * ck + (c - cp.c_begin[k])));
            }
            dst[o] = v;
        }
    }
}

template <int V>
__global__ void __launch_bounds__(256)
upsample2x_reflect_pad_cat_bwd_kernel(const float* __restrict__ gy, const __grid_constant__ CatParts cp,
                                      int64_t N, int C, int H, int W) {
    using T = typename VecT<V>::T;
    const int Ho = 2 * H + 2, Wo = 2 * W + 2, CV = C / V;
    const int CS = cp.c_begin[cp.n_parts] / V;               // channel vectors that have a destination
    const float inv_cs = 1.f / (float)CS;
    const int row_items = W * CS;
    for (int64_t row = blockIdx.x; row < N * H; row += gridDim.x) {
        const int64_t n = row / H;
        const int h = (int)(row - n * H);
        int ri[4], nr = 0;
        if (h == 0) ri[nr++] = 0;
        ri[nr++] = 2 * h + 1;
        ri[nr++] = 2 * h + 2;
        if (h == H - 1) ri[nr++] = 2 * H + 1;
        const T* base = reinterpret_cast<const T*>(gy + n * (int64_t)Ho * Wo * C);
        for (int o = threadIdx.x; o < row_items; o += 256) {
            const int w = (int)(((float)o + 0.5f) * inv_cs), cv = o - w * CS, c = cv * V;
            const int k = cat_part(cp, c);
            if (!cp.p[k]) continue;
            T acc = T();
            for (int q = 0; q < nr; ++q) {
                const T* r = base + (int64_t)ri[q] * Wo * CV + cv;
                T s = w == 0 ? __ldg(r) : T();
                s = vadd(s, __ldg(r + (2 * w + 1) * CV));
                s = vadd(s, __ldg(r + (2 * w + 2) * CV));
                if (w == W - 1) s = vadd(s, __ldg(r + (2 * W + 1) * CV));
                acc = vadd(acc, s);
            }
            const int ck = cp.c_begin[k + 1] - cp.c_begin[k];
            *reinterpret_cast<T*>(cp.p[k] + (row * W + w) * ck + (c - cp.c_begin[k])) = acc;
        }
    }
}

static int glue_grid(int64_t rows) {
    int64_t g = (rows + kGlueWarps - 1) / kGlueWarps;
    const int64_t cap = (int64_t)kNumSMs * 32;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}


// ---------------------------------------------------------------------------------------
// ReLU -> BatchNorm2d (training mode), the tail of every conv block of the network
// (EV_FlowNet/net.py:36,54: the reference normalises AFTER the activation).  cuDNN's NCHW
// batch-norm kernels run ONE CTA per channel (32 CTAs on 148 SMs for the widest activation) and
// the ReLU is a separate pass: 8.7 ms of a 28 ms step.  Here a channel is spread over `slices`
// CTAs (each takes whole (n, c) planes), partial sums are combined per channel in fp64, and the
// ReLU is applied on the fly in every pass: forward reads z twice and writes y once, backward
// reads z and grad_y twice and writes grad_z once.
//   y  = (relu(z) - mean) * invstd * gamma + beta,  mean/var over (N, H, W) of relu(z), var biased;
//   running_mean/var updated with momentum (unbiased var), like torch.nn.BatchNorm2d.
//   gz = [z > 0] * gamma * invstd * (gy - mean(gy) - xhat * mean(gy * xhat))
// ---------------------------------------------------------------------------------------
constexpr int kBnThreads = 256;
constexpr int kFwdSums = 2;       // sum x, sum x^2                       (x = relu(z + conv_bias))
constexpr int kBwdSums = 5;       // sum gy, sum gy*xhat; over x > 0: sum gy, count, sum xhat

struct BnShape { int64_t N; int C; int64_t HW; int slices; };

// CTA-wide sums of K per-thread values, fp64 from the warp level up; result valid in thread 0
template <int K>
__device__ __forceinline__ void bn_block_sum(const float (&v)[K], double (&out)[K]) {
    __shared__ double sm[K][kBnThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        double d = v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if (lane == 0) sm[k][warp] = d;
    }
    __syncthreads();
    if (threadIdx.x == 0)
#pragma unroll
        for (int k = 0; k < K; ++k) {
            double t = 0;
            for (int w = 0; w < kBnThreads / 32; ++w) t += sm[k][w];
            out[k] = t;
        }
}

// visit the planes (n, c), n = slice, slice + slices, ... ; f(value index in the tensor, count 1 or 4)
template <class F4, class F1>
__device__ __forceinline__ void bn_for_each(const BnShape& d, int c, int slice, F4 f4, F1 f1) {
    const bool vec = (d.HW & 3) == 0;
    for (int64_t n = slice; n < d.N; n += d.slices) {
        const int64_t base = (n * d.C + c) * d.HW;
        if (vec) for (int64_t i = (int64_t)threadIdx.x * 4; i < d.HW; i += kBnThreads * 4) f4(base + i);
        else for (int64_t i = threadIdx.x; i < d.HW; i += kBnThreads) f1(base + i);
    }
}

__device__ __forceinline__ float relu_b(float z, float cb) { return fmaxf(z + cb, 0.f); }

__global__ void __launch_bounds__(kBnThreads)
relu_bn_stats_kernel(const float* __restrict__ z, const float* __restrict__ conv_bias, BnShape d,
                     double* __restrict__ partial) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float cb = conv_bias ? conv_bias[c] : 0.f;
    float v[kFwdSums] = {0.f, 0.f};
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 q = *reinterpret_cast<const float4*>(z + i);
            const float a = relu_b(q.x, cb), b = relu_b(q.y, cb), e = relu_b(q.z, cb), f = relu_b(q.w, cb);
            v[0] += (a + b) + (e + f);
            v[1] += (a * a + b * b) + (e * e + f * f);
        },
        [&](int64_t i) { const float a = relu_b(z[i], cb); v[0] += a; v[1] += a * a; });
    double t[kFwdSums];
    bn_block_sum(v, t);
    if (threadIdx.x == 0)
        for (int k = 0; k < kFwdSums; ++k) partial[((size_t)c * d.slices + slice) * kFwdSums + k] = t[k];
}

// one WARP per channel: lanes stride over the slices, fixed-order combination.
// partial[(c * cs + k * ks) * K + j]: cs, ks = layout of the stats kernel
template <int K>
__device__ __forceinline__ void bn_channel_sum(const double* __restrict__ partial, int c, int cs, int ks,
                                               int slices, double (&out)[K]) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < K; ++j) out[j] = 0;
    for (int k = lane; k < slices; k += 32)
#pragma unroll
        for (int j = 0; j < K; ++j) out[j] += partial[((size_t)c * cs + (size_t)k * ks) * K + j];
#pragma unroll
    for (int j = 0; j < K; ++j)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) out[j] += __shfl_xor_sync(0xffffffffu, out[j], o);
}

__global__ void __launch_bounds__(128)
relu_bn_finalize_kernel(BnShape d, int cs, int ks, const double* __restrict__ partial, float momentum, float eps,
                        float* running_mean, float* running_var, float* __restrict__ save_mean,
                        float* __restrict__ save_invstd) {
    const int c = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (c >= d.C) return;
    double t[kFwdSums];
    bn_channel_sum(partial, c, cs, ks, d.slices, t);
    if ((threadIdx.x & 31) != 0) return;
    const double M = (double)d.N * (double)d.HW;
    const double mean = t[0] / M;
    double var = t[1] / M - mean * mean;
    var = var > 0 ? var : 0;
    save_mean[c] = (float)mean;
    save_invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
    if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)mean;
    if (running_var) {
        const double unbiased = M > 1 ? var * M / (M - 1) : var;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
    }
}

__global__ void __launch_bounds__(kBnThreads)
relu_bn_apply_kernel(const float* __restrict__ z, const float* __restrict__ conv_bias, BnShape d,
                     const float* __restrict__ weight, const float* __restrict__ bias,
                     const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                     float* __restrict__ y) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float cb = conv_bias ? conv_bias[c] : 0.f;
    const float mean = save_mean[c];
    const float a = save_invstd[c] * (weight ? weight[c] : 1.f), b = bias ? bias[c] : 0.f;
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(z + i));
            float4 o;
            o.x = (relu_b(v.x, cb) - mean) * a + b; o.y = (relu_b(v.y, cb) - mean) * a + b;
            o.z = (relu_b(v.z, cb) - mean) * a + b; o.w = (relu_b(v.w, cb) - mean) * a + b;
            *reinterpret_cast<float4*>(y + i) = o;
        },
        [&](int64_t i) { y[i] = (relu_b(z[i], cb) - mean) * a + b; });
}

// the five backward sums of one element
__device__ __forceinline__ void bwd_acc(float (&v)[kBwdSums], float zv, float g, float cb, float mean, float invstd) {
    const float zb = zv + cb;
    const float xhat = (fmaxf(zb, 0.f) - mean) * invstd;
    v[0] += g;
    v[1] += g * xhat;
    if (zb > 0.f) { v[2] += g; v[3] += 1.f; v[4] += xhat; }
}

__global__ void __launch_bounds__(kBnThreads)
relu_bn_bwd_stats_kernel(const float* __restrict__ z, const float* __restrict__ conv_bias,
                         const float* __restrict__ gy, BnShape d, const float* __restrict__ save_mean,
                         const float* __restrict__ save_invstd, double* __restrict__ partial) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float cb = conv_bias ? conv_bias[c] : 0.f;
    const float mean = save_mean[c], invstd = save_invstd[c];
    float v[kBwdSums] = {0.f, 0.f, 0.f, 0.f, 0.f};
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 q = *reinterpret_cast<const float4*>(z + i), g = *reinterpret_cast<const float4*>(gy + i);
            bwd_acc(v, q.x, g.x, cb, mean, invstd); bwd_acc(v, q.y, g.y, cb, mean, invstd);
            bwd_acc(v, q.z, g.z, cb, mean, invstd); bwd_acc(v, q.w, g.w, cb, mean, invstd);
        },
        [&](int64_t i) { bwd_acc(v, z[i], gy[i], cb, mean, invstd); });
    double t[kBwdSums];
    bn_block_sum(v, t);
    if (threadIdx.x == 0)
        for (int k = 0; k < kBwdSums; ++k) partial[((size_t)c * d.slices + slice) * kBwdSums + k] = t[k];
}

// grad_weight = sum gy*xhat, grad_bias = sum gy, and the gradient of the folded convolution bias:
// sum over the channel of gz = gamma*invstd * (A - mean(gy)*count - mean(gy*xhat)*X) over x > 0
__global__ void __launch_bounds__(128)
relu_bn_bwd_finalize_kernel(BnShape d, int cs, int ks, const double* __restrict__ partial,
                            const float* __restrict__ weight, const float* __restrict__ save_invstd,
                            float* __restrict__ grad_weight, float* __restrict__ grad_bias,
                            float* __restrict__ grad_conv_bias,
                            float* __restrict__ coef /* [C][2]: mean(gy), mean(gy xhat) */) {
    const int c = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (c >= d.C) return;
    double t[kBwdSums];
    bn_channel_sum(partial, c, cs, ks, d.slices, t);
    if ((threadIdx.x & 31) != 0) return;
    const double M = (double)d.N * (double)d.HW;
    const double m1 = t[0] / M, m2 = t[1] / M;
    if (grad_bias) grad_bias[c] = (float)t[0];
    if (grad_weight) grad_weight[c] = (float)t[1];
    if (grad_conv_bias) {
        const double k = (double)save_invstd[c] * (weight ? (double)weight[c] : 1.0);
        grad_conv_bias[c] = (float)(k * (t[2] - m1 * t[3] - m2 * t[4]));
    }
    coef[2 * c] = (float)m1;
    coef[2 * c + 1] = (float)m2;
}

__global__ void __launch_bounds__(kBnThreads)
relu_bn_bwd_apply_kernel(const float* __restrict__ z, const float* __restrict__ conv_bias,
                         const float* __restrict__ gy, BnShape d, const float* __restrict__ weight,
                         const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                         const float* __restrict__ coef, float* __restrict__ gz) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float cb = conv_bias ? conv_bias[c] : 0.f;
    const float mean = save_mean[c], invstd = save_invstd[c];
    const float k = invstd * (weight ? weight[c] : 1.f), m1 = coef[2 * c], m2 = coef[2 * c + 1];
    auto one = [&](float zv, float g) {
        const float zb = zv + cb;
        const float xhat = (fmaxf(zb, 0.f) - mean) * invstd;
        return zb > 0.f ? k * (g - m1 - xhat * m2) : 0.f;
    };
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(z + i)), g = __ldcs(reinterpret_cast<const float4*>(gy + i));
            float4 o;
            o.x = one(v.x, g.x); o.y = one(v.y, g.y); o.z = one(v.z, g.z); o.w = one(v.w, g.w);
            *reinterpret_cast<float4*>(gz + i) = o;
        },
        [&](int64_t i) { gz[i] = one(z[i], gy[i]); });
}

// channels_last: the tensor is (M = N*HW rows) x C.  A thread owns 4 consecutive channels (C % 4 == 0)
// and walks rows; 256 / (C/4) rows are in flight per CTA iteration (C <= 1024), a CTA takes a
// contiguous chunk of rows; the threads that share a channel quad are combined through shared
// memory.  Partials: [slices][C][K] fp64, slices = number of CTAs.
struct BnRows { int64_t M; int C; int slices; int64_t rows_per_cta; };

template <class Body>
__device__ __forceinline__ void bn_rows_for_each(const BnRows& d, Body body) {
    const int CV = d.C >> 2, rpi = 256 / CV;                 // rows per iteration
    const int r0 = threadIdx.x / CV, cv = threadIdx.x - r0 * CV;
    if (r0 >= rpi) return;
    const int64_t beg = (int64_t)blockIdx.x * d.rows_per_cta;
    const int64_t end = beg + d.rows_per_cta < d.M ? beg + d.rows_per_cta : d.M;
    for (int64_t r = beg + r0; r < end; r += rpi) body((r * d.C >> 2) + cv);     // float4 index
}

// combine the K float4 sums of the threads sharing a channel quad and store them:
// partial[(slice * C + c) * K + j]
template <int K>
__device__ __forceinline__ void bn_rows_combine_store(const BnRows& d, const float4 (&v)[K], double* partial) {
    __shared__ float4 sm[K][256];
    const int CV = d.C >> 2, rpi = 256 / CV;
#pragma unroll
    for (int j = 0; j < K; ++j) sm[j][threadIdx.x] = v[j];
    __syncthreads();
    if ((int)threadIdx.x >= CV) return;                      // r0 == 0 threads: cv = threadIdx.x
#pragma unroll
    for (int j = 0; j < K; ++j) {
        double a[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
        for (int k = 1; k < rpi; ++k) {
            const float4 u = sm[j][k * CV + threadIdx.x];
            a[0] += u.x; a[1] += u.y; a[2] += u.z; a[3] += u.w;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
            partial[((size_t)blockIdx.x * d.C + (size_t)threadIdx.x * 4 + q) * K + j] = a[q];
    }
}

__device__ __forceinline__ float4 load_cb(const float* conv_bias, int cv) {
    return conv_bias ? *reinterpret_cast<const float4*>(conv_bias + cv * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
}
__device__ __forceinline__ float4 relu4(float4 v, float4 cb) {
    return make_float4(relu_b(v.x, cb.x), relu_b(v.y, cb.y), relu_b(v.z, cb.z), relu_b(v.w, cb.w));
}

__global__ void __launch_bounds__(256)
relu_bn_stats_nhwc_kernel(const float4* __restrict__ z, const float* __restrict__ conv_bias, BnRows d,
                          double* __restrict__ partial) {
    const float4 cb = load_cb(conv_bias, threadIdx.x % (d.C >> 2));
    float4 v[kFwdSums] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
    bn_rows_for_each(d, [&](int64_t i) {
        const float4 x = relu4(z[i], cb);
        v[0].x += x.x; v[0].y += x.y; v[0].z += x.z; v[0].w += x.w;
        v[1].x += x.x * x.x; v[1].y += x.y * x.y; v[1].z += x.z * x.z; v[1].w += x.w * x.w;
    });
    bn_rows_combine_store(d, v, partial);
}

__global__ void __launch_bounds__(256)
relu_bn_apply_nhwc_kernel(const float4* __restrict__ z, const float* __restrict__ conv_bias, BnRows d,
                          const float* __restrict__ weight, const float* __restrict__ bias,
                          const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                          float4* __restrict__ y) {
    const int CV = d.C >> 2, cv = threadIdx.x % CV;
    const float4 cb = load_cb(conv_bias, cv);
    float mean[4], a[4], b[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = cv * 4 + k;
        mean[k] = save_mean[c];
        a[k] = save_invstd[c] * (weight ? weight[c] : 1.f);
        b[k] = bias ? bias[c] : 0.f;
    }
    bn_rows_for_each(d, [&](int64_t i) {
        const float4 v = relu4(__ldcs(z + i), cb);
        y[i] = make_float4((v.x - mean[0]) * a[0] + b[0], (v.y - mean[1]) * a[1] + b[1],
                           (v.z - mean[2]) * a[2] + b[2], (v.w - mean[3]) * a[3] + b[3]);
    });
}

__global__ void __launch_bounds__(256)
relu_bn_bwd_stats_nhwc_kernel(const float4* __restrict__ z, const float* __restrict__ conv_bias,
                              const float4* __restrict__ gy, BnRows d, const float* __restrict__ save_mean,
                              const float* __restrict__ save_invstd, double* __restrict__ partial) {
    const int CV = d.C >> 2, cv = threadIdx.x % CV;
    const float4 cb = load_cb(conv_bias, cv);
    float mean[4], inv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { mean[k] = save_mean[cv * 4 + k]; inv[k] = save_invstd[cv * 4 + k]; }
    float s[4][kBwdSums];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < kBwdSums; ++j) s[q][j] = 0.f;
    bn_rows_for_each(d, [&](int64_t i) {
        const float4 q = z[i], g = gy[i];
        bwd_acc(s[0], q.x, g.x, cb.x, mean[0], inv[0]); bwd_acc(s[1], q.y, g.y, cb.y, mean[1], inv[1]);
        bwd_acc(s[2], q.z, g.z, cb.z, mean[2], inv[2]); bwd_acc(s[3], q.w, g.w, cb.w, mean[3], inv[3]);
    });
    float4 v[kBwdSums];
#pragma unroll
    for (int j = 0; j < kBwdSums; ++j) v[j] = make_float4(s[0][j], s[1][j], s[2][j], s[3][j]);
    bn_rows_combine_store(d, v, partial);
}

__global__ void __launch_bounds__(256)
relu_bn_bwd_apply_nhwc_kernel(const float4* __restrict__ z, const float* __restrict__ conv_bias,
                              const float4* __restrict__ gy, BnRows d, const float* __restrict__ weight,
                              const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                              const float* __restrict__ coef, float4* __restrict__ gz) {
    const int CV = d.C >> 2, cv = threadIdx.x % CV;
    const float4 cb4 = load_cb(conv_bias, cv);
    const float cb[4] = {cb4.x, cb4.y, cb4.z, cb4.w};
    float mean[4], inv[4], k4[4], m1[4], m2[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = cv * 4 + k;
        mean[k] = save_mean[c]; inv[k] = save_invstd[c];
        k4[k] = inv[k] * (weight ? weight[c] : 1.f);
        m1[k] = coef[2 * c]; m2[k] = coef[2 * c + 1];
    }
    auto one = [&](float zv, float g, int k) {
        const float zb = zv + cb[k];
        const float xhat = (fmaxf(zb, 0.f) - mean[k]) * inv[k];
        return zb > 0.f ? k4[k] * (g - m1[k] - xhat * m2[k]) : 0.f;
    };
    bn_rows_for_each(d, [&](int64_t i) {
        const float4 v = __ldcs(z + i), g = __ldcs(gy + i);
        gz[i] = make_float4(one(v.x, g.x, 0), one(v.y, g.y, 1), one(v.z, g.z, 2), one(v.w, g.w, 3));
    });
}

static int bn_slices(int64_t N, int C) {
    int64_t s = (8 * (int64_t)kNumSMs + C - 1) / C;
    s = s < 1 ? 1 : s;
    return (int)(s > N ? N : s);
}

}  // namespace dvsof

using namespace dvsof;

static int nhwc_vec(int C, const void* a, const void* b) {
    const bool al = aligned16(a) && aligned16(b);
    return (C % 4 == 0 && al) ? 4 : ((C % 2 == 0 && al) ? 2 : 1);
}
static int row_grid(int64_t rows) {
    const int64_t cap = (int64_t)kNumSMs * 16;
    return (int)(rows < 1 ? 1 : (rows > cap ? cap : rows));
}

extern "C" int dvsof_upsample2x_reflect_pad_forward(const float* x, int64_t N, int32_t C, int32_t H, int32_t W,
                                                    float* y, int32_t channels_last, void* stream_) {
    if (N < 0 || C < 1 || H < 1 || W < 1 || (N > 0 && (!x || !y)) || H > (1 << 20) || W > (1 << 20) ||
        (channels_last && (2 * (int64_t)W + 2) * C >= (1 << 21)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_forward: bad arguments");
    if (N == 0) return DVSOF_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!channels_last) {
        const int64_t planes = N * C;
        upsample2x_reflect_pad_fwd_kernel<<<glue_grid(planes * (2 * (int64_t)H + 2)), kGlueWarps * 32, 0, stream>>>(
            x, y, planes, H, W);
    } else {
        const int grid = row_grid(N * (2 * (int64_t)H + 2));
        switch (nhwc_vec(C, x, y)) {
            case 4: upsample2x_reflect_pad_fwd_nhwc_kernel<4><<<grid, 256, 0, stream>>>(x, y, N, C, H, W); break;
            case 2: upsample2x_reflect_pad_fwd_nhwc_kernel<2><<<grid, 256, 0, stream>>>(x, y, N, C, H, W); break;
            default: upsample2x_reflect_pad_fwd_nhwc_kernel<1><<<grid, 256, 0, stream>>>(x, y, N, C, H, W);
        }
    }
    return check_launch("upsample2x_reflect_pad_fwd");
}

extern "C" int dvsof_upsample2x_reflect_pad_backward(const float* grad_y, int64_t N, int32_t C, int32_t H,
                                                     int32_t W, float* grad_x, int32_t channels_last,
                                                     void* stream_) {
    if (N < 0 || C < 1 || H < 1 || W < 1 || (N > 0 && (!grad_y || !grad_x)) || H > (1 << 20) || W > (1 << 20) ||
        (channels_last && (2 * (int64_t)W + 2) * C >= (1 << 21)))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_backward: bad arguments");
    if (N == 0) return DVSOF_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!channels_last) {
        const int64_t planes = N * C;
        upsample2x_reflect_pad_bwd_kernel<<<glue_grid(planes * (int64_t)H), kGlueWarps * 32, 0, stream>>>(
            grad_y, grad_x, planes, H, W);
    } else {
        const int grid = row_grid(N * (int64_t)H);
        switch (nhwc_vec(C, grad_y, grad_x)) {
            case 4: upsample2x_reflect_pad_bwd_nhwc_kernel<4><<<grid, 256, 0, stream>>>(grad_y, grad_x, N, C, H, W); break;
            case 2: upsample2x_reflect_pad_bwd_nhwc_kernel<2><<<grid, 256, 0, stream>>>(grad_y, grad_x, N, C, H, W); break;
            default: upsample2x_reflect_pad_bwd_nhwc_kernel<1><<<grid, 256, 0, stream>>>(grad_y, grad_x, N, C, H, W);
        }
    }
    return check_launch("upsample2x_reflect_pad_bwd");
}

static int cat_setup(float* const* parts, const int32_t* part_channels, int32_t n_parts, int32_t pad_channels,
                     bool null_ok, CatParts* cp, int* C, int* V) {
    if (n_parts < 1 || n_parts > kMaxParts || !parts || !part_channels || pad_channels < 0) return -1;
    int c = 0, v = 4;
    cp->n_parts = n_parts;
    for (int k = 0; k < kMaxParts; ++k) { cp->p[k] = nullptr; cp->c_begin[k + 1] = 0; }
    cp->c_begin[0] = 0;
    for (int k = 0; k < n_parts; ++k) {
        if (part_channels[k] < 1 || (!parts[k] && !null_ok)) return -1;
        cp->p[k] = parts[k];
        c += part_channels[k];
        cp->c_begin[k + 1] = c;
        while (v > 1 && (part_channels[k] % v || (reinterpret_cast<uintptr_t>(parts[k]) & (4 * v - 1)))) v >>= 1;
    }
    for (int k = n_parts + 1; k <= kMaxParts; ++k) cp->c_begin[k] = c;
    while (v > 1 && pad_channels % v) v >>= 1;
    *C = c + pad_channels;
    *V = v;
    return 0;
}

extern "C" int dvsof_upsample2x_reflect_pad_cat_forward(const float* const* parts, const int32_t* part_channels,
                                                        int32_t n_parts, int32_t pad_channels, int64_t N,
                                                        int32_t H, int32_t W, float* y, void* stream_) {
    CatParts cp;
    int C = 0, V = 1;
    if (cat_setup(const_cast<float* const*>(reinterpret_cast<const float* const*>(parts)), part_channels, n_parts,
                  pad_channels, false, &cp, &C, &V) || N < 0 || H < 1 || W < 1 || (N > 0 && !y) || H > (1 << 20) ||
        W > (1 << 20) || (2 * (int64_t)W + 2) * C >= (1 << 21))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_cat_forward: bad arguments");
    if (N == 0) return DVSOF_OK;
    while (V > 1 && (reinterpret_cast<uintptr_t>(y) & (4 * V - 1))) V >>= 1;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int grid = row_grid(N * (2 * (int64_t)H + 2));
    switch (V) {
        case 4: upsample2x_reflect_pad_cat_fwd_kernel<4><<<grid, 256, 0, stream>>>(cp, y, N, C, H, W); break;
        case 2: upsample2x_reflect_pad_cat_fwd_kernel<2><<<grid, 256, 0, stream>>>(cp, y, N, C, H, W); break;
        default: upsample2x_reflect_pad_cat_fwd_kernel<1><<<grid, 256, 0, stream>>>(cp, y, N, C, H, W);
    }
    return check_launch("upsample2x_reflect_pad_cat_fwd");
}

extern "C" int dvsof_upsample2x_reflect_pad_cat_backward(const float* grad_y, float* const* grad_parts,
                                                         const int32_t* part_channels, int32_t n_parts,
                                                         int32_t pad_channels, int64_t N, int32_t H, int32_t W,
                                                         void* stream_) {
    CatParts cp;
    int C = 0, V = 1;
    if (cat_setup(grad_parts, part_channels, n_parts, pad_channels, true, &cp, &C, &V) || N < 0 || H < 1 || W < 1 ||
        (N > 0 && !grad_y) || H > (1 << 20) || W > (1 << 20) || (2 * (int64_t)W + 2) * C >= (1 << 21))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_cat_backward: bad arguments");
    if (N == 0) return DVSOF_OK;
    while (V > 1 && (reinterpret_cast<uintptr_t>(grad_y) & (4 * V - 1))) V >>= 1;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int grid = row_grid(N * (int64_t)H);
    switch (V) {
        case 4: upsample2x_reflect_pad_cat_bwd_kernel<4><<<grid, 256, 0, stream>>>(grad_y, cp, N, C, H, W); break;
        case 2: upsample2x_reflect_pad_cat_bwd_kernel<2><<<grid, 256, 0, stream>>>(grad_y, cp, N, C, H, W); break;
        default: upsample2x_reflect_pad_cat_bwd_kernel<1><<<grid, 256, 0, stream>>>(grad_y, cp, N, C, H, W);
    }
    return check_launch("upsample2x_reflect_pad_cat_bwd");
}

// workspace: partial sums [C x slices][2] fp64 | coefficients [C][2] fp32
static BnRows bn_rows(int64_t N, int C, int64_t HW) {
    BnRows d;
    d.M = N * HW; d.C = C;
    const int rpi = 256 / (C / 4);
    int64_t s = (d.M + (int64_t)rpi * 8 - 1) / ((int64_t)rpi * 8);       // >= 8 iterations per CTA
    s = s < 1 ? 1 : (s > 8 * kNumSMs ? 8 * kNumSMs : s);
    d.rows_per_cta = (d.M + s - 1) / s;
    d.rows_per_cta = (d.rows_per_cta + rpi - 1) / rpi * rpi;
    d.slices = (int)((d.M + d.rows_per_cta - 1) / d.rows_per_cta);
    return d;
}
static bool bn_nhwc_ok(int C) { return C % 4 == 0 && C >= 4 && C <= 1024; }

extern "C" size_t dvsof_relu_bn_workspace_bytes(int64_t N, int32_t C, int64_t HW, int32_t channels_last) {
    if (N < 1 || C < 1 || HW < 1 || (channels_last && !bn_nhwc_ok(C))) return 0;
    const size_t slices = channels_last ? (size_t)bn_rows(N, C, HW).slices : (size_t)bn_slices(N, C);
    return round_up((size_t)C * slices * kBwdSums * sizeof(double), 256) + round_up((size_t)C * 2 * sizeof(float), 256);
}

static bool bn_args_ok(int64_t N, int32_t C, int64_t HW, const void* a, const void* b) {
    return N >= 1 && C >= 1 && C <= 65535 && HW >= 1 && a && b && (reinterpret_cast<uintptr_t>(a) & 15) == 0 &&
           (reinterpret_cast<uintptr_t>(b) & 15) == 0;
}
static float* bn_coef(void* workspace, int C, int slices) {
    return reinterpret_cast<float*>(static_cast<char*>(workspace) +
                                    round_up((size_t)C * slices * kBwdSums * sizeof(double), 256));
}

extern "C" int dvsof_relu_bn_forward(const float* z, const float* conv_bias, int64_t N, int32_t C, int64_t HW,
                                     const float* weight, const float* bias, float* running_mean,
                                     float* running_var, float momentum, float eps, float* y,
                                     float* save_mean, float* save_invstd, void* workspace,
                                     int32_t channels_last, void* stream_) {
    if (!bn_args_ok(N, C, HW, z, y) || !save_mean || !save_invstd || !workspace || (channels_last && !bn_nhwc_ok(C)) ||
        (channels_last && conv_bias && (reinterpret_cast<uintptr_t>(conv_bias) & 15) != 0))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "relu_bn_forward: bad arguments (fp32, 16-byte aligned; "
                                                "channels_last needs C % 4 == 0, C <= 1024)");
    cudaStream_t stream = (cudaStream_t)stream_;
    double* partial = static_cast<double*>(workspace);
    if (channels_last) {
        const BnRows r = bn_rows(N, C, HW);
        const BnShape d{N, C, HW, r.slices};
        const float4* z4 = reinterpret_cast<const float4*>(z);
        relu_bn_stats_nhwc_kernel<<<r.slices, 256, 0, stream>>>(z4, conv_bias, r, partial);
        relu_bn_finalize_kernel<<<(C + 3) / 4, 128, 0, stream>>>(d, 1, C, partial, momentum, eps, running_mean,
                                                               running_var, save_mean, save_invstd);
        relu_bn_apply_nhwc_kernel<<<r.slices, 256, 0, stream>>>(z4, conv_bias, r, weight, bias, save_mean, save_invstd,
                                                                reinterpret_cast<float4*>(y));
        return check_launch("relu_bn_forward(nhwc)");
    }
    const BnShape d{N, C, HW, bn_slices(N, C)};
    const dim3 grid((unsigned)d.slices, (unsigned)C);
    relu_bn_stats_kernel<<<grid, kBnThreads, 0, stream>>>(z, conv_bias, d, partial);
    relu_bn_finalize_kernel<<<(C + 3) / 4, 128, 0, stream>>>(d, d.slices, 1, partial, momentum, eps, running_mean,
                                                           running_var, save_mean, save_invstd);
    relu_bn_apply_kernel<<<grid, kBnThreads, 0, stream>>>(z, conv_bias, d, weight, bias, save_mean, save_invstd, y);
    return check_launch("relu_bn_forward");
}

extern "C" int dvsof_relu_bn_backward(const float* z, const float* conv_bias, const float* grad_y, int64_t N,
                                      int32_t C, int64_t HW, const float* weight, const float* save_mean,
                                      const float* save_invstd, float* grad_z, float* grad_weight,
                                      float* grad_bias, float* grad_conv_bias, void* workspace,
                                      int32_t channels_last, void* stream_) {
    if (!bn_args_ok(N, C, HW, z, grad_y) || !grad_z || (reinterpret_cast<uintptr_t>(grad_z) & 15) != 0 ||
        !save_mean || !save_invstd || !workspace || (channels_last && !bn_nhwc_ok(C)) ||
        (channels_last && conv_bias && (reinterpret_cast<uintptr_t>(conv_bias) & 15) != 0))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "relu_bn_backward: bad arguments (fp32, 16-byte aligned; "
                                                "channels_last needs C % 4 == 0, C <= 1024)");
    cudaStream_t stream = (cudaStream_t)stream_;
    double* partial = static_cast<double*>(workspace);
    if (channels_last) {
        const BnRows r = bn_rows(N, C, HW);
        const BnShape d{N, C, HW, r.slices};
        float* coef = bn_coef(workspace, C, r.slices);
        const float4* z4 = reinterpret_cast<const float4*>(z);
        const float4* g4 = reinterpret_cast<const float4*>(grad_y);
        relu_bn_bwd_stats_nhwc_kernel<<<r.slices, 256, 0, stream>>>(z4, conv_bias, g4, r, save_mean, save_invstd, partial);
        relu_bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, stream>>>(d, 1, C, partial, weight, save_invstd, grad_weight,
                                                                   grad_bias, grad_conv_bias, coef);
        relu_bn_bwd_apply_nhwc_kernel<<<r.slices, 256, 0, stream>>>(z4, conv_bias, g4, r, weight, save_mean, save_invstd,
                                                                    coef, reinterpret_cast<float4*>(grad_z));
        return check_launch("relu_bn_backward(nhwc)");
    }
    const BnShape d{N, C, HW, bn_slices(N, C)};
    float* coef = bn_coef(workspace, C, d.slices);
    const dim3 grid((unsigned)d.slices, (unsigned)C);
    relu_bn_bwd_stats_kernel<<<grid, kBnThreads, 0, stream>>>(z, conv_bias, grad_y, d, save_mean, save_invstd, partial);
    relu_bn_bwd_finalize_kernel<<<(C + 3) / 4, 128, 0, stream>>>(d, d.slices, 1, partial, weight, save_invstd,
                                                               grad_weight, grad_bias, grad_conv_bias, coef);
    relu_bn_bwd_apply_kernel<<<grid, kBnThreads, 0, stream>>>(z, conv_bias, grad_y, d, weight, save_mean, save_invstd,
                                                             coef, grad_z);
    return check_launch("relu_bn_backward");
}
