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

struct BnShape { int64_t N; int C; int64_t HW; int slices; };

__device__ __forceinline__ double2 bn_block_sum(float a, float b) {
    __shared__ double sa[kBnThreads / 32], sb[kBnThreads / 32];
    double da = a, db = b;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        da += __shfl_xor_sync(0xffffffffu, da, o);
        db += __shfl_xor_sync(0xffffffffu, db, o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sa[warp] = da; sb[warp] = db; }
    __syncthreads();
    double ta = 0, tb = 0;
    if (threadIdx.x == 0)
        for (int w = 0; w < kBnThreads / 32; ++w) { ta += sa[w]; tb += sb[w]; }
    return make_double2(ta, tb);                           // valid in thread 0
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

__global__ void __launch_bounds__(kBnThreads)
relu_bn_stats_kernel(const float* __restrict__ z, BnShape d, double* __restrict__ partial) {
    const int c = blockIdx.y, slice = blockIdx.x;
    float s = 0.f, q = 0.f;
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 v = *reinterpret_cast<const float4*>(z + i);
            const float a = fmaxf(v.x, 0.f), b = fmaxf(v.y, 0.f), e = fmaxf(v.z, 0.f), f = fmaxf(v.w, 0.f);
            s += (a + b) + (e + f);
            q += (a * a + b * b) + (e * e + f * f);
        },
        [&](int64_t i) { const float a = fmaxf(z[i], 0.f); s += a; q += a * a; });
    const double2 t = bn_block_sum(s, q);
    if (threadIdx.x == 0) {
        partial[((size_t)c * d.slices + slice) * 2] = t.x;
        partial[((size_t)c * d.slices + slice) * 2 + 1] = t.y;
    }
}

__global__ void __launch_bounds__(128)
relu_bn_finalize_kernel(BnShape d, const double* __restrict__ partial, float momentum, float eps,
                        float* running_mean, float* running_var, float* __restrict__ save_mean,
                        float* __restrict__ save_invstd) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d.C) return;
    double s = 0, q = 0;
    for (int k = 0; k < d.slices; ++k) {
        s += partial[((size_t)c * d.slices + k) * 2];
        q += partial[((size_t)c * d.slices + k) * 2 + 1];
    }
    const double M = (double)d.N * (double)d.HW;
    const double mean = s / M;
    double var = q / M - mean * mean;
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
relu_bn_apply_kernel(const float* __restrict__ z, BnShape d, const float* __restrict__ weight,
                     const float* __restrict__ bias, const float* __restrict__ save_mean,
                     const float* __restrict__ save_invstd, float* __restrict__ y) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float mean = save_mean[c];
    const float a = save_invstd[c] * (weight ? weight[c] : 1.f), b = bias ? bias[c] : 0.f;
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(z + i));
            float4 o;
            o.x = (fmaxf(v.x, 0.f) - mean) * a + b; o.y = (fmaxf(v.y, 0.f) - mean) * a + b;
            o.z = (fmaxf(v.z, 0.f) - mean) * a + b; o.w = (fmaxf(v.w, 0.f) - mean) * a + b;
            *reinterpret_cast<float4*>(y + i) = o;
        },
        [&](int64_t i) { y[i] = (fmaxf(z[i], 0.f) - mean) * a + b; });
}

__global__ void __launch_bounds__(kBnThreads)
relu_bn_bwd_stats_kernel(const float* __restrict__ z, const float* __restrict__ gy, BnShape d,
                         const float* __restrict__ save_mean, const float* __restrict__ save_invstd,
                         double* __restrict__ partial) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float mean = save_mean[c], invstd = save_invstd[c];
    float s1 = 0.f, s2 = 0.f;
    bn_for_each(d, c, slice,
        [&](int64_t i) {
            const float4 v = *reinterpret_cast<const float4*>(z + i), g = *reinterpret_cast<const float4*>(gy + i);
            s1 += (g.x + g.y) + (g.z + g.w);
            s2 += (g.x * ((fmaxf(v.x, 0.f) - mean) * invstd) + g.y * ((fmaxf(v.y, 0.f) - mean) * invstd)) +
                  (g.z * ((fmaxf(v.z, 0.f) - mean) * invstd) + g.w * ((fmaxf(v.w, 0.f) - mean) * invstd));
        },
        [&](int64_t i) { s1 += gy[i]; s2 += gy[i] * ((fmaxf(z[i], 0.f) - mean) * invstd); });
    const double2 t = bn_block_sum(s1, s2);
    if (threadIdx.x == 0) {
        partial[((size_t)c * d.slices + slice) * 2] = t.x;
        partial[((size_t)c * d.slices + slice) * 2 + 1] = t.y;
    }
}

__global__ void __launch_bounds__(128)
relu_bn_bwd_finalize_kernel(BnShape d, const double* __restrict__ partial, float* __restrict__ grad_weight,
                            float* __restrict__ grad_bias, float* __restrict__ coef /* [C][2]: mean(gy), mean(gy xhat) */) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d.C) return;
    double s1 = 0, s2 = 0;
    for (int k = 0; k < d.slices; ++k) {
        s1 += partial[((size_t)c * d.slices + k) * 2];
        s2 += partial[((size_t)c * d.slices + k) * 2 + 1];
    }
    const double M = (double)d.N * (double)d.HW;
    if (grad_bias) grad_bias[c] = (float)s1;
    if (grad_weight) grad_weight[c] = (float)s2;
    coef[2 * c] = (float)(s1 / M);
    coef[2 * c + 1] = (float)(s2 / M);
}

__global__ void __launch_bounds__(kBnThreads)
relu_bn_bwd_apply_kernel(const float* __restrict__ z, const float* __restrict__ gy, BnShape d,
                         const float* __restrict__ weight, const float* __restrict__ save_mean,
                         const float* __restrict__ save_invstd, const float* __restrict__ coef,
                         float* __restrict__ gz) {
    const int c = blockIdx.y, slice = blockIdx.x;
    const float mean = save_mean[c], invstd = save_invstd[c];
    const float k = invstd * (weight ? weight[c] : 1.f), m1 = coef[2 * c], m2 = coef[2 * c + 1];
    auto one = [&](float zv, float g) {
        const float xhat = (fmaxf(zv, 0.f) - mean) * invstd;
        return zv > 0.f ? k * (g - m1 - xhat * m2) : 0.f;
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

static int bn_slices(int64_t N, int C) {
    int64_t s = (8 * (int64_t)kNumSMs + C - 1) / C;
    s = s < 1 ? 1 : s;
    return (int)(s > N ? N : s);
}

}  // namespace dvsof

using namespace dvsof;

extern "C" int dvsof_upsample2x_reflect_pad_forward(const float* x, int64_t planes, int32_t H, int32_t W,
                                                    float* y, void* stream) {
    if (planes < 0 || H < 1 || W < 1 || (planes > 0 && (!x || !y)) || H > (1 << 28) || W > (1 << 28))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_forward: bad arguments");
    if (planes == 0) return DVSOF_OK;
    upsample2x_reflect_pad_fwd_kernel<<<glue_grid(planes * (2 * (int64_t)H + 2)), kGlueWarps * 32, 0,
                                        (cudaStream_t)stream>>>(x, y, planes, H, W);
    return check_launch("upsample2x_reflect_pad_fwd_kernel");
}

extern "C" int dvsof_upsample2x_reflect_pad_backward(const float* grad_y, int64_t planes, int32_t H,
                                                     int32_t W, float* grad_x, void* stream) {
    if (planes < 0 || H < 1 || W < 1 || (planes > 0 && (!grad_y || !grad_x)) || H > (1 << 28) || W > (1 << 28))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "upsample2x_reflect_pad_backward: bad arguments");
    if (planes == 0) return DVSOF_OK;
    upsample2x_reflect_pad_bwd_kernel<<<glue_grid(planes * (int64_t)H), kGlueWarps * 32, 0,
                                        (cudaStream_t)stream>>>(grad_y, grad_x, planes, H, W);
    return check_launch("upsample2x_reflect_pad_bwd_kernel");
}

// workspace: partial sums [C][slices][2] fp64 | coefficients [C][2] fp32
extern "C" size_t dvsof_relu_bn_workspace_bytes(int64_t N, int32_t C) {
    if (N < 1 || C < 1) return 0;
    return round_up((size_t)C * bn_slices(N, C) * 2 * sizeof(double), 256) + round_up((size_t)C * 2 * sizeof(float), 256);
}

static bool bn_args_ok(int64_t N, int32_t C, int64_t HW, const void* a, const void* b) {
    return N >= 1 && C >= 1 && C <= 65535 && HW >= 1 && a && b && (reinterpret_cast<uintptr_t>(a) & 15) == 0 &&
           (reinterpret_cast<uintptr_t>(b) & 15) == 0;
}

extern "C" int dvsof_relu_bn_forward(const float* z, int64_t N, int32_t C, int64_t HW, const float* weight,
                                     const float* bias, float* running_mean, float* running_var,
                                     float momentum, float eps, float* y, float* save_mean,
                                     float* save_invstd, void* workspace, void* stream_) {
    if (!bn_args_ok(N, C, HW, z, y) || !save_mean || !save_invstd || !workspace)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "relu_bn_forward: bad arguments (fp32 NCHW, 16-byte aligned)");
    cudaStream_t stream = (cudaStream_t)stream_;
    const BnShape d{N, C, HW, bn_slices(N, C)};
    double* partial = static_cast<double*>(workspace);
    const dim3 grid((unsigned)d.slices, (unsigned)C);
    relu_bn_stats_kernel<<<grid, kBnThreads, 0, stream>>>(z, d, partial);
    relu_bn_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(d, partial, momentum, eps, running_mean, running_var,
                                                               save_mean, save_invstd);
    relu_bn_apply_kernel<<<grid, kBnThreads, 0, stream>>>(z, d, weight, bias, save_mean, save_invstd, y);
    return check_launch("relu_bn_forward");
}

extern "C" int dvsof_relu_bn_backward(const float* z, const float* grad_y, int64_t N, int32_t C, int64_t HW,
                                      const float* weight, const float* save_mean, const float* save_invstd,
                                      float* grad_z, float* grad_weight, float* grad_bias, void* workspace,
                                      void* stream_) {
    if (!bn_args_ok(N, C, HW, z, grad_y) || !grad_z || (reinterpret_cast<uintptr_t>(grad_z) & 15) != 0 ||
        !save_mean || !save_invstd || !workspace)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "relu_bn_backward: bad arguments (fp32 NCHW, 16-byte aligned)");
    cudaStream_t stream = (cudaStream_t)stream_;
    const BnShape d{N, C, HW, bn_slices(N, C)};
    double* partial = static_cast<double*>(workspace);
    float* coef = reinterpret_cast<float*>(static_cast<char*>(workspace) +
                                           round_up((size_t)C * d.slices * 2 * sizeof(double), 256));
    const dim3 grid((unsigned)d.slices, (unsigned)C);
    relu_bn_bwd_stats_kernel<<<grid, kBnThreads, 0, stream>>>(z, grad_y, d, save_mean, save_invstd, partial);
    relu_bn_bwd_finalize_kernel<<<(C + 127) / 128, 128, 0, stream>>>(d, partial, grad_weight, grad_bias, coef);
    relu_bn_bwd_apply_kernel<<<grid, kBnThreads, 0, stream>>>(z, grad_y, d, weight, save_mean, save_invstd, coef, grad_z);
    return check_launch("relu_bn_backward");
}
