// eval.cu -- evaluation step of the reference on device: utils/eval.py:6-50 `flow_error_dense`
// (average endpoint error of a predicted flow against ground truth, masked by the event-count
// image and by GT validity), called once per evaluation window by utils/testing.py:84-92 right
// after get_count_image (csrc/transforms.cu).
//
// One pass over the (rows, W) pixels: 16 B (fp32 flows) or 24-32 B of flow + 8 B of count per
// pixel, HBM bound; per-CTA partials are summed in CTA order by the last CTA in fp64, so the
// result does not depend on scheduling.  Arithmetic follows numpy for the given dtypes:
//   mask  = count > 0  and  gt.x, gt.y not inf  and  sqrt(gt.x^2 + gt.y^2) > 0   (in gt's dtype)
//   EE    = sqrt(dx^2 + dy^2),  d = gt - pred  in the promoted dtype (fp32 only if both are fp32),
//           products and sum rounded separately (numpy does not contract to FMA)
//   below = EE < 3 in that dtype.
#include "common.cuh"

namespace dvsof {

constexpr int kEvalThreads = 256;

struct EvalPartial { double sum; unsigned long long n, below; };

template <typename TG, typename TP>
__global__ void __launch_bounds__(kEvalThreads)
flow_error_kernel(const TG* __restrict__ gt, const TP* __restrict__ pred,
                  const unsigned long long* __restrict__ count, int64_t n_pix, int dense,
                  EvalPartial* __restrict__ partials, unsigned* ticket, double* __restrict__ out) {
    __shared__ double s_sum[kEvalThreads / 32];
    __shared__ unsigned long long s_n[kEvalThreads / 32], s_b[kEvalThreads / 32];
    __shared__ bool is_last;
    double sum = 0.0;
    unsigned long long n = 0, below = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pix; i += stride) {
        const TG gx = gt[2 * i], gy = gt[2 * i + 1];
        const bool ev = dense || count[i] > 0;                                    // eval.py:26-29
        // eval.py:32-35: not inf, norm > 0 (norm evaluated in gt's dtype: squares may underflow)
        const TG nrm2 = sizeof(TG) == 4 ? (TG)__fadd_rn(__fmul_rn((float)gx, (float)gx), __fmul_rn((float)gy, (float)gy))
                                        : (TG)__dadd_rn(__dmul_rn((double)gx, (double)gx), __dmul_rn((double)gy, (double)gy));
        const bool ok = ev && !isinf(gx) && !isinf(gy) && nrm2 > (TG)0;
        if (!ok) continue;
        double ee;
        bool lt3;
        if (sizeof(TG) == 4 && sizeof(TP) == 4) {
            const float dx = __fsub_rn((float)gx, (float)pred[2 * i]), dy = __fsub_rn((float)gy, (float)pred[2 * i + 1]);
            const float e = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));   // eval.py:41
            ee = (double)e;
            lt3 = e < 3.f;
        } else {
            const double dx = (double)gx - (double)pred[2 * i], dy = (double)gy - (double)pred[2 * i + 1];
            const double e = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
            ee = e;
            lt3 = e < 3.0;
        }
        sum += ee;
        ++n;
        below += lt3 ? 1 : 0;                                                     // eval.py:46-47
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
        below += __shfl_xor_sync(0xffffffffu, below, o);
    }
    if (lane == 0) { s_sum[warp] = sum; s_n[warp] = n; s_b[warp] = below; }
    __syncthreads();
    if (threadIdx.x == 0) {
        EvalPartial p{0.0, 0, 0};
        for (int w = 0; w < kEvalThreads / 32; ++w) { p.sum += s_sum[w]; p.n += s_n[w]; p.below += s_b[w]; }
        partials[blockIdx.x] = p;
        __threadfence();
        is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last || threadIdx.x != 0) return;
    __threadfence();
    EvalPartial t{0.0, 0, 0};
    for (unsigned b = 0; b < gridDim.x; ++b) {                   // fixed order: deterministic
        const volatile EvalPartial* p = partials + b;
        t.sum += p->sum; t.n += p->n; t.below += p->below;
    }
    out[0] = t.sum;
    out[1] = (double)t.n;
    out[2] = (double)t.below;
    *ticket = 0u;
}

}  // namespace dvsof

using namespace dvsof;

static int eval_grid(int64_t n_pix) {
    int64_t blocks = (n_pix + kEvalThreads * 4 - 1) / (kEvalThreads * 4);
    if (blocks < 1) blocks = 1;
    if (blocks > kNumSMs * 4) blocks = kNumSMs * 4;
    return (int)blocks;
}

extern "C" size_t dvsof_flow_error_workspace_bytes(void) {
    return round_up((size_t)kNumSMs * 4 * sizeof(EvalPartial), 256) + 256;
}

extern "C" int dvsof_flow_error_dense(const void* flow_gt, int32_t gt_elem_size, const void* flow_pred,
                                      int32_t pred_elem_size, const unsigned long long* count_image,
                                      int32_t rows, int32_t W, int32_t is_dense, double* out,
                                      void* workspace, void* stream_) {
    if (rows < 0 || W < 0 || !out || !workspace || (!is_dense && !count_image) ||
        ((int64_t)rows * W > 0 && (!flow_gt || !flow_pred)) ||
        (gt_elem_size != 4 && gt_elem_size != 8) || (pred_elem_size != 4 && pred_elem_size != 8))
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "flow_error_dense: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t n_pix = (int64_t)rows * W;
    EvalPartial* partials = reinterpret_cast<EvalPartial*>(workspace);
    unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) +
                                                   round_up((size_t)kNumSMs * 4 * sizeof(EvalPartial), 256));
    cudaError_t e = cudaMemsetAsync(ticket, 0, 4, stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "flow_error memset: %s", cudaGetErrorString(e));
    const int grid = eval_grid(n_pix);
    if (gt_elem_size == 4 && pred_elem_size == 4)
        flow_error_kernel<float, float><<<grid, kEvalThreads, 0, stream>>>(
            (const float*)flow_gt, (const float*)flow_pred, count_image, n_pix, is_dense, partials, ticket, out);
    else if (gt_elem_size == 8 && pred_elem_size == 4)
        flow_error_kernel<double, float><<<grid, kEvalThreads, 0, stream>>>(
            (const double*)flow_gt, (const float*)flow_pred, count_image, n_pix, is_dense, partials, ticket, out);
    else if (gt_elem_size == 4 && pred_elem_size == 8)
        flow_error_kernel<float, double><<<grid, kEvalThreads, 0, stream>>>(
            (const float*)flow_gt, (const double*)flow_pred, count_image, n_pix, is_dense, partials, ticket, out);
    else
        flow_error_kernel<double, double><<<grid, kEvalThreads, 0, stream>>>(
            (const double*)flow_gt, (const double*)flow_pred, count_image, n_pix, is_dense, partials, ticket, out);
    return check_launch("flow_error_kernel");
}
