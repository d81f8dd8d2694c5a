// loss.cu -- multiscale self-supervised warp loss (utils/loss.py of the reference), forward and
// analytic backward.  One launch covers ALL scales and predictions: a CTA owns one 32x32 pixel
// tile of one (scale, prediction), stages the flow tile (+1 halo) in shared memory, gathers the
// bilinear taps of the `next` pyramid level, and reduces the Charbonnier terms to one partial
// record per CTA.  A single-CTA-per-scale finalize kernel sums the records in a fixed order in
// fp64 => deterministic results, and the per-sample out-of-border counts n_i the backward needs.
//
// HBM traffic per pixel: forward 8 B flow + 4 B prev + 4 B next (taps hit L1/L2); backward the
// same + 8 B grad_flow written once.  Arithmetic follows the reference's CPU path: the sampling
// grid is g = (pix + flow) / ((S-1)/2) - 1 with IEEE division (utils/loss.py:150-156), then
// ix = (g + 1) * ((S-1)/2) as ATen's CPU grid_sampler does for align_corners=True.
#include "common.cuh"
#include "exact_div.h"
#include <stdlib.h>

namespace dvsof {

constexpr float kAlpha = 0.45f;    // utils/loss.py:25
constexpr float kEps2 = 1e-3f * 1e-3f;  // utils/loss.py:26,31
constexpr int kTile = 32;          // tile edge (pixels)
constexpr int kLossThreads = 256;  // 32 x 8 threads, 4 rows per thread
constexpr int kRec = 8;            // floats per CTA partial record

// Charbonnier c(d) = (d^2 + eps^2)^alpha and c'(d) = 2 alpha d (d^2+eps^2)^(alpha-1).
// lg2/ex2 on the SFU: relative error ~3e-7, inside the 1e-5 budget of the path.
__device__ __forceinline__ float ex2_approx(float x) {   // MUFU.EX2 (arguments here are > -20: no denormals)
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lg2_approx(float x) {   // MUFU.LG2 (arguments here are >= 1e-6: normal)
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float charb(float d) {
    const float s = fmaf(d, d, kEps2);
    return ex2_approx(kAlpha * lg2_approx(s));
}
__device__ __forceinline__ float charb_grad(float d) {
    const float s = fmaf(d, d, kEps2);
    return (2.f * kAlpha) * d * ex2_approx((kAlpha - 1.f) * lg2_approx(s));
}

struct ScaleTable {
    int n_scales, n_pred;
    int H[DVSOF_MAX_SCALES], W[DVSOF_MAX_SCALES];
    int tiles_x[DVSOF_MAX_SCALES], tiles_y[DVSOF_MAX_SCALES];
    int cta_begin[DVSOF_MAX_SCALES + 1];
    float rx[DVSOF_MAX_SCALES], ry[DVSOF_MAX_SCALES];   // RN(1/((W-1)/2)), RN(1/((H-1)/2)); 0 = plain division
    float coef[3 * DVSOF_MAX_SCALES];                   // fused pass: d(loss)/d(term[k][s]) at [k * MAX + s]
    const float* flow[DVSOF_MAX_SCALES];
    const float* images[DVSOF_MAX_SCALES];
    float* grad_flow[DVSOF_MAX_SCALES];
};

struct TileCoord { int s, n, y0, x0; };
__device__ __forceinline__ TileCoord decode_cta(const ScaleTable& tab) {
    int s = 0;
    while (s + 1 < tab.n_scales && (int)blockIdx.x >= tab.cta_begin[s + 1]) ++s;
    const int local = blockIdx.x - tab.cta_begin[s];
    const int per = tab.tiles_x[s] * tab.tiles_y[s];
    const int n = local / per, tile = local - n * per;
    const int ty = tile / tab.tiles_x[s];
    return TileCoord{s, n, ty * kTile, (tile - ty * tab.tiles_x[s]) * kTile};
}

// bilinear sample of `img` (H x W) at the position the reference's grid_sample visits for
// pixel (x,y) displaced by (u,v); also returns the normalised grid (for the border mask) and the
// spatial derivatives of the sample.
struct Warp { float val, ddx, ddy, gx, gy; };
struct GridNorm { float hx, hy, rx, ry; };   // half extents (S-1)/2 and their correctly rounded reciprocals

__device__ __forceinline__ Warp warp_sample(const float* __restrict__ img, int H, int W,
                                            const GridNorm& gn, int x, int y, float u, float v,
                                            bool valid = true) {
    Warp r;
    r.gx = __fsub_rn(dvsof_div_by_const(__fadd_rn((float)x, u), gn.hx, gn.rx), 1.f);   // utils/loss.py:150-156
    r.gy = __fsub_rn(dvsof_div_by_const(__fadd_rn((float)y, v), gn.hy, gn.ry), 1.f);
    const float ix = __fmul_rn(__fadd_rn(r.gx, 1.f), gn.hx);            // ATen CPU unnormalize
    const float iy = __fmul_rn(__fadd_rn(r.gy, 1.f), gn.hy);
    const float xw = floorf(ix), yn = floorf(iy);
    const float wx = ix - xw, wy = iy - yn;                          // weight of east / south
    const float ex = 1.f - wx, ey = 1.f - wy;
    // clamp in float (huge / non-finite coordinates never reach the int conversion), then the
    // four zero-padded taps with unsigned range tests; 32-bit offsets inside one image plane
    const int xi = (int)fminf(fmaxf(xw, -2.f), (float)W);
    const int yi = (int)fminf(fmaxf(yn, -2.f), (float)H);
    const bool x0ok = (unsigned)xi < (unsigned)W, x1ok = (unsigned)(xi + 1) < (unsigned)W;
    const bool y0ok = valid && (unsigned)yi < (unsigned)H, y1ok = valid && (unsigned)(yi + 1) < (unsigned)H;
    const int o = yi * W + xi;
    const float nw = (x0ok && y0ok) ? __ldg(img + o) : 0.f;
    const float ne = (x1ok && y0ok) ? __ldg(img + o + 1) : 0.f;
    const float sw = (x0ok && y1ok) ? __ldg(img + o + W) : 0.f;
    const float se = (x1ok && y1ok) ? __ldg(img + o + W + 1) : 0.f;
    // same operations as ATen's vectorised CPU grid_sampler (weights s*e, s*w, n*e, n*w; the sum
    // nw*w0 + ne*w1 + sw*w2 + se*w3 is compiled there to mul + 3 chained FMAs -- pinned by
    // tests/test_oracle_golden.py::test_aten_bilinear_operation_order).  Where warped ~= prev the
    // Charbonnier derivative amplifies a 1-ulp difference of `val` ~1e3 times, so the
    // interpolation is kept bit-compatible with the reference's CPU path.
    r.val = __fmaf_rn(se, __fmul_rn(wy, wx),
                      __fmaf_rn(sw, __fmul_rn(wy, ex),
                                __fmaf_rn(ne, __fmul_rn(ey, wx), __fmul_rn(nw, __fmul_rn(ey, ex)))));
    r.ddx = (ne - nw) * ey + (se - sw) * wy;
    r.ddy = (sw - nw) * ex + (se - ne) * wx;
    return r;
}

// sum K per-thread values over the warp with the halving butterfly: 4+2+1+1+1 = 9 shuffles for
// K = 7 instead of 35.  Afterwards lane l (l & 3 == 0) holds the total of value index
// ((l >> 4) & 1) * 4 + ((l >> 3) & 1) * 2 + ((l >> 2) & 1) in v[0].
__device__ __forceinline__ void warp_sum8(float (&v)[8], int lane) {
    const bool up16 = lane & 16, up8 = lane & 8, up4 = lane & 4;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float send = up16 ? v[i] : v[i + 4];
        const float keep = up16 ? v[i + 4] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const float send = up8 ? v[i] : v[i + 2];
        const float keep = up8 ? v[i + 2] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
    }
    {
        const float send = up4 ? v[0] : v[1];
        const float keep = up4 ? v[1] : v[0];
        v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
    }
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 2);
    v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
}

// ------------------------------ timestamp matching -------------------------------------
__global__ void match_timestamps_kernel(const float* __restrict__ flow_ts,
                                        const int64_t* __restrict__ flow_sample_idx, int P,
                                        const float* __restrict__ timestamps,
                                        const int64_t* __restrict__ sample_idx, int D,
                                        int32_t* __restrict__ match_idx,
                                        uint32_t* __restrict__ status, unsigned* ticket) {
    pdl_launch_dependents();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;   // i = k * P + p
    if (i == 0 && ticket) *ticket = 0u;                    // arms the forward kernel's finalize
    if (i >= 2 * P) return;
    const int k = i / P, p = i - k * P;
    const float want = flow_ts[2 * p + k];
    const int64_t smp = flow_sample_idx[p];
    int hits = 0, idx = 0;
    for (int d = 0; d < D; ++d)
        if (timestamps[d] == want && sample_idx[d] == smp) { ++hits; idx = d; }   // loss.py:187-191
    if (hits != 1) atomicOr(status, DVSOF_ST_NO_MATCH);                           // loss.py:193
    match_idx[i] = idx;
}

// ------------------------------ bilinear resize (align_corners) ------------------------
__global__ void __launch_bounds__(256)
resize_bilinear_kernel(const float* __restrict__ src, int Hin, int Win, float* __restrict__ dst,
                       int Hout, int Wout, int64_t total) {
    pdl_launch_dependents();
    pdl_wait();                                  // src may be the previous level of the chain
    const float sy = Hout > 1 ? (float)(Hin - 1) / (float)(Hout - 1) : 0.f;
    const float sx = Wout > 1 ? (float)(Win - 1) / (float)(Wout - 1) : 0.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int x = (int)(i % Wout);
        const int64_t r = i / Wout;
        const int y = (int)(r % Hout);
        const int64_t d = r / Hout;
        const float fy = sy * (float)y, fx = sx * (float)x;
        int y0 = (int)fy, x0 = (int)fx;
        y0 = y0 < Hin - 1 ? y0 : Hin - 1;
        x0 = x0 < Win - 1 ? x0 : Win - 1;
        const int y1 = y0 + (y0 < Hin - 1), x1 = x0 + (x0 < Win - 1);
        const float ly = fminf(fmaxf(fy - (float)y0, 0.f), 1.f), lx = fminf(fmaxf(fx - (float)x0, 0.f), 1.f);
        const float hy = 1.f - ly, hx = 1.f - lx;
        const float* im = src + (size_t)d * Hin * Win;
        const float a = __ldg(im + (size_t)y0 * Win + x0), b = __ldg(im + (size_t)y0 * Win + x1);
        const float c = __ldg(im + (size_t)y1 * Win + x0), e = __ldg(im + (size_t)y1 * Win + x1);
        // operation order of ATen's CPU upsample_bilinear2d as compiled (two FMAs per row pair, one
        // for the rows): bit-identical to the reference's F.interpolate on the upsampling levels of
        // the chain (pinned on CPU by tests/test_oracle_golden.py::test_aten_upsample_operation_order)
        const float top = __fmaf_rn(hx, a, __fmul_rn(lx, b)), bot = __fmaf_rn(hx, c, __fmul_rn(lx, e));
        dst[i] = __fmaf_rn(hy, top, __fmul_rn(ly, bot));
    }
}

// Stage the 2-channel flow tile (EDGE x EDGE pixels starting HALO before the tile origin, zero
// outside the image) into shared memory [2][EDGE][PITCH].  All global loads of a thread are
// issued before its first shared store (fixed trip count, fully unrolled): one DRAM round trip
// per CTA instead of one per row -- ncu showed the previous row-per-warp loop (load -> store ->
// next load) holding 60 % of all stall samples.
template <int EDGE, int PITCH, int HALO>
__device__ __forceinline__ void stage_flow_tile(float* __restrict__ Fs, const float* __restrict__ flow,
                                                size_t plane, int H, int W, int y0, int x0) {
    constexpr int kElems = 2 * EDGE * EDGE;
    constexpr int kIters = (kElems + kLossThreads - 1) / kLossThreads;
    float v[kIters];
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
        const int i = threadIdx.x + k * kLossThreads;
        const int ch = i / (EDGE * EDGE), rem = i - ch * (EDGE * EDGE);
        const int ly = rem / EDGE, lx = rem - ly * EDGE;
        const int y = y0 + ly - HALO, x = x0 + lx - HALO;
        const bool ok = i < kElems && y >= 0 && y < H && x >= 0 && x < W;
        v[k] = ok ? __ldg(flow + ch * plane + y * W + x) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
        const int i = threadIdx.x + k * kLossThreads;
        const int ch = i / (EDGE * EDGE), rem = i - ch * (EDGE * EDGE);
        const int ly = rem / EDGE, lx = rem - ly * EDGE;
        if (i < kElems) Fs[(ch * EDGE + ly) * PITCH + lx] = v[k];
    }
}

// ------------------------------ forward ------------------------------------------------
// record: [photo, smooth_h, smooth_v, smooth_diag_a, smooth_diag_b, outborder_sum, n_bad, -]
constexpr int kMaxPairs = 256;     // (scale, prediction) pairs reduced per finalize pass
__device__ void finalize_terms(const ScaleTable& tab, const float* partials, float* terms,
                               int32_t* n_bad, float* loss_out, double (*pair_acc)[6]);

__global__ void __launch_bounds__(kLossThreads)
warp_loss_fwd_kernel(const __grid_constant__ ScaleTable tab, const int32_t* __restrict__ match_idx,
                     float* partials, unsigned* ticket, float* __restrict__ terms,
                     int32_t* __restrict__ n_bad) {
    __shared__ float F[2][kTile + 1][kTile + 2];
    __shared__ float red[kLossThreads / 32][kRec];
    __shared__ double wacc[kMaxPairs][6];
    __shared__ bool is_last;
    pdl_launch_dependents();
    pdl_wait();                       // pyramid levels + match_idx come from the preceding kernels
    const TileCoord tc = decode_cta(tab);
    const int H = tab.H[tc.s], W = tab.W[tc.s], P = tab.n_pred;
    const size_t plane = (size_t)H * W;
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    stage_flow_tile<kTile + 1, kTile + 2, 0>(&F[0][0][0], flow, plane, H, W, tc.y0, tc.x0);
    __syncthreads();
    const float* prev = tab.images[tc.s] + (size_t)match_idx[tc.n] * plane;
    const float* next = tab.images[tc.s] + (size_t)match_idx[P + tc.n] * plane;
    const GridNorm gn{(float)(W - 1) / 2.f, (float)(H - 1) / 2.f, tab.rx[tc.s], tab.ry[tc.s]};
    const int lx = threadIdx.x & 31, lyb = threadIdx.x >> 5;
    const int x = tc.x0 + lx;
    const float m_r = (x + 1 < W) ? 1.f : 0.f;            // pixel has a right neighbour
    const int xc = min(x, W - 1);
    // acc: [photo, smooth_h, smooth_v, smooth_diag_a, smooth_diag_b, outborder_sum, n_bad, -]
    float acc[kRec] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // branch-free over the tile (off-image lanes/rows contribute through 0/1 masks, F is 0
    // there): the four rows' loads and MUFU chains interleave freely
#pragma unroll
    for (int r = 0; r < kTile / 8; ++r) {
        const int ly = lyb + 8 * r, y = tc.y0 + ly;
        const bool valid = (x < W) && (y < H);
        const float m_v = valid ? 1.f : 0.f;
        const float u = F[0][ly][lx], v = F[1][ly][lx];
        const Warp wp = warp_sample(next, H, W, gn, x, y, u, v, valid);
        const float pv = __ldg(prev + min(y, H - 1) * W + xc);
        acc[0] = fmaf(m_v, charb(wp.val - pv), acc[0]);                     // loss.py:72-74
        const float m_d = (y + 1 < H) ? m_v : 0.f, m_rv = m_r * m_v, m_rd = m_r * m_d;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {                                    // loss.py:76-90
            const float f = F[ch][ly][lx], fr = F[ch][ly][lx + 1];
            const float fd = F[ch][ly + 1][lx], fdr = F[ch][ly + 1][lx + 1];
            acc[1] = fmaf(m_rv, charb(fr - f), acc[1]);
            acc[2] = fmaf(m_d, charb(fd - f), acc[2]);
            acc[3] = fmaf(m_rd, charb(fdr - f), acc[3]);
            acc[4] = fmaf(m_rd, charb(fr - fd), acc[4]);
        }
        const bool bad = valid && (wp.gx < -1.f || wp.gx > 1.f || wp.gy < -1.f || wp.gy > 1.f);   // loss.py:92-94
        if (__any_sync(0xffffffffu, bad)) {                                 // warp-uniform, rare
            const float m_b = bad ? 1.f : 0.f;
            acc[5] = fmaf(m_b, charb(u) + charb(v), acc[5]);
            acc[6] += m_b;
        }
    }
    warp_sum8(acc, lx);
    if ((lx & 3) == 0) red[lyb][((lx >> 4) & 1) * 4 + ((lx >> 3) & 1) * 2 + ((lx >> 2) & 1)] = acc[0];
    __syncthreads();
    if (threadIdx.x < 7) {
        float t = 0.f;
        for (int w = 0; w < kLossThreads / 32; ++w) t += red[w][threadIdx.x];
        __stcg(partials + (size_t)blockIdx.x * kRec + threadIdx.x, t);
    }
    // the last CTA to arrive reduces all records in a FIXED order (deterministic, fp64)
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(ticket, 1u);
        is_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    finalize_terms(tab, partials, terms, n_bad, nullptr, wacc);
    if (threadIdx.x == 0) *ticket = 0u;   // self-resetting for the next call on this workspace
}

// run by ONE CTA (any whole number of warps): warp w sums the tile records of the (scale, prediction)
// pairs w, w+8, ... (all scales in flight at once: one round of loads, one barrier) in a fixed
// order in fp64; thread s then combines scale s.
__device__ void finalize_terms(const ScaleTable& tab, const float* partials, float* terms,
                               int32_t* n_bad, float* loss_out, double (*pair_acc)[6]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int P = tab.n_pred, S = tab.n_scales;
    double tot[6] = {0, 0, 0, 0, 0, 0};            // thread s < S: running totals of scale s
    for (int p0 = 0; p0 < P; p0 += kMaxPairs / DVSOF_MAX_SCALES) {
        const int pn = min(P - p0, kMaxPairs / DVSOF_MAX_SCALES);
        for (int q = warp; q < S * pn; q += n_warps) {
            const int s = q / pn, n = p0 + (q - s * pn);
            const int per = tab.tiles_x[s] * tab.tiles_y[s];
            double a[7] = {0, 0, 0, 0, 0, 0, 0};
            const float* rec = partials + ((size_t)tab.cta_begin[s] + (size_t)n * per) * kRec;
            for (int t = lane; t < per; t += 32) {
                const float4 r0 = __ldcg(reinterpret_cast<const float4*>(rec + (size_t)t * kRec));
                const float4 r1 = __ldcg(reinterpret_cast<const float4*>(rec + (size_t)t * kRec) + 1);
                a[0] += (double)r0.x; a[1] += (double)r0.y; a[2] += (double)r0.z; a[3] += (double)r0.w;
                a[4] += (double)r1.x; a[5] += (double)r1.y; a[6] += (double)r1.z;
            }
#pragma unroll
            for (int k = 0; k < 7; ++k)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
            if (lane == 0) {
                const double nb = a[6];
                for (int k = 0; k < 5; ++k) pair_acc[q][k] = a[k];
                pair_acc[q][5] = nb > 0 ? a[5] / (2.0 * nb * (double)P) : 0.0;   // loss.py:99-113
                if (n_bad) n_bad[s * P + n] = (int32_t)nb;
            }
        }
        __syncthreads();
        if ((int)threadIdx.x < S)
            for (int j = 0; j < pn; ++j)
                for (int k = 0; k < 6; ++k) tot[k] += pair_acc[threadIdx.x * pn + j][k];
        __syncthreads();
    }
    if ((int)threadIdx.x < S) {
        const int s = threadIdx.x, H = tab.H[s], W = tab.W[s];
        const double N2 = 2.0 * P;
        const double ch = N2 * H * (W - 1), cv = N2 * (H - 1) * W, cd = N2 * (H - 1) * (W - 1);
        double sm = 0;
        if (ch > 0) sm += tot[1] / ch;
        if (cv > 0) sm += tot[2] / cv;
        if (cd > 0) sm += (tot[3] + tot[4]) / cd;
        terms[s] = (float)(sm / 4.0);                                        // loss.py:87-90
        terms[S + s] = (float)(tot[0] / ((double)P * H * W));                // loss.py:33
        terms[2 * S + s] = (float)tot[5];
    }
    if (loss_out) {                       // fused pass: loss = sum_k sum_s coef[k][s] * term[k][s]
        __syncthreads();
        if (threadIdx.x == 0) {
            float l = 0.f;
            for (int k = 0; k < 3; ++k)
                for (int s = 0; s < S; ++s) l = fmaf(tab.coef[k * DVSOF_MAX_SCALES + s], terms[k * S + s], l);
            *loss_out = l;
        }
    }
}

// ------------------------------ backward -----------------------------------------------
// d(terms)/d(flow), contracted with grad_terms.  Every pixel gathers its own gradient (no
// atomics, one write): photometric via the bilinear tap differences, out-of-border via
// c'(flow)/(2 n_i N), smoothness from the 8 neighbour pairs it belongs to.  Each neighbour pair
// is owned by the top-left pixel ("cell") of its 2x2 block, its Charbonnier derivative is
// evaluated ONCE (4 pairs x 2 channels per cell) into shared memory and read by both ends:
// 9 instead of 17 lg2/ex2 pairs per pixel -- the kernel is SFU/issue bound, not HBM bound.
constexpr int kCell = kTile + 1;   // cells (x0-1 .. x0+31) x (y0-1 .. y0+31)

__global__ void __launch_bounds__(kLossThreads)
warp_loss_bwd_kernel(const __grid_constant__ ScaleTable tab, const int32_t* __restrict__ match_idx,
                     const float* __restrict__ grad_terms, const int32_t* __restrict__ n_bad) {
    __shared__ float F[2][kTile + 2][kTile + 3];   // flow, halo of 1: F[.][ly+1][lx+1]
    __shared__ float G[2][4][kCell][kCell + 1];    // pair gradients [ch][h,v,da,db][cy+1][cx+1]
    const TileCoord tc = decode_cta(tab);
    const int S = tab.n_scales, P = tab.n_pred;
    const int H = tab.H[tc.s], W = tab.W[tc.s];
    const size_t plane = (size_t)H * W;
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    float* grad = tab.grad_flow[tc.s] + (size_t)tc.n * 2 * plane;
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    stage_flow_tile<kTile + 2, kTile + 3, 1>(&F[0][0][0], flow, plane, H, W, tc.y0, tc.x0);
    const float g_sm = grad_terms[tc.s], g_ph = grad_terms[S + tc.s], g_ob = grad_terms[2 * S + tc.s];
    const float N2 = 2.f * (float)P;
    const float ch_n = N2 * (float)H * (float)(W - 1), cv_n = N2 * (float)(H - 1) * (float)W;
    const float cd_n = N2 * (float)(H - 1) * (float)(W - 1);
    const float k_h = ch_n > 0.f ? 0.25f * g_sm / ch_n : 0.f;
    const float k_v = cv_n > 0.f ? 0.25f * g_sm / cv_n : 0.f;
    const float k_d = cd_n > 0.f ? 0.25f * g_sm / cd_n : 0.f;
    __syncthreads();
    // pair gradients of the 33x33 cells: cell (cx,cy) = pixel (x0-1+cx, y0-1+cy)
    for (int i = threadIdx.x; i < kCell * kCell; i += kLossThreads) {
        const int cy = i / kCell, cx = i - cy * kCell;
        const int x = tc.x0 - 1 + cx, y = tc.y0 - 1 + cy;
        const bool in = x >= 0 && y >= 0 && x < W && y < H;
        const bool R = in && x + 1 < W, Dn = in && y + 1 < H;
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            const float f = F[ch][cy][cx], fr = F[ch][cy][cx + 1];
            const float fd = F[ch][cy + 1][cx], fdr = F[ch][cy + 1][cx + 1];
            G[ch][0][cy][cx] = R ? k_h * charb_grad(fr - f) : 0.f;
            G[ch][1][cy][cx] = Dn ? k_v * charb_grad(fd - f) : 0.f;
            G[ch][2][cy][cx] = (R && Dn) ? k_d * charb_grad(fdr - f) : 0.f;
            G[ch][3][cy][cx] = (R && Dn) ? k_d * charb_grad(fr - fd) : 0.f;
        }
    }
    const float k_ph = g_ph / ((float)P * (float)H * (float)W);
    const int nb = n_bad[tc.s * P + tc.n];
    const float k_ob = nb > 0 ? g_ob / (2.f * (float)nb * (float)P) : 0.f;
    const float* prev = tab.images[tc.s] + (size_t)match_idx[tc.n] * plane;
    const float* next = tab.images[tc.s] + (size_t)match_idx[P + tc.n] * plane;
    const GridNorm gn{(float)(W - 1) / 2.f, (float)(H - 1) / 2.f, tab.rx[tc.s], tab.ry[tc.s]};
    __syncthreads();
    const int lx = lane + 1;
    const int x = tc.x0 + lane;
    const int xc = min(x, W - 1);
#pragma unroll
    for (int r = 0; r < kTile / 8; ++r) {
        const int ly = wrp + 8 * r + 1, y = tc.y0 + ly - 1;
        const bool valid = (x < W) && (y < H);
        const float u = F[0][ly][lx], v = F[1][ly][lx];
        const Warp wp = warp_sample(next, H, W, gn, x, y, u, v, valid);
        const float cp = k_ph * charb_grad(wp.val - __ldg(prev + min(y, H - 1) * W + xc));
        float g[2] = {cp * wp.ddx, cp * wp.ddy};
        const bool bad = valid && (wp.gx < -1.f || wp.gx > 1.f || wp.gy < -1.f || wp.gy > 1.f);
        if (__any_sync(0xffffffffu, bad)) {                                 // warp-uniform, rare
            const float kb = bad ? k_ob : 0.f;
            g[0] = fmaf(kb, charb_grad(u), g[0]);
            g[1] = fmaf(kb, charb_grad(v), g[1]);
        }
        // own cell = (lx, ly) in G coordinates; left = lx-1, up = ly-1
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
            float a = G[ch][0][ly][lx - 1] - G[ch][0][ly][lx];         // horizontal pairs
            a += G[ch][1][ly - 1][lx] - G[ch][1][ly][lx];              // vertical
            a += G[ch][2][ly - 1][lx - 1] - G[ch][2][ly][lx];          // diagonal "\\"
            a += G[ch][3][ly][lx - 1] - G[ch][3][ly - 1][lx];          // diagonal "/"
            g[ch] += a;
        }
        if (valid) {
            grad[y * W + x] = g[0];
            grad[plane + y * W + x] = g[1];
        }
    }
}

// d(loss)/d(flow) *= factor, in place, for all scales in one launch; nothing is touched when the
// factor is exactly 1 (loss.backward() of the unscaled training loss).
__global__ void __launch_bounds__(256)
scale_gradients_kernel(const __grid_constant__ ScaleTable tab, const float* __restrict__ factor) {
    const float f = *factor;
    if (f == 1.f) return;
    for (int s = 0; s < tab.n_scales; ++s) {
        float* g = tab.grad_flow[s];
        const size_t n = (size_t)tab.n_pred * 2 * tab.H[s] * tab.W[s];
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            g[i] *= f;
    }
}

}  // namespace dvsof

#include "loss_strip.cuh"

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
using namespace dvsof;

// DVSOF_LOSS_LEGACY=1 selects the first-generation tile kernels (kept for A/B measurements)
static bool legacy_loss() {
    static const bool v = [] { const char* e = getenv("DVSOF_LOSS_LEGACY"); return e && e[0] == '1'; }();
    return v;
}

static int build_table(const dvsof_loss_scales_t* sc, ScaleTable* tab, bool need_grad) {
    if (!sc) return fail(DVSOF_ERR_INVALID_ARGUMENT, "loss: null scale table");
    if (sc->n_scales < 1 || sc->n_scales > DVSOF_MAX_SCALES || sc->n_pred < 1)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "loss: n_scales=%d n_pred=%d", sc->n_scales, sc->n_pred);
    tab->n_scales = sc->n_scales;
    tab->n_pred = sc->n_pred;
    int64_t begin = 0;
    for (int s = 0; s < sc->n_scales; ++s) {
        if (sc->H[s] < 1 || sc->W[s] < 1 || !sc->flow[s] || !sc->images[s] ||
            (need_grad && !sc->grad_flow[s]))
            return fail(DVSOF_ERR_INVALID_ARGUMENT, "loss: bad scale %d", s);
        tab->H[s] = sc->H[s];
        tab->W[s] = sc->W[s];
        tab->tiles_x[s] = (sc->W[s] + kTile - 1) / kTile;
        tab->tiles_y[s] = (sc->H[s] + kTile - 1) / kTile;
        const float hx = (float)(sc->W[s] - 1) / 2.f, hy = (float)(sc->H[s] - 1) / 2.f;
        // a 1-pixel side makes the reference divide by zero (NaN/inf grid); here: NaN grid as well
        tab->rx[s] = dvsof_div_const_ok(hx) ? dvsof_recip_rn(hx) : NAN;
        tab->ry[s] = dvsof_div_const_ok(hy) ? dvsof_recip_rn(hy) : NAN;
        tab->flow[s] = sc->flow[s];
        tab->images[s] = sc->images[s];
        tab->grad_flow[s] = sc->grad_flow[s];
        tab->cta_begin[s] = (int)begin;
        begin += (int64_t)tab->tiles_x[s] * tab->tiles_y[s] * sc->n_pred;
        // 2^23 tiles: the float tile decode of the strip kernels stays exact; 2*H*W < 2^31: 32-bit
        // offsets inside one prediction's flow block
        if (begin > (1 << 23)) return fail(DVSOF_ERR_INVALID_ARGUMENT, "loss: too many tiles");
        if ((int64_t)sc->H[s] * sc->W[s] >= (1 << 30))
            return fail(DVSOF_ERR_INVALID_ARGUMENT, "loss: scale %d too large", s);
    }
    for (int s = sc->n_scales; s <= DVSOF_MAX_SCALES; ++s) tab->cta_begin[s] = (int)begin;
    for (int i = 0; i < 3 * DVSOF_MAX_SCALES; ++i) tab->coef[i] = 0.f;
    return DVSOF_OK;
}

static size_t partial_bytes(const ScaleTable& tab) {
    return round_up((size_t)tab.cta_begin[DVSOF_MAX_SCALES] * kRec * sizeof(float), 256);
}

extern "C" size_t dvsof_loss_workspace_bytes(const dvsof_loss_scales_t* sc) {
    ScaleTable tab;
    if (build_table(sc, &tab, false)) return 0;
    return partial_bytes(tab) + 256;   // tile records + the finalize ticket
}

static int launch_forward(const ScaleTable& tab, const int32_t* match_idx, float* terms,
                          int32_t* n_bad, void* workspace, cudaStream_t stream, bool pdl) {
    float* partials = reinterpret_cast<float*>(workspace);
    unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + partial_bytes(tab));
    if (legacy_loss())
        return launch_kernel("warp_loss_fwd_kernel", warp_loss_fwd_kernel, tab.cta_begin[DVSOF_MAX_SCALES],
                             kLossThreads, 0, stream, pdl, tab, match_idx, partials, ticket, terms, n_bad);
    return launch_kernel("warp_loss_strip_kernel<fwd>", warp_loss_strip_kernel<true, false>,
                         tab.cta_begin[DVSOF_MAX_SCALES], kSThreads, 0, stream, pdl, tab, match_idx,
                         (const float*)nullptr, (const int32_t*)nullptr, partials, ticket, terms, n_bad,
                         (float*)nullptr);
}

extern "C" int dvsof_warp_loss_forward(const dvsof_loss_scales_t* sc, const int32_t* match_idx,
                                       float* terms, int32_t* n_bad, void* workspace,
                                       void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, false)) return rc;
    if (!match_idx || !terms || !n_bad || !workspace)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "warp_loss_forward: null buffer");
    cudaStream_t stream = (cudaStream_t)stream_;
    cudaError_t e = cudaMemsetAsync(static_cast<char*>(workspace) + partial_bytes(tab), 0, 4, stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "loss ticket memset: %s", cudaGetErrorString(e));
    return launch_forward(tab, match_idx, terms, n_bad, workspace, stream, false);
}

extern "C" int dvsof_warp_loss_backward(const dvsof_loss_scales_t* sc, const int32_t* match_idx,
                                        const float* grad_terms, const int32_t* n_bad,
                                        void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, true)) return rc;
    if (!match_idx || !grad_terms || !n_bad)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "warp_loss_backward: null buffer");
    if (legacy_loss())
        return launch_kernel("warp_loss_bwd_kernel", warp_loss_bwd_kernel, tab.cta_begin[DVSOF_MAX_SCALES],
                             kLossThreads, 0, (cudaStream_t)stream_, false, tab, match_idx, grad_terms, n_bad);
    return launch_kernel("warp_loss_strip_kernel<bwd>", warp_loss_strip_kernel<false, true>,
                         tab.cta_begin[DVSOF_MAX_SCALES], kSThreads, 0, (cudaStream_t)stream_, false, tab,
                         match_idx, grad_terms, n_bad, (float*)nullptr, (unsigned*)nullptr, (float*)nullptr,
                         (int32_t*)nullptr, (float*)nullptr);
}

static int launch_match(const float* flow_ts, const int64_t* flow_sample_idx, int64_t P,
                        const float* timestamps, const int64_t* sample_idx, int64_t D,
                        int32_t* match_idx, uint32_t* status, unsigned* ticket, cudaStream_t stream) {
    if (P < 1 || D < 1 || P > (1 << 24) || D > (1 << 24) || !flow_ts || !flow_sample_idx ||
        !timestamps || !sample_idx || !match_idx || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "match_timestamps: bad arguments");
    const int threads = 128, blocks = (int)((2 * P + threads - 1) / threads);
    return launch_kernel("match_timestamps_kernel", match_timestamps_kernel, blocks, threads, 0, stream,
                         false, flow_ts, flow_sample_idx, (int)P, timestamps, sample_idx, (int)D,
                         match_idx, status, ticket);
}

extern "C" int dvsof_match_timestamps(const float* flow_ts, const int64_t* flow_sample_idx,
                                      int64_t P, const float* timestamps,
                                      const int64_t* sample_idx, int64_t D, int32_t* match_idx,
                                      uint32_t* status, void* stream_) {
    return launch_match(flow_ts, flow_sample_idx, P, timestamps, sample_idx, D, match_idx, status,
                        nullptr, (cudaStream_t)stream_);
}

static int launch_resize(const float* src, int64_t D, int32_t Hin, int32_t Win, float* dst,
                         int32_t Hout, int32_t Wout, cudaStream_t stream, bool pdl) {
    if (D < 1 || Hin < 1 || Win < 1 || Hout < 1 || Wout < 1 || !src || !dst)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "resize_bilinear: bad arguments");
    const int64_t total = D * Hout * Wout;
    int64_t blocks = (total + 255) / 256;
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    return launch_kernel("resize_bilinear_kernel", resize_bilinear_kernel, (int)blocks, 256, 0, stream,
                         pdl, src, (int)Hin, (int)Win, dst, (int)Hout, (int)Wout, total);
}

extern "C" int dvsof_resize_bilinear(const float* src, int64_t D, int32_t Hin, int32_t Win,
                                     float* dst, int32_t Hout, int32_t Wout, void* stream_) {
    return launch_resize(src, D, Hin, Win, dst, Hout, Wout, (cudaStream_t)stream_, false);
}

// matching + chained pyramid (utils/loss.py:182-210), PDL-chained on the stream
static int launch_match_and_pyramid(const dvsof_loss_scales_t* sc, const dvsof_loss_frames_t* fr,
                                    int32_t* match_idx, uint32_t* status, unsigned* ticket,
                                    cudaStream_t stream) {
    if (int rc = launch_match(fr->flow_ts, fr->flow_sample_idx, sc->n_pred, fr->timestamps,
                              fr->sample_idx, fr->n_images, match_idx, status, ticket, stream))
        return rc;
    const float* src = fr->images;
    int Hin = fr->H0, Win = fr->W0;
    for (int s = 0; s < sc->n_scales; ++s) {                               // loss.py:207-210
        float* dst = const_cast<float*>(sc->images[s]);
        if (int rc = launch_resize(src, fr->n_images, Hin, Win, dst, sc->H[s], sc->W[s], stream, true))
            return rc;
        src = dst; Hin = sc->H[s]; Win = sc->W[s];
    }
    return DVSOF_OK;
}

static bool frames_ok(const dvsof_loss_frames_t* fr) {
    return fr && fr->images && fr->n_images >= 1 && fr->H0 >= 1 && fr->W0 >= 1;
}

// Whole forward of utils/loss.py:179-214 in one call: timestamp matching, the chained pyramid
// (each level resized from the previous one, line 210) into the caller's level buffers
// scales->images[s], the fused per-scale terms.  Kernels are chained with programmatic dependent
// launch, so there is no host work and no idle gap between them.
extern "C" int dvsof_multiscale_loss_forward(const dvsof_loss_scales_t* sc,
                                             const dvsof_loss_frames_t* fr, int32_t* match_idx,
                                             float* terms, int32_t* n_bad, void* workspace,
                                             uint32_t* status, void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, false)) return rc;
    if (!frames_ok(fr) || !match_idx || !terms || !n_bad || !workspace || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "multiscale_loss_forward: bad arguments");
    cudaStream_t stream = (cudaStream_t)stream_;
    unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + partial_bytes(tab));
    if (int rc = launch_match_and_pyramid(sc, fr, match_idx, status, ticket, stream)) return rc;
    return launch_forward(tab, match_idx, terms, n_bad, workspace, stream, true);
}

// The training pass: the same forward plus d(loss)/d(flow) of
// loss = sum_k sum_s coef[k*n_scales + s] * term[k][s] (utils/training.py:12-24 gives
// coef = w_k / n_scales) written to scales->grad_flow in the SAME sweep over the flow.
extern "C" int dvsof_multiscale_loss_fused(const dvsof_loss_scales_t* sc,
                                           const dvsof_loss_frames_t* fr, const float* coef,
                                           int32_t* match_idx, float* terms, float* loss,
                                           int32_t* n_bad, void* workspace, uint32_t* status,
                                           void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, true)) return rc;
    if (!frames_ok(fr) || !coef || !match_idx || !terms || !loss || !n_bad || !workspace || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "multiscale_loss_fused: bad arguments");
    for (int k = 0; k < 3; ++k)
        for (int s = 0; s < sc->n_scales; ++s) tab.coef[k * DVSOF_MAX_SCALES + s] = coef[k * sc->n_scales + s];
    cudaStream_t stream = (cudaStream_t)stream_;
    float* partials = reinterpret_cast<float*>(workspace);
    unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + partial_bytes(tab));
    cudaError_t e = cudaMemsetAsync(n_bad, 0, sizeof(int32_t) * sc->n_scales * sc->n_pred, stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "n_bad memset: %s", cudaGetErrorString(e));
    if (int rc = launch_match_and_pyramid(sc, fr, match_idx, status, ticket, stream)) return rc;
    const int grid = tab.cta_begin[DVSOF_MAX_SCALES];
    if (int rc = launch_kernel("outborder_count_kernel", outborder_count_kernel, grid, kSThreads, 0,
                               stream, true, tab, n_bad))
        return rc;
    return launch_kernel("warp_loss_strip_kernel<fused>", warp_loss_strip_kernel<true, true>, grid,
                         kSThreads, 0, stream, true, tab, (const int32_t*)match_idx,
                         (const float*)nullptr, (const int32_t*)n_bad, partials, ticket, terms,
                         (int32_t*)nullptr, loss);
}

// The fused sweep on ready-made pyramid levels (single-scale Loss objects, kernel-level timing):
// count_pass != 0 runs the out-of-border counting pre-pass into n_bad first; 0 reuses n_bad.
extern "C" int dvsof_warp_loss_fused(const dvsof_loss_scales_t* sc, const int32_t* match_idx,
                                     const float* coef, float* terms, float* loss, int32_t* n_bad,
                                     void* workspace, int32_t count_pass, void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, true)) return rc;
    if (!match_idx || !coef || !terms || !loss || !n_bad || !workspace)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "warp_loss_fused: null buffer");
    for (int k = 0; k < 3; ++k)
        for (int s = 0; s < sc->n_scales; ++s) tab.coef[k * DVSOF_MAX_SCALES + s] = coef[k * sc->n_scales + s];
    cudaStream_t stream = (cudaStream_t)stream_;
    float* partials = reinterpret_cast<float*>(workspace);
    unsigned* ticket = reinterpret_cast<unsigned*>(static_cast<char*>(workspace) + partial_bytes(tab));
    const int grid = tab.cta_begin[DVSOF_MAX_SCALES];
    cudaError_t e = cudaMemsetAsync(ticket, 0, 4, stream);
    if (e == cudaSuccess && count_pass)
        e = cudaMemsetAsync(n_bad, 0, sizeof(int32_t) * sc->n_scales * sc->n_pred, stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "warp_loss_fused memset: %s", cudaGetErrorString(e));
    if (count_pass)
        if (int rc = launch_kernel("outborder_count_kernel", outborder_count_kernel, grid, kSThreads, 0,
                                   stream, false, tab, n_bad))
            return rc;
    return launch_kernel("warp_loss_strip_kernel<fused>", warp_loss_strip_kernel<true, true>, grid,
                         kSThreads, 0, stream, false, tab, match_idx, (const float*)nullptr,
                         (const int32_t*)n_bad, partials, ticket, terms, (int32_t*)nullptr, loss);
}

// grad_flow[s] *= *factor for all scales (backward of the fused pass when the incoming gradient
// of the loss is not 1); a no-op launch when it is.
extern "C" int dvsof_scale_gradients(const dvsof_loss_scales_t* sc, const float* factor, void* stream_) {
    ScaleTable tab;
    if (!sc || !factor) return fail(DVSOF_ERR_INVALID_ARGUMENT, "scale_gradients: null argument");
    dvsof_loss_scales_t tmp = *sc;
    for (int s = 0; s < tmp.n_scales && s < DVSOF_MAX_SCALES; ++s) {   // flow / images are not read here
        if (!tmp.flow[s]) tmp.flow[s] = tmp.grad_flow[s];
        if (!tmp.images[s]) tmp.images[s] = tmp.grad_flow[s];
    }
    if (int rc = build_table(&tmp, &tab, true)) return rc;
    return launch_kernel("scale_gradients_kernel", scale_gradients_kernel, kNumSMs * 8, 256, 0,
                         (cudaStream_t)stream_, false, tab, factor);
}
