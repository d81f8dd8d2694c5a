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
struct ScaleTable {
    int n_scales, n_pred;
    int H[DVSOF_MAX_SCALES], W[DVSOF_MAX_SCALES];
    int tiles_x[DVSOF_MAX_SCALES], tiles_y[DVSOF_MAX_SCALES];
    float inv_tx[DVSOF_MAX_SCALES], inv_per[DVSOF_MAX_SCALES];   // 1 / tiles_x, 1 / (tiles_x * tiles_y)
    int cta_begin[DVSOF_MAX_SCALES + 1];
    float hx[DVSOF_MAX_SCALES], hy[DVSOF_MAX_SCALES];   // half extents (W-1)/2, (H-1)/2
    float rx[DVSOF_MAX_SCALES], ry[DVSOF_MAX_SCALES];   // their correctly rounded reciprocals (NaN: degenerate side)
    // gradient normalisers 2 alpha / count: [0..2] smoothness h / v / diagonal (incl. the 1/4),
    // [3] photometric 1/(P H W), [4] border 1/(2 P) (divided by n_i in the kernel)
    float norm[5][DVSOF_MAX_SCALES];
    float lk[3][DVSOF_MAX_SCALES];                      // log2(norm[0..2]); -inf for a direction without pairs
    float coef[3 * DVSOF_MAX_SCALES];                   // fused pass: d(loss)/d(term[k][s]) at [k * MAX + s]
    const float* flow[DVSOF_MAX_SCALES];
    const float* images[DVSOF_MAX_SCALES];
    float* grad_flow[DVSOF_MAX_SCALES];
    // the sweep's work units: segments of 32 columns x seg_rows[s] rows, one warp (= CTA) each;
    // seg_per[s] segments per (scale, prediction), seg_begin[s] = first block of scale s
    int seg_rows[DVSOF_MAX_SCALES], seg_per[DVSOF_MAX_SCALES];
    float inv_seg_per[DVSOF_MAX_SCALES];
    int seg_begin[DVSOF_MAX_SCALES + 1];
    int use_tma[DVSOF_MAX_SCALES];                      // flow / level of scale s have tensor maps
};

struct TileCoord { int s, n, y0, x0; };
// Bilinear sampling notes (the code is request_taps + phase 3 of the strip kernel, loss_strip.cuh):
// the position is the one the reference's grid_sample visits for pixel (x,y) displaced by (u,v):
//   g  = (pix + flow) / ((S-1)/2) - 1        IEEE division (utils/loss.py:150-156), exact_div.h
//   ix = (g + 1) * ((S-1)/2)                 ATen CPU unnormalize for align_corners=True
// coordinates are clamped in float (huge / non-finite values never reach the int conversion), the
// four zero-padded taps use unsigned range tests and 32-bit offsets inside one image plane, and
// the weighted sum is evaluated with the same operations as ATen's vectorised CPU grid_sampler
// (weights s*e, s*w, n*e, n*w; nw*w0 + ne*w1 + sw*w2 + se*w3 is compiled there to mul + 3 chained
// FMAs -- pinned by tests/test_oracle_golden.py::test_aten_bilinear_operation_order).  Where
// warped ~= prev the Charbonnier derivative amplifies a 1-ulp difference of the sample ~1e3
// times, so the interpolation is kept bit-compatible with the reference's CPU path.
struct GridNorm { float hx, hy, rx, ry; };   // half extents (S-1)/2 and their correctly rounded reciprocals

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
                                        uint32_t* __restrict__ status, unsigned* tickets,
                                        int n_tickets) {
    pdl_launch_dependents();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;   // i = k * P + p
    if (tickets)                                           // arms the reduction of the loss kernel
        for (int k = i; k < n_tickets; k += gridDim.x * blockDim.x) tickets[k] = 0u;
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

// ------------------------------ deterministic reduction of the segment records ---------
// record of one sweep segment (warp): [photo, smoothness, outborder_sum, n_bad].
// Two levels, both in loss_reduce_kernel (launched behind the sweep):
//   pair level  -- one warp per (scale, prediction) pair sums that pair's records in a fixed order
//                  in fp64 and writes the pair record + n_i;
//   final level -- the CTA that arrives LAST at the ticket (release/acquire at gpu scope, issued
//                  by one thread; the data is read back with L2 loads) sums the pair records per
//                  scale, again in a fixed order, and writes the 3*S terms (+ the weighted loss).
// Results do not depend on arrival order; the ticket resets itself for the next call.
constexpr int kRec = 4;            // floats per segment record
constexpr int kPairRec = 4;        // doubles per pair record: [photo, smoothness, outborder/(2 n_i P), -]

__device__ __forceinline__ unsigned ticket_arrive(unsigned* t) {
    unsigned before;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(before) : "l"(t) : "memory");
    return before;
}
__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// one warp: the records of pair (s, n) in a fixed order (lane l takes records l, l + 32, ...; eight
// of them in flight: every record is an L2 round trip), fp64, then a butterfly over the lanes
__device__ void reduce_pair(const ScaleTable& tab, int s, int n, const float* partials,
                            double* pair_rec, int32_t* n_bad) {
    const int lane = threadIdx.x & 31;
    const int P = tab.n_pred, per = tab.seg_per[s];
    const float* rec = partials + ((size_t)tab.seg_begin[s] + (size_t)n * per) * kRec;
    double a0 = 0, a1 = 0, a5 = 0, a6 = 0;         // photo, smoothness, border sum, border count
    constexpr int kInFlight = 8;
    for (int t0 = lane; t0 < per; t0 += kInFlight * 32) {
        float4 r[kInFlight];
#pragma unroll
        for (int j = 0; j < kInFlight; ++j) {
            const int t = t0 + j * 32;
            r[j] = t < per ? __ldcg(reinterpret_cast<const float4*>(rec + (size_t)t * kRec)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int j = 0; j < kInFlight; ++j) {
            a0 += (double)r[j].x; a1 += (double)r[j].y; a5 += (double)r[j].z; a6 += (double)r[j].w;
        }
    }
    a0 = warp_sum_f64(a0); a1 = warp_sum_f64(a1); a5 = warp_sum_f64(a5); a6 = warp_sum_f64(a6);
    if (lane == 0) {
        double* out = pair_rec + (size_t)(s * P + n) * kPairRec;
        __stcg(out + 0, a0); __stcg(out + 1, a1);
        __stcg(out + 2, a6 > 0 ? a5 / (2.0 * a6 * (double)P) : 0.0);                 // loss.py:99-113
        if (n_bad) n_bad[s * P + n] = (int32_t)a6;
    }
}

// one CTA (whole warps): per scale the pair records in a fixed order -> the 3*S terms (+ the loss)
__device__ void reduce_final(const ScaleTable& tab, const double* pair_rec, float* terms,
                             float* loss_out, float* term_sm /* shared, 3 * DVSOF_MAX_SCALES */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int P = tab.n_pred, S = tab.n_scales;
    for (int s = warp; s < S; s += n_warps) {
        double a[3] = {0, 0, 0};
        for (int n = lane; n < P; n += 32) {
            const double* r = pair_rec + (size_t)(s * P + n) * kPairRec;
#pragma unroll
            for (int k = 0; k < 3; ++k) a[k] += __ldcg(r + k);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) a[k] = warp_sum_f64(a[k]);
        if (lane == 0) {
            const int H = tab.H[s], W = tab.W[s];
            // the sweep accumulates k c(d) with k = 2 alpha / (4 x count) per direction (loss.py:87-90)
            const float sm = (float)(a[1] / (2.0 * (double)kAlpha));
            const float ph = (float)(a[0] / ((double)P * H * W));                  // loss.py:33
            const float ob = (float)a[2];
            terms[s] = sm; terms[S + s] = ph; terms[2 * S + s] = ob;
            term_sm[s] = sm; term_sm[DVSOF_MAX_SCALES + s] = ph; term_sm[2 * DVSOF_MAX_SCALES + s] = ob;
        }
    }
    if (loss_out) {                       // fused pass: loss = sum_k sum_s coef[k][s] * term[k][s]
        __syncthreads();
        if (threadIdx.x == 0) {
            float l = 0.f;
            for (int k = 0; k < 3; ++k)
                for (int s = 0; s < S; ++s)
                    l = fmaf(tab.coef[k * DVSOF_MAX_SCALES + s], term_sm[k * DVSOF_MAX_SCALES + s], l);
            *loss_out = l;
        }
    }
}

// workspace carved by the host: segment records | pair records | ticket | per-tile border counts
struct LossWorkspace {
    float* partials;
    double* pair_rec;
    unsigned* tickets;     // [0] = pairs finished (loss_reduce_kernel)
    int32_t* tile_bad;     // out-of-border count of every tile (multiscale fused pass)
};

// The reduction runs as its own small kernel behind the sweep (programmatic dependent launch: it is
// resident and waiting when the sweep's last warp retires).  A ticket per sweep warp -- a gpu-scope
// release fence + atomic at the end of every 1-warp CTA -- was measured at 13-25 us of a 30-100 us
// sweep (gpurun_out/r02h: every warp holds its SM slot while its gradient stores drain); a kernel
// boundary orders the records for free.  One CTA per (scale, prediction) pair sums that pair's
// segment records; the last of these S*P CTAs (one ticket each) writes the terms and the loss.
constexpr int kReduceThreads = 128;                        // 4 warps = 4 pairs per CTA
__global__ void __launch_bounds__(kReduceThreads)
loss_reduce_kernel(const __grid_constant__ ScaleTable tab, const LossWorkspace ws, float* terms,
                   int32_t* __restrict__ n_bad, float* __restrict__ loss_out) {
    __shared__ float term_sm[3 * DVSOF_MAX_SCALES];
    __shared__ bool is_last;
    pdl_launch_dependents();
    pdl_wait();                                            // the sweep's records are complete and visible
    const int P = tab.n_pred, n_pairs = tab.n_scales * P;
    const int pair = blockIdx.x * (kReduceThreads / 32) + (threadIdx.x >> 5);
    if (pair < n_pairs) {                                  // warp-uniform
        const int s = pair / P;
        reduce_pair(tab, s, pair - s * P, ws.partials, ws.pair_rec, n_bad);
    }
    __syncthreads();
    if (threadIdx.x == 0) is_last = ticket_arrive(ws.tickets) == gridDim.x - 1u;
    __syncthreads();
    if (!is_last) return;
    __threadfence();                                       // acquire side (the finishing CTA only)
    reduce_final(tab, ws.pair_rec, terms, loss_out, term_sm);
    if (threadIdx.x == 0) ws.tickets[0] = 0u;              // self-resetting
}

// d(loss)/d(flow) *= factor, in place, for all scales in one launch; nothing is touched when the
// factor is exactly 1 (loss.backward() of the unscaled training loss).
__global__ void __launch_bounds__(256)
scale_gradients_kernel(const __grid_constant__ ScaleTable tab, const float* __restrict__ factor) {
    // launched with programmatic serialization: `factor` is an input of the call, so the common
    // case (d/d(loss) == 1) returns while the kernels before it are still running
    const float f = *factor;
    if (f == 1.f) return;
    pdl_wait();                                   // the gradients come from the sweep
    for (int s = 0; s < tab.n_scales; ++s) {
        float* g = tab.grad_flow[s];
        const size_t n = (size_t)tab.n_pred * 2 * tab.H[s] * tab.W[s];
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
            g[i] *= f;
    }
}

}  // namespace dvsof

#include "loss_ring.cuh"

namespace dvsof {
// ------------------------------ bilinear resize (align_corners) ------------------------
// A thread produces 4 consecutive pixels of kRY consecutive output rows: the column taps and
// weights are computed once per thread, a row costs two base pointers, 16 loads and one 16-byte
// store (when the row pitch allows).  A CTA covers 128 columns x 32 rows of the (D*Hout) x Wout
// output; 32-bit index arithmetic, one integer division per thread.
constexpr int kRX = 4, kRY = 4, kRWarps = 8;
//
// Side job (the fused training pass, last pyramid level only): the first `side_ctas` CTAs of the
// launch count the out-of-border pixels of the flow instead (outborder_count_tile) -- work that
// depends on nothing in the chain, so it runs before / while the resize CTAs wait for the
// previous level, and the launch as a whole still completes only after that level has.
template <bool kSide>
__global__ void __launch_bounds__(32 * kRWarps)
resize_bilinear_kernel(const float* __restrict__ src, int Hin, int Win, float* __restrict__ dst,
                       int Hout, int Wout, int rows /* D * Hout */, int col_ctas,
                       const __grid_constant__ ScaleTable tab, int32_t* __restrict__ n_bad, int side_ctas,
                       int bad_per_tile) {
    pdl_launch_dependents();
    int cta = blockIdx.x;
    if (kSide) {
        __shared__ int cnt[kRWarps];
        if (cta < side_ctas) {
            outborder_count_tile<kRWarps>(tab, n_bad, bad_per_tile != 0, cta, cnt);
            return;
        }
        cta -= side_ctas;
    }
    pdl_wait_for(src);                           // src may be the previous level of the chain
    const int cta_y = cta / col_ctas, cta_x = cta - cta_y * col_ctas;
    const int row0 = (cta_y * kRWarps + (threadIdx.x >> 5)) * kRY;
    const int xb = (cta_x * 32 + (threadIdx.x & 31)) * kRX;
    if (row0 >= rows || xb >= Wout) return;
    const float sy = Hout > 1 ? (float)(Hin - 1) / (float)(Hout - 1) : 0.f;
    const float sx = Wout > 1 ? (float)(Win - 1) / (float)(Wout - 1) : 0.f;
    int x0[kRX], x1[kRX];
    float lx[kRX], hx[kRX];
#pragma unroll
    for (int k = 0; k < kRX; ++k) {
        const int x = min(xb + k, Wout - 1);     // lanes past the row end recompute its last pixel
        const float fx = sx * (float)x;
        int xi = (int)fx;
        xi = xi < Win - 1 ? xi : Win - 1;
        x0[k] = xi;
        x1[k] = xi + (xi < Win - 1);
        lx[k] = fminf(fmaxf(fx - (float)xi, 0.f), 1.f);
        hx[k] = 1.f - lx[k];
    }
    const bool vec = (Wout & 3) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
    int d = row0 / Hout, y = row0 - d * Hout;
#pragma unroll
    for (int j = 0; j < kRY; ++j) {
        const int row = row0 + j;
        if (row >= rows) break;
        const float fy = sy * (float)y;
        int y0 = (int)fy;
        y0 = y0 < Hin - 1 ? y0 : Hin - 1;
        const int y1 = y0 + (y0 < Hin - 1);
        const float ly = fminf(fmaxf(fy - (float)y0, 0.f), 1.f), hy = 1.f - ly;
        const float* r0 = src + (size_t)(d * Hin + y0) * Win;
        const float* r1 = src + (size_t)(d * Hin + y1) * Win;
        float* out = dst + (size_t)row * Wout + xb;
        asm volatile("" : "+l"(r0), "+l"(r1), "+l"(out));   // finished row pointers: tap = base + 32-bit column
        float o[kRX];
#pragma unroll
        for (int k = 0; k < kRX; ++k) {
            const float a = __ldg(r0 + x0[k]), b = __ldg(r0 + x1[k]);
            const float c = __ldg(r1 + x0[k]), e = __ldg(r1 + x1[k]);
            // operation order of ATen's CPU upsample_bilinear2d as compiled (two FMAs per row pair,
            // one for the rows): bit-identical to the reference's F.interpolate on the upsampling
            // levels of the chain (pinned on CPU by
            // tests/test_oracle_golden.py::test_aten_upsample_operation_order)
            const float top = __fmaf_rn(hx[k], a, __fmul_rn(lx[k], b)), bot = __fmaf_rn(hx[k], c, __fmul_rn(lx[k], e));
            o[k] = __fmaf_rn(hy, top, __fmul_rn(ly, bot));
        }
        if (vec) {
            *reinterpret_cast<float4*>(out) = make_float4(o[0], o[1], o[2], o[3]);
        } else {
#pragma unroll
            for (int k = 0; k < kRX; ++k)
                if (xb + k < Wout) out[k] = o[k];
        }
        if (++y == Hout) { y = 0; ++d; }
    }
}

// ------------------------------ pyramid chain in one launch ----------------------------
// utils/loss.py:207-210 resizes level s from level s-1, so the levels of ONE frame form a
// dependency chain, but frames -- and row slabs of a frame -- are independent: `slabs` CTAs
// (1024 threads each) share a frame.  CTA g owns rows [H*g/slabs, H*(g+1)/slabs) of every level;
// walking the chain backwards gives the rows of each earlier level those depend on, which the CTA
// computes for itself (a few halo rows are computed twice, never stored twice).  A level stays in
// shared memory (two alternating buffers) for the next one to read -- the same per-pixel
// arithmetic as resize_bilinear_kernel on the same values, hence the same bits -- while the owned
// rows also go to global memory for the loss kernel.  One launch replaces the first `n_levels`
// resize launches (normally all of them) AND the matching kernel: CTA 0 matches flows to frames
// (utils/loss.py:182-206) and arms the tickets.  Side CTAs (last in the grid) run the border
// count when this launch is the last of the pyramid.
struct PyramidArgs {
    const float* src; int D, H0, W0;
    int n_levels;                                  // levels produced by this launch (>= 1)
    int slabs;                                     // CTAs per frame
    int H[DVSOF_MAX_SCALES], W[DVSOF_MAX_SCALES];
    float sy[DVSOF_MAX_SCALES], sx[DVSOF_MAX_SCALES];   // source step per output row / column (host-computed)
    float* dst[DVSOF_MAX_SCALES];
    int buf_b_offset;                              // floats: buffer A = [0, off), buffer B = [off, ...)
    const float* flow_ts; const int64_t* flow_sample_idx; int P;
    const float* timestamps; const int64_t* sample_idx;
    int32_t* match_idx; uint32_t* status; unsigned* tickets; int n_tickets;
};
constexpr int kPyrThreads = 1024;                   // at most; the host launches fewer warps for small slabs


struct RowRange { int lo, hi; };                   // [lo, hi); empty when lo >= hi

// source row of output row y (align_corners=True), exactly as the resize arithmetic picks it
__host__ __device__ inline float pyr_step(int n_in, int n_out) {
    return n_out > 1 ? (float)(n_in - 1) / (float)(n_out - 1) : 0.f;
}
__host__ __device__ inline int pyr_src_row(int Hin, float sy, int y) {
    int y0 = (int)(sy * (float)y);
    return y0 < Hin - 1 ? y0 : Hin - 1;
}

// rows of every level CTA `g` of a frame has to produce: need[l] (superset of its own rows)
__host__ __device__ inline void pyr_plan_rows(const PyramidArgs& a, int g, RowRange* own, RowRange* need) {
    for (int l = 0; l < a.n_levels; ++l) {
        own[l].lo = (int)((unsigned)(a.H[l] * g) / (unsigned)a.slabs);          // H < 2^26, slabs <= 16
        own[l].hi = (int)((unsigned)(a.H[l] * (g + 1)) / (unsigned)a.slabs);
    }
    need[a.n_levels - 1] = own[a.n_levels - 1];
    for (int l = a.n_levels - 1; l > 0; --l) {
        RowRange r = own[l - 1];
        if (need[l].lo < need[l].hi) {
            const int Hin = a.H[l - 1];
            const int lo = pyr_src_row(Hin, a.sy[l], need[l].lo);
            int hi = pyr_src_row(Hin, a.sy[l], need[l].hi - 1) + 2;      // y1 = y0 + 1, exclusive
            hi = hi < Hin ? hi : Hin;
            if (r.lo >= r.hi) { r.lo = lo; r.hi = hi; }
            else { r.lo = lo < r.lo ? lo : r.lo; r.hi = hi > r.hi ? hi : r.hi; }
        }
        need[l - 1] = r;
    }
}

// Rows [nd.lo, nd.hi) of one level for one frame, from `in` (row in_row0 of the source level is
// its first row).  A warp task is kPyrRows rows x 128 columns: a lane owns 4 consecutive columns and
// computes their taps and weights once per task.  The horizontal interpolation of a SOURCE row is
// kept while consecutive output rows keep using it (when upsampling by 2, 8 output rows touch 5
// source rows, not 16) -- it is the same expression on the same operands whichever output row
// consumes it, so the bits do not change.  Owned rows go to `out` (global), all rows to `keep`
// (shared, relative to nd.lo) when there is a next level.
constexpr int kPyrRows = 8;
__device__ __forceinline__ void pyr_hlerp(const float* row, const int* ox0, const int* ox1,
                                          const float* lx, const float* hx, float* h) {
    const char* base = reinterpret_cast<const char*>(row);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float p = *reinterpret_cast<const float*>(base + ox0[j]);
        const float q = *reinterpret_cast<const float*>(base + ox1[j]);
        h[j] = __fmaf_rn(hx[j], p, __fmul_rn(lx[j], q));
    }
}

__device__ __forceinline__ void pyr_level_rows(const float* in, int Hin, int Win, int in_row0,
                                               float* __restrict__ out, float* keep, int Wout,
                                               float sy, float sx, RowRange nd, RowRange ow) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vec = (Wout & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0;
    // rows per task (a power of two): up to kPyrRows, fewer when that would leave warps without a task
    const int n_rows = nd.hi - nd.lo, n_warps = blockDim.x >> 5;
    const int col_chunks = (Wout + 127) >> 7;
    int rpt_log = 3;
    while (rpt_log > 1 && (n_rows >> rpt_log) * col_chunks < n_warps) --rpt_log;
    const int rpt = 1 << rpt_log;
    const int n_tasks = ((n_rows + rpt - 1) >> rpt_log) * col_chunks;
    const float inv_cc = 1.f / (float)col_chunks;
    asm volatile("" : "+l"(out));
    for (int task = warp; task < n_tasks; task += n_warps) {
        // task = row task * col_chunks + column chunk; the float quotient is exact below 2^21 tasks
        const int rt = n_tasks < (1 << 21) ? (int)(((float)task + 0.5f) * inv_cc) : task / col_chunks;
        const int ybeg = nd.lo + (rt << rpt_log);
        const int xq = ((task - rt * col_chunks) * 32 + lane) * 4;
        if (xq >= Wout) continue;
        int ox0[4], ox1[4];                                   // byte offsets of the taps in a source row
        float lx[4], hx[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float fx = sx * (float)(xq + j);
            int t = (int)fx;
            t = t < Win - 1 ? t : Win - 1;
            ox0[j] = t * 4;
            ox1[j] = (t + (t < Win - 1)) * 4;
            lx[j] = fminf(fmaxf(fx - (float)t, 0.f), 1.f);
            hx[j] = 1.f - lx[j];
        }
        const int yend = min(ybeg + rpt, nd.hi);
        // walk the SOURCE row pairs: all output rows between the same two source rows are emitted
        // from one pair of horizontal interpolations, and the lower one becomes the next upper one
        int y = ybeg, have = -1;
        int off = y * Wout + xq;                              // element offset of (y, xq) in the level
        float* keep_rel = keep ? keep - nd.lo * Wout : nullptr;   // keep_rel + off = the kept copy
        const unsigned own_rows = (unsigned)(ow.hi > ow.lo ? ow.hi - ow.lo : 0);
        float fy = sy * (float)y;
        int y0 = (int)fy;
        y0 = y0 < Hin - 1 ? y0 : Hin - 1;
        float top[4], bot[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1                                      // compact code: the kernel starts with a cold i-cache
        while (y < yend) {                                    // warp-uniform control flow throughout
            const int y1 = y0 + (y0 < Hin - 1);
            if (y0 == have) {
#pragma unroll
                for (int j = 0; j < 4; ++j) top[j] = bot[j];
            } else {
                pyr_hlerp(in + (y0 - in_row0) * Win, ox0, ox1, lx, hx, top);
            }
            if (y1 == y0) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bot[j] = top[j];
            } else {
                pyr_hlerp(in + (y1 - in_row0) * Win, ox0, ox1, lx, hx, bot);
            }
            have = y1;
            const float y0f = (float)y0;
            int yn;
#pragma unroll 1
            do {
                const float ly = fminf(fmaxf(fy - y0f, 0.f), 1.f), hy = 1.f - ly;
                float o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    o[j] = __fmaf_rn(hy, top[j], __fmul_rn(ly, bot[j]));       // ATen's operation order
                const bool mine = (unsigned)(y - ow.lo) < own_rows;
                if (vec) {
                    const float4 v = make_float4(o[0], o[1], o[2], o[3]);
                    if (mine) *reinterpret_cast<float4*>(out + off) = v;
                    if (keep) *reinterpret_cast<float4*>(keep_rel + off) = v;
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (xq + j >= Wout) break;
                        if (mine) out[off + j] = o[j];
                        if (keep) keep_rel[off + j] = o[j];
                    }
                }
                ++y;
                off += Wout;
                fy = sy * (float)y;
                yn = (int)fy;
                yn = yn < Hin - 1 ? yn : Hin - 1;
            } while (y < yend && yn == y0);
            y0 = yn;
        }
    }
}

template <bool kSide>
__global__ void __launch_bounds__(kPyrThreads)
pyramid_chain_kernel(const __grid_constant__ PyramidArgs a, const __grid_constant__ ScaleTable tab,
                     int32_t* __restrict__ n_bad, int side_ctas, int bad_per_tile) {
    extern __shared__ __align__(16) float levels_smem[];
    pdl_launch_dependents();
    const int cta = blockIdx.x;
    if (kSide) {                                   // after the pyramid CTAs (the critical path) in the grid
        const int n_pyr = a.D * a.slabs;
        if (cta >= n_pyr) {
            outborder_count_warp_tile(tab, n_bad, bad_per_tile != 0,      // a tile per warp
                                      (cta - n_pyr) * (int)(blockDim.x >> 5) + (int)(threadIdx.x >> 5));
            return;
        }
    }
    if (cta == 0) {                                // utils/loss.py:182-206 (see match_timestamps_kernel)
        for (int i = threadIdx.x; i < 2 * a.P; i += blockDim.x) {
            const int k = i / a.P, p = i - k * a.P;
            const float want = a.flow_ts[2 * p + k];
            const int64_t smp = a.flow_sample_idx[p];
            int hits = 0, idx = 0;
#pragma unroll 2
            for (int q = 0; q < a.D; ++q)
                if (a.timestamps[q] == want && a.sample_idx[q] == smp) { ++hits; idx = q; }
            if (hits != 1) atomicOr(a.status, DVSOF_ST_NO_MATCH);
            a.match_idx[i] = idx;
        }
        for (int k = threadIdx.x; k < a.n_tickets; k += blockDim.x) a.tickets[k] = 0u;
    }
    const int d = cta / a.slabs, g = cta - d * a.slabs;
    __shared__ RowRange own[DVSOF_MAX_SCALES], need[DVSOF_MAX_SCALES];
    if (threadIdx.x == 0) pyr_plan_rows(a, g, own, need);      // once per CTA, not once per warp
    __syncthreads();
#pragma unroll 1
    for (int l = 0; l < a.n_levels; ++l) {
        const int Hout = a.H[l], Wout = a.W[l];
        float* keep = (l + 1 < a.n_levels) ? levels_smem + ((l & 1) ? a.buf_b_offset : 0) : nullptr;
        float* out = a.dst[l] + (size_t)d * Hout * Wout;
        if (l == 0)                                // from the frame in global memory
            pyr_level_rows(a.src + (size_t)d * a.H0 * a.W0, a.H0, a.W0, 0, out, keep, Wout, a.sy[0], a.sx[0],
                           need[0], own[0]);
        else                                       // from the previous level's rows in shared memory
            pyr_level_rows(levels_smem + ((l & 1) ? 0 : a.buf_b_offset), a.H[l - 1], a.W[l - 1], need[l - 1].lo,
                           out, keep, Wout, a.sy[l], a.sx[l], need[l], own[l]);
        __syncthreads();
    }
}

}  // namespace dvsof

// ---------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------
using namespace dvsof;

// Rows per sweep segment (one warp walks 32 columns x R rows): the taller the segment, the better
// its fixed costs (decode, pointer set-up, halo pairs, first TMA round trip, record) are amortised;
// the shorter, the more warps there are to fill 148 SMs x kSweepWarpsPerSM resident warps and the
// finer the tail.  R = the largest of 32 / 16 / 8 that leaves four waves of warps -- measured
// (gpurun_out/r02o): cfg2 22.3 us at R = 8 / 24.7 at 16; cfg3 41.2 / 42.9; cfg4 77.2 / 73.9 / 91.6 (R = 32).
// DVSOF_LOSS_SEG_ROWS overrides (tuning).
constexpr int kSweepWarpsPerSM = DVSOF_LOSS_MIN_CTAS;
static void plan_segments(const dvsof_loss_scales_t* sc, ScaleTable* tab) {
    int64_t warp_rows = 0;
    for (int s = 0; s < sc->n_scales; ++s) warp_rows += (int64_t)sc->n_pred * tab->tiles_x[s] * sc->H[s];
    int R = 8;
    for (int r = 32; r > 8; r >>= 1)
        if (warp_rows / r >= 4LL * kNumSMs * kSweepWarpsPerSM) { R = r; break; }
    if (const char* e = getenv("DVSOF_LOSS_SEG_ROWS")) {
        const int v = atoi(e);
        if (v >= 8 && v <= 4096 && v % 8 == 0) R = v;
    }
    int64_t begin = 0;
    for (int s = 0; s < sc->n_scales; ++s) {
        tab->seg_rows[s] = R;
        tab->seg_per[s] = tab->tiles_x[s] * ((sc->H[s] + R - 1) / R);
        tab->inv_seg_per[s] = 1.f / (float)tab->seg_per[s];
        tab->seg_begin[s] = (int)begin;
        begin += (int64_t)tab->seg_per[s] * sc->n_pred;
        // a tensor map needs 16-byte aligned bases and row pitches
        // bit 0: the flow box, bit 1: the prev-level box
        tab->use_tma[s] = sc->W[s] % 4 == 0 ? (aligned16(sc->flow[s]) ? 1 : 0) |
                                                  (aligned16(sc->images[s]) && sc->n_images > 0 ? 2 : 0)
                                            : 0;
    }
    for (int s = sc->n_scales; s <= DVSOF_MAX_SCALES; ++s) tab->seg_begin[s] = (int)begin;
    for (int s = sc->n_scales; s < DVSOF_MAX_SCALES; ++s) tab->use_tma[s] = 0;
}

// ---- tensor maps (cuTensorMapEncodeTiled through the runtime's driver entry point: no libcuda link) ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
// fp32 tensor (W, H, N) row-major with a (bw, bh, bn) box, zero fill outside
static bool encode_map3(CUtensorMap* m, const float* base, int W, int H, int64_t N, int bw, int bh, int bn) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    const cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
    const cuuint32_t box[3] = {(cuuint32_t)bw, (cuuint32_t)bh, (cuuint32_t)bn};
    const cuuint32_t elem[3] = {1, 1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, elem,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static void build_maps(const dvsof_loss_scales_t* sc, ScaleTable* tab, LossMaps* maps) {
    for (int s = 0; s < sc->n_scales; ++s) {
        // no driver entry point / unencodable: plain loads for that part
        if ((tab->use_tma[s] & 1) &&
            !encode_map3(&maps->flow[s], sc->flow[s], sc->W[s], sc->H[s], 2LL * sc->n_pred, kBoxW, kBoxH, 2))
            tab->use_tma[s] &= ~1;
        if ((tab->use_tma[s] & 2) &&
            !encode_map3(&maps->prev[s], sc->images[s], sc->W[s], sc->H[s], sc->n_images, 32, kStageRows, 1))
            tab->use_tma[s] &= ~2;
    }
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
        tab->inv_tx[s] = 1.f / (float)tab->tiles_x[s];
        tab->inv_per[s] = 1.f / ((float)tab->tiles_x[s] * (float)tab->tiles_y[s]);
        const float hx = (float)(sc->W[s] - 1) / 2.f, hy = (float)(sc->H[s] - 1) / 2.f;
        // a 1-pixel side makes the reference divide by zero (NaN/inf grid); here: NaN grid as well
        tab->hx[s] = hx;
        tab->hy[s] = hy;
        tab->rx[s] = dvsof_div_const_ok(hx) ? dvsof_recip_rn(hx) : NAN;
        tab->ry[s] = dvsof_div_const_ok(hy) ? dvsof_recip_rn(hy) : NAN;
        const double a2 = 2.0 * (double)kAlpha, N2 = 2.0 * sc->n_pred, Hd = sc->H[s], Wd = sc->W[s];
        const double cnt[3] = {N2 * Hd * (Wd - 1), N2 * (Hd - 1) * Wd, N2 * (Hd - 1) * (Wd - 1)};
        for (int k = 0; k < 3; ++k) {
            tab->norm[k][s] = cnt[k] > 0 ? (float)(a2 * 0.25 / cnt[k]) : 0.f;
            tab->lk[k][s] = cnt[k] > 0 ? (float)log2(a2 * 0.25 / cnt[k]) : -INFINITY;
        }
        tab->norm[3][s] = (float)(a2 / ((double)sc->n_pred * Hd * Wd));
        tab->norm[4][s] = (float)(a2 / N2);
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
    plan_segments(sc, tab);
    return DVSOF_OK;
}

// workspace layout: tile records | pair records | tickets | per-tile out-of-border counts
static size_t partial_bytes(const ScaleTable& tab) {
    // one record per segment; sized for the shortest segments (8 rows = 4 per 32x32 tile) so that a
    // workspace stays valid whatever DVSOF_LOSS_SEG_ROWS says at launch time
    return round_up((size_t)tab.cta_begin[DVSOF_MAX_SCALES] * 4 * kRec * sizeof(float), 256);
}
static size_t pair_bytes(const ScaleTable& tab) {
    return round_up((size_t)tab.n_scales * tab.n_pred * kPairRec * sizeof(double), 256);
}
static int n_tickets(const ScaleTable&) { return 1; }      // the reduce kernel's "pairs finished" counter
static size_t ticket_bytes(const ScaleTable& tab) { return round_up((size_t)n_tickets(tab) * 4, 256); }
static LossWorkspace carve(const ScaleTable& tab, void* workspace) {
    char* p = static_cast<char*>(workspace);
    LossWorkspace ws;
    ws.partials = reinterpret_cast<float*>(p);
    ws.pair_rec = reinterpret_cast<double*>(p + partial_bytes(tab));
    ws.tickets = reinterpret_cast<unsigned*>(p + partial_bytes(tab) + pair_bytes(tab));
    ws.tile_bad = reinterpret_cast<int32_t*>(p + partial_bytes(tab) + pair_bytes(tab) + ticket_bytes(tab));
    return ws;
}

extern "C" size_t dvsof_loss_workspace_bytes(const dvsof_loss_scales_t* sc) {
    ScaleTable tab;
    if (build_table(sc, &tab, false)) return 0;
    return partial_bytes(tab) + pair_bytes(tab) + ticket_bytes(tab) +
           round_up((size_t)tab.cta_begin[DVSOF_MAX_SCALES] * sizeof(int32_t), 256);
}

static int zero_tickets(const ScaleTable& tab, const LossWorkspace& ws, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(ws.tickets, 0, (size_t)n_tickets(tab) * 4, stream);
    if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "loss ticket memset: %s", cudaGetErrorString(e));
    return DVSOF_OK;
}

template <bool kVal, bool kGrad>
static int launch_sweep(const char* what, const dvsof_loss_scales_t* sc, ScaleTable& tab, const int32_t* match_idx,
                        const float* grad_terms, const int32_t* n_bad_in, const LossWorkspace& ws,
                        float* terms, int32_t* n_bad_out, float* loss, cudaStream_t stream, bool pdl,
                        int bad_per_tile = 0) {
    LossMaps maps;
    if (getenv("DVSOF_LOSS_NO_TMA"))                 // tests: force the plain-load staging
        for (int s = 0; s < DVSOF_MAX_SCALES; ++s) tab.use_tma[s] = 0;
    if (const char* e = getenv("DVSOF_LOSS_TMA_MASK"))   // debugging: 1 = flow boxes only, 2 = prev boxes only
        for (int s = 0; s < DVSOF_MAX_SCALES; ++s) tab.use_tma[s] &= atoi(e);
    if (const char* e = getenv("DVSOF_LOSS_TMA_MINW"))   // debugging: tensor maps only for scales at least this wide
        for (int s = 0; s < sc->n_scales; ++s)
            if (sc->W[s] < atoi(e) || sc->H[s] < atoi(e)) tab.use_tma[s] = 0;
    build_maps(sc, &tab, &maps);
    if (int rc = launch_kernel(what, warp_loss_ring_kernel<kVal, kGrad>, tab.seg_begin[DVSOF_MAX_SCALES], 32, 0,
                               stream, pdl, tab, maps, match_idx, grad_terms, n_bad_in, bad_per_tile, ws))
        return rc;
    if (!kVal) return DVSOF_OK;
    const int pairs_per_cta = kReduceThreads / 32;
    return launch_kernel("loss_reduce_kernel", loss_reduce_kernel,
                         (tab.n_scales * tab.n_pred + pairs_per_cta - 1) / pairs_per_cta, kReduceThreads, 0, stream,
                         true, tab, ws, terms, n_bad_out, loss);
}

extern "C" int dvsof_warp_loss_forward(const dvsof_loss_scales_t* sc, const int32_t* match_idx,
                                       float* terms, int32_t* n_bad, void* workspace,
                                       void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, false)) return rc;
    if (!match_idx || !terms || !n_bad || !workspace)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "warp_loss_forward: null buffer");
    cudaStream_t stream = (cudaStream_t)stream_;
    const LossWorkspace ws = carve(tab, workspace);
    if (int rc = zero_tickets(tab, ws, stream)) return rc;
    return launch_sweep<true, false>("warp_loss_ring_kernel<fwd>", sc, tab, match_idx, nullptr, nullptr, ws,
                                     terms, n_bad, nullptr, stream, false);
}

extern "C" int dvsof_warp_loss_backward(const dvsof_loss_scales_t* sc, const int32_t* match_idx,
                                        const float* grad_terms, const int32_t* n_bad,
                                        void* stream_) {
    ScaleTable tab;
    if (int rc = build_table(sc, &tab, true)) return rc;
    if (!match_idx || !grad_terms || !n_bad)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "warp_loss_backward: null buffer");
    return launch_sweep<false, true>("warp_loss_ring_kernel<bwd>", sc, tab, match_idx, grad_terms, n_bad,
                                     LossWorkspace{nullptr, nullptr, nullptr}, nullptr, nullptr, nullptr,
                                     (cudaStream_t)stream_, false);
}

static int launch_match(const float* flow_ts, const int64_t* flow_sample_idx, int64_t P,
                        const float* timestamps, const int64_t* sample_idx, int64_t D,
                        int32_t* match_idx, uint32_t* status, unsigned* tickets, int n_tick,
                        cudaStream_t stream) {
    if (P < 1 || D < 1 || P > (1 << 24) || D > (1 << 24) || !flow_ts || !flow_sample_idx ||
        !timestamps || !sample_idx || !match_idx || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "match_timestamps: bad arguments");
    const int threads = 128, blocks = (int)((2 * P + threads - 1) / threads);
    return launch_kernel("match_timestamps_kernel", match_timestamps_kernel, blocks, threads, 0, stream,
                         false, flow_ts, flow_sample_idx, (int)P, timestamps, sample_idx, (int)D,
                         match_idx, status, tickets, n_tick);
}

extern "C" int dvsof_match_timestamps(const float* flow_ts, const int64_t* flow_sample_idx,
                                      int64_t P, const float* timestamps,
                                      const int64_t* sample_idx, int64_t D, int32_t* match_idx,
                                      uint32_t* status, void* stream_) {
    return launch_match(flow_ts, flow_sample_idx, P, timestamps, sample_idx, D, match_idx, status,
                        nullptr, 0, (cudaStream_t)stream_);
}

// side != nullptr: the launch also runs the out-of-border count of `side` (tiles first) into n_bad
static int launch_resize(const float* src, int64_t D, int32_t Hin, int32_t Win, float* dst,
                         int32_t Hout, int32_t Wout, cudaStream_t stream, bool pdl,
                         const ScaleTable* side = nullptr, int32_t* n_bad = nullptr, int bad_per_tile = 0) {
    if (D < 1 || Hin < 1 || Win < 1 || Hout < 1 || Wout < 1 || !src || !dst)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "resize_bilinear: bad arguments");
    const int64_t rows = D * Hout;
    if (rows > INT32_MAX / 2 || (int64_t)Hin * Win > INT32_MAX / 2)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "resize_bilinear: too large");
    const int cta_rows = kRWarps * kRY;
    const int64_t row_ctas = (rows + cta_rows - 1) / cta_rows, col_ctas = (Wout + 32 * kRX - 1) / (32 * kRX);
    const int side_ctas = side ? side->cta_begin[DVSOF_MAX_SCALES] : 0;
    if (row_ctas * col_ctas + side_ctas > INT32_MAX)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "resize_bilinear: too large");
    static const ScaleTable none = {};
    const int grid = (int)(row_ctas * col_ctas) + side_ctas;
    if (side)
        return launch_kernel("resize_bilinear_kernel<side>", resize_bilinear_kernel<true>, grid, 32 * kRWarps,
                             0, stream, pdl, src, (int)Hin, (int)Win, dst, (int)Hout, (int)Wout, (int)rows,
                             (int)col_ctas, *side, n_bad, side_ctas, bad_per_tile);
    return launch_kernel("resize_bilinear_kernel", resize_bilinear_kernel<false>, grid, 32 * kRWarps, 0,
                         stream, pdl, src, (int)Hin, (int)Win, dst, (int)Hout, (int)Wout, (int)rows,
                         (int)col_ctas, none, (int32_t*)nullptr, 0, 0);
}

extern "C" int dvsof_resize_bilinear(const float* src, int64_t D, int32_t Hin, int32_t Win,
                                     float* dst, int32_t Hout, int32_t Wout, void* stream_) {
    return launch_resize(src, D, Hin, Win, dst, Hout, Wout, (cudaStream_t)stream_, false);
}

// Plan of one pyramid_chain_kernel launch: the number of leading levels it produces (level l+1 is
// computed from the rows of level l held in shared memory, so the kept slabs of levels 0..k-2 must
// fit), the CTAs per frame, and the shared-memory layout.
constexpr size_t kPyrSmemMax = 200 * 1024;
static bool plan_chain(const dvsof_loss_scales_t* sc, const dvsof_loss_frames_t* fr, PyramidArgs* a,
                       size_t* smem_bytes, int* threads) {
    const int last_rows = sc->H[sc->n_scales - 1];
    int slabs = kNumSMs / (fr->n_images > 0 ? fr->n_images : 1);
    slabs = slabs > 16 ? 16 : slabs;
    if (const char* e = getenv("DVSOF_PYR_SLABS")) {            // tuning: slabs per frame
        const int v = atoi(e);
        if (v >= 1 && v <= 64) slabs = v;
    }
    slabs = slabs > last_rows / 4 ? last_rows / 4 : slabs;
    a->slabs = slabs < 1 ? 1 : slabs;
    // warps per CTA: about four 2-row x 128-column tasks of the last level each (every warp pays the
    // per-level set-up, so idle warps are not free)
    const int64_t tasks = (int64_t)((last_rows / a->slabs + 1) / 2 + 1) * ((sc->W[sc->n_scales - 1] + 127) / 128);
    int warps = (int)((tasks + 3) / 4 < 4 ? 4 : ((tasks + 3) / 4 > kPyrThreads / 32 ? kPyrThreads / 32 : (tasks + 3) / 4));
    *threads = warps * 32;
    for (int l = 0; l < sc->n_scales; ++l) {
        if (sc->H[l] >= (1 << 26)) return false;             // 32-bit slab arithmetic in the kernel
        a->H[l] = sc->H[l]; a->W[l] = sc->W[l];
        a->sy[l] = pyr_step(l ? sc->H[l - 1] : fr->H0, sc->H[l]);
        a->sx[l] = pyr_step(l ? sc->W[l - 1] : fr->W0, sc->W[l]);
    }
    for (int k = sc->n_scales; k >= 1; --k) {
        a->n_levels = k;
        size_t buf[2] = {0, 0};
        for (int g = 0; g < a->slabs; ++g) {
            RowRange own[DVSOF_MAX_SCALES], need[DVSOF_MAX_SCALES];
            pyr_plan_rows(*a, g, own, need);
            for (int l = 0; l + 1 < k; ++l) {
                const int rows = need[l].hi > need[l].lo ? need[l].hi - need[l].lo : 0;
                const size_t n = (size_t)rows * sc->W[l];
                if (n > buf[l & 1]) buf[l & 1] = n;
            }
        }
        if ((buf[0] + 4 + buf[1]) * sizeof(float) <= kPyrSmemMax) {
            buf[0] = (buf[0] + 3) & ~(size_t)3;                 // 16-byte aligned second buffer
            a->buf_b_offset = (int)buf[0];
            *smem_bytes = (buf[0] + buf[1]) * sizeof(float);
            return true;
        }
    }
    return false;
}

// matching + chained pyramid (utils/loss.py:182-210), PDL-chained on the stream.  side != nullptr:
// the last launch of the pyramid also counts the out-of-border pixels into n_bad (per tile).
static int launch_match_and_pyramid(const dvsof_loss_scales_t* sc, const dvsof_loss_frames_t* fr,
                                    int32_t* match_idx, uint32_t* status, unsigned* tickets,
                                    int n_tick, cudaStream_t stream,
                                    const ScaleTable* side = nullptr, int32_t* n_bad = nullptr) {
    size_t smem = 0;
    int threads = kPyrThreads;
    PyramidArgs a = {};
    const bool chain = sc->n_pred <= (1 << 20) && (int64_t)fr->H0 * fr->W0 < (1 << 30) &&
                       fr->n_images <= (1 << 20) && plan_chain(sc, fr, &a, &smem, &threads);
    const int k = a.n_levels;
    int first = 0;
    const float* src = fr->images;
    int Hin = fr->H0, Win = fr->W0;
    if (chain) {
        a.src = fr->images; a.D = fr->n_images; a.H0 = fr->H0; a.W0 = fr->W0;
        for (int l = 0; l < k; ++l) a.dst[l] = const_cast<float*>(sc->images[l]);
        a.flow_ts = fr->flow_ts; a.flow_sample_idx = fr->flow_sample_idx; a.P = sc->n_pred;
        a.timestamps = fr->timestamps; a.sample_idx = fr->sample_idx;
        a.match_idx = match_idx; a.status = status; a.tickets = tickets; a.n_tickets = n_tick;
        if (!a.flow_ts || !a.flow_sample_idx || !a.timestamps || !a.sample_idx)
            return fail(DVSOF_ERR_INVALID_ARGUMENT, "multiscale loss: null matching input");
        static unsigned long long attr_done = 0;             // per device; a race only repeats the call
        int dev = 0;
        if (needs_device_setup(attr_done, &dev)) {
            cudaFuncSetAttribute(pyramid_chain_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPyrSmemMax);
            cudaFuncSetAttribute(pyramid_chain_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPyrSmemMax);
            device_setup_done(attr_done, dev);
        }
        const bool last = k == sc->n_scales;
        static const ScaleTable none = {};
        if (last && side) {
            const int side_ctas = (side->cta_begin[DVSOF_MAX_SCALES] + threads / 32 - 1) / (threads / 32);
            if (int rc = launch_kernel("pyramid_chain_kernel<side>", pyramid_chain_kernel<true>,
                                       fr->n_images * a.slabs + side_ctas, threads, smem, stream, false, a, *side,
                                       n_bad, side_ctas, 1))
                return rc;
        } else if (int rc = launch_kernel("pyramid_chain_kernel", pyramid_chain_kernel<false>, fr->n_images * a.slabs,
                                          threads, smem, stream, false, a, none, (int32_t*)nullptr, 0, 0)) {
            return rc;
        }
        first = k;
        if (k > 0) { src = sc->images[k - 1]; Hin = sc->H[k - 1]; Win = sc->W[k - 1]; }
    } else if (int rc = launch_match(fr->flow_ts, fr->flow_sample_idx, sc->n_pred, fr->timestamps,
                                     fr->sample_idx, fr->n_images, match_idx, status, tickets, n_tick, stream)) {
        return rc;
    }
    for (int s = first; s < sc->n_scales; ++s) {                           // loss.py:207-210
        float* dst = const_cast<float*>(sc->images[s]);
        const bool last = s + 1 == sc->n_scales;
        if (int rc = launch_resize(src, fr->n_images, Hin, Win, dst, sc->H[s], sc->W[s], stream, true,
                                   last ? side : nullptr, last ? n_bad : nullptr, 1))
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
    if (!sc || !frames_ok(fr) || !match_idx || !terms || !n_bad || !workspace || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "multiscale_loss_forward: bad arguments");
    dvsof_loss_scales_t with_frames = *sc;
    with_frames.n_images = fr->n_images;
    sc = &with_frames;
    if (int rc = build_table(sc, &tab, false)) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    const LossWorkspace ws = carve(tab, workspace);
    if (int rc = launch_match_and_pyramid(sc, fr, match_idx, status, ws.tickets, n_tickets(tab), stream))
        return rc;
    return launch_sweep<true, false>("warp_loss_ring_kernel<fwd>", sc, tab, match_idx, nullptr, nullptr, ws,
                                     terms, n_bad, nullptr, stream, true);
}

static void set_coef(ScaleTable* tab, const float* coef) {
    for (int k = 0; k < 3; ++k)
        for (int s = 0; s < tab->n_scales; ++s) tab->coef[k * DVSOF_MAX_SCALES + s] = coef[k * tab->n_scales + s];
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
    if (!sc || !frames_ok(fr) || !coef || !match_idx || !terms || !loss || !n_bad || !workspace || !status)
        return fail(DVSOF_ERR_INVALID_ARGUMENT, "multiscale_loss_fused: bad arguments");
    dvsof_loss_scales_t with_frames = *sc;
    with_frames.n_images = fr->n_images;
    sc = &with_frames;
    if (int rc = build_table(sc, &tab, true)) return rc;
    set_coef(&tab, coef);
    cudaStream_t stream = (cudaStream_t)stream_;
    const LossWorkspace ws = carve(tab, workspace);
    // the last pyramid launch leaves one out-of-border count per tile; the sweep sums its pair's
    // counts itself and reports the pair totals in n_bad (no memset, no atomics)
    if (int rc = launch_match_and_pyramid(sc, fr, match_idx, status, ws.tickets, n_tickets(tab), stream,
                                          &tab, ws.tile_bad))
        return rc;
    return launch_sweep<true, true>("warp_loss_ring_kernel<fused>", sc, tab, match_idx, nullptr, ws.tile_bad, ws,
                                    terms, n_bad, loss, stream, true, 1);
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
    set_coef(&tab, coef);
    cudaStream_t stream = (cudaStream_t)stream_;
    const LossWorkspace ws = carve(tab, workspace);
    if (int rc = zero_tickets(tab, ws, stream)) return rc;
    if (count_pass) {
        cudaError_t e = cudaMemsetAsync(n_bad, 0, sizeof(int32_t) * sc->n_scales * sc->n_pred, stream);
        if (e != cudaSuccess) return fail(DVSOF_ERR_CUDA, "n_bad memset: %s", cudaGetErrorString(e));
        if (int rc = launch_kernel("outborder_count_kernel", outborder_count_kernel,
                                   tab.cta_begin[DVSOF_MAX_SCALES], 32 * kCountWarps, 0, stream, false, tab, n_bad))
            return rc;
    }
    return launch_sweep<true, true>("warp_loss_ring_kernel<fused>", sc, tab, match_idx, nullptr, n_bad, ws,
                                    terms, nullptr, loss, stream, false);
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
                         (cudaStream_t)stream_, true, tab, factor);
}
