// loss_strip.cuh -- register-blocked warp-loss kernel (included by loss.cu).
//
// One launch covers all scales and predictions of utils/loss.py:121-171.  A CTA owns one 32x32
// pixel tile of one (scale, prediction); each of its 4 warps walks a STRIP of 32 columns x 8
// consecutive rows top to bottom, one pixel column per lane.  Every neighbour pair of the
// smoothness term (utils/loss.py:76-90) is owned by the top-left pixel ("cell") of its 2x2 block
// and its Charbonnier value AND derivative are evaluated exactly once, from one lg2 + one ex2:
//     s = d^2 + eps^2,  t = s^(alpha-1),  c(d) = t * s,  c'(d) = 2 alpha d t.
// The two flow channels travel as one packed fp32x2 value (FADD2 / FMUL2 / FFMA2, sm_100), so
// the FMA-pipe work of the pair phase is issued once for both.  The derivatives never leave
// registers: a pixel's gradient is
//     g(x,y) = L(x-1,y) - Gh(x,y) + Gv(x,y-1) - Gv(x,y) - Gda(x,y) - Gdb(x,y-1)
//     L(x,y) = Gh(x,y) + Gda(x,y-1) + Gdb(x,y)            (everything the right neighbour needs)
// so the row above is carried in registers and the left neighbour arrives with one shuffle per
// channel.  Strip borders are recomputed (one extra cell row per strip, one extra cell column per
// tile, spread over the lanes): +15 % pair evaluations, no shared-memory gradient planes, no
// inter-warp synchronisation besides the flow-tile staging.
//
// Latency: a strip is processed in groups of kGroup rows.  Per group the sampling coordinates
// are computed and ALL bilinear taps are requested first, the (load-free) pair phase of the group
// runs while they are in flight, and only then are the taps consumed.
//
// Modes (template):  <kVal, !kGrad> forward terms only;  <!kVal, kGrad> backward for arbitrary
// d(loss)/d(terms) read from device memory;  <kVal, kGrad> the fused training pass: terms and
// d(sum_k coef_k term_k)/d(flow) in a single sweep -- 16 B/pixel read, 8 B/pixel written.
#pragma once

namespace dvsof {

#ifndef DVSOF_LOSS_GROUP
#define DVSOF_LOSS_GROUP 2                // tuning knobs (scripts/exp_loss_variants.sh builds the sweep)
#endif
#ifndef DVSOF_LOSS_MIN_CTAS
#define DVSOF_LOSS_MIN_CTAS 7
#endif
#ifndef DVSOF_LOSS_STRIP_ROWS
#define DVSOF_LOSS_STRIP_ROWS 8
#endif
constexpr int kSR = DVSOF_LOSS_STRIP_ROWS;  // rows per warp strip
constexpr int kGroup = DVSOF_LOSS_GROUP;  // rows whose taps are in flight together
constexpr int kSWarps = kTile / kSR;      // 4 warps per CTA
constexpr int kSThreads = 32 * kSWarps;   // 128
constexpr int kFH = kTile + 2;            // staged flow rows  y0-1 .. y0+32
constexpr int kFP = kTile + 2;            // staged flow columns x0-1 .. x0+32 (float2 = (u, v))
static_assert(kSR % kGroup == 0, "a strip is a whole number of groups");

// ---- packed fp32x2 helpers (both flow channels in one FMA-pipe instruction) ----
__device__ __forceinline__ float2 f2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 f2add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 f2sub(float2 a, float2 b) { return __ffma2_rn(b, f2(-1.f), a); }
__device__ __forceinline__ float2 f2mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 f2fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }

struct Charb2 { float2 val, der; };       // c(d) and c'(d) / (2 alpha), both channels
template <bool kVal, bool kGrad>
__device__ __forceinline__ Charb2 charb2(float2 p, float2 q) {     // d = p - q
    const float2 d = f2sub(p, q);
    const float2 s = f2fma(d, d, f2(kEps2));
    const float2 l = make_float2(lg2_approx(s.x), lg2_approx(s.y));
    Charb2 r;
    if (kGrad) {
        const float2 e = f2mul(l, f2(kAlpha - 1.f));
        const float2 t = make_float2(ex2_approx(e.x), ex2_approx(e.y));
        r.der = f2mul(d, t);
        r.val = f2mul(t, s);
    } else {
        const float2 e = f2mul(l, f2(kAlpha));
        r.der = f2(0.f);
        r.val = make_float2(ex2_approx(e.x), ex2_approx(e.y));
    }
    return r;
}
struct Charb1 { float val, der; };
template <bool kVal, bool kGrad>
__device__ __forceinline__ Charb1 charb1(float d) {
    const float s = fmaf(d, d, kEps2);
    const float l = lg2_approx(s);
    Charb1 r;
    if (kGrad) {
        const float t = ex2_approx((kAlpha - 1.f) * l);
        r.der = d * t;
        r.val = t * s;
    } else {
        r.der = 0.f;
        r.val = ex2_approx(kAlpha * l);
    }
    return r;
}

// normalised sampling coordinate of utils/loss.py:150-156 (IEEE division by the half extent)
__device__ __forceinline__ float grid_coord(float pix, float disp, float h, float r) {
    return __fsub_rn(dvsof_div_by_const_grid(__fadd_rn(pix, disp), h, r), 1.f);
}
__device__ __forceinline__ bool out_of_border(float gx, float gy) {               // loss.py:92-94
    return gx < -1.f || gx > 1.f || gy < -1.f || gy > 1.f;
}
// The counting jobs only need the predicate, not the coordinates: pixel + flow against two
// thresholds per axis decides it without the division (exact_div.h: dvsof_grid_outside).
struct BorderTest { float hi_x, lo_x, hi_y, lo_y; };
__device__ __forceinline__ BorderTest border_test(const ScaleTable& tab, int s) {
    return BorderTest{dvsof_grid_outside_hi(tab.hx[s], tab.rx[s]), dvsof_grid_outside_lo(tab.hx[s], tab.rx[s]),
                      dvsof_grid_outside_hi(tab.hy[s], tab.ry[s]), dvsof_grid_outside_lo(tab.hy[s], tab.ry[s])};
}
__device__ __forceinline__ bool outside_grid(float ax, float ay, const BorderTest& bt) {
    return dvsof_grid_outside(ax, bt.hi_x, bt.lo_x) || dvsof_grid_outside(ay, bt.hi_y, bt.lo_y);
}

// (scale, prediction, tile) of this CTA without integer division: at most 2^23 tiles (checked on
// the host), so a float product with the host's reciprocal and one correction step is exact.
__device__ __forceinline__ TileCoord decode_cta_fast(const ScaleTable& tab, int cta) {
    int s = 0;
#pragma unroll 1
    for (int k = 1; k < tab.n_scales; ++k) s += cta >= tab.cta_begin[k] ? 1 : 0;
    const int local = cta - tab.cta_begin[s];
    const int tx = tab.tiles_x[s], per = tx * tab.tiles_y[s];
    int n = (int)(((float)local + 0.5f) * tab.inv_per[s]);
    n += (n + 1) * per <= local ? 1 : 0;
    n -= n * per > local ? 1 : 0;
    const int tile = local - n * per;
    int ty = (int)(((float)tile + 0.5f) * tab.inv_tx[s]);
    ty += (ty + 1) * tx <= tile ? 1 : 0;
    ty -= ty * tx > tile ? 1 : 0;
    return TileCoord{s, n, ty * kTile, (tile - ty * tx) * kTile};
}

// Stage the flow tile rows y0-1..y0+32, columns x0-1..x0+32 (zero outside the image) into
// F[kFH][kFP] as (u, v) pairs; tile column c <-> x = x0 - 1 + c.  Lane l loads the aligned main
// columns c = 1..32 (one 128-byte line per warp, row and channel), the two edge columns go to
// the first 68 threads.  All global loads are issued before the first shared store; offsets are
// 32-bit (one (prediction, channel-pair) block holds < 2^31 elements, checked on the host).
__device__ __forceinline__ void stage_flow_strip_tile(float2 (*F)[kFP], const float* __restrict__ flow,
                                                      int plane, int H, int W, int y0, int x0) {
    const float* flow1 = flow + plane;
    asm volatile("" : "+l"(flow), "+l"(flow1));           // finished channel bases: address = base + 32-bit offset
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kPer = (kFH + kSWarps - 1) / kSWarps;   // 9 row slots per warp (34 rows)
    const int x = x0 + lane;
    const bool xok = x < W;
    float u[kPer], v[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int ty = warp + k * kSWarps;
        const int y = y0 - 1 + ty;
        const bool ok = xok && ty < kFH && y >= 0 && y < H;
        const int off = y * W + x;
        u[k] = ok ? __ldg(flow + off) : 0.f;
        v[k] = ok ? __ldg(flow1 + off) : 0.f;
    }
    float eu = 0.f, ev = 0.f;
    const int ety = threadIdx.x >> 1, eright = threadIdx.x & 1;          // threads 0..67: edge columns
    {
        const int y = y0 - 1 + ety, xe = eright ? x0 + kTile : x0 - 1;
        const bool ok = ety < kFH && y >= 0 && y < H && xe >= 0 && xe < W;
        const int off = y * W + xe;
        eu = ok ? __ldg(flow + off) : 0.f;
        ev = ok ? __ldg(flow1 + off) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int ty = warp + k * kSWarps;
        if (ty < kFH) F[ty][lane + 1] = make_float2(u[k], v[k]);
    }
    if (ety < kFH) F[ety][eright ? kTile + 1 : 0] = make_float2(eu, ev);
}

// One row's bilinear taps in flight: requested in phase 1, consumed in phase 3.
struct Taps { float nw, ne, sw, se, pv, wx, wy; bool bad; };

// phase 1 of a row: sampling position exactly as the reference computes it (see warp_sample in
// loss.cu for the operation-order notes), then the four zero-padded taps of `next` and the pixel
// of `prev` are requested.  32-bit offsets inside one image plane.
__device__ __forceinline__ Taps request_taps(const float* __restrict__ next, const float* __restrict__ prev,
                                             int H, int W, const GridNorm& gn, float xf, float yf,
                                             int off_px, float u, float v, bool valid) {
    Taps t;
    const float gx = grid_coord(xf, u, gn.hx, gn.rx), gy = grid_coord(yf, v, gn.hy, gn.ry);
    t.bad = valid && out_of_border(gx, gy);
    const float ix = __fmul_rn(__fadd_rn(gx, 1.f), gn.hx);            // ATen CPU unnormalize
    const float iy = __fmul_rn(__fadd_rn(gy, 1.f), gn.hy);
    const float xw = floorf(ix), yn = floorf(iy);
    t.wx = ix - xw;                                                  // weight of east / south
    t.wy = iy - yn;
    const int xi = (int)fminf(fmaxf(xw, -2.f), (float)W);
    const int yi = (int)fminf(fmaxf(yn, -2.f), (float)H);
    const bool x0ok = (unsigned)xi < (unsigned)W, x1ok = (unsigned)(xi + 1) < (unsigned)W;
    const bool y0ok = valid && (unsigned)yi < (unsigned)H, y1ok = valid && (unsigned)(yi + 1) < (unsigned)H;
    const float* p0 = next + (yi * W + xi);
    const float* p1 = p0 + W;
    t.nw = (x0ok && y0ok) ? __ldg(p0) : 0.f;
    t.ne = (x1ok && y0ok) ? __ldg(p0 + 1) : 0.f;
    t.sw = (x0ok && y1ok) ? __ldg(p1) : 0.f;
    t.se = (x1ok && y1ok) ? __ldg(p1 + 1) : 0.f;
    t.pv = valid ? __ldg(prev + off_px) : 0.f;
    return t;
}

template <bool kVal, bool kGrad>
__global__ void __launch_bounds__(kSThreads, DVSOF_LOSS_MIN_CTAS)
warp_loss_strip_kernel(const __grid_constant__ ScaleTable tab, const int32_t* __restrict__ match_idx,
                       const float* __restrict__ grad_terms, const int32_t* __restrict__ n_bad_in,
                       int bad_per_tile, const LossWorkspace ws, float* terms, int32_t* __restrict__ n_bad_out,
                       float* __restrict__ loss_out) {
    __shared__ __align__(16) float2 F[kFH][kFP];
    __shared__ float red[kSWarps][kRec];
    __shared__ double dred[kVal ? kSWarps : 1][7];
    __shared__ bool is_last;
    pdl_launch_dependents();
    const TileCoord tc = decode_cta_fast(tab, blockIdx.x);
    const int S = tab.n_scales, P = tab.n_pred;
    const int H = tab.H[tc.s], W = tab.W[tc.s];
    const int plane = H * W;
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    // the flow is an input of the call: its tile is staged while the prelude (pyramid levels,
    // match_idx, tile counts, tickets) may still be running
    stage_flow_strip_tile(F, flow, plane, H, W, tc.y0, tc.x0);
    pdl_wait_for(match_idx, n_bad_in);             // both come from the prelude (see pdl_acquire)

    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);   // provably warp-uniform
    const int x = tc.x0 + lane, lx = lane + 1;
    const int ys = tc.y0 + warp * kSR;            // first pixel row of this warp's strip
    const int ty0 = 1 + warp * kSR;               // its tile row
    const bool x_in = x < W, x_r = x + 1 < W;

    // d(loss)/d(term) of this scale: device vector (generic backward) or host coefficients (fused)
    float kh = 0.f, kv = 0.f, kd = 0.f, k_ph = 0.f, k_ob = 0.f;
    if (kGrad) {
        const float g_sm = grad_terms ? grad_terms[tc.s] : tab.coef[tc.s];
        const float g_ph = grad_terms ? grad_terms[S + tc.s] : tab.coef[DVSOF_MAX_SCALES + tc.s];
        const float g_ob = grad_terms ? grad_terms[2 * S + tc.s] : tab.coef[2 * DVSOF_MAX_SCALES + tc.s];
        // 2 alpha / (element count of the mean), per direction, precomputed on the host
        kh = g_sm * tab.norm[0][tc.s];
        kv = g_sm * tab.norm[1][tc.s];
        kd = g_sm * tab.norm[2][tc.s];
        k_ph = g_ph * tab.norm[3][tc.s];
        int nb;
        if (bad_per_tile) {                       // per-tile counts of this (scale, prediction): sum them
            const int per = tab.tiles_x[tc.s] * tab.tiles_y[tc.s];
            const int32_t* tb = n_bad_in + tab.cta_begin[tc.s] + tc.n * per;
            nb = 0;
            for (int t = lane; t < per; t += 32) nb += __ldcg(tb + t);
            nb = __reduce_add_sync(0xffffffffu, nb);
        } else {
            nb = n_bad_in[tc.s * P + tc.n];
        }
        k_ob = nb > 0 ? __fdividef(g_ob * tab.norm[4][tc.s], (float)nb) : 0.f;   // loss.py:99-113
    }
    const float khx = x_r ? kh : 0.f, kvx = x_in ? kv : 0.f, kdx = x_r ? kd : 0.f;
    const float mhx = x_r ? 1.f : 0.f, mvx = x_in ? 1.f : 0.f;
    const float* prev = tab.images[tc.s] + (size_t)match_idx[tc.n] * plane;
    const float* next = tab.images[tc.s] + (size_t)match_idx[P + tc.n] * plane;
    const GridNorm gn{tab.hx[tc.s], tab.hy[tc.s], tab.rx[tc.s], tab.ry[tc.s]};
    float* grad = kGrad ? tab.grad_flow[tc.s] + (size_t)tc.n * 2 * plane : nullptr;
    // finished base pointers in registers: every access below is base + 32-bit offset (IMAD.WIDE)
    asm volatile("" : "+l"(prev), "+l"(next), "+l"(grad));
    const float xf = (float)x;
    __syncthreads();

    // packed accumulators (x + y summed at the end): smooth_h, smooth_v, smooth_diag; scalars:
    // photometric, outborder sum, outborder count
    float2 acc_h = f2(0.f), acc_v = f2(0.f), acc_d = f2(0.f);
    float acc_p = 0.f, acc_o = 0.f, acc_n = 0.f;
    if (ys < H) {                                         // warp-uniform
        float2 lh = f2(0.f);                               // lane j: L of the cell left of the tile, row ys+j
        float2 gv_p = f2(0.f), gda_p = f2(0.f), gdb_p = f2(0.f);   // pair gradients of the row above
        if (kGrad && tc.x0 > 0 && lane < kSR) {
            const int ty = ty0 + lane, y = ys + lane;
            const float k_h = y < H ? kh : 0.f;
            const float k_a = (y >= 1 && y < H) ? kd : 0.f;
            const float k_b = y + 1 < H ? kd : 0.f;
            const float2 p = F[ty][0], q = F[ty][1], up = F[ty - 1][0], dn = F[ty + 1][0];
            lh = f2mul(f2(k_h), charb2<false, true>(q, p).der);
            lh = f2fma(f2(k_a), charb2<false, true>(q, up).der, lh);
            lh = f2fma(f2(k_b), charb2<false, true>(q, dn).der, lh);
        }
        float2 a, b;                                       // flow at (x, y), (x+1, y) of the current row
        if (kGrad) {                                       // cell row ys-1: the pairs reaching into row ys
            const int ty = ty0 - 1;
            a = F[ty][lx]; b = F[ty][lx + 1];
            const float2 c = F[ty + 1][lx], d = F[ty + 1][lx + 1];
            if (ys > 0) {                                  // ys < H holds: every such pair exists in y
                gv_p = f2mul(f2(kvx), charb2<false, true>(c, a).der);
                gda_p = f2mul(f2(kdx), charb2<false, true>(d, a).der);
                gdb_p = f2mul(f2(kdx), charb2<false, true>(b, c).der);
            }
            a = c; b = d;
        } else {
            a = F[ty0][lx]; b = F[ty0][lx + 1];
        }
#pragma unroll
        for (int g0 = 0; g0 < kSR; g0 += kGroup) {
            // ---- phase 1: sampling positions of the group's rows, all taps requested ----
            Taps tp[kGroup];
            float2 uv[kGroup];
#pragma unroll
            for (int k = 0; k < kGroup; ++k) {
                const int y = ys + g0 + k;
                uv[k] = F[ty0 + g0 + k][lx];
                tp[k] = request_taps(next, prev, H, W, gn, xf, (float)y, y * W + x, uv[k].x, uv[k].y,
                                     x_in && y < H);
            }
            // ---- phase 2: neighbour pairs of the group's rows (no global loads) ----
            float2 gs[kGroup];
#pragma unroll
            for (int k = 0; k < kGroup; ++k) {
                const int r = g0 + k, ty = ty0 + r, y = ys + r;
                const bool row_in = y < H, row_d = y + 1 < H;                  // warp-uniform
                const float2 c = F[ty + 1][lx], d = F[ty + 1][lx + 1];
                const Charb2 ph = charb2<kVal, kGrad>(b, a);                   // loss.py:76-90
                const Charb2 pv = charb2<kVal, kGrad>(c, a);
                const Charb2 pa = charb2<kVal, kGrad>(d, a);
                const Charb2 pb = charb2<kVal, kGrad>(b, c);
                if (kVal) {
                    const float mh_r = row_in ? mhx : 0.f, mv_r = row_d ? mvx : 0.f, md_r = row_d ? mhx : 0.f;
                    acc_h = f2fma(f2(mh_r), ph.val, acc_h);
                    acc_v = f2fma(f2(mv_r), pv.val, acc_v);
                    acc_d = f2fma(f2(md_r), f2add(pa.val, pb.val), acc_d);
                }
                gs[k] = f2(0.f);
                if (kGrad) {
                    const float kh_r = row_in ? khx : 0.f, kv_r = row_d ? kvx : 0.f, kd_r = row_d ? kdx : 0.f;
                    const float2 gh = f2mul(f2(kh_r), ph.der), gv = f2mul(f2(kv_r), pv.der);
                    const float2 gda = f2mul(f2(kd_r), pa.der), gdb = f2mul(f2(kd_r), pb.der);
                    const float2 L = f2add(f2add(gh, gda_p), gdb);
                    float2 Ll = make_float2(__shfl_up_sync(0xffffffffu, L.x, 1), __shfl_up_sync(0xffffffffu, L.y, 1));
                    const float2 Lh = make_float2(__shfl_sync(0xffffffffu, lh.x, r), __shfl_sync(0xffffffffu, lh.y, r));
                    if (lane == 0) Ll = Lh;
                    gs[k] = f2sub(f2add(f2sub(Ll, gh), f2sub(gv_p, gv)), f2add(gda, gdb_p));
                    gv_p = gv; gda_p = gda; gdb_p = gdb;
                }
                a = c; b = d;
            }
            // ---- phase 3: consume the taps: photometric term, border term, gradient ----
#pragma unroll
            for (int k = 0; k < kGroup; ++k) {
                const int y = ys + g0 + k;
                const bool valid = x_in && y < H;
                const Taps& t = tp[k];
                const float ex = 1.f - t.wx, ey = 1.f - t.wy;
                // operation order of ATen's CPU grid_sampler (see warp_sample in loss.cu)
                const float val = __fmaf_rn(t.se, __fmul_rn(t.wy, t.wx),
                                            __fmaf_rn(t.sw, __fmul_rn(t.wy, ex),
                                                      __fmaf_rn(t.ne, __fmul_rn(ey, t.wx),
                                                                __fmul_rn(t.nw, __fmul_rn(ey, ex)))));
                const Charb1 pp = charb1<kVal, kGrad>(val - t.pv);             // loss.py:72-74
                if (kVal) acc_p = fmaf(valid ? 1.f : 0.f, pp.val, acc_p);
                float2 gout = gs[k];
                if (kGrad) {
                    const float ddx = (t.ne - t.nw) * ey + (t.se - t.sw) * t.wy;
                    const float ddy = (t.sw - t.nw) * ex + (t.se - t.ne) * t.wx;
                    gout = f2fma(f2(k_ph * pp.der), make_float2(ddx, ddy), gout);
                }
                if (__any_sync(0xffffffffu, t.bad)) {                          // warp-uniform, rare
                    const Charb2 bo = charb2<kVal, kGrad>(uv[k], f2(0.f));
                    const float m_b = t.bad ? 1.f : 0.f;
                    if (kVal) {
                        acc_o = fmaf(m_b, bo.val.x + bo.val.y, acc_o);
                        acc_n += m_b;
                    }
                    if (kGrad) gout = f2fma(f2(t.bad ? k_ob : 0.f), bo.der, gout);
                }
                if (kGrad && valid) {
                    float* gp = grad + (y * W + x);
                    gp[0] = gout.x;
                    gp[plane] = gout.y;
                }
            }
        }
    }
    if (!kVal) return;
    // record: [photo, smooth_h, smooth_v, smooth_diag (both), -, outborder_sum, n_bad, -]
    float acc[kRec] = {acc_p, acc_h.x + acc_h.y, acc_v.x + acc_v.y, acc_d.x + acc_d.y, 0.f, acc_o, acc_n, 0.f};
    warp_sum8(acc, lane);
    if ((lane & 3) == 0) red[warp][((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)] = acc[0];
    __syncthreads();
    if (threadIdx.x < 7) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kSWarps; ++w) t += red[w][threadIdx.x];
        __stcg(ws.partials + (size_t)blockIdx.x * kRec + threadIdx.x, t);
    }
    finish_tile(tab, tc, ws, terms, n_bad_out, loss_out, dred, &is_last);
}

// Out-of-border pixel count n_i per (scale, prediction) (utils/loss.py:92-113) ahead of the fused
// pass, which needs it for the gradient of the border term: same coordinates, same comparison.
// One 32x32 tile per CTA of NW warps (tile index `cta`).  per_tile: the tile's count is stored at
// n_bad[cta] (no atomics, nothing to zero; the fused kernel sums a pair's tiles itself); otherwise
// it is added to n_bad[s * P + n], which must be zero beforehand.
template <int NW, int NT = 1>
__device__ __forceinline__ void outborder_count_tile(const ScaleTable& tab, int32_t* __restrict__ n_bad,
                                                     bool per_tile, int first_tile,
                                                     int* cnt /* shared, NW * NT ints */) {
    constexpr int kRows = kTile / NW;
    const int lane = threadIdx.x & 31, warp_all = threadIdx.x >> 5;
    const int group = NT > 1 ? warp_all / NW : 0, warp = warp_all - group * NW;
    const int cta = first_tile + group;
    const bool active = NT == 1 || cta < tab.cta_begin[DVSOF_MAX_SCALES];
    const TileCoord tc = decode_cta_fast(tab, active ? cta : 0);
    const int H = tab.H[tc.s], W = tab.W[tc.s];
    const int plane = H * W;
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    const float* flow1 = flow + plane;
    asm volatile("" : "+l"(flow), "+l"(flow1));
    const BorderTest bt = border_test(tab, tc.s);
    const int x = tc.x0 + lane, ys = tc.y0 + warp * kRows;
    const float xf = (float)x;
    int n = 0;
    float u[kRows], v[kRows];
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
        const bool ok = active && x < W && ys + r < H;
        const int off = (ys + r) * W + x;
        u[r] = ok ? __ldg(flow + off) : 0.f;             // 0 flow is never out of border
        v[r] = ok ? __ldg(flow1 + off) : 0.f;
    }
#pragma unroll
    for (int r = 0; r < kRows; ++r) {
        const bool ok = active && x < W && ys + r < H;
        n += (ok && outside_grid(__fadd_rn(xf, u[r]), __fadd_rn((float)(ys + r), v[r]), bt)) ? 1 : 0;
    }
    n = warp_sum(n);
    if (lane == 0) cnt[warp_all] = n;
    __syncthreads();
    if (warp == 0 && lane == 0 && active) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) t += cnt[group * NW + w];
        if (per_tile) n_bad[cta] = t;
        else if (t) atomicAdd(n_bad + tc.s * tab.n_pred + tc.n, t);
    }
}

// The same count with a whole 32x32 tile per WARP (8 rows in flight at a time): the tile decode
// and pointer set-up are paid once per 32 rows and no cross-warp reduction is needed.  Used by the
// 1024-thread pyramid kernel, where a CTA takes 32 tiles.
__device__ __forceinline__ void outborder_count_warp_tile(const ScaleTable& tab, int32_t* __restrict__ n_bad,
                                                          bool per_tile, int cta) {
    constexpr int kRows = 8;
    if (cta >= tab.cta_begin[DVSOF_MAX_SCALES]) return;          // warp-uniform
    const TileCoord tc = decode_cta_fast(tab, cta);
    const int H = tab.H[tc.s], W = tab.W[tc.s];
    const int plane = H * W;
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    const float* flow1 = flow + plane;
    asm volatile("" : "+l"(flow), "+l"(flow1));
    const BorderTest bt = border_test(tab, tc.s);
    const int lane = threadIdx.x & 31;
    const int x = tc.x0 + lane;
    const float xf = (float)x;
    int n = 0;
    if (x < W) {
#pragma unroll 1
        for (int ys = tc.y0; ys < min(tc.y0 + kTile, H); ys += kRows) {
            float u[kRows], v[kRows];
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                const bool ok = ys + r < H;
                const int off = (ys + r) * W + x;
                u[r] = ok ? __ldg(flow + off) : 0.f;             // 0 flow is never out of border
                v[r] = ok ? __ldg(flow1 + off) : 0.f;
            }
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                n += (ys + r < H && outside_grid(__fadd_rn(xf, u[r]), __fadd_rn((float)(ys + r), v[r]), bt)) ? 1 : 0;
            }
        }
    }
    n = warp_sum(n);
    if (lane == 0) {
        if (per_tile) n_bad[cta] = n;
        else if (n) atomicAdd(n_bad + tc.s * tab.n_pred + tc.n, n);
    }
}

// stand-alone launch (dvsof_warp_loss_fused with count_pass; the multiscale call runs the same
// tiles as a side job of the last pyramid level's resize launch, see resize_bilinear_kernel)
__global__ void __launch_bounds__(kSThreads)
outborder_count_kernel(const __grid_constant__ ScaleTable tab, int32_t* __restrict__ n_bad) {
    __shared__ int cnt[kSWarps];
    outborder_count_tile<kSWarps>(tab, n_bad, false, blockIdx.x, cnt);
}

}  // namespace dvsof
