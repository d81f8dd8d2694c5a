// loss_ring.cuh -- the warp-loss sweep (included by loss.cu): utils/loss.py:121-171 for all scales
// and predictions in one launch, forward terms and analytic backward.
//
// Work unit = one WARP = one CTA: a SEGMENT of 32 pixel columns x R rows (R = 8, 16 or 32, chosen
// per launch) of one (scale, prediction).  The warp walks its segment top to bottom in stages of
// 8 rows, one pixel column per lane.  Each stage's inputs arrive through TMA
// (cp.async.bulk.tensor.3d, completion on an mbarrier) into a two-stage ring in shared memory:
//     flow box   2 channels x 10 rows (y-1 .. y+8) x 40 columns (x0-4 .. x0+35), zero outside the image
//     prev box   8 rows x 32 columns of the matched `prev` pyramid level
// while the previous stage is being processed, so staging costs one elected lane a handful of
// instructions per 8 rows instead of ~25 per row, and there is no CTA-wide barrier anywhere.
// (W % 4 != 0 or unaligned bases cannot be described by a tensor map: those scales fill the same
// ring with plain loads.)  One warp per CTA: a persistent one-wave grid with the ring running
// across segments was measured slower (static stride vs the hardware CTA scheduler, and a third
// more instructions in the compiler's stage body; commits 1099f6e / 67b5e1a, gpurun_out/r02l-n).
// The bilinear taps of `next` are a data-dependent gather and go through L1 (LDG).  The arithmetic:
//   * every neighbour pair of the smoothness term (utils/loss.py:76-90) is owned by the top-left
//     pixel of its 2x2 block; value AND derivative come from one lg2 + one ex2:
//         s = d^2 + eps^2,  t = s^(alpha-1),  c(d) = t * s,  c'(d) = 2 alpha d t;
//   * the two flow channels travel as one packed fp32x2 value (FADD2 / FMUL2 / FFMA2);
//   * derivatives never leave registers: a pixel's gradient is
//         g(x,y) = L(x-1,y) - Gh(x,y) + Gv(x,y-1) - Gv(x,y) - Gda(x,y) - Gdb(x,y-1)
//         L(x,y) = Gh(x,y) + Gda(x,y-1) + Gdb(x,y)          (everything the right neighbour needs)
//     the row above is carried, the left neighbour arrives with one shuffle per channel; lane 0
//     takes the left halo's L (evaluated for the 8 rows of a stage at once, spread over the lanes)
//     with a predicated shared-memory load;
//   * rows are processed in groups of 2: sampling positions and all taps of the group are
//     requested first, the load-free pair phase runs while they are in flight.
// Stages whose 8 rows and the row below lie inside the image (all but the last of a column) take
// a path without per-row masks.
//
// Modes (template):  <kVal, !kGrad> forward terms;  <!kVal, kGrad> backward for arbitrary
// d(loss)/d(terms);  <kVal, kGrad> the fused training pass: terms and d(sum_k coef_k term_k)/d(flow)
// in a single sweep -- 16 B/pixel read, 8 B/pixel written.
#pragma once
#include <cuda.h>

namespace dvsof {

#ifndef DVSOF_LOSS_GROUP
#define DVSOF_LOSS_GROUP 2                // rows whose taps are in flight together
#endif
#ifndef DVSOF_LOSS_MIN_CTAS
#define DVSOF_LOSS_MIN_CTAS 20            // 1-warp CTAs per SM the register budget is held to
#endif
constexpr int kGroup = DVSOF_LOSS_GROUP;
constexpr int kStageRows = 8;                               // rows per ring stage
constexpr int kRingStages = 2;
// The flow box starts 4 columns left of the segment although only 1 is needed: the innermost TMA
// coordinate has to be a multiple of 16 bytes (x0 - 1 raised "illegal instruction" on B200,
// gpurun_out/r02c vs r02d), and x0 is a multiple of 32.
constexpr int kBoxLeft = 4;
constexpr int kBoxW = 40, kBoxH = kStageRows + 2;           // flow box (floats per row, rows)
constexpr int kPlane = kBoxW * kBoxH;                       // one channel of the flow box
constexpr int kPrevOff = (2 * kPlane + 31) / 32 * 32;       // floats; 128-byte aligned
constexpr int kStageFloats = (kPrevOff + kStageRows * 32 + 31) / 32 * 32;
constexpr unsigned kFlowBoxBytes = 2u * kPlane * 4u, kPrevBoxBytes = kStageRows * 32u * 4u;
static_assert(kStageRows % kGroup == 0, "a stage is a whole number of groups");
static_assert(2 * kPlane <= kPrevOff && kPrevOff + kStageRows * 32 <= kStageFloats, "stage layout");

struct LossMaps {                         // tensor maps of the launch (kernel parameter space)
    CUtensorMap flow[DVSOF_MAX_SCALES];   // (W, H, 2P) fp32, box (kBoxW, 10, 2)
    CUtensorMap prev[DVSOF_MAX_SCALES];   // (W, H, D) fp32, box (32, 8, 1)
};

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

// q = a / b for 0 <= a < 2^23 and the host's float reciprocal of b, exact after one correction
__device__ __forceinline__ int div_small(int a, int b, float inv_b) {
    int q = (int)(((float)a + 0.5f) * inv_b);
    q += (q + 1) * b <= a ? 1 : 0;
    q -= q * b > a ? 1 : 0;
    return q;
}

// (scale, prediction, 32x32 count tile) of border-count job `cta` (tiles_x / tiles_y / cta_begin)
__device__ __forceinline__ TileCoord decode_cta_fast(const ScaleTable& tab, int cta) {
    int s = 0;
#pragma unroll 1
    for (int k = 1; k < tab.n_scales; ++k) s += cta >= tab.cta_begin[k] ? 1 : 0;
    const int local = cta - tab.cta_begin[s];
    const int tx = tab.tiles_x[s], per = tx * tab.tiles_y[s];
    const int n = div_small(local, per, tab.inv_per[s]);
    const int tile = local - n * per;
    const int ty = div_small(tile, tx, tab.inv_tx[s]);
    return TileCoord{s, n, ty * kTile, (tile - ty * tx) * kTile};
}

// (scale, prediction, first row, first column) of sweep segment `blk` (seg_* tables)
__device__ __forceinline__ TileCoord decode_segment(const ScaleTable& tab, int blk) {
    int s = 0;
#pragma unroll 1
    for (int k = 1; k < tab.n_scales; ++k) s += blk >= tab.seg_begin[k] ? 1 : 0;
    const int local = blk - tab.seg_begin[s];
    const int tx = tab.tiles_x[s], per = tab.seg_per[s];
    const int n = div_small(local, per, tab.inv_seg_per[s]);
    const int seg = local - n * per;
    const int cy = div_small(seg, tx, tab.inv_tx[s]);
    return TileCoord{s, n, cy * tab.seg_rows[s], (seg - cy * tx) * kTile};
}

// ---- mbarrier / TMA (one elected lane issues, the whole warp waits) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// the executing CTA's shared window is part of its (1-CTA) cluster window: same 32-bit address
__device__ __forceinline__ void tma_load_3d(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// the flow box of a stage: columns x0 - kBoxLeft .., rows y0 - 1 .., both channels of prediction n
__device__ __forceinline__ void tma_load_flow(unsigned dst, const CUtensorMap* map, unsigned bar, int x0, int y0, int n) {
    tma_load_3d(dst, map, bar, x0 - kBoxLeft, y0 - 1, 2 * n);
}

// Charbonnier pair with the term's normaliser folded into the exponent: with lk = log2(k),
//     t' = 2^((alpha-1) lg2(s) + lk) = k s^(alpha-1),   der' = d t' = k c'(d) / (2 alpha),   val' = t' s = k c(d)
// so neither the gradient nor the value needs a multiply by k afterwards, and a pair that does not
// exist (image border) is switched off with lk = -inf (ex2(-inf) = +0: der' = val' = 0).
template <bool kVal, bool kGrad>
__device__ __forceinline__ Charb2 charb2k(float2 p, float2 q, float lk) {
    const float2 d = f2sub(p, q);
    const float2 s = f2fma(d, d, f2(kEps2));
    const float2 l = make_float2(lg2_approx(s.x), lg2_approx(s.y));
    Charb2 r;
    if (kGrad) {
        const float2 e = f2fma(l, f2(kAlpha - 1.f), f2(lk));
        const float2 t = make_float2(ex2_approx(e.x), ex2_approx(e.y));
        r.der = f2mul(d, t);
        r.val = f2mul(t, s);
    } else {
        const float2 e = f2fma(l, f2(kAlpha), f2(lk));
        r.der = f2(0.f);
        r.val = make_float2(ex2_approx(e.x), ex2_approx(e.y));
    }
    return r;
}

// Everything a warp carries from row to row and stage to stage.
struct SweepState {
    float2 a, b;                          // flow at (x, y), (x+1, y) of the current row
    float2 gv_p, gda_p, gdb_p;            // normalised pair derivatives of the row above
    float2 acc_s;                         // smoothness: sum of k c(d) over all four directions (both channels)
    float2 acc_p;                         // photometric sum (even / odd rows of the row pairs)
    float acc_o, acc_n;                   // border sum, border count
};
struct SweepConst {
    int H, W;
    float Hf, Wf, xf;
    bool x_in;
    float lkh, lkv, lkd;                  // log2 of the smoothness normalisers, -inf where this lane has no such pair
    float lk_h0, lk_d0;                   // the same for the halo column left of the segment (unmasked)
    float g_sm, k_ph, k_ob;               // d(loss)/d(term) x normaliser of the remaining terms
    GridNorm gn;
    const float* next;
    float* grad0;                         // this lane's column of the gradient, channel 0 / 1:
    float* grad1;                         // pixel (x, y) at [y * W]
};

__device__ __forceinline__ float2 f2pair(float lo, float hi) { return make_float2(lo, hi); }

// 8 rows y0 .. y0+7 from stage buffer F (flow box rows y0-1 .. y0+8, prev box rows y0 .. y0+7).
// kEdge = false: all 8 rows and row y0+8 lie inside the image (no per-row masks).
//
// The smoothness pairs are evaluated per pixel with the two flow CHANNELS packed (u, v); the
// photometric path (sampling position, weights, bilinear sum, Charbonnier, d/dx d/dy) is evaluated
// for the two ROWS of a group packed (row k, row k+1): the same FADD2 / FMUL2 / FFMA2 do two pixels.
template <bool kVal, bool kGrad, bool kEdge>
__device__ __forceinline__ void sweep_stage(const float* F, float2* lh_sm, const SweepConst& c, SweepState& st,
                                            int y0, bool has_left, int lane) {
    static_assert(kGroup == 2, "the photometric path packs the two rows of a group");
    const int lx = lane + kBoxLeft;
    const float ninf = -INFINITY;
    // ---- L of the cell left of the segment for the 8 rows: lane j evaluates row y0 + j ----
    if (kGrad && has_left) {              // warp-uniform
        const int j = lane & 7, y = y0 + j;
        const float* fu = F + (j + 1) * kBoxW + (kBoxLeft - 1);
        const float* fv = fu + kPlane;
        const float2 p = make_float2(fu[0], fv[0]), q = make_float2(fu[1], fv[1]);
        const float2 up = make_float2(fu[-kBoxW], fv[-kBoxW]), dn = make_float2(fu[kBoxW], fv[kBoxW]);
        // the halo cell's pairs all end in column x0 >= 1: they exist in x; rows decide
        const float l_h = (!kEdge || y < c.H) ? c.lk_h0 : ninf;
        const float l_a = (y >= 1 && (!kEdge || y < c.H)) ? c.lk_d0 : ninf;
        const float l_b = (!kEdge || y + 1 < c.H) ? c.lk_d0 : ninf;
        float2 lh = charb2k<false, true>(q, p, l_h).der;
        lh = f2add(lh, charb2k<false, true>(q, up, l_a).der);
        lh = f2add(lh, charb2k<false, true>(q, dn, l_b).der);
        __syncwarp();                     // the previous stage's reads of lh_sm are done
        if (lane < kStageRows) lh_sm[lane] = lh;
        __syncwarp();
    }
    const unsigned lh_addr = smem_u32(lh_sm);
    int goff = y0 * c.W;                  // warp-uniform element offset of row y in this lane's column
    const float yf0 = (float)y0;
    const float2 hx2 = f2(c.gn.hx), hy2 = f2(c.gn.hy), rx2 = f2(c.gn.rx), ry2 = f2(c.gn.ry);
#pragma unroll
    for (int g0 = 0; g0 < kStageRows; g0 += 2) {
        const float* fr0 = F + (g0 + 1) * kBoxW + lx;      // box row of image row y0 + g0, this lane's column
        const bool v0 = c.x_in && (!kEdge || y0 + g0 < c.H), v1 = c.x_in && (!kEdge || y0 + g0 + 1 < c.H);
        // ---- phase 1 (rows packed): sampling positions of both rows, all eight taps requested ----
        // utils/loss.py:150-156 + ATen's CPU unnormalize, operation for operation (notes in loss.cu);
        // the position is clamped to [-2, S] in float before the conversion (out there no tap lies
        // inside the image: sample and derivative are 0 whatever the weights), so one F2I.FLOOR
        // replaces floor + F2I
        const float2 U = f2pair(fr0[0], fr0[kBoxW]), V = f2pair(fr0[kPlane], fr0[kPlane + kBoxW]);
        const float2 ax = f2add(f2(c.xf), U), ay = f2add(f2pair(yf0 + (float)g0, yf0 + (float)(g0 + 1)), V);
        const float2 qx0 = f2mul(ax, rx2), qy0 = f2mul(ay, ry2);
        const float2 qx = f2fma(f2fma(qx0, f2pair(-c.gn.hx, -c.gn.hx), ax), rx2, qx0);   // exact_div.h: a / h
        const float2 qy = f2fma(f2fma(qy0, f2pair(-c.gn.hy, -c.gn.hy), ay), ry2, qy0);
        const float2 gx = f2add(qx, f2(-1.f)), gy = f2add(qy, f2(-1.f));
        const bool bad0 = v0 && fmaxf(fabsf(gx.x), fabsf(gy.x)) > 1.f;                  // loss.py:92-94
        const bool bad1 = v1 && fmaxf(fabsf(gx.y), fabsf(gy.y)) > 1.f;
        const float2 ix = f2mul(f2add(gx, f2(1.f)), hx2), iy = f2mul(f2add(gy, f2(1.f)), hy2);
        const float cx0 = fminf(fmaxf(ix.x, -2.f), c.Wf), cx1 = fminf(fmaxf(ix.y, -2.f), c.Wf);
        const float cy0 = fminf(fmaxf(iy.x, -2.f), c.Hf), cy1 = fminf(fmaxf(iy.y, -2.f), c.Hf);
        const int xi0 = __float2int_rd(cx0), xi1 = __float2int_rd(cx1);
        const int yi0 = __float2int_rd(cy0), yi1 = __float2int_rd(cy1);
        const float2 wx = f2sub(f2pair(cx0, cx1), f2pair((float)xi0, (float)xi1));       // weight of east
        const float2 wy = f2sub(f2pair(cy0, cy1), f2pair((float)yi0, (float)yi1));       // weight of south
        float2 nw, ne, sw, se;
        {
            const bool xa = (unsigned)xi0 < (unsigned)c.W, xb = (unsigned)(xi0 + 1) < (unsigned)c.W;
            const bool ya = (unsigned)yi0 < (unsigned)c.H, yb = (unsigned)(yi0 + 1) < (unsigned)c.H;
            const int idx = yi0 * c.W + xi0;
            const float* p0 = c.next + idx;
            const float* p1 = c.next + (idx + c.W);
            nw.x = (xa && ya) ? __ldg(p0) : 0.f;
            ne.x = (xb && ya) ? __ldg(p0 + 1) : 0.f;
            sw.x = (xa && yb) ? __ldg(p1) : 0.f;
            se.x = (xb && yb) ? __ldg(p1 + 1) : 0.f;
        }
        {
            const bool xa = (unsigned)xi1 < (unsigned)c.W, xb = (unsigned)(xi1 + 1) < (unsigned)c.W;
            const bool ya = (unsigned)yi1 < (unsigned)c.H, yb = (unsigned)(yi1 + 1) < (unsigned)c.H;
            const int idx = yi1 * c.W + xi1;
            const float* p0 = c.next + idx;
            const float* p1 = c.next + (idx + c.W);
            nw.y = (xa && ya) ? __ldg(p0) : 0.f;
            ne.y = (xb && ya) ? __ldg(p0 + 1) : 0.f;
            sw.y = (xa && yb) ? __ldg(p1) : 0.f;
            se.y = (xb && yb) ? __ldg(p1 + 1) : 0.f;
        }
        // ---- phase 2 (channels packed): neighbour pairs of both rows (no global loads) ----
        float2 gs[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int r = g0 + k, y = y0 + r;
            const bool row_in = !kEdge || y < c.H, row_d = !kEdge || y + 1 < c.H;    // warp-uniform
            const float* fn = fr0 + (k + 1) * kBoxW;                                  // the row below
            const float2 cn = make_float2(fn[0], fn[kPlane]), dn = make_float2(fn[1], fn[kPlane + 1]);
            const float l_h = row_in ? c.lkh : ninf, l_v = row_d ? c.lkv : ninf, l_d = row_d ? c.lkd : ninf;
            const Charb2 ph = charb2k<kVal, kGrad>(st.b, st.a, l_h);                 // loss.py:76-90
            const Charb2 pvv = charb2k<kVal, kGrad>(cn, st.a, l_v);
            const Charb2 pa = charb2k<kVal, kGrad>(dn, st.a, l_d);
            const Charb2 pb = charb2k<kVal, kGrad>(st.b, cn, l_d);
            if (kVal) st.acc_s = f2add(st.acc_s, f2add(f2add(ph.val, pvv.val), f2add(pa.val, pb.val)));
            gs[k] = f2(0.f);
            if (kGrad) {
                const float2 L = f2add(f2add(ph.der, st.gda_p), pb.der);
                float2 Ll = make_float2(__shfl_up_sync(0xffffffffu, L.x, 1), __shfl_up_sync(0xffffffffu, L.y, 1));
                // lane 0: the left halo's L of this row (0 in the first column block)
                asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %2, 0;\n\t"
                             "@p ld.shared.v2.f32 {%0, %1}, [%3];\n\t}"
                             : "+f"(Ll.x), "+f"(Ll.y) : "r"(lane), "r"(lh_addr + 8u * r) : "memory");
                gs[k] = f2sub(f2add(f2sub(Ll, ph.der), f2sub(st.gv_p, pvv.der)), f2add(pa.der, st.gdb_p));
                st.gv_p = pvv.der; st.gda_p = pa.der; st.gdb_p = pb.der;
            }
            st.a = cn; st.b = dn;
        }
        // ---- phase 3 (rows packed): consume the taps: photometric term and its gradient ----
        const float2 pv = f2pair(F[kPrevOff + g0 * 32 + lane], F[kPrevOff + (g0 + 1) * 32 + lane]);
        const float2 ex = f2sub(f2(1.f), wx), ey = f2sub(f2(1.f), wy);
        // operation order of ATen's CPU grid_sampler: nw*(ey*ex) then fma ne*(ey*wx), sw*(wy*ex), se*(wy*wx)
        const float2 val = f2fma(se, f2mul(wy, wx), f2fma(sw, f2mul(wy, ex), f2fma(ne, f2mul(ey, wx), f2mul(nw, f2mul(ey, ex)))));
        const float2 d = f2sub(val, pv);                                             // loss.py:72-74
        const float2 s2 = f2fma(d, d, f2(kEps2));
        const float2 l2 = make_float2(lg2_approx(s2.x), lg2_approx(s2.y));
        float2 pder = f2(0.f);
        if (kGrad) {
            const float2 t = make_float2(ex2_approx((kAlpha - 1.f) * l2.x), ex2_approx((kAlpha - 1.f) * l2.y));
            pder = f2mul(d, t);
            if (kVal) st.acc_p = f2fma(f2pair(v0 ? 1.f : 0.f, v1 ? 1.f : 0.f), f2mul(t, s2), st.acc_p);
        } else {
            const float2 cval = make_float2(ex2_approx(kAlpha * l2.x), ex2_approx(kAlpha * l2.y));
            st.acc_p = f2fma(f2pair(v0 ? 1.f : 0.f, v1 ? 1.f : 0.f), cval, st.acc_p);
        }
        float ox0 = 0.f, oy0 = 0.f, ox1 = 0.f, oy1 = 0.f;
        if (kGrad) {
            const float2 ddx = f2fma(f2sub(se, sw), wy, f2mul(f2sub(ne, nw), ey));
            const float2 ddy = f2fma(f2sub(se, ne), wx, f2mul(f2sub(sw, nw), ex));
            const float2 kder = f2mul(f2(c.k_ph), pder);
            const float2 px = f2mul(kder, ddx), py = f2mul(kder, ddy);
            ox0 = fmaf(c.g_sm, gs[0].x, px.x); oy0 = fmaf(c.g_sm, gs[0].y, py.x);
            ox1 = fmaf(c.g_sm, gs[1].x, px.y); oy1 = fmaf(c.g_sm, gs[1].y, py.y);
        }
        if (__any_sync(0xffffffffu, bad0 || bad1)) {                                 // warp-uniform, rare
            // border term (loss.py:96-119): Charbonnier of the flow itself on out-of-border pixels
            const Charb2 b0 = charb2<kVal, kGrad>(make_float2(U.x, V.x), f2(0.f));
            const Charb2 b1 = charb2<kVal, kGrad>(make_float2(U.y, V.y), f2(0.f));
            const float m0 = bad0 ? 1.f : 0.f, m1 = bad1 ? 1.f : 0.f;
            if (kVal) {
                st.acc_o = fmaf(m0, b0.val.x + b0.val.y, st.acc_o);
                st.acc_o = fmaf(m1, b1.val.x + b1.val.y, st.acc_o);
                st.acc_n += m0 + m1;
            }
            if (kGrad) {
                const float k0 = bad0 ? c.k_ob : 0.f, k1 = bad1 ? c.k_ob : 0.f;
                ox0 = fmaf(k0, b0.der.x, ox0); oy0 = fmaf(k0, b0.der.y, oy0);
                ox1 = fmaf(k1, b1.der.x, ox1); oy1 = fmaf(k1, b1.der.y, oy1);
            }
        }
        if (kGrad) {
            if (v0) { c.grad0[goff] = ox0; c.grad1[goff] = oy0; }
            if (v1) { c.grad0[goff + c.W] = ox1; c.grad1[goff + c.W] = oy1; }
        }
        goff += 2 * c.W;
    }
}

// A stage filled with plain loads (scales a tensor map cannot describe): the same box layout,
// zero outside the image.  All lanes; the caller separates it from the readers with __syncwarp.
__device__ __forceinline__ void stage_manual_flow(float* F, const float* __restrict__ flow, int H, int W, int y0,
                                                  int x0, int lane) {
    const int plane = H * W;
#pragma unroll 1
    for (int r = 0; r < kBoxH; ++r) {
        const int y = y0 - 1 + r;
        const bool yok = y >= 0 && y < H;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int cidx = lane + 32 * h, x = x0 - kBoxLeft + cidx;
            if (cidx < kBoxW) {
                const bool ok = yok && x >= 0 && x < W;
                F[r * kBoxW + cidx] = ok ? __ldg(flow + y * W + x) : 0.f;
                F[kPlane + r * kBoxW + cidx] = ok ? __ldg(flow + plane + y * W + x) : 0.f;
            }
        }
    }
}
__device__ __forceinline__ void stage_manual_prev(float* F, const float* __restrict__ prev, int H, int W, int y0,
                                                  int x0, int lane) {
#pragma unroll
    for (int r = 0; r < kStageRows; ++r) {
        const int y = y0 + r, x = x0 + lane;
        F[kPrevOff + r * 32 + lane] = (y < H && x < W) ? __ldg(prev + y * W + x) : 0.f;
    }
}

template <bool kVal, bool kGrad>
__global__ void __launch_bounds__(32, DVSOF_LOSS_MIN_CTAS)
warp_loss_ring_kernel(const __grid_constant__ ScaleTable tab, const __grid_constant__ LossMaps maps,
                      const int32_t* __restrict__ match_idx, const float* __restrict__ grad_terms,
                      const int32_t* __restrict__ n_bad_in, int bad_per_tile, const LossWorkspace ws) {
    __shared__ __align__(128) float ring[kRingStages][kStageFloats];
    __shared__ __align__(8) uint64_t bars[kRingStages];
    __shared__ __align__(8) float2 lh_sm[kStageRows];
    pdl_launch_dependents();
    const int lane = threadIdx.x;
    const TileCoord tc = decode_segment(tab, blockIdx.x);
    const int S = tab.n_scales, P = tab.n_pred;
    const int H = tab.H[tc.s], W = tab.W[tc.s];
    const int plane = H * W;
    const int rows = min(tab.seg_rows[tc.s], H - tc.y0);
    const int n_stages = (rows + kStageRows - 1) / kStageRows;
    const bool tma_flow = (tab.use_tma[tc.s] & 1) != 0, tma_prev = (tab.use_tma[tc.s] & 2) != 0;
    const bool tma = tma_flow || tma_prev;
    const unsigned stage_bytes = (tma_flow ? kFlowBoxBytes : 0u) + (tma_prev ? kPrevBoxBytes : 0u);
    const float* flow = tab.flow[tc.s] + (size_t)tc.n * 2 * plane;
    const unsigned bar0 = smem_u32(&bars[0]), ring0 = smem_u32(&ring[0][0]);
    if (lane < kStageRows) lh_sm[lane] = make_float2(0.f, 0.f);
    if (tma && lane == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // the flow is an input of the call: stage 0's box is requested while the prelude (pyramid
        // levels, match_idx, tile counts, tickets) may still be running
        mbar_expect_tx(bar0, stage_bytes);
        if (tma_flow) tma_load_flow(ring0, &maps.flow[tc.s], bar0, tc.x0, tc.y0, tc.n);
    }
    pdl_wait_for(match_idx, n_bad_in);             // both come from the prelude (see pdl_acquire)
    const int i_prev = match_idx[tc.n], i_next = match_idx[P + tc.n];
    if (tma_prev && lane == 0)
        tma_load_3d(ring0 + 4u * kPrevOff, &maps.prev[tc.s], bar0, tc.x0, tc.y0, i_prev);
    __syncwarp();

    SweepConst c;
    c.H = H; c.W = W;
    c.Hf = (float)H; c.Wf = (float)W;
    const int x = tc.x0 + lane;
    c.xf = (float)x;
    c.x_in = x < W;
    const bool x_r = x + 1 < W;
    // log2 of 2 alpha / (4 x element count of the mean) per direction, precomputed on the host
    // (-inf for a direction without pairs); lanes whose pixel has no such pair: -inf
    c.lk_h0 = tab.lk[0][tc.s]; c.lk_d0 = tab.lk[2][tc.s];
    c.lkh = x_r ? c.lk_h0 : -INFINITY;
    c.lkv = c.x_in ? tab.lk[1][tc.s] : -INFINITY;
    c.lkd = x_r ? c.lk_d0 : -INFINITY;
    c.g_sm = c.k_ph = c.k_ob = 0.f;
    if (kGrad) {
        // d(loss)/d(term) of this scale: device vector (generic backward) or host coefficients (fused)
        c.g_sm = grad_terms ? grad_terms[tc.s] : tab.coef[tc.s];
        const float g_ph = grad_terms ? grad_terms[S + tc.s] : tab.coef[DVSOF_MAX_SCALES + tc.s];
        const float g_ob = grad_terms ? grad_terms[2 * S + tc.s] : tab.coef[2 * DVSOF_MAX_SCALES + tc.s];
        c.k_ph = g_ph * tab.norm[3][tc.s];
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
        c.k_ob = nb > 0 ? __fdividef(g_ob * tab.norm[4][tc.s], (float)nb) : 0.f;   // loss.py:99-113
    }
    c.gn = GridNorm{tab.hx[tc.s], tab.hy[tc.s], tab.rx[tc.s], tab.ry[tc.s]};
    const float* prev = tab.images[tc.s] + (size_t)i_prev * plane;
    c.next = tab.images[tc.s] + (size_t)i_next * plane;
    c.grad0 = kGrad ? tab.grad_flow[tc.s] + (size_t)tc.n * 2 * plane + x : nullptr;
    c.grad1 = kGrad ? c.grad0 + plane : nullptr;
    asm volatile("" : "+l"(c.next), "+l"(c.grad0), "+l"(c.grad1));  // finished bases: access = base + 32-bit offset

    SweepState st;
    st.a = st.b = f2(0.f);
    st.acc_s = st.acc_p = f2(0.f);
    st.acc_o = st.acc_n = 0.f;
    st.gv_p = st.gda_p = st.gdb_p = f2(0.f);
    const bool has_left = tc.x0 > 0;
#pragma unroll 1
    for (int j = 0; j < n_stages; ++j) {
        const int y0 = tc.y0 + j * kStageRows;
        float* F = ring[j & 1];
        if (tma) {
            if (j + 1 < n_stages) {               // request the next stage, then wait for this one
                __syncwarp();                     // every lane is done with the buffer being refilled
                if (lane == 0) {
                    const unsigned bar = bar0 + 8u * ((j + 1) & 1), dst = ring0 + 4u * kStageFloats * ((j + 1) & 1);
                    mbar_expect_tx(bar, stage_bytes);
                    if (tma_flow) tma_load_flow(dst, &maps.flow[tc.s], bar, tc.x0, y0 + kStageRows, tc.n);
                    if (tma_prev) tma_load_3d(dst + 4u * kPrevOff, &maps.prev[tc.s], bar, tc.x0, y0 + kStageRows, i_prev);
                }
            }
        }
        if (!tma_flow || !tma_prev) {             // the parts no tensor map describes: plain loads
            __syncwarp();
            if (!tma_flow) stage_manual_flow(F, flow, H, W, y0, tc.x0, lane);
            if (!tma_prev) stage_manual_prev(F, prev, H, W, y0, tc.x0, lane);
            __syncwarp();
        }
        if (tma) mbar_wait(bar0 + 8u * (j & 1), (unsigned)(j >> 1) & 1u);
        if (j == 0) {                             // first row of the segment and the cell row above it
            const float* fu = F + kBoxW + lane + kBoxLeft;
            st.a = make_float2(fu[0], fu[kPlane]);
            st.b = make_float2(fu[1], fu[kPlane + 1]);
            if (kGrad && tc.y0 > 0) {             // the pairs reaching into row y0 from above
                const float2 ua = make_float2(fu[-kBoxW], fu[kPlane - kBoxW]);
                const float2 ub = make_float2(fu[1 - kBoxW], fu[kPlane + 1 - kBoxW]);
                st.gv_p = charb2k<false, true>(st.a, ua, c.lkv).der;
                st.gda_p = charb2k<false, true>(st.b, ua, c.lkd).der;
                st.gdb_p = charb2k<false, true>(ub, st.a, c.lkd).der;
            }
        }
        if (y0 + kStageRows < H) sweep_stage<kVal, kGrad, false>(F, lh_sm, c, st, y0, has_left, lane);
        else sweep_stage<kVal, kGrad, true>(F, lh_sm, c, st, y0, has_left, lane);
    }
    if (!kVal) return;
    // record: [photo, smoothness (normalised, all directions), outborder_sum, n_bad]; summed by
    // loss_reduce_kernel (loss.cu) after this launch -- no ticket, no fence here
    float acc[8] = {st.acc_p.x + st.acc_p.y, st.acc_s.x + st.acc_s.y, st.acc_o, st.acc_n, 0.f, 0.f, 0.f, 0.f};
    warp_sum8(acc, lane);
    if ((lane & 19) == 0)                         // lanes 0, 4, 8, 12 hold the totals of values 0, 1, 2, 3
        __stcg(ws.partials + (size_t)blockIdx.x * kRec + (((lane >> 3) & 1) * 2 + ((lane >> 2) & 1)), acc[0]);
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
// tiles as a side job of the pyramid launch)
constexpr int kCountWarps = 4;
__global__ void __launch_bounds__(32 * kCountWarps)
outborder_count_kernel(const __grid_constant__ ScaleTable tab, int32_t* __restrict__ n_bad) {
    __shared__ int cnt[kCountWarps];
    outborder_count_tile<kCountWarps>(tab, n_bad, false, blockIdx.x, cnt);
}

}  // namespace dvsof
