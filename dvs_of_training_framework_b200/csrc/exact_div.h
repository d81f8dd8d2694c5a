// exact_div.h -- correctly rounded fp32 division by a loop-invariant divisor without the
// per-element reciprocal: q = RN(a / h) from r = RN(1 / h) (computed once on the host) in three
// fp32 operations (Markstein's correction step):
//     q0  = RN(a * r)            faithful quotient
//     rem = fma(-q0, h, a)       exact remainder (representable: |rem| < ulp(q0) * h)
//     q   = fma(rem, r, q0)      correctly rounded for every a in the normal range when r is the
//                                correctly rounded reciprocal and the significand of h is not all ones
// The warp-loss kernels divide every pixel coordinate by (W-1)/2 and (H-1)/2 (utils/loss.py:152-154,
// true division on the reference's CPU path); the selected bilinear cell depends on the last bit
// of that quotient, so the division must be IEEE-exact -- this does it in 3 FMA-pipe instructions
// instead of MUFU.RCP + 8.  Shared by the CUDA kernels and by tests/exact_div_check.c (gcc),
// which compares it with `a / h` over ~10^8 random operands and every divisor the configs use.
#ifndef DVSOF_EXACT_DIV_H_
#define DVSOF_EXACT_DIV_H_

#include <math.h>

#ifdef __CUDACC__
#define DVSOF_HD __host__ __device__ __forceinline__
#else
#define DVSOF_HD static inline
#endif

// |a| above this (or non-finite) and tiny quotients go through the plain division
#define DVSOF_EXACT_DIV_MAX 1.0e30f
#define DVSOF_EXACT_DIV_MIN 1.0e-30f

DVSOF_HD float dvsof_div_by_const(float a, float h, float r) {
    const float mag = fabsf(a);
#if defined(__GNUC__) || defined(__CUDACC__)
    if (__builtin_expect(!(mag < DVSOF_EXACT_DIV_MAX) || (mag < DVSOF_EXACT_DIV_MIN && mag != 0.f), 0)) {
#else
    if (!(mag < DVSOF_EXACT_DIV_MAX) || (mag < DVSOF_EXACT_DIV_MIN && mag != 0.f)) {
#endif
#ifdef __CUDA_ARCH__
        return __fdiv_rn(a, h);
#else
        return a / h;
#endif
    }
#ifdef __CUDA_ARCH__
    const float q0 = __fmul_rn(a, r);
    const float rem = __fmaf_rn(-q0, h, a);
    return __fmaf_rn(rem, r, q0);
#else
    const float q0 = a * r;
    const float rem = fmaf(-q0, h, a);
    return fmaf(rem, r, q0);
#endif
}

// Branch-free variant for the sampling grid g = a / h - 1 of the warp-loss kernels (5 instructions,
// no slow path): bit-identical to a / h for 1e-30 <= |a| < 1e30 and a == 0.  Outside that range
// the callers do not need the last bit: for tinier |a| the quotient only has to vanish against the
// "- 1" that follows (it does: |q| < 1e-29), for larger or non-finite |a| it only has to keep
// sign, magnitude class (huge / inf / NaN) -- RN(a * r) does.
DVSOF_HD float dvsof_div_by_const_grid(float a, float h, float r) {
#ifdef __CUDA_ARCH__
    const float q0 = __fmul_rn(a, r);
    const float rem = __fmaf_rn(-q0, h, a);
    const float q = __fmaf_rn(rem, r, q0);
#else
    const float q0 = a * r;
    const float rem = fmaf(-q0, h, a);
    const float q = fmaf(rem, r, q0);
#endif
    return fabsf(a) < DVSOF_EXACT_DIV_MAX ? q : q0;
}

// Out-of-border test of the grid coordinate g = RN(RN(a / h) - 1) without computing it:
//   g >  1  <=>  RN(a / h) > 2       (floats above 1 that round back to 1 under "- 1" do not exist
//                                     for a float q: q > 2 means q >= 2 + 2^-22, so q - 1 > 1)
//           <=>  a > 2h              (2h is a float; the next float above it divides to more than
//                                     the rounding midpoint 2 + 2^-23, 2h itself gives exactly 2)
//   g < -1  <=>  RN(a / h) < -2^-24  (quotients in [-2^-24, 0) are absorbed by the "- 1")
//           <=>  a < -h * 2^-24      (same argument one binade down; h * 2^-24 is exact)
// A NaN or missing reciprocal (1-pixel side: the reference divides by zero) never tests outside,
// like the comparisons on a NaN grid.  Enumerated against the division in
// tests/csrc/exact_div_check.c.
DVSOF_HD float dvsof_grid_outside_hi(float h, float r) { return r == r ? 2.f * h : INFINITY; }
DVSOF_HD float dvsof_grid_outside_lo(float h, float r) { return r == r ? -h * 0x1p-24f : -INFINITY; }
DVSOF_HD int dvsof_grid_outside(float a, float hi, float lo) { return a > hi || a < lo; }

// host: correctly rounded fp32 reciprocal of h (h > 0, normal), and whether the fast path may be
// used for this divisor (significand not all ones, sane exponent)
static inline float dvsof_recip_rn(float h) {
    const double exact = 1.0 / (double)h;       // 53-bit; the float rounding below is the only
    float r = (float)exact;                     // inexact step unless `exact` sits on a float tie
    // fix a possible double rounding: pick the neighbour whose residual |1 - h*r| is smallest
    const float lo = nextafterf(r, 0.f), hi = nextafterf(r, INFINITY);
    double best = fabs(1.0 - (double)h * (double)r);
    const double elo = fabs(1.0 - (double)h * (double)lo), ehi = fabs(1.0 - (double)h * (double)hi);
    if (elo < best) { best = elo; r = lo; }
    if (ehi < best) { r = hi; }
    return r;
}

static inline int dvsof_div_const_ok(float h) {
    if (!(h >= 0x1p-60f && h <= 0x1p60f)) return 0;
    int e;
    const float m = frexpf(h, &e);              // m in [0.5, 1)
    return m != 0x1.fffffep-1f;                 // all-ones significand: use the plain division
}

#endif  // DVSOF_EXACT_DIV_H_
