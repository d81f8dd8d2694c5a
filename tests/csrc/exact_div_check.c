/* exact_div_check.c -- CPU proof-by-enumeration for csrc/exact_div.h: dvsof_div_by_const(a, h, r)
 * must equal the IEEE division a / h bit for bit.  Divisors: (S-1)/2 for every image side
 * S in [2, 4096] (every grid normaliser utils/loss.py:152-154 can produce for realistic frames);
 * numerators: pixel + flow, dense around the integers and half-integers where the bilinear cell
 * changes, plus random magnitudes.  Prints the number of mismatches (expected 0).
 * Build: gcc -O2 -ffp-contract=off -o exact_div_check exact_div_check.c -lm */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../dvs_of_training_framework_b200/csrc/exact_div.h"

static uint64_t s = 0x9E3779B97F4A7C15ull;
static uint32_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 32); }
static float bits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

int main(int argc, char** argv) {
    const long per = argc > 1 ? atol(argv[1]) : 20000;
    long bad = 0, total = 0, skipped = 0;
    for (int S = 2; S <= 4096; ++S) {
        const float h = (float)(S - 1) / 2.f;
        if (!dvsof_div_const_ok(h)) { ++skipped; continue; }
        const float r = dvsof_recip_rn(h);
        for (long i = 0; i < per; ++i) {
            float a;
            const uint32_t k = rnd();
            switch (k & 3) {
                case 0: a = (float)(rnd() % (uint32_t)S) + ((float)(rnd() & 0xffff) - 32768.f) * 0x1p-12f; break;
                case 1: a = (float)(rnd() % (uint32_t)S) + bits(0x33000000u + (rnd() & 0x0fffffffu)); break;  /* tiny offsets */
                case 2: a = bits((rnd() & 0x8fffffffu) | ((uint32_t)(100 + rnd() % 60) << 23)); break;      /* wide range */
                default: a = ((float)rnd() / 4294967296.f - 0.5f) * 4.f * (float)S; break;
            }
            const float want = a / h, got = dvsof_div_by_const(a, h, r);
            ++total;
            if (memcmp(&want, &got, 4) != 0 && !(want != want && got != got)) {
                if (bad < 5) printf("mismatch S=%d a=%a want=%a got=%a\n", S, a, want, got);
                ++bad;
            }
            /* the branch-free grid variant: exact in [1e-30, 1e30), harmless outside (see header) */
            const float gg = dvsof_div_by_const_grid(a, h, r), mag = fabsf(a);
            int ok;
            if (mag == 0.f || (mag >= DVSOF_EXACT_DIV_MIN && mag < DVSOF_EXACT_DIV_MAX))
                ok = memcmp(&want, &gg, 4) == 0;
            else if (mag < DVSOF_EXACT_DIV_MIN)
                ok = (gg - 1.f) == (want - 1.f) && (gg + 1.f) == (want + 1.f);
            else
                ok = (gg != gg) == (want != want) && (gg != gg || ((gg < 0) == (want < 0) && fabsf(gg) > 1e20f));
            if (!ok) {
                if (bad < 5) printf("grid variant mismatch S=%d a=%a want=%a got=%a\n", S, a, want, gg);
                ++bad;
            }
        }
    }
    /* the division-free out-of-border test against g = a / h - 1 compared with +-1: every float
     * within 256 ulps of both thresholds and of 0, plus random numerators */
    long border = 0;
    for (int S = 2; S <= 4096; ++S) {
        const float h = (float)(S - 1) / 2.f;
        if (!dvsof_div_const_ok(h)) continue;
        const float r = dvsof_recip_rn(h);
        const float hi = dvsof_grid_outside_hi(h, r), lo = dvsof_grid_outside_lo(h, r);
        const float centres[4] = {hi, lo, 0.f, -0.f};
        for (int c = 0; c < 4; ++c) {
            float up = centres[c], dn = centres[c];
            for (int k = 0; k <= 256; ++k) {
                const float cand[2] = {up, dn};
                for (int m = 0; m < 2; ++m) {
                    const float a = cand[m], g = dvsof_div_by_const_grid(a, h, r) - 1.f, g2 = a / h - 1.f;
                    const int want = g < -1.f || g > 1.f, want2 = g2 < -1.f || g2 > 1.f;
                    const int got = dvsof_grid_outside(a, hi, lo);
                    ++border;
                    if (want != got || want2 != got) {
                        if (bad < 5) printf("border mismatch S=%d a=%a want=%d/%d got=%d\n", S, a, want, want2, got);
                        ++bad;
                    }
                }
                up = nextafterf(up, INFINITY);
                dn = nextafterf(dn, -INFINITY);
            }
        }
        for (long i = 0; i < per / 4; ++i) {
            float a;
            switch (rnd() & 3) {
                case 0: a = bits(rnd()); break;                                        /* any bit pattern */
                case 1: a = ((float)rnd() / 4294967296.f - 0.25f) * 4.f * (float)S; break;
                case 2: a = bits((rnd() & 0x807fffffu) | ((uint32_t)(90 + rnd() % 20) << 23)); break;   /* around -h 2^-24 */
                default: a = 2.f * h + ((float)(rnd() & 0xffff) - 32768.f) * 0x1p-18f * h; break;
            }
            const float g = dvsof_div_by_const_grid(a, h, r) - 1.f;
            const int want = g < -1.f || g > 1.f, got = dvsof_grid_outside(a, hi, lo);
            ++border;
            if (want != got) {
                if (bad < 5) printf("border mismatch S=%d a=%a want=%d got=%d\n", S, a, want, got);
                ++bad;
            }
        }
    }
    printf("checked=%ld border_checked=%ld mismatches=%ld skipped_divisors=%ld\n", total, border, bad, skipped);
    return bad != 0;
}
