/* Property check for the division-free segment-crossing test of the contact resolver (csrc/kernels.cuh: segments_intersect)
 * against the reference's rounded-quotient form (PhysicsEngine::get_intersecting, PhysicsEngine.cpp:1094-1121):
 *
 *     reference:  t = n / det;  0 <= t && t <= 1            (IEEE division, round to nearest even)
 *     engine:     det < 0 ? (n = -n, det = -det) : 0;  0 <= n && n <= det
 *
 * Claim: for every finite n and every det with |det| >= 1e-4 (smaller determinants are rejected before) the two agree,
 * except when n / det underflows to -0.0 (n < 0, |n| < 2^-1075 |det|: impossible for |det| < 1, and the lengths of two
 * voxel edges give |det| <= 0.03).  Proof sketch for the upper end: n > det > 0 means n >= det + ulp(det), i.e.
 * n / det >= 1 + 1/M with det = M 2^e, 2^52 <= M < 2^53; 1/M > 2^-53 puts the quotient strictly above the midpoint of
 * 1 and 1 + 2^-52, so it rounds to a value > 1; the lower end is the sign of n.  This program hammers the boundary cases:
 * n = det +- k ulp, n = +-0, +-denormals, random n / det of all magnitudes, both signs of det.
 *
 *     gcc -O2 -ffp-contract=off tools/check_segment_test.c -lm -o /tmp/cst && /tmp/cst       (expects "mismatches 0")
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t s = 0x9E3779B97F4A7C15ull;
static inline uint64_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static inline double from_bits(uint64_t b) { double d; memcpy(&d, &b, 8); return d; }
static inline uint64_t to_bits(double d) { uint64_t b; memcpy(&b, &d, 8); return b; }

static int ref_in(double n, double det) { const double t = n / det; return 0.0 <= t && t <= 1.0; }
static int eng_in(double n, double det) {
    if (det < 0) { n = -n; det = -det; }
    return 0.0 <= n && n <= det;
}

int main(int argc, char** argv) {
    long bad = 0, underflow_cases = 0, underflow_small_det = 0, total = 0;
    const long cases = argc > 1 ? atol(argv[1]) : 200000000L;
    for (long i = 0; i < cases; i++) {
        /* det: magnitude 1e-4 .. 1e3 (the engine's edges give <= 0.03), random sign and mantissa */
        const int e = -14 + (int)(rnd() % 24);
        double det = ldexp(1.0 + (double)(rnd() >> 12) / 4503599627370496.0, e);
        if (fabs(det) < 1e-4) det = 1e-4 + fabs(det);
        if (rnd() & 1) det = -det;
        double n;
        switch (i & 7) {
            case 0: n = from_bits(to_bits(det) + (rnd() % 5) - 2); break;                 /* det +- 2 ulp (magnitude) */
            case 1: n = (rnd() & 1) ? 0.0 : -0.0; break;
            case 2: n = from_bits((rnd() % 64) | ((rnd() & 1) << 63)); break;              /* +- tiny denormals */
            case 3: n = det * (1.0 + ldexp((double)((long)(rnd() % 9) - 4), -52)); break;  /* products near det */
            case 4: n = ldexp((double)(rnd() >> 11) / 9007199254740992.0, -(int)(rnd() % 1000)) * ((rnd() & 1) ? 1 : -1); break;
            case 5: n = from_bits(to_bits(-det) + (rnd() % 5) - 2); break;                 /* -det +- 2 ulp */
            default: n = det * ((double)(rnd() >> 11) / 9007199254740992.0 * 2.5 - 0.75); break;   /* n / det in [-0.75, 1.75] */
        }
        total++;
        const int r = ref_in(n, det), g = eng_in(n, det);
        if (r != g) {
            const double t = n / det;
            if (t == 0.0 && n != 0.0) {                                      /* the documented underflow exception */
                underflow_cases++;
                if (fabs(det) < 1.0) underflow_small_det++;                  /* must never happen: |n / det| > |n| there */
                continue;
            }
            if (bad < 10) printf("  mismatch n=%a det=%a t=%a ref=%d engine=%d\n", n, det, t, r, g);
            bad++;
        }
    }
    printf("cases %ld  mismatches %ld  underflow exceptions %ld  of which with |det| < 1: %ld\n", total, bad, underflow_cases, underflow_small_det);
    return bad != 0 || underflow_small_det != 0;
}
