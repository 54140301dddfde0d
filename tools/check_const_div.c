#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
static uint64_t s = 88172645463325252ull;
static inline uint64_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static inline double mk(uint64_t mant, int e, int sign) {
    uint64_t bits = ((uint64_t)sign << 63) | ((uint64_t)(e + 1023) << 52) | (mant & 0xfffffffffffffull);
    double d; memcpy(&d, &bits, 8); return d;
}
static long test(double b, long n, int emin, int emax) {
    const double rb = 1.0 / b;
    long bad = 0;
    for (long i = 0; i < n; i++) {
        uint64_t r = rnd();
        int e = emin + (int)((r >> 53) % (uint64_t)(emax - emin + 1));
        double a = mk(rnd(), e, (int)(r & 1));
        // also hit structured mantissas
        if ((i & 7) == 0) a = mk((rnd() & 0xff) << 44, e, 0);
        if ((i & 7) == 1) a = mk(~(rnd() & 0xffff), e, 0);
        double q = a * rb;
        double rr = fma(-q, b, a);
        double q2 = fma(rr, rb, q);
        double ref = a / b;
        if (q2 != ref) { if (bad < 5) printf("  mismatch a=%a b=%a got %a want %a\n", a, b, q2, ref); bad++; }
    }
    return bad;
}
// gcc -O2 -march=native -ffp-contract=off check_const_div.c -lm && ./a.out   (expects "bad 0" twice)
int main(void) {
    printf("b=6: bad %ld\n", test(6.0, 400000000L, -60, 60));
    printf("b=1e-4: bad %ld\n", test(0.0001, 400000000L, -60, 60));
    return 0;
}
