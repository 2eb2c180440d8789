/* Host-side check of the constant-division sequence used by csrc/cg_math.cuh (div_const):
 * q = a*rc; r = fma(-q, c, a); q + r*rc  must equal IEEE a/c.  Prints the mismatch count. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
static double mdiv(double a, double c, double rc) { double q = a * rc; double r = fma(-q, c, a); return fma(r, rc, q); }
int main(void) {
    const double cs[] = {1000.0, 45.0, 60.0, 90.0, 30.0, 300.0, 22.5, 7.5};
    long bad = 0;
    for (int k = 0; k < 8; ++k) {
        const double c = cs[k], rc = 1.0 / c;
        for (long i = -300000; i <= 4000000; ++i) { double a = (double)i; if (mdiv(a, c, rc) != a / c) ++bad; }
        uint64_t s = 88172645463325252ULL + k;
        for (long i = 0; i < 4000000; ++i) {
            s ^= s << 13; s ^= s >> 7; s ^= s << 17;
            uint64_t bits = (s & 0x000fffffffffffffULL) | ((uint64_t)(1023 - 60 + (s >> 52) % 120) << 52);
            double a; *(uint64_t*)&a = bits; if (s & 1) a = -a;
            if (mdiv(a, c, rc) != a / c) ++bad;
        }
    }
    printf("%ld\n", bad);
    return 0;
}
