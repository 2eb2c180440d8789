// Host-side check of the division sequences of csrc/cg_math.cuh (the functions themselves, compiled
// as plain C++): div_const (Markstein, constant divisor) and ddiv_pos (Newton + two residual
// corrections) must return the correctly rounded quotient, i.e. exactly IEEE a / b.
// Usage: div_check [millions of random operand pairs]   -> prints "<bad div_const> <bad ddiv_pos>".
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include "../research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200/csrc/cg_math.cuh"

static uint64_t s = 88172645463325252ULL;
static inline uint64_t rnd() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static inline double mk(uint64_t mant, int e, bool neg) {
    uint64_t bits = (mant & 0x000fffffffffffffULL) | ((uint64_t)(1023 + e) << 52) | ((uint64_t)neg << 63);
    double v;
    memcpy(&v, &bits, 8);
    return v;
}

int main(int argc, char** argv) {
    const long M = (argc > 1 ? atol(argv[1]) : 4) * 1000000L;
    long bad_c = 0, bad_p = 0;
    const double cs[] = {1000.0, 45.0, 60.0, 90.0, 30.0, 300.0, 22.5, 7.5};
    for (int k = 0; k < 8; ++k) {
        const double c = cs[k], rc = 1.0 / c;
        for (long i = -300000; i <= 4000000; ++i) { const double a = (double)i; if (cg::div_const(a, c, rc) != a / c) ++bad_c; }
        for (long i = 0; i < M; ++i) {
            const uint64_t r = rnd();
            const double a = mk(r, (int)((r >> 52) % 120) - 60, r & 1);
            if (cg::div_const(a, c, rc) != a / c) ++bad_c;
        }
        const double sp[] = {0.0, -0.0, 1e-320, 1e308, -1e308, INFINITY, -INFINITY, NAN, 5e-324, 1e-300};
        for (double a : sp) { const double x = cg::div_const(a, c, rc), y = a / c; if (!(x == y || (x != x && y != y))) ++bad_c; }
    }
    // ddiv_pos: random operands over wide exponent ranges, integer-valued operands (counters), and the
    // classic hard cases for reciprocal-based division (divisor significands of all ones / just above a
    // power of two, quotients next to a rounding boundary)
    for (long i = 0; i < 8 * M; ++i) {
        const uint64_t r1 = rnd(), r2 = rnd();
        double a, b;
        switch (i & 7) {
            case 0: a = mk(r1, (int)((r1 >> 52) % 200) - 100, r1 & 1); b = mk(r2, (int)((r2 >> 52) % 120) - 60, false); break;
            case 1: a = (double)(r1 % 100000000ULL); b = (double)(1 + r2 % 100000000ULL); break;
            case 2: a = (double)(r1 % 4000000ULL) * 0.001; b = (double)(1 + r2 % 5000ULL); break;
            case 3: a = mk(r1, 3, false); b = mk(~0ULL << (r2 % 40), 0, false); break;                   // 1.111..1000
            case 4: a = mk(r1, 0, false); b = mk(r2 % 1024, 2, false); break;                           // 1.000..0xxx
            case 5: { b = mk(r2, (int)(r2 % 9), false); const double q = mk(r1, 1, false); a = q * b; } break;   // a/b near q
            case 6:
                if (r1 & 2) { a = mk(r1, -700 + (int)((r1 >> 52) % 1400), r1 & 1); b = mk(r2, -90 + (int)((r2 >> 52) % 180), false); }
                else {   // divisor significand close to 2 (n leading ones, then random), dividend a small integer or a power of two
                    const int n = 1 + (int)((r2 >> 8) % 52);
                    b = mk((~0ULL << (52 - n)) | (r2 & ((1ULL << (52 - n)) - 1) & ((r1 & 4) ? ~0ULL : 0ULL)), (int)(r2 % 12), false);
                    a = (r1 & 8) ? (double)(1 + (r1 >> 16) % 4096) : mk(0, (int)((r1 >> 16) % 24) - 4, false);
                }
                break;
            default: a = (double)(int64_t)(r1 % 2000001ULL) - 1000000.0; b = mk(r2, (int)(r2 % 30) - 10, false); break;
        }
        const double x = cg::ddiv_pos(a, b), y = a / b;
        if (!(x == y || (x != x && y != y))) { if (bad_p < 5) fprintf(stderr, "ddiv_pos(%a, %a) = %a, want %a\n", a, b, x, y); ++bad_p; }
    }
    const double sa[] = {0.0, -0.0, 1e-320, 1e308, -1e308, INFINITY, -INFINITY, NAN, 5e-324, 1.0, 3.0, 1e-250, 1e250};
    const double sb[] = {1e-320, 1e308, INFINITY, 5e-324, 1.0, 3.0, 1e-40, 1e40, 1e-200, 1e200, 0.1};
    for (double a : sa) for (double b : sb) { const double x = cg::ddiv_pos(a, b), y = a / b; if (!(x == y || (x != x && y != y))) ++bad_p; }
    printf("%ld %ld\n", bad_c, bad_p);
    return 0;
}
