// cg_math.cuh -- per-(series, tick) arithmetic of the canary path, device side.
//
// Everything here is IEEE fp64, one rounding per written operation (the library is
// compiled with -fmad=false), in the same operation order as SURVEY.md Appendix A
// (Prometheus 2.x extrapolatedRate / bucketQuantile) and as the reference's Python
// (/root/reference/mlflow_operator.py:387-404, :440-454), so results are meant to be
// bit-identical to the CPU restatement, not merely within 1e-6.
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>
#include "../../include/canary_gate.h"

// The per-lane arithmetic (this file and cg_warp.cuh's lane functions) also compiles as plain
// C++ so that tests/host_emul can run it against the oracle without a GPU.  That build is test
// infrastructure; the product is the sm_100a build.
#ifdef __CUDACC__
#define CG_DEVINL __device__ __forceinline__
#define CG_DEVNOINL static __device__ __noinline__
#define CG_DEVMEM __device__ __forceinline__
#else
#include <cmath>
#define CG_DEVINL static inline
#define CG_DEVNOINL static __attribute__((noinline))
#define CG_DEVMEM inline
using std::isinf;
using std::isnan;
static inline int __ffs(unsigned x) { return x ? __builtin_ctz(x) + 1 : 0; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) {
    sh &= 31u;
    return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
}
#endif

namespace cg {

CG_DEVINL int64_t floor_div(int64_t a, int64_t b) {
    int64_t q = a / b, r = a % b;
    return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
}
CG_DEVINL int64_t ceil_div(int64_t a, int64_t b) { return -floor_div(-a, b); }

// Window of a regular-grid series (sample i at phase + i*period) at one tick, plus the
// extrapolation terms that depend only on timestamps (Appendix A step 3 and 5).
struct TickGeom {
    int    i0;        // first sample index in the window
    int    n;         // samples in the window (< 2: the series yields no output)
    double sampled;   // (ts_last - ts_first)/1000
    double d_start;   // (ts_first - (t - r))/1000
    double avg;       // sampled/(n-1)
    double thr;       // avg*1.1
    double end_term;  // d_end < thr ? d_end : avg/2
};

CG_DEVINL TickGeom tick_geom(const cg_params& p, int32_t phase, int tick) {
    TickGeom g;
    const int64_t t  = (int64_t)p.tick0_ms + (int64_t)tick * p.tick_step_ms;
    const int64_t lo = t - p.window_ms;
    int64_t a = p.left_open ? floor_div(lo - phase, p.period_ms) + 1
                            : ceil_div(lo - phase, p.period_ms);
    int64_t b = floor_div(t - phase, p.period_ms);
    if (a < 0) a = 0;
    if (b > p.n_samples - 1) b = p.n_samples - 1;
    g.i0 = (int)a;
    g.n  = (int)(b - a + 1);
    if (g.n < 2) { g.n = g.n < 0 ? 0 : g.n; g.sampled = g.d_start = g.avg = g.thr = g.end_term = 0.0; return g; }
    const int64_t first_t = (int64_t)phase + a * p.period_ms;
    const int64_t last_t  = (int64_t)phase + b * p.period_ms;
    g.d_start = (double)(first_t - lo) / 1000.0;
    const double d_end = (double)(t - last_t) / 1000.0;
    g.sampled = (double)(last_t - first_t) / 1000.0;
    g.avg = g.sampled / (double)(g.n - 1);
    g.thr = g.avg * 1.1;
    g.end_term = (d_end < g.thr) ? d_end : g.avg / 2;
    return g;
}

// high word of a double (sign, exponent, top of the significand)
CG_DEVINL uint32_t hi_word(double v) {
#ifdef __CUDACC__
    return (uint32_t)__double2hiint(v);
#else
    uint64_t u;
    memcpy(&u, &v, 8);
    return (uint32_t)(u >> 32);
#endif
}
// biased exponent in [lo, lo + span): one integer subtract and compare (false for 0, denormals, Inf, NaN
// when lo >= 1 and lo + span <= 2047)
CG_DEVINL bool exp_in(double v, uint32_t lo, uint32_t span) { return ((hi_word(v) >> 20) & 0x7ffu) - lo < span; }

// Out-of-line IEEE division: one copy of the generic routine (with its special-case path) instead
// of one per use.
CG_DEVNOINL double ddiv_call(double a, double b) { return a / b; }

// a / c for a constant c with rc = 1/c correctly rounded: the Markstein sequence
// (q = a*rc; r = fma(-q, c, a); q + r*rc) gives the correctly rounded quotient, i.e. exactly what
// IEEE division gives, in three operations instead of the ~45-instruction division routine.  Checked
// against true division on 2e8 random doubles over 80 binades and every integer in [-3e5, 4e6]
// for each constant used here (tests/test_div_const.py keeps a smaller version of that check).
// c is within a few binades of 1 here (1000, the window and sample spans in seconds).  Zero goes
// straight through; magnitudes near the overflow / underflow ends, Inf and NaN take the true division.
CG_DEVINL double div_const(double a, double c, double rc) {
    const double q = a * rc;
    const double r = fma(-q, c, a);
    const double q2 = fma(r, rc, q);
    if (exp_in(a, 60u, 1960u)) return q2;          // 2^-963 <= |a| < 2^997
    if (a == 0.0) return q2;                        // +-0 / c = +-0 = q2
    return ddiv_call(a, c);
}

// a / b for b > 0, both in a safe range: float reciprocal seed, two Newton steps, Markstein
// quotient with two residual corrections.  The value before the last rounding is within about
// 2^-54 ulp of a/b, and a quotient of doubles stays at least 2^-53/m ulp away from a rounding
// boundary (m in [1,2) the divisor's significand), so the rounding can only go wrong for
// significands next to 2 (4 / 0x1.fffffffffffffp+4 does): divisors whose top 20 fraction bits are
// all ones take the true division, as does anything outside the exponent range.
// tests/test_div_const.py checks it against IEEE division on random and adversarial operands.
CG_DEVINL double ddiv_pos(double a, double b) {
    if (exp_in(b, 923u, 200u) && exp_in(a, 223u, 1600u) &&     // 2^-100 <= b < 2^100, 2^-800 <= |a| < 2^800
        (hi_word(b) & 0xfffffu) != 0xfffffu) {
#ifdef __CUDACC__
        double y = (double)__frcp_rn((float)b);
#else
        double y = (double)(1.0f / (float)b);
#endif
        double e = fma(-b, y, 1.0);
        y = fma(y, e, y);
        e = fma(-b, y, 1.0);
        y = fma(y, e, y);
        double q = a * y;
        double r = fma(-b, q, a);
        q = fma(r, y, q);
        r = fma(-b, q, a);
        return fma(r, y, q);
    }
    if (a == 0.0 && exp_in(b, 1u, 2046u)) return a;            // +-0 / (finite b > 0) = +-0
    return ddiv_call(a, b);
}

// Same as tick_geom with 32-bit integer arithmetic, for callers that have checked that every
// time stamp involved fits in 31 bits (the streaming kernel's eligibility test).
CG_DEVINL int floor_div32(int a, int b) {   // b > 0
    int q = a / b;
    return (a % b != 0 && a < 0) ? q - 1 : q;
}
CG_DEVNOINL int floor_div32_call(int a, int b) { return floor_div32(a, b); }   // one copy where speed does not matter
CG_DEVINL TickGeom tick_geom32(const cg_params& p, int32_t phase, int tick) {
    TickGeom g;
    const int t  = p.tick0_ms + tick * p.tick_step_ms;
    const int lo = t - p.window_ms;
    int a = p.left_open ? floor_div32(lo - phase, p.period_ms) + 1 : -floor_div32(phase - lo, p.period_ms);
    int b = floor_div32(t - phase, p.period_ms);
    if (a < 0) a = 0;
    if (b > p.n_samples - 1) b = p.n_samples - 1;
    g.i0 = a;
    g.n  = b - a + 1;
    if (g.n < 2) { g.n = g.n < 0 ? 0 : g.n; g.sampled = g.d_start = g.avg = g.thr = g.end_term = 0.0; return g; }
    const int first_t = phase + a * p.period_ms;
    const int last_t  = phase + b * p.period_ms;
    g.d_start = (double)(first_t - lo) / 1000.0;
    const double d_end = (double)(t - last_t) / 1000.0;
    g.sampled = (double)(last_t - first_t) / 1000.0;
    g.avg = g.sampled / (double)(g.n - 1);
    g.thr = g.avg * 1.1;
    g.end_term = (d_end < g.thr) ? d_end : g.avg / 2;
    return g;
}

// Appendix A steps 4-6 for one series once `result` (step 2) is known.
CG_DEVINL double extrapolate(double result, double first_v, const TickGeom& g) {
    double d_start = g.d_start;
    if (result > 0 && first_v >= 0) {
        const double d_zero = g.sampled * (first_v / result);
        if (d_zero < d_start) d_start = d_zero;
    }
    double ext = g.sampled;
    ext += (d_start < g.thr) ? d_start : g.avg / 2;
    ext += g.end_term;
    return result * (ext / g.sampled);
}

// Appendix A step 2 over a window held in memory (`row` points at sample i0).
template <typename T>
CG_DEVINL double window_result(const T* __restrict__ row, int n, double& first_v) {
    first_v = (double)row[0];
    double result = (double)row[n - 1] - first_v;
    double prev = first_v;
    for (int i = 1; i < n; ++i) {
        const double v = (double)row[i];
        if (v < prev) result += prev;      // counter reset
        prev = v;
    }
    return result;
}

template <typename T>
CG_DEVINL double series_increase(const T* __restrict__ row, const TickGeom& g) {
    double first_v;
    const double result = window_result(row + g.i0, g.n, first_v);
    return extrapolate(result, first_v, g);
}

// ---- should_promote_model (mlflow_operator.py:419-460) -------------------------
// nm/om: [q_gate, error_rate, latency_avg] of the new / old version.
CG_DEVINL uint8_t gate_eval(const double nm[3], uint8_t nv, const double om[3], uint8_t ov,
                            const cg_params& p) {
    const bool need_q = p.checks & CG_G_P95_OK, need_e = p.checks & CG_G_ERR_OK,
               need_a = p.checks & CG_G_AVG_OK;
    bool avail = true;                                               // :430-434
    if (need_q && !((nv & CG_V_QUANTILE) && (ov & CG_V_QUANTILE))) avail = false;
    if (need_e && !((nv & CG_V_ERROR_RATE) && (ov & CG_V_ERROR_RATE))) avail = false;
    if (need_a && !((nv & CG_V_LATENCY_AVG) && (ov & CG_V_LATENCY_AVG))) avail = false;
    if (!avail) return 0;
    uint8_t d = CG_G_AVAILABLE;
    bool promote = true;
    if (need_q) { if (nm[0] <= om[0] * p.one_plus_thr[0]) d |= CG_G_P95_OK; else promote = false; }  // :440
    if (need_e) { if (nm[1] <= om[1] * p.one_plus_thr[1]) d |= CG_G_ERR_OK; else promote = false; }  // :447
    if (need_a) { if (nm[2] <= om[2] * p.one_plus_thr[2]) d |= CG_G_AVG_OK; else promote = false; }  // :454
    if (promote) d |= CG_G_PROMOTE;                                  // :460
    return d;
}

// ---- the traffic step loop (mlflow_operator.py:296-361) ------------------------
template <typename BitFn>
CG_DEVINL cg_decision replay_schedule(BitFn gate_bit, bool has_previous, int start_tick,
                                      const cg_params& p) {
    cg_decision d;
    if (!has_previous) {                                             // :188-191, :289
        d.traffic_current = p.max_traffic; d.traffic_prev = 0; d.state = CG_STATE_NO_CANARY;
        d.n_promotions = 0; d.n_evals = 0; d.last_tick = -1;
        return d;
    }
    int cur = p.init_current, prev = p.max_traffic - p.init_current; // :186-187
    int tick = start_tick, n_prom = 0, n_eval = 0, last = -1;
    int state = CG_STATE_IN_PROGRESS;
    bool out_of_data = false;
    while (cur < p.max_traffic && !out_of_data) {                    // :296
        int attempt = 0;
        bool success = false;
        while (attempt < p.max_attempts) {                           // :300
            if (tick >= p.n_ticks || tick < 0) { out_of_data = true; break; }
            ++n_eval; last = tick;
            if (gate_bit(tick)) {                                    // :310
                cur += p.step; prev -= p.step;                       // :312-313
                if (cur > p.max_traffic) cur = p.max_traffic;        // :316
                if (prev < 0) prev = 0;                              // :317
                ++n_prom; success = true;
                break;
            }
            ++attempt;                                               // :337
            if (attempt < p.max_attempts) tick += p.attempt_delay_ticks;   // :340
        }
        if (out_of_data) break;
        if (!success) { state = CG_STATE_ABORTED; break; }           // :347-349
        tick += p.interval_ticks;                                    // :352
    }
    if (cur == p.max_traffic) state = CG_STATE_COMPLETE;             // :355
    if (state == CG_STATE_ABORTED && p.on_abort == CG_ABORT_ROLLBACK) {   // opt-in, not in the reference (:345)
        cur = 0; prev = p.max_traffic; state = CG_STATE_ROLLED_BACK;
    }
    d.traffic_current = cur; d.traffic_prev = prev; d.state = state;
    d.n_promotions = n_prom; d.n_evals = n_eval; d.last_tick = last;
    return d;
}

// The same loop over a packed bit-vector (bit j%32 of word j/32 = promote at tick j).  When
// retries are one tick apart (attempt_delay = tick step, the reference's 10 s / 10 s) a whole
// inner `while attempt < max_attempts` loop (:300-345) collapses to one find-first-set over the
// next max_attempts bits, so a canary costs ~9 short steps instead of up to 90 dependent reads.
CG_DEVINL cg_decision replay_schedule_bits(const uint32_t* bits, int n_words, bool has_previous,
                                           int start_tick, const cg_params& p) {
    if (!has_previous || p.attempt_delay_ticks != 1 || p.max_attempts > 32 || start_tick < 0)
        return replay_schedule([&](int t) { return (bits[t >> 5] >> (t & 31)) & 1u; }, has_previous,
                               start_tick, p);
    int cur = p.init_current, prev = p.max_traffic - p.init_current;   // :186-187
    int tick = start_tick, n_prom = 0, n_eval = 0, last = -1;
    int state = CG_STATE_IN_PROGRESS;
    while (cur < p.max_traffic) {                                      // :296
        const int avail = p.n_ticks - tick;
        if (avail <= 0) break;                                         // ran out of data
        const int n_try = avail < p.max_attempts ? avail : p.max_attempts;
        const int w = tick >> 5;
        const uint32_t lo = bits[w], hi = (w + 1 < n_words) ? bits[w + 1] : 0u;
        uint32_t win = __funnelshift_r(lo, hi, tick & 31);
        if (n_try < 32) win &= (1u << n_try) - 1u;
        if (win) {                                                     // :310 succeeds at attempt k
            const int k = __ffs(win) - 1;
            n_eval += k + 1; last = tick + k;
            cur += p.step; prev -= p.step;                             // :312-313
            if (cur > p.max_traffic) cur = p.max_traffic;              // :316
            if (prev < 0) prev = 0;                                    // :317
            ++n_prom;
            tick += k + p.interval_ticks;                              // :352
        } else {
            n_eval += n_try; last = tick + n_try - 1;
            if (n_try == p.max_attempts) state = CG_STATE_ABORTED;     // :341-349
            break;                                                     // else: data ended mid-retry
        }
    }
    if (cur == p.max_traffic) state = CG_STATE_COMPLETE;               // :355
    if (state == CG_STATE_ABORTED && p.on_abort == CG_ABORT_ROLLBACK) {   // opt-in, not in the reference (:345)
        cur = 0; prev = p.max_traffic; state = CG_STATE_ROLLED_BACK;
    }
    cg_decision d;
    d.traffic_current = cur; d.traffic_prev = prev; d.state = state;
    d.n_promotions = n_prom; d.n_evals = n_eval; d.last_tick = last;
    return d;
}

}  // namespace cg
