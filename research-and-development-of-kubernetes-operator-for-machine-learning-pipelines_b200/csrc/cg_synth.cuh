// cg_synth.cuh -- on-device generator of synthetic Prometheus-style series (benchmark and
// test utility; the reference has no counterpart).  Statistical model of SURVEY.md 8(d):
// per-scrape requests ~ Poisson(lambda) (normal approximation), errors ~ Binomial(req,
// p_err), log-normal latencies binned into cumulative `le` buckets, random counter
// offsets, per-series counter resets, a share of adversarial pairs.  Every value is a
// pure function of (seed, global pair index, version, series, sample), so shards
// generated on different GPUs or in different chunks are consistent.
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/canary_gate.h"

namespace cg {

struct SynthArgs {
    void*     scalars;
    void*     buckets;
    const double* le;
    int32_t*  phase;
    int64_t   n_pairs;
    int64_t   pair_offset;      // global index of local pair 0
    uint64_t  seed;
    double    reset_prob;       // per series per scrape
    double    adversarial;      // share of adversarial pairs
    int32_t   phase_avoid;      // gcd(period, tick step): phases never land on a multiple
    cg_params p;
};

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__device__ __forceinline__ uint64_t key(uint64_t seed, uint64_t a, uint64_t b, uint64_t c) {
    return mix64(mix64(mix64(seed ^ (a * 0xD6E8FEB86659FD93ull)) + b) + c * 0xA24BAED4963EE407ull);
}
__device__ __forceinline__ double u01(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }
__device__ __forceinline__ double gauss(uint64_t h) {
    const double u1 = u01(h) + 1e-300, u2 = u01(mix64(h));
    return sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}

struct PairModel {
    double lam, p_err[2], mu[2];
    int adv_kind;   // -1 none; 0 zero traffic; 1 empty histogram; 2 non-monotonic buckets; 3 old er = 0, new > 0
    int bad_bucket;
};

__device__ inline PairModel pair_model(const SynthArgs& a, int64_t gp) {
    PairModel m;
    const uint64_t s = a.seed;
    m.lam = exp(log(5.0) + u01(key(s, gp, 1, 0)) * (log(500.0) - log(5.0)));
    const double pe_old = u01(key(s, gp, 2, 0)) < 0.3 ? 0.0
                        : exp(log(1e-4) + u01(key(s, gp, 3, 0)) * (log(5e-2) - log(1e-4)));
    m.p_err[1] = pe_old;
    m.p_err[0] = fmin(pe_old * exp(0.15 * gauss(key(s, gp, 4, 0))), 1.0);
    m.mu[1] = log(0.01) + u01(key(s, gp, 5, 0)) * (log(0.5) - log(0.01));
    m.mu[0] = m.mu[1] + 0.04 * gauss(key(s, gp, 6, 0));
    m.adv_kind = u01(key(s, gp, 7, 0)) < a.adversarial ? (int)(key(s, gp, 8, 0) & 3) : -1;
    m.bad_bucket = a.p.n_buckets > 2 ? 1 + (int)(key(s, gp, 9, 0) % (uint64_t)(a.p.n_buckets - 2)) : 0;
    if (m.adv_kind == 3) { m.p_err[1] = 0.0; m.p_err[0] = fmax(m.p_err[0], 0.02); }
    return m;
}

// requests of (pair, version) at scrape i and that scrape's latency location
__device__ inline void scrape_model(const SynthArgs& a, const PairModel& m, int64_t gp, int ver, int i,
                                    double& req, double& mu_t) {
    const uint64_t h = key(a.seed, gp, 100 + ver, i);
    req = floor(fmax(m.lam + sqrt(m.lam) * gauss(h), 0.0) + 0.5);
    if (m.adv_kind == 0) req = 0.0;
    mu_t = m.mu[ver] + 0.02 * gauss(mix64(h ^ 0x5bd1e995ull));
}

template <typename T>
__global__ void synth_kernel(const __grid_constant__ SynthArgs a) {
    const cg_params& p = a.p;
    const int S = p.n_scalar + p.n_buckets;
    const int64_t total = a.n_pairs * 2 * S;
    const double sigma = 0.5;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int s = (int)(idx % S);
        const int64_t pv = idx / S;
        const int ver = (int)(pv & 1);
        const int64_t lp = pv >> 1, gp = lp + a.pair_offset;
        const PairModel m = pair_model(a, gp);
        if (s == 0) {
            int ph = 1 + (int)(key(a.seed, gp, 10 + ver, 0) % (uint64_t)(p.period_ms > 2 ? p.period_ms - 2 : 1));
            if (a.phase_avoid > 1 && ph % a.phase_avoid == 0) ph += 1;   // keep samples off tick boundaries
            a.phase[pv] = ph;
        }
        T* row = s < p.n_scalar
                     ? reinterpret_cast<T*>(a.scalars) + ((size_t)pv * p.n_scalar + s) * p.row_pitch
                     : reinterpret_cast<T*>(a.buckets) + ((size_t)pv * p.n_buckets + (s - p.n_scalar)) * p.row_pitch;
        double run = floor(u01(key(a.seed, gp, 200 + ver, s)) * 1e6);
        const int b = s - p.n_scalar;
        const double le_b = b >= 0 ? a.le[b] : 0.0;
        for (int i = 0; i < p.n_samples; ++i) {
            double req, mu_t;
            scrape_model(a, m, gp, ver, i, req, mu_t);
            const uint64_t hs = key(a.seed ^ 0xABCDEFull, (uint64_t)(gp * 2 + ver), s, i);
            double inc;
            if (s < p.n_scalar) {
                const unsigned role = p.scalar_role[s];
                const double mean = req * m.p_err[ver];
                double err = floor(mean + sqrt(mean * (1.0 - m.p_err[ver])) * gauss(key(a.seed, gp, 300 + ver, i)) +
                                   u01(key(a.seed, gp, 400 + ver, i)));
                err = fmin(fmax(err, 0.0), req);
                if (role & CG_ROLE_FEEDBACK) inc = floor(0.05 * req * u01(hs) * 2.0);
                else if (role & CG_ROLE_SERVER_ERROR) inc = err;
                else if (role & CG_ROLE_SERVER_TOTAL) inc = req - err;
                else if (role & CG_ROLE_CLIENT_SUM) inc = req * exp(mu_t + 0.5 * sigma * sigma) * (0.95 + 0.1 * u01(hs));
                else inc = req;                               // client count
            } else {
                double cum;
                if (isinf(le_b)) cum = req;
                else cum = fmin(floor(req * normcdf((log(le_b) - mu_t) / sigma) + 0.5), req);
                if (m.adv_kind == 1) cum = 0.0;
                inc = cum;
            }
            if (i > 0 && u01(mix64(hs)) < a.reset_prob) run = 0.0;     // counter reset
            run += inc;
            double v = run;
            if (b >= 0 && m.adv_kind == 2 && b == m.bad_bucket) v = run * 0.5;
            row[i] = (T)v;
        }
        for (int i = p.n_samples; i < p.row_pitch; ++i) row[i] = (T)0;
    }
}

}  // namespace cg
