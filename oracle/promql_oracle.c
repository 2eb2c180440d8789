/*
 * promql_oracle.c -- CPU restatement (plain C, fp64) of the canary metric-validation
 * path, batched over the same SoA layout the CUDA library uses.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product never does.
 *
 * What it follows:
 *   - the reference's own Python logic, /root/reference/mlflow_operator.py:
 *       get_model_metrics defaults + derived divisions   :369-415
 *       should_promote_model                             :419-460
 *       thresholds / initial weights / schedule constants :175-191, :290-294
 *       the inline step loop                             :296-361
 *   - for rate()/increase()/sum/histogram_quantile, which the reference delegates to
 *     an external, UNPINNED Prometheus server (:47, :368-411; upstream
 *     github.com/prometheus/prometheus promql/functions.go extrapolatedRate and
 *     promql/quantile.go bucketQuantile, not vendored): SURVEY.md Appendix A.
 *     PARITY UNPINNED at that boundary: the reference holds no test or golden vector
 *     for it.  Pinned here by SURVEY.md Appendix B known-answer vectors
 *     (tests/golden/kat_promql.json) and by agreement with oracle/pyoracle.py, which
 *     the unmodified reference runs on top of in oracle/ref_loader.py.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -pthread).  No FMA
 * contraction and no fast-math: every operation is one IEEE fp64 rounding, in the
 * order written, so results are reproducible bit for bit.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include "../include/canary_gate.h"

#include <pthread.h>
#include <unistd.h>

/* ---- a minimal static-partition parallel-for over pthreads (no OpenMP in this image) */
typedef void (*cgo_range_fn)(int64_t begin, int64_t end, void* ctx);
typedef struct { cgo_range_fn fn; void* ctx; int64_t begin, end; } cgo_job;
static void* cgo_job_main(void* arg) { cgo_job* j = (cgo_job*)arg; j->fn(j->begin, j->end, j->ctx); return 0; }
static void parallel_for(int64_t n, int n_threads, cgo_range_fn fn, void* ctx)
{
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    if ((int64_t)n_threads > n) n_threads = n > 0 ? (int)n : 1;
    if (n_threads == 1) { fn(0, n, ctx); return; }
    pthread_t th[256];
    cgo_job jobs[256];
    for (int i = 0; i < n_threads; ++i) {
        jobs[i].fn = fn; jobs[i].ctx = ctx;
        jobs[i].begin = n * i / n_threads; jobs[i].end = n * (i + 1) / n_threads;
        if (i + 1 < n_threads) pthread_create(&th[i], 0, cgo_job_main, &jobs[i]);
    }
    cgo_job_main(&jobs[n_threads - 1]);
    for (int i = 0; i + 1 < n_threads; ++i) pthread_join(th[i], 0);
}

/* ---- single-series primitives (also exported for the KAT tests) ------------ */

/* SURVEY.md Appendix A steps 2-6 given the samples already selected.
 * `first_t`/`last_t` are the timestamps (ms) of v[0] and v[n-1]. Returns 0 when the
 * series yields no output (n < 2), else 1 and *increase. */
static int increase_core(const double* v, int n, int64_t first_t, int64_t last_t,
                         int64_t t_ms, int64_t range_ms, double* increase)
{
    if (n < 2) return 0;                                   /* step 1 */
    double first_v = v[0];
    double result = v[n - 1] - first_v;                    /* step 2 */
    double prev = first_v;
    for (int i = 1; i < n; ++i) {
        if (v[i] < prev) result += prev;                   /* counter reset */
        prev = v[i];
    }
    int64_t range_start = t_ms - range_ms;
    double d_start = (double)(first_t - range_start) / 1000.0;   /* step 3 */
    double d_end   = (double)(t_ms - last_t) / 1000.0;
    double sampled = (double)(last_t - first_t) / 1000.0;
    double avg = sampled / (double)(n - 1);
    if (result > 0 && first_v >= 0) {                      /* step 4 */
        double d_zero = sampled * (first_v / result);
        if (d_zero < d_start) d_start = d_zero;
    }
    double thr = avg * 1.1;                                /* step 5 */
    double ext = sampled;
    ext += (d_start < thr) ? d_start : avg / 2;
    ext += (d_end < thr) ? d_end : avg / 2;
    *increase = result * (ext / sampled);                  /* step 6 */
    return 1;
}

/* Arbitrary timestamps; selects the window itself. */
int cgo_extrapolated_increase(const int64_t* ts, const double* v, int n_all,
                              int64_t t_ms, int64_t range_ms, int left_open,
                              double* increase)
{
    int64_t lo = t_ms - range_ms;
    int a = 0, b = n_all - 1;
    while (a < n_all && (left_open ? ts[a] <= lo : ts[a] < lo)) ++a;
    while (b >= 0 && ts[b] > t_ms) --b;
    int n = b - a + 1;
    if (n < 2) return 0;
    return increase_core(v + a, n, ts[a], ts[b], t_ms, range_ms, increase);
}

/* SURVEY.md Appendix A histogram_quantile steps 2-7; `cnt` is modified (monotonic
 * fix).  le must be sorted ascending. */
double cgo_bucket_quantile(double q, const double* le, double* cnt, int n)
{
    if (isnan(q)) return NAN;
    if (q < 0) return -INFINITY;
    if (q > 1) return INFINITY;
    if (n < 1 || !(isinf(le[n - 1]) && le[n - 1] > 0)) return NAN;   /* step 2 */
    if (n < 2) return NAN;
    double mx = cnt[0];                                    /* step 3 */
    for (int i = 1; i < n; ++i) {
        if (cnt[i] > mx) mx = cnt[i];
        else if (cnt[i] < mx) cnt[i] = mx;
    }
    double obs = cnt[n - 1];
    if (obs == 0) return NAN;                              /* step 4 */
    double rank = q * obs;                                 /* step 5 */
    int b = n - 1;
    for (int i = 0; i < n - 1; ++i)
        if (cnt[i] >= rank) { b = i; break; }
    if (b == n - 1) return le[n - 2];                      /* step 6 */
    if (b == 0 && le[0] <= 0) return le[0];
    double start = 0.0, end = le[b], c = cnt[b];           /* step 7 */
    if (b > 0) {
        start = le[b - 1];
        c -= cnt[b - 1];
        rank -= cnt[b - 1];
    }
    return start + (end - start) * (rank / c);
}

/* ---- helpers --------------------------------------------------------------- */
static int64_t floor_div(int64_t a, int64_t b) { int64_t q = a / b, r = a % b; return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q; }
static int64_t ceil_div(int64_t a, int64_t b) { return -floor_div(-a, b); }

static inline double load_elem(const void* base, int dtype, int64_t idx)
{
    return dtype == CG_F32 ? (double)((const float*)base)[idx] : ((const double*)base)[idx];
}

#define CGO_MAX_WINDOW 64

/* window of one regular-grid series at tick time t: indices [i0, i1] */
static void window_bounds(const cg_params* p, int64_t phase, int64_t t, int* i0, int* i1)
{
    int64_t lo = t - p->window_ms;
    int64_t a = p->left_open ? floor_div(lo - phase, p->period_ms) + 1
                             : ceil_div(lo - phase, p->period_ms);
    int64_t b = floor_div(t - phase, p->period_ms);
    if (a < 0) a = 0;
    if (b > p->n_samples - 1) b = p->n_samples - 1;
    *i0 = (int)a; *i1 = (int)b;
}

int cgo_metric_width(const cg_params* p) { return p->n_q + CG_M_EXTRA; }

/* ---- get_model_metrics on the tick grid (mlflow_operator.py:363-417) -------- */
static void eval_pv_tick(const void* sc, const void* bk, const double* le, int64_t phase,
                         const cg_params* p, int tick, double* m, uint8_t* valid_out)
{
    const int W = cgo_metric_width(p);
    const int nq = p->n_q;
    int64_t t = (int64_t)p->tick0_ms + (int64_t)tick * p->tick_step_ms;
    int i0, i1;
    window_bounds(p, phase, t, &i0, &i1);
    int n = i1 - i0 + 1;
    int64_t first_t = phase + (int64_t)i0 * p->period_ms;
    int64_t last_t = phase + (int64_t)i1 * p->period_ms;
    double win[CGO_MAX_WINDOW];
    uint8_t valid = 0;

    /* Q2..Q6: sum(increase(...)) or on() vector(0)  (:375, :383, :393, :397, :410) */
    double sums[5] = {0, 0, 0, 0, 0};
    int have[5] = {0, 0, 0, 0, 0};
    if (n >= 2) {
        for (int k = 0; k < p->n_scalar; ++k) {
            for (int i = 0; i < n; ++i)
                win[i] = load_elem(sc, p->dtype, (int64_t)k * p->row_pitch + i0 + i);
            double inc;
            if (!increase_core(win, n, first_t, last_t, t, p->window_ms, &inc)) continue;
            for (int r = 0; r < 5; ++r)
                if (p->scalar_role[k] & (1u << r)) {
                    sums[r] = have[r] ? sums[r] + inc : inc;
                    have[r] = 1;
                }
        }
    }
    double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];

    /* Q1: histogram_quantile(q, sum(rate(..._bucket[r])) by (le))  (:367) */
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    if (p->n_buckets > 0 && n >= 2) {
        double cnt0[1024], cnt[1024];
        double range_s = (double)p->window_ms / 1000.0;
        for (int b = 0; b < p->n_buckets; ++b) {
            for (int i = 0; i < n; ++i)
                win[i] = load_elem(bk, p->dtype, (int64_t)b * p->row_pitch + i0 + i);
            double inc = 0;
            increase_core(win, n, first_t, last_t, t, p->window_ms, &inc);
            cnt0[b] = inc / range_s;                        /* rate, Appendix A step 6 */
        }
        for (int j = 0; j < nq; ++j) {
            memcpy(cnt, cnt0, sizeof(double) * p->n_buckets);
            m[j] = cgo_bucket_quantile(p->q[j], le, cnt, p->n_buckets);
        }
        valid |= CG_V_QUANTILE;                             /* non-empty vector (:369-370) */
    }
    m[nq + CG_M_ERROR_RESPONSES] = err;                     /* :378 */
    if (total > 0) {                                        /* :387-390 */
        m[nq + CG_M_ERROR_RATE] = err / total;
        valid |= CG_V_ERROR_RATE;
    } else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) {                                      /* :401-404 */
        m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt;
        valid |= CG_V_LATENCY_AVG;
    } else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;                   /* :407 */
    m[nq + CG_M_FEEDBACK_COUNT] = fb;                       /* :413 */
    (void)W;
    *valid_out = valid;
}

/* ---- should_promote_model (mlflow_operator.py:419-460) ---------------------- */
static uint8_t gate_one(const double* nm, uint8_t nv, const double* om, uint8_t ov,
                        const cg_params* p)
{
    const int nq = p->n_q;
    uint8_t d = 0;
    int need_q = (p->checks & CG_G_P95_OK) != 0, need_e = (p->checks & CG_G_ERR_OK) != 0,
        need_a = (p->checks & CG_G_AVG_OK) != 0;
    int avail = 1;                                          /* :430-434 */
    if (need_q && !((nv & CG_V_QUANTILE) && (ov & CG_V_QUANTILE))) avail = 0;
    if (need_e && !((nv & CG_V_ERROR_RATE) && (ov & CG_V_ERROR_RATE))) avail = 0;
    if (need_a && !((nv & CG_V_LATENCY_AVG) && (ov & CG_V_LATENCY_AVG))) avail = 0;
    if (!avail) return 0;
    d |= CG_G_AVAILABLE;
    int promote = 1;
    if (need_q) {                                           /* :440 */
        if (nm[p->gate_q] <= om[p->gate_q] * p->one_plus_thr[0]) d |= CG_G_P95_OK; else promote = 0;
    }
    if (need_e) {                                           /* :447 */
        if (nm[nq + CG_M_ERROR_RATE] <= om[nq + CG_M_ERROR_RATE] * p->one_plus_thr[1]) d |= CG_G_ERR_OK; else promote = 0;
    }
    if (need_a) {                                           /* :454 */
        if (nm[nq + CG_M_LATENCY_AVG] <= om[nq + CG_M_LATENCY_AVG] * p->one_plus_thr[2]) d |= CG_G_AVG_OK; else promote = 0;
    }
    if (promote) d |= CG_G_PROMOTE;                         /* :460 */
    return d;
}

/* ---- the step loop (mlflow_operator.py:296-361) ----------------------------- */
static void replay_one(const uint32_t* bits, int has_previous, int start_tick,
                       const cg_params* p, cg_decision* out)
{
    if (!has_previous) {                                    /* :188-191, :289 */
        out->traffic_current = p->max_traffic; out->traffic_prev = 0;
        out->state = CG_STATE_NO_CANARY; out->n_promotions = 0; out->n_evals = 0;
        out->last_tick = -1;
        return;
    }
    int cur = p->init_current, prev = p->max_traffic - p->init_current;  /* :186-187 */
    int tick = start_tick, n_prom = 0, n_eval = 0, last = -1;
    int state = CG_STATE_IN_PROGRESS, out_of_data = 0;
    while (cur < p->max_traffic && !out_of_data) {          /* :296 */
        int attempt = 0, success = 0;
        while (attempt < p->max_attempts) {                 /* :300 */
            if (tick >= p->n_ticks || tick < 0) { out_of_data = 1; break; }
            ++n_eval; last = tick;
            if ((bits[tick >> 5] >> (tick & 31)) & 1u) {    /* :310 */
                cur += p->step; prev -= p->step;            /* :312-313 */
                if (cur > p->max_traffic) cur = p->max_traffic;     /* :316 */
                if (prev < 0) prev = 0;                     /* :317 */
                ++n_prom; success = 1;
                break;
            }
            ++attempt;                                      /* :337 */
            if (attempt < p->max_attempts) tick += p->attempt_delay_ticks;  /* :340 */
        }
        if (out_of_data) break;
        if (!success) { state = CG_STATE_ABORTED; break; }  /* :347-349 */
        tick += p->interval_ticks;                          /* :352 */
    }
    if (cur == p->max_traffic) state = CG_STATE_COMPLETE;   /* :355 */
    if (state == CG_STATE_ABORTED && p->on_abort == CG_ABORT_ROLLBACK) {   /* opt-in; the reference has a placeholder at :345 */
        cur = 0; prev = p->max_traffic; state = CG_STATE_ROLLED_BACK;
    }
    out->traffic_current = cur; out->traffic_prev = prev; out->state = state;
    out->n_promotions = n_prom; out->n_evals = n_eval; out->last_tick = last;
}

/* ---- batched entry points (HOST pointers, same layout as the device library) - */
static int params_ok(const cg_params* p)
{
    if (!p || p->struct_bytes != (int32_t)sizeof(cg_params)) return 0;
    if (p->n_buckets > 1024 || p->n_scalar > CG_MAX_SCALAR || p->n_q > CG_MAX_Q) return 0;
    if (p->period_ms <= 0 || p->window_ms <= 0 || p->tick_step_ms <= 0) return 0;
    if ((int64_t)p->window_ms / p->period_ms + 2 > CGO_MAX_WINDOW) return 0;
    return 1;
}

typedef struct {
    const void* scalars; const void* buckets; const double* le; const int32_t* phase;
    const uint8_t* has_previous; const int32_t* start_tick; const cg_params* p;
    cg_decision* decisions; uint32_t* gate_bits; double* metrics; uint8_t* valid;
} cgo_ctx;

static void grid_range(int64_t begin, int64_t end, void* ctx)
{
    const cgo_ctx* c = (const cgo_ctx*)ctx;
    const cg_params* p = c->p;
    const int W = cgo_metric_width(p);
    const size_t esz = p->dtype == CG_F32 ? 4 : 8;
    for (int64_t pv = begin; pv < end; ++pv) {
        const char* sc = (const char*)c->scalars + (size_t)pv * p->n_scalar * p->row_pitch * esz;
        const char* bk = c->buckets ? (const char*)c->buckets + (size_t)pv * p->n_buckets * p->row_pitch * esz : 0;
        for (int j = 0; j < p->n_ticks; ++j) {
            double m[CG_MAX_Q + CG_M_EXTRA];
            uint8_t v;
            eval_pv_tick(sc, bk, c->le, c->phase[pv], p, j, m, &v);
            if (c->metrics) memcpy(c->metrics + ((size_t)pv * p->n_ticks + j) * W, m, sizeof(double) * W);
            if (c->valid) c->valid[(size_t)pv * p->n_ticks + j] = v;
        }
    }
}

int cgo_eval_grid(const void* scalars, const void* buckets, const double* le,
                  const int32_t* phase, int64_t n_pairs, const cg_params* p,
                  double* metrics, uint8_t* valid, int n_threads)
{
    if (!params_ok(p)) return CG_EINVAL;
    cgo_ctx c = {scalars, buckets, le, phase, 0, 0, p, 0, 0, metrics, valid};
    parallel_for(n_pairs * 2, n_threads, grid_range, &c);
    return CG_OK;
}

int cgo_gate(const double* new_m, const uint8_t* new_v, const double* old_m,
             const uint8_t* old_v, int64_t n, const cg_params* p, uint8_t* detail)
{
    if (!params_ok(p)) return CG_EINVAL;
    const int W = cgo_metric_width(p);
    for (int64_t i = 0; i < n; ++i)
        detail[i] = gate_one(new_m + i * W, new_v[i], old_m + i * W, old_v[i], p);
    return CG_OK;
}

int cgo_replay(const uint32_t* gate_bits, const uint8_t* has_previous,
               const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
               cg_decision* decisions)
{
    if (!params_ok(p)) return CG_EINVAL;
    const int nw = (p->n_ticks + 31) / 32;
    for (int64_t i = 0; i < n_pairs; ++i)
        replay_one(gate_bits + i * nw, has_previous ? has_previous[i] : 1,
                   start_tick ? start_tick[i] : p->start_tick, p, decisions + i);
    return CG_OK;
}

/* get_model_metrics x2 -> should_promote_model -> step loop, per pair (:302-317). */
static void fused_range(int64_t begin, int64_t end, void* ctx)
{
    const cgo_ctx* c = (const cgo_ctx*)ctx;
    const cg_params* p = c->p;
    const size_t esz = p->dtype == CG_F32 ? 4 : 8;
    const int nw = (p->n_ticks + 31) / 32;
    for (int64_t pr = begin; pr < end; ++pr) {
        uint32_t bits[256];
        uint32_t* bw = c->gate_bits ? c->gate_bits + pr * nw : bits;
        memset(bw, 0, sizeof(uint32_t) * nw);
        for (int j = 0; j < p->n_ticks; ++j) {
            double m[2][CG_MAX_Q + CG_M_EXTRA];
            uint8_t v[2];
            for (int ver = 0; ver < 2; ++ver) {
                int64_t pv = pr * 2 + ver;
                const char* sc = (const char*)c->scalars + (size_t)pv * p->n_scalar * p->row_pitch * esz;
                const char* bk = c->buckets ? (const char*)c->buckets + (size_t)pv * p->n_buckets * p->row_pitch * esz : 0;
                eval_pv_tick(sc, bk, c->le, c->phase[pv], p, j, m[ver], &v[ver]);
            }
            if (gate_one(m[0], v[0], m[1], v[1], p) & CG_G_PROMOTE) bw[j >> 5] |= 1u << (j & 31);
        }
        replay_one(bw, c->has_previous ? c->has_previous[pr] : 1,
                   c->start_tick ? c->start_tick[pr] : p->start_tick, p, c->decisions + pr);
    }
}

int cgo_eval_fused(const void* scalars, const void* buckets, const double* le,
                   const int32_t* phase, const uint8_t* has_previous,
                   const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
                   cg_decision* decisions, uint32_t* gate_bits, int n_threads)
{
    if (!params_ok(p)) return CG_EINVAL;
    if (!gate_bits && (p->n_ticks + 31) / 32 > 256) return CG_EINVAL;
    cgo_ctx c = {scalars, buckets, le, phase, has_previous, start_tick, p, decisions, gate_bits, 0, 0};
    parallel_for(n_pairs, n_threads, fused_range, &c);
    return CG_OK;
}

/* The same loop the way the reference actually runs it: metrics are evaluated only at
 * the ticks the schedule visits (:300-310), at most 9*10 of them.  Same decisions as
 * cgo_eval_fused; used as a cross-check and as the "lazy" CPU figure in bench.py. */
static void lazy_range(int64_t begin, int64_t end, void* ctx)
{
    const cgo_ctx* c = (const cgo_ctx*)ctx;
    const cg_params* p = c->p;
    const void* scalars = c->scalars; const void* buckets = c->buckets; const double* le = c->le;
    const int32_t* phase = c->phase; const uint8_t* has_previous = c->has_previous;
    const int32_t* start_tick = c->start_tick; cg_decision* decisions = c->decisions;
    const size_t esz = p->dtype == CG_F32 ? 4 : 8;
    for (int64_t pr = begin; pr < end; ++pr) {
        cg_decision* out = decisions + pr;
        if (has_previous && !has_previous[pr]) { replay_one(0, 0, 0, p, out); continue; }
        int cur = p->init_current, prev = p->max_traffic - p->init_current;
        int tick = start_tick ? start_tick[pr] : p->start_tick;
        int n_prom = 0, n_eval = 0, last = -1, state = CG_STATE_IN_PROGRESS, out_of_data = 0;
        while (cur < p->max_traffic && !out_of_data) {
            int attempt = 0, success = 0;
            while (attempt < p->max_attempts) {
                if (tick >= p->n_ticks || tick < 0) { out_of_data = 1; break; }
                ++n_eval; last = tick;
                double m[2][CG_MAX_Q + CG_M_EXTRA];
                uint8_t v[2];
                for (int ver = 0; ver < 2; ++ver) {
                    int64_t pv = pr * 2 + ver;
                    const char* sc = (const char*)scalars + (size_t)pv * p->n_scalar * p->row_pitch * esz;
                    const char* bk = buckets ? (const char*)buckets + (size_t)pv * p->n_buckets * p->row_pitch * esz : 0;
                    eval_pv_tick(sc, bk, le, phase[pv], p, tick, m[ver], &v[ver]);
                }
                if (gate_one(m[0], v[0], m[1], v[1], p) & CG_G_PROMOTE) {
                    cur += p->step; prev -= p->step;
                    if (cur > p->max_traffic) cur = p->max_traffic;
                    if (prev < 0) prev = 0;
                    ++n_prom; success = 1;
                    break;
                }
                ++attempt;
                if (attempt < p->max_attempts) tick += p->attempt_delay_ticks;
            }
            if (out_of_data) break;
            if (!success) { state = CG_STATE_ABORTED; break; }
            tick += p->interval_ticks;
        }
        if (cur == p->max_traffic) state = CG_STATE_COMPLETE;
        if (state == CG_STATE_ABORTED && p->on_abort == CG_ABORT_ROLLBACK) {
            cur = 0; prev = p->max_traffic; state = CG_STATE_ROLLED_BACK;
        }
        out->traffic_current = cur; out->traffic_prev = prev; out->state = state;
        out->n_promotions = n_prom; out->n_evals = n_eval; out->last_tick = last;
    }
}

int cgo_eval_lazy(const void* scalars, const void* buckets, const double* le,
                  const int32_t* phase, const uint8_t* has_previous,
                  const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
                  cg_decision* decisions, int n_threads)
{
    if (!params_ok(p)) return CG_EINVAL;
    cgo_ctx c = {scalars, buckets, le, phase, has_previous, start_tick, p, decisions, 0, 0, 0};
    parallel_for(n_pairs, n_threads, lazy_range, &c);
    return CG_OK;
}

/* ---- irregular series: per-series timestamps, ragged lengths --------
 * Same queries as eval_pv_tick, but every series carries its own time stamps (strictly increasing,
 * ms) and its own sample count; missed scrapes are simply absent and stale markers have been
 * dropped by the ingest layer, as Prometheus drops them from range vectors.  Each series selects
 * its own window (Appendix A "Range selection") and may yield no output (< 2 samples); a bucket
 * series without output is absent from the `sum(...) by (le)` vector that histogram_quantile sees.
 *   ts_sc [K][row_pitch] int32, cnt_sc [K];  ts_bk [B][row_pitch] int32, cnt_bk [B]   (one version) */
static int irregular_increase(const void* vals, int dtype, int64_t row_off, const int32_t* ts, int cnt,
                              int64_t t, const cg_params* p, double* win, int win_cap, double* inc)
{
    const int64_t lo = t - p->window_ms;
    int a = 0, b = cnt - 1;
    while (a < cnt && (p->left_open ? (int64_t)ts[a] <= lo : (int64_t)ts[a] < lo)) ++a;
    while (b >= 0 && (int64_t)ts[b] > t) --b;
    const int n = b - a + 1;
    if (n < 2 || n > win_cap) return 0;
    for (int i = 0; i < n; ++i) win[i] = load_elem(vals, dtype, row_off + a + i);
    return increase_core(win, n, ts[a], ts[b], t, p->window_ms, inc);
}

static void eval_pv_tick_irregular(const void* sc, const void* bk, const int32_t* ts_sc, const int32_t* cnt_sc,
                                   const int32_t* ts_bk, const int32_t* cnt_bk, const double* le,
                                   const cg_params* p, int tick, double* win, double* m, uint8_t* valid_out)
{
    const int nq = p->n_q;
    const int64_t t = (int64_t)p->tick0_ms + (int64_t)tick * p->tick_step_ms;
    uint8_t valid = 0;
    double sums[5] = {0, 0, 0, 0, 0};
    int have[5] = {0, 0, 0, 0, 0};
    for (int k = 0; k < p->n_scalar; ++k) {               /* Q2..Q6 (:375, :383, :393, :397, :410) */
        double inc;
        if (!irregular_increase(sc, p->dtype, (int64_t)k * p->row_pitch, ts_sc + (int64_t)k * p->row_pitch,
                                cnt_sc[k], t, p, win, p->row_pitch, &inc)) continue;
        for (int r = 0; r < 5; ++r)
            if (p->scalar_role[k] & (1u << r)) { sums[r] = have[r] ? sums[r] + inc : inc; have[r] = 1; }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    {                                                      /* Q1 (:367) over the buckets that have output */
        double le_p[1024], cnt_p[1024], tmp[1024];
        int np = 0;
        const double range_s = (double)p->window_ms / 1000.0;
        for (int b = 0; b < p->n_buckets; ++b) {
            double inc;
            if (!irregular_increase(bk, p->dtype, (int64_t)b * p->row_pitch, ts_bk + (int64_t)b * p->row_pitch,
                                    cnt_bk[b], t, p, win, p->row_pitch, &inc)) continue;
            le_p[np] = le[b];
            cnt_p[np] = inc / range_s;
            ++np;
        }
        if (np > 0) {
            for (int j = 0; j < nq; ++j) {
                memcpy(tmp, cnt_p, sizeof(double) * np);
                m[j] = cgo_bucket_quantile(p->q[j], le_p, tmp, np);
            }
            valid |= CG_V_QUANTILE;                         /* non-empty vector (:369-370) */
        }
    }
    m[nq + CG_M_ERROR_RESPONSES] = err;
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; valid |= CG_V_ERROR_RATE; }
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; valid |= CG_V_LATENCY_AVG; }
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
    m[nq + CG_M_FEEDBACK_COUNT] = fb;
    *valid_out = valid;
}

typedef struct {
    const void* scalars; const void* buckets;
    const int32_t* ts_sc; const int32_t* ts_bk; const int32_t* cnt_sc; const int32_t* cnt_bk;
    const double* le; const uint8_t* has_previous; const int32_t* start_tick; const cg_params* p;
    cg_decision* decisions; uint32_t* gate_bits; double* metrics; uint8_t* valid;
} cgo_irr_ctx;

static void irregular_range(int64_t begin, int64_t end, void* ctx)
{
    const cgo_irr_ctx* c = (const cgo_irr_ctx*)ctx;
    const cg_params* p = c->p;
    const size_t esz = p->dtype == CG_F32 ? 4 : 8;
    const int nw = (p->n_ticks + 31) / 32, W = cgo_metric_width(p);
    double* win = (double*)__builtin_alloca(sizeof(double) * (size_t)p->row_pitch);
    for (int64_t pr = begin; pr < end; ++pr) {
        uint32_t bits[256];
        uint32_t* bw = c->gate_bits ? c->gate_bits + pr * nw : bits;
        memset(bw, 0, sizeof(uint32_t) * nw);
        for (int j = 0; j < p->n_ticks; ++j) {
            double m[2][CG_MAX_Q + CG_M_EXTRA];
            uint8_t v[2];
            for (int ver = 0; ver < 2; ++ver) {
                const int64_t pv = pr * 2 + ver;
                const char* sc = (const char*)c->scalars + (size_t)pv * p->n_scalar * p->row_pitch * esz;
                const char* bk = c->buckets ? (const char*)c->buckets + (size_t)pv * p->n_buckets * p->row_pitch * esz : 0;
                eval_pv_tick_irregular(sc, bk, c->ts_sc + (size_t)pv * p->n_scalar * p->row_pitch,
                                       c->cnt_sc + (size_t)pv * p->n_scalar,
                                       c->ts_bk ? c->ts_bk + (size_t)pv * p->n_buckets * p->row_pitch : 0,
                                       c->cnt_bk ? c->cnt_bk + (size_t)pv * p->n_buckets : 0,
                                       c->le, p, j, win, m[ver], &v[ver]);
                if (c->metrics) memcpy(c->metrics + ((size_t)pv * p->n_ticks + j) * W, m[ver], sizeof(double) * W);
                if (c->valid) c->valid[(size_t)pv * p->n_ticks + j] = v[ver];
            }
            if (gate_one(m[0], v[0], m[1], v[1], p) & CG_G_PROMOTE) bw[j >> 5] |= 1u << (j & 31);
        }
        if (c->decisions)
            replay_one(bw, c->has_previous ? c->has_previous[pr] : 1,
                       c->start_tick ? c->start_tick[pr] : p->start_tick, p, c->decisions + pr);
    }
}

int cgo_eval_irregular(const void* scalars, const void* buckets, const int32_t* ts_scalars,
                       const int32_t* ts_buckets, const int32_t* cnt_scalars, const int32_t* cnt_buckets,
                       const double* le, const uint8_t* has_previous, const int32_t* start_tick,
                       int64_t n_pairs, const cg_params* p, cg_decision* decisions, uint32_t* gate_bits,
                       double* metrics, uint8_t* valid, int n_threads)
{
    if (p->n_buckets > 1024 || p->n_q > CG_MAX_Q || p->n_ticks < 0) return CG_EINVAL;
    if (!gate_bits && (p->n_ticks + 31) / 32 > 256) return CG_EINVAL;
    cgo_irr_ctx c = {scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le,
                     has_previous, start_tick, p, decisions, gate_bits, metrics, valid};
    parallel_for(n_pairs, n_threads, irregular_range, &c);
    return CG_OK;
}

int cgo_max_threads(void)
{
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}
