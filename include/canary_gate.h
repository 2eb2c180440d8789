/*
 * canary_gate.h -- C ABI of libcanarygate.so, the B200 (sm_100a) implementation of
 * the canary metric-validation step of the MLflow->Seldon operator.
 *
 * Plain C: pointers, sizes and POD structs only; no torch / C++ types.  Every
 * function returns 0 on success or a CG_E* code; cg_last_error() gives the text
 * (thread-local).  No C++ exception crosses this boundary.  Unless a parameter is
 * documented as HOST memory, every data pointer is DEVICE memory owned by the
 * caller.  All work is enqueued on the `stream` argument (a cudaStream_t passed as
 * void*, NULL = the legacy default stream) of the CURRENT device; nothing here
 * synchronises unless documented.  The library keeps no mutable global state (a launch
 * counter for reporting aside), reads no environment variable and allocates no device
 * memory: every kernel works in its inputs, its outputs and shared memory.
 *
 * Reference citations are file:line in /root/reference/mlflow_operator.py.
 *
 * Data layout (SoA, sample axis contiguous):
 *   scalars [n_pairs][2][n_scalar ][row_pitch]   counter series
 *   buckets [n_pairs][2][n_buckets][row_pitch]   cumulative histogram counters
 *   le      [n_buckets]  fp64 upper bounds, strictly increasing, last = +Inf
 *   phase   [n_pairs][2] int32: timestamp (ms) of sample 0 of that version's series;
 *           sample i was scraped at phase + i*period_ms
 * Version index 0 is the NEW (current) model, 1 the OLD (previous) one
 * (:302-303).  Elements are fp64 (CG_F64) or fp32 (CG_F32); arithmetic is always
 * IEEE fp64 with no FMA contraction, because the reference computes in Python
 * floats (:387-404, :440-454).
 */
#ifndef CANARY_GATE_H
#define CANARY_GATE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CG_ABI_VERSION 1

/* status codes */
#define CG_OK          0
#define CG_EINVAL      1   /* bad argument / unsupported geometry */
#define CG_ECUDA       2   /* a CUDA runtime call failed */
#define CG_ENOSPACE    3   /* caller workspace / static scratch too small */

/* element types of the series arrays */
#define CG_F64 0
#define CG_F32 1

/* role bits of a scalar counter series: which of the reference's `sum(increase())`
 * queries the series' labels match.  A series may match several. */
#define CG_ROLE_SERVER_TOTAL 0x01  /* ..server_requests_seconds_count{}              (Q3, :383) */
#define CG_ROLE_SERVER_ERROR 0x02  /* ..server_requests_seconds_count{code!="200"}   (Q2, :375) */
#define CG_ROLE_CLIENT_SUM   0x04  /* ..client_requests_seconds_sum                  (Q4, :393) */
#define CG_ROLE_CLIENT_COUNT 0x08  /* ..client_requests_seconds_count                (Q5, :397) */
#define CG_ROLE_FEEDBACK     0x10  /* ..server_requests_seconds_count{service="feedback"} (Q6, :410) */

#define CG_MAX_SCALAR 16
#define CG_MAX_Q      4

/* per-tick metric record: n_q quantiles then these five (all fp64) */
#define CG_M_ERROR_RESPONSES 0   /* + n_q  ('error_responses',  :378) */
#define CG_M_ERROR_RATE      1   /* + n_q  ('error_rate',       :388) */
#define CG_M_LATENCY_AVG     2   /* + n_q  ('latency_avg',      :402) */
#define CG_M_REQUEST_COUNT   3   /* + n_q  ('request_count',    :407) */
#define CG_M_FEEDBACK_COUNT  4   /* + n_q  ('feedback_request_count', :413) */
#define CG_M_EXTRA           5

/* validity bits of a metric record (a cleared bit = Python None in the reference) */
#define CG_V_QUANTILE   0x01  /* histogram_quantile query returned a sample (:369-372) */
#define CG_V_ERROR_RATE 0x02  /* total_responses > 0 (:387-390) */
#define CG_V_LATENCY_AVG 0x04 /* latency_count > 0 (:401-404) */

/* gate detail bits (one byte per evaluation) */
#define CG_G_P95_OK     0x01  /* :440 */
#define CG_G_ERR_OK     0x02  /* :447 */
#define CG_G_AVG_OK     0x04  /* :454 */
#define CG_G_AVAILABLE  0x08  /* all required metrics present on both sides (:430-434) */
#define CG_G_PROMOTE    0x80  /* return value of should_promote_model (:460) */

/* decision states */
#define CG_STATE_IN_PROGRESS 0 /* ran out of ticks before the schedule finished */
#define CG_STATE_COMPLETE    1 /* reached max_traffic; old predictor dropped (:355-361) */
#define CG_STATE_ABORTED     2 /* max_attempts consecutive failures; weights frozen (:341-349) */
#define CG_STATE_NO_CANARY   3 /* no previous version: 100/0, no loop (:188-191, :289) */
#define CG_STATE_ROLLED_BACK 4 /* only with on_abort = CG_ABORT_ROLLBACK: as ABORTED, then all traffic back to the old version */

/* what the schedule does after max_attempts consecutive failures */
#define CG_ABORT_FREEZE   0    /* the reference: log, event, weights stay where they are (:341-349) */
#define CG_ABORT_ROLLBACK 1    /* opt-in extension (the reference leaves a placeholder at :345):
                                * traffic_current = 0, traffic_prev = max_traffic */

typedef struct cg_params {
    int32_t struct_bytes;        /* = sizeof(cg_params); checked by every entry point */
    int32_t dtype;               /* CG_F64 | CG_F32 */
    /* series geometry */
    int32_t n_samples;           /* T: samples per series */
    int32_t row_pitch;           /* elements between rows; >= T; pitch bytes % 16 == 0 */
    int32_t n_scalar;            /* K scalar counter series per (pair, version), <= CG_MAX_SCALAR */
    int32_t n_buckets;           /* B histogram buckets per (pair, version); 0 = none */
    uint8_t scalar_role[CG_MAX_SCALAR];
    /* time base, all in ms relative to the store's epoch */
    int32_t period_ms;           /* scrape period */
    int32_t window_ms;           /* range of rate()/increase(): elapsed_time*1000 (:363) */
    int32_t tick0_ms;            /* evaluation time of tick 0 */
    int32_t tick_step_ms;        /* spacing of the evaluation grid (attempt_delay, :294) */
    int32_t n_ticks;
    int32_t left_open;           /* 0: [t-r, t] (Prometheus 2.x)   1: (t-r, t] (3.x) */
    /* quantiles */
    int32_t n_q;                 /* 1..CG_MAX_Q (0 allowed when n_buckets == 0) */
    int32_t gate_q;              /* which quantile feeds the latency_95th check (:367 uses 0.95) */
    double  q[CG_MAX_Q];
    /* gate: the reference evaluates (1 + thresholds[k]) in fp64 at run time (:440,447,454) */
    double  one_plus_thr[3];     /* latency_95th, error_rate, latency_avg (:175-179) */
    int32_t checks;              /* CG_G_P95_OK|CG_G_ERR_OK|CG_G_AVG_OK to require (7 = reference) */
    /* traffic schedule (:184-191, :290-294) */
    int32_t init_current;        /* 10  (:187) */
    int32_t max_traffic;         /* 100 (:290) */
    int32_t step;                /* 10  (:291) */
    int32_t interval_ticks;      /* interval / tick_step = 6 (:292) */
    int32_t max_attempts;        /* 10  (:293) */
    int32_t attempt_delay_ticks; /* attempt_delay / tick_step = 1 (:294) */
    int32_t start_tick;          /* tick at which the canary loop starts (:296) */
    int32_t reserved;            /* must be 0 (every entry point rejects anything else) */
    int32_t on_abort;            /* CG_ABORT_FREEZE (reference, default) | CG_ABORT_ROLLBACK */
} cg_params;

/* Decision record, 24 bytes. */
typedef struct cg_decision {
    int32_t traffic_current;     /* :312-317 */
    int32_t traffic_prev;
    int32_t state;               /* CG_STATE_* */
    int32_t n_promotions;
    int32_t n_evals;             /* gate evaluations consumed (each = 12 queries in the reference, :302-303) */
    int32_t last_tick;           /* tick index of the last evaluation, -1 if none */
} cg_decision;

/* ---- housekeeping -------------------------------------------------------- */
int         cg_abi_version(void);
const char* cg_last_error(void);

/* Fill *p with the reference's constants: window 60 s (:363), q = 0.95 (:367),
 * thresholds 0.05/0.02/0.05 (:175-179), 10/100/10, interval 60 s, 10 attempts,
 * delay 10 s (:186-187, :290-294); 15 s scrapes, ticks every 10 s from t = 60 s,
 * K = 4 scalar series (server 200, server 500, client sum, client count), B = 64. */
int cg_params_default(cg_params* p);

/* Validate a parameter block without touching the GPU. */
int cg_params_check(const cg_params* p);

/* Number of fp64 values per metric record for these params (n_q + CG_M_EXTRA). */
int cg_metric_width(const cg_params* p);

/* Algorithmic HBM bytes one pair costs:
 * 2*(K+B)*T*sizeof(elem) + sizeof(cg_decision) + 8. */
int64_t cg_bytes_per_pair(const cg_params* p);

/* ---- the hot path -------------------------------------------------------- */

/* Batched get_model_metrics (:363-417) on a dense tick grid: for every pair,
 * both versions and ticks [0, n_ticks) evaluate the six queries (:367-410), the
 * None/0.0 defaults (:369-415) and the two derived divisions (:387-404).
 *   metrics [n_pairs][2][n_ticks][cg_metric_width]  fp64   (may be NULL)
 *   valid   [n_pairs][2][n_ticks]                   uint8  CG_V_* (may be NULL)
 * Replaces: 6 x PrometheusConnect.custom_query + parsing per (version, tick). */
int cg_eval_grid(const void* scalars, const void* buckets, const double* le,
                 const int32_t* phase, int64_t n_pairs, const cg_params* p,
                 double* metrics, uint8_t* valid, void* stream);

/* Batched should_promote_model (:419-460) over n records:
 *   new_m/old_m [n][cg_metric_width] fp64, new_v/old_v [n] uint8 -> detail [n] uint8 CG_G_*. */
int cg_gate(const double* new_m, const uint8_t* new_v, const double* old_m,
            const uint8_t* old_v, int64_t n, const cg_params* p, uint8_t* detail,
            void* stream);

/* The traffic step loop (:296-352, completion :355-361) as a pure function of
 * gate bits:  gate_bits [n_pairs][ceil(n_ticks/32)] uint32 (bit j%32 of word j/32
 * = promote at tick j); has_previous [n_pairs] uint8 or NULL (= all 1, :184-191);
 * start_tick [n_pairs] int32 or NULL (= p->start_tick). */
int cg_replay(const uint32_t* gate_bits, const uint8_t* has_previous,
              const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
              cg_decision* decisions, void* stream);

/* Everything fused in one pass over the series: metrics -> gate -> replay, no
 * intermediate in HBM.  Optional extra outputs (NULL to skip):
 *   gate_bits [n_pairs][ceil(n_ticks/32)] uint32,
 *   metrics/valid as in cg_eval_grid.
 * Replaces :302-317 for n_pairs MlflowModel resources at once. */
int cg_eval_fused(const void* scalars, const void* buckets, const double* le,
                  const int32_t* phase, const uint8_t* has_previous,
                  const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
                  cg_decision* decisions, uint32_t* gate_bits,
                  double* metrics, uint8_t* valid, void* stream);

/* The same fused step for IRREGULAR series: every series has its own time
 * stamps and its own number of samples -- jittered or missed scrapes, series that start late or
 * end early.  Stale markers must already be dropped, as Prometheus drops them from range vectors
 * (canary_b200.ingest does it).  A series with fewer than 2 samples in a window yields no output
 * for that tick; a bucket without output is absent from the vector histogram_quantile sees.
 *   scalars / buckets        values, laid out as for cg_eval_fused (only the first cnt samples of a row count)
 *   ts_scalars / ts_buckets  [n_pairs][2][K|B][row_pitch] int32 ms, strictly increasing within a row,
 *                            same epoch as tick0_ms; 16-byte aligned; row_pitch a multiple of 4
 *   cnt_scalars/ cnt_buckets [n_pairs][2][K|B] int32 samples present (0..n_samples)
 * period_ms is ignored.  Outputs as in cg_eval_fused (any may be NULL, not all).
 * Replaces :302-317 against a Prometheus whose scrapes are not a perfect grid. */
int cg_eval_irregular(const void* scalars, const void* buckets, const int32_t* ts_scalars,
                      const int32_t* ts_buckets, const int32_t* cnt_scalars,
                      const int32_t* cnt_buckets, const double* le, const uint8_t* has_previous,
                      const int32_t* start_tick, int64_t n_pairs, const cg_params* p,
                      cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                      uint8_t* valid, void* stream);

/* Same as cg_eval_irregular with an explicit kernel choice: path 0 = automatic, 1 = values and time
 * stamps of a (pair, version) resident in shared memory (at most about 220 KB of them), 3 = rows
 * and their time-stamp rows through a ring (any K + B).  Automatic: 3 for grids of 32 ticks or
 * more or when the resident tile does not fit, else 1.  Both give bit-identical results.
 * Replaces the same lines as cg_eval_irregular (:302-317). */
int cg_eval_irregular_path(int path, const void* scalars, const void* buckets,
                           const int32_t* ts_scalars, const int32_t* ts_buckets,
                           const int32_t* cnt_scalars, const int32_t* cnt_buckets, const double* le,
                           const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                           const cg_params* p, cg_decision* decisions, uint32_t* gate_bits,
                           double* metrics, uint8_t* valid, void* stream);

/* Benchmark / test utility with no counterpart in the reference: fill device arrays with
 * seeded synthetic series (SURVEY.md 8(d) model).  Values are a pure function of
 * (seed, pair_offset + local pair, version, series, sample), so shards agree. */
int cg_synth_fill(void* scalars, void* buckets, const double* le, int32_t* phase,
                  int64_t n_pairs, int64_t pair_offset, const cg_params* p, uint64_t seed,
                  double reset_prob, double adversarial, void* stream);

/* Device-resident scrape history (the reference keeps none: every attempt asks Prometheus for
 * "now", :302-303, and sleeps 10 s or 60 s in between, :340, :352; Prometheus' TSDB holds the
 * hour).  Rows are allocated with slack after the newest sample (row_pitch > samples in use).
 * cg_store_append writes one new scrape -- new_scalars [n_pairs][2][n_scalar], new_buckets
 * [n_pairs][2][n_buckets], DEVICE memory, element type p->dtype -- into column `column` of every
 * row.  cg_store_shift moves samples [shift, shift+count) of every row to [0, count) (the oldest
 * `shift` samples fall off; count * element size <= 48 KB); the caller adds shift*period_ms to
 * its phases.  Only p->dtype, n_scalar, n_buckets and row_pitch are used. */
int cg_store_append(void* scalars, void* buckets, const void* new_scalars, const void* new_buckets,
                    int64_t n_pairs, const cg_params* p, int32_t column, void* stream);
int cg_store_shift(void* scalars, void* buckets, int64_t n_pairs, const cg_params* p, int32_t shift,
                   int32_t count, void* stream);

/* Same as cg_eval_fused with an explicit kernel choice: path 0 = automatic, 1 = the general
 * resident-tile kernel, 2 = the streaming kernel (CG_EINVAL if the geometry is not covered:
 * it needs window = 4 scrape periods and tick step = 2/3 period, the reference's constants
 * (:363, :294) with 15 s scrapes, at least 8 samples per series, at most 64 buckets and at
 * least one scalar series), 3 = the general row-streaming kernel (any window, tick grid and
 * scrape period; rows must be a multiple of 16 bytes).  Automatic: 2 where it applies, else 3
 * for grids of 32 ticks or more (or whenever decisions are asked for), else 1.  All kernels
 * give bit-identical results. */
int cg_eval_fused_path(int path, const void* scalars, const void* buckets, const double* le,
                       const int32_t* phase, const uint8_t* has_previous, const int32_t* start_tick,
                       int64_t n_pairs, const cg_params* p, cg_decision* decisions,
                       uint32_t* gate_bits, double* metrics, uint8_t* valid, void* stream);

/* 1 if cg_eval_fused would take the streaming kernel for these params, else 0. */
int cg_stream_eligible(const cg_params* p);

/* Launch geometry the library would use (for reporting): CTAs, threads, dynamic smem. */
int cg_launch_info(const cg_params* p, int64_t n_pairs, int32_t* grid, int32_t* block,
                   int32_t* smem_bytes);

/* Number of kernel launches the library has enqueued from this process so far. */
int64_t cg_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* CANARY_GATE_H */
