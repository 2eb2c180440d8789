// cg_irregular.cuh -- the canary step for IRREGULAR series: every series
// carries its own time stamps and its own sample count.  Missed scrapes are simply absent; stale
// markers have been dropped by the ingest layer, as Prometheus drops them from range vectors.
//
// Same structure as the general resident-tile kernel (cg_general.cuh): one CTA walks pairs; for
// each (pair, version) the value tile [K+B rows x pitch], the time-stamp tile of the same shape
// (int32 ms) and the per-row sample counts are brought into shared memory (TMA bulk copies on one
// mbarrier), then one thread per tick evaluates the six queries of get_model_metrics
// (mlflow_operator.py:367-410).  What differs is the window: each series finds its own
// [first, last] samples by a short walk from a hint in its time stamps, the extrapolation terms
// (Appendix A steps 3-5) are computed from those time stamps -- so the avg*1.1 threshold branch
// is live -- and a series with fewer than 2 samples in the window yields no output: it drops out
// of its sum, and a bucket without output is absent from the vector histogram_quantile sees.
#pragma once
#include "cg_math.cuh"
#include "cg_ptx.cuh"

namespace cg {

struct IrregularArgs {
    const void*    scalars;
    const void*    buckets;
    const int32_t* ts_scalars;     // [n_pairs][2][K][pitch] ms, strictly increasing within a row
    const int32_t* ts_buckets;     // [n_pairs][2][B][pitch]
    const int32_t* cnt_scalars;    // [n_pairs][2][K] samples present
    const int32_t* cnt_buckets;    // [n_pairs][2][B]
    const double*  le;
    const uint8_t* has_previous;   // may be null
    const int32_t* start_tick;     // may be null
    cg_decision*   decisions;      // may be null
    uint32_t*      gate_bits;      // may be null
    double*        metrics;        // may be null
    uint8_t*       valid;          // may be null
    int64_t        n_pairs;
    cg_params      p;
    uint32_t off_ts, off_cnt, off_le, off_stash, off_stash_valid, off_bits, off_bar, smem_bytes;
};

// Extrapolation terms of a window (Appendix A steps 3 and 5).  They depend on the time stamps alone,
// and the series of one target are scraped together, so a thread keeps the terms of the last window
// it saw and reuses them while (first, last, count) stay the same: same operations, done once.
struct WindowTerms {
    int first_t, last_t, n;      // the window they belong to; n = 0: none yet
    int a;                       // index of its first sample (search hint for the next row)
    double d_start, sampled, thr, half_avg, end_term;
    double f;                    // ext / sampled when the zero-point clamp does not fire
    double gchk;                 // first_v * gchk >= result  ==>  the zero-point clamp cannot fire
};

// increase() of one series at one tick from its own time stamps (Appendix A steps 1-6, the same
// operations in the same order as oracle/promql_oracle.c increase_core).  false: no output.
template <typename T>
__device__ bool irregular_increase(const T* __restrict__ v, const int32_t* __restrict__ ts, int cnt, int t0, float inv,
                                   int t, int lo, bool left_open, WindowTerms& wt, double& inc) {
    if (cnt < 2) return false;
    // a = first sample inside the window.  Start where the previous row had it (same scrape times), or
    // where a regular grid would put it (t0 = first time stamp of the row, inv = samples per ms), and
    // walk: O(1) for near-regular scrapes, O(gap) otherwise.
    int a = wt.n ? wt.a : (int)((float)(lo - t0) * inv);
    a = a < 0 ? 0 : (a > cnt ? cnt : a);
    // (the hint is right for all but a few rows: each loop tests once and falls through; not unrolled)
#pragma unroll 1
    while (a > 0 && !(left_open ? ts[a - 1] <= lo : ts[a - 1] < lo)) --a;
#pragma unroll 1
    while (a < cnt && (left_open ? ts[a] <= lo : ts[a] < lo)) ++a;
    // b = last sample at or before the tick: start from the length of the previous row's window
    int b = a - 1 + (wt.n ? wt.n : 0);
    if (b > cnt - 1) b = cnt - 1;
#pragma unroll 1
    while (b >= a && ts[b] > t) --b;
#pragma unroll 1
    while (b + 1 < cnt && ts[b + 1] <= t) ++b;
    const int n = b - a + 1;
    if (n < 2) return false;                                   // step 1
    const double first_v = (double)v[a];
    double result = (double)v[b] - first_v;                    // step 2
    double prev = first_v;
#pragma unroll 1
    for (int i = a + 1; i <= b; ++i) {
        const double x = (double)v[i];
        if (x < prev) result += prev;                          // counter reset
        prev = x;
    }
    const int first_t = ts[a], last_t = ts[b];
    if (!(wt.n == n && wt.first_t == first_t && wt.last_t == last_t)) {
        // (division by a constant: div_const gives the correctly rounded quotient, as `/` does)
        wt.d_start = div_const((double)(first_t - lo), 1000.0, 1e-3);          // step 3
        const double d_end = div_const((double)(t - last_t), 1000.0, 1e-3);
        wt.sampled = div_const((double)(last_t - first_t), 1000.0, 1e-3);
        const double avg = wt.sampled / (double)(n - 1);
        wt.thr = avg * 1.1;                                    // step 5
        wt.half_avg = avg / 2;
        wt.end_term = (d_end < wt.thr) ? d_end : wt.half_avg;
        double ext = wt.sampled;
        ext += (wt.d_start < wt.thr) ? wt.d_start : wt.half_avg;
        ext += wt.end_term;
        wt.f = ext / wt.sampled;
        wt.gchk = (wt.sampled / wt.d_start) * (1.0 - 1e-6);   // (d_start = 0: +Inf, or NaN: always the exact path)
        wt.first_t = first_t; wt.last_t = last_t; wt.n = n;
    }
    wt.a = a;
    if (result > 0 && first_v >= 0) {                          // step 4: the zero-point clamp
        // d_zero = sampled * (first_v / result) < d_start needs first_v * (sampled / d_start) < result: with gchk a
        // slight under-estimate of sampled / d_start, first_v * gchk >= result proves that the clamp cannot fire
        // (one multiply instead of a division; the streaming kernel's test)
        if (!(first_v * wt.gchk >= result)) {
            const double d_zero = wt.sampled * (first_v / result);
            if (d_zero < wt.d_start) {
                double ext = wt.sampled;
                ext += (d_zero < wt.thr) ? d_zero : wt.half_avg;
                ext += wt.end_term;
                inc = result * (ext / wt.sampled);             // step 6
                return true;
            }
        }
    }
    inc = result * wt.f;                                       // step 6
    return true;
}

template <typename T>
__device__ void eval_tick_irregular(const T* __restrict__ sc_rows, const T* __restrict__ bk_rows,
                                    const int32_t* __restrict__ ts_sc, const int32_t* __restrict__ ts_bk,
                                    const int32_t* __restrict__ cnt_sc, const int32_t* __restrict__ cnt_bk,
                                    const int32_t* __restrict__ t0_row, const float* __restrict__ inv_row,
                                    const double* __restrict__ le, const cg_params& p, int tick,
                                    double* __restrict__ m, uint8_t& valid_out) {
    const int nq = p.n_q, pitch = p.row_pitch, B = p.n_buckets, K = p.n_scalar;
    const int t = p.tick0_ms + tick * p.tick_step_ms, lo = t - p.window_ms;
    const bool lopen = p.left_open != 0;
    uint8_t valid = 0;
    WindowTerms wt;
    wt.n = 0;
    double sums[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    unsigned have = 0;
    for (int k = 0; k < p.n_scalar; ++k) {      // Q2..Q6: sum(increase(...)) or on() vector(0)
        double inc;
        if (!irregular_increase(sc_rows + (size_t)k * pitch, ts_sc + (size_t)k * pitch, cnt_sc[k], t0_row[k], inv_row[k],
                                t, lo, lopen, wt, inc))
            continue;
        const unsigned role = p.scalar_role[k];
#pragma unroll
        for (int r = 0; r < 5; ++r)
            if (role & (1u << r)) {
                sums[r] = (have & (1u << r)) ? sums[r] + inc : inc;
                have |= 1u << r;
            }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];

    // Q1: histogram_quantile over the buckets that have output, in `le` order
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    const double range_s = (double)p.window_ms / 1000.0, rc_range = 1.0 / range_s;
    // Usual shape first, in ONE pass: the +Inf bucket has output, and (speculation, checked at the end)
    // it holds the observations, i.e. no lower bucket exceeds it after the monotonic fix.
    bool done = false;
    if (B >= 1 && isinf(le[B - 1]) && le[B - 1] > 0) {
        double inc_inf;
        if (irregular_increase(bk_rows + (size_t)(B - 1) * pitch, ts_bk + (size_t)(B - 1) * pitch, cnt_bk[B - 1],
                               t0_row[K + B - 1], inv_row[K + B - 1], t, lo, lopen, wt, inc_inf)) {
            const double c_inf = div_const(inc_inf, range_s, rc_range);
            bool found[CG_MAX_Q];
            double rank[CG_MAX_Q], cand[CG_MAX_Q];
            for (int j = 0; j < nq; ++j) { found[j] = false; rank[j] = p.q[j] * c_inf; cand[j] = NAN; }   // step 5
            double prevc = 0.0, prev_le = 0.0, mx = 0.0;
            int idx = 0, second_b = -1;
            for (int b = 0; b < B - 1; ++b) {
                double inc;
                if (!irregular_increase(bk_rows + (size_t)b * pitch, ts_bk + (size_t)b * pitch, cnt_bk[b], t0_row[K + b],
                                        inv_row[K + b], t, lo, lopen, wt, inc))
                    continue;
                double c = div_const(inc, range_s, rc_range);
                if (idx == 0) mx = c;
                else if (c > mx) mx = c;
                else if (c < mx) c = mx;
                for (int j = 0; j < nq; ++j) {
                    if (!found[j] && c >= rank[j]) {           // (never true for a NaN rank)
                        found[j] = true;
                        if (idx == 0 && le[b] <= 0) cand[j] = le[b];              // step 6
                        else {
                            double start = 0.0, end = le[b], cc = c, rk = rank[j];   // step 7
                            if (idx > 0) { start = prev_le; cc -= prevc; rk -= prevc; }
                            cand[j] = start + (end - start) * (rk / cc);
                        }
                    }
                }
                prevc = c;
                prev_le = le[b];
                second_b = b;
                ++idx;
            }
            if (!(idx > 0 && mx > c_inf)) {                    // the +Inf bucket is the maximum: observations = c_inf
                done = true;
                valid |= CG_V_QUANTILE;                        // non-empty vector (:369-370)
                for (int j = 0; j < nq; ++j) {
                    const double q = p.q[j];
                    if (isnan(q)) m[j] = NAN;
                    else if (q < 0) m[j] = -INFINITY;
                    else if (q > 1) m[j] = INFINITY;
                    else if (idx < 1) m[j] = NAN;              // fewer than 2 buckets (Appendix A step 2)
                    else if (c_inf == 0) m[j] = NAN;           // step 4
                    else m[j] = found[j] ? cand[j] : le[second_b];   // not found: bound of the bucket below +Inf
                }
            }
        }
    }
    if (!done) {
        // pass 1: which buckets are present, the fixed count of the highest one (the observations)
        int n_present = 0, last_b = -1, second_b = -1;
        double mx = 0.0, obs = 0.0;
        for (int b = 0; b < B; ++b) {
            double inc;
            if (!irregular_increase(bk_rows + (size_t)b * pitch, ts_bk + (size_t)b * pitch, cnt_bk[b], t0_row[K + b],
                                    inv_row[K + b], t, lo, lopen, wt, inc))
                continue;
            double c = div_const(inc, range_s, rc_range);
            if (n_present == 0) mx = c;
            else if (c > mx) mx = c;
            else if (c < mx) c = mx;
            obs = c;
            second_b = last_b;
            last_b = b;
            ++n_present;
        }
        if (n_present > 0) {
            valid |= CG_V_QUANTILE;                                // non-empty vector (:369-370)
            const bool top_inf = isinf(le[last_b]) && le[last_b] > 0;   // highest bound present must be +Inf
            bool need_scan = false;
            bool found[CG_MAX_Q];
            double rank[CG_MAX_Q];
            for (int j = 0; j < nq; ++j) {
                const double q = p.q[j];
                found[j] = true;
                if (isnan(q)) m[j] = NAN;
                else if (q < 0) m[j] = -INFINITY;
                else if (q > 1) m[j] = INFINITY;
                else if (!top_inf || n_present < 2) m[j] = NAN;    // Appendix A step 2
                else if (obs == 0) m[j] = NAN;                     // step 4
                else { found[j] = false; need_scan = true; rank[j] = q * obs; }   // step 5
            }
            if (need_scan) {
                // pass 2: first present bucket below the highest one whose fixed count reaches the rank
                double prevc = 0.0, prev_le = 0.0;
                int idx = 0;
                mx = 0.0;
                for (int b = 0; b < last_b; ++b) {
                    double inc;
                    if (!irregular_increase(bk_rows + (size_t)b * pitch, ts_bk + (size_t)b * pitch, cnt_bk[b], t0_row[K + b],
                                            inv_row[K + b], t, lo, lopen, wt, inc))
                        continue;
                    double c = div_const(inc, range_s, rc_range);
                    if (idx == 0) mx = c;
                    else if (c > mx) mx = c;
                    else if (c < mx) c = mx;
                    bool all = true;
                    for (int j = 0; j < nq; ++j) {
                        if (!found[j] && c >= rank[j]) {
                            found[j] = true;
                            if (idx == 0 && le[b] <= 0) m[j] = le[b];             // step 6
                            else {
                                double start = 0.0, end = le[b], cc = c, rk = rank[j];   // step 7
                                if (idx > 0) { start = prev_le; cc -= prevc; rk -= prevc; }
                                m[j] = start + (end - start) * (rk / cc);
                            }
                        }
                        all = all && found[j];
                    }
                    prevc = c;
                    prev_le = le[b];
                    ++idx;
                    if (all) break;
                }
                for (int j = 0; j < nq; ++j) if (!found[j]) m[j] = le[second_b];   // b == n-1: bound of the bucket below +Inf
            }
        }
    }
    m[nq + CG_M_ERROR_RESPONSES] = err;
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; valid |= CG_V_ERROR_RATE; }
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; valid |= CG_V_LATENCY_AVG; }
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
    m[nq + CG_M_FEEDBACK_COUNT] = fb;
    valid_out = valid;
}

template <typename T>
__global__ void __launch_bounds__(512) irregular_kernel(const __grid_constant__ IrregularArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA, R = K + B;
    T* tile_sc = reinterpret_cast<T*>(smem);
    T* tile_bk = tile_sc + (size_t)K * p.row_pitch;
    int32_t* ts_sc = reinterpret_cast<int32_t*>(smem + a.off_ts);
    int32_t* ts_bk = ts_sc + (size_t)K * p.row_pitch;
    int32_t* s_cnt = reinterpret_cast<int32_t*>(smem + a.off_cnt);      // [K] then [B]
    int32_t* s_t0 = s_cnt + R;                                          // first time stamp of each row
    float*   s_inv = reinterpret_cast<float*>(s_t0 + R);                // samples per ms of each row (search hint)
    double*   s_le    = reinterpret_cast<double*>(smem + a.off_le);
    double*   s_stash = reinterpret_cast<double*>(smem + a.off_stash);
    uint8_t*  s_sv    = smem + a.off_stash_valid;
    uint32_t* s_bits  = reinterpret_cast<uint32_t*>(smem + a.off_bits);
    uint64_t* bar     = reinterpret_cast<uint64_t*>(smem + a.off_bar);

    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) { mbar_init(bar, 1); mbar_init_fence(); }
    for (int b = tid; b < B; b += blockDim.x) s_le[b] = a.le[b];
    __syncthreads();

    const uint32_t sc_bytes = (uint32_t)((size_t)K * p.row_pitch * sizeof(T));
    const uint32_t bk_bytes = (uint32_t)((size_t)B * p.row_pitch * sizeof(T));
    const uint32_t tsc_bytes = (uint32_t)((size_t)K * p.row_pitch * 4);
    const uint32_t tbk_bytes = (uint32_t)((size_t)B * p.row_pitch * 4);
    uint32_t parity = 0;
    const int n_words = (p.n_ticks + 31) >> 5;

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        for (int ver = 0; ver < 2; ++ver) {
            const int64_t pv = pair * 2 + ver;
            if (tid == 0) {
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the tiles were read by generic loads
                mbar_arrive_expect_tx(bar, sc_bytes + bk_bytes + tsc_bytes + tbk_bytes);
                auto copy = [&](void* dst, const void* src, uint32_t bytes) {   // chunks of <= 64 KB
                    for (uint32_t o = 0; o < bytes; o += 65536u)
                        bulk_g2s(reinterpret_cast<char*>(dst) + o, reinterpret_cast<const char*>(src) + o,
                                 min(65536u, bytes - o), bar);
                };
                copy(tile_sc, reinterpret_cast<const char*>(a.scalars) + (size_t)pv * sc_bytes, sc_bytes);
                copy(tile_bk, reinterpret_cast<const char*>(a.buckets) + (size_t)pv * bk_bytes, bk_bytes);
                copy(ts_sc, reinterpret_cast<const char*>(a.ts_scalars) + (size_t)pv * tsc_bytes, tsc_bytes);
                copy(ts_bk, reinterpret_cast<const char*>(a.ts_buckets) + (size_t)pv * tbk_bytes, tbk_bytes);
            }
            for (int r = tid; r < R; r += blockDim.x) {
                int c = r < K ? a.cnt_scalars[pv * K + r] : a.cnt_buckets[pv * B + (r - K)];
                s_cnt[r] = c < 0 ? 0 : (c > p.n_samples ? p.n_samples : c);
            }
            mbar_wait(bar, parity);
            parity ^= 1u;
            for (int r = tid; r < R; r += blockDim.x) {    // search hints (they steer the walk, never its result)
                const int c = s_cnt[r];
                const int32_t* row = ts_sc + (size_t)r * p.row_pitch;
                const int span = c > 1 ? row[c - 1] - row[0] : 0;
                s_t0[r] = c > 0 ? row[0] : 0;
                s_inv[r] = span > 0 ? (float)(c - 1) / (float)span : 0.0f;
            }
            __syncthreads();       // counts and hints visible

            for (int base = (tid & ~31); base < p.n_ticks; base += blockDim.x) {
                const int j = base + lane;
                bool promote = false;
                if (j < p.n_ticks) {
                    double m[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t v;
                    eval_tick_irregular<T>(tile_sc, tile_bk, ts_sc, ts_bk, s_cnt, s_cnt + K, s_t0, s_inv, s_le, p, j, m, v);
                    if (a.metrics) {
                        double* o = a.metrics + ((size_t)pv * p.n_ticks + j) * W;
                        for (int c = 0; c < W; ++c) o[c] = m[c];
                    }
                    if (a.valid) a.valid[(size_t)pv * p.n_ticks + j] = v;
                    const double gm[3] = {p.n_q > 0 ? m[p.gate_q] : NAN, m[p.n_q + CG_M_ERROR_RATE],
                                          m[p.n_q + CG_M_LATENCY_AVG]};
                    if (ver == 0) {
                        s_stash[3 * j + 0] = gm[0]; s_stash[3 * j + 1] = gm[1]; s_stash[3 * j + 2] = gm[2];
                        s_sv[j] = v;
                    } else {
                        const double nm[3] = {s_stash[3 * j + 0], s_stash[3 * j + 1], s_stash[3 * j + 2]};
                        promote = (gate_eval(nm, s_sv[j], gm, v, p) & CG_G_PROMOTE) != 0;
                    }
                }
                if (ver == 1) {
                    const uint32_t word = __ballot_sync(0xffffffffu, promote);
                    if (lane == 0) s_bits[base >> 5] = word;
                }
            }
            __syncthreads();   // tiles fully consumed (and, for ver 1, gate bits complete)
        }
        if (a.gate_bits)
            for (int w = tid; w < n_words; w += blockDim.x) a.gate_bits[pair * n_words + w] = s_bits[w];
        if (tid == 0 && a.decisions) {
            const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
            const int st = a.start_tick ? a.start_tick[pair] : p.start_tick;
            a.decisions[pair] = replay_schedule_bits(s_bits, n_words, hp, st, p);
        }
        __syncthreads();       // s_bits / stash free for the next pair
    }
}

}  // namespace cg
