// cg_general.cuh -- the GENERAL canary kernel: any window length, any tick grid, any
// K/B that fits shared memory.  One CTA walks pairs; for each (pair, version) the whole
// [K+B rows x T samples] tile is brought into shared memory by the TMA bulk-copy engine
// (two cp.async.bulk, completion on an mbarrier), then one thread per tick evaluates
// the six queries of get_model_metrics (mlflow_operator.py:367-410), the derived
// metrics (:387-404), the gate (:419-460); thread 0 replays the step loop (:296-361).
//
// It is the reference implementation of the semantics on the device and the path taken
// by cg_eval_grid (per-tick metric output) and by geometries the streaming kernel
// (cg_stream.cuh) does not cover.
#pragma once
#include "cg_math.cuh"
#include "cg_ptx.cuh"

namespace cg {

struct GeneralArgs {
    const void*    scalars;
    const void*    buckets;
    const double*  le;
    const int32_t* phase;
    const uint8_t* has_previous;   // may be null
    const int32_t* start_tick;     // may be null
    cg_decision*   decisions;      // may be null (grid-only mode)
    uint32_t*      gate_bits;      // may be null
    double*        metrics;        // may be null
    uint8_t*       valid;          // may be null
    int64_t        n_pairs;
    cg_params      p;
    // shared-memory layout (bytes from the base)
    uint32_t off_le, off_stash, off_stash_valid, off_bits, off_bar, smem_bytes;
};

// increase() of one series with the tick's extrapolation factor computed once (f = ext / sampled
// for the unclamped d_start: it depends on the time stamps alone) and one multiply test in front of
// the zero-point division: d_zero = sampled * (first_v / result) < d_start is impossible when
// first_v * sampled exceeds d_start * result by more than the roundings involved.  Same operations
// as extrapolate() whenever the clamp can fire; the same value otherwise.
template <typename T>
CG_DEVINL double series_increase_f(const T* __restrict__ row, const TickGeom& g, double f) {
    double first_v;
    const double result = window_result(row + g.i0, g.n, first_v);
    if (result > 0 && first_v >= 0) {
        const double lhs = first_v * g.sampled, rhs = (g.d_start * result) * (1.0 + 1e-12);
        if (!(lhs >= rhs && lhs < 1e300 && rhs > 1e-290)) return extrapolate(result, first_v, g);
    }
    return result * f;
}

// Everything get_model_metrics returns for one (pair, version, tick), from a resident tile.
template <typename T>
__device__ void eval_tick_resident(const T* __restrict__ sc_rows, const T* __restrict__ bk_rows,
                                   const double* __restrict__ le, const cg_params& p,
                                   const TickGeom& g, double* __restrict__ m, uint8_t& valid_out) {
    const int nq = p.n_q;
    uint8_t valid = 0;
    double sums[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    unsigned have = 0;
    double f = 0.0;               // Appendix A steps 5-6 without the clamp, once per tick
    if (g.n >= 2) {
        double ext = g.sampled;
        ext += (g.d_start < g.thr) ? g.d_start : g.avg / 2;
        ext += g.end_term;
        f = ext / g.sampled;
    }
    if (g.n >= 2) {   // Q2..Q6: sum(increase(...)) or on() vector(0)  (:375,:383,:393,:397,:410)
        for (int k = 0; k < p.n_scalar; ++k) {
            const double inc = series_increase_f(sc_rows + (size_t)k * p.row_pitch, g, f);
            const unsigned role = p.scalar_role[k];
#pragma unroll
            for (int r = 0; r < 5; ++r)
                if (role & (1u << r)) {
                    sums[r] = (have & (1u << r)) ? sums[r] + inc : inc;
                    have |= 1u << r;
                }
        }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];

    // Q1: histogram_quantile(q, sum(rate(..._bucket[r])) by (le))  (:367)
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    const int B = p.n_buckets;
    if (B > 0 && g.n >= 2) {
        valid |= CG_V_QUANTILE;                       // non-empty vector (:369-370)
        const double range_s = (double)p.window_ms / 1000.0, rc_range = 1.0 / range_s;
        const bool top_inf = isinf(le[B - 1]) && le[B - 1] > 0;
        bool need_scan = false;
        double rank[CG_MAX_Q];
        bool found[CG_MAX_Q];
        for (int j = 0; j < nq; ++j) {
            const double q = p.q[j];
            found[j] = true;
            if (isnan(q)) m[j] = NAN;
            else if (q < 0) m[j] = -INFINITY;
            else if (q > 1) m[j] = INFINITY;
            else if (!top_inf || B < 2) m[j] = NAN;   // Appendix A step 2
            else { found[j] = false; need_scan = true; }
        }
        if (need_scan) {
            // One pass in the usual case: the +Inf bucket first, speculating that it holds the
            // observations (no lower bucket exceeds it after the monotonic fix, step 3-4); checked below.
            const double c_inf = div_const(series_increase_f(bk_rows + (size_t)(B - 1) * p.row_pitch, g, f), range_s, rc_range);
            double cand[CG_MAX_Q];
            bool hit[CG_MAX_Q];
            for (int j = 0; j < nq; ++j) { rank[j] = p.q[j] * c_inf; cand[j] = NAN; hit[j] = found[j]; }   // step 5
            double prevc = 0.0, mx = 0.0;
            for (int b = 0; b < B - 1; ++b) {
                double c = div_const(series_increase_f(bk_rows + (size_t)b * p.row_pitch, g, f), range_s, rc_range);
                if (b == 0) mx = c;
                else if (c > mx) mx = c;
                else if (c < mx) c = mx;
                for (int j = 0; j < nq; ++j) {
                    if (!hit[j] && c >= rank[j]) {
                        hit[j] = true;
                        if (b == 0 && le[0] <= 0) cand[j] = le[0];            // step 6
                        else {
                            double start = 0.0, end = le[b], cc = c, rk = rank[j];   // step 7
                            if (b > 0) { start = le[b - 1]; cc -= prevc; rk -= prevc; }
                            cand[j] = start + (end - start) * (rk / cc);
                        }
                    }
                }
                prevc = c;
            }
            if (!(mx > c_inf)) {      // the +Inf bucket is the maximum: observations = c_inf
                for (int j = 0; j < nq; ++j)
                    if (!found[j]) m[j] = (c_inf == 0) ? NAN : hit[j] ? cand[j] : le[B - 2];   // step 4; found; b == n-1
            } else {
                // pass 1: observations = last count after the monotonic fix (step 3-4)
                double mx = 0.0, obs = 0.0;
                for (int b = 0; b < B; ++b) {
                    double c = div_const(series_increase_f(bk_rows + (size_t)b * p.row_pitch, g, f), range_s, rc_range);
                    if (b == 0) mx = c;
                    else if (c > mx) mx = c;
                    else if (c < mx) c = mx;
                    obs = c;
                }
                if (obs == 0) {
                    for (int j = 0; j < nq; ++j) if (!found[j]) { m[j] = NAN; found[j] = true; }
                } else {
                    for (int j = 0; j < nq; ++j) rank[j] = p.q[j] * obs;          // step 5
                    // pass 2: first bucket in [0, B-2] whose fixed count reaches the rank
                    double prevc = 0.0;
                    mx = 0.0;
                    for (int b = 0; b < B - 1; ++b) {
                        double c = div_const(series_increase_f(bk_rows + (size_t)b * p.row_pitch, g, f), range_s, rc_range);
                        if (b == 0) mx = c;
                        else if (c > mx) mx = c;
                        else if (c < mx) c = mx;
                        bool all = true;
                        for (int j = 0; j < nq; ++j) {
                            if (!found[j] && c >= rank[j]) {
                                found[j] = true;
                                if (b == 0 && le[0] <= 0) m[j] = le[0];           // step 6
                                else {
                                    double start = 0.0, end = le[b], cc = c, rk = rank[j];   // step 7
                                    if (b > 0) { start = le[b - 1]; cc -= prevc; rk -= prevc; }
                                    m[j] = start + (end - start) * (rk / cc);
                                }
                            }
                            all = all && found[j];
                        }
                        prevc = c;
                        if (all) break;
                    }
                    for (int j = 0; j < nq; ++j) if (!found[j]) m[j] = le[B - 2];  // b == n-1
                }
            }
        }
    }
    m[nq + CG_M_ERROR_RESPONSES] = err;                                    // :378
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; valid |= CG_V_ERROR_RATE; }   // :387-390
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; valid |= CG_V_LATENCY_AVG; }  // :401-404
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;                                  // :407
    m[nq + CG_M_FEEDBACK_COUNT] = fb;                                      // :413
    valid_out = valid;
}

template <typename T>
__global__ void __launch_bounds__(512) general_kernel(const __grid_constant__ GeneralArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    T* tile_sc = reinterpret_cast<T*>(smem);
    T* tile_bk = tile_sc + (size_t)K * p.row_pitch;
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
    uint32_t parity = 0;
    const int n_words = (p.n_ticks + 31) >> 5;

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        for (int ver = 0; ver < 2; ++ver) {
            const int64_t pv = pair * 2 + ver;
            if (tid == 0) {
                mbar_arrive_expect_tx(bar, sc_bytes + bk_bytes);
                const char* gs = reinterpret_cast<const char*>(a.scalars) + (size_t)pv * sc_bytes;
                const char* gb = reinterpret_cast<const char*>(a.buckets) + (size_t)pv * bk_bytes;
                // chunks of <= 64 KB keep each bulk request modest
                for (uint32_t o = 0; o < sc_bytes; o += 65536u)
                    bulk_g2s(reinterpret_cast<char*>(tile_sc) + o, gs + o, min(65536u, sc_bytes - o), bar);
                for (uint32_t o = 0; o < bk_bytes; o += 65536u)
                    bulk_g2s(reinterpret_cast<char*>(tile_bk) + o, gb + o, min(65536u, bk_bytes - o), bar);
            }
            const int32_t phase = a.phase[pv];
            mbar_wait(bar, parity);
            parity ^= 1u;

            for (int base = (tid & ~31); base < p.n_ticks; base += blockDim.x) {
                const int j = base + lane;
                bool promote = false;
                if (j < p.n_ticks) {
                    const TickGeom g = tick_geom(p, phase, j);
                    double m[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t v;
                    eval_tick_resident<T>(tile_sc, tile_bk, s_le, p, g, m, v);
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
            __syncthreads();   // tile fully consumed (and, for ver 1, gate bits complete)
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
