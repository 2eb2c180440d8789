// cg_rows.cuh -- the ROW-STREAMING general kernel: any window length, any tick grid, any scrape
// period (what cg_general.cuh covers), but the series rows pass through a small shared-memory ring
// instead of sitting there as one [K+B rows x T samples] tile.
//
// Why: the resident tile of BASELINE.json's layout is 130 KB, so one CTA fits an SM, its 12 warps
// load, wait, compute, and HBM idles meanwhile (8 % of the copy peak); larger layouts do not fit at
// all.  Here a ring slot holds a few rows (15 KB), a CTA needs about 55 KB whatever K + B is, two
// CTAs of 12 warps fit an SM by registers, and the TMA bulk-copy engine refills a slot while the
// others are read.  Measured on configs[3]: 843 GB/s against 526 GB/s (13 % / 8 % of the copy
// peak): the thread-per-tick scan itself (about 200 k warp instructions per pair) is what limits it.
//
//   * One thread per tick (a CTA covers up to blockDim.x ticks per pass over the pair's rows; longer
//     grids take more passes).  The thread keeps the state of its tick's scan in registers: the
//     window geometry (tick_geom), five role sums, the rank(s), the running maximum of the bucket
//     rates (monotonic fix of histogram_quantile) and the captured quantile(s).
//   * Row order per (pair, version): the K scalar rows, the +Inf bucket row, bucket rows 0..B-2 --
//     the +Inf bucket first so that the rank is known before the search (speculating that it holds
//     the observations; a tick where a lower bucket exceeds it is re-evaluated from global memory
//     with the two-pass scan of the resident-tile kernel, eval_tick_resident).
//   * The CTA is its own producer: after a __syncthreads at the end of a slot, thread 0 issues the
//     bulk copies (cp.async.bulk + mbarrier complete_tx) for the slot use that comes n_slots later.
//     (Letting the warp that leaves a slot last issue them, with no CTA barrier, was measured: no gain.)
//
// Same operations in the same order as eval_tick_resident (mlflow_operator.py:367-410, :387-404,
// :419-460, :296-361): results are bit-identical to the other kernels and to the oracle.
#pragma once
#include "cg_general.cuh"

namespace cg {

struct RowsArgs {
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
    int32_t  n_slots;              // ring depth
    int32_t  rows_per_slot;
    int32_t  uses_per_ver;         // slot uses for the K+B rows of one version
    int32_t  n_passes;             // passes over a pair's rows: ceil(n_ticks / blockDim.x)
    uint32_t row_bytes, slot_bytes;
    uint32_t off_bars, off_le, off_stash, off_stash_valid, off_bits, smem_bytes;   // the ring sits at offset 0
};

// what one thread carries through the rows of one (pair, version) for its tick
template <int MQ>                  // MQ: quantiles the instance has room for (1, or CG_MAX_Q)
struct TickScan {
    double   sums[5];
    double   f;                    // ext / sampled without the zero-point clamp (0 when the window is empty)
    double   c_inf;                // rate of the +Inf bucket: the speculated observations
    double   rank[MQ], cand[MQ];
    double   prevc, mx;
    unsigned have;                 // roles that had a series
    unsigned search;               // quantiles that need the bucket scan
    unsigned hit;                  // ... and have found their bucket
};

// increase() of one series over the tick's window (Appendix A steps 2-6), the window's first sample at
// shared-memory address win_s.  Same operations as series_increase_f; the test in front of the
// zero-point clamp is the streaming kernel's: d_zero = sampled * (first / result) < d_start needs
// first * (sampled / d_start) < result, so first * gchk >= result with gchk a slight under-estimate of
// sampled / d_start proves that the clamp cannot fire (anything else takes extrapolate(), which decides).
// n_warp is the longest window among the warp's ticks: every lane walks that many samples, so the loop
// does not diverge; a lane whose window is shorter re-reads its last sample, which changes nothing (a
// sample equal to its predecessor is no reset).
template <typename T>
__device__ __forceinline__ double window_increase(uint32_t win_s, const TickGeom& g, int n_warp, double f, double gchk) {
    const T* __restrict__ w = reinterpret_cast<const T*>(__cvta_shared_to_generic((size_t)win_s));
    const int last = g.n - 1;
    const double first_v = (double)w[0];
    double result = (double)w[last] - first_v, prev = first_v;
#pragma unroll 1
    for (int i = 1; i < n_warp; ++i) {
        const double v = (double)w[min(i, last)];
        if (v < prev) result += prev;      // counter reset
        prev = v;
    }
    if (!(first_v * gchk >= result)) return extrapolate(result, first_v, g);
    return result * f;
}

// The same when the warp has already looked for counter resets in the stretch of the row that its 32
// windows span (one compare per lane and a ballot): `resets` = the reset flags of this lane's window.
// Without one the walk is not needed: result = last - first.
template <typename T>
__device__ __forceinline__ double window_increase_flagged(uint32_t win_s, const TickGeom& g, unsigned resets, double f,
                                                          double gchk) {
    const T* __restrict__ w = reinterpret_cast<const T*>(__cvta_shared_to_generic((size_t)win_s));
    const int last = g.n - 1;
    const double first_v = (double)w[0];
    double result = (double)w[last] - first_v;
    if (resets) {
        double prev = first_v;
#pragma unroll 1
        for (int i = 1; i <= last; ++i) {
            const double v = (double)w[i];
            if (v < prev) result += prev;      // counter reset
            prev = v;
        }
    }
    if (!(first_v * gchk >= result)) return extrapolate(result, first_v, g);
    return result * f;
}

template <typename T, int MQ>
__global__ void __launch_bounds__(384, 2) rows_kernel(const __grid_constant__ RowsArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets, R = K + B, W = p.n_q + CG_M_EXTRA, nq = p.n_q;
    const int G = a.rows_per_slot, n_slots = a.n_slots, pitch = p.row_pitch;
    double*   s_le    = reinterpret_cast<double*>(smem + a.off_le);
    double*   s_stash = reinterpret_cast<double*>(smem + a.off_stash);
    uint8_t*  s_sv    = smem + a.off_stash_valid;
    uint32_t* s_bits  = reinterpret_cast<uint32_t*>(smem + a.off_bits);
    uint64_t* bars    = reinterpret_cast<uint64_t*>(smem + a.off_bars);
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < n_slots; ++s) mbar_init(bars + s, 1);
        mbar_init_fence();
    }
    for (int b = tid; b < B; b += blockDim.x) s_le[b] = a.le[b];
    __syncthreads();
    const uint32_t ring_s = smem_u32(smem), bars_s = ring_s + a.off_bars;
    const size_t ver_bytes_sc = (size_t)K * a.row_bytes, ver_bytes_bk = (size_t)B * a.row_bytes;
    const int n_words = (p.n_ticks + 31) >> 5;
    const double range_s = (double)p.window_ms / 1000.0, rc_range = 1.0 / range_s;
    const bool top_inf = B > 0 && isinf(s_le[B - 1]) && s_le[B - 1] > 0;

    // ---- producer side (thread 0): a cursor over (pair, pass, version, slot use)
    int64_t f_pair = blockIdx.x;
    int f_pass = 0, f_ver = 0, f_use = 0;
    auto fill_next = [&]() {
        if (++f_use == a.uses_per_ver) {
            f_use = 0;
            if (++f_ver == 2) {
                f_ver = 0;
                if (++f_pass == a.n_passes) { f_pass = 0; f_pair += gridDim.x; }
            }
        }
    };
    auto issue_fill = [&](uint32_t slot) {      // thread 0: stream rows [r0, r1) of (f_pair, f_ver) into `slot`
        const int r0 = f_use * G, r1 = min(R, r0 + G);
        const int64_t pv = f_pair * 2 + f_ver;
        const char* gs = reinterpret_cast<const char*>(a.scalars) + pv * ver_bytes_sc;
        const char* gb = reinterpret_cast<const char*>(a.buckets) + pv * ver_bytes_bk;
        const uint32_t dst = ring_s + slot * a.slot_bytes, bar = bars_s + 8u * slot;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // reads of the slot before the async write
        mbar_arrive_expect_tx_s(bar, (uint32_t)(r1 - r0) * a.row_bytes);
        // up to three runs of rows that are contiguous in global memory
        const int s1 = min(r1, K);                                      // scalar rows [r0, s1)
        if (r0 < s1) bulk_g2s_s(dst, gs + (size_t)r0 * a.row_bytes, (uint32_t)(s1 - r0) * a.row_bytes, bar);
        if (B > 0 && r0 <= K && K < r1)                                 // the +Inf bucket row (stream row K)
            bulk_g2s_s(dst + (uint32_t)(K - r0) * a.row_bytes, gb + (size_t)(B - 1) * a.row_bytes, a.row_bytes, bar);
        const int b0 = max(r0, K + 1);                                  // bucket rows b0-K-1 .. r1-K-2
        if (b0 < r1)
            bulk_g2s_s(dst + (uint32_t)(b0 - r0) * a.row_bytes, gb + (size_t)(b0 - K - 1) * a.row_bytes,
                       (uint32_t)(r1 - b0) * a.row_bytes, bar);
    };
    for (int s = 0; s < n_slots; ++s) {          // prologue: fill the whole ring
        if (tid == 0 && f_pair < a.n_pairs) issue_fill((uint32_t)s);
        fill_next();
    }
    uint32_t head = 0, head_parity = 0;

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
#pragma unroll 1
        for (int pass = 0; pass < a.n_passes; ++pass) {
            const int j = pass * (int)blockDim.x + tid;
            const bool active = j < p.n_ticks;
#pragma unroll 1
            for (int ver = 0; ver < 2; ++ver) {
                const int64_t pv = pair * 2 + ver;
                TickGeom g;
                g.n = 0; g.i0 = 0; g.sampled = g.d_start = g.avg = g.thr = g.end_term = 0.0;
                if (active) g = tick_geom(p, a.phase[pv], j);
                const bool live = g.n >= 2;
                const int n_warp = __reduce_max_sync(0xffffffffu, live ? g.n : 0);
                // the stretch of a row that the warp's windows span: samples [base, base + span)
                const int base = __reduce_min_sync(0xffffffffu, live ? g.i0 : 0x7fffffff);
                const int span = __reduce_max_sync(0xffffffffu, live ? g.i0 + g.n : 0) - base;
                const bool flagged = n_warp >= 2 && n_warp <= 32 && span <= 64;     // reset flags fit two ballots
                const int off = live ? g.i0 + 1 - base : 0;                        // flag of sample i sits at bit i - base
                const unsigned wmask = live ? (g.n > 32 ? 0xffffffffu : (1u << (g.n - 1)) - 1u) : 0u;
                double gchk = 0.0;
                TickScan<MQ> sc;
#pragma unroll
                for (int r = 0; r < 5; ++r) sc.sums[r] = 0.0;
                sc.have = 0u; sc.search = 0u; sc.hit = 0u;
                sc.f = 0.0; sc.c_inf = 0.0; sc.prevc = 0.0; sc.mx = 0.0;
#pragma unroll
                for (int q = 0; q < MQ; ++q) { sc.rank[q] = NAN; sc.cand[q] = NAN; }
                if (live) {                       // Appendix A steps 5-6 without the clamp, once per tick
                    double ext = g.sampled;
                    ext += (g.d_start < g.thr) ? g.d_start : g.avg / 2;
                    ext += g.end_term;
                    sc.f = ext / g.sampled;
                    gchk = (g.sampled / g.d_start) * (1.0 - 1e-6);   // (d_start = 0: +Inf, or NaN -> always the exact path)
                    if (top_inf && B >= 2)
                        for (int q = 0; q < nq; ++q) {
                            const double qq = p.q[q];
                            if (!(qq != qq) && !(qq < 0) && !(qq > 1)) sc.search |= 1u << q;
                        }
                }
#pragma unroll 1
                for (int u = 0; u < a.uses_per_ver; ++u) {
                    mbar_wait_s(bars_s + 8u * head, head_parity);
                    const int r0 = u * G, r1 = min(R, r0 + G);
                    if (n_warp >= 2) {
                        uint32_t win_s = ring_s + head * a.slot_bytes + (uint32_t)g.i0 * (uint32_t)sizeof(T);
                        uint32_t row_s = ring_s + head * a.slot_bytes;
#pragma unroll 1
                        for (int r = r0; r < r1; ++r, win_s += a.row_bytes, row_s += a.row_bytes) {
                            double inc;
                            if (flagged) {        // (warp-uniform)
                                const T* __restrict__ x = reinterpret_cast<const T*>(__cvta_shared_to_generic((size_t)row_s));
                                const int i0s = base + lane, i1s = i0s + 32;
                                const unsigned m0 = __ballot_sync(0xffffffffu, lane > 0 && lane < span && x[i0s] < x[i0s - 1]);
                                unsigned m1 = 0u;
                                if (span > 32) m1 = __ballot_sync(0xffffffffu, lane + 32 < span && x[i1s] < x[i1s - 1]);
                                if (!live) continue;
                                const unsigned lo = off >= 32 ? m1 : m0, hi = off >= 32 ? 0u : m1;
                                inc = window_increase_flagged<T>(win_s, g, __funnelshift_r(lo, hi, off & 31) & wmask, sc.f, gchk);
                            } else {
                                if (!live) continue;
                                inc = window_increase<T>(win_s, g, n_warp, sc.f, gchk);
                            }
                            if (r > K) {          // bucket b: monotonic fix, first bucket reaching each rank
                                const int b = r - K - 1;
                                double c = div_const(inc, range_s, rc_range);
                                if (b == 0) sc.mx = c;
                                else if (c > sc.mx) sc.mx = c;
                                else if (c < sc.mx) c = sc.mx;
                                const unsigned todo = sc.search & ~sc.hit;
                                if (todo) {
#pragma unroll
                                    for (int q = 0; q < MQ; ++q) {
                                        if (((todo >> q) & 1u) && c >= sc.rank[q]) {
                                            sc.hit |= 1u << q;
                                            if (b == 0 && s_le[0] <= 0) sc.cand[q] = s_le[0];          // step 6
                                            else {
                                                double start = 0.0, end = s_le[b], cc = c, rk = sc.rank[q];   // step 7
                                                if (b > 0) { start = s_le[b - 1]; cc -= sc.prevc; rk -= sc.prevc; }
                                                sc.cand[q] = start + (end - start) * ddiv_call(rk, cc);
                                            }
                                        }
                                    }
                                }
                                sc.prevc = c;
                            } else if (r == K) {  // the +Inf bucket: speculated observations, ranks (step 5)
                                sc.c_inf = div_const(inc, range_s, rc_range);
#pragma unroll
                                for (int q = 0; q < MQ; ++q)
                                    if (q < nq) sc.rank[q] = p.q[q] * sc.c_inf;
                            } else {              // Q2..Q6: sum(increase(...)) per role (:375-410)
                                const unsigned role = p.scalar_role[r];
#pragma unroll
                                for (int q = 0; q < 5; ++q)
                                    if (role & (1u << q)) {
                                        sc.sums[q] = (sc.have & (1u << q)) ? sc.sums[q] + inc : inc;
                                        sc.have |= 1u << q;
                                    }
                            }
                        }
                    }
                    __syncthreads();              // every thread is done with the slot
                    if (tid == 0 && f_pair < a.n_pairs) issue_fill(head);
                    fill_next();
                    const bool wrap = ++head == (uint32_t)n_slots;
                    head = wrap ? 0u : head;
                    head_parity ^= wrap ? 1u : 0u;
                }

                // ---- the tick's metrics (the tail of eval_tick_resident), gate input, output
                bool promote = false;
                if (active) {
                    double m[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t v = 0;
                    bool exact = false;
                    if (B > 0 && live) {
                        if (sc.search && sc.mx > sc.c_inf) exact = true;     // a lower bucket exceeds +Inf: two-pass scan
                        v |= CG_V_QUANTILE;                                   // non-empty vector (:369-370)
                    }
                    if (exact) {
                        const T* gsc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * pitch;
                        const T* gbk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * pitch;
                        eval_tick_resident<T>(gsc, gbk, s_le, p, g, m, v);
                    } else {
#pragma unroll
                        for (int q = 0; q < MQ; ++q) {
                            if (q >= nq) continue;
                            double r = NAN;
                            if (B > 0 && live) {
                                const double qq = p.q[q];
                                if (qq != qq) r = NAN;
                                else if (qq < 0) r = -INFINITY;
                                else if (qq > 1) r = INFINITY;
                                else if (!top_inf || B < 2) r = NAN;                          // Appendix A step 2
                                else r = (sc.c_inf == 0) ? NAN : ((sc.hit >> q) & 1u) ? sc.cand[q] : s_le[B - 2];   // step 4; found; b == n-1
                            }
                            m[q] = r;
                        }
                        const double total = sc.sums[0], err = sc.sums[1], lat_sum = sc.sums[2], lat_cnt = sc.sums[3];
                        m[nq + CG_M_ERROR_RESPONSES] = err;                                    // :378
                        if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; v |= CG_V_ERROR_RATE; }   // :387-390
                        else m[nq + CG_M_ERROR_RATE] = NAN;
                        if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; v |= CG_V_LATENCY_AVG; }  // :401-404
                        else m[nq + CG_M_LATENCY_AVG] = NAN;
                        m[nq + CG_M_REQUEST_COUNT] = lat_cnt;                                  // :407
                        m[nq + CG_M_FEEDBACK_COUNT] = sc.sums[4];                              // :413
                    }
                    if (a.metrics) {
                        double* o = a.metrics + ((size_t)pv * p.n_ticks + j) * W;
#pragma unroll 1
                        for (int c = 0; c < W; ++c) o[c] = m[c];
                    }
                    if (a.valid) a.valid[(size_t)pv * p.n_ticks + j] = v;
                    const double gm[3] = {nq > 0 ? m[p.gate_q] : NAN, m[nq + CG_M_ERROR_RATE], m[nq + CG_M_LATENCY_AVG]};
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
                    const int base = pass * (int)blockDim.x + (tid & ~31);
                    if (lane == 0 && base < p.n_ticks) s_bits[base >> 5] = word;
                }
            }
        }
        __syncthreads();       // gate bits complete
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
