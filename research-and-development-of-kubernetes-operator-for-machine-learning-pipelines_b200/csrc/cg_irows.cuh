// cg_irows.cuh -- IRREGULAR series (per-series time stamps) with the rows
// passing through a shared-memory ring: what cg_rows.cuh is to cg_general.cuh, this is to
// cg_irregular.cuh.
//
// The resident-tile kernel needs the values AND the time stamps of a whole (pair, version) in shared
// memory (196 KB for BASELINE.json's layout: one CTA of 12 warps per SM, nothing larger fits).  Here a
// ring slot holds a few value rows and their time-stamp rows (TMA bulk copies, one mbarrier per slot,
// thread 0 refills a slot after a __syncthreads), a CTA needs about 60 KB whatever K + B is, and two
// CTAs fit an SM.  Measured on the full layout: 407 GB/s against 362 GB/s for the resident tile.
//
//   * One thread per tick.  It carries through the rows: the extrapolation terms of its last window
//     (WindowTerms: the series of a target are scraped together, so they are reused row after row), the
//     five role sums, and the one-pass bucket scan of eval_tick_irregular -- the +Inf bucket first,
//     speculating that it has output and holds the observations.
//   * Row order per (pair, version): K scalar rows, the +Inf bucket row, bucket rows 0..B-2.
//   * A tick the one-pass scan cannot settle (the +Inf bucket has no output in its window, the highest
//     bound is not +Inf, a lower bucket exceeds the +Inf bucket) is evaluated by eval_tick_irregular
//     itself from global memory: same function, same result as the resident-tile kernel.
//
// Results are bit-identical to cg_irregular.cuh and to the oracle (oracle/promql_oracle.c
// cgo_eval_irregular): the per-series arithmetic IS irregular_increase.
#pragma once
#include "cg_irregular.cuh"

namespace cg {

struct IRowsArgs {
    const void*    scalars;
    const void*    buckets;
    const int32_t* ts_scalars;
    const int32_t* ts_buckets;
    const int32_t* cnt_scalars;
    const int32_t* cnt_buckets;
    const double*  le;
    const uint8_t* has_previous;   // may be null
    const int32_t* start_tick;     // may be null
    cg_decision*   decisions;      // may be null
    uint32_t*      gate_bits;      // may be null
    double*        metrics;        // may be null
    uint8_t*       valid;          // may be null
    int64_t        n_pairs;
    cg_params      p;
    int32_t  n_slots, rows_per_slot, uses_per_ver, n_passes;
    uint32_t vrow_bytes, trow_bytes;          // one row of values / of time stamps
    uint32_t slot_bytes;                      // rows_per_slot * (vrow_bytes + trow_bytes): value rows, then stamp rows
    uint32_t off_bars, off_cnt, off_le, off_stash, off_stash_valid, off_bits, smem_bytes;   // the ring sits at offset 0
};

// one out-of-line copy of the full evaluation for the ticks the one-pass scan cannot settle
template <typename T>
__device__ __noinline__ void eval_tick_irregular_cold(const T* sc_rows, const T* bk_rows, const int32_t* ts_sc,
                                                       const int32_t* ts_bk, const int32_t* cnt, int K, const int32_t* t0_row,
                                                       const float* inv_row, const double* le, const cg_params& p, int tick,
                                                       double* m, uint8_t* valid_out) {
    uint8_t v;
    eval_tick_irregular<T>(sc_rows, bk_rows, ts_sc, ts_bk, cnt, cnt + K, t0_row, inv_row, le, p, tick, m, v);
    *valid_out = v;
}

template <typename T, int MQ>      // MQ: quantiles the instance has room for (1, or CG_MAX_Q)
__global__ void __launch_bounds__(384, 2) irows_kernel(const __grid_constant__ IRowsArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets, R = K + B, W = p.n_q + CG_M_EXTRA, nq = p.n_q;
    const int G = a.rows_per_slot, n_slots = a.n_slots, pitch = p.row_pitch;
    int32_t* s_cnt = reinterpret_cast<int32_t*>(smem + a.off_cnt);      // [K] then [B], memory order
    int32_t* s_t0 = s_cnt + R;                                          // first time stamp of each row
    float*   s_inv = reinterpret_cast<float*>(s_t0 + R);                // samples per ms of each row (search hint)
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
    const size_t ver_v_sc = (size_t)K * a.vrow_bytes, ver_v_bk = (size_t)B * a.vrow_bytes;
    const size_t ver_t_sc = (size_t)K * a.trow_bytes, ver_t_bk = (size_t)B * a.trow_bytes;
    const uint32_t stamps_off = (uint32_t)G * a.vrow_bytes;             // stamp rows follow the value rows of a slot
    const int n_words = (p.n_ticks + 31) >> 5;
    const double range_s = (double)p.window_ms / 1000.0, rc_range = 1.0 / range_s;
    const bool top_inf = B > 0 && isinf(s_le[B - 1]) && s_le[B - 1] > 0;
    const bool lopen = p.left_open != 0;

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
    auto issue_fill = [&](uint32_t slot) {      // thread 0: stream rows [r0, r1) of (f_pair, f_ver), values and stamps
        const int r0 = f_use * G, r1 = min(R, r0 + G);
        const int64_t pv = f_pair * 2 + f_ver;
        const uint32_t dst = ring_s + slot * a.slot_bytes, bar = bars_s + 8u * slot;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // reads of the slot before the async write
        mbar_arrive_expect_tx_s(bar, (uint32_t)(r1 - r0) * (a.vrow_bytes + a.trow_bytes));
        const int s1 = min(r1, K), b0 = max(r0, K + 1);
#pragma unroll
        for (int part = 0; part < 2; ++part) {   // 0: values, 1: time stamps -- up to three contiguous runs each
            const uint32_t rb = part ? a.trow_bytes : a.vrow_bytes;
            const uint32_t d = dst + (part ? stamps_off : 0u);
            const char* gs = part ? reinterpret_cast<const char*>(a.ts_scalars) + pv * ver_t_sc
                                  : reinterpret_cast<const char*>(a.scalars) + pv * ver_v_sc;
            const char* gb = part ? reinterpret_cast<const char*>(a.ts_buckets) + pv * ver_t_bk
                                  : reinterpret_cast<const char*>(a.buckets) + pv * ver_v_bk;
            if (r0 < s1) bulk_g2s_s(d, gs + (size_t)r0 * rb, (uint32_t)(s1 - r0) * rb, bar);
            if (B > 0 && r0 <= K && K < r1) bulk_g2s_s(d + (uint32_t)(K - r0) * rb, gb + (size_t)(B - 1) * rb, rb, bar);
            if (b0 < r1) bulk_g2s_s(d + (uint32_t)(b0 - r0) * rb, gb + (size_t)(b0 - K - 1) * rb, (uint32_t)(r1 - b0) * rb, bar);
        }
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
            const int t = p.tick0_ms + j * p.tick_step_ms, lo = t - p.window_ms;
#pragma unroll 1
            for (int ver = 0; ver < 2; ++ver) {
                const int64_t pv = pair * 2 + ver;
                // sample counts and search hints of this version's rows (memory order), from global memory:
                // the exact path needs them for every row, the scan for the row at hand
                __syncthreads();                  // the previous version's readers are done with them
                for (int r = tid; r < R; r += blockDim.x) {
                    int c = r < K ? a.cnt_scalars[pv * K + r] : a.cnt_buckets[pv * B + (r - K)];
                    c = c < 0 ? 0 : (c > p.n_samples ? p.n_samples : c);
                    const int32_t* row = r < K ? a.ts_scalars + ((size_t)pv * K + r) * pitch
                                               : a.ts_buckets + ((size_t)pv * B + (r - K)) * pitch;
                    const int first = c > 0 ? row[0] : 0, span = c > 1 ? row[c - 1] - first : 0;
                    s_cnt[r] = c;
                    s_t0[r] = first;
                    s_inv[r] = span > 0 ? (float)(c - 1) / (float)span : 0.0f;
                }
                __syncthreads();

                WindowTerms wt;
                wt.n = 0;
                double sums[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
                unsigned have = 0u, found = 0u;
                bool inf_ok = false;
                double c_inf = 0.0, prevc = 0.0, prev_le = 0.0, mx = 0.0;
                double rank[MQ], cand[MQ];
#pragma unroll
                for (int q = 0; q < MQ; ++q) { rank[q] = NAN; cand[q] = NAN; }
                int idx = 0, second_b = -1;
#pragma unroll 1
                for (int u = 0; u < a.uses_per_ver; ++u) {
                    mbar_wait_s(bars_s + 8u * head, head_parity);
                    const int r0 = u * G, r1 = min(R, r0 + G);
                    if (active) {
                        const T* vrow = reinterpret_cast<const T*>(smem + (size_t)head * a.slot_bytes);
                        const int32_t* trow = reinterpret_cast<const int32_t*>(smem + (size_t)head * a.slot_bytes + stamps_off);
#pragma unroll 1
                        for (int r = r0; r < r1; ++r, vrow += pitch, trow += pitch) {
                            const int mr = r < K ? r : (r == K ? R - 1 : r - 1);   // the row's place in memory order
                            if (r > K && !inf_ok) continue;                      // this tick goes to the exact path anyway
                            double inc;
                            if (!irregular_increase(vrow, trow, s_cnt[mr], s_t0[mr], s_inv[mr], t, lo, lopen, wt, inc)) continue;
                            if (r > K) {          // bucket b, present: monotonic fix, first bucket reaching each rank
                                const int b = r - K - 1;
                                double c = div_const(inc, range_s, rc_range);
                                if (idx == 0) mx = c;
                                else if (c > mx) mx = c;
                                else if (c < mx) c = mx;
#pragma unroll
                                for (int q = 0; q < MQ; ++q) {
                                    if (q < nq && !((found >> q) & 1u) && c >= rank[q]) {     // (never true for a NaN rank)
                                        found |= 1u << q;
                                        if (idx == 0 && s_le[b] <= 0) cand[q] = s_le[b];        // step 6
                                        else {
                                            double start = 0.0, end = s_le[b], cc = c, rk = rank[q];   // step 7
                                            if (idx > 0) { start = prev_le; cc -= prevc; rk -= prevc; }
                                            cand[q] = start + (end - start) * ddiv_call(rk, cc);
                                        }
                                    }
                                }
                                prevc = c;
                                prev_le = s_le[b];
                                second_b = b;
                                ++idx;
                            } else if (r == K) {  // the +Inf bucket has output: speculated observations, ranks (step 5)
                                if (top_inf) {
                                    inf_ok = true;
                                    c_inf = div_const(inc, range_s, rc_range);
#pragma unroll
                                    for (int q = 0; q < MQ; ++q)
                                        if (q < nq) rank[q] = p.q[q] * c_inf;
                                }
                            } else {              // Q2..Q6: sum(increase(...)) or on() vector(0)
                                const unsigned role = p.scalar_role[r];
#pragma unroll
                                for (int q = 0; q < 5; ++q)
                                    if (role & (1u << q)) {
                                        sums[q] = (have & (1u << q)) ? sums[q] + inc : inc;
                                        have |= 1u << q;
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

                // ---- the tick's metrics, gate input, output
                bool promote = false;
                if (active) {
                    double m[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t v = 0;
                    const bool settled = B == 0 || (inf_ok && !(idx > 0 && mx > c_inf));
                    if (!settled) {
                        const T* gsc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * pitch;
                        const T* gbk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * pitch;
                        const int32_t* gts = a.ts_scalars + (size_t)pv * K * pitch;
                        const int32_t* gtb = a.ts_buckets + (size_t)pv * B * pitch;
                        eval_tick_irregular_cold<T>(gsc, gbk, gts, gtb, s_cnt, K, s_t0, s_inv, s_le, p, j, m, &v);
                    } else {
#pragma unroll
                        for (int q = 0; q < MQ; ++q) {
                            if (q >= nq) continue;
                            double r = NAN;
                            if (B > 0) {
                                const double qq = p.q[q];
                                if (qq != qq) r = NAN;
                                else if (qq < 0) r = -INFINITY;
                                else if (qq > 1) r = INFINITY;
                                else if (idx < 1) r = NAN;              // fewer than 2 buckets (Appendix A step 2)
                                else if (c_inf == 0) r = NAN;           // step 4
                                else r = ((found >> q) & 1u) ? cand[q] : s_le[second_b];   // not found: bound of the bucket below +Inf
                            }
                            m[q] = r;
                        }
                        if (B > 0) v |= CG_V_QUANTILE;                  // non-empty vector (:369-370)
                        const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3];
                        m[nq + CG_M_ERROR_RESPONSES] = err;
                        if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; v |= CG_V_ERROR_RATE; }
                        else m[nq + CG_M_ERROR_RATE] = NAN;
                        if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; v |= CG_V_LATENCY_AVG; }
                        else m[nq + CG_M_LATENCY_AVG] = NAN;
                        m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
                        m[nq + CG_M_FEEDBACK_COUNT] = sums[4];
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
