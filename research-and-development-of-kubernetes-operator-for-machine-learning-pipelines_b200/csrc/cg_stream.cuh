// cg_stream.cuh -- the STREAMING canary kernel: the fast path for the reference's own
// time constants (60 s window = 4 scrape periods, 10 s evaluation grid = 2/3 period;
// mlflow_operator.py:363, :294 with Prometheus' usual 15 s scrapes).
//
// Layout of the work
//   * One CTA = CW consumer warps + 1 producer warp, persistent over model pairs.
//   * The producer's elected lane streams series ROWS through a shared-memory ring with
//     the TMA bulk-copy engine (cp.async.bulk + mbarrier complete_tx): per (pair,
//     version) the K scalar rows, then the +Inf bucket row, then bucket rows 0..B-2 in
//     groups of 4.  Every sample is read from HBM exactly once; the ring keeps several
//     KB per CTA in flight so the copy engine never waits for the math.
//   * A consumer thread owns one PERIOD of the tick grid: 3 consecutive ticks whose
//     4-sample windows live in 6 consecutive samples (three 128-bit shared loads per
//     row).  Which of the 6 each tick uses depends only on the series phase, so it is a
//     CTA-uniform pattern (O1, O2) and the row loop is instantiated per pattern.
//   * Per row and tick: reset-aware window delta, zero-point-clamp test, extrapolation
//     factor, then the histogram search state (first bucket reaching the rank) and a
//     monotonicity flag.  After the last row each thread finishes histogram_quantile,
//     the derived ratios (:387-404) and, once both versions are done, the gate
//     (:419-460).  Gate bits are OR-ed into shared memory and thread 0 replays the step
//     loop (:296-361).
//   * Anything the fast path cannot decide exactly (series edges, coincident phases,
//     non-monotonic bucket rates, a search landing within an ulp of the rank) is
//     re-evaluated for that tick by the whole warp from global memory with the plain
//     sequential semantics (coop_eval_tick).  Results are bit-identical either way.
#pragma once
#include "cg_math.cuh"
#include "cg_ptx.cuh"

namespace cg {

constexpr int kRowsPerSlot = 4;
constexpr int kG = 3;            // ticks per thread (one period of the tick grid)
constexpr unsigned kFull = 0xffffffffu;

struct StreamArgs {
    const void*    scalars;
    const void*    buckets;
    const double*  le;
    const int32_t* phase;
    const uint8_t* has_previous;   // may be null
    const int32_t* start_tick;     // may be null
    cg_decision*   decisions;
    uint32_t*      gate_bits;      // may be null
    double*        metrics;        // may be null
    uint8_t*       valid;          // may be null
    int64_t        n_pairs;
    cg_params      p;
    int32_t  n_groups;             // ceil(n_ticks / 3)
    int32_t  consumer_warps;
    int32_t  n_slots;
    uint32_t slot_bytes;           // kRowsPerSlot * row_pitch * sizeof(elem)
    uint32_t off_le, off_bits, off_bars, smem_bytes;
};

// ---- 6-sample register tile ---------------------------------------------------------
template <typename T> struct Tile6;
template <> struct Tile6<double> {
    static __device__ __forceinline__ void load(const double* p, double (&x)[6]) {
        const double2* q = reinterpret_cast<const double2*>(p);
        const double2 a = q[0], b = q[1], c = q[2];
        x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y; x[4] = c.x; x[5] = c.y;
    }
};
template <> struct Tile6<float> {
    static __device__ __forceinline__ void load(const float* p, double (&x)[6]) {
        const float2* q = reinterpret_cast<const float2*>(p);
        const float2 a = q[0], b = q[1], c = q[2];
        x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y; x[4] = c.x; x[5] = c.y;
    }
};

// The three 4-sample windows a thread can need, at tile offsets 0, 1, 2 (Appendix A step 2:
// last - first, then + the value before each counter reset, in time order).  Counter resets are
// rare, so the corrections are skipped when no lane of the warp saw one.  All 32 lanes call this.
template <bool NEED_W2>
__device__ __forceinline__ void windows3(const double (&x)[6], double& w0, double& w1, double& w2) {
    w0 = x[3] - x[0];
    w1 = x[4] - x[1];
    w2 = NEED_W2 ? x[5] - x[2] : 0.0;
    const bool r1 = x[1] < x[0], r2 = x[2] < x[1], r3 = x[3] < x[2], r4 = x[4] < x[3],
               r5 = NEED_W2 && x[5] < x[4];
    if (__any_sync(kFull, r1 | r2 | r3 | r4 | r5)) {
        if (r1) w0 += x[0];
        if (r2) { w0 += x[1]; w1 += x[1]; }
        if (r3) { w0 += x[2]; w1 += x[2]; if (NEED_W2) w2 += x[2]; }
        if (r4) { w1 += x[3]; if (NEED_W2) w2 += x[3]; }
        if (r5) w2 += x[4];
    }
}

// Per-thread state kept in REGISTERS across the rows of one (pair, version): only what every
// row touches.
template <int NQ>
struct Hot {
    double f[kG];        // ext/sampled with the unclamped d_start (Appendix A step 5-6)
    double gchk[kG];     // first_v * gchk >= result  ==>  the zero-point clamp cannot fire
    double prev[kG];     // running max of the bucket increases seen so far (the monotonic fix)
    double x0[kG][NQ];   // search threshold in increase space (rank * range_s); NaN once found
};

// Per-thread state touched rarely: lives in local memory (indexed dynamically on purpose).
template <int NQ>
struct Cold {
    double d_start, end_term;        // for the exact clamp path
    double obs_inc;                  // increase of the +Inf bucket
    double cb[NQ], cprev[NQ];        // fixed counts at / before the bucket that reaches the rank
    int    bstar[NQ];
    double sums[5];                  // Q3, Q2, Q4, Q5, Q6 in role-bit order
    unsigned have;
    unsigned flags;                  // kRegular | kFallback
    int    tick;
};

constexpr unsigned kRegular = 1u, kFallback = 2u;

// ---- warp-cooperative exact evaluation of one (pair, version, tick) from global memory
template <typename T>
__device__ __noinline__ void coop_eval_tick(const StreamArgs& a, int64_t pv, int32_t phase, int tick, int lane,
                               const double* __restrict__ s_le, double* m, uint8_t& valid_out) {
    const cg_params& p = a.p;
    const int nq = p.n_q, K = p.n_scalar, B = p.n_buckets;
    const TickGeom g = tick_geom(p, phase, tick);
    const T* sc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * p.row_pitch;
    const T* bk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * p.row_pitch;
    uint8_t valid = 0;
    double sums[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    unsigned have = 0;
    if (g.n >= 2) {
        double mine = 0.0;
        if (lane < K) mine = series_increase(sc + (size_t)lane * p.row_pitch, g);
        for (int k = 0; k < K; ++k) {
            const double inc = __shfl_sync(kFull, mine, k);
            const unsigned role = p.scalar_role[k];
#pragma unroll
            for (int r = 0; r < 5; ++r)
                if (role & (1u << r)) {
                    sums[r] = (have & (1u << r)) ? sums[r] + inc : inc;
                    have |= 1u << r;
                }
        }
    }
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    if (B > 0 && g.n >= 2) {
        valid |= CG_V_QUANTILE;
        const double range_s = (double)p.window_ms / 1000.0;
        const bool top_inf = isinf(s_le[B - 1]) && s_le[B - 1] > 0;
        bool scan = false;
        bool found[CG_MAX_Q];
        for (int j = 0; j < nq; ++j) {
            const double q = p.q[j];
            found[j] = true;
            if (isnan(q)) m[j] = NAN;
            else if (q < 0) m[j] = -INFINITY;
            else if (q > 1) m[j] = INFINITY;
            else if (!top_inf || B < 2) m[j] = NAN;
            else { found[j] = false; scan = true; }
        }
        if (scan) {
            const int chunks = (B + 31) >> 5;
            // pass 1: observations = last count after the running-max fix
            double carry = -INFINITY, obs = 0.0;
            for (int c = 0; c < chunks; ++c) {
                const int b = (c << 5) + lane;
                double v = -INFINITY;
                if (b < B) v = series_increase(bk + (size_t)b * p.row_pitch, g) / range_s;
                for (int d = 1; d < 32; d <<= 1) {
                    const double o = __shfl_up_sync(kFull, v, d);
                    if (lane >= d && o > v) v = o;
                }
                if (carry > v) v = carry;
                const int last = min(31, B - 1 - (c << 5));
                carry = __shfl_sync(kFull, v, last);
                obs = carry;
            }
            if (obs == 0) {
                for (int j = 0; j < nq; ++j) if (!found[j]) { m[j] = NAN; found[j] = true; }
            } else {
                double rank[CG_MAX_Q];
                for (int j = 0; j < nq; ++j) rank[j] = p.q[j] * obs;
                carry = -INFINITY;
                for (int c = 0; c < chunks; ++c) {
                    const int b = (c << 5) + lane;
                    double v = -INFINITY;
                    if (b < B) v = series_increase(bk + (size_t)b * p.row_pitch, g) / range_s;
                    for (int d = 1; d < 32; d <<= 1) {
                        const double o = __shfl_up_sync(kFull, v, d);
                        if (lane >= d && o > v) v = o;
                    }
                    if (carry > v) v = carry;
                    for (int j = 0; j < nq; ++j) {
                        if (found[j]) continue;
                        const unsigned hit = __ballot_sync(kFull, b < B - 1 && v >= rank[j]);
                        if (hit) {
                            const int src = __ffs(hit) - 1;
                            const int bb = (c << 5) + src;
                            const double cbv = __shfl_sync(kFull, v, src);
                            const double up = __shfl_sync(kFull, v, src > 0 ? src - 1 : 0);
                            const double cpv = src > 0 ? up : carry;
                            found[j] = true;
                            if (bb == 0 && s_le[0] <= 0) m[j] = s_le[0];
                            else {
                                double start = 0.0, end = s_le[bb], cc = cbv, rk = rank[j];
                                if (bb > 0) { start = s_le[bb - 1]; cc -= cpv; rk -= cpv; }
                                m[j] = start + (end - start) * (rk / cc);
                            }
                        }
                    }
                    const int last = min(31, B - 1 - (c << 5));
                    carry = __shfl_sync(kFull, v, last);
                }
                for (int j = 0; j < nq; ++j) if (!found[j]) m[j] = s_le[B - 2];
            }
        }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];
    m[nq + CG_M_ERROR_RESPONSES] = err;
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = err / total; valid |= CG_V_ERROR_RATE; }
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = lat_sum / lat_cnt; valid |= CG_V_LATENCY_AVG; }
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
    m[nq + CG_M_FEEDBACK_COUNT] = fb;
    valid_out = valid;
}

// Rare path of a fast row: the zero-point clamp may fire -> exact Appendix A step 4-6 for a
// regular 4-sample window (sampled interval = 3 scrape periods).
static __device__ __noinline__ double slow_increase(double result, double first_v, double d_start,
                                                    double end_term, double sampled) {
    TickGeom g;
    g.i0 = 0; g.n = 4;
    g.sampled = sampled;
    g.d_start = d_start;
    g.avg = sampled / 3.0;
    g.thr = g.avg * 1.1;
    g.end_term = end_term;
    return extrapolate(result, first_v, g);
}

static __device__ __noinline__ TickGeom tick_geom_call(const cg_params& p, int32_t phase, int tick) {
    return tick_geom32(p, phase, tick);
}

// ---- the rows of one (pair, version) -------------------------------------------------
// O1 / O2: tile offsets of the windows of tick 1 and tick 2 (tick 0 is at offset 0).  The three
// patterns that occur are (0,1), (1,1), (1,2); the choice depends only on the series phase.
template <typename T, int NQ, int O1, int O2>
__device__ __forceinline__ void consume_pv(const StreamArgs& a, unsigned char* ring, uint64_t* full_bar,
                                           uint64_t* empty_bar, uint32_t& it, int base, Hot<NQ>& h,
                                           Cold<NQ>* cold, bool active, int lane) {
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets;
    const size_t pitch = p.row_pitch;
    const double range_s = (double)p.window_ms / 1000.0;
    const double sampled = (double)(3 * p.period_ms) / 1000.0;
    const int n_sc = (K + kRowsPerSlot - 1) / kRowsPerSlot;
    const int n_use = n_sc + (B > 0 ? 1 + (B - 1 + kRowsPerSlot - 1) / kRowsPerSlot : 0);

#define CG_ROW_INCREASES(ROWPTR)                                                                   \
    double x[6];                                                                                   \
    Tile6<T>::load((ROWPTR) + base, x);                                                            \
    double w0, w1, w2;                                                                             \
    windows3<O2 == 2>(x, w0, w1, w2);                                                              \
    const double res[kG] = {w0, O1 == 0 ? w0 : w1, O2 == 1 ? w1 : w2};                             \
    const double fst[kG] = {x[0], x[O1], x[O2]};                                                   \
    double inc[kG];                                                                                \
    bool slow[kG];                                                                                 \
    _Pragma("unroll") for (int s = 0; s < kG; ++s) {                                               \
        slow[s] = !(fst[s] * h.gchk[s] >= res[s]);                                                 \
        inc[s] = res[s] * h.f[s];                                                                  \
    }                                                                                              \
    if (__any_sync(kFull, slow[0] | slow[1] | slow[2])) {   /* rare: zero-point clamp may fire */  \
        _Pragma("unroll") for (int s = 0; s < kG; ++s)                                             \
            if (slow[s]) inc[s] = slow_increase(res[s], fst[s], cold[s].d_start, cold[s].end_term, sampled); \
    }

    for (int u = 0; u < n_use; ++u) {
        const uint32_t slot = it % (uint32_t)a.n_slots;
        mbar_wait(full_bar + slot, (it / (uint32_t)a.n_slots) & 1u);
        const T* rows = reinterpret_cast<const T*>(ring + (size_t)slot * a.slot_bytes);
        if (!(p.reserved & 1)) {   // (reserved bit 0: debugging aid, skip the row math)
            // every lane runs the row code (neutral factors for lanes without regular ticks)
            if (u < n_sc) {
                // scalar counter rows: sum(increase(...)) per role, in series order
                const int k0 = u * kRowsPerSlot, nr = min(kRowsPerSlot, K - k0);
#pragma unroll 1
                for (int k = 0; k < nr; ++k) {
                    CG_ROW_INCREASES(rows + k * pitch)
                    const unsigned role = p.scalar_role[k0 + k];
#pragma unroll
                    for (int s = 0; s < kG; ++s)
#pragma unroll
                        for (int q = 0; q < 5; ++q)
                            if (role & (1u << q)) {
                                cold[s].sums[q] = (cold[s].have & (1u << q)) ? cold[s].sums[q] + inc[s] : inc[s];
                                cold[s].have |= 1u << q;
                            }
                }
            } else if (u == n_sc) {
                // the +Inf bucket first: observations -> rank -> search threshold
                CG_ROW_INCREASES(rows)
#pragma unroll
                for (int s = 0; s < kG; ++s) {
                    cold[s].obs_inc = inc[s];
                    const double obs = inc[s] / range_s;
#pragma unroll
                    for (int q = 0; q < NQ; ++q) h.x0[s][q] = active ? (p.q[q] * obs) * range_s : NAN;
                }
            } else {
                const int b0 = (u - n_sc - 1) * kRowsPerSlot, nr = min(kRowsPerSlot, B - 1 - b0);
#pragma unroll 1
                for (int k = 0; k < nr; ++k) {
                    CG_ROW_INCREASES(rows + k * pitch)
                    double c[kG];
                    bool any_hit = false;
#pragma unroll
                    for (int s = 0; s < kG; ++s) {
                        c[s] = inc[s] < h.prev[s] ? h.prev[s] : inc[s];   // monotonic fix (Appendix A step 3)
#pragma unroll
                        for (int q = 0; q < NQ; ++q) any_hit |= c[s] >= h.x0[s][q];
                    }
                    if (__any_sync(kFull, any_hit)) {       // some lane's first bucket reaching the rank
#pragma unroll
                        for (int s = 0; s < kG; ++s)
#pragma unroll
                            for (int q = 0; q < NQ; ++q)
                                if (c[s] >= h.x0[s][q]) {
                                    cold[s].bstar[q] = b0 + k;
                                    cold[s].cb[q] = c[s];
                                    cold[s].cprev[q] = h.prev[s];
                                    h.x0[s][q] = NAN;       // never true again
                                }
                    }
#pragma unroll
                    for (int s = 0; s < kG; ++s) h.prev[s] = c[s];
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_bar + slot);
        ++it;
    }
#undef CG_ROW_INCREASES
}

// Everything one consumer thread does for one (pair, version): tick geometry, the row stream,
// the per-tick finish and the exact fallback.  Kept out of line so that its registers are not
// live across the two versions of a pair.  -> m[s][*], v[s] for the thread's kG ticks.
template <typename T, int NQ>
static __device__ __forceinline__ void run_version(const StreamArgs& a, unsigned char* ring, uint64_t* full_bar,
                                                uint64_t* empty_bar, const double* s_le, uint32_t& it,
                                                int64_t pv, int grp, bool has_ticks, int lane,
                                                double (*m)[CG_MAX_Q + CG_M_EXTRA], uint8_t* v) {
    const cg_params& p = a.p;
    const int B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    const double range_s = (double)p.window_ms / 1000.0;
    const int32_t phase = a.phase[pv];
    {
            Hot<NQ> h;
            Cold<NQ> cold[kG];
            int base = 0;
            // ---- per-thread tick geometry and the CTA-uniform window pattern
            int off[kG];
#pragma unroll
            for (int s = 0; s < kG; ++s) {
                Cold<NQ>& c = cold[s];
                c.tick = grp * kG + s;
                c.flags = 0; c.have = 0; c.obs_inc = 0.0; c.d_start = 0.0; c.end_term = 0.0;
                h.prev[s] = -INFINITY; h.f[s] = 0.0; h.gchk[s] = INFINITY;
#pragma unroll
                for (int q = 0; q < 5; ++q) c.sums[q] = 0.0;
#pragma unroll
                for (int q = 0; q < NQ; ++q) { h.x0[s][q] = NAN; c.cb[q] = c.cprev[q] = 0.0; c.bstar[q] = -1; }
                off[s] = -1;
                if (has_ticks && c.tick < p.n_ticks) {
                    const TickGeom g = tick_geom_call(p, phase, c.tick);
                    if (s == 0) base = g.i0;
                    off[s] = g.i0 - base;
                    if (g.n == 4) {
                        c.flags |= kRegular;
                        double ext = g.sampled;
                        ext += (g.d_start < g.thr) ? g.d_start : g.avg / 2;
                        ext += g.end_term;
                        h.f[s] = ext / g.sampled;
                        h.gchk[s] = (g.sampled / g.d_start) * (1.0 - 9.094947017729282e-13);   // 1 - 2^-40
                        c.d_start = g.d_start;
                        c.end_term = g.end_term;
                    }
                }
            }
            // A thread is self-regular if every tick it owns has a 4-sample window starting at an
            // even sample; the window pattern (O1, O2) of such threads depends only on the phase.
            bool self_ok = has_ticks && cold[0].tick < p.n_ticks && (cold[0].flags & kRegular) &&
                           (base & 1) == 0 && base >= 0;
            bool full_group = true;
#pragma unroll
            for (int s = 1; s < kG; ++s) {
                if (has_ticks && cold[s].tick < p.n_ticks) { if (!(cold[s].flags & kRegular)) self_ok = false; }
                else full_group = false;
            }
            int O1 = 0, O2 = 1;
            {   // the warp adopts the pattern of its first self-regular lane with a full group
                const unsigned cand = __ballot_sync(kFull, self_ok && full_group);
                const int src = cand ? __ffs(cand) - 1 : 0;
                O1 = __shfl_sync(kFull, off[1], src);
                O2 = __shfl_sync(kFull, off[2], src);
                if (!cand) { O1 = 0; O2 = 1; }
            }
            const bool pat_ok = (O1 == 0 && O2 == 1) || (O1 == 1 && O2 == 1) || (O1 == 1 && O2 == 2);
            bool active = self_ok && pat_ok;
#pragma unroll
            for (int s = 1; s < kG; ++s)
                if (has_ticks && cold[s].tick < p.n_ticks && off[s] != (s == 1 ? O1 : O2)) active = false;
            if (!active) {
                // neutral row code: loads stay inside the row, nothing is captured; the ticks
                // (if any) are evaluated exactly by the whole warp afterwards
                base = 0;
#pragma unroll
                for (int s = 0; s < kG; ++s) {
                    h.f[s] = 0.0; h.gchk[s] = INFINITY;
                    if (has_ticks && cold[s].tick < p.n_ticks) cold[s].flags |= kFallback;
                }
            } else {
                // slots of a partial last group reuse tick 0's factors; their results are never read
#pragma unroll
                for (int s = 1; s < kG; ++s)
                    if (!(cold[s].tick < p.n_ticks)) { h.f[s] = h.f[0]; h.gchk[s] = INFINITY; }
            }
            if (!pat_ok) { O1 = 0; O2 = 1; }

            // ---- stream the rows
            if (O1 == 0) consume_pv<T, NQ, 0, 1>(a, ring, full_bar, empty_bar, it, base, h, cold, active, lane);
            else if (O2 == 1) consume_pv<T, NQ, 1, 1>(a, ring, full_bar, empty_bar, it, base, h, cold, active, lane);
            else consume_pv<T, NQ, 1, 2>(a, ring, full_bar, empty_bar, it, base, h, cold, active, lane);

            if (p.reserved & 2) return;   // (reserved bit 1: debugging aid, skip the per-tick finish)
            // ---- finish each tick: histogram_quantile, ratios, validity
#pragma unroll 1
            for (int s = 0; s < kG; ++s) {
                Cold<NQ>& t = cold[s];
                v[s] = 0;
                if (!(has_ticks && t.tick < p.n_ticks) || (t.flags & kFallback)) continue;
                uint8_t valid = 0;
                if (B > 0) {
                    valid |= CG_V_QUANTILE;
                    // observations after the monotonic fix; the search assumed the +Inf bucket itself
                    const bool spec_ok = !(t.obs_inc < h.prev[s]);
                    const double obs = t.obs_inc / range_s;
                    const bool top_inf = isinf(s_le[B - 1]) && s_le[B - 1] > 0;
                    for (int q = 0; q < p.n_q; ++q) {
                        const double qq = p.q[q];
                        double r;
                        if (isnan(qq)) r = NAN;
                        else if (qq < 0) r = -INFINITY;
                        else if (qq > 1) r = INFINITY;
                        else if (!top_inf || B < 2) r = NAN;
                        else if (!spec_ok) { t.flags |= kFallback; r = NAN; }
                        else if (obs == 0) r = NAN;
                        else if (t.bstar[q] < 0) r = s_le[B - 2];
                        else {
                            const int bb = t.bstar[q];
                            const double rank = qq * obs;
                            const double cbr = t.cb[q] / range_s;
                            const double cpr = bb > 0 ? t.cprev[q] / range_s : 0.0;
                            // the search ran in increase space; confirm it in rate space
                            if (!(cbr >= rank) || (bb > 0 && !(cpr < rank))) { t.flags |= kFallback; r = NAN; }
                            else if (bb == 0 && s_le[0] <= 0) r = s_le[0];
                            else {
                                double start = 0.0, end = s_le[bb], cc = cbr, rk = rank;
                                if (bb > 0) { start = s_le[bb - 1]; cc -= cpr; rk -= cpr; }
                                r = start + (end - start) * (rk / cc);
                            }
                        }
                        m[s][q] = r;
                    }
                }
                const double total = t.sums[0], err = t.sums[1], lsum = t.sums[2], lcnt = t.sums[3], fb = t.sums[4];
                const int nq = p.n_q;
                m[s][nq + CG_M_ERROR_RESPONSES] = err;
                if (total > 0) { m[s][nq + CG_M_ERROR_RATE] = err / total; valid |= CG_V_ERROR_RATE; }
                else m[s][nq + CG_M_ERROR_RATE] = NAN;
                if (lcnt > 0) { m[s][nq + CG_M_LATENCY_AVG] = lsum / lcnt; valid |= CG_V_LATENCY_AVG; }
                else m[s][nq + CG_M_LATENCY_AVG] = NAN;
                m[s][nq + CG_M_REQUEST_COUNT] = lcnt;
                m[s][nq + CG_M_FEEDBACK_COUNT] = fb;
                v[s] = valid;
            }
            // ---- exact re-evaluation of whatever the fast path could not decide
#pragma unroll 1
            for (int s = 0; s < kG; ++s) {
                const bool need = has_ticks && cold[s].tick < p.n_ticks && (cold[s].flags & kFallback);
                unsigned todo = __ballot_sync(kFull, need);
                while (todo) {
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int tk = __shfl_sync(kFull, cold[s].tick, src);
                    double mm[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t vv;
                    coop_eval_tick<T>(a, pv, phase, tk, lane, s_le, mm, vv);
                    if (lane == src) {
                        for (int c = 0; c < W; ++c) m[s][c] = mm[c];
                        v[s] = vv;
                    }
                }
            }
    }
}

template <typename T, int NQ, int CW>
__global__ void __launch_bounds__(CW * 32 + 32, (CW <= 4) ? 3 : 1)
stream_kernel(const __grid_constant__ StreamArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    unsigned char* ring = smem;
    double*   s_le   = reinterpret_cast<double*>(smem + a.off_le);
    uint32_t* s_bits = reinterpret_cast<uint32_t*>(smem + a.off_bits);
    uint64_t* full_bar  = reinterpret_cast<uint64_t*>(smem + a.off_bars);
    uint64_t* empty_bar = full_bar + a.n_slots;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_cons = a.consumer_warps * 32;
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    const int n_words = (p.n_ticks + 31) >> 5;
    if (tid == 0) {
        for (int s = 0; s < a.n_slots; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, a.consumer_warps); }
        mbar_init_fence();
    }
    for (int b = tid; b < B; b += blockDim.x) s_le[b] = a.le[b];
    for (int w = tid; w < 3 * n_words; w += blockDim.x) s_bits[w] = 0u;
    __syncthreads();

    const size_t row_bytes = (size_t)p.row_pitch * sizeof(T);

    if (warp == a.consumer_warps) {
        // ===== producer: one elected lane drives the TMA bulk-copy engine =====
        if (lane == 0) {
            uint32_t it = 0;
            auto push = [&](const char* src, uint32_t bytes) {
                const uint32_t slot = it % (uint32_t)a.n_slots;
                mbar_wait_backoff(empty_bar + slot, ((it / (uint32_t)a.n_slots) & 1u) ^ 1u);
                mbar_arrive_expect_tx(full_bar + slot, bytes);
                bulk_g2s(ring + (size_t)slot * a.slot_bytes, src, bytes, full_bar + slot);
                ++it;
            };
            for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
                if (a.has_previous && !a.has_previous[pair]) continue;
                for (int ver = 0; ver < 2; ++ver) {
                    const int64_t pv = pair * 2 + ver;
                    const char* gs = reinterpret_cast<const char*>(a.scalars) + (size_t)pv * K * row_bytes;
                    const char* gb = reinterpret_cast<const char*>(a.buckets) + (size_t)pv * B * row_bytes;
                    for (int k0 = 0; k0 < K; k0 += kRowsPerSlot)
                        push(gs + (size_t)k0 * row_bytes, (uint32_t)(min(kRowsPerSlot, K - k0) * row_bytes));
                    if (B > 0) {
                        push(gb + (size_t)(B - 1) * row_bytes, (uint32_t)row_bytes);
                        for (int b0 = 0; b0 < B - 1; b0 += kRowsPerSlot)
                            push(gb + (size_t)b0 * row_bytes, (uint32_t)(min(kRowsPerSlot, B - 1 - b0) * row_bytes));
                    }
                }
            }
        }
        return;
    }

    // ===== consumers =====
    const int grp = tid;                       // thread <-> period of the tick grid
    const bool has_ticks = grp < a.n_groups;
    const double range_s = (double)p.window_ms / 1000.0;
    uint32_t it = 0;
    int bit_buf = 0;                           // which of the 3 gate-bit buffers this pair uses

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
        const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
        if (!hp) {
            if (tid == 0) a.decisions[pair] = replay_schedule([](int) { return false; }, false, st_tick, p);
            if (a.gate_bits) for (int w = tid; w < n_words; w += n_cons) a.gate_bits[pair * n_words + w] = 0u;
            continue;
        }
        uint32_t* bits = s_bits + bit_buf * n_words;
        {   // Clear the buffer the NEXT pair will use.  It was last read right after the barrier
            // two pairs ago, and every thread has passed the barrier one pair ago since, so no
            // reader is left; the next pair's atomicOr comes after this pair's barrier.
            uint32_t* nxt = s_bits + ((bit_buf + 1) % 3) * n_words;
            for (int w = tid; w < n_words; w += n_cons) nxt[w] = 0u;
        }
        double nm[kG][3];          // the new version's gate inputs, kept for the old version's pass
        uint8_t nv[kG];
#pragma unroll
        for (int s = 0; s < kG; ++s) { nm[s][0] = nm[s][1] = nm[s][2] = NAN; nv[s] = 0; }

        for (int ver = 0; ver < 2; ++ver) {
            const int64_t pv = pair * 2 + ver;
            double m[kG][CG_MAX_Q + CG_M_EXTRA];
            uint8_t v[kG];
#pragma unroll
            for (int s = 0; s < kG; ++s) v[s] = 0;
            run_version<T, NQ>(a, ring, full_bar, empty_bar, s_le, it, pv, grp, has_ticks, lane, m, v);
            int tick_of[kG];
#pragma unroll
            for (int s = 0; s < kG; ++s) tick_of[s] = grp * kG + s;
            // ---- outputs of this version; gate once both are known
#pragma unroll
            for (int s = 0; s < kG; ++s) {
                const int tk = tick_of[s];
                if (!(has_ticks && tk < p.n_ticks)) continue;
                if (a.metrics) {
                    double* o = a.metrics + ((size_t)pv * p.n_ticks + tk) * W;
                    for (int c = 0; c < W; ++c) o[c] = m[s][c];
                }
                if (a.valid) a.valid[(size_t)pv * p.n_ticks + tk] = v[s];
                const double gm[3] = {p.n_q > 0 ? m[s][p.gate_q] : NAN, m[s][p.n_q + CG_M_ERROR_RATE],
                                      m[s][p.n_q + CG_M_LATENCY_AVG]};
                if (ver == 0) {
                    nm[s][0] = gm[0]; nm[s][1] = gm[1]; nm[s][2] = gm[2]; nv[s] = v[s];
                } else if (gate_eval(nm[s], nv[s], gm, v[s], p) & CG_G_PROMOTE) {
                    atomicOr(bits + (tk >> 5), 1u << (tk & 31));
                }
            }
        }
        // ---- all consumer threads have OR-ed their gate bits: replay the step loop
        asm volatile("bar.sync 1, %0;" ::"r"(n_cons) : "memory");
        if (a.gate_bits) for (int w = tid; w < n_words; w += n_cons) a.gate_bits[pair * n_words + w] = bits[w];
        if (tid == 0)
            a.decisions[pair] = replay_schedule([&](int t) { return (bits[t >> 5] >> (t & 31)) & 1u; }, true, st_tick, p);
        bit_buf = (bit_buf + 1) % 3;
    }
}

// Host-side launcher for one (element type, quantile count) instance; CW is picked at run time.
template <typename T, int NQ>
cudaError_t launch_stream_tn(const StreamArgs& a, int grid, int block, int cw_template, cudaStream_t s) {
    cudaError_t e;
    if (cw_template <= 4) {
        auto kern = stream_kernel<T, NQ, 4>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a.smem_bytes);
        if (e != cudaSuccess) return e;
        kern<<<grid, block, a.smem_bytes, s>>>(a);
    } else {
        auto kern = stream_kernel<T, NQ, 15>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a.smem_bytes);
        if (e != cudaSuccess) return e;
        kern<<<grid, block, a.smem_bytes, s>>>(a);
    }
    return cudaGetLastError();
}

}  // namespace cg
