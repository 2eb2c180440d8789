// cg_stream.cuh -- the STREAMING canary kernel: the fast path for the reference's own
// time constants (60 s window = 4 scrape periods, 10 s evaluation grid = 2/3 period;
// mlflow_operator.py:363, :294 with Prometheus' usual 15 s scrapes).
//
// Layout of the work
//   * One CTA = 4 warps, 4 CTAs per SM, persistent over model pairs.  Every warp is a consumer;
//     there is no producer warp.
//   * Series ROWS stream through a shared-memory ring filled by the TMA bulk-copy engine
//     (cp.async.bulk + mbarrier complete_tx).  The warp that leaves a ring slot last issues the
//     fill that comes n_slots later in the stream, so nobody polls for free slots.  Order per
//     (pair, version): the K scalar rows, the +Inf bucket row, then bucket rows 0..B-2 in
//     groups of 4.  Every sample is read from HBM exactly once.
//   * A thread owns one PERIOD of the tick grid: 3 consecutive ticks whose 4-sample windows live
//     in 6 consecutive samples (three 128-bit shared loads per row).  Which of the 6 each tick
//     uses depends only on the series phase: a warp-uniform pattern.  Because 3 tick steps are
//     exactly 2 scrape periods, the extrapolation terms depend on the tick slot alone and are
//     computed once per (pair, version) by three lanes.
//   * Per row: window deltas with reset corrections, a one-multiply proof that the zero-point
//     clamp cannot fire, increase = result * f; for bucket rows the monotonic fix as a running
//     max and the search for the rank in increase space.  Each rare event (a reset in the tile,
//     a possible clamp, a thread reaching its rank or needing the fix) is one warp vote, so the
//     common path is straight-line.  After the last row each thread finishes
//     histogram_quantile, the ratios (:387-404) and, once both versions are done, the gate
//     (:419-460); gate bits are OR-ed into shared memory and the warp that finishes a pair last
//     replays the step loop (:296-361).
//   * Registers hold only what every row touches; everything else is a shared-memory record per
//     thread (L1 is too small to back local arrays once the rings take their share), and code
//     that runs once per (pair, version) is kept compact and out of line for the instruction cache.
//   * Anything the fast path cannot decide exactly (series edges, phases that put a sample on a
//     window edge, a mis-speculated observation count, a search landing within an ulp of the
//     rank) is re-evaluated for that tick by the whole warp from global memory with the plain
//     sequential semantics (coop_eval_tick).  Results are bit-identical either way.
#pragma once
#include "cg_math.cuh"
#include "cg_ptx.cuh"

namespace cg {

constexpr int kRowsPerSlot = 4;
constexpr int kG = 3;            // ticks per thread (one period of the tick grid)
constexpr unsigned kFull = 0xffffffffu;

// developer aid (build with CG_TIMING=1): per-phase warp-cycle counters, read by tools/profile_kernel.py
#ifdef CG_TIMING
__device__ unsigned long long cg_timing[16];   // wait, rows, setup, finish, barrier, total, warps, long waits
#define CG_T(var) const long long var = clock64()
#define CG_TADD(i, d) do { if (lane == 0) atomicAdd(&cg_timing[i], (unsigned long long)(d)); } while (0)
#else
#define CG_T(var)
#define CG_TADD(i, d)
#endif

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
    uint32_t off_le, off_bits, off_bars, off_rec, smem_bytes;
    // host-computed constants of a regular 4-sample window (all exact IEEE operations)
    double   range_s, rc_range;        // window_ms/1000 and its reciprocal
    double   sampled, rc_sampled;      // 3*period_ms/1000 and its reciprocal
    double   thr, half_avg;            // (sampled/3)*1.1 and (sampled/3)/2
    int32_t  rec_doubles;              // doubles per thread in the shared-memory record
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

// ---- per-thread state ------------------------------------------------------------------
// Registers hold only what every row touches.  Everything touched rarely (captured search
// results, observations, the derived ratios, the other version's gate inputs) lives in SHARED
// memory as structure-of-arrays records [field][thread]: with most of the 228 KB of an SM carved
// out for the row ring, L1 is too small to back local-memory arrays, so none are used.
template <int NQ>
struct Hot {
    double f[kG];        // ext/sampled with the unclamped d_start (Appendix A step 5-6)
    double gchk[kG];     // first_v * gchk >= result  ==>  the zero-point clamp cannot fire
    double prev[kG];     // running max of the bucket increases seen so far (the monotonic fix)
    double x0[kG][NQ];   // search threshold in increase space (rank * range_s); NaN once found
};

template <int NQ>
struct Rec {                       // field indices of the shared-memory record (in doubles)
    static constexpr int kObs   = 0;                   // [kG]      increase of the +Inf bucket
    static constexpr int kCb    = kObs + kG;           // [kG][NQ]  fixed count at the bucket reaching the rank
    static constexpr int kCprev = kCb + kG * NQ;       // [kG][NQ]  fixed count just before it
    static constexpr int kEr    = kCprev + kG * NQ;    // [kG]      error_rate of this version
    static constexpr int kAvg   = kEr + kG;            // [kG]      latency_avg of this version
    static constexpr int kQv    = kAvg + kG;           // [kG]      gate quantile of this version
    static constexpr int kStash = kQv + kG;            // [kG][3]   the NEW version's gate inputs
    static constexpr int kInts  = kStash + kG * 3;     // then (kG*NQ + 2) ints: bstar, valid, stash valid
    static constexpr int kDoubles = kInts + (kG * NQ + 2 + 1) / 2;
};

static __device__ __noinline__ double extrapolate_call(double result, double first_v, const TickGeom& g) {
    return extrapolate(result, first_v, g);
}

// ---- warp-cooperative exact evaluation of one (pair, version, tick) from global memory
template <typename T>
__device__ __noinline__ void coop_eval_tick(const StreamArgs& a, int64_t pv, int32_t phase, int tick, int lane,
                               const double* __restrict__ s_le, double* m, uint8_t& valid_out) {
    const cg_params& p = a.p;
    const int nq = p.n_q, K = p.n_scalar, B = p.n_buckets;
    const TickGeom g = tick_geom32(p, phase, tick);   // the eligibility test bounds every time stamp
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
        if (scan) {   // B <= 64 (eligibility test of the streaming kernel)
            // one pass: lane l holds buckets l and l+32; both windows are loaded before any
            // arithmetic so the global loads overlap
            double v0 = -INFINITY, v1 = -INFINITY;
            {
                double x0[8], x1[8];
                const bool h0 = lane < B, h1 = lane + 32 < B;   // g.n <= 5: the window is 4 scrape periods
                const T* r0 = bk + (size_t)lane * p.row_pitch + g.i0;
                const T* r1 = bk + (size_t)(lane + 32) * p.row_pitch + g.i0;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    x0[i] = (h0 && i < g.n) ? (double)r0[i] : 0.0;
                    x1[i] = (h1 && i < g.n) ? (double)r1[i] : 0.0;
                }
                double res0 = 0.0, res1 = 0.0;
#pragma unroll
                for (int i = 0; i < 8; ++i)           // last - first (+ value before each reset, in order)
                    if (i == g.n - 1) { res0 = x0[i] - x0[0]; res1 = x1[i] - x1[0]; }
#pragma unroll
                for (int i = 1; i < 8; ++i)
                    if (i < g.n) {
                        if (x0[i] < x0[i - 1]) res0 += x0[i - 1];
                        if (x1[i] < x1[i - 1]) res1 += x1[i - 1];
                    }
                if (h0) v0 = ddiv_call(extrapolate_call(res0, x0[0], g), range_s);
                if (h1) v1 = ddiv_call(extrapolate_call(res1, x1[0], g), range_s);
            }
            for (int d = 1; d < 32; d <<= 1) {        // running max in bucket order (monotonic fix)
                const double o0 = __shfl_up_sync(kFull, v0, d), o1 = __shfl_up_sync(kFull, v1, d);
                if (lane >= d && o0 > v0) v0 = o0;
                if (lane >= d && o1 > v1) v1 = o1;
            }
            const double carry = __shfl_sync(kFull, v0, 31);
            if (B > 32 && carry > v1) v1 = carry;
            const double obs = B <= 32 ? __shfl_sync(kFull, v0, B - 1) : __shfl_sync(kFull, v1, B - 33);
            if (obs == 0) {
                for (int j = 0; j < nq; ++j) if (!found[j]) { m[j] = NAN; found[j] = true; }
            } else {
                for (int j = 0; j < nq; ++j) {
                    if (found[j]) continue;
                    const double rank = p.q[j] * obs;
                    const unsigned hit0 = __ballot_sync(kFull, lane < B - 1 && v0 >= rank);
                    const unsigned hit1 = __ballot_sync(kFull, lane + 32 < B - 1 && v1 >= rank);
                    if (!hit0 && !hit1) { m[j] = s_le[B - 2]; continue; }
                    const int src = hit0 ? __ffs(hit0) - 1 : __ffs(hit1) - 1;
                    const int bb = hit0 ? src : src + 32;
                    const double cbv = __shfl_sync(kFull, hit0 ? v0 : v1, src);
                    const double up = __shfl_sync(kFull, hit0 ? v0 : v1, src > 0 ? src - 1 : 0);
                    const double cpv = src > 0 ? up : carry;     // (bb == 32: the last bucket of the lower half)
                    if (bb == 0 && s_le[0] <= 0) m[j] = s_le[0];
                    else {
                        double start = 0.0, end = s_le[bb], cc = cbv, rk = rank;
                        if (bb > 0) { start = s_le[bb - 1]; cc -= cpv; rk -= cpv; }
                        m[j] = start + (end - start) * ddiv_call(rk, cc);
                    }
                }
            }
        }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fb = sums[4];
    m[nq + CG_M_ERROR_RESPONSES] = err;
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = ddiv_call(err, total); valid |= CG_V_ERROR_RATE; }
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = ddiv_call(lat_sum, lat_cnt); valid |= CG_V_LATENCY_AVG; }
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
    m[nq + CG_M_FEEDBACK_COUNT] = fb;
    valid_out = valid;
}

// Rare path of a fast row: the zero-point clamp may fire -> exact Appendix A step 4-6 for a
// regular 4-sample window; its two tick-dependent terms are recomputed from the time stamps.
static __device__ __noinline__ double slow_increase(const StreamArgs& a, int32_t phase, int tick, double result,
                                                    double first_v) {
    const cg_params& p = a.p;
    const int t = p.tick0_ms + tick * p.tick_step_ms, lo = t - p.window_ms;
    int i0 = p.left_open ? floor_div32(lo - phase, p.period_ms) + 1 : -floor_div32(phase - lo, p.period_ms);
    int i1 = floor_div32(t - phase, p.period_ms);
    if (i0 < 0) i0 = 0;
    if (i1 > p.n_samples - 1) i1 = p.n_samples - 1;
    TickGeom g;
    g.i0 = i0; g.n = 4;
    g.sampled = a.sampled;
    g.d_start = div_const((double)(phase + i0 * p.period_ms - lo), 1000.0, 1e-3);
    const double d_end = div_const((double)(t - (phase + i1 * p.period_ms)), 1000.0, 1e-3);
    g.avg = a.half_avg + a.half_avg;
    g.thr = a.thr;
    g.end_term = (d_end < a.thr) ? d_end : a.half_avg;
    return extrapolate(result, first_v, g);
}

// Window deltas and fast increases of one row for the thread's kG ticks.  o1 / o2 (warp-uniform):
// tick 1 uses the window at tile offset (o1 ? 1 : 0), tick 2 the one at (o2 ? 2 : 1); tick 0 is at
// offset 0.  The pattern depends only on the series phase.  One copy of this code serves every
// pattern and every kind of row: the kernel's hot loop has to stay small for the instruction cache.
// All 32 lanes call this.
template <typename T, int NQ>
__device__ __forceinline__ void row_increases(const T* __restrict__ row, const Hot<NQ>& h, const StreamArgs& a,
                                              const int32_t phase, const int tick0, const bool lane_live,
                                              const bool o1, const bool o2, double (&inc)[kG]) {
    double x[6];
    Tile6<T>::load(row, x);
    double w0 = x[3] - x[0], w1 = x[4] - x[1], w2 = x[5] - x[2];
    const bool r1 = x[1] < x[0], r2 = x[2] < x[1], r3 = x[3] < x[2], r4 = x[4] < x[3], r5 = o2 && x[5] < x[4];
    // idle lanes (no ticks, or ticks left to the exact path) must not drag the warp into the rare paths
    if (__any_sync(kFull, (r1 | r2 | r3 | r4 | r5) && lane_live)) {   // counter resets are rare (Appendix A step 2)
        // value before each reset (0 where there is none), added to each window in time order;
        // adding +0.0 leaves a sum unchanged, so this is the oracle's conditional add
        const double d1 = r1 ? x[0] : 0.0, d2 = r2 ? x[1] : 0.0, d3 = r3 ? x[2] : 0.0, d4 = r4 ? x[3] : 0.0,
                     d5 = r5 ? x[4] : 0.0;
        w0 = ((w0 + d1) + d2) + d3;
        w1 = ((w1 + d2) + d3) + d4;
        w2 = ((w2 + d3) + d4) + d5;
    }
    const double res[kG] = {w0, o1 ? w1 : w0, o2 ? w2 : w1};
    const double fst[kG] = {x[0], o1 ? x[1] : x[0], o2 ? x[2] : x[1]};
    // One multiply proves that the zero-point clamp cannot fire (first_v * gchk >= result, with
    // gchk a slight under-estimate of sampled/d_start); lanes that fail it take the exact path.
    bool slow[kG];
#pragma unroll
    for (int s = 0; s < kG; ++s) {
        // gchk is finite for evaluated ticks, so a NaN product can only come from a NaN sample, for
        // which the fast formula is already what the exact path gives.  Tick slots this thread does
        // not evaluate (idle lanes, a partial last group) carry gchk = NaN: never "slow".
        slow[s] = fst[s] * h.gchk[s] < res[s];
        inc[s] = res[s] * h.f[s];
    }
    if (__any_sync(kFull, slow[0] | slow[1] | slow[2])) {   // rare: a counter that restarted one scrape ago
#pragma unroll
        for (int s = 0; s < kG; ++s)
            if (slow[s]) inc[s] = slow_increase(a, phase, tick0 + s, res[s], fst[s]);
    }
}

// Position in a CTA's stream of ring fills: (pair, version, slot-use within the version).
struct FillCursor {
    int64_t pair;
    int ver, u;
};

template <typename T, int NQ, int CW>
__global__ void __launch_bounds__(CW * 32, (CW <= 4) ? 4 : 1)
stream_kernel(const __grid_constant__ StreamArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    typedef Rec<NQ> R;
    const cg_params& p = a.p;
    unsigned char* ring = smem;
    double*   s_le   = reinterpret_cast<double*>(smem + a.off_le);
    uint32_t* s_bits = reinterpret_cast<uint32_t*>(smem + a.off_bits);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + a.off_bars);
    uint32_t* s_arr = reinterpret_cast<uint32_t*>(full_bar + a.n_slots);   // per-slot arrival counters

    const int tid = threadIdx.x, lane = tid & 31;
    const int n_cons = a.consumer_warps * 32;
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    const int n_words = (p.n_ticks + 31) >> 5;
    uint32_t* s_cnt = s_bits + 3 * n_words;     // arrival counters, one per gate-bit buffer
    if (tid == 0) {
        for (int s = 0; s < a.n_slots; ++s) { mbar_init(full_bar + s, 1); s_arr[s] = 0u; }
        mbar_init_fence();
    }
    for (int b = tid; b < B; b += blockDim.x) s_le[b] = a.le[b];
    for (int w = tid; w < 3 * n_words + 3; w += blockDim.x) s_bits[w] = 0u;
    __syncthreads();

    const size_t row_bytes = (size_t)p.row_pitch * sizeof(T);
    const int n_sc = (K + kRowsPerSlot - 1) / kRowsPerSlot;
    const int n_use = n_sc + (B > 0 ? 1 + (B - 1 + kRowsPerSlot - 1) / kRowsPerSlot : 0);

    // ---- the ring is refilled by the consumers themselves: the warp that releases a slot last
    // issues the TMA bulk copy of the fill that comes n_slots later in the stream.  No producer
    // warp, no polling on "empty" barriers.  Order per (pair, version): the K scalar rows, the
    // +Inf bucket row, then bucket rows 0..B-2 in groups of kRowsPerSlot; every sample leaves
    // HBM exactly once.
    auto cursor_valid = [&](const FillCursor& c) { return c.pair < a.n_pairs; };
    auto skip_pairs = [&](FillCursor& c) {
        while (c.pair < a.n_pairs && a.has_previous && !a.has_previous[c.pair]) c.pair += gridDim.x;
    };
    auto cursor_next = [&](FillCursor& c) {
        if (++c.u == n_use) {
            c.u = 0;
            if (++c.ver == 2) { c.ver = 0; c.pair += gridDim.x; skip_pairs(c); }
        }
    };
    auto issue_fill = [&](const FillCursor& c, uint32_t slot) {      // one lane
        const int64_t pv = c.pair * 2 + c.ver;
        const char* src;
        uint32_t bytes;
        if (c.u < n_sc) {
            const int k0 = c.u * kRowsPerSlot;
            src = reinterpret_cast<const char*>(a.scalars) + ((size_t)pv * K + k0) * row_bytes;
            bytes = (uint32_t)(min(kRowsPerSlot, K - k0) * row_bytes);
        } else if (c.u == n_sc) {
            src = reinterpret_cast<const char*>(a.buckets) + ((size_t)pv * B + (B - 1)) * row_bytes;
            bytes = (uint32_t)row_bytes;
        } else {
            const int b0 = (c.u - n_sc - 1) * kRowsPerSlot;
            src = reinterpret_cast<const char*>(a.buckets) + ((size_t)pv * B + b0) * row_bytes;
            bytes = (uint32_t)(min(kRowsPerSlot, B - 1 - b0) * row_bytes);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // reads of the slot before the async write
        mbar_arrive_expect_tx(full_bar + slot, bytes);
        bulk_g2s(ring + (size_t)slot * a.slot_bytes, src, bytes, full_bar + slot);
    };
    FillCursor fill;
    fill.pair = blockIdx.x; fill.ver = 0; fill.u = 0;
    skip_pairs(fill);
    for (int s = 0; s < a.n_slots; ++s) {          // prologue: fill the whole ring
        if (tid == 0 && cursor_valid(fill)) issue_fill(fill, (uint32_t)s);
        cursor_next(fill);
    }

    // ===== consumers =====
    const int grp = tid;                       // thread <-> period of the tick grid
    const int tick0 = grp * kG;
    const bool has_ticks = grp < a.n_groups;
    int n_mine = 0;                            // how many of the kG ticks exist
    if (has_ticks) n_mine = min(kG, p.n_ticks - tick0);
    double* rec = reinterpret_cast<double*>(smem + a.off_rec) + tid;          // field f at rec[f * n_cons]
    int* rec_i = reinterpret_cast<int*>(reinterpret_cast<double*>(smem + a.off_rec) + (size_t)R::kInts * n_cons) + tid;
#define CG_REC(f) rec[(f) * n_cons]
#define CG_RECI(f) rec_i[(f) * n_cons]        // ints: [0, kG*NQ) bstar, kG*NQ valid bytes, kG*NQ+1 stash valid bytes
    const int pitch = p.row_pitch;
    uint32_t slot = 0, ring_phase = 0;         // position in the ring and the parity of its mbarrier
    int bit_buf = 0;                           // which of the 3 gate-bit buffers this pair uses

#ifdef CG_TIMING
    long long t_acq = 0;
#endif
    auto acquire = [&]() -> const T* {
        CG_T(tw0);
        mbar_wait(full_bar + slot, ring_phase);
#ifdef CG_TIMING
        t_acq = clock64();
        CG_TADD(0, t_acq - tw0);
#endif
        return reinterpret_cast<const T*>(ring + (size_t)slot * a.slot_bytes);
    };
    auto release = [&]() {
        __syncwarp();
#ifdef CG_TIMING
        CG_TADD(1, clock64() - t_acq);
#endif
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(s_arr + slot, 1u) == (unsigned)a.consumer_warps - 1u) {   // last warp out refills
                s_arr[slot] = 0u;
                if (cursor_valid(fill)) issue_fill(fill, slot);
            }
        }
        cursor_next(fill);
        if (++slot == (uint32_t)a.n_slots) { slot = 0; ring_phase ^= 1u; }   // no division in the row stream
    };
    CG_T(tk0);

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
        if (!hp) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            if (tid == 0 && a.decisions) a.decisions[pair] = replay_schedule([](int) { return false; }, false, st_tick, p);
            if (a.gate_bits) for (int w = tid; w < n_words; w += n_cons) a.gate_bits[pair * n_words + w] = 0u;
            continue;
        }
        uint32_t* bits = s_bits + bit_buf * n_words;
        {   // Clear the buffer the NEXT pair will use.  It was last read right after the barrier
            // two pairs ago, and every thread has passed the barrier one pair ago since, so no
            // reader is left; the next pair's atomicOr comes after this pair's barrier.
            uint32_t* nxt = s_bits + ((bit_buf + 1) % 3) * n_words;
            for (int w = tid; w < n_words; w += n_cons) nxt[w] = 0u;
            if (tid == 0) s_cnt[(bit_buf + 1) % 3] = 0u;
        }
#pragma unroll 1
        for (int ver = 0; ver < 2; ++ver) {
            const int64_t pv = pair * 2 + ver;
            const int32_t phase = a.phase[pv];
            CG_T(ts0);
            // ---- tick geometry.  Three tick steps are exactly two scrape periods, so the window of
            // tick 3g+s is the window of tick s shifted by 2g samples and its extrapolation terms
            // (d_start, d_end -> f, gchk) depend on s alone: lanes 0..2 compute them once per
            // (pair, version) and the warp shares them.  32-bit integers throughout (the eligibility
            // test bounds every time stamp); every division is by a constant (div_const).
            Hot<NQ> h;
            int i0r = 0, i1r = 0;            // unclamped window bounds of tick s = lane (lanes 0..2)
            double fs = 0.0, gs = 1e300;
            if (lane < kG) {
                const int t = p.tick0_ms + lane * p.tick_step_ms, lo = t - p.window_ms;
                i0r = p.left_open ? floor_div32(lo - phase, p.period_ms) + 1 : -floor_div32(phase - lo, p.period_ms);
                i1r = floor_div32(t - phase, p.period_ms);
                if (i1r - i0r + 1 == 4) {
                    const double d_start = div_const((double)(phase + i0r * p.period_ms - lo), 1000.0, 1e-3);
                    const double d_end = div_const((double)(t - (phase + i1r * p.period_ms)), 1000.0, 1e-3);
                    double ext = a.sampled;
                    ext += (d_start < a.thr) ? d_start : a.half_avg;
                    ext += (d_end < a.thr) ? d_end : a.half_avg;
                    fs = div_const(ext, a.sampled, a.rc_sampled);
                    // an under-estimate of sampled/d_start only sends more rows to the exact path
                    gs = fmin((a.sampled * (double)__frcp_rd(__double2float_ru(d_start))) * (1.0 - 1e-6), 1e300);
                }
            }
            int i0s[kG], i1s[kG];
#pragma unroll
            for (int s = 0; s < kG; ++s) {
                i0s[s] = __shfl_sync(kFull, i0r, s);
                i1s[s] = __shfl_sync(kFull, i1r, s);
                h.f[s] = __shfl_sync(kFull, fs, s);
                h.gchk[s] = __shfl_sync(kFull, gs, s);
                h.prev[s] = -INFINITY;
#pragma unroll
                for (int q = 0; q < NQ; ++q) h.x0[s][q] = NAN;
            }
            const int off1 = i0s[1] - i0s[0], off2 = i0s[2] - i0s[0];
            int pat = (off1 == 0 && off2 == 1) ? 0 : (off1 == 1 && off2 == 1) ? 1 : (off1 == 1 && off2 == 2) ? 2 : -1;
            int base = i0s[0] + 2 * grp;
            // this thread's ticks are regular if each has an unclipped 4-sample window
            bool regular = n_mine > 0 && pat >= 0 && (base & 1) == 0;
#pragma unroll
            for (int s = 0; s < kG; ++s)
                if (s < n_mine && !(i1s[s] - i0s[s] + 1 == 4 && i0s[s] + 2 * grp >= 0 && i1s[s] + 2 * grp <= p.n_samples - 1))
                    regular = false;
            if (pat < 0) pat = 0;
            const bool active = regular;
            unsigned fb_mask = 0;                // ticks that need the exact warp-cooperative path
            if (!active) {
                // neutral row code: loads stay inside the row, nothing is captured
                base = 0;
                fb_mask = (1u << n_mine) - 1u;
            }
#pragma unroll
            for (int s = 0; s < kG; ++s)      // tick slots the fast path does not evaluate
                if (!active || s >= n_mine) { h.f[s] = 0.0; h.gchk[s] = NAN; }
#pragma unroll
            for (int q = 0; q < kG * NQ; ++q) CG_RECI(q) = -1;

            CG_TADD(2, clock64() - ts0);
            // ---- the row stream: K scalar rows (sum(increase(...)) per role, :375-410), the +Inf
            // bucket (observations -> rank -> search threshold), then bucket rows 0..B-2 (monotonic
            // fix and the search for the rank, in bucket order).  One loop, one copy of the row code.
            const bool o1 = pat != 0, o2 = pat == 2;
            {   // scalar rows; the role sums live in registers only inside this block
                double sums[kG][5];
                unsigned have = 0;
#pragma unroll
                for (int s = 0; s < kG; ++s)
#pragma unroll
                    for (int q = 0; q < 5; ++q) sums[s][q] = 0.0;
#pragma unroll 1
                for (int u = 0; u < n_sc; ++u) {
                    const T* rows = acquire();
                    const int first = u * kRowsPerSlot, nr = min(kRowsPerSlot, K - first);
                    if (!(p.reserved & 1))
#pragma unroll 1
                    for (int k = 0; k < nr; ++k) {
                        double inc[kG];
                        row_increases<T, NQ>(rows + k * pitch + base, h, a, phase, tick0, active, o1, o2, inc);
                        const unsigned role = p.scalar_role[first + k];
#pragma unroll
                        for (int q = 0; q < 5; ++q)
                            if (role & (1u << q)) {
#pragma unroll
                                for (int s = 0; s < kG; ++s) sums[s][q] = (have & (1u << q)) ? sums[s][q] + inc[s] : inc[s];
                                have |= 1u << q;
                            }
                    }
                    release();
                }
                // derived metrics (:387-404), validity, optional record output
                unsigned vbytes = 0;
#pragma unroll
                for (int s = 0; s < kG; ++s) {
                    const double total = sums[s][0], err = sums[s][1], lsum = sums[s][2], lcnt = sums[s][3];
                    unsigned valid = (B > 0) ? CG_V_QUANTILE : 0u;
                    double er = NAN, avg = NAN;
                    if (total > 0) { er = ddiv_call(err, total); valid |= CG_V_ERROR_RATE; }
                    if (lcnt > 0) { avg = ddiv_call(lsum, lcnt); valid |= CG_V_LATENCY_AVG; }
                    CG_REC(R::kEr + s) = er;
                    CG_REC(R::kAvg + s) = avg;
                    vbytes |= valid << (8 * s);
                    if (a.metrics && s < n_mine && active) {
                        double* o = a.metrics + ((size_t)pv * p.n_ticks + tick0 + s) * W + p.n_q;
                        o[CG_M_ERROR_RESPONSES] = err; o[CG_M_ERROR_RATE] = er; o[CG_M_LATENCY_AVG] = avg;
                        o[CG_M_REQUEST_COUNT] = lcnt; o[CG_M_FEEDBACK_COUNT] = sums[s][4];
                    }
                }
                CG_RECI(kG * NQ) = (int)vbytes;
            }
            // ---- the +Inf bucket (observations -> rank -> search threshold), then bucket rows
            // 0..B-2 (monotonic fix and the search for the rank, in bucket order): one loop
#pragma unroll 1
            for (int u = n_sc; u < n_use; ++u) {
                const T* rows = acquire();
                const bool is_inf = u == n_sc;
                const int first = is_inf ? B - 1 : (u - n_sc - 1) * kRowsPerSlot;
                const int nr = is_inf ? 1 : min(kRowsPerSlot, B - 1 - first);
                if (!(p.reserved & 1))
#pragma unroll 1
                for (int k = 0; k < nr; ++k) {
                    double inc[kG];
                    row_increases<T, NQ>(rows + k * pitch + base, h, a, phase, tick0, active, o1, o2, inc);
                    if (!is_inf) {
                        // common case: increases are monotone in the bucket index and below the
                        // search threshold -> prev = inc.  One vote decides whether any lane needs
                        // the monotonic fix (Appendix A step 3) or has reached its rank.
                        bool rare = false;
#pragma unroll
                        for (int s = 0; s < kG; ++s) {
                            rare |= inc[s] < h.prev[s];
#pragma unroll
                            for (int q = 0; q < NQ; ++q) rare |= inc[s] >= h.x0[s][q];
                        }
                        if (__any_sync(kFull, rare)) {
#pragma unroll
                            for (int s = 0; s < kG; ++s) {
                                const double c = inc[s] < h.prev[s] ? h.prev[s] : inc[s];
#pragma unroll
                                for (int q = 0; q < NQ; ++q)
                                    if (c >= h.x0[s][q]) {      // first bucket reaching the rank
                                        CG_RECI(s * NQ + q) = first + k;
                                        CG_REC(R::kCb + s * NQ + q) = c;
                                        CG_REC(R::kCprev + s * NQ + q) = h.prev[s];
                                        h.x0[s][q] = NAN;       // never true again
                                    }
                                inc[s] = c;
                            }
                        }
#pragma unroll
                        for (int s = 0; s < kG; ++s) h.prev[s] = inc[s];
                    } else {
#pragma unroll
                        for (int s = 0; s < kG; ++s) {
                            CG_REC(R::kObs + s) = inc[s];
                            const double obs = div_const(inc[s], a.range_s, a.rc_range);
#pragma unroll
                            for (int q = 0; q < NQ; ++q)   // a candidate threshold; the finish confirms it in rate space
                                h.x0[s][q] = active ? fmin((p.q[q] * obs) * a.range_s, inc[s]) : NAN;
                        }
                    }
                }
                release();
            }
            if (p.reserved & 2) continue;   // (reserved bit 1: debugging aid, skip the per-tick finish)
            CG_T(tf0);
            // ---- finish each tick: histogram_quantile (Appendix A) from the captured search state
            if (B > 0 && active && !(p.reserved & 8)) {
                // Straight-line per (tick, quantile): every candidate value is computed, then one
                // chain of selects picks the case, so the three ticks overlap in the pipelines.
                const bool top_inf = isinf(s_le[B - 1]) && s_le[B - 1] > 0;
                const bool shape_ok = top_inf && B >= 2;
                const double le_lo = s_le[0], le_top = B >= 2 ? s_le[B - 2] : NAN;
                // observations after the monotonic fix; the search assumed the +Inf bucket itself
                unsigned spec_bad = 0;
#pragma unroll
                for (int s = 0; s < kG; ++s) {
                    if (CG_REC(R::kObs + s) < h.prev[s]) spec_bad |= 1u << s;
#pragma unroll
                    for (int q = 0; q < NQ; ++q)     // search never fired: remember the largest fixed count
                        if (CG_RECI(s * NQ + q) < 0) CG_REC(R::kCprev + s * NQ + q) = h.prev[s];
                }
#pragma unroll 1
                for (int s = 0; s < n_mine; ++s) {
                    const double obs_inc = CG_REC(R::kObs + s);
                    const bool spec_ok = !((spec_bad >> s) & 1u);
                    const double obs = div_const(obs_inc, a.range_s, a.rc_range);
#pragma unroll
                    for (int q = 0; q < NQ; ++q) {
                        if (q >= p.n_q) continue;
                        const double qq = p.q[q];
                        const int bb = CG_RECI(s * NQ + q);
                        const int bi = bb > 0 ? bb : 0;
                        const double rank = qq * obs;
                        const double cbr = div_const(CG_REC(R::kCb + s * NQ + q), a.range_s, a.rc_range);
                        const double cpr = div_const(CG_REC(R::kCprev + s * NQ + q), a.range_s, a.rc_range);
                        const double start = bb > 0 ? s_le[bi - 1] : 0.0, end = s_le[bi];
                        const double cc = bb > 0 ? cbr - cpr : cbr, rk = bb > 0 ? rank - cpr : rank;
                        const double interp = start + (end - start) * ddiv_call(rk, cc);
                        // the search ran in increase space; confirm it in rate space
                        // (no bucket reached the threshold: then none below the +Inf one may reach the rank)
                        const bool confirmed = bb < 0 ? (cpr < rank) : ((cbr >= rank) && (bb == 0 || cpr < rank));
                        const bool q_ok = !isnan(qq) && qq >= 0 && qq <= 1;
                        const bool searched = q_ok && shape_ok && obs != 0;
                        double r = interp;
                        if (bb == 0 && le_lo <= 0) r = le_lo;
                        if (bb < 0) r = le_top;
                        if (!searched) r = NAN;                       // bad shape, q is NaN, or no observations
                        if (qq < 0) r = -INFINITY;
                        if (qq > 1) r = INFINITY;
                        if (searched && (!spec_ok || !confirmed)) fb_mask |= 1u << s;
                        if (q == p.gate_q) CG_REC(R::kQv + s) = r;
                        if (a.metrics) a.metrics[((size_t)pv * p.n_ticks + tick0 + s) * W + q] = r;
                    }
                }
            }
            if (a.valid && active)
                for (int s = 0; s < n_mine; ++s)
                    a.valid[(size_t)pv * p.n_ticks + tick0 + s] = (uint8_t)((CG_RECI(kG * NQ) >> (8 * s)) & 0xff);
            CG_T(tq1);
            CG_TADD(8, tq1 - tf0);
            // ---- exact re-evaluation of whatever the fast path could not decide
#pragma unroll 1
            for (int s = 0; s < kG; ++s) {
                unsigned todo = __ballot_sync(kFull, ((fb_mask >> s) & 1u) && !(p.reserved & 4));
                while (todo) {
                    CG_TADD(7, 1);
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int tk = __shfl_sync(kFull, tick0 + s, src);
                    double mm[CG_MAX_Q + CG_M_EXTRA];
                    uint8_t vv;
                    coop_eval_tick<T>(a, pv, phase, tk, lane, s_le, mm, vv);
                    if (lane == src) {
                        CG_REC(R::kQv + s) = p.n_q > 0 ? mm[p.gate_q] : NAN;
                        CG_REC(R::kEr + s) = mm[p.n_q + CG_M_ERROR_RATE];
                        CG_REC(R::kAvg + s) = mm[p.n_q + CG_M_LATENCY_AVG];
                        CG_RECI(kG * NQ) = (CG_RECI(kG * NQ) & ~(0xff << (8 * s))) | ((int)vv << (8 * s));
                        if (a.metrics) {
                            double* o = a.metrics + ((size_t)pv * p.n_ticks + tk) * W;
                            for (int c = 0; c < W; ++c) o[c] = mm[c];
                        }
                        if (a.valid) a.valid[(size_t)pv * p.n_ticks + tk] = vv;
                    }
                }
            }
            CG_T(tq2);
            CG_TADD(9, tq2 - tq1);
            // ---- the new version's gate inputs wait for the old version's pass; then the gate
            if (p.reserved & 16) continue;
            if (ver == 0) {
#pragma unroll
                for (int s = 0; s < kG; ++s) {
                    CG_REC(R::kStash + s * 3 + 0) = (p.n_q > 0) ? CG_REC(R::kQv + s) : NAN;
                    CG_REC(R::kStash + s * 3 + 1) = CG_REC(R::kEr + s);
                    CG_REC(R::kStash + s * 3 + 2) = CG_REC(R::kAvg + s);
                }
                CG_RECI(kG * NQ + 1) = CG_RECI(kG * NQ);
            } else {
                const int nvb = CG_RECI(kG * NQ + 1), ovb = CG_RECI(kG * NQ);
#pragma unroll 1
                for (int s = 0; s < n_mine; ++s) {
                    const int tk = tick0 + s;
                    const double nm[3] = {CG_REC(R::kStash + s * 3 + 0), CG_REC(R::kStash + s * 3 + 1), CG_REC(R::kStash + s * 3 + 2)};
                    const double om[3] = {(p.n_q > 0) ? CG_REC(R::kQv + s) : NAN, CG_REC(R::kEr + s), CG_REC(R::kAvg + s)};
                    if (gate_eval(nm, (uint8_t)((nvb >> (8 * s)) & 0xff), om, (uint8_t)((ovb >> (8 * s)) & 0xff), p) & CG_G_PROMOTE)
                        atomicOr(bits + (tk >> 5), 1u << (tk & 31));
                }
            }
            CG_TADD(3, clock64() - tf0);
        }
        // ---- the warp that finishes this pair last replays the step loop; nobody waits.
        // (Consumer warps of a CTA are never more than the ring depth apart, far less than a pair,
        // so the triple-buffered gate bits and arrival counters cannot be overtaken.)
        __syncwarp();
        int last = 0;
        if (lane == 0) {
            __threadfence_block();
            last = (atomicAdd(s_cnt + bit_buf, 1u) == (unsigned)a.consumer_warps - 1u) ? 1 : 0;
        }
        last = __shfl_sync(kFull, last, 0);
        if (last) {
            __threadfence_block();
            if (a.gate_bits) for (int w = lane; w < n_words; w += 32) a.gate_bits[pair * n_words + w] = bits[w];
            if (lane == 0 && a.decisions) {
                const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
                a.decisions[pair] = replay_schedule_bits(bits, n_words, true, st_tick, p);
            }
        }
        bit_buf = (bit_buf + 1) % 3;
    }
#ifdef CG_TIMING
    CG_TADD(5, clock64() - tk0);
    CG_TADD(6, 1);
#endif
#undef CG_REC
#undef CG_RECI
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
