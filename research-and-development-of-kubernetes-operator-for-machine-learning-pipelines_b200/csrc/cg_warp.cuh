// cg_warp.cuh -- the STREAMING canary kernel (fast path for the reference's own time constants:
// 60 s window = 4 scrape periods, 10 s evaluation grid = 2/3 period; mlflow_operator.py:363, :294
// with Prometheus' usual 15 s scrapes).
//
// Layout of the work
//   * One CTA = the kWarps (two) warps of one model pair; 8 CTAs per SM; persistent over pairs.
//     The pair's series rows stream through the CTA's shared-memory ring, filled by the TMA
//     bulk-copy engine (cp.async.bulk + mbarrier complete_tx, 2 rows = 3.8 KB per copy).  The CTA
//     is its own producer: when both warps are done with a slot, thread 0 issues the copy that
//     comes n_slots later (a per-pass table in shared memory says which rows).  Row order per
//     pair: scalar rows of the new version, scalar rows of the old version, then per version the
//     +Inf bucket row followed by bucket rows 0..B-2.  Every sample leaves HBM once.
//   * A lane owns kG = 2 periods of the tick grid = 6 consecutive ticks (3 tick steps are exactly
//     2 scrape periods), whose 4-sample windows start at 5 consecutive samples: one 8-sample
//     register tile per row (four 128-bit shared loads), 5 window deltas.  60 of the pair's 64
//     lanes cover the 355 ticks of an hour; longer grids take more passes over the rows.  The
//     tick grid is rotated by 0..2 ticks per (pair, version) so that tiles start on even samples,
//     and edge lanes slide their tile back inside the series.
//   * Bucket rows are processed in WINDOW space.  A tick's increase is delta * f with f > 0
//     depending on the tick alone, and x -> fl(fl(x*f)/range) is monotone, so the monotonic fix
//     of histogram_quantile is a running max of window deltas (one compare-select per window and
//     row) and "this bucket reaches the rank" is a compare of the delta against a per-window
//     threshold (an under-estimate, so it never misses).  Only when a compare fires is anything
//     evaluated per tick, exactly, in rate space, by an event loop over the flagged windows; it
//     captures the interpolated quantile on the spot and, for the old version, decides the
//     quantile part of the gate against the new version's value.
//   * Counter resets are corrected branch-free in the register tile when any lane sees one.  The
//     zero-point clamp of extrapolatedRate makes a row's factor tick- and row-specific: one
//     multiply per tile proves it cannot fire; rows where it does are kept out of the window
//     running max and folded, exactly, into a per-tick rate-space maximum kept in shared memory.
//   * Scalar rows (sum(increase(...)) per role): all K rows of a version are resident in the ring
//     at once and each lane walks its ticks over them with five running sums; the two versions'
//     ratios are compared as soon as the old version's scalar rows are through, so only pass/fail
//     bits wait for the quantiles.
//   * Divisions: by constants with the 3-operation Markstein sequence, by data with a Newton
//     sequence that is proven (tests/div_check.cpp) to give the correctly rounded quotient.
//   * Ticks the fast path cannot decide exactly (series edges, a sample on a window edge, a
//     non-finite sample, a mis-speculated observation count) are re-evaluated at the end by a
//     whole warp from global memory with the plain sequential semantics (coop_eval_tick).
//     Results are bit-identical to the oracle either way.
//
// The per-lane arithmetic in the first half of this file is plain C++ as well (CG_DEVINL), so
// tests/host_emul compiles it with g++ and checks it against the oracle without a GPU.
#pragma once
#include "cg_math.cuh"
#ifdef __CUDACC__
#include "cg_ptx.cuh"
#endif

namespace cg {

// A lane owns kG periods of the tick grid (3 ticks and 2 new samples each).  4/kG warps share one
// pair, so that 128 lane-periods = 384 ticks are covered in one pass whatever kG is: kG = 4 is one
// warp with 12 ticks per lane, kG = 2 two warps with 6 ticks per lane (half the per-warp state and
// ring, twice the resident warps).
#ifndef CG_G
#define CG_G 2
#endif
constexpr int kG = CG_G;
static_assert(kG == 1 || kG == 2 || kG == 4, "kG must be 1, 2 or 4");
constexpr int kWarps = 4 / kG;             // warps per pair (= per CTA)
constexpr int kLanes = 32 * kWarps;        // lanes per pair
constexpr int kLaneTicks = 3 * kG;         // ticks per lane
constexpr int kLaneWin = 2 * kG + 1;       // window start positions per lane tile
constexpr int kTile = 2 * kG + 4;          // samples per lane tile
constexpr int kChunkTicks = 382;           // ticks one pass over the rows covers (384 minus the rotation slack)
#ifndef CG_ROWS_PER_USE
#define CG_ROWS_PER_USE 2
#endif
constexpr int kRowsPerUse = CG_ROWS_PER_USE;   // rows per ring slot
constexpr unsigned kFullMask = 0xffffffffu;

// Developer build with bounds checks of our own (build.py with CG_CHECKED=1; tools/run_checked.py reads the
// counters): every shared-memory tile, ring fill and bit-vector word the kernel touches is tested against its
// array; a failed test bumps a counter instead of touching memory it should not.
#if defined(CG_CHECKED) && defined(__CUDACC__)
static __device__ unsigned int cg_check_failures[8];
#define CG_CHECK(i, cond) do { if (!(cond)) atomicAdd(&cg_check_failures[i], 1u); } while (0)
#else
#define CG_CHECK(i, cond)
#endif

#ifndef __CUDACC__
// host emulation only: what a lane's last bucket row did (warp-level statistics in tests/host_emul)
struct EmulRowStat { int reset, event, windows, captures, clamps, cfl, hfl; };
static EmulRowStat g_emul_row;
#define CG_EMUL_STAT(field, v) (g_emul_row.field += (v))
#else
#define CG_EMUL_STAT(field, v)
#endif

// per-tick state arrays in shared memory, [field][local tick][lane]
constexpr int kFA = 0;    // new version's error_rate until the old one's scalar rows are done; then observations (rate of the +Inf bucket)
constexpr int kFB = 1;    // new version's latency_avg; then the per-tick rate-space maximum of excluded rows
constexpr int kFSQ = 2;   // new version's gate quantile, in the NEW version's lane coordinates
constexpr int kFields = 3;

struct WarpArgs {
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
    int32_t  n_slots;              // ring depth in slots
    int32_t  n_sc_uses;            // slot uses for the K scalar rows of one version
    int32_t  n_bk_uses;            // slot uses for bucket rows 0..B-2 of one version (the +Inf row has its own)
    int32_t  uses_per_pass;        // slot uses of one pass over a pair's rows
    int32_t  n_chunks;             // passes per pair: ceil(n_ticks / kChunkTicks)
    uint32_t row_bytes, slot_bytes;
    uint32_t off_bars, off_state, off_scratch, off_geom, off_bits, off_tab, smem_bytes;   // the ring sits at offset 0
    // host-computed constants of a regular 4-sample window (all exact IEEE operations)
    double   range_s, rc_range;    // window_ms/1000 and its reciprocal
    double   sampled, rc_sampled;  // 3*period_ms/1000 and its reciprocal
    double   thr, half_avg;        // (sampled/3)*1.1 and (sampled/3)/2
    int32_t  qkind[CG_MAX_Q];      // 0: searched; 1: q is NaN; 2: q < 0; 3: q > 1  (Appendix A step 1)
};

// Host side: the constants above from the parameter block (used by the launcher in cg_api.cu and by
// the host emulation in tests/host_emul).
inline void fill_warp_constants(WarpArgs& a) {
    const cg_params& p = a.p;
    a.range_s = (double)p.window_ms / 1000.0;
    a.rc_range = 1.0 / a.range_s;
    a.sampled = (double)(3 * (int64_t)p.period_ms) / 1000.0;
    a.rc_sampled = 1.0 / a.sampled;
    const double avg = a.sampled / 3.0;     // sampled / (n - 1), n = 4
    a.thr = avg * 1.1;
    a.half_avg = avg / 2;
    for (int q = 0; q < CG_MAX_Q; ++q) {    // Appendix A step 1, in the oracle's order
        const double v = p.q[q];
        a.qkind[q] = v != v ? 1 : v < 0 ? 2 : v > 1 ? 3 : 0;
    }
    a.n_sc_uses = (p.n_scalar + kRowsPerUse - 1) / kRowsPerUse;
    a.n_bk_uses = p.n_buckets > 1 ? (p.n_buckets - 1 + kRowsPerUse - 1) / kRowsPerUse : 0;
    a.uses_per_pass = 2 * a.n_sc_uses + (p.n_buckets > 0 ? 2 * (1 + a.n_bk_uses) : 0);
    a.n_chunks = (p.n_ticks + kChunkTicks - 1) / kChunkTicks;
}

// Extrapolation terms of the three tick slots of one (pair, version, pass).  Tick j0 + 3g + s has
// its window at sample i00 + off[s] + 2g and the slot's terms, for every g.
struct SlotGeom {
    double f[3];        // ext/sampled with the unclamped d_start (Appendix A step 5-6)
    double gchk[3];     // first_v * gchk >= result  ==>  the zero-point clamp cannot fire
    double dstart[3];   // d_start
    double endt[3];     // d_end < thr ? d_end : avg/2
    double rinv[3];     // under-estimate of range_s / f: rank * rinv <= smallest delta reaching the rank
    double gmin;        // min of gchk over the usable slots
    int    off[3];      // window start of slot s relative to slot 0
    int    n4[3];       // the slot's windows hold exactly 4 samples (else: exact fallback)
    int    i00;         // window start of tick j0 (even, so that lane tiles are 16-byte aligned)
    int    j0;          // tick of lane 0, local tick 0 (chunk start minus a rotation of 0..2 ticks)
};

struct SlotTerms { int i0, n4; double f, gchk, dstart, endt, rinv; };

// first sample of the window of tick j (any integer, also negative)
CG_DEVINL int slot_first_sample(const WarpArgs& a, int32_t phase, int j) {
    const cg_params& p = a.p;
    const int lo = p.tick0_ms + j * p.tick_step_ms - p.window_ms;
    return p.left_open ? floor_div32_call(lo - phase, p.period_ms) + 1 : -floor_div32_call(phase - lo, p.period_ms);
}

// Appendix A step 3 and 5 for the 4-sample window of tick j (any integer, also negative).
CG_DEVINL SlotTerms slot_terms(const WarpArgs& a, int32_t phase, int j) {
    const cg_params& p = a.p;
    SlotTerms r;
    const int t = p.tick0_ms + j * p.tick_step_ms, lo = t - p.window_ms;
    const int i0 = slot_first_sample(a, phase, j);
    const int i1 = floor_div32_call(t - phase, p.period_ms);
    r.i0 = i0;
    r.n4 = (i1 - i0 + 1 == 4) ? 1 : 0;
    r.f = 0.0; r.gchk = 1.0; r.dstart = 0.0; r.endt = 0.0; r.rinv = 0.0;
    if (r.n4) {
        const double d_start = div_const((double)(phase + i0 * p.period_ms - lo), 1000.0, 1e-3);
        const double d_end = div_const((double)(t - (phase + i1 * p.period_ms)), 1000.0, 1e-3);
        if (!(d_start > 0.0)) { r.n4 = 0; return r; }
        double ext = a.sampled;
        ext += (d_start < a.thr) ? d_start : a.half_avg;
        const double endt = (d_end < a.thr) ? d_end : a.half_avg;
        ext += endt;
        r.f = div_const(ext, a.sampled, a.rc_sampled);
        r.dstart = d_start;
        r.endt = endt;
        // under-estimates only send more rows to the exact path
        r.gchk = ddiv_call(a.sampled, d_start) * (1.0 - 1e-6);
        r.rinv = ddiv_call(a.range_s, r.f) * (1.0 - 0x1p-40);
    }
    return r;
}

// Where one lane's tile and ticks sit (registers; one per version).
struct LaneGeo {
    int bt;          // first sample of the tile: even, and the tile lies inside the series
    int dl;          // local tick tt uses window 2*(tt/3) + off[tt%3] + dl of the tile
    unsigned act;    // local ticks the fast path evaluates
    unsigned host;   // windows that host at least one of them
};

CG_DEVINL LaneGeo lane_setup(const WarpArgs& a, const SlotGeom& g, int lane, int chunk0, int chunk_end,
                             unsigned& valid_ticks) {
    LaneGeo lg;
    const int b = g.i00 + 2 * kG * lane;
    // The tile starts on an even sample and lies inside the ROW; with an odd number of samples its
    // last sample may be the one just after the series (the pitch is even, so it exists): windows
    // that would use it host no tick, and a reset seen against it only costs a needless correction.
    const int lim = (a.p.n_samples - kTile + 1) & ~1;
    int bt = b < 0 ? 0 : b;
    if (bt > lim) bt = lim;
    lg.bt = bt; lg.dl = b - bt; lg.act = 0u; lg.host = 0u;
    valid_ticks = 0u;
    const int tick_base = g.j0 + kLaneTicks * lane;
#pragma unroll 1
    for (int s = 0; s < 3; ++s) {
        const int c = g.off[s] + lg.dl;
        const bool n4 = g.n4[s] != 0;
#pragma unroll
        for (int kk = 0; kk < kG; ++kk) {
            const int tt = 3 * kk + s, j = tick_base + tt, k = 2 * kk + c;
            if (j >= chunk0 && j < chunk_end) {
                valid_ticks |= 1u << tt;
                if (n4 && k >= 0 && k < kLaneWin && bt + k + 3 < a.p.n_samples) { lg.act |= 1u << tt; lg.host |= 1u << k; }
            }
        }
    }
    return lg;
}

CG_DEVINL int lane_window(const SlotGeom& g, const LaneGeo& lg, int tt) {
    return 2 * (tt / 3) + g.off[tt % 3] + lg.dl;
}

// this lane's column of the shared-memory state
struct Lane {
    int lane;     // lane of the pair, 0 .. kLanes-1
    double* st;   // per-tick state:  st[(field*kLaneTicks + tt)*32]
    double* sa;   // scratch:         sa[k*32] window thresholds (after the last bucket row: the final window maxima)
};
#define CG_ST(L, f, tt) (L).st[((f) * kLaneTicks + (tt)) * kLanes]
#define CG_SY(L, k) (L).sa[(k) * kLanes]

// kernel-wide constants derived from `le` and the quantile list
struct QuantConst {
    double le0, le_top;     // le[0], le[B-2]
    double qdef[CG_MAX_Q];  // result when q is not searched (NaN, -Inf, +Inf, bad bucket shape)
    bool   search[CG_MAX_Q];
};

CG_DEVINL QuantConst quant_const(const WarpArgs& a) {
    QuantConst c;
    const int B = a.p.n_buckets;
    const bool shape_ok = B >= 2 && isinf(a.le[B - 1]) && a.le[B - 1] > 0;   // Appendix A step 2
    c.le0 = B > 0 ? a.le[0] : NAN;
    c.le_top = B >= 2 ? a.le[B - 2] : NAN;
#pragma unroll
    for (int q = 0; q < CG_MAX_Q; ++q) {
        const int kind = a.qkind[q];
        c.search[q] = kind == 0 && shape_ok;
        c.qdef[q] = kind == 2 ? -INFINITY : kind == 3 ? INFINITY : NAN;
    }
    return c;
}

// ---- a position in a series row ----------------------------------------------------------
// On the device the rows sit in the shared-memory ring: the position is a 32-bit shared-memory
// address, so that moving along a row or to the next row is one integer add.  The host emulation
// reads the caller's arrays through a plain pointer.
template <typename T> struct Row {
#ifdef __CUDACC__
    uint32_t s;
    __device__ __forceinline__ Row operator+(int n) const { return Row{s + (uint32_t)(n * (int)sizeof(T))}; }
    __device__ __forceinline__ const T* ptr() const { return reinterpret_cast<const T*>(__cvta_shared_to_generic((size_t)s)); }
#else
    const T* p;
    Row(const T* q) : p(q) {}
    Row operator+(int n) const { return Row(p + n); }
    const T* ptr() const { return p; }
#endif
};

// ---- register tile ---------------------------------------------------------------------
template <typename T> struct TileLoad;
template <> struct TileLoad<double> {
    static CG_DEVMEM void load(const double* p, double (&x)[kTile]) {
#ifdef __CUDACC__
        const double2* q = reinterpret_cast<const double2*>(p);
#pragma unroll
        for (int i = 0; i < kTile / 2; ++i) { const double2 v = q[i]; x[2 * i] = v.x; x[2 * i + 1] = v.y; }
#else
        for (int i = 0; i < kTile; ++i) x[i] = p[i];
#endif
    }
};
template <> struct TileLoad<float> {
    static CG_DEVMEM void load(const float* p, double (&x)[kTile]) {
#ifdef __CUDACC__
        const float2* q = reinterpret_cast<const float2*>(p);
#pragma unroll
        for (int i = 0; i < kTile / 2; ++i) { const float2 v = q[i]; x[2 * i] = v.x; x[2 * i + 1] = v.y; }
#else
        for (int i = 0; i < kTile; ++i) x[i] = p[i];
#endif
    }
};

CG_DEVINL void window_deltas(const double (&x)[kTile], double (&w)[kLaneWin]) {
#pragma unroll
    for (int k = 0; k < kLaneWin; ++k) w[k] = x[k + 3] - x[k];
}

// a counter reset (Appendix A step 2) or an unordered pair (NaN) anywhere in the tile; three
// independent compare chains, so the pipeline is not one serial run of dependent predicates
CG_DEVINL bool tile_has_reset(const double (&x)[kTile]) {
    bool r[3] = {false, false, false};
#pragma unroll
    for (int i = 0; i + 1 < kTile; ++i) r[i % 3] |= !(x[i + 1] >= x[i]);
    return r[0] | r[1] | r[2];
}

// value before each reset (0 where there is none), added to each window in time order; adding
// +0.0 leaves a sum unchanged, so this is the oracle's conditional add
CG_DEVINL void fix_resets(const double (&x)[kTile], double (&w)[kLaneWin]) {
    double d[kTile - 1];
#pragma unroll
    for (int i = 0; i + 1 < kTile; ++i) d[i] = (x[i + 1] < x[i]) ? x[i] : 0.0;
#pragma unroll
    for (int k = 0; k < kLaneWin; ++k) w[k] = ((w[k] + d[k]) + d[k + 1]) + d[k + 2];
}

// v[k] for a run-time k without sending the array to local memory
CG_DEVINL double pick_window(const double (&v)[kLaneWin], int k) {
#ifdef __CUDACC__
    double r = v[0];
#pragma unroll
    for (int i = 1; i < kLaneWin; ++i) r = (k == i) ? v[i] : r;
    return r;
#else
    return v[k];
#endif
}

// Appendix A step 2 for the 4 samples at t4; also returns the first sample
template <typename T>
CG_DEVINL double window4(const Row<T> w4, double& first) {
    const T* __restrict__ t4 = w4.ptr();
    const double x0 = (double)t4[0], x1 = (double)t4[1], x2 = (double)t4[2], x3 = (double)t4[3];
    double res = x3 - x0;
    if (x1 < x0) res += x0;
    if (x2 < x1) res += x1;
    if (x3 < x2) res += x2;
    first = x0;
    return res;
}

// Appendix A steps 4-6 with the clamp, for a regular 4-sample window of a slot
CG_DEVNOINL double exact_increase(double res, double first, double dstart, double endt, double sampled,
                                  double rc_sampled, double thr, double half_avg) {
    if (res > 0 && first >= 0) {
        const double d_zero = sampled * ddiv_pos(first, res);
        if (d_zero < dstart) dstart = d_zero;
    }
    double ext = sampled;
    ext += (dstart < thr) ? dstart : half_avg;
    ext += endt;
    return res * div_const(ext, sampled, rc_sampled);
}

CG_DEVINL bool finite_d(double v) { return exp_in(v, 0u, 2047u); }   // false for Inf and NaN

// smallest window delta that can reach `rank` in rate space, under-estimated; -Inf = look at every row
CG_DEVINL double win_threshold(double rank, double rinv) {
    return (rank > 0.0 && exp_in(rank, 100u, 1800u)) ? rank * rinv : -INFINITY;
}

// increase of one (row, tick): the slot's factor unless the zero-point clamp may fire
CG_DEVINL double tick_increase(const WarpArgs& a, const SlotGeom& g, int s, double res, double first) {
    if (!(first * g.gchk[s] >= res))
        return exact_increase(res, first, g.dstart[s], g.endt[s], a.sampled, a.rc_sampled, a.thr, a.half_avg);
    return res * g.f[s];
}

// where the new version keeps the values of the tick that is local tick tt of this (old-version) lane
CG_DEVINL const double* new_version_column(const Lane& L, int tt, int delta, int& tn) {
    tn = tt + delta;
    int ln = L.lane;
    if (tn < 0) { tn += kLaneTicks; --ln; }
    if (tn >= kLaneTicks) { tn -= kLaneTicks; ++ln; }
    return (ln >= 0 && ln < kLanes) ? L.st + (ln - L.lane) : nullptr;
}

// ---- scalar rows: sum(increase(...)) per role and tick (:375-410), then the derived metrics
// (:387-404).  All K scalar rows of the version are resident (row_ptr(r) -> row r in the ring), so
// the loop is tick-outer and needs five running sums, not sixty.
// New version (ver 0): error_rate and latency_avg wait in the state arrays.  Old version: compared
// on the spot against the new version's values of the same tick, which sit at the tick's
// new-version coordinates; the result is one pass bit per tick.  A missing value (None in the
// reference) is NaN here: both make the gate fail (:430-434, :447, :454).
template <typename T, typename RowFn>
CG_DEVINL void scalar_phase(const WarpArgs& a, const SlotGeom& g, const LaneGeo& lg, const Lane& L, int ver,
                            int64_t pv, int tick_base, int delta /* j0(old) - j0(new) */, RowFn row_ptr,
                            unsigned& pass_bits) {
    const cg_params& p = a.p;
    const int K = p.n_scalar, W = p.n_q + CG_M_EXTRA;
    const bool need_e = (p.checks & CG_G_ERR_OK) != 0, need_a = (p.checks & CG_G_AVG_OK) != 0;
    unsigned have = 0u;
    for (int r = 0; r < K; ++r) have |= p.scalar_role[r];
#pragma unroll 1
    for (int kk = 0; kk < kG; ++kk) {
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const int tt = 3 * kk + s;
            if (!((lg.act >> tt) & 1u)) continue;
            const int off = lg.bt + 2 * kk + g.off[s] + lg.dl;
            // a sum seeded with -0.0 equals "the first term, then + each further term" bit for bit
            double sums[5] = {-0.0, -0.0, -0.0, -0.0, -0.0};
#pragma unroll 1
            for (int r = 0; r < K; ++r) {
                double first;
                const double res = window4<T>(row_ptr(r) + off, first);
                const double inc = tick_increase(a, g, s, res, first);
                const unsigned role = p.scalar_role[r];
#pragma unroll
                for (int q = 0; q < 5; ++q)
                    if (role & (1u << q)) sums[q] += inc;
            }
#pragma unroll
            for (int q = 0; q < 5; ++q)
                if (!(have & (1u << q))) sums[q] = 0.0;            // `or on() vector(0)` / no series with that role
            const double total = sums[0], err = sums[1], lsum = sums[2], lcnt = sums[3];
            double er = NAN, avg = NAN;
            if (total > 0) er = ddiv_pos(err, total);
            if (lcnt > 0) avg = ddiv_pos(lsum, lcnt);
            if (ver == 0) {
                CG_ST(L, kFA, tt) = er;
                CG_ST(L, kFB, tt) = avg;
            } else {
                int tn;
                const double* o = new_version_column(L, tt, delta, tn);
                if (o) {
                    const double er_n = o[(kFA * kLaneTicks + tn) * kLanes], avg_n = o[(kFB * kLaneTicks + tn) * kLanes];
                    const bool ok = (!need_e || er_n <= er * p.one_plus_thr[1]) &&      // :447
                                    (!need_a || avg_n <= avg * p.one_plus_thr[2]);      // :454
                    if (ok) pass_bits |= 1u << tt;
                }
            }
            if (a.metrics) {
                double* o = a.metrics + ((size_t)pv * p.n_ticks + tick_base + tt) * W + p.n_q;
                o[CG_M_ERROR_RESPONSES] = err; o[CG_M_ERROR_RATE] = er; o[CG_M_LATENCY_AVG] = avg;
                o[CG_M_REQUEST_COUNT] = lcnt; o[CG_M_FEEDBACK_COUNT] = sums[4];
            }
            if (a.valid)
                a.valid[(size_t)pv * p.n_ticks + tick_base + tt] =
                    (uint8_t)((p.n_buckets > 0 ? CG_V_QUANTILE : 0) | (total > 0 ? CG_V_ERROR_RATE : 0) |
                              (lcnt > 0 ? CG_V_LATENCY_AVG : 0));
        }
    }
}

// The gate quantile of a tick is known (captured, or the default when no bucket reaches the rank).
// New version: it waits in the state array.  Old version: the quantile part of the gate (:440) is
// decided on the spot against the new version's value of the same tick.
CG_DEVINL void gate_quantile_known(const WarpArgs& a, const Lane& L, int ver, int tt, int delta, double r,
                                   unsigned& pass_bits) {
    if (ver == 0) {
        CG_ST(L, kFSQ, tt) = r;
    } else {
        int tn;
        const double* o = new_version_column(L, tt, delta, tn);
        const bool ok = o && o[(kFSQ * kLaneTicks + tn) * kLanes] <= r * a.p.one_plus_thr[0];
        pass_bits = ok ? pass_bits | (1u << tt) : pass_bits & ~(1u << tt);
    }
}

// ---- the +Inf bucket row: observations, rank, window thresholds -----------------------------
// Speculates that the +Inf bucket holds the observations (checked in bucket_finish).
template <typename T, int NQ>
CG_DEVINL void inf_row_setup(const WarpArgs& a, const SlotGeom& g, LaneGeo& lg, const Lane& L,
                             const QuantConst& qc, const Row<T> tile, int ver, int delta, int64_t pv,
                             int tick_base, uint64_t& found, unsigned& fb, unsigned& pass_bits) {
    const cg_params& p = a.p;
    const int W = p.n_q + CG_M_EXTRA;
#pragma unroll
    for (int k = 0; k < kLaneWin; ++k) CG_SY(L, k) = INFINITY;
    found = ~0ull;
#pragma unroll 1
    for (int kk = 0; kk < kG; ++kk) {
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const int tt = 3 * kk + s;
            if (!((lg.act >> tt) & 1u)) continue;
            const int k = 2 * kk + g.off[s] + lg.dl;
            double x0;
            const double wv = window4<T>(tile + k, x0);
            double obs = NAN;
            if (finite_d(wv) && finite_d(x0)) obs = div_const(tick_increase(a, g, s, wv, x0), a.range_s, a.rc_range);
            if (!finite_d(obs)) { fb |= 1u << tt; lg.act &= ~(1u << tt); continue; }
            CG_ST(L, kFA, tt) = obs;
            CG_ST(L, kFB, tt) = -INFINITY;
            double y = CG_SY(L, k);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const bool searching = qc.search[q] && obs != 0;                      // Appendix A step 4
                const double def = qc.search[q] ? (obs != 0 ? qc.le_top : NAN) : qc.qdef[q];   // step 6: no bucket reaches the rank
                if (q == p.gate_q) gate_quantile_known(a, L, ver, tt, delta, def, pass_bits);
                if (a.metrics) a.metrics[((size_t)pv * p.n_ticks + tick_base + tt) * W + q] = def;
                if (searching) {
                    found &= ~(1ull << (tt * NQ + q));
                    y = fmin(y, win_threshold(p.q[q] * obs, g.rinv[s]));               // step 5
                }
            }
            CG_SY(L, k) = y;
        }
    }
}

// histogram_quantile steps 6-7 once bucket b is the first to reach the rank
CG_DEVNOINL double interpolate_quantile(const double* __restrict__ le, int b, double c, double fixed_prev, double rank) {
    if (b == 0 && le[0] <= 0) return le[0];               // step 6
    double start = 0.0, end = le[b], cc = c, rk = rank;    // step 7
    if (b > 0) { start = le[b - 1]; cc -= fixed_prev; rk -= fixed_prev; }
    return start + (end - start) * ddiv_pos(rk, cc);
}

// ---- bucket rows: the event loop --------------------------------------------------------------
// Exact, per tick and in rate space, for the flagged windows of bucket row b.  prevw holds the
// windows' running maxima before this row.  Returns the windows that must be kept out of the
// running maximum for this row (a clamped tick: the row's values went into the per-tick maxima
// instead).
template <typename T, int NQ>
CG_DEVINL unsigned bucket_events(const WarpArgs& a, const SlotGeom& g, LaneGeo& lg, const Lane& L,
                                 const Row<T> tile, int b, unsigned mask, const double (&prevw)[kLaneWin],
                                 int ver, int delta, int64_t pv, int tick_base, uint64_t& found, unsigned& fb,
                                 unsigned& pass_bits) {
    const cg_params& p = a.p;
    const int W = p.n_q + CG_M_EXTRA;
    unsigned excl = 0u;
#pragma unroll 1
    while (mask) {
        const int k = __ffs(mask) - 1;
        mask &= mask - 1;
        CG_EMUL_STAT(windows, 1);
        double x0;
        const double res = window4<T>(tile + k, x0);
        const bool fin = finite_d(res) && finite_d(x0);
        unsigned hosted = 0u, clamped = 0u;
        double y = INFINITY;
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const int num = k - lg.dl - g.off[s];
            if ((num & ~(2 * (kG - 1))) != 0) continue;           // num in {0, 2, .., 2(kG-1)}: the window hosts a tick of slot s
            const int tt = 3 * (num >> 1) + s;
            if (!((lg.act >> tt) & 1u)) continue;
            if (!fin) { fb |= 1u << tt; lg.act &= ~(1u << tt); continue; }
            hosted |= 1u << s;
            const double fs = g.f[s];
            double inc = res * fs;
            if (!(x0 * g.gchk[s] >= res)) {
                const double e = exact_increase(res, x0, g.dstart[s], g.endt[s], a.sampled, a.rc_sampled, a.thr, a.half_avg);
                if (e != inc) { clamped |= 1u << s; inc = e; }
            }
            const double c = div_const(inc, a.range_s, a.rc_range);
            const double obs = CG_ST(L, kFA, tt);
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                const uint64_t bit = 1ull << (tt * NQ + q);
                if (found & bit) continue;
                const double rank = p.q[q] * obs;                 // step 5
                if (c >= rank) {                                  // first bucket reaching the rank
                    double fixed_prev = -INFINITY;
                    if (b > 0) {
                        fixed_prev = CG_ST(L, kFB, tt);
                        const double pw = pick_window(prevw, k);
                        if (pw > -INFINITY) fixed_prev = fmax(div_const(pw * fs, a.range_s, a.rc_range), fixed_prev);
                    }
                    const double r = interpolate_quantile(a.le, b, c, fixed_prev, rank);
                    found |= bit;
                    CG_EMUL_STAT(captures, 1);
                    if (q == p.gate_q) gate_quantile_known(a, L, ver, tt, delta, r, pass_bits);
                    if (a.metrics) a.metrics[((size_t)pv * p.n_ticks + tick_base + tt) * W + q] = r;
                } else {
                    y = fmin(y, win_threshold(rank, g.rinv[s]));
                }
            }
            if ((clamped >> s) & 1u) CG_ST(L, kFB, tt) = fmax(CG_ST(L, kFB, tt), c);
        }
        if (clamped) {
            // the window stays out of the running maximum for this row: ticks of the window that
            // were not clamped themselves put their value into their per-tick maximum, too
            excl |= 1u << k;
            const unsigned rest = hosted & ~clamped;
#pragma unroll 1
            for (int s = 0; s < 3; ++s) {
                if (!((rest >> s) & 1u)) continue;
                const int tt = 3 * ((k - lg.dl - g.off[s]) >> 1) + s;
                CG_ST(L, kFB, tt) = fmax(CG_ST(L, kFB, tt), div_const(res * g.f[s], a.range_s, a.rc_range));
            }
        }
        CG_EMUL_STAT(clamps, clamped ? 1 : 0);
        CG_SY(L, k) = y;
    }
    return excl;
}

// After the last bucket row: the speculation "the +Inf bucket holds the observations" is checked
// per tick (Appendix A step 3-4: observations = the running maximum over ALL buckets).  The caller
// has left the final window maxima in the scratch column.
CG_DEVINL void bucket_finish(const WarpArgs& a, const SlotGeom& g, LaneGeo& lg, const Lane& L, unsigned& fb) {
#pragma unroll 1
    for (int kk = 0; kk < kG; ++kk) {
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const int tt = 3 * kk + s;
            if (!((lg.act >> tt) & 1u)) continue;
            const double pvw = CG_SY(L, 2 * kk + g.off[s] + lg.dl);
            double fixed = CG_ST(L, kFB, tt);
            if (pvw > -INFINITY) fixed = fmax(div_const(pvw * g.f[s], a.range_s, a.rc_range), fixed);
            if (!(fixed <= CG_ST(L, kFA, tt))) { fb |= 1u << tt; lg.act &= ~(1u << tt); }
        }
    }
}

// one bucket row in window space; prevw = running maxima of the window deltas (monotonic fix),
// ymin = smallest delta that may reach a rank still searched
template <typename T, int NQ, typename AnyFn>
CG_DEVINL void bucket_row(const WarpArgs& a, const SlotGeom& g, LaneGeo& lg, const Lane& L,
                          const Row<T> tile, int b, int ver, int delta, int64_t pv, int tick_base,
                          double (&prevw)[kLaneWin], double (&ymin)[kLaneWin], uint64_t& found, unsigned& fb,
                          unsigned& pass_bits, AnyFn any) {
    double x[kTile], w[kLaneWin];
    TileLoad<T>::load(tile.ptr(), x);
    window_deltas(x, w);
    const double gmin = g.gmin;
    bool cflag;
    const bool rflag = tile_has_reset(x);
    CG_EMUL_STAT(reset, rflag ? 1 : 0);
    if (any(rflag)) {
        fix_resets(x, w);
        bool c[3] = {false, false, false};
#pragma unroll
        for (int k = 0; k < kLaneWin; ++k) c[k % 3] |= !(x[k] * gmin >= w[k]);
        cflag = c[0] | c[1] | c[2];
    } else {
        // no reset in the tile: samples are non-decreasing, so every window delta is at most
        // x[last]-x[0] and every first sample at least x[0]
        cflag = !(x[0] * gmin >= x[kTile - 1] - x[0]);
    }
    bool h[3] = {false, false, false};
#pragma unroll
    for (int k = 0; k < kLaneWin; ++k) h[k % 3] |= !(w[k] < ymin[k]);
    const bool hflag = h[0] | h[1] | h[2];
    const bool ev = (cflag | hflag) && lg.host != 0u;
    CG_EMUL_STAT(event, ev ? 1 : 0);
    CG_EMUL_STAT(cfl, (cflag && lg.host != 0u) ? 1 : 0);
    CG_EMUL_STAT(hfl, (hflag && lg.host != 0u) ? 1 : 0);
    if (any(ev)) {
        unsigned mask = 0u;
#pragma unroll
        for (int k = 0; k < kLaneWin; ++k)
            if (!(x[k] * gmin >= w[k]) || !(w[k] < ymin[k])) mask |= 1u << k;
        const unsigned excl = bucket_events<T, NQ>(a, g, lg, L, tile, b, mask & lg.host, prevw, ver, delta, pv, tick_base,
                                                   found, fb, pass_bits);
#pragma unroll
        for (int k = 0; k < kLaneWin; ++k) {
            if ((excl >> k) & 1u) w[k] = -INFINITY;
            ymin[k] = CG_SY(L, k);
        }
    }
#pragma unroll
    for (int k = 0; k < kLaneWin; ++k) prevw[k] = w[k] > prevw[k] ? w[k] : prevw[k];   // NaN deltas never enter
}

// 12 pass bits of a lane (local tick order) -> positions tick_base.. of a packed bit vector
struct BitField { int word0; uint32_t lo, hi; };
CG_DEVINL BitField place_bits(unsigned field, int tick_base) {
    BitField r;
    if (tick_base < 0) { field >>= -tick_base; tick_base = 0; }
    r.word0 = tick_base >> 5;
    const int sh = tick_base & 31;
    r.lo = field << sh;
    r.hi = sh > 32 - kLaneTicks ? field >> (32 - sh) : 0u;
    return r;
}

// one entry of the per-pass fill table: where a ring slot's rows come from
struct FillEntry {
    uint32_t off;     // byte offset from the pair's first scalar / bucket row
    uint32_t meta;    // bytes, bit 31: bucket rows
};
CG_DEVINL FillEntry fill_entry(const WarpArgs& a, int u) {
    const int K = a.p.n_scalar, B = a.p.n_buckets, n_sc = a.n_sc_uses, n_bk = a.n_bk_uses;
    FillEntry e;
    int rows = 1;
    if (u < 2 * n_sc) {
        const int ver = u >= n_sc ? 1 : 0, k0 = (u - ver * n_sc) * kRowsPerUse;
        rows = K - k0 < kRowsPerUse ? K - k0 : kRowsPerUse;
        e.off = (uint32_t)(ver * K + k0) * a.row_bytes;
        e.meta = (uint32_t)rows * a.row_bytes;
    } else {
        u -= 2 * n_sc;
        const int ver = u >= 1 + n_bk ? 1 : 0, uu = u - ver * (1 + n_bk);
        int b0 = B - 1;                                          // the +Inf bucket goes first
        if (uu > 0) { b0 = (uu - 1) * kRowsPerUse; rows = B - 1 - b0 < kRowsPerUse ? B - 1 - b0 : kRowsPerUse; }
        e.off = (uint32_t)(ver * B + b0) * a.row_bytes;
        e.meta = ((uint32_t)rows * a.row_bytes) | 0x80000000u;
    }
    return e;
}

#ifdef __CUDACC__
// =============================================================================================
// device side
// =============================================================================================

// ---- warp-cooperative exact evaluation of one (pair, version, tick) from global memory: the
// plain sequential semantics (tick_geom, series_increase, bucketQuantile), two buckets per lane.
// one out-of-line copy of the window walk for the three places below (this path is cold; small code matters more)
template <typename T>
__device__ __noinline__ double series_increase_cold(const T* __restrict__ row, const TickGeom& g) {
    double first_v = (double)row[g.i0];
    double result = (double)row[g.i0 + g.n - 1] - first_v, prev = first_v;
#pragma unroll 1
    for (int i = 1; i < g.n; ++i) {
        const double v = (double)row[g.i0 + i];
        if (v < prev) result += prev;      // counter reset
        prev = v;
    }
    return extrapolate(result, first_v, g);
}

template <typename T>
__device__ __noinline__ void coop_eval_tick(const WarpArgs& a, int64_t pv, int32_t phase, int tick, int lane,
                                            double* m, uint8_t& valid_out) {
    const cg_params& p = a.p;
    const int nq = p.n_q, K = p.n_scalar, B = p.n_buckets;
    const TickGeom g = tick_geom(p, phase, tick);
    const T* sc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * p.row_pitch;
    const T* bk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * p.row_pitch;
    const double* le = a.le;
    uint8_t valid = 0;
    double sums[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    unsigned have = 0;
    if (g.n >= 2) {
        double mine = 0.0;
        if (lane < K) mine = series_increase_cold<T>(sc + (size_t)lane * p.row_pitch, g);
        for (int k = 0; k < K; ++k) {
            const double inc = __shfl_sync(kFullMask, mine, k);
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
        const bool top_inf = isinf(le[B - 1]) && le[B - 1] > 0;
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
        if (scan) {   // B <= 64 (eligibility test): lane l holds buckets l and l+32
            double v0 = -INFINITY, v1 = -INFINITY;
            if (lane < B) v0 = ddiv_call(series_increase_cold<T>(bk + (size_t)lane * p.row_pitch, g), range_s);
            if (lane + 32 < B) v1 = ddiv_call(series_increase_cold<T>(bk + (size_t)(lane + 32) * p.row_pitch, g), range_s);
#pragma unroll 1
            for (int d = 1; d < 32; d <<= 1) {        // running max in bucket order (monotonic fix)
                const double o0 = __shfl_up_sync(kFullMask, v0, d), o1 = __shfl_up_sync(kFullMask, v1, d);
                if (lane >= d && o0 > v0) v0 = o0;
                if (lane >= d && o1 > v1) v1 = o1;
            }
            const double carry = __shfl_sync(kFullMask, v0, 31);
            if (B > 32 && carry > v1) v1 = carry;
            const double obs = B <= 32 ? __shfl_sync(kFullMask, v0, B - 1) : __shfl_sync(kFullMask, v1, B - 33);
            if (obs == 0) {
                for (int j = 0; j < nq; ++j) if (!found[j]) { m[j] = NAN; found[j] = true; }
            } else {
                for (int j = 0; j < nq; ++j) {
                    if (found[j]) continue;
                    const double rank = p.q[j] * obs;
                    const unsigned hit0 = __ballot_sync(kFullMask, lane < B - 1 && v0 >= rank);
                    const unsigned hit1 = __ballot_sync(kFullMask, lane + 32 < B - 1 && v1 >= rank);
                    if (!hit0 && !hit1) { m[j] = le[B - 2]; continue; }
                    const int src = hit0 ? __ffs(hit0) - 1 : __ffs(hit1) - 1;
                    const int bb = hit0 ? src : src + 32;
                    const double cbv = __shfl_sync(kFullMask, hit0 ? v0 : v1, src);
                    const double up = __shfl_sync(kFullMask, hit0 ? v0 : v1, src > 0 ? src - 1 : 0);
                    const double cpv = src > 0 ? up : carry;     // (bb == 32: the last bucket of the lower half)
                    if (bb == 0 && le[0] <= 0) m[j] = le[0];
                    else {
                        double start = 0.0, end = le[bb], cc = cbv, rk = rank;
                        if (bb > 0) { start = le[bb - 1]; cc -= cpv; rk -= cpv; }
                        m[j] = start + (end - start) * ddiv_call(rk, cc);
                    }
                }
            }
        }
    }
    const double total = sums[0], err = sums[1], lat_sum = sums[2], lat_cnt = sums[3], fbk = sums[4];
    m[nq + CG_M_ERROR_RESPONSES] = err;
    if (total > 0) { m[nq + CG_M_ERROR_RATE] = ddiv_call(err, total); valid |= CG_V_ERROR_RATE; }
    else m[nq + CG_M_ERROR_RATE] = NAN;
    if (lat_cnt > 0) { m[nq + CG_M_LATENCY_AVG] = ddiv_call(lat_sum, lat_cnt); valid |= CG_V_LATENCY_AVG; }
    else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = lat_cnt;
    m[nq + CG_M_FEEDBACK_COUNT] = fbk;
    valid_out = valid;
}

// Extrapolation terms of the three tick slots, computed by lanes 0..2.  The origin tick j0 is the
// chunk start moved back by 0..2 ticks so that its window starts on an even sample.
__device__ __forceinline__ void warp_geometry(const WarpArgs& a, int32_t phase, int chunk0, int lane, SlotGeom* g) {
    const bool phase_ok = phase > -(1 << 30) && phase < (1 << 30);
    SlotTerms t;
    t.i0 = 1; t.n4 = 0; t.f = 0.0; t.gchk = 1.0; t.dstart = 0.0; t.endt = 0.0; t.rinv = 0.0;
    if (lane < 3 && phase_ok) t.i0 = slot_first_sample(a, phase, chunk0 - lane);
    const unsigned even = __ballot_sync(kFullMask, lane < 3 && !(t.i0 & 1));
    const int c = even ? __ffs(even) - 1 : 0;
    const int j0 = chunk0 - c;
    if (lane < 3 && phase_ok) t = slot_terms(a, phase, j0 + lane);
    const int i00 = __shfl_sync(kFullMask, t.i0, 0);
    const bool usable = lane < 3 && t.n4 && phase_ok && !(i00 & 1);
    double gm = usable ? t.gchk : INFINITY;
    gm = fmin(gm, __shfl_xor_sync(kFullMask, gm, 1));
    gm = fmin(gm, __shfl_xor_sync(kFullMask, gm, 2));
    if (lane < 3) {
        g->f[lane] = t.f; g->gchk[lane] = t.gchk; g->dstart[lane] = t.dstart; g->endt[lane] = t.endt;
        g->rinv[lane] = t.rinv; g->off[lane] = t.i0 - i00; g->n4[lane] = usable ? 1 : 0;
    }
    if (lane == 0) { g->i00 = i00; g->j0 = j0; g->gmin = gm < INFINITY ? gm : 1.0; }
}

struct WarpAny {
    __device__ __forceinline__ bool operator()(bool f) const { return __any_sync(kFullMask, f) != 0; }
};

// position in the warp's stream of ring fills
struct FillCursor {
    int64_t pair;
    int chunk, u;
    const char* sc;     // first scalar / bucket row of the pair
    const char* bk;
};

// CTA-wide barrier that a single-warp CTA does not need
__device__ __forceinline__ void pair_sync() {
    if (kWarps > 1) __syncthreads();
    else __syncwarp();
}

#ifndef CG_MIN_CTAS
#define CG_MIN_CTAS (kG == 4 ? 7 : kG == 2 ? 8 : 4)
#endif
#ifdef CG_MAXNREG
#define CG_REG_BUDGET __maxnreg__(CG_MAXNREG)      // developer build: explicit register cap
#else
#define CG_REG_BUDGET __launch_bounds__(kLanes, CG_MIN_CTAS)
#endif
template <typename T, int NQ>
__global__ void CG_REG_BUDGET
warp_kernel(const __grid_constant__ WarpArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    const cg_params& p = a.p;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + a.off_bars);
    SlotGeom* geo = reinterpret_cast<SlotGeom*>(smem + a.off_geom) + 2 * warp;   // each warp keeps its own copy
    const int n_words = (p.n_ticks + 31) >> 5;
    uint32_t* s_bits = reinterpret_cast<uint32_t*>(smem + a.off_bits);   // promote bits of the pair
    uint32_t* s_fb = s_bits + n_words;                                   // ticks for the exact fallback
    const FillEntry* s_tab = reinterpret_cast<const FillEntry*>(smem + a.off_tab);
    Lane L;
    L.lane = tid;
    L.st = reinterpret_cast<double*>(smem + a.off_state) + tid;
    L.sa = reinterpret_cast<double*>(smem + a.off_scratch) + tid;
    const int B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    const int n_sc = a.n_sc_uses, n_bk = a.n_bk_uses, n_slots = a.n_slots;
    const WarpAny any;
    if (tid == 0) {
        for (int s = 0; s < n_slots; ++s) mbar_init(bars + s, 1);
        mbar_init_fence();
    }
    for (int u = tid; u < a.uses_per_pass; u += kLanes) reinterpret_cast<FillEntry*>(smem + a.off_tab)[u] = fill_entry(a, u);
    pair_sync();
    const QuantConst qc = quant_const(a);

    // ---- the ring: the pair's warps are its consumers, thread 0 its producer -----------------
    const size_t pair_bytes_sc = (size_t)2 * p.n_scalar * a.row_bytes, pair_bytes_bk = (size_t)2 * B * a.row_bytes;
    auto skip_pairs = [&](FillCursor& c) {
        while (c.pair < a.n_pairs && a.has_previous && !a.has_previous[c.pair]) c.pair += gridDim.x;
        c.sc = reinterpret_cast<const char*>(a.scalars) + c.pair * pair_bytes_sc;
        c.bk = reinterpret_cast<const char*>(a.buckets) + c.pair * pair_bytes_bk;
    };
    auto cursor_next = [&](FillCursor& c) {
        if (++c.u == a.uses_per_pass) {
            c.u = 0;
            if (++c.chunk == a.n_chunks) { c.chunk = 0; c.pair += gridDim.x; skip_pairs(c); }
        }
    };
    const uint32_t ring_s = smem_u32(smem), bars_s = ring_s + a.off_bars;   // the ring sits at offset 0
    auto issue_fill = [&](const FillCursor& c, uint32_t slot_s, uint32_t bar_s) {      // one thread
        const FillEntry e = s_tab[c.u];
        const char* src = ((e.meta >> 31) ? c.bk : c.sc) + e.off;
        const uint32_t bytes = e.meta & 0x7fffffffu;
        CG_CHECK(0, c.u >= 0 && c.u < a.uses_per_pass && slot_s - ring_s < (uint32_t)n_slots * a.slot_bytes && bytes <= a.slot_bytes && bytes % 16 == 0);
        CG_CHECK(1, (e.meta >> 31) ? (src >= (const char*)a.buckets && src + bytes <= (const char*)a.buckets + a.n_pairs * pair_bytes_bk)
                                   : (src >= (const char*)a.scalars && src + bytes <= (const char*)a.scalars + a.n_pairs * pair_bytes_sc));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // reads of the slot before the async write
        mbar_arrive_expect_tx_s(bar_s, bytes);
        bulk_g2s_s(slot_s, src, bytes, bar_s);
    };
    FillCursor fill;
    fill.pair = blockIdx.x; fill.chunk = 0; fill.u = 0;
    skip_pairs(fill);
    for (int s = 0; s < n_slots; ++s) {          // prologue: fill the whole ring
        if (tid == 0 && fill.pair < a.n_pairs) issue_fill(fill, ring_s + (uint32_t)s * a.slot_bytes, bars_s + 8u * (uint32_t)s);
        cursor_next(fill);
    }
    // oldest slot in use: its index and the parity of its mbarrier
    uint32_t head = 0, head_parity = 0;
    auto slot_wait = [&](int i) -> uint32_t {     // i-th slot from the head; returns the shared address of its rows
        uint32_t s = head + (uint32_t)i, par = head_parity;
        if (s >= (uint32_t)n_slots) { s -= n_slots; par ^= 1u; }
        mbar_wait_s(bars_s + 8u * s, par);
        return ring_s + s * a.slot_bytes;
    };
    auto head_wait = [&]() -> uint32_t {
        mbar_wait_s(bars_s + 8u * head, head_parity);
        return ring_s + head * a.slot_bytes;
    };
    auto slot_release = [&]() {                   // the head slot; callers have synchronised the pair's warps
        if (tid == 0 && fill.pair < a.n_pairs) issue_fill(fill, ring_s + head * a.slot_bytes, bars_s + 8u * head);
        cursor_next(fill);
        const bool wrap = ++head == (uint32_t)n_slots;
        head = wrap ? 0u : head;
        head_parity ^= wrap ? 1u : 0u;
    };

    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
        if (!hp) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            if (tid == 0 && a.decisions) a.decisions[pair] = replay_schedule([](int) { return false; }, false, st_tick, p);
            if (a.gate_bits) {
#pragma unroll 1
                for (int w = tid; w < n_words; w += kLanes) a.gate_bits[pair * n_words + w] = 0u;
            }
            continue;
        }
#pragma unroll 1
        for (int w = tid; w < 2 * n_words; w += kLanes) s_bits[w] = 0u;
        const int32_t phase0 = a.phase[pair * 2], phase1 = a.phase[pair * 2 + 1];
#pragma unroll 1
        for (int chunk = 0; chunk < a.n_chunks; ++chunk) {
            const int chunk0 = chunk * kChunkTicks, chunk_end = min(p.n_ticks, chunk0 + kChunkTicks);
            __syncwarp();
            warp_geometry(a, phase0, chunk0, lane, geo);
            warp_geometry(a, phase1, chunk0, lane, geo + 1);
            __syncwarp();
            const int delta = geo[1].j0 - geo[0].j0;
            unsigned pass_sc = 0u, pass_q = 0u, act_old = 0u;
            unsigned fbm[2];
            LaneGeo lg[2];
#pragma unroll
            for (int ver = 0; ver < 2; ++ver) {
                unsigned vt;
                lg[ver] = lane_setup(a, geo[ver], tid, chunk0, chunk_end, vt);
                fbm[ver] = vt & ~lg[ver].act;
                CG_CHECK(3, lg[ver].bt >= 0 && lg[ver].bt + kTile <= p.row_pitch && !(lg[ver].bt & 1) && !(lg[ver].act & ~vt));
            }

            // ---- scalar rows of both versions: all K rows of a version resident, then released together
#pragma unroll 1
            for (int ver = 0; ver < 2; ++ver) {
                for (int i = 0; i < n_sc; ++i) slot_wait(i);
                const uint32_t h = head;
                auto row_ptr = [&](int r) -> Row<T> {
                    uint32_t s = h + (uint32_t)(r / kRowsPerUse);
                    if (s >= (uint32_t)n_slots) s -= n_slots;
                    CG_CHECK(2, r >= 0 && r < p.n_scalar && s < (uint32_t)n_slots);
                    return Row<T>{ring_s + s * a.slot_bytes + (uint32_t)(r % kRowsPerUse) * a.row_bytes};
                };
                scalar_phase<T>(a, geo[ver], lg[ver], L, ver, pair * 2 + ver, geo[ver].j0 + kLaneTicks * tid, delta,
                                row_ptr, pass_sc);
                pair_sync();
                for (int i = 0; i < n_sc; ++i) slot_release();
            }

            // ---- bucket rows: per version the +Inf row, then rows 0..B-2
            if (B > 0) {
#pragma unroll 1
                for (int ver = 0; ver < 2; ++ver) {
                    const SlotGeom& g = geo[ver];
                    const int64_t pv = pair * 2 + ver;
                    const int tick_base = g.j0 + kLaneTicks * tid;
                    LaneGeo l = lg[ver];
                    unsigned fb = fbm[ver];
                    double prevw[kLaneWin], ymin[kLaneWin];
                    uint64_t found = ~0ull;
                    const uint32_t tile_off = (uint32_t)l.bt * (uint32_t)sizeof(T);
                    {
                        const Row<T> tile{head_wait() + tile_off};
                        inf_row_setup<T, NQ>(a, g, l, L, qc, tile, ver, delta, pv, tick_base, found, fb, pass_q);
#pragma unroll
                        for (int k = 0; k < kLaneWin; ++k) { ymin[k] = CG_SY(L, k); prevw[k] = -INFINITY; }
                        pair_sync();
                        slot_release();
                    }
#pragma unroll 1
                    for (int u = 0; u < n_bk; ++u) {
                        Row<T> tile{head_wait() + tile_off};
                        const int first = u * kRowsPerUse, nr = min(kRowsPerUse, B - 1 - first);
#pragma unroll 1
                        for (int r = 0; r < nr; ++r) {
                            CG_CHECK(4, tile.s >= ring_s && tile.s + kTile * sizeof(T) <= ring_s + (uint32_t)n_slots * a.slot_bytes &&
                                        first + r < B - 1);
                            bucket_row<T, NQ>(a, g, l, L, tile, first + r, ver, delta, pv, tick_base,
                                              prevw, ymin, found, fb, pass_q, any);
                            tile.s += a.row_bytes;
                        }
                        pair_sync();
                        slot_release();
                    }
#pragma unroll
                    for (int k = 0; k < kLaneWin; ++k) CG_SY(L, k) = prevw[k];
                    bucket_finish(a, g, l, L, fb);
                    fbm[ver] = fb;
                    if (ver == 1) act_old = l.act;
                    pair_sync();
                }
            } else {
                act_old = lg[1].act;
            }
            if (!(p.checks & CG_G_P95_OK)) pass_q = (1u << kLaneTicks) - 1u;

            // ---- this pass's bits: promote bits from the old version's lanes, fallback ticks from both
            {
                const BitField pb = place_bits(pass_sc & pass_q & act_old, geo[1].j0 + kLaneTicks * tid);
                CG_CHECK(5, (!pb.lo || (pb.word0 >= 0 && pb.word0 < n_words)) && (!pb.hi || pb.word0 + 1 < n_words));
                if (pb.lo) atomicOr(s_bits + pb.word0, pb.lo);
                if (pb.hi) atomicOr(s_bits + pb.word0 + 1, pb.hi);
#pragma unroll
                for (int ver = 0; ver < 2; ++ver) {
                    const BitField f = place_bits(fbm[ver], geo[ver].j0 + kLaneTicks * tid);
                    CG_CHECK(6, (!f.lo || (f.word0 >= 0 && f.word0 < n_words)) && (!f.hi || f.word0 + 1 < n_words));
                    if (f.lo) atomicOr(s_fb + f.word0, f.lo);
                    if (f.hi) atomicOr(s_fb + f.word0 + 1, f.hi);
                }
            }
        }
        pair_sync();
        // ---- exact re-evaluation of whatever the fast path could not decide (the pair's warps take
        // the fallback ticks in turn)
        int fb_seq = 0;
#pragma unroll 1
        for (int wd = 0; wd < n_words; ++wd) {
            uint32_t todo = s_fb[wd];
            while (todo) {
                const int bit = __ffs(todo) - 1;
                todo &= todo - 1;
                if ((fb_seq++ % kWarps) != warp) continue;
                const int tick = wd * 32 + bit;
                double mn[CG_MAX_Q + CG_M_EXTRA], mo[CG_MAX_Q + CG_M_EXTRA];
                uint8_t vn, vo;
                coop_eval_tick<T>(a, pair * 2, phase0, tick, lane, mn, vn);
                coop_eval_tick<T>(a, pair * 2 + 1, phase1, tick, lane, mo, vo);
                if (lane == 0) {
                    const double nm[3] = {p.n_q > 0 ? mn[p.gate_q] : NAN, mn[p.n_q + CG_M_ERROR_RATE], mn[p.n_q + CG_M_LATENCY_AVG]};
                    const double om[3] = {p.n_q > 0 ? mo[p.gate_q] : NAN, mo[p.n_q + CG_M_ERROR_RATE], mo[p.n_q + CG_M_LATENCY_AVG]};
                    if (gate_eval(nm, vn, om, vo, p) & CG_G_PROMOTE) atomicOr(s_bits + wd, 1u << bit);
                    else atomicAnd(s_bits + wd, ~(1u << bit));
                    if (a.metrics) {
                        double* on = a.metrics + ((size_t)(pair * 2) * p.n_ticks + tick) * W;
                        double* oo = a.metrics + ((size_t)(pair * 2 + 1) * p.n_ticks + tick) * W;
#pragma unroll 1
                        for (int c = 0; c < W; ++c) { on[c] = mn[c]; oo[c] = mo[c]; }
                    }
                    if (a.valid) {
                        a.valid[(size_t)(pair * 2) * p.n_ticks + tick] = vn;
                        a.valid[(size_t)(pair * 2 + 1) * p.n_ticks + tick] = vo;
                    }
                }
            }
        }
        pair_sync();
        // ---- the step loop (:296-361) over the pair's promote bits
        if (a.gate_bits) {
#pragma unroll 1
            for (int w = tid; w < n_words; w += kLanes) a.gate_bits[pair * n_words + w] = s_bits[w];
        }
        if (tid == 0 && a.decisions) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            a.decisions[pair] = replay_schedule_bits(s_bits, n_words, true, st_tick, p);
        }
        pair_sync();
    }
}

// Host-side launcher for one (element type, quantile count) instance.
template <typename T, int NQ>
cudaError_t launch_warp_tn(const WarpArgs& a, int grid, cudaStream_t s) {
    auto kern = warp_kernel<T, NQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a.smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<grid, kLanes, a.smem_bytes, s>>>(a);
    return cudaGetLastError();
}
#endif  // __CUDACC__

}  // namespace cg
