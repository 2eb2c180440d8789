// cg_pairstream.cuh -- the PAIR-PARALLEL streaming kernel: both versions of a model pair are
// evaluated at the same time by one CTA, and a thread owns TWO periods of the tick grid.
//
// Same arithmetic, same ring, same exact fallback as cg_stream.cuh (which it supersedes for tick
// grids of up to 384 ticks); what changes is the mapping of work to threads:
//   * CTA = 4 warps.  Warps 0-1 evaluate the NEW version (array index 0), warps 2-3 the OLD one.
//     A ring slot carries kRowsPerVer rows of each version, so the two halves of the CTA stream
//     the same row indices of their own series side by side; the warp that leaves a slot last
//     issues the two bulk copies that refill it.
//   * A thread owns 6 consecutive ticks (two periods of 3): their 4-sample windows live in 8
//     consecutive samples, four 128-bit shared loads per row.  The per-row overheads (address
//     arithmetic, the warp votes for the rare paths, loop control) are paid once per 6 ticks
//     instead of once per 3, which is what the previous kernel was bound by.
//   * When a thread reaches its rank it finishes histogram_quantile on the spot (a rare block),
//     so only the result is kept, not the captured counts.
//   * After the last row both halves post their per-tick gate inputs in a shared exchange area;
//     the warp that finishes the pair last evaluates should_promote_model for every tick
//     (:419-460), builds the gate bit-vector and replays the step loop (:296-361).  No CTA
//     barrier in the steady state.
#pragma once
#include "cg_stream.cuh"

namespace cg {

constexpr int kP6 = 2 * kG;          // ticks per thread
constexpr int kRowsPerVer = 2;       // rows of each version per ring slot
constexpr int kHalf = 64;            // threads per version
#ifndef CG_PAIR_MINB
#define CG_PAIR_MINB 3
#endif

struct PairArgs {
    const void*    scalars;
    const void*    buckets;
    const double*  le;
    const int32_t* phase;
    const uint8_t* has_previous;
    const int32_t* start_tick;
    cg_decision*   decisions;
    uint32_t*      gate_bits;
    double*        metrics;
    uint8_t*       valid;
    int64_t        n_pairs;
    cg_params      p;
    int32_t  n_slots;
    uint32_t slot_bytes;           // 2 * kRowsPerVer * row bytes
    uint32_t off_le, off_bits, off_bars, off_rec, off_xchg, smem_bytes;
    double   range_s, rc_range, sampled, rc_sampled, thr, half_avg;
};

// ---- 8-sample register tile ---------------------------------------------------------
template <typename T> struct Tile8;
template <> struct Tile8<double> {
    // off67: element offset of the last pair (6, or 4 for the thread whose tile would leave the row)
    static __device__ __forceinline__ void load(const double* p, int off67, double (&x)[8]) {
        const double2* q = reinterpret_cast<const double2*>(p);
        const double2 a = q[0], b = q[1], c = q[2], d = *reinterpret_cast<const double2*>(p + off67);
        x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y; x[4] = c.x; x[5] = c.y; x[6] = d.x; x[7] = d.y;
    }
};
template <> struct Tile8<float> {
    static __device__ __forceinline__ void load(const float* p, int off67, double (&x)[8]) {
        const float2* q = reinterpret_cast<const float2*>(p);
        const float2 a = q[0], b = q[1], c = q[2], d = *reinterpret_cast<const float2*>(p + off67);
        x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y; x[4] = c.x; x[5] = c.y; x[6] = d.x; x[7] = d.y;
    }
};

// extrapolation terms of the thread's ticks (they depend on the tick slot mod 3 only)
template <int NQ>
struct Hot6 {
    double f[kG];        // ext/sampled with the unclamped d_start (Appendix A step 5-6)
    double gchk[kG];     // first_v * gchk >= result  ==>  the zero-point clamp cannot fire
};

// State of the bucket rows, kept per WINDOW (5 per thread), not per tick.  The 6 ticks of a thread
// read only 4-5 distinct 4-sample windows, and a tick's increase is its window's delta times a
// factor f > 0 that does not change from row to row.  x -> round(x * f) is monotone, so
//   * the running maximum of the increases (the monotonic fix, Appendix A step 3) is the rounded
//     product of the running maximum of the window deltas, as long as no row took the zero-point
//     clamp, which only makes an increase smaller;
//   * an increase can reach its rank only if the window delta reaches y0, a slight under-estimate
//     of rank / f.
// A row where every live window is monotone, below y0, and provably unclamped changes nothing but
// wprev: that is the common case, decided by one warp vote.  Any other row is evaluated per tick
// with the exact arithmetic (and leaves the exact per-tick maxima in prevx for the next row).
template <int NQ>
struct WinState {
    double wprev[5];      // running max of the window deltas over the bucket rows so far
    double y0[5][NQ];     // candidate threshold per window; NaN once every tick of the window has its bucket
    double gminw[5];      // smallest gchk among the live ticks of the window (+inf: none)
    double prevx[kP6];    // exact running max per tick after row lastB
    int    lastB;
};

// per-thread shared-memory record, structure of arrays [field][thread]
template <int NQ>
struct Rec6 {
    static constexpr int kObs = 0;                     // [6]      increase of the +Inf bucket
    static constexpr int kQv  = kObs + kP6;            // [6][NQ]  count at the bucket that reached the rank, later the quantile
    static constexpr int kCprev = kQv + kP6 * NQ;      // [6][NQ]  count of the bucket below it
    static constexpr int kEr  = kCprev + kP6 * NQ;     // [6]
    static constexpr int kAvg = kEr + kP6;             // [6]
    static constexpr int kInts = kAvg + kP6;           // then 2 ints (spare, validity bits) and [6][NQ] bytes (bucket index)
    static constexpr int kDoubles = kInts + 1 + (kP6 * NQ + 7) / 8;
    static_assert(kInts >= kP6 * 5, "the scalar role sums borrow the first 30 fields");
};

static __device__ __noinline__ double slow_increase6(const PairArgs& a, int32_t phase, int tick, double result,
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

// histogram_quantile for one (tick, q) at the moment the search reaches the rank (Appendix A
// steps 5-7), with the confirmation in rate space.  Returns NaN and sets *bad when the candidate
// is not confirmed (the tick then goes to the exact path).
static __device__ __noinline__ double finish_quantile(const PairArgs& a, const double* s_le, int b, double cb,
                                                      double cprev, double obs_inc, double qq, int* bad) {
    const double obs = div_const(obs_inc, a.range_s, a.rc_range);
    const double rank = qq * obs;
    const double cbr = div_const(cb, a.range_s, a.rc_range);
    const double cpr = div_const(cprev, a.range_s, a.rc_range);
    if (!(cbr >= rank) || (b > 0 && !(cpr < rank))) { *bad = 1; return NAN; }
    if (b == 0 && s_le[0] <= 0) return s_le[0];
    double start = 0.0, end = s_le[b], cc = cbr, rk = rank;
    if (b > 0) { start = s_le[b - 1]; cc -= cpr; rk -= cpr; }
    return start + (end - start) * ddiv_call(rk, cc);
}

// Window deltas and fast increases of one row for the thread's 6 ticks.  o1 / o2 (uniform within
// a version's warps): window offsets of ticks (1, 2) and (4, 5) inside their period; ticks 0 and 3
// sit at tile offsets 0 and 2.  All 32 lanes call this.
template <typename T, int NQ>
__device__ __forceinline__ void row_increases6(const T* __restrict__ row, const Hot6<NQ>& h, const PairArgs& a,
                                               const int32_t phase, const int tick0, const bool lane_live,
                                               const bool k6ok, const bool k7ok, const int off67, const bool o1,
                                               const bool o2, double (&inc)[kP6], double (&w)[5], unsigned& slow_mask) {
    double x[8];
    Tile8<T>::load(row, off67, x);
#pragma unroll
    for (int j = 0; j < 5; ++j) w[j] = x[j + 3] - x[j];
    bool r[8];
#pragma unroll
    for (int k = 1; k < 8; ++k) r[k] = x[k] < x[k - 1];
    r[6] = r[6] && k6ok;     // samples no live window of this thread reads (they may lie past the row end)
    r[7] = r[7] && k7ok;
    const bool any = ((r[1] || r[2]) || (r[3] || r[4])) || ((r[5] || r[6]) || r[7]);
    // idle lanes (no ticks, or ticks left to the exact path) must not drag the warp into the rare paths
    if (__any_sync(kFull, any && lane_live)) {   // counter resets are rare (Appendix A step 2)
        // value before each reset (0 where there is none), added to each window in time order;
        // adding +0.0 leaves a sum unchanged, so this is the oracle's conditional add
        double d[8];
#pragma unroll
        for (int k = 1; k < 8; ++k) d[k] = r[k] ? x[k - 1] : 0.0;
#pragma unroll
        for (int j = 0; j < 5; ++j) w[j] = ((w[j] + d[j + 1]) + d[j + 2]) + d[j + 3];
    }
    const double res[kP6] = {w[0], o1 ? w[1] : w[0], o2 ? w[2] : w[1], w[2], o1 ? w[3] : w[2], o2 ? w[4] : w[3]};
    const double fst[kP6] = {x[0], o1 ? x[1] : x[0], o2 ? x[2] : x[1], x[2], o1 ? x[3] : x[2], o2 ? x[4] : x[3]};
    // One multiply proves that the zero-point clamp cannot fire (first_v * gchk >= result, gchk a
    // slight under-estimate of sampled/d_start, finite for evaluated ticks and NaN for the others).
    bool sl[kP6];
#pragma unroll
    for (int s = 0; s < kP6; ++s) {
        sl[s] = fst[s] * h.gchk[s % kG] < res[s];
        inc[s] = res[s] * h.f[s % kG];
    }
    const bool slow = ((sl[0] || sl[1]) || (sl[2] || sl[3])) || (sl[4] || sl[5]);
    slow_mask = 0;
    if (__any_sync(kFull, slow)) {   // rare: a counter that restarted one scrape ago
#pragma unroll
        for (int s = 0; s < kP6; ++s)
            if (fst[s] * h.gchk[s % kG] < res[s]) {
                inc[s] = slow_increase6(a, phase, tick0 + s, res[s], fst[s]);
                slow_mask |= 1u << s;
            }
    }
}

template <typename T, int NQ>
__global__ void __launch_bounds__(2 * kHalf, CG_PAIR_MINB) pair_kernel(const __grid_constant__ PairArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    typedef Rec6<NQ> R;
    const cg_params& p = a.p;
    unsigned char* ring = smem;
    double*   s_le   = reinterpret_cast<double*>(smem + a.off_le);
    uint32_t* s_bits = reinterpret_cast<uint32_t*>(smem + a.off_bits);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + a.off_bars);
    uint32_t* s_arr = reinterpret_cast<uint32_t*>(full_bar + a.n_slots);      // per-slot arrival counters
    double*   s_x   = reinterpret_cast<double*>(smem + a.off_xchg);           // [2][3][kHalf*6] gate inputs
    uint8_t*  s_xv  = reinterpret_cast<uint8_t*>(s_x + 2 * 3 * kHalf * kP6);   // [2][kHalf*6] validity

    const int tid = threadIdx.x, lane = tid & 31;
    const int ver = tid >> 6, g2 = tid & (kHalf - 1);
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA;
    const int n_words = (p.n_ticks + 31) >> 5;
    uint32_t* s_cnt = s_bits + n_words;         // pair-completion counter
    if (tid == 0) {
        for (int s = 0; s < a.n_slots; ++s) { mbar_init(full_bar + s, 1); s_arr[s] = 0u; }
        mbar_init_fence();
    }
    for (int b = tid; b < B; b += blockDim.x) s_le[b] = a.le[b];
    for (int w = tid; w < n_words + 1; w += blockDim.x) s_bits[w] = 0u;
    __syncthreads();

    const size_t row_bytes = (size_t)p.row_pitch * sizeof(T);
    const uint32_t half_bytes = (uint32_t)(kRowsPerVer * row_bytes);
    const int n_sc = (K + kRowsPerVer - 1) / kRowsPerVer;
    const int n_use = n_sc + (B > 0 ? 1 + (B - 1 + kRowsPerVer - 1) / kRowsPerVer : 0);

    // ---- ring fills: both versions' rows of the same indices, one mbarrier per slot
    auto cursor_valid = [&](const FillCursor& c) { return c.pair < a.n_pairs; };
    auto skip_pairs = [&](FillCursor& c) {
        while (c.pair < a.n_pairs && a.has_previous && !a.has_previous[c.pair]) c.pair += gridDim.x;
    };
    auto cursor_next = [&](FillCursor& c) {
        if (++c.u == n_use) { c.u = 0; c.pair += gridDim.x; skip_pairs(c); }
    };
    auto issue_fill = [&](const FillCursor& c, uint32_t slot) {      // one lane
        const char* base;
        size_t ver_stride, off;
        uint32_t bytes;
        if (c.u < n_sc) {
            const int k0 = c.u * kRowsPerVer;
            base = reinterpret_cast<const char*>(a.scalars); ver_stride = (size_t)K * row_bytes;
            off = (size_t)k0 * row_bytes; bytes = (uint32_t)(min(kRowsPerVer, K - k0) * row_bytes);
        } else if (c.u == n_sc) {
            base = reinterpret_cast<const char*>(a.buckets); ver_stride = (size_t)B * row_bytes;
            off = (size_t)(B - 1) * row_bytes; bytes = (uint32_t)row_bytes;
        } else {
            const int b0 = (c.u - n_sc - 1) * kRowsPerVer;
            base = reinterpret_cast<const char*>(a.buckets); ver_stride = (size_t)B * row_bytes;
            off = (size_t)b0 * row_bytes; bytes = (uint32_t)(min(kRowsPerVer, B - 1 - b0) * row_bytes);
        }
        const char* src0 = base + (size_t)c.pair * 2 * ver_stride + off;
        unsigned char* dst = ring + (size_t)slot * a.slot_bytes;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // reads of the slot before the async write
        mbar_arrive_expect_tx(full_bar + slot, 2 * bytes);
        bulk_g2s(dst, src0, bytes, full_bar + slot);
        bulk_g2s(dst + half_bytes, src0 + ver_stride, bytes, full_bar + slot);
    };
    FillCursor fill;
    fill.pair = blockIdx.x; fill.ver = 0; fill.u = 0;
    skip_pairs(fill);
    for (int s = 0; s < a.n_slots; ++s) {          // prologue: fill the whole ring
        if (tid == 0 && cursor_valid(fill)) issue_fill(fill, (uint32_t)s);
        cursor_next(fill);
    }

    // ---- ticks of this thread.  A thread owns two consecutive periods (m, m+1) of the tick grid.
    // Full periods are covered pairwise; an odd one out is covered by a thread that overlaps the
    // previous one and keeps only its second period; a partial last period (n_ticks % 3 != 0) gets a
    // thread of its own that overlaps the last full period and silences the tick slots past the end
    // (factor 0), because those would read beyond the row.
    const int n_per = (p.n_ticks + kG - 1) / kG, rem = p.n_ticks % kG;
    const int n_fullp = rem ? n_per - 1 : n_per;
    const int n_thr_full = (n_fullp + 1) / 2;
    int m = -1;                 // first period, -1: idle thread
    unsigned own = 0;           // tick slots whose results this thread delivers
    int quiet_from = kG;        // slots j >= quiet_from (mod 3) are silenced
    if (g2 < n_thr_full) {
        if (2 * g2 + 1 < n_fullp) { m = 2 * g2; own = 0x3fu; }
        else { m = n_fullp - 2; own = 0x38u; }
    } else if (g2 == n_thr_full && rem) {
        m = n_per - 2; own = ((1u << rem) - 1u) << kG; quiet_from = rem;
    }
    const int tick0 = kG * (m < 0 ? 0 : m);
    double* rec = reinterpret_cast<double*>(smem + a.off_rec) + tid;          // field f at rec[f * 128]
    int* rec_i = reinterpret_cast<int*>(reinterpret_cast<double*>(smem + a.off_rec) + (size_t)R::kInts * 2 * kHalf) + tid;
#define CG_REC(f) rec[(f) * (2 * kHalf)]
#define CG_RECI(f) rec_i[(f) * (2 * kHalf)]        // 1: validity, 3 bits per tick
    uint8_t* rec_b = reinterpret_cast<uint8_t*>(reinterpret_cast<double*>(smem + a.off_rec) + (size_t)(R::kInts + 1) * 2 * kHalf) + tid;
#define CG_RECB(f) rec_b[(f) * (2 * kHalf)]        // bucket index where the search for (tick, q) stopped, 255: not yet
    const int pitch = p.row_pitch;
    uint32_t slot = 0, ring_phase = 0;      // position in the ring and the parity of its mbarrier
    const unsigned char* my_half = ring + (size_t)ver * half_bytes;

    auto acquire = [&]() -> const T* {
        CG_T(tw0);
        mbar_wait(full_bar + slot, ring_phase);
        CG_TADD(ver, clock64() - tw0);
        return reinterpret_cast<const T*>(my_half + (size_t)slot * a.slot_bytes);
    };
    auto release = [&]() {
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(s_arr + slot, 1u) == 3u) {   // last of the 4 warps out refills
                s_arr[slot] = 0u;
                if (cursor_valid(fill)) issue_fill(fill, slot);
            }
        }
        cursor_next(fill);
        if (++slot == (uint32_t)a.n_slots) { slot = 0; ring_phase ^= 1u; }
    };

    CG_T(tk0);
    for (int64_t pair = blockIdx.x; pair < a.n_pairs; pair += gridDim.x) {
        CG_T(tp0);
        const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
        if (!hp) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            if (tid == 0 && a.decisions) a.decisions[pair] = replay_schedule([](int) { return false; }, false, st_tick, p);
            if (a.gate_bits) for (int w = tid; w < n_words; w += 2 * kHalf) a.gate_bits[pair * n_words + w] = 0u;
            continue;
        }
        const int64_t pv = pair * 2 + ver;
        const int32_t phase = a.phase[pv];
        // ---- tick geometry (see cg_stream.cuh): lanes 0..2 compute the three slots of a period
        Hot6<NQ> h;
        int i0r = 0, i1r = 0;
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
        }
        const int off1 = i0s[1] - i0s[0], off2 = i0s[2] - i0s[0];
        int pat = (off1 == 0 && off2 == 1) ? 0 : (off1 == 1 && off2 == 1) ? 1 : (off1 == 1 && off2 == 2) ? 2 : -1;
        int base = i0s[0] + 2 * m;
        bool regular = m >= 0 && pat >= 0 && (base & 1) == 0;
#pragma unroll
        for (int s = 0; s < kP6; ++s) {
            const int shift = 2 * (m + s / kG);
            if (s % kG < quiet_from && !(i1s[s % kG] - i0s[s % kG] + 1 == 4 && i0s[s % kG] + shift >= 0 &&
                                         i1s[s % kG] + shift <= p.n_samples - 1))
                regular = false;
        }
        if (pat < 0) pat = 0;
        const bool active = regular;
        unsigned fb_mask = 0;                // ticks that need the exact warp-cooperative path
        if (!active) { base = 0; fb_mask = own; }
#pragma unroll
        for (int s = 0; s < kG; ++s)
            if (!active || s >= quiet_from) { h.f[s] = 0.0; h.gchk[s] = NAN; }
#pragma unroll
        for (int i = 0; i < kP6 * NQ; ++i) CG_RECB(i) = (i % NQ < p.n_q) ? 255 : 0;
        const bool o1 = pat != 0, o2 = pat == 2;
        const unsigned live = active ? own : 0u;           // tick slots the fast path is responsible for
        // window (tile offset 0..4) of each tick slot
        const int wj[kP6] = {0, o1 ? 1 : 0, o2 ? 2 : 1, 2, o1 ? 3 : 2, o2 ? 4 : 3};
        // Samples 6 and 7 of the tile are read by windows 3 and 4 only.  The thread of the partial
        // last period has no live tick there and would read past the row end: it re-reads samples
        // 4 and 5 instead (valid, monotone data) and ignores "resets" among them.
        const bool k6ok = quiet_from == kG || (quiet_from == 2 && o1);
        const bool k7ok = quiet_from == kG && o2;
        const int off67 = k6ok ? 6 : 4;
        CG_T(tp1);
        CG_TADD(4, tp1 - tp0);
        // ---- scalar counter rows: sum(increase(...)) per role, in series order (:375-410)
        {
            // The role sums live in the shared-memory record (fields 0..29, free until the bucket
            // rows start): 30 doubles per thread would not fit the register budget of 3 CTAs per SM.
            unsigned have = 0;
#pragma unroll 1
            for (int u = 0; u < n_sc; ++u) {
                const T* rows = acquire();
                const int first = u * kRowsPerVer, nr = min(kRowsPerVer, K - first);
                const T* rowp = rows + base;
#pragma unroll 1
                for (int k = 0; k < nr; ++k, rowp += pitch) {
                    double inc[kP6], wrow[5];
                    unsigned slow_mask;
                    row_increases6<T, NQ>(rowp, h, a, phase, tick0, active, k6ok, k7ok, off67, o1, o2, inc, wrow, slow_mask);
                    const unsigned role = p.scalar_role[first + k];
#pragma unroll 1
                    for (int q = 0; q < 5; ++q)
                        if (role & (1u << q)) {
                            const bool first_term = !(have & (1u << q));
#pragma unroll
                            for (int s = 0; s < kP6; ++s) CG_REC(s * 5 + q) = first_term ? inc[s] : CG_REC(s * 5 + q) + inc[s];
                            have |= 1u << q;
                        }
                }
                release();
            }
            // derived metrics (:387-404), validity, optional record output
            unsigned vbits = 0;
            double er6[kP6], avg6[kP6];
#pragma unroll
            for (int s = 0; s < kP6; ++s) {
                const double total = (have & 1u) ? CG_REC(s * 5 + 0) : 0.0, err = (have & 2u) ? CG_REC(s * 5 + 1) : 0.0;
                const double lsum = (have & 4u) ? CG_REC(s * 5 + 2) : 0.0, lcnt = (have & 8u) ? CG_REC(s * 5 + 3) : 0.0;
                unsigned valid = (B > 0) ? CG_V_QUANTILE : 0u;
                double er = NAN, avg = NAN;
                if (total > 0) { er = ddiv_call(err, total); valid |= CG_V_ERROR_RATE; }
                if (lcnt > 0) { avg = ddiv_call(lsum, lcnt); valid |= CG_V_LATENCY_AVG; }
                er6[s] = er; avg6[s] = avg;
                vbits |= valid << (3 * s);
                if (a.metrics && ((own >> s) & 1u) && active) {
                    double* o = a.metrics + ((size_t)pv * p.n_ticks + tick0 + s) * W + p.n_q;
                    o[CG_M_ERROR_RESPONSES] = err; o[CG_M_ERROR_RATE] = er; o[CG_M_LATENCY_AVG] = avg;
                    o[CG_M_REQUEST_COUNT] = lcnt; o[CG_M_FEEDBACK_COUNT] = (have & 16u) ? CG_REC(s * 5 + 4) : 0.0;
                }
            }
#pragma unroll
            for (int s = 0; s < kP6; ++s) { CG_REC(R::kEr + s) = er6[s]; CG_REC(R::kAvg + s) = avg6[s]; }   // (the sums are dead now)
            CG_RECI(1) = (int)vbits;
        }
        // ---- the +Inf bucket row (observations -> rank -> candidate thresholds), then bucket rows
        // 0..B-2 in window space (see WinState): the monotonic fix and the search for the rank
        WinState<NQ> ws;
        ws.lastB = -5;
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            ws.wprev[j] = -INFINITY;
            double g = INFINITY;
#pragma unroll
            for (int s = 0; s < kP6; ++s)
                if (((live >> s) & 1u) && wj[s] == j) g = fmin(g, h.gchk[s % kG]);
            ws.gminw[j] = g;
#pragma unroll
            for (int q = 0; q < NQ; ++q) ws.y0[j][q] = NAN;
        }
#pragma unroll
        for (int s = 0; s < kP6; ++s) ws.prevx[s] = -INFINITY;
        // a candidate threshold in increase space; finish_quantile confirms every use of it in rate space
        auto x0_of = [&](double obs_inc, int q) {
            return fmin((p.q[q] * div_const(obs_inc, a.range_s, a.rc_range)) * a.range_s, obs_inc);
        };
#pragma unroll 1
        for (int u = n_sc; u < n_use; ++u) {
            const T* rows = acquire();
            if (u == n_sc) {
                double inc[kP6], wI[5];
                unsigned slow_mask;
                row_increases6<T, NQ>(rows + base, h, a, phase, tick0, active, k6ok, k7ok, off67, o1, o2, inc, wI, slow_mask);
#pragma unroll
                for (int s = 0; s < kP6; ++s) CG_REC(R::kObs + s) = inc[s];
                // y0 = min(q, 1) * (window delta of the +Inf row), shaved by 1e-12: for every tick of
                // the window, delta < y0 implies round(delta * f) < x0 (a few roundings of 2^-53 each)
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    bool alive = false;
#pragma unroll
                    for (int s = 0; s < kP6; ++s) alive = alive || (((live >> s) & 1u) && wj[s] == j);
                    const double wv = wI[j];
#pragma unroll
                    for (int q = 0; q < NQ; ++q) {
                        double y = (fmin(p.q[q], 1.0) * wv) * (1.0 - 1e-12);
                        if (!(wv >= 1e-290)) y = (wv == 0.0) ? 0.0 : -INFINITY;   // tiny / negative / NaN: always a candidate
                        ws.y0[j][q] = (alive && q < p.n_q) ? y : NAN;
                    }
                }
                // ticks whose +Inf increase took the zero-point clamp have a smaller rank: their own threshold
                if (slow_mask & live) {
#pragma unroll 1
                    for (int s = 0; s < kP6; ++s) {
                        if (!(((slow_mask & live) >> s) & 1u)) continue;
                        for (int q = 0; q < p.n_q; ++q) {
                            const double x0 = x0_of(CG_REC(R::kObs + s), q);
                            const int sj = s % kG, wsj = (sj == 0 ? 0 : sj == 1 ? (o1 ? 1 : 0) : (o2 ? 2 : 1)) + 2 * (s / kG);
                            const double fsj = sj == 0 ? h.f[0] : sj == 1 ? h.f[1] : h.f[2];
                            double y = ddiv_call(x0, fsj) * (1.0 - 1e-12);
                            if (!(y >= 1e-290)) y = (x0 == 0.0) ? 0.0 : -INFINITY;
#pragma unroll
                            for (int j = 0; j < 5; ++j)
#pragma unroll
                                for (int qq = 0; qq < NQ; ++qq)
                                    if (j == wsj && qq == q) ws.y0[j][qq] = fmin(ws.y0[j][qq], y);
                        }
                    }
                }
                release();
                continue;
            }
            const int first = (u - n_sc - 1) * kRowsPerVer;
            const int nr = min(kRowsPerVer, B - 1 - first);
            const T* rowp = rows + base;
#pragma unroll 1
            for (int k = 0; k < nr; ++k, rowp += pitch) {
                double x[8], w[5];
                Tile8<T>::load(rowp, off67, x);
#pragma unroll
                for (int j = 0; j < 5; ++j) w[j] = x[j + 3] - x[j];
                bool r[8];
#pragma unroll
                for (int kk = 1; kk < 8; ++kk) r[kk] = x[kk] < x[kk - 1];
                r[6] = r[6] && k6ok;
                r[7] = r[7] && k7ok;
                const bool any = ((r[1] || r[2]) || (r[3] || r[4])) || ((r[5] || r[6]) || r[7]);
                if (__any_sync(kFull, any && live != 0u)) {   // counter resets (Appendix A step 2)
                    double d[8];
#pragma unroll
                    for (int kk = 1; kk < 8; ++kk) d[kk] = r[kk] ? x[kk - 1] : 0.0;
#pragma unroll
                    for (int j = 0; j < 5; ++j) w[j] = ((w[j] + d[j + 1]) + d[j + 2]) + d[j + 3];
                }
                // Is this row uneventful for every live window?  (a) the zero-point clamp cannot
                // fire: first * gminw >= delta, first >= 0; (b) delta >= running max; (c) delta < y0.
                bool cs[5], cm[5], ck[5];
#pragma unroll
                for (int j = 0; j < 5; ++j) {
                    cs[j] = x[j] * ws.gminw[j] < w[j];
                    cm[j] = w[j] < ws.wprev[j];
                    ck[j] = false;
#pragma unroll
                    for (int q = 0; q < NQ; ++q) ck[j] = ck[j] || w[j] >= ws.y0[j][q];
                }
                cm[4] = cm[4] && o2;        // window 4 exists for one phase pattern only
                const bool neg = ((__double2hiint(x[0]) | __double2hiint(x[1])) | (__double2hiint(x[2]) | __double2hiint(x[3])) |
                                  __double2hiint(x[4])) < 0;
                const bool kc = ((ck[0] || ck[1]) || (ck[2] || ck[3])) || ck[4];
                const bool cand = (((cs[0] || cs[1]) || (cs[2] || cs[3])) || (cs[4] || neg)) ||
                                  (((cm[0] || cm[1]) || (cm[2] || cm[3])) || (cm[4] || kc));
#ifdef CG_TIMING
                {
                    const bool lvn = live != 0u;
                    CG_TADD(10, 1);
                    if (__any_sync(kFull, lvn && (((cs[0] || cs[1]) || (cs[2] || cs[3])) || cs[4]))) CG_TADD(11, 1);
                    if (__any_sync(kFull, lvn && neg)) CG_TADD(12, 1);
                    if (__any_sync(kFull, lvn && (((cm[0] || cm[1]) || (cm[2] || cm[3])) || cm[4]))) CG_TADD(13, 1);
                    if (__any_sync(kFull, lvn && kc)) CG_TADD(14, 1);
                    if (__any_sync(kFull, cand && lvn)) CG_TADD(15, 1);
                }
#endif
                if (!__any_sync(kFull, cand && live != 0u)) {
#pragma unroll
                    for (int j = 0; j < 5; ++j) ws.wprev[j] = w[j];
                    continue;
                }
                // ---- eventful row: the exact per-tick arithmetic (same operations as row_increases6)
                const int b = first + k;
                double wn[5];
#pragma unroll
                for (int j = 0; j < 5; ++j) wn[j] = w[j] < ws.wprev[j] ? ws.wprev[j] : w[j];
                const double res[kP6] = {w[0], o1 ? w[1] : w[0], o2 ? w[2] : w[1], w[2], o1 ? w[3] : w[2], o2 ? w[4] : w[3]};
                const double fst[kP6] = {x[0], o1 ? x[1] : x[0], o2 ? x[2] : x[1], x[2], o1 ? x[3] : x[2], o2 ? x[4] : x[3]};
                const double wp[kP6] = {ws.wprev[0], o1 ? ws.wprev[1] : ws.wprev[0], o2 ? ws.wprev[2] : ws.wprev[1],
                                        ws.wprev[2], o1 ? ws.wprev[3] : ws.wprev[2], o2 ? ws.wprev[4] : ws.wprev[3]};
                const bool usex = ws.lastB == b - 1;
                bool fired = false;
#pragma unroll
                for (int s = 0; s < kP6; ++s) {
                    const bool lv = (live >> s) & 1u;
                    double inc = res[s] * h.f[s % kG];
                    if (lv && fst[s] * h.gchk[s % kG] < res[s]) inc = slow_increase6(a, phase, tick0 + s, res[s], fst[s]);
                    // exact running max before this row: kept from the previous row if that one was
                    // eventful too, else the rounded product of the window's running max
                    const double pv_ = usex ? ws.prevx[s] : wp[s] * h.f[s % kG];
                    const double c = inc < pv_ ? pv_ : inc;      // the monotonic fix (Appendix A step 3)
                    if (lv && kc) {
                        const double oi = CG_REC(R::kObs + s);
#pragma unroll
                        for (int q = 0; q < NQ; ++q)
                            if (CG_RECB(s * NQ + q) == 255 && c >= x0_of(oi, q)) {   // first bucket reaching the rank
                                CG_REC(R::kQv + s * NQ + q) = c;
                                CG_REC(R::kCprev + s * NQ + q) = pv_;
                                CG_RECB(s * NQ + q) = (uint8_t)b;
                                fired = true;
                            }
                    }
                    ws.prevx[s] = c;
                }
                ws.lastB = b;
#pragma unroll
                for (int j = 0; j < 5; ++j) ws.wprev[j] = wn[j];
                if (fired) {      // windows all of whose ticks have their bucket stop being candidates
#pragma unroll
                    for (int j = 0; j < 5; ++j)
#pragma unroll
                        for (int q = 0; q < NQ; ++q) {
                            bool alive = false;
#pragma unroll
                            for (int s = 0; s < kP6; ++s)
                                alive = alive || (((live >> s) & 1u) && wj[s] == j && CG_RECB(s * NQ + q) == 255);
                            if (!alive) ws.y0[j][q] = NAN;
                        }
                }
            }
            release();
        }
        // exact running max of each tick after the last bucket row
        double prevf[kP6];
        {
            const double wp[kP6] = {ws.wprev[0], o1 ? ws.wprev[1] : ws.wprev[0], o2 ? ws.wprev[2] : ws.wprev[1],
                                    ws.wprev[2], o1 ? ws.wprev[3] : ws.wprev[2], o2 ? ws.wprev[4] : ws.wprev[3]};
#pragma unroll
            for (int s = 0; s < kP6; ++s) prevf[s] = (ws.lastB == B - 2) ? ws.prevx[s] : wp[s] * h.f[s % kG];
        }
        CG_T(tp2);
        CG_TADD(2 + ver, tp2 - tp1);       // includes the waits
        // ---- finish: quantiles whose search never fired, special cases, the speculation check
        if (B > 0 && active) {
            const bool top_inf = isinf(s_le[B - 1]) && s_le[B - 1] > 0;
            const bool shape_ok = top_inf && B >= 2;
            const double le_top = B >= 2 ? s_le[B - 2] : NAN;
            // observations after the monotonic fix; the search assumed the +Inf bucket itself
            unsigned spec_bad = 0;
#pragma unroll
            for (int s = 0; s < kP6; ++s)
                if (CG_REC(R::kObs + s) < prevf[s]) spec_bad |= 1u << s;
#pragma unroll 1
            for (int s = 0; s < kP6; ++s) {
                if (!((own >> s) & 1u)) continue;
                const double obs_inc = CG_REC(R::kObs + s);
                const double obs = div_const(obs_inc, a.range_s, a.rc_range);
                double top = 0.0;
#pragma unroll
                for (int t = 0; t < kP6; ++t) if (t == s) top = prevf[t];
                top = div_const(top, a.range_s, a.rc_range);     // highest count below the +Inf bucket
                for (int q = 0; q < p.n_q; ++q) {
                    const double qq = p.q[q];
                    const bool q_ok = !isnan(qq) && qq >= 0 && qq <= 1;
                    const bool searched = q_ok && shape_ok && obs != 0;
                    const int bst = CG_RECB(s * NQ + q);
                    int bad = (spec_bad >> s) & 1u;
                    double r;
                    if (bst != 255) {
                        r = finish_quantile(a, s_le, bst, CG_REC(R::kQv + s * NQ + q), CG_REC(R::kCprev + s * NQ + q),
                                            obs_inc, qq, &bad);
                    } else {          // search never fired: then no bucket below the +Inf one may reach the rank
                        r = le_top;
                        if (!(top < qq * obs)) bad = 1;
                    }
                    if (!searched) r = NAN;                       // bad shape, q is NaN, or no observations
                    if (qq < 0) r = -INFINITY;
                    if (qq > 1) r = INFINITY;
                    if (searched && bad) fb_mask |= 1u << s;
                    CG_REC(R::kQv + s * NQ + q) = r;
                    if (a.metrics) a.metrics[((size_t)pv * p.n_ticks + tick0 + s) * W + q] = r;
                }
            }
        }
        if (B == 0 && active)
            for (int s = 0; s < kP6; ++s)
                if ((own >> s) & 1u)
                    for (int q = 0; q < p.n_q; ++q) {
                        CG_REC(R::kQv + s * NQ + q) = NAN;
                        if (a.metrics) a.metrics[((size_t)pv * p.n_ticks + tick0 + s) * W + q] = NAN;
                    }
        if (a.valid && active)
            for (int s = 0; s < kP6; ++s)
                if ((own >> s) & 1u)
                    a.valid[(size_t)pv * p.n_ticks + tick0 + s] = (uint8_t)((CG_RECI(1) >> (3 * s)) & 7);
        CG_T(tp3);
        CG_TADD(5, tp3 - tp2);
        // ---- exact re-evaluation of whatever the fast path could not decide
#pragma unroll 1
        for (int s = 0; s < kP6; ++s) {
            unsigned todo = __ballot_sync(kFull, (fb_mask >> s) & 1u);
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const int tk = __shfl_sync(kFull, tick0 + s, src);
                double mm[CG_MAX_Q + CG_M_EXTRA];
                uint8_t vv;
                coop_eval_tick<T>(a, pv, phase, tk, lane, s_le, mm, vv);
                if (lane == src) {
                    for (int q = 0; q < p.n_q; ++q) CG_REC(R::kQv + s * NQ + q) = mm[q];
                    CG_REC(R::kEr + s) = mm[p.n_q + CG_M_ERROR_RATE];
                    CG_REC(R::kAvg + s) = mm[p.n_q + CG_M_LATENCY_AVG];
                    CG_RECI(1) = (CG_RECI(1) & ~(7 << (3 * s))) | ((int)vv << (3 * s));
                    if (a.metrics) {
                        double* o = a.metrics + ((size_t)pv * p.n_ticks + tk) * W;
                        for (int c = 0; c < W; ++c) o[c] = mm[c];
                    }
                    if (a.valid) a.valid[(size_t)pv * p.n_ticks + tk] = vv;
                }
            }
        }
        CG_T(tp4);
        CG_TADD(6, tp4 - tp3);
        // ---- post this version's gate inputs; the warp that finishes the pair last runs the gate
        // (:419-460) for every tick and replays the step loop.  (Warps of a CTA are never more than
        // the ring depth apart, far less than a pair, so the exchange area and the counter of the
        // previous pair are no longer in use.)
        {
            const int vb = CG_RECI(1);
#pragma unroll 1
            for (int s = 0; s < kP6; ++s) {
                if (!((own >> s) & 1u)) continue;
                const int t = tick0 + s;
                s_x[(ver * 3 + 0) * (kHalf * kP6) + t] = (p.n_q > 0) ? CG_REC(R::kQv + s * NQ + p.gate_q) : NAN;
                s_x[(ver * 3 + 1) * (kHalf * kP6) + t] = CG_REC(R::kEr + s);
                s_x[(ver * 3 + 2) * (kHalf * kP6) + t] = CG_REC(R::kAvg + s);
                s_xv[ver * (kHalf * kP6) + t] = (uint8_t)((vb >> (3 * s)) & 7);
            }
        }
        __syncwarp();
        int last = 0;
        if (lane == 0) {
            __threadfence_block();
            last = (atomicAdd(s_cnt, 1u) == 3u) ? 1 : 0;
        }
        last = __shfl_sync(kFull, last, 0);
        if (last) {
            __threadfence_block();
            if (lane == 0) *s_cnt = 0u;
            for (int w = 0; w < n_words; ++w) {
                const int t = w * 32 + lane;
                bool promote = false;
                if (t < p.n_ticks) {
                    const double nm[3] = {s_x[0 * (kHalf * kP6) + t], s_x[1 * (kHalf * kP6) + t], s_x[2 * (kHalf * kP6) + t]};
                    const double om[3] = {s_x[3 * (kHalf * kP6) + t], s_x[4 * (kHalf * kP6) + t], s_x[5 * (kHalf * kP6) + t]};
                    promote = (gate_eval(nm, s_xv[t], om, s_xv[kHalf * kP6 + t], p) & CG_G_PROMOTE) != 0;
                }
                const uint32_t word = __ballot_sync(kFull, promote);
                if (lane == 0) s_bits[w] = word;
            }
            __syncwarp();
            if (a.gate_bits) for (int w = lane; w < n_words; w += 32) a.gate_bits[pair * n_words + w] = s_bits[w];
            if (lane == 0 && a.decisions) {
                const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
                a.decisions[pair] = replay_schedule_bits(s_bits, n_words, true, st_tick, p);
            }
        }
        CG_TADD(7, clock64() - tp4);
    }
    CG_TADD(8, clock64() - tk0);
    CG_TADD(9, 1);
#undef CG_REC
#undef CG_RECI
#undef CG_RECB
}

template <typename T, int NQ>
cudaError_t launch_pair_tn(const PairArgs& a, int grid, cudaStream_t s) {
    auto kern = pair_kernel<T, NQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)a.smem_bytes);
    if (e != cudaSuccess) return e;
    kern<<<grid, 2 * kHalf, a.smem_bytes, s>>>(a);
    return cudaGetLastError();
}

}  // namespace cg
