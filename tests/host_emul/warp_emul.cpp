// warp_emul.cpp -- HOST emulation of the streaming kernel's per-lane arithmetic (csrc/cg_warp.cuh).
//
// TEST INFRASTRUCTURE ONLY.  The lane functions of cg_warp.cuh are plain C++ as well; this file
// drives them with the same sequence of phases the CUDA kernel uses (the pair's lanes emulated one after
// the other, warp votes replaced by "my own flag", or by "always" to prove that every rare path is
// neutral), so that the arithmetic can be compared with the oracle on a machine without a GPU.
// Ticks the fast path hands to the exact fallback are evaluated here by a plain sequential
// routine (the device uses the warp-cooperative coop_eval_tick).  The product never loads this.
//
// Build: g++ -O1 -ffp-contract=off -shared -fPIC (tests/test_host_emul.py does it).
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

#include "../../research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200/csrc/cg_warp.cuh"

namespace {

using namespace cg;

struct HostAny {
    bool force;
    bool operator()(bool f) const { return f || force; }
};

// the kernel's warp_geometry, serially
void host_geometry(const WarpArgs& a, int32_t phase, int chunk0, SlotGeom* g) {
    const bool phase_ok = phase > -(1 << 30) && phase < (1 << 30);
    int c = 0;
    bool have_even = false;
    if (phase_ok)
        for (int l = 0; l < 3 && !have_even; ++l)
            if (!(slot_terms(a, phase, chunk0 - l).i0 & 1)) { c = l; have_even = true; }
    const int j0 = chunk0 - c;
    SlotTerms t[3];
    for (int s = 0; s < 3; ++s) {
        t[s].i0 = 1; t[s].n4 = 0; t[s].f = 0; t[s].gchk = 1; t[s].dstart = 0; t[s].endt = 0; t[s].rinv = 0;
        if (phase_ok) t[s] = slot_terms(a, phase, j0 + s);
    }
    const int i00 = t[0].i0;
    double gm = INFINITY;
    for (int s = 0; s < 3; ++s) {
        const bool usable = t[s].n4 && phase_ok && !(i00 & 1);
        g->f[s] = t[s].f; g->gchk[s] = t[s].gchk; g->dstart[s] = t[s].dstart; g->endt[s] = t[s].endt;
        g->rinv[s] = t[s].rinv; g->off[s] = t[s].i0 - i00; g->n4[s] = usable ? 1 : 0;
        if (usable) gm = fmin(gm, t[s].gchk);
    }
    g->i00 = i00; g->j0 = j0; g->gmin = gm < INFINITY ? gm : 1.0;
}

// plain sequential semantics of one (pair, version, tick): what coop_eval_tick computes on the device
template <typename T>
void exact_tick(const WarpArgs& a, int64_t pv, int32_t phase, int tick, double* m, uint8_t& valid_out) {
    const cg_params& p = a.p;
    const int nq = p.n_q, K = p.n_scalar, B = p.n_buckets;
    const TickGeom g = tick_geom(p, phase, tick);
    const T* sc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * p.row_pitch;
    const T* bk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * p.row_pitch;
    uint8_t valid = 0;
    double sums[5] = {0, 0, 0, 0, 0};
    unsigned have = 0;
    if (g.n >= 2)
        for (int k = 0; k < K; ++k) {
            const double inc = series_increase(sc + (size_t)k * p.row_pitch, g);
            for (int r = 0; r < 5; ++r)
                if (p.scalar_role[k] & (1u << r)) { sums[r] = (have & (1u << r)) ? sums[r] + inc : inc; have |= 1u << r; }
        }
    for (int j = 0; j < nq; ++j) m[j] = NAN;
    if (B > 0 && g.n >= 2) {
        valid |= CG_V_QUANTILE;
        const double range_s = (double)p.window_ms / 1000.0;
        std::vector<double> cnt0(B), cnt(B);
        for (int b = 0; b < B; ++b) cnt0[b] = series_increase(bk + (size_t)b * p.row_pitch, g) / range_s;
        for (int j = 0; j < nq; ++j) {
            const double q = p.q[j];
            if (q != q) { m[j] = NAN; continue; }
            if (q < 0) { m[j] = -INFINITY; continue; }
            if (q > 1) { m[j] = INFINITY; continue; }
            if (!(std::isinf(a.le[B - 1]) && a.le[B - 1] > 0) || B < 2) { m[j] = NAN; continue; }
            cnt = cnt0;
            double mx = cnt[0];
            for (int i = 1; i < B; ++i) { if (cnt[i] > mx) mx = cnt[i]; else if (cnt[i] < mx) cnt[i] = mx; }
            const double obs = cnt[B - 1];
            if (obs == 0) { m[j] = NAN; continue; }
            double rank = q * obs;
            int b = B - 1;
            for (int i = 0; i < B - 1; ++i) if (cnt[i] >= rank) { b = i; break; }
            if (b == B - 1) { m[j] = a.le[B - 2]; continue; }
            if (b == 0 && a.le[0] <= 0) { m[j] = a.le[0]; continue; }
            double start = 0.0, end = a.le[b], c = cnt[b];
            if (b > 0) { start = a.le[b - 1]; c -= cnt[b - 1]; rank -= cnt[b - 1]; }
            m[j] = start + (end - start) * (rank / c);
        }
    }
    m[nq + CG_M_ERROR_RESPONSES] = sums[1];
    if (sums[0] > 0) { m[nq + CG_M_ERROR_RATE] = sums[1] / sums[0]; valid |= CG_V_ERROR_RATE; } else m[nq + CG_M_ERROR_RATE] = NAN;
    if (sums[3] > 0) { m[nq + CG_M_LATENCY_AVG] = sums[2] / sums[3]; valid |= CG_V_LATENCY_AVG; } else m[nq + CG_M_LATENCY_AVG] = NAN;
    m[nq + CG_M_REQUEST_COUNT] = sums[3];
    m[nq + CG_M_FEEDBACK_COUNT] = sums[4];
    valid_out = valid;
}

struct Counters { long long fb_ticks, event_rows, rows, reset_rows, event_iters, captures, clamps, windows, cfl_rows, hfl_rows; };

template <typename T, int NQ>
void emulate(const WarpArgs& a, bool force_rare, Counters* cnt) {
    const cg_params& p = a.p;
    const int K = p.n_scalar, B = p.n_buckets, W = p.n_q + CG_M_EXTRA, pitch = p.row_pitch;
    const int n_words = (p.n_ticks + 31) >> 5;
    std::vector<double> st((size_t)kFields * kLaneTicks * kLanes), sa((size_t)kLaneWin * kLanes);
    std::vector<uint32_t> s_bits(n_words), s_fb(n_words);
    const QuantConst qc = quant_const(a);
    const HostAny any{force_rare};
    SlotGeom geo[2];
    for (int64_t pair = 0; pair < a.n_pairs; ++pair) {
        const bool hp = a.has_previous ? a.has_previous[pair] != 0 : true;
        if (!hp) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            if (a.decisions) a.decisions[pair] = replay_schedule([](int) { return false; }, false, st_tick, p);
            if (a.gate_bits) for (int w = 0; w < n_words; ++w) a.gate_bits[pair * n_words + w] = 0u;
            continue;
        }
        std::fill(s_bits.begin(), s_bits.end(), 0u);
        std::fill(s_fb.begin(), s_fb.end(), 0u);
        const int32_t phase0 = a.phase[pair * 2], phase1 = a.phase[pair * 2 + 1];
        for (int chunk = 0; chunk < a.n_chunks; ++chunk) {
            const int chunk0 = chunk * kChunkTicks, chunk_end = std::min(p.n_ticks, chunk0 + kChunkTicks);
            host_geometry(a, phase0, chunk0, geo);
            host_geometry(a, phase1, chunk0, geo + 1);
            LaneGeo lg[2][kLanes];
            unsigned fbm[2][kLanes], pass_sc[kLanes], pass_q[kLanes];
            Lane L[kLanes];
            for (int lane = 0; lane < kLanes; ++lane) {
                L[lane].lane = lane; L[lane].st = st.data() + lane; L[lane].sa = sa.data() + lane;
                pass_sc[lane] = pass_q[lane] = 0u;
                for (int ver = 0; ver < 2; ++ver) {
                    unsigned vt;
                    lg[ver][lane] = lane_setup(a, geo[ver], lane, chunk0, chunk_end, vt);
                    fbm[ver][lane] = vt & ~lg[ver][lane].act;
                }
            }
            const int delta = geo[1].j0 - geo[0].j0;
            for (int ver = 0; ver < 2; ++ver) {
                const SlotGeom& g = geo[ver];
                const int64_t pv = pair * 2 + ver;
                const T* sc = reinterpret_cast<const T*>(a.scalars) + (size_t)pv * K * pitch;
                auto row_ptr = [&](int r) -> const T* { return sc + (size_t)r * pitch; };
                for (int lane = 0; lane < kLanes; ++lane)
                    scalar_phase<T>(a, g, lg[ver][lane], L[lane], ver, pv, g.j0 + kLaneTicks * lane, delta, row_ptr, pass_sc[lane]);
            }
            if (B > 0) {
                for (int ver = 0; ver < 2; ++ver) {
                    const SlotGeom& g = geo[ver];
                    const int64_t pv = pair * 2 + ver;
                    const T* bk = reinterpret_cast<const T*>(a.buckets) + (size_t)pv * B * pitch;
                    // row-outer, lane-inner, like the warp: lets the statistics see warp-level votes
                    double prevw[kLanes][kLaneWin], ymin[kLanes][kLaneWin];
                    uint64_t found[kLanes];
                    for (int lane = 0; lane < kLanes; ++lane) {
                        LaneGeo& l = lg[ver][lane];
                        const T* tile = bk + (size_t)(B - 1) * pitch + l.bt;
                        found[lane] = ~0ull;
                        inf_row_setup<T, NQ>(a, g, l, L[lane], qc, tile, ver, delta, pv, g.j0 + kLaneTicks * lane, found[lane],
                                             fbm[ver][lane], pass_q[lane]);
                        for (int k = 0; k < kLaneWin; ++k) { ymin[lane][k] = CG_SY(L[lane], k); prevw[lane][k] = -INFINITY; }
                    }
                    for (int b = 0; b < B - 1; ++b) {
                        int w_reset = 0, w_event = 0, w_iter = 0, w_cfl = 0, w_hfl = 0;
                        for (int lane = 0; lane < kLanes; ++lane) {
                            LaneGeo& l = lg[ver][lane];
                            // the event loop works from the scratch columns: thresholds live there per lane
                            for (int k = 0; k < kLaneWin; ++k) CG_SY(L[lane], k) = ymin[lane][k];
                            g_emul_row = EmulRowStat{0, 0, 0, 0, 0, 0, 0};
                            bucket_row<T, NQ>(a, g, l, L[lane], bk + (size_t)b * pitch + l.bt, b, ver, delta, pv,
                                              g.j0 + kLaneTicks * lane, prevw[lane], ymin[lane], found[lane], fbm[ver][lane],
                                              pass_q[lane], any);
                            w_reset |= g_emul_row.reset; w_event |= g_emul_row.event; w_cfl |= g_emul_row.cfl; w_hfl |= g_emul_row.hfl;
                            w_iter = std::max(w_iter, g_emul_row.windows);
                            if (cnt) { cnt->captures += g_emul_row.captures; cnt->clamps += g_emul_row.clamps; cnt->windows += g_emul_row.windows; }
                        }
                        if (cnt) { ++cnt->rows; cnt->reset_rows += w_reset; cnt->event_rows += w_event; cnt->event_iters += w_iter; cnt->cfl_rows += w_cfl; cnt->hfl_rows += w_hfl; }
                    }
                    for (int lane = 0; lane < kLanes; ++lane) {
                        for (int k = 0; k < kLaneWin; ++k) CG_SY(L[lane], k) = prevw[lane][k];
                        bucket_finish(a, g, lg[ver][lane], L[lane], fbm[ver][lane]);
                    }
                }
            }
            for (int lane = 0; lane < kLanes; ++lane) {
                if (!(p.checks & CG_G_P95_OK)) pass_q[lane] = (1u << kLaneTicks) - 1u;
                const BitField pb = place_bits(pass_sc[lane] & pass_q[lane] & lg[1][lane].act, geo[1].j0 + kLaneTicks * lane);
                if (pb.lo) s_bits[pb.word0] |= pb.lo;
                if (pb.hi) s_bits[pb.word0 + 1] |= pb.hi;
                for (int ver = 0; ver < 2; ++ver) {
                    const BitField f = place_bits(fbm[ver][lane], geo[ver].j0 + kLaneTicks * lane);
                    if (f.lo) s_fb[f.word0] |= f.lo;
                    if (f.hi) s_fb[f.word0 + 1] |= f.hi;
                }
            }
        }
        for (int wd = 0; wd < n_words; ++wd) {
            uint32_t todo = s_fb[wd];
            while (todo) {
                const int bit = __ffs(todo) - 1;
                todo &= todo - 1;
                const int tick = wd * 32 + bit;
                if (cnt) ++cnt->fb_ticks;
                double mn[CG_MAX_Q + CG_M_EXTRA], mo[CG_MAX_Q + CG_M_EXTRA];
                uint8_t vn, vo;
                exact_tick<T>(a, pair * 2, phase0, tick, mn, vn);
                exact_tick<T>(a, pair * 2 + 1, phase1, tick, mo, vo);
                const double nm[3] = {p.n_q > 0 ? mn[p.gate_q] : NAN, mn[p.n_q + CG_M_ERROR_RATE], mn[p.n_q + CG_M_LATENCY_AVG]};
                const double om[3] = {p.n_q > 0 ? mo[p.gate_q] : NAN, mo[p.n_q + CG_M_ERROR_RATE], mo[p.n_q + CG_M_LATENCY_AVG]};
                uint32_t wv = s_bits[wd] & ~(1u << bit);
                if (gate_eval(nm, vn, om, vo, p) & CG_G_PROMOTE) wv |= 1u << bit;
                s_bits[wd] = wv;
                if (a.metrics) {
                    double* on = a.metrics + ((size_t)(pair * 2) * p.n_ticks + tick) * W;
                    double* oo = a.metrics + ((size_t)(pair * 2 + 1) * p.n_ticks + tick) * W;
                    for (int c = 0; c < W; ++c) { on[c] = mn[c]; oo[c] = mo[c]; }
                }
                if (a.valid) {
                    a.valid[(size_t)(pair * 2) * p.n_ticks + tick] = vn;
                    a.valid[(size_t)(pair * 2 + 1) * p.n_ticks + tick] = vo;
                }
            }
        }
        if (a.gate_bits) for (int w = 0; w < n_words; ++w) a.gate_bits[pair * n_words + w] = s_bits[w];
        if (a.decisions) {
            const int st_tick = a.start_tick ? a.start_tick[pair] : p.start_tick;
            a.decisions[pair] = replay_schedule_bits(s_bits.data(), n_words, true, st_tick, p);
        }
    }
}

}  // namespace

// Same arguments as cg_eval_fused, HOST pointers.  force_rare != 0: every warp vote answers "yes".
// counters (may be NULL, 10 entries): [0] ticks handed to the exact fallback, [1] warp-rows that entered the event
// loop, [2] bucket warp-rows, [3] warp-rows with a reset correction, [4] event-loop iterations (max over lanes, summed
// over rows), [5] quantile captures, [6] clamped windows, [7] windows looked at,
// [8] warp-rows where the clamp pre-test fired, [9] warp-rows where a rank threshold fired.
extern "C" int emul_eval_fused(const void* scalars, const void* buckets, const double* le, const int32_t* phase,
                               const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                               const cg_params* p, cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                               uint8_t* valid, int force_rare, long long* counters) {
    if (!p || p->struct_bytes != (int32_t)sizeof(cg_params)) return CG_EINVAL;
    if ((int64_t)p->window_ms != 4 * (int64_t)p->period_ms || 3 * (int64_t)p->tick_step_ms != 2 * (int64_t)p->period_ms ||
        p->n_samples < cg::kTile || p->n_scalar < 1 || p->n_buckets > 64)
        return CG_EINVAL;
    cg::WarpArgs a;
    memset(&a, 0, sizeof(a));
    a.scalars = scalars; a.buckets = buckets; a.le = le; a.phase = phase; a.has_previous = has_previous;
    a.start_tick = start_tick; a.decisions = decisions; a.gate_bits = gate_bits; a.metrics = metrics; a.valid = valid;
    a.n_pairs = n_pairs;
    a.p = *p;
    cg::fill_warp_constants(a);
    Counters c = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const bool f = force_rare != 0;
    const int nq = p->n_q < 1 ? 1 : p->n_q;
    if (p->dtype == CG_F32) {
        switch (nq) {
            case 1: emulate<float, 1>(a, f, &c); break;
            case 2: emulate<float, 2>(a, f, &c); break;
            case 3: emulate<float, 3>(a, f, &c); break;
            default: emulate<float, 4>(a, f, &c); break;
        }
    } else {
        switch (nq) {
            case 1: emulate<double, 1>(a, f, &c); break;
            case 2: emulate<double, 2>(a, f, &c); break;
            case 3: emulate<double, 3>(a, f, &c); break;
            default: emulate<double, 4>(a, f, &c); break;
        }
    }
    if (counters) { counters[0] = c.fb_ticks; counters[1] = c.event_rows; counters[2] = c.rows; counters[3] = c.reset_rows;
                    counters[4] = c.event_iters; counters[5] = c.captures; counters[6] = c.clamps; counters[7] = c.windows;
                    counters[8] = c.cfl_rows; counters[9] = c.hfl_rows; }
    return CG_OK;
}
