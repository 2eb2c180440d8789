// cg_api.cu -- the C ABI of libcanarygate.so (see include/canary_gate.h) and the small
// standalone kernels (gate, replay).  Build: nvcc -gencode arch=compute_100a,code=sm_100a
// -fmad=false (no FMA contraction: the reference computes in Python floats).
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include <atomic>

#include "../../include/canary_gate.h"
#include "cg_general.cuh"
#include "cg_rows.cuh"
#include "cg_warp.cuh"
#include "cg_irregular.cuh"
#include "cg_irows.cuh"
#include "cg_synth.cuh"
#include "cg_store.cuh"

namespace {

thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

int cuda_fail(cudaError_t e, const char* what) {
    return fail(CG_ECUDA, "%s: %s", what, cudaGetErrorString(e));
}

constexpr int kMaxSmem = 232448;   // 227 KB opt-in limit per CTA on sm_100

size_t elem_size(const cg_params* p) { return p->dtype == CG_F32 ? 4 : 8; }

int check_params(const cg_params* p) {
    if (!p) return fail(CG_EINVAL, "params is NULL");
    if (p->struct_bytes != (int32_t)sizeof(cg_params))
        return fail(CG_EINVAL, "cg_params.struct_bytes = %d, library expects %d (ABI mismatch)",
                    p->struct_bytes, (int)sizeof(cg_params));
    if (p->dtype != CG_F64 && p->dtype != CG_F32) return fail(CG_EINVAL, "dtype must be CG_F64 or CG_F32");
    if (p->n_samples < 1) return fail(CG_EINVAL, "n_samples must be >= 1");
    if (p->row_pitch < p->n_samples) return fail(CG_EINVAL, "row_pitch < n_samples");
    if (((size_t)p->row_pitch * elem_size(p)) % 16 != 0)
        return fail(CG_EINVAL, "row pitch (%d elements) must be a multiple of 16 bytes", p->row_pitch);
    if (p->n_scalar < 0 || p->n_scalar > CG_MAX_SCALAR) return fail(CG_EINVAL, "n_scalar out of range [0,%d]", CG_MAX_SCALAR);
    if (p->n_buckets < 0 || p->n_buckets > 1024) return fail(CG_EINVAL, "n_buckets out of range [0,1024]");
    if (p->period_ms <= 0 || p->window_ms <= 0 || p->tick_step_ms <= 0)
        return fail(CG_EINVAL, "period_ms, window_ms and tick_step_ms must be positive");
    if (p->n_ticks < 0) return fail(CG_EINVAL, "n_ticks must be >= 0");
    if (p->n_q < 0 || p->n_q > CG_MAX_Q) return fail(CG_EINVAL, "n_q out of range [0,%d]", CG_MAX_Q);
    if (p->n_buckets > 0 && p->n_q < 1) return fail(CG_EINVAL, "n_q must be >= 1 when n_buckets > 0");
    if (p->n_buckets == 0 && p->n_q != 0) return fail(CG_EINVAL, "n_q must be 0 when n_buckets == 0");
    if (p->n_q > 0 && (p->gate_q < 0 || p->gate_q >= p->n_q)) return fail(CG_EINVAL, "gate_q out of range");
    if (p->checks & ~7) return fail(CG_EINVAL, "checks has unknown bits");
    if (p->max_attempts < 1 || p->step < 1 || p->max_traffic < 1 || p->interval_ticks < 0 ||
        p->attempt_delay_ticks < 0 || p->init_current < 0)
        return fail(CG_EINVAL, "bad schedule constants");
    if (p->reserved != 0) return fail(CG_EINVAL, "cg_params.reserved must be 0 (got %d): zero the struct or start from cg_params_default", p->reserved);
    if (p->on_abort != CG_ABORT_FREEZE && p->on_abort != CG_ABORT_ROLLBACK) return fail(CG_EINVAL, "on_abort must be CG_ABORT_FREEZE or CG_ABORT_ROLLBACK");
    if (p->attempt_delay_ticks == 0 && p->interval_ticks == 0)
        return fail(CG_EINVAL, "schedule never advances (both delays are 0 ticks)");
    return CG_OK;
}

int check_aligned(const void* ptr, const char* name) {
    if (((uintptr_t)ptr) & 15) return fail(CG_EINVAL, "%s must be 16-byte aligned", name);
    return CG_OK;
}

struct GeneralPlan {
    cg::GeneralArgs args;
    int grid, block;
};

int plan_general(const cg_params* p, int64_t n_pairs, GeneralPlan* plan) {
    cg::GeneralArgs& a = plan->args;
    memset(&a, 0, sizeof(a));
    a.p = *p;
    a.n_pairs = n_pairs;
    const size_t esz = elem_size(p);
    size_t off = (size_t)(p->n_scalar + p->n_buckets) * p->row_pitch * esz;
    off = (off + 15) & ~(size_t)15;
    a.off_le = (uint32_t)off;           off += (size_t)p->n_buckets * 8;
    a.off_stash = (uint32_t)off;        off += (size_t)p->n_ticks * 3 * 8;
    a.off_stash_valid = (uint32_t)off;  off += ((size_t)p->n_ticks + 7) & ~(size_t)7;
    a.off_bits = (uint32_t)off;         off += (((size_t)p->n_ticks + 31) / 32) * 4;
    off = (off + 7) & ~(size_t)7;
    a.off_bar = (uint32_t)off;          off += 8;
    if (off > (size_t)kMaxSmem)
        return fail(CG_EINVAL, "geometry needs %zu bytes of shared memory per CTA (limit %d): "
                    "(n_scalar+n_buckets)*row_pitch too large for the resident-tile kernel", off, kMaxSmem);
    if ((size_t)(p->n_scalar + p->n_buckets) * p->row_pitch * esz >= (1u << 20))
        return fail(CG_EINVAL, "tile exceeds the mbarrier transaction-count range");
    a.smem_bytes = (uint32_t)off;
    int threads = ((p->n_ticks + 31) / 32) * 32;
    if (threads < 32) threads = 32;
    if (threads > 512) threads = 512;
    plan->block = threads;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int per_sm = (int)(kMaxSmem / (off + 1024)) > 0 ? (int)(kMaxSmem / (off + 1024)) : 1;
    int64_t g = (int64_t)sms * per_sm;
    if (g > n_pairs) g = n_pairs;
    if (g < 1) g = 1;
    plan->grid = (int)g;
    return CG_OK;
}

template <typename T>
int launch_general(const GeneralPlan& plan, cudaStream_t s) {
    auto kern = cg::general_kernel<T>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)plan.args.smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(general_kernel)");
    kern<<<plan.grid, plan.block, plan.args.smem_bytes, s>>>(plan.args);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch general_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

// ---- row-streaming general kernel (cg_rows.cuh): eligibility, shared-memory plan, launch ------
struct RowsPlan {
    cg::RowsArgs args;
    int grid, block;
};

int plan_rows(const cg_params* p, int64_t n_pairs, RowsPlan* plan, bool quiet) {
    cg::RowsArgs& a = plan->args;
    memset(&a, 0, sizeof(a));
    a.p = *p;
    a.n_pairs = n_pairs;
    const size_t esz = elem_size(p), row_bytes = (size_t)p->row_pitch * esz;
    const int rows = p->n_scalar + p->n_buckets;
    if (row_bytes % 16 || row_bytes == 0 || rows < 1)
        return quiet ? CG_EINVAL : fail(CG_EINVAL, "row-streaming kernel: a row must be a non-zero multiple of 16 bytes");
    int g = (int)(16384 / row_bytes);           // about 16 KB per ring slot: a CTA barrier every 8 rows of an hour of scrapes
    g = g < 1 ? 1 : (g > 16 ? 16 : g);
    if (g > rows) g = rows;
    const int slots = 3;
    a.rows_per_slot = g;
    a.n_slots = slots;
    a.row_bytes = (uint32_t)row_bytes;
    a.slot_bytes = (uint32_t)(g * row_bytes);
    a.uses_per_ver = (rows + g - 1) / g;
    int threads = ((p->n_ticks + 31) / 32) * 32;
    threads = threads < 32 ? 32 : (threads > 384 ? 384 : threads);
    plan->block = threads;
    a.n_passes = (p->n_ticks + threads - 1) / threads;
    size_t off = (size_t)slots * a.slot_bytes;
    a.off_bars = (uint32_t)off;         off += (size_t)slots * 8;
    off = (off + 15) & ~(size_t)15;
    a.off_le = (uint32_t)off;           off += (size_t)p->n_buckets * 8;
    a.off_stash = (uint32_t)off;        off += (size_t)p->n_ticks * 3 * 8;
    a.off_stash_valid = (uint32_t)off;  off += ((size_t)p->n_ticks + 7) & ~(size_t)7;
    a.off_bits = (uint32_t)off;         off += (((size_t)p->n_ticks + 31) / 32) * 4;
    off = (off + 15) & ~(size_t)15;
    if (off > (size_t)kMaxSmem || a.slot_bytes >= (1u << 20))
        return quiet ? CG_EINVAL : fail(CG_EINVAL, "row-streaming kernel: ring and per-tick state need %zu bytes of shared memory (limit %d)", off, kMaxSmem);
    a.smem_bytes = (uint32_t)off;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int per_sm = (int)((kMaxSmem + 1024) / (off + 1024));
    per_sm = per_sm < 1 ? 1 : per_sm;
    const int by_threads = 768 / threads;       // two CTAs of 384 threads by registers (__launch_bounds__(384, 2))
    if (per_sm > by_threads) per_sm = by_threads < 1 ? 1 : by_threads;
    int64_t gr = (int64_t)sms * per_sm;
    if (gr > n_pairs) gr = n_pairs;
    if (gr < 1) gr = 1;
    plan->grid = (int)gr;
    return CG_OK;
}

template <typename T, int MQ>
int launch_rows(const RowsPlan& plan, cudaStream_t s) {
    auto kern = cg::rows_kernel<T, MQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.args.smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(rows_kernel)");
    kern<<<plan.grid, plan.block, plan.args.smem_bytes, s>>>(plan.args);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch rows_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

// ---- streaming kernel (cg_warp.cuh): eligibility, shared-memory plan, launch -----------
struct StreamPlan {
    cg::WarpArgs args;
    int grid;
};

constexpr size_t kWarpState = (size_t)cg::kFields * cg::kLaneTicks * cg::kLanes * 8;   // per-tick state arrays
constexpr size_t kWarpScratch = (size_t)cg::kLaneWin * cg::kLanes * 8;                 // window thresholds

size_t warp_fixed_smem(const cg_params* p) {
    const size_t n_words = ((size_t)p->n_ticks + 31) / 32;
    const size_t uses = 2 * (((size_t)p->n_scalar + cg::kRowsPerUse - 1) / cg::kRowsPerUse) +
                        (p->n_buckets > 0 ? 2 * (1 + ((size_t)p->n_buckets - 1 + cg::kRowsPerUse - 1) / cg::kRowsPerUse) : 0);
    return kWarpState + kWarpScratch + 2 * cg::kWarps * sizeof(cg::SlotGeom) + 2 * n_words * 4 + uses * sizeof(cg::FillEntry) +
           8 * 16 + 64;
}

bool stream_eligible(const cg_params* p) {
    if ((int64_t)p->window_ms != 4 * (int64_t)p->period_ms) return false;
    if (3 * (int64_t)p->tick_step_ms != 2 * (int64_t)p->period_ms) return false;
    if (p->n_ticks < 1 || p->n_ticks > 16384) return false;
    if (p->n_buckets > 64 || p->n_scalar < 1) return false;   // exact fallback holds two buckets per lane; ratios come from the scalar rows
    if (p->n_samples < cg::kTile) return false;               // a lane's 12-sample tile must fit the series
    const size_t slot = (size_t)cg::kRowsPerUse * p->row_pitch * elem_size(p);
    if (slot >= (1u << 20)) return false;
    // 32-bit time arithmetic in the kernel
    if ((int64_t)p->tick0_ms + (int64_t)p->n_ticks * p->tick_step_ms >= (1ll << 30)) return false;
    if ((int64_t)p->n_samples * p->period_ms >= (1ll << 30) || p->tick0_ms < -(1 << 30)) return false;
    // all scalar rows of a version are resident at once; one more slot keeps the stream going
    const size_t need = ((size_t)p->n_scalar + cg::kRowsPerUse - 1) / cg::kRowsPerUse + 1;
    return (need < 2 ? 2 : need) * slot + warp_fixed_smem(p) <= (size_t)kMaxSmem;
}

int plan_stream(const cg_params* p, int64_t n_pairs, StreamPlan* plan) {
    cg::WarpArgs& a = plan->args;
    memset(&a, 0, sizeof(a));
    a.p = *p;
    a.n_pairs = n_pairs;
    a.row_bytes = (uint32_t)((size_t)p->row_pitch * elem_size(p));
    a.slot_bytes = (uint32_t)cg::kRowsPerUse * a.row_bytes;
    cg::fill_warp_constants(a);
    const size_t n_words = ((size_t)p->n_ticks + 31) / 32;
    const size_t fixed = warp_fixed_smem(p);
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // a CTA is the kWarps warps of one pair: as many CTAs per SM as still leave each a ring of `want` slots (fewer if forced)
    int per_sm = 0, slots = 0;
    const int floor_slots = a.n_sc_uses + 1 > 2 ? a.n_sc_uses + 1 : 2;     // scalar rows of a version are resident together
    int c_max = cg::kG == 1 ? 4 : 8, want = floor_slots > 3 ? floor_slots : 3;
#ifdef CG_ABLATE
    if (const char* e = getenv("CG_WARP_PER_SM")) { int v = atoi(e); if (v >= 1 && v <= 16) c_max = v; }
    if (const char* e = getenv("CG_WARP_SLOTS")) { int v = atoi(e); if (v >= 2 && v <= 16) want = v; }
#endif
    for (int need = want; need >= floor_slots && !per_sm; --need)
        for (int c = c_max; c >= 1; --c) {
            const size_t budget = (size_t)(kMaxSmem + 1024) / c - 1024;
            if (budget <= fixed) continue;
            const int sl = (int)((budget - fixed) / a.slot_bytes);
            if (sl >= need) { per_sm = c; slots = sl; break; }
        }
    if (!per_sm) return fail(CG_EINVAL, "streaming kernel: ring does not fit shared memory");
    if (slots > 8 && slots > floor_slots) slots = floor_slots > 8 ? floor_slots : 8;
#ifdef CG_ABLATE
    if (const char* e = getenv("CG_WARP_SLOTS")) { int v = atoi(e); if (v >= floor_slots && v < slots) slots = v; }
#endif
    a.n_slots = slots;
    size_t off = (size_t)slots * a.slot_bytes;          // the ring sits at offset 0
    a.off_bars = (uint32_t)off;    off += (size_t)slots * 8;
    off = (off + 15) & ~(size_t)15;
    a.off_state = (uint32_t)off;   off += kWarpState;
    a.off_scratch = (uint32_t)off; off += kWarpScratch;
    a.off_geom = (uint32_t)off;    off += 2 * cg::kWarps * sizeof(cg::SlotGeom);
    a.off_bits = (uint32_t)off;    off += 2 * n_words * 4;
    off = (off + 7) & ~(size_t)7;
    a.off_tab = (uint32_t)off;     off += (size_t)a.uses_per_pass * sizeof(cg::FillEntry);
    off = (off + 15) & ~(size_t)15;
    a.smem_bytes = (uint32_t)off;
    int64_t g = (int64_t)sms * per_sm;
    if (g > n_pairs) g = n_pairs;
    if (g < 1) g = 1;
    plan->grid = (int)g;
    return CG_OK;
}

}  // namespace
// one per translation unit cg_warp_inst.cu (see build.py)
#define CG_DECL_INST(n) cudaError_t n(const cg::WarpArgs&, int, cudaStream_t)
CG_DECL_INST(cg_launch_warp_d1); CG_DECL_INST(cg_launch_warp_d2); CG_DECL_INST(cg_launch_warp_d3);
CG_DECL_INST(cg_launch_warp_d4); CG_DECL_INST(cg_launch_warp_f1); CG_DECL_INST(cg_launch_warp_f2);
CG_DECL_INST(cg_launch_warp_f3); CG_DECL_INST(cg_launch_warp_f4);
namespace {

int launch_stream(const StreamPlan& plan, cudaStream_t s) {
    typedef cudaError_t (*fn_t)(const cg::WarpArgs&, int, cudaStream_t);
    static const fn_t table[2][4] = {
        {cg_launch_warp_d1, cg_launch_warp_d2, cg_launch_warp_d3, cg_launch_warp_d4},
        {cg_launch_warp_f1, cg_launch_warp_f2, cg_launch_warp_f3, cg_launch_warp_f4}};
    int nq = plan.args.p.n_q < 1 ? 1 : plan.args.p.n_q;
    const cudaError_t e = table[plan.args.p.dtype == CG_F32 ? 1 : 0][nq - 1](plan.args, plan.grid, s);
    if (e != cudaSuccess) return cuda_fail(e, "launch warp_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

// ---- standalone gate / replay kernels ------------------------------------------
__global__ void gate_kernel(const double* __restrict__ new_m, const uint8_t* __restrict__ new_v,
                            const double* __restrict__ old_m, const uint8_t* __restrict__ old_v,
                            int64_t n, const __grid_constant__ cg_params p, uint8_t* __restrict__ detail) {
    const int W = p.n_q + CG_M_EXTRA;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double* a = new_m + i * W;
        const double* b = old_m + i * W;
        const double nm[3] = {p.n_q > 0 ? a[p.gate_q] : NAN, a[p.n_q + CG_M_ERROR_RATE], a[p.n_q + CG_M_LATENCY_AVG]};
        const double om[3] = {p.n_q > 0 ? b[p.gate_q] : NAN, b[p.n_q + CG_M_ERROR_RATE], b[p.n_q + CG_M_LATENCY_AVG]};
        detail[i] = cg::gate_eval(nm, new_v[i], om, old_v[i], p);
    }
}

__global__ void replay_kernel(const uint32_t* __restrict__ gate_bits, const uint8_t* __restrict__ has_previous,
                              const int32_t* __restrict__ start_tick, int64_t n_pairs,
                              const __grid_constant__ cg_params p, cg_decision* __restrict__ decisions) {
    const int n_words = (p.n_ticks + 31) >> 5;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n_pairs; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t* bits = gate_bits + i * n_words;
        const bool hp = has_previous ? has_previous[i] != 0 : true;
        const int st = start_tick ? start_tick[i] : p.start_tick;
        decisions[i] = cg::replay_schedule_bits(bits, n_words, hp, st, p);
    }
}

int grid_for(int64_t n, int block) {
    int64_t g = (n + block - 1) / block;
    if (g > 148 * 16) g = 148 * 16;
    if (g < 1) g = 1;
    return (int)g;
}

int run_general(const void* scalars, const void* buckets, const double* le, const int32_t* phase,
                const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                const cg_params* p, cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                uint8_t* valid, void* stream, int path = 0) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (n_pairs == 0) return CG_OK;
    if (p->n_scalar > 0 && !scalars) return fail(CG_EINVAL, "scalars is NULL but n_scalar > 0");
    if (p->n_buckets > 0 && (!buckets || !le)) return fail(CG_EINVAL, "buckets/le is NULL but n_buckets > 0");
    if (!phase) return fail(CG_EINVAL, "phase is NULL");
    if ((rc = check_aligned(scalars, "scalars"))) return rc;
    if ((rc = check_aligned(buckets, "buckets"))) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // per-tick metric output alone (cg_eval_grid) also streams once the grid is worth a CTA
    if ((decisions || p->n_ticks >= 32) && path != 1 && path != 3 && stream_eligible(p)) {
        StreamPlan sp;
        if ((rc = plan_stream(p, n_pairs, &sp))) return rc;
        cg::WarpArgs& a = sp.args;
        a.scalars = scalars; a.buckets = buckets; a.le = le; a.phase = phase;
        a.has_previous = has_previous; a.start_tick = start_tick; a.decisions = decisions;
        a.gate_bits = gate_bits; a.metrics = metrics; a.valid = valid;
        return launch_stream(sp, s);
    }
    if (path == 2) return fail(CG_EINVAL, "geometry not covered by the streaming kernel");
    // any other geometry: the rows pass through a ring (cg_rows.cuh) once the grid is worth a CTA;
    // the resident-tile kernel keeps tiny grids and is the explicit path 1
    if (path == 3 || (path == 0 && (decisions || p->n_ticks >= 32))) {
        RowsPlan rp;
        rc = plan_rows(p, n_pairs, &rp, path != 3);
        if (rc == CG_OK) {
            cg::RowsArgs& a = rp.args;
            a.scalars = scalars; a.buckets = buckets; a.le = le; a.phase = phase;
            a.has_previous = has_previous; a.start_tick = start_tick; a.decisions = decisions;
            a.gate_bits = gate_bits; a.metrics = metrics; a.valid = valid;
            if (p->n_q <= 1) return p->dtype == CG_F32 ? launch_rows<float, 1>(rp, s) : launch_rows<double, 1>(rp, s);
            return p->dtype == CG_F32 ? launch_rows<float, CG_MAX_Q>(rp, s) : launch_rows<double, CG_MAX_Q>(rp, s);
        }
        if (path == 3) return rc;
    }
    GeneralPlan plan;
    if ((rc = plan_general(p, n_pairs, &plan))) return rc;
    cg::GeneralArgs& a = plan.args;
    a.scalars = scalars; a.buckets = buckets; a.le = le; a.phase = phase;
    a.has_previous = has_previous; a.start_tick = start_tick; a.decisions = decisions;
    a.gate_bits = gate_bits; a.metrics = metrics; a.valid = valid;
    return p->dtype == CG_F32 ? launch_general<float>(plan, s) : launch_general<double>(plan, s);
}


// ---- irregular series: shared-memory plan and launch -----------------------------------
struct IrregularPlan {
    cg::IrregularArgs args;
    int grid, block;
};

int plan_irregular(const cg_params* p, int64_t n_pairs, IrregularPlan* plan) {
    cg::IrregularArgs& a = plan->args;
    memset(&a, 0, sizeof(a));
    a.p = *p;
    a.n_pairs = n_pairs;
    const size_t esz = elem_size(p);
    const size_t rows = (size_t)p->n_scalar + p->n_buckets;
    if (p->row_pitch % 4) return fail(CG_EINVAL, "irregular series: row_pitch must be a multiple of 4 samples");
    size_t off = rows * p->row_pitch * esz;
    off = (off + 15) & ~(size_t)15;
    a.off_ts = (uint32_t)off;           off += rows * p->row_pitch * 4;
    a.off_cnt = (uint32_t)off;          off += rows * 12;     // counts, first time stamps, search hints
    off = (off + 7) & ~(size_t)7;
    a.off_le = (uint32_t)off;           off += (size_t)p->n_buckets * 8;
    a.off_stash = (uint32_t)off;        off += (size_t)p->n_ticks * 3 * 8;
    a.off_stash_valid = (uint32_t)off;  off += ((size_t)p->n_ticks + 7) & ~(size_t)7;
    a.off_bits = (uint32_t)off;         off += (((size_t)p->n_ticks + 31) / 32) * 4;
    off = (off + 7) & ~(size_t)7;
    a.off_bar = (uint32_t)off;          off += 8;
    if (off > (size_t)kMaxSmem)
        return fail(CG_EINVAL, "irregular series: values + time stamps of one (pair, version) need %zu bytes of shared "
                    "memory (limit %d): (n_scalar+n_buckets)*row_pitch too large", off, kMaxSmem);
    if (rows * p->row_pitch * (esz + 4) >= (1u << 20))
        return fail(CG_EINVAL, "tile exceeds the mbarrier transaction-count range");
    a.smem_bytes = (uint32_t)off;
    int threads = ((p->n_ticks + 31) / 32) * 32;
    if (threads < 32) threads = 32;
    if (threads > 512) threads = 512;
    plan->block = threads;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int per_sm = (int)(kMaxSmem / (off + 1024)) > 0 ? (int)(kMaxSmem / (off + 1024)) : 1;
    int64_t g = (int64_t)sms * per_sm;
    if (g > n_pairs) g = n_pairs;
    if (g < 1) g = 1;
    plan->grid = (int)g;
    return CG_OK;
}

struct IRowsPlan {
    cg::IRowsArgs args;
    int grid, block;
};

// row-streaming kernel for irregular series (cg_irows.cuh); quiet: no error text when it does not apply
int plan_irows(const cg_params* p, int64_t n_pairs, IRowsPlan* plan, bool quiet) {
    cg::IRowsArgs& a = plan->args;
    memset(&a, 0, sizeof(a));
    a.p = *p;
    a.n_pairs = n_pairs;
    const size_t esz = elem_size(p), vrow = (size_t)p->row_pitch * esz, trow = (size_t)p->row_pitch * 4;
    const int rows = p->n_scalar + p->n_buckets;
    if (p->row_pitch % 4 || vrow == 0 || rows < 1)
        return quiet ? CG_EINVAL : fail(CG_EINVAL, "irregular series: row_pitch must be a non-zero multiple of 4 samples");
    int g = (int)(16384 / (vrow + trow));       // about 16 KB per ring slot
    g = g < 1 ? 1 : (g > 16 ? 16 : g);
    if (g > rows) g = rows;
    const int slots = 3;
    a.rows_per_slot = g;
    a.n_slots = slots;
    a.vrow_bytes = (uint32_t)vrow;
    a.trow_bytes = (uint32_t)trow;
    a.slot_bytes = (uint32_t)(g * (vrow + trow));
    a.uses_per_ver = (rows + g - 1) / g;
    int threads = ((p->n_ticks + 31) / 32) * 32;
    threads = threads < 32 ? 32 : (threads > 384 ? 384 : threads);
    plan->block = threads;
    a.n_passes = (p->n_ticks + threads - 1) / threads;
    size_t off = (size_t)slots * a.slot_bytes;
    a.off_bars = (uint32_t)off;         off += (size_t)slots * 8;
    off = (off + 15) & ~(size_t)15;
    a.off_cnt = (uint32_t)off;          off += (size_t)rows * 12;       // counts, first time stamps, search hints
    off = (off + 7) & ~(size_t)7;
    a.off_le = (uint32_t)off;           off += (size_t)p->n_buckets * 8;
    a.off_stash = (uint32_t)off;        off += (size_t)p->n_ticks * 3 * 8;
    a.off_stash_valid = (uint32_t)off;  off += ((size_t)p->n_ticks + 7) & ~(size_t)7;
    a.off_bits = (uint32_t)off;         off += (((size_t)p->n_ticks + 31) / 32) * 4;
    off = (off + 15) & ~(size_t)15;
    if (off > (size_t)kMaxSmem || a.slot_bytes >= (1u << 20))
        return quiet ? CG_EINVAL : fail(CG_EINVAL, "irregular series: ring and per-tick state need %zu bytes of shared memory (limit %d)", off, kMaxSmem);
    a.smem_bytes = (uint32_t)off;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int per_sm = (int)((kMaxSmem + 1024) / (off + 1024));
    per_sm = per_sm < 1 ? 1 : per_sm;
    const int by_threads = 768 / threads;       // two CTAs of 384 threads by registers (__launch_bounds__(384, 2))
    if (per_sm > by_threads) per_sm = by_threads < 1 ? 1 : by_threads;
    int64_t gr = (int64_t)sms * per_sm;
    if (gr > n_pairs) gr = n_pairs;
    if (gr < 1) gr = 1;
    plan->grid = (int)gr;
    return CG_OK;
}

template <typename T, int MQ>
int launch_irows(const IRowsPlan& plan, cudaStream_t s) {
    auto kern = cg::irows_kernel<T, MQ>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.args.smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(irows_kernel)");
    kern<<<plan.grid, plan.block, plan.args.smem_bytes, s>>>(plan.args);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch irows_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

template <typename T>
int launch_irregular(const IrregularPlan& plan, cudaStream_t s) {
    auto kern = cg::irregular_kernel<T>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan.args.smem_bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(irregular_kernel)");
    kern<<<plan.grid, plan.block, plan.args.smem_bytes, s>>>(plan.args);
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch irregular_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

}  // namespace

extern "C" {

int cg_abi_version(void) { return CG_ABI_VERSION; }
const char* cg_last_error(void) { return g_err; }
int64_t cg_launch_count(void) { return g_launches.load(); }

int cg_params_default(cg_params* p) {
    if (!p) return fail(CG_EINVAL, "params is NULL");
    memset(p, 0, sizeof(*p));
    p->struct_bytes = (int32_t)sizeof(cg_params);
    p->dtype = CG_F64;
    p->n_samples = 240; p->row_pitch = 240;
    p->n_scalar = 4; p->n_buckets = 64;
    p->scalar_role[0] = CG_ROLE_SERVER_TOTAL;
    p->scalar_role[1] = CG_ROLE_SERVER_TOTAL | CG_ROLE_SERVER_ERROR;
    p->scalar_role[2] = CG_ROLE_CLIENT_SUM;
    p->scalar_role[3] = CG_ROLE_CLIENT_COUNT;
    p->period_ms = 15000; p->window_ms = 60000;            // elapsed_time=60 (:363)
    p->tick0_ms = 60000; p->tick_step_ms = 10000; p->n_ticks = 355;
    p->left_open = 0;
    p->n_q = 1; p->gate_q = 0; p->q[0] = 0.95;             // :367
    p->one_plus_thr[0] = 1 + 0.05;                         // :176, :440
    p->one_plus_thr[1] = 1 + 0.02;                         // :177, :447
    p->one_plus_thr[2] = 1 + 0.05;                         // :178, :454
    p->checks = CG_G_P95_OK | CG_G_ERR_OK | CG_G_AVG_OK;
    p->init_current = 10; p->max_traffic = 100; p->step = 10;   // :187, :290-291
    p->interval_ticks = 6; p->max_attempts = 10; p->attempt_delay_ticks = 1;   // :292-294
    p->start_tick = 0;
    return CG_OK;
}

int cg_params_check(const cg_params* p) { return check_params(p); }

int cg_metric_width(const cg_params* p) { return p ? p->n_q + CG_M_EXTRA : -1; }

int64_t cg_bytes_per_pair(const cg_params* p) {
    if (!p) return -1;
    return 2 * (int64_t)(p->n_scalar + p->n_buckets) * p->n_samples * (int64_t)elem_size(p) +
           (int64_t)sizeof(cg_decision) + 8;
}

int cg_eval_grid(const void* scalars, const void* buckets, const double* le, const int32_t* phase,
                 int64_t n_pairs, const cg_params* p, double* metrics, uint8_t* valid, void* stream) {
    return run_general(scalars, buckets, le, phase, nullptr, nullptr, n_pairs, p, nullptr, nullptr,
                       metrics, valid, stream);
}

int cg_eval_fused(const void* scalars, const void* buckets, const double* le, const int32_t* phase,
                  const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                  const cg_params* p, cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                  uint8_t* valid, void* stream) {
    if (n_pairs > 0 && !decisions) return fail(CG_EINVAL, "decisions is NULL");
    return run_general(scalars, buckets, le, phase, has_previous, start_tick, n_pairs, p, decisions,
                       gate_bits, metrics, valid, stream);
}

int cg_eval_fused_path(int path, const void* scalars, const void* buckets, const double* le,
                       const int32_t* phase, const uint8_t* has_previous, const int32_t* start_tick,
                       int64_t n_pairs, const cg_params* p, cg_decision* decisions, uint32_t* gate_bits,
                       double* metrics, uint8_t* valid, void* stream) {
    if (n_pairs > 0 && !decisions) return fail(CG_EINVAL, "decisions is NULL");
    if (path < 0 || path > 3) return fail(CG_EINVAL, "path must be 0 (auto), 1 (resident tile), 2 (streaming) or 3 (row streaming)");
    return run_general(scalars, buckets, le, phase, has_previous, start_tick, n_pairs, p, decisions,
                       gate_bits, metrics, valid, stream, path);
}

int cg_eval_irregular(const void* scalars, const void* buckets, const int32_t* ts_scalars,
                      const int32_t* ts_buckets, const int32_t* cnt_scalars, const int32_t* cnt_buckets,
                      const double* le, const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                      const cg_params* p, cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                      uint8_t* valid, void* stream) {
    return cg_eval_irregular_path(0, scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le, has_previous,
                                  start_tick, n_pairs, p, decisions, gate_bits, metrics, valid, stream);
}

int cg_eval_irregular_path(int path, const void* scalars, const void* buckets, const int32_t* ts_scalars,
                           const int32_t* ts_buckets, const int32_t* cnt_scalars, const int32_t* cnt_buckets,
                           const double* le, const uint8_t* has_previous, const int32_t* start_tick, int64_t n_pairs,
                           const cg_params* p, cg_decision* decisions, uint32_t* gate_bits, double* metrics,
                           uint8_t* valid, void* stream) {
    if (path != 0 && path != 1 && path != 3) return fail(CG_EINVAL, "path must be 0 (auto), 1 (resident tile) or 3 (row streaming)");
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (n_pairs == 0) return CG_OK;
    if (!decisions && !gate_bits && !metrics && !valid) return fail(CG_EINVAL, "no output requested");
    if (p->n_scalar > 0 && (!scalars || !ts_scalars || !cnt_scalars))
        return fail(CG_EINVAL, "scalars / ts_scalars / cnt_scalars is NULL but n_scalar > 0");
    if (p->n_buckets > 0 && (!buckets || !ts_buckets || !cnt_buckets || !le))
        return fail(CG_EINVAL, "buckets / ts_buckets / cnt_buckets / le is NULL but n_buckets > 0");
    if ((rc = check_aligned(scalars, "scalars")) || (rc = check_aligned(buckets, "buckets")) ||
        (rc = check_aligned(ts_scalars, "ts_scalars")) || (rc = check_aligned(ts_buckets, "ts_buckets")))
        return rc;
    if ((int64_t)p->tick0_ms + (int64_t)p->n_ticks * p->tick_step_ms >= (1ll << 31) ||
        (int64_t)p->tick0_ms - p->window_ms <= -(1ll << 31))
        return fail(CG_EINVAL, "irregular series: tick times must fit int32 milliseconds");
    // the rows through a ring once the grid is worth a CTA (and whenever the resident tile would not fit)
    IrregularPlan plan;
    const bool tile_fits = path != 3 && plan_irregular(p, n_pairs, &plan) == CG_OK;
    if (path == 3 || (path == 0 && (p->n_ticks >= 32 || !tile_fits))) {
        IRowsPlan rp;
        rc = plan_irows(p, n_pairs, &rp, path != 3 && tile_fits);
        if (rc == CG_OK) {
            cg::IRowsArgs& ra = rp.args;
            ra.scalars = scalars; ra.buckets = buckets; ra.ts_scalars = ts_scalars; ra.ts_buckets = ts_buckets;
            ra.cnt_scalars = cnt_scalars; ra.cnt_buckets = cnt_buckets; ra.le = le;
            ra.has_previous = has_previous; ra.start_tick = start_tick; ra.decisions = decisions;
            ra.gate_bits = gate_bits; ra.metrics = metrics; ra.valid = valid;
            cudaStream_t st = (cudaStream_t)stream;
            if (p->n_q <= 1) return p->dtype == CG_F32 ? launch_irows<float, 1>(rp, st) : launch_irows<double, 1>(rp, st);
            return p->dtype == CG_F32 ? launch_irows<float, CG_MAX_Q>(rp, st) : launch_irows<double, CG_MAX_Q>(rp, st);
        }
        if (path == 3 || !tile_fits) return rc;
    }
    if ((rc = plan_irregular(p, n_pairs, &plan))) return rc;
    cg::IrregularArgs& a = plan.args;
    a.scalars = scalars; a.buckets = buckets; a.ts_scalars = ts_scalars; a.ts_buckets = ts_buckets;
    a.cnt_scalars = cnt_scalars; a.cnt_buckets = cnt_buckets; a.le = le;
    a.has_previous = has_previous; a.start_tick = start_tick; a.decisions = decisions;
    a.gate_bits = gate_bits; a.metrics = metrics; a.valid = valid;
    return p->dtype == CG_F32 ? launch_irregular<float>(plan, (cudaStream_t)stream)
                              : launch_irregular<double>(plan, (cudaStream_t)stream);
}

int cg_stream_eligible(const cg_params* p) {
    if (check_params(p)) return 0;
    return stream_eligible(p) ? 1 : 0;
}

int cg_gate(const double* new_m, const uint8_t* new_v, const double* old_m, const uint8_t* old_v,
            int64_t n, const cg_params* p, uint8_t* detail, void* stream) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n < 0) return fail(CG_EINVAL, "n < 0");
    if (n == 0) return CG_OK;
    if (!new_m || !new_v || !old_m || !old_v || !detail) return fail(CG_EINVAL, "NULL argument");
    gate_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(new_m, new_v, old_m, old_v, n, *p, detail);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch gate_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

int cg_replay(const uint32_t* gate_bits, const uint8_t* has_previous, const int32_t* start_tick,
              int64_t n_pairs, const cg_params* p, cg_decision* decisions, void* stream) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (n_pairs == 0) return CG_OK;
    if (!gate_bits || !decisions) return fail(CG_EINVAL, "NULL argument");
    replay_kernel<<<grid_for(n_pairs, 128), 128, 0, (cudaStream_t)stream>>>(gate_bits, has_previous, start_tick,
                                                                         n_pairs, *p, decisions);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch replay_kernel");
    g_launches.fetch_add(1);
    return CG_OK;
}

int cg_synth_fill(void* scalars, void* buckets, const double* le, int32_t* phase, int64_t n_pairs,
                  int64_t pair_offset, const cg_params* p, uint64_t seed, double reset_prob,
                  double adversarial, void* stream) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (n_pairs == 0) return CG_OK;
    if (!phase || (p->n_scalar && !scalars) || (p->n_buckets && (!buckets || !le)))
        return fail(CG_EINVAL, "NULL argument");
    cg::SynthArgs a;
    a.scalars = scalars; a.buckets = buckets; a.le = le; a.phase = phase; a.n_pairs = n_pairs;
    a.pair_offset = pair_offset; a.seed = seed; a.reset_prob = reset_prob; a.adversarial = adversarial;
    int g = p->period_ms, h = p->tick_step_ms;
    while (h) { int t = g % h; g = h; h = t; }
    a.phase_avoid = g;
    a.p = *p;
    const int64_t total = n_pairs * 2 * (p->n_scalar + p->n_buckets);
    const int grid = grid_for(total, 128);
    if (p->dtype == CG_F32) cg::synth_kernel<float><<<grid, 128, 0, (cudaStream_t)stream>>>(a);
    else cg::synth_kernel<double><<<grid, 128, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "launch synth_kernel");
    return CG_OK;
}

int cg_store_append(void* scalars, void* buckets, const void* new_scalars, const void* new_buckets,
                    int64_t n_pairs, const cg_params* p, int32_t column, void* stream) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (column < 0 || column >= p->row_pitch) return fail(CG_EINVAL, "column %d outside the row (row_pitch %d)", column, p->row_pitch);
    if (n_pairs == 0) return CG_OK;
    if ((p->n_scalar && (!scalars || !new_scalars)) || (p->n_buckets && (!buckets || !new_buckets)))
        return fail(CG_EINVAL, "NULL argument");
    cudaStream_t s = (cudaStream_t)stream;
    for (int part = 0; part < 2; ++part) {
        const int64_t n_series = n_pairs * 2 * (part ? p->n_buckets : p->n_scalar);
        if (n_series == 0) continue;
        void* rows = part ? buckets : scalars;
        const void* fresh = part ? new_buckets : new_scalars;
        const int grid = grid_for(n_series, 256);
        if (p->dtype == CG_F32)
            cg::store_append_kernel<float><<<grid, 256, 0, s>>>((float*)rows, (const float*)fresh, n_series, p->row_pitch, column);
        else
            cg::store_append_kernel<double><<<grid, 256, 0, s>>>((double*)rows, (const double*)fresh, n_series, p->row_pitch, column);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "launch store_append_kernel");
        g_launches.fetch_add(1);
    }
    return CG_OK;
}

int cg_store_shift(void* scalars, void* buckets, int64_t n_pairs, const cg_params* p, int32_t shift, int32_t count,
                   void* stream) {
    int rc = check_params(p);
    if (rc) return rc;
    if (n_pairs < 0) return fail(CG_EINVAL, "n_pairs < 0");
    if (shift < 0 || count < 0 || (int64_t)shift + count > p->row_pitch)
        return fail(CG_EINVAL, "shift %d + count %d exceed the row (row_pitch %d)", shift, count, p->row_pitch);
    if (n_pairs == 0 || shift == 0 || count == 0) return CG_OK;
    const size_t smem = (size_t)count * elem_size(p);
    if (smem > 48 * 1024) return fail(CG_EINVAL, "count %d: a row segment of more than 48 KB cannot be staged", count);
    cudaStream_t s = (cudaStream_t)stream;
    for (int part = 0; part < 2; ++part) {
        const int64_t n_rows = n_pairs * 2 * (part ? p->n_buckets : p->n_scalar);
        void* rows = part ? buckets : scalars;
        if (n_rows == 0) continue;
        if (!rows) return fail(CG_EINVAL, "NULL argument");
        int64_t g = n_rows < 148 * 16 ? n_rows : 148 * 16;
        if (p->dtype == CG_F32)
            cg::store_shift_kernel<float><<<(int)g, 128, smem, s>>>((float*)rows, n_rows, p->row_pitch, shift, count);
        else
            cg::store_shift_kernel<double><<<(int)g, 128, smem, s>>>((double*)rows, n_rows, p->row_pitch, shift, count);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "launch store_shift_kernel");
        g_launches.fetch_add(1);
    }
    return CG_OK;
}

int cg_launch_info(const cg_params* p, int64_t n_pairs, int32_t* grid, int32_t* block, int32_t* smem_bytes) {
    int rc = check_params(p);
    if (rc) return rc;
    if (stream_eligible(p)) {
        StreamPlan sp;
        if ((rc = plan_stream(p, n_pairs, &sp))) return rc;
        if (grid) *grid = sp.grid;
        if (block) *block = cg::kLanes;
        if (smem_bytes) *smem_bytes = (int32_t)sp.args.smem_bytes;
        return CG_OK;
    }
    RowsPlan rp;
    if (p->n_ticks >= 32 && plan_rows(p, n_pairs, &rp, true) == CG_OK) {
        if (grid) *grid = rp.grid;
        if (block) *block = rp.block;
        if (smem_bytes) *smem_bytes = (int32_t)rp.args.smem_bytes;
        return CG_OK;
    }
    GeneralPlan plan;
    if ((rc = plan_general(p, n_pairs, &plan))) return rc;
    if (grid) *grid = plan.grid;
    if (block) *block = plan.block;
    if (smem_bytes) *smem_bytes = (int32_t)plan.args.smem_bytes;
    return CG_OK;
}

}  // extern "C"
