#!/usr/bin/env python
"""Benchmark of the canary metric-validation path (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W          # this library (CUDA, sm_100a)
  python bench.py --impl reference --gpus N --steps K ... # CPU arm: the oracle port on host cores

A "step" is one pass of the hot path over one batch: every model pair's full series
(dense tick grid -> gate -> traffic step loop).  Workload at N GPUs: BASELINE.json
configs[3] per GPU (100k pairs x (4 scalar + 64 bucket) series x 240 samples fp64 with
counter resets, full promote/hold/abort schedule), i.e. weak scaling; at N > 1 each step
ends with the NCCL all-gather of the 24-byte decision records.  Prints ONE JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "canary_pair_evals_per_sec"
UNIT = "pairs/s"
WORKLOAD = ("BASELINE.json configs[3]: 100k model pairs x 1 h of 15 s scrapes (68 series x 240 samples, "
            "fp64) with counter resets, dense 10 s tick grid, full promote/hold/abort schedule")


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def traffic_from_profile():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f)
    except Exception:  # noqa: BLE001
        return None


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md): NVML polled every
    ~10 ms from a thread (the timed region is only ~0.1-0.2 s); nvidia-smi as a fallback."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.sm, self.reasons, self.max_mhz = index, [], set(), None
        self.stop = threading.Event()
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001
            self.nvml = None

    def _poll_nvml(self):
        n = self.nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
        r = n.nvmlDeviceGetCurrentClocksEventReasons(self.handle) if hasattr(
            n, "nvmlDeviceGetCurrentClocksEventReasons") else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("sw_thermal_slowdown", 0x20),
                          ("hw_thermal_slowdown", 0x40)):
            if r & bit:
                self.reasons.add(name)

    def _poll_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True,
                             timeout=5).stdout.strip()
        r = [c.strip() for c in out.split(",")]
        self.sm.append(float(r[0]))
        self.max_mhz = float(r[1])
        for nm, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
            if val.lower().startswith("active"):
                self.reasons.add(nm)

    def _run(self):
        while not self.stop.is_set():
            try:
                self._poll_nvml() if self.nvml else self._poll_smi()
            except Exception:  # noqa: BLE001
                pass
            self.stop.wait(0.01 if self.nvml else 0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(sm), "source": "nvml" if self.nvml else "nvidia-smi"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


# ------------------------------------------------------------------------------------
def reference_arm(args):
    """The CPU arm: the reference's algorithm for this path (oracle/promql_oracle.c, a
    restatement; the reference itself is one Python file that cannot travel to the GPU box
    and delegates the arithmetic to an external Prometheus) with all host threads, on a
    bounded sample of the same workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import numpy as np
    import canary_b200  # noqa: F401
    from canary_b200 import params as P, synth
    from oracle import c_oracle as co
    p = P.make_params()
    cores = host_cores()
    base_n, reps = 256, max(1, args.ref_pairs // 256)
    b = synth.generate(p, base_n, seed=0xC0FFEE + 3, adversarial=0.01)
    n = base_n * reps
    tile = lambda a: np.ascontiguousarray(np.concatenate([a] * reps, axis=0))  # noqa: E731
    scal, buck, phase = tile(b.scalars), tile(b.buckets), tile(b.phase)
    phase[:] = (phase + (np.arange(n)[:, None] // base_n) * 37) % p.period_ms
    phase[phase % 5000 == 0] += 1
    for _ in range(args.warmup):
        co.eval_fused(scal, buck, b.le, phase, p, n_threads=cores, want_bits=False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dec, _ = co.eval_fused(scal, buck, b.le, phase, p, n_threads=cores, want_bits=False)
    dt = time.perf_counter() - t0
    t1 = time.perf_counter()
    lazy = co.eval_lazy(scal, buck, b.le, phase, p, n_threads=cores)
    dt_lazy = time.perf_counter() - t1
    assert (lazy == dec).all()
    value = n * args.steps / dt
    sample = (f"{n} pairs per step ({base_n} seeded pairs x {reps} phase-shifted copies) of the same "
              f"workload, dense tick grid, {args.steps} steps")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": {"workload": WORKLOAD, "sample_pairs_per_step": n},
        "series_samples_per_sec": value * 2 * (p.n_scalar + p.n_buckets) * p.n_samples,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "lazy_value": n / dt_lazy,
                         "lazy_note": "same decisions evaluating only the ticks the schedule visits, "
                                      "as the reference's loop does (mlflow_operator.py:300-310)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------
def _timed_ms(fn, reps=5):
    """min over `reps` of the CUDA-event time of fn() on the current stream, after one warm-up call"""
    import torch
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))


def secondary_configs(dev, peak, n_main, seed):
    """The other BASELINE.json configs and storage modes, outside the headline's timed region:
    kernel time with inputs resident in HBM, pairs/s, algorithmic GB/s and its fraction of the HBM peak."""
    import torch
    from canary_b200 import engine, params as P, synth
    out = {}

    def entry(name, n, p, fn, bytes_per_pair=None, note=None):
        ms = _timed_ms(fn)
        bpp = p.bytes_per_pair() if bytes_per_pair is None else bytes_per_pair
        gbs = n * bpp / (ms * 1e-3) / 1e9
        out[name] = {"pairs": n, "ms": ms, "pairs_per_s": n / (ms * 1e-3), "gb_per_s": gbs, "frac": gbs / peak,
                     "bytes_per_pair": bpp, "streaming_kernel": bool(engine.stream_eligible(p))}
        if note:
            out[name]["note"] = note

    # configs[1]: 1k pairs, error-rate gate only, two server counters, no histogram
    p1 = P.make_params(n_scalar=2, n_buckets=0, checks=P.G_ERR_OK)
    le0 = torch.zeros(0, dtype=torch.float64, device=dev)
    sc, bk, ph = engine.synth_fill(1000, p1, le0, seed=seed + 1, device=dev)
    entry("configs1_1k_error_rate", 1000, p1, lambda: engine.eval_fused(sc, bk, le0, ph, p1), note="launch-latency bound")
    # configs[2]: 10k pairs, 64 buckets, p50 / p95 / p99
    p2 = P.make_params(quantiles=(0.5, 0.95, 0.99), gate_q=1)
    le = torch.from_numpy(synth.default_le(p2.n_buckets)).to(dev)
    sc2, bk2, ph2 = engine.synth_fill(10000, p2, le, seed=seed + 2, device=dev)
    entry("configs2_10k_p50_p95_p99", 10000, p2, lambda: engine.eval_fused(sc2, bk2, le, ph2, p2))
    del sc2, bk2
    # configs[3] in fp32 storage (fp64 arithmetic on fp32-rounded samples: half the bytes)
    p3 = P.make_params(dtype=P.CG_F32)
    sc3, bk3, ph3 = engine.synth_fill(n_main, p3, le, seed=seed + 3, device=dev)
    entry("configs3_fp32_storage", n_main, p3, lambda: engine.eval_fused(sc3, bk3, le, ph3, p3))
    # ... and the cold end-to-end path in that storage mode: half the bytes cross PCIe (same API and timed
    # region as the headline `e2e`: pinned host series -> chunks -> kernel -> decisions in pinned host memory)
    try:
        import time as _time
        dec3 = engine.eval_fused(sc3, bk3, le, ph3, p3)
        pipe3 = engine.HostPipeline(p3, chunk_pairs=4096)
        hs3 = torch.empty(sc3.shape, dtype=sc3.dtype, pin_memory=True)
        hb3 = torch.empty(bk3.shape, dtype=bk3.dtype, pin_memory=True)
        hp3 = torch.empty(ph3.shape, dtype=ph3.dtype, pin_memory=True)
        hs3.copy_(sc3); hb3.copy_(bk3); hp3.copy_(ph3)
        out3 = torch.empty((n_main, 6), dtype=torch.int32, pin_memory=True)
        torch.cuda.synchronize()
        pipe3.run(hs3, hb3, hp3, le, out3)                                   # warm-up
        same3 = bool((out3.numpy() == dec3.cpu().numpy()).all())
        steps3 = 2
        t0 = _time.perf_counter()
        for _ in range(steps3):
            pipe3.run(hs3, hb3, hp3, le, out3)
        torch.cuda.synchronize()
        ms3 = 1e3 * (_time.perf_counter() - t0) / steps3
        out["e2e_fp32_storage"] = {"pairs": n_main, "ms_per_step": ms3, "pairs_per_s": n_main / (ms3 * 1e-3),
                                   "h2d_bytes_per_step": pipe3.h2d_bytes, "d2h_bytes_per_step": pipe3.d2h_bytes,
                                   "matches_device_path": same3, "steps": steps3,
                                   "note": "engine.HostPipeline.run with fp32 series: the cold upload at half the bytes"}
        del hs3, hb3, hp3, pipe3
    except Exception as e:  # noqa: BLE001
        out["e2e_fp32_storage"] = {"error": str(e)[:200]}
    del sc3, bk3
    # irregular series (per-series time stamps): 4096 pairs of the full layout
    p4 = P.make_params()
    ni = 4096
    sc4, bk4, ph4 = engine.synth_fill(ni, p4, le, seed=seed + 3, device=dev)
    grid = ph4[:, :, None, None].long() + torch.arange(p4.row_pitch, device=dev).long() * p4.period_ms
    ts_s = grid.expand(ni, 2, p4.n_scalar, p4.row_pitch).to(torch.int32).contiguous()
    ts_b = grid.expand(ni, 2, p4.n_buckets, p4.row_pitch).to(torch.int32).contiguous()
    cnt_s = torch.full((ni, 2, p4.n_scalar), p4.n_samples, dtype=torch.int32, device=dev)
    cnt_b = torch.full((ni, 2, p4.n_buckets), p4.n_samples, dtype=torch.int32, device=dev)
    bpp_i = p4.bytes_per_pair() + 2 * (p4.n_scalar + p4.n_buckets) * (p4.row_pitch * 4 + 4)
    entry("irregular_4096", ni, p4, lambda: engine.eval_irregular(sc4, bk4, ts_s, ts_b, cnt_s, cnt_b, le, p4),
          bytes_per_pair=bpp_i, note="cg_eval_irregular: values + int32 time stamps + counts")
    out["irregular_4096"]["streaming_kernel"] = False
    return out


def ingest_rate():
    """Host-side ingest: samples per second from query_range JSON text (string
    values) and from raw int64/fp64 arrays (remote-read shaped) into the library's SoA layout."""
    import json as _json
    import numpy as np
    from canary_b200 import ingest, params as P, synth
    p = P.make_params(n_samples=240, n_buckets=16)
    b = synth.generate(p, 8, seed=4)
    epoch = 1_700_000_000_000
    answers = {ingest.SERVER_COUNT: [], ingest.CLIENT_SUM: [], ingest.CLIENT_COUNT: [], ingest.CLIENT_BUCKET: []}
    versions = {}
    for i in range(8):
        key = (f"model-{i:03d}", "prod")
        versions[key] = ("v2", "v1")
        for v, pred in enumerate(("v2", "v1")):
            base = {"deployment_name": key[0], "namespace": key[1], "predictor_name": pred}
            ts = [(epoch + int(b.phase[i, v]) + k * p.period_ms) / 1000.0 for k in range(p.n_samples)]
            mk = lambda labels, row: {"metric": labels, "values": [[t, repr(float(x))] for t, x in zip(ts, row)]}  # noqa: E731
            answers[ingest.SERVER_COUNT] += [mk(dict(base, code="200"), b.scalars[i, v, 0]), mk(dict(base, code="500"), b.scalars[i, v, 1])]
            answers[ingest.CLIENT_SUM].append(mk(dict(base), b.scalars[i, v, 2]))
            answers[ingest.CLIENT_COUNT].append(mk(dict(base), b.scalars[i, v, 3]))
            for k in range(p.n_buckets):
                le = "+Inf" if np.isinf(b.le[k]) else repr(float(b.le[k]))
                answers[ingest.CLIENT_BUCKET].append(mk(dict(base, le=le), b.buckets[i, v, k]))
    wire = {n_: _json.dumps({"status": "success", "data": {"resultType": "matrix", "result": r}}) for n_, r in answers.items()}
    samples = 8 * 2 * (p.n_scalar + p.n_buckets) * p.n_samples
    t0 = time.perf_counter()
    ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)
    t_json = time.perf_counter() - t0
    raw = {n_: ingest.parse_matrix(a_) for n_, a_ in wire.items()}
    t0 = time.perf_counter()
    ingest.build_host_batch(raw, versions, period_ms=p.period_ms, epoch_ms=epoch)
    t_raw = time.perf_counter() - t0
    return {"samples": samples, "json_samples_per_s": samples / t_json, "raw_samples_per_s": samples / t_raw,
            "note": "one host core; JSON text with string values vs int64/fp64 arrays (remote-read shaped)"}


# ------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs-per-gpu", type=int, default=100_000)
    ap.add_argument("--ref-pairs", type=int, default=4096)
    ap.add_argument("--cpu-pairs", type=int, default=8192)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--ref-python-pairs-per-core", type=int, default=16)
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = max(args.warmup, 1)
    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch
    import canary_b200  # noqa: F401
    from canary_b200 import _lib, engine, params as P, sharding, synth
    from oracle import c_oracle as co      # checker + cpu_baseline leg only

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    p = P.make_params()
    n = args.pairs_per_gpu
    le = torch.from_numpy(synth.default_le(p.n_buckets)).to(dev)
    SEED = 0xC0FFEE + 3
    scalars, buckets, phase = engine.synth_fill(n, p, le, seed=SEED, pair_offset=rank * n)
    torch.cuda.synchronize()
    bytes_per_pair = p.bytes_per_pair()

    # decisions live in one gather buffer; each rank's kernel writes its own slice in place
    n_total = world * n
    gathered = sharding.alloc_gather_buffer(n_total, world, dev)
    mine = sharding.my_slice(gathered, n_total, rank, world)

    def step():
        engine.eval_fused(scalars, buckets, le, phase, p, out=mine)
        if dist is not None:
            sharding.gather_decisions(gathered, n_total)

    # ---- parity spot check against the oracle before timing anything
    step()
    torch.cuda.synchronize()
    idx = torch.arange(0, n, max(1, n // 64), device=dev)[:64]
    sub = [t[idx].contiguous().cpu().numpy() for t in (scalars, buckets, phase)]
    want, _ = co.eval_fused(sub[0], sub[1], le.cpu().numpy(), sub[2], p, n_threads=host_cores(), want_bits=False)
    parity_ok = bool((mine[idx].cpu().numpy() == want).all())

    # ---- N > 1: pairs that OTHER ranks computed must be right in this rank's gathered buffer, and
    # the collective is timed by itself (the generator is a pure function of the global pair index)
    gather_parity, gather_ms = None, None
    if dist is not None:
        if rank == 0:
            rng = np.random.default_rng(7)
            gidx = np.sort(rng.choice(np.arange(n, n_total), size=min(64, n_total - n), replace=False))
            parts = [engine.synth_fill(1, p, le, seed=SEED, pair_offset=int(g)) for g in gidx]
            rs, rb, rp = (torch.cat([t[k] for t in parts]).cpu().numpy() for k in range(3))
            want_r, _ = co.eval_fused(rs, rb, le.cpu().numpy(), rp, p, n_threads=host_cores(), want_bits=False)
            got_r = gathered[torch.from_numpy(gidx).to(dev)].cpu().numpy()
            gather_parity = bool((got_r == want_r).all())
        torch.cuda.synchronize()
        dist.barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(10):
            sharding.gather_decisions(gathered, n_total)
        g1.record()
        torch.cuda.synchronize()
        gt = torch.tensor([g0.elapsed_time(g1) / 10], dtype=torch.float64, device=dev)
        dist.all_reduce(gt, op=dist.ReduceOp.MAX)
        gather_ms = float(gt[0])

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 2)]
    launches0 = _lib.launch_count()
    with ClockSampler(local) as clocks:
        ev[0].record(stream)
        for i in range(args.steps):
            ev[1 + 2 * i].record(stream)
            engine.eval_fused(scalars, buckets, le, phase, p, out=mine)
            ev[2 + 2 * i].record(stream)
            if dist is not None:
                sharding.gather_decisions(gathered, n_total)
        ev[-1].record(stream)
        torch.cuda.synchronize()
    launches = _lib.launch_count() - launches0
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    total_ms = ev[0].elapsed_time(ev[-1])
    kern_ms = sum(ev[1 + 2 * i].elapsed_time(ev[2 + 2 * i]) for i in range(args.steps)) / args.steps
    if dist is not None:
        t = torch.tensor([total_ms, kern_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kern_ms = float(t[0]), float(t[1])
    ms_per_step = total_ms / args.steps
    value = world * n / (ms_per_step * 1e-3)
    peak, peak_src = measured_peak()
    achieved = n * bytes_per_pair / (kern_ms * 1e-3) / 1e9
    traffic = traffic_from_profile()

    # ---- end to end through the public host-buffer API: pinned host -> GPU -> decisions on host
    e2e = None
    if not args.no_e2e:
        pipe = engine.HostPipeline(p, chunk_pairs=4096)
        hs = torch.empty(scalars.shape, dtype=scalars.dtype, pin_memory=True)
        hb = torch.empty(buckets.shape, dtype=buckets.dtype, pin_memory=True)
        hp = torch.empty(phase.shape, dtype=phase.dtype, pin_memory=True)
        hs.copy_(scalars); hb.copy_(buckets); hp.copy_(phase)
        out_h = torch.empty((n, 6), dtype=torch.int32, pin_memory=True)
        torch.cuda.synchronize()
        pipe.run(hs, hb, hp, le, out_h)                    # warm-up
        same = bool((out_h.numpy() == mine.cpu().numpy()).all())
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(args.e2e_steps):
            pipe.run(hs, hb, hp, le, out_h)
        e1.record(stream)
        torch.cuda.synchronize()
        wall_ms = 1e3 * (time.perf_counter() - t0)
        e2e_ms = max(e0.elapsed_time(e1), wall_ms) / args.e2e_steps
        if dist is not None:
            t = torch.tensor([e2e_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t[0])
        e2e = {"value": world * n / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
               "d2h_bytes_per_step": pipe.d2h_bytes, "ms_per_step": e2e_ms, "steps": args.e2e_steps,
               "matches_device_path": same,
               "api": "engine.HostPipeline.run (pinned host series -> cg_eval_fused per 4096-pair chunk -> decisions in pinned host memory)"}
        del hs, hb, hp

    # ---- end to end the way the operator lives: the hour stays in HBM (engine.RollingStore), every
    # evaluation brings ONE new scrape per series from pinned host memory, then the decisions go back
    e2e_inc = None
    if not args.no_e2e:
        steps_i = max(args.e2e_steps, 6)
        extra = steps_i + 2
        extra += extra & 1
        # the same series continued: the generator is a pure function of (seed, pair, version, series, sample)
        p_long = P.make_params(n_samples=p.n_samples + extra, n_ticks=p.n_ticks)
        sc_l, bk_l, ph_l = engine.synth_fill(n, p_long, le, seed=SEED, pair_offset=rank * n)
        store = engine.RollingStore(p, n, le, slack=16)
        store.load(sc_l[..., :p.n_samples], bk_l[..., :p.n_samples], ph_l)
        K, B = p.n_scalar, p.n_buckets
        fresh_sc = [sc_l[..., p.n_samples + j].contiguous().cpu().pin_memory() for j in range(extra)]
        fresh_bk = [bk_l[..., p.n_samples + j].contiguous().cpu().pin_memory() for j in range(extra)]
        del sc_l, bk_l
        out_i = torch.empty((n, 6), dtype=torch.int32, pin_memory=True)
        dec_i = torch.empty((n, 6), dtype=torch.int32, device=dev)

        def inc_step(j):
            store.append_scrape(fresh_sc[j], fresh_bk[j])
            store.evaluate(out=dec_i)
            out_i.copy_(dec_i, non_blocking=True)

        for j in range(2):
            inc_step(j)                                     # warm-up (also an odd and an even view offset)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        i0.record(stream)
        for j in range(steps_i):
            inc_step(2 + j)
        i1.record(stream)
        torch.cuda.synchronize()
        inc_ms = max(i0.elapsed_time(i1), 1e3 * (time.perf_counter() - t0)) / steps_i
        if dist is not None:
            t = torch.tensor([inc_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            inc_ms = float(t[0])
        # the store's view after the last append, against the oracle on a strided sample of pairs
        off, n_s, pp_v, _ = store.view()
        sidx = torch.arange(0, n, max(1, n // 32), device=dev)[:32]
        q = pp_v.copy()
        q.row_pitch = n_s
        vs = store.scalars[sidx][..., off:off + n_s].contiguous().cpu().numpy()
        vb = store.buckets[sidx][..., off:off + n_s].contiguous().cpu().numpy()
        vp = (store.phase[sidx] + off * p.period_ms).to(torch.int32).cpu().numpy()
        want_i, _ = co.eval_fused(vs, vb, le.cpu().numpy(), vp, q, n_threads=host_cores(), want_bits=False)
        e2e_inc = {"value": world * n / (inc_ms * 1e-3), "unit": UNIT, "ms_per_step": inc_ms, "steps": steps_i,
                   "h2d_bytes_per_step": n * 2 * (K + B) * p.elem_bytes, "d2h_bytes_per_step": n * 24,
                   "matches_oracle_on_sample": bool((out_i[sidx.cpu()].numpy() == want_i).all()),
                   "api": "engine.RollingStore.append_scrape (one new scrape per series, pinned host -> HBM on the store's copy "
                          "stream, cg_store_append) + RollingStore.evaluate (cg_eval_fused over the newest hour) + decisions to "
                          "pinned host memory; the upload of step j+1 overlaps the evaluation of step j"}
        del store, fresh_sc, fresh_bk
        torch.cuda.empty_cache()

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only): the oracle port
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        m = min(args.cpu_pairs, n)
        sub = [t[:m].contiguous().cpu().numpy() for t in (scalars, buckets, phase)]
        cores = host_cores()
        le_h = le.cpu().numpy()
        co.eval_fused(sub[0][:256], sub[1][:256], le_h, sub[2][:256], p, n_threads=cores, want_bits=False)
        t0 = time.perf_counter()
        cdec, _ = co.eval_fused(sub[0], sub[1], le_h, sub[2], p, n_threads=cores, want_bits=False)
        dt = time.perf_counter() - t0
        t1 = time.perf_counter()
        co.eval_lazy(sub[0], sub[1], le_h, sub[2], p, n_threads=cores)
        dt_lazy = time.perf_counter() - t1
        parity_ok = parity_ok and bool((cdec == mine[:m].cpu().numpy()).all())
        cpu = {"value": m / dt, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"first {m} pairs of the same workload, dense tick grid, one pass ({dt:.1f} s); "
                         f"all {m} decisions equal the GPU's",
               "lazy_value": m / dt_lazy}
        # the reference's OWN Python (unmodified file, oracle/_ref copy on this box) on the
        # host cores, one process per core, on a small sample of the same workload
        try:
            res = subprocess.run([sys.executable, "-m", "oracle.ref_bench", str(args.ref_python_pairs_per_core), str(cores)],
                                 cwd=ROOT, capture_output=True, text=True, timeout=600)
            rp = json.loads(res.stdout.strip().splitlines()[-1])
        except Exception as e:  # noqa: BLE001
            rp = {"unavailable": f"{type(e).__name__}: {e}"}
        if "pairs_per_s" in rp:
            cpu["reference_python"] = {"pairs_per_s": rp["pairs_per_s"], "cores": rp["cores"],
                                       "sample": f"{rp['pairs']} pairs of the same workload in {rp['seconds']:.1f} s",
                                       "extrapolated": True, "decisions_equal_oracle": rp["matches_oracle"], "what": rp["what"]}
        else:
            cpu["reference_python"] = rp

    # ---- the other BASELINE configs and storage modes (outside the headline's timed region)
    secondary = None
    if rank == 0 and not args.no_secondary:
        del scalars, buckets
        torch.cuda.empty_cache()
        secondary = secondary_configs(dev, peak, n, 0xC0FFEE)
        secondary["ingest_host"] = ingest_rate()
    elif not args.no_secondary:
        del scalars, buckets
        torch.cuda.empty_cache()

    # ---- BASELINE.json configs[4] as written: 1M pairs sharded over the N GPUs, decisions gathered
    if dist is not None and not args.no_secondary:
        dist.barrier()
        n4_total = 1_000_000
        n4 = sharding.shard_size(n4_total, world)
        lo4, hi4 = sharding.shard_range(n4_total, rank, world)
        sc4, bk4, ph4 = engine.synth_fill(hi4 - lo4, p, le, seed=SEED + 1, pair_offset=lo4)
        buf4 = sharding.alloc_gather_buffer(n4_total, world, dev)
        mine4 = sharding.my_slice(buf4, n4_total, rank, world)

        def step4():
            engine.eval_fused(sc4, bk4, le, ph4, p, out=mine4)
            sharding.gather_decisions(buf4, n4_total)

        for _ in range(2):
            step4()
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        for _ in range(5):
            step4()
        a1.record(stream)
        torch.cuda.synchronize()
        t4 = torch.tensor([a0.elapsed_time(a1) / 5], dtype=torch.float64, device=dev)
        dist.all_reduce(t4, op=dist.ReduceOp.MAX)
        ok4 = None
        if rank == 0:      # a few pairs from the LAST rank's slice, checked against the oracle
            g4 = np.arange(n4_total - 24, n4_total)
            parts = [engine.synth_fill(1, p, le, seed=SEED + 1, pair_offset=int(g)) for g in g4]
            rs, rb, rp4 = (torch.cat([t[k] for t in parts]).cpu().numpy() for k in range(3))
            w4, _ = co.eval_fused(rs, rb, le.cpu().numpy(), rp4, p, n_threads=host_cores(), want_bits=False)
            ok4 = bool((buf4[torch.from_numpy(g4).to(dev)].cpu().numpy() == w4).all())
            secondary = secondary or {}
            secondary["configs4_1M"] = {"pairs_total": n4_total, "pairs_per_gpu": n4, "ms_per_step": float(t4[0]),
                                        "pairs_per_s": n4_total / (float(t4[0]) * 1e-3), "scaling": "strong (1M pairs over N GPUs)",
                                        "gather_parity": ok4, "includes": "kernel + NCCL all-gather of the decision records"}
        del sc4, bk4

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_gpu": n, "series_per_version": p.n_scalar + p.n_buckets,
                       "samples": p.n_samples, "ticks": p.n_ticks, "bytes_per_pair": bytes_per_pair,
                       "l2": f"inputs are {n * bytes_per_pair / 1e9:.1f} GB per GPU, far larger than the 126 MB L2; no flush",
                       "collective": "all_gather of 24 B decision records (NCCL)" if world > 1 else "none"},
            "series_samples_per_sec": value * 2 * (p.n_scalar + p.n_buckets) * p.n_samples,
            "kernel_ms": kern_ms,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak,
                         "traffic": None if traffic is None else traffic["bytes_per_pair"] * n,
                         "traffic_source": None if traffic is None else traffic["source"] + "; scaled per pair",
                         "peak_source": peak_src, "kernel": "cg::warp_kernel<double,1> behind cg_eval_fused",
                         "algorithmic_bytes_per_launch": n * bytes_per_pair},
            "cpu_baseline": cpu, "e2e": e2e, "e2e_incremental": e2e_inc, "secondary": secondary,
            "gather_parity": gather_parity, "gather_ms": gather_ms, "gpu_launches": int(launches),
            "clocks": clocks.summary(), "parity_spot_check": parity_ok,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
