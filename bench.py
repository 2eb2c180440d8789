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
    scalars, buckets, phase = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3, pair_offset=rank * n)
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
                         "peak_source": peak_src, "kernel": "cg_eval_fused",
                         "algorithmic_bytes_per_launch": n * bytes_per_pair},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks.summary(), "parity_spot_check": parity_ok,
        }
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
