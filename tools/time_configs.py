"""Developer aid: kernel time of BASELINE.json configs[1] and configs[2] (the bench line is configs[3])."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))


for name, n, p in (("configs[1] 1k pairs, error-rate gate, K=2, no histogram", 1000,
                    P.make_params(n_scalar=2, n_buckets=0, checks=P.G_ERR_OK)),
                   ("configs[2] 10k pairs, 64 buckets, p50/p95/p99", 10000,
                    P.make_params(quantiles=(0.5, 0.95, 0.99), gate_q=1))):
    le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda() if p.n_buckets else torch.zeros(0, dtype=torch.float64, device="cuda")
    sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 1)
    ms = timed(lambda: engine.eval_fused(sc, bk, le, ph, p))
    gb = n * p.bytes_per_pair() / 1e9
    print(f"{name}: {ms:.3f} ms, {n / ms * 1e3:.3e} pairs/s, {gb / ms * 1e3:.0f} GB/s, streaming={engine.stream_eligible(p)}")
