"""Time the UNMODIFIED reference on this machine's host cores.

TEST / BENCHMARK INFRASTRUCTURE ONLY (bench.py's cpu_baseline leg runs it as a subprocess; the
product never imports oracle/).  What is timed is the reference's own code: `get_model_metrics`
(mlflow_operator.py:363-417), `should_promote_model` (:419-460) and the step loop (:296-352),
imported untouched by oracle/ref_loader.py, with its PrometheusConnect answered by the pure-Python
fake Prometheus of oracle/pyoracle.py and its sleeps by a virtual clock.  The reference evaluates
only the ticks its schedule visits (about 20 per pair); one process per core.

  python -m oracle.ref_bench [pairs per core] [cores]   -> one JSON line
"""
from __future__ import annotations

import json
import multiprocessing as mp
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_BATCH = None


def _work(span):
    from oracle import soa_bridge as sb
    lo, hi = span
    out = []
    for i in range(lo, hi):
        r = sb.run_reference_on_pair(_BATCH, i)
        out.append((r["cur"], r["prev"], r["state"], r["n_promotions"], r["n_evals"]))
    return out


def main():
    global _BATCH
    per_core = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    cores = int(sys.argv[2]) if len(sys.argv) > 2 else len(os.sched_getaffinity(0))
    from oracle import ref_loader as rl
    if not rl.reference_available():
        print(json.dumps({"unavailable": f"no reference file at {rl.REFERENCE_FILE}"}))
        return 0
    import numpy as np
    import canary_b200  # noqa: F401
    from canary_b200 import params as P, synth
    from oracle import c_oracle as co
    p = P.make_params()
    n = per_core * cores
    _BATCH = synth.generate(p, n, seed=0xC0FFEE + 3, adversarial=0.01)      # BASELINE.json configs[3] shape
    rl.load_reference()                                                       # import once, before forking
    spans = [(c * per_core, (c + 1) * per_core) for c in range(cores)]
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        pool.map(_work, [(0, 1)] * cores)                                     # warm the workers
        t0 = time.perf_counter()
        parts = pool.map(_work, spans)
        dt = time.perf_counter() - t0
    got = np.array([row for part in parts for row in part], dtype=np.int32)
    want, _ = co.eval_fused(_BATCH.scalars, _BATCH.buckets, _BATCH.le, _BATCH.phase, p, n_threads=cores, want_bits=False)
    print(json.dumps({"pairs_per_s": n / dt, "cores": cores, "pairs": n, "seconds": dt,
                      "matches_oracle": bool(np.array_equal(got, want[:, :5])),
                      "what": "unmodified mlflow_operator.py (get_model_metrics, should_promote_model, step loop) over the "
                              "pure-Python fake Prometheus, virtual clock, one process per core; evaluates only the ticks "
                              "its schedule visits",
                      "reference_file": rl.REFERENCE_FILE}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
