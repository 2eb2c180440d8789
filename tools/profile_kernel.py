"""Small fixed workload for ncu captures: `python tools/profile_kernel.py [pairs] [path] [reps]`.
Generates BASELINE configs[3]-shaped series on the GPU and runs cg_eval_fused `reps` times."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
path = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n_samples = int(sys.argv[4]) if len(sys.argv) > 4 else 240      # shorter series: less shared memory per CTA
p = P.make_params(n_samples=n_samples, n_ticks=(n_samples * 15 - 60) // 10 + 1) if n_samples != 240 else P.make_params()
le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
ev[0].record()
for i in range(reps):
    dec = engine.eval_fused(sc, bk, le, ph, p, path=path)
    ev[i + 1].record()
torch.cuda.synchronize()
ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]
gb = n * p.bytes_per_pair() / 1e9
print(f"pairs={n} path={path} ms={['%.3f' % m for m in ms]} GB/s={gb / (min(ms) * 1e-3):.1f} states={torch.bincount(dec[:, 2], minlength=4).tolist()}")

