"""Developer aid: rare-path rates of the streaming kernel, from the host emulation, on series made by the
library's own device generator (run on the GPU box) or by the numpy generator (anywhere).
  python tools/emul_stats.py [pairs] [device|numpy]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import params as P, synth  # noqa: E402
import test_host_emul as T  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
src = sys.argv[2] if len(sys.argv) > 2 else "numpy"
p = P.make_params()
if src == "device":
    import torch
    from canary_b200 import engine
    le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
    sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3)
    b = synth.SynthBatch(sc.cpu().numpy(), bk.cpu().numpy(), le.cpu().numpy(), ph.cpu().numpy(), p)
else:
    b = synth.generate(p, n, seed=1234, adversarial=0.01)
dec, bits, m, v, c = T.emulate(b, p)
rows = max(int(c[2]), 1)
print(f"{src}: pairs {n} warp-rows {rows}  reset_rows {c[3] / rows:.3f}  event_rows {c[1] / rows:.3f}  iters/row {c[4] / rows:.3f}  windows/row {c[7] / rows:.2f}  clamped windows/row {c[6] / rows:.3f}  "
      f"captures/pair {c[5] / n:.0f}  fallback ticks/pair {c[0] / n:.2f}")
d = np.diff(b.buckets[:, :, :, :p.n_samples], axis=-1)
print("resets per bucket row:", float((d < 0).sum()) / (n * 2 * p.n_buckets), " first sample median", float(np.median(b.buckets[..., 0])),
      " min", float(b.buckets[..., 0].min()))
print(f"warp-rows with a clamp flag {c[8] / rows:.3f}  with a threshold flag {c[9] / rows:.3f}  warp-level event iterations per pair {c[4] / n:.0f}  "
      f"lane-windows per pair {c[7] / n:.0f}  clamped lane-windows per pair {c[6] / n:.0f}")
