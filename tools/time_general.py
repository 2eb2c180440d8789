"""Developer aid: time the general resident-tile kernel (path 1) next to the streaming kernel
(`python tools/time_general.py [pairs]`)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
p = P.make_params()
le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        out = fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(reps)), out


ms1, d1 = timed(lambda: engine.eval_fused(sc, bk, le, ph, p, path=1))
ms2, d2 = timed(lambda: engine.eval_fused(sc, bk, le, ph, p, path=2))
assert torch.equal(d1, d2)
gb = n * p.bytes_per_pair() / 1e9
print(f"pairs={n} general {ms1:.3f} ms ({gb / ms1 * 1e3:.0f} GB/s) | streaming {ms2:.3f} ms ({gb / ms2 * 1e3:.0f} GB/s); decisions identical")
