"""Developer aid: fixed irregular workload (grid series spelled with explicit time stamps) for ncu captures:
`python tools/profile_irregular.py [pairs] [path] [reps]`."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
path = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
p = P.make_params()
le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3)
grid = ph[:, :, None, None].long() + torch.arange(p.row_pitch, device="cuda").long() * p.period_ms
ts_s = grid.expand(n, 2, p.n_scalar, p.row_pitch).to(torch.int32).contiguous()
ts_b = grid.expand(n, 2, p.n_buckets, p.row_pitch).to(torch.int32).contiguous()
cnt_s = torch.full((n, 2, p.n_scalar), p.n_samples, dtype=torch.int32, device="cuda")
cnt_b = torch.full((n, 2, p.n_buckets), p.n_samples, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
for _ in range(reps):
    dec = engine.eval_irregular(sc, bk, ts_s, ts_b, cnt_s, cnt_b, le, p, path=path)
torch.cuda.synchronize()
print("done", torch.bincount(dec[:, 2], minlength=4).tolist())
