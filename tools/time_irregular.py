"""Developer aid: time cg_eval_irregular on grid series spelled with explicit time stamps
(`python tools/time_irregular.py [pairs]`), next to the streaming kernel on the same series."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
p = P.make_params()
le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3)
grid = ph[:, :, None, None].long() + torch.arange(p.row_pitch, device="cuda").long() * p.period_ms
ts_s = grid.expand(n, 2, p.n_scalar, p.row_pitch).to(torch.int32).contiguous()
ts_b = grid.expand(n, 2, p.n_buckets, p.row_pitch).to(torch.int32).contiguous()
cnt_s = torch.full((n, 2, p.n_scalar), p.n_samples, dtype=torch.int32, device="cuda")
cnt_b = torch.full((n, 2, p.n_buckets), p.n_samples, dtype=torch.int32, device="cuda")


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


ms_g, dec_g = timed(lambda: engine.eval_fused(sc, bk, le, ph, p))
bytes_i = n * (p.bytes_per_pair() + 2 * (p.n_scalar + p.n_buckets) * (p.row_pitch * 4 + 4))
for path, what in ((1, "resident tile"), (3, "rows through the ring")):
    ms_i, dec_i = timed(lambda: engine.eval_irregular(sc, bk, ts_s, ts_b, cnt_s, cnt_b, le, p, path=path))
    assert torch.equal(dec_i, dec_g), "irregular and grid kernels disagree on grid series"
    print(f"pairs={n} irregular, {what}: {ms_i:.3f} ms ({n / ms_i * 1e3:.0f} pairs/s, {bytes_i / ms_i / 1e6:.0f} GB/s of values+stamps)")
print(f"streaming kernel on the same series: {ms_g:.3f} ms ({n / ms_g * 1e3:.0f} pairs/s); decisions identical")
