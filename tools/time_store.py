"""Developer aid: where the time of one RollingStore step goes (python tools/time_store.py [pairs] [slack])."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import engine, params as P, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
slack = int(sys.argv[2]) if len(sys.argv) > 2 else 64
p = P.make_params()
le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
p_long = P.make_params(n_samples=p.n_samples + 16, n_ticks=p.n_ticks)
sc, bk, ph = engine.synth_fill(n, p_long, le, seed=0xC0FFEE + 3)
store = engine.RollingStore(p, n, le, slack=slack)
store.load(sc[..., :240], bk[..., :240], ph)
nxt = iter(range(240, 256))
if os.environ.get('ALIGN'):
    store.align = int(os.environ['ALIGN'])
fs = sc[..., 240].contiguous().cpu().pin_memory()
fb = bk[..., 240].contiguous().cpu().pin_memory()
dec = torch.empty((n, 6), dtype=torch.int32, device="cuda")


def timed(fn, reps=4):
    fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))


print(f"pairs={n} slack={slack} pitch={store.cap}")
print("evaluate        ms", timed(lambda: store.evaluate(out=dec)))
store.append_scrape(fs, fb)
print("append (pinned) done, count now", store.count)
print("evaluate        ms", timed(lambda: store.evaluate(out=dec)), " view", store.view()[:2])
for _ in range(4):
    j = store.count
    store.append_scrape(sc[..., j].contiguous(), bk[..., j].contiguous())
    print("evaluate        ms", timed(lambda: store.evaluate(out=dec)), " view", store.view()[:2], "count", store.count)
