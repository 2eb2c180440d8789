"""Developer aid: run the streaming kernel's parity cases on a CG_CHECKED build (our own device-side bounds
checks; compute-sanitizer is not available on this GPU pool) and print the failure counters.
  CG_CHECKED=1 python <package>/build.py --force; cp <package>/libcanarygate.so variants/lib_checked.so
  CG_LIB_PATH=variants/lib_checked.so python tools/run_checked.py"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import canary_b200  # noqa: E402,F401
from canary_b200 import _lib, engine, params as P, synth  # noqa: E402
from oracle import c_oracle as co  # noqa: E402

CASES = {   # fp64, one quantile: the instance that carries the counters
    "cfg4_full_resets": (dict(), 512, dict(adversarial=0.1)),
    "k16_b64": (dict(n_samples=48, n_scalar=16, n_buckets=64, roles=[1, 3, 4, 8] + [1, 3, 17, 19] * 3), 64, dict(adversarial=0.2)),
    "two_passes_800_ticks": (dict(n_samples=600, n_buckets=5, n_ticks=800), 32, dict(adversarial=0.2)),
    "twelve_samples": (dict(n_samples=12, n_buckets=4, n_ticks=14), 64, {}),
    "odd_samples_241": (dict(n_samples=241, row_pitch=242, tick0_ms=75000, n_ticks=355), 128, {}),
    "ragged_pitch": (dict(n_samples=101, row_pitch=104, n_buckets=10), 64, dict(adversarial=0.1)),
    "late_tick0_many_ticks": (dict(n_samples=240, n_buckets=6, tick0_ms=45000, n_ticks=400), 32, {}),
    "left_open_coincident": (dict(left_open=True, n_samples=120, n_buckets=16), 64, dict(coincident_phase=True)),
    "no_buckets": (dict(n_scalar=2, n_buckets=0, checks=P.G_ERR_OK), 256, {}),
}
names = ["fill entry", "fill source range", "scalar row slot", "lane tile", "bucket tile", "promote-bit word", "fallback-bit word", "-"]
lib = _lib.lib
assert hasattr(lib, "cg_debug_check_failures"), "not a CG_CHECKED build (set CG_LIB_PATH)"
buf = (C.c_uint * 8)()
lib.cg_debug_check_failures(buf, 1)
total = 0
for name, (mk, n, gen) in CASES.items():
    p = P.make_params(**mk)
    b = synth.generate(p, n, seed=99, **gen)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    rng = np.random.default_rng(1)
    hp = (rng.random(n) < 0.9).astype(np.uint8)
    dec, bits, m, v = engine.eval_fused(t(b.scalars), t(b.buckets), t(b.le), t(b.phase), p, has_previous=t(hp),
                                        want_bits=True, want_metrics=True, path=2)
    torch.cuda.synchronize()
    want, wbits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, has_previous=hp)
    ok_d = np.array_equal(dec.cpu().numpy(), want)
    ok_b = np.array_equal(bits.cpu().numpy().view(np.uint32)[hp != 0], wbits[hp != 0])   # pairs without a previous version run no canary
    ok = ok_d and ok_b
    if not ok:
        print("   decisions equal:", ok_d, " gate bits equal:", ok_b, " first bad rows:", np.argwhere((dec.cpu().numpy() != want).any(1))[:4].ravel())
    lib.cg_debug_check_failures(buf, 1)
    fails = list(buf)
    total += sum(fails) + (0 if ok else 1)
    print(f"{name:26s} pairs {n:4d}  parity {'ok' if ok else 'FAIL'}  failed checks {dict((k, c) for k, c in zip(names, fails) if c) or 0}")
print("TOTAL failed checks / parity failures:", total)
sys.exit(1 if total else 0)
