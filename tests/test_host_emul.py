"""CPU test of the streaming kernel's per-lane arithmetic (csrc/cg_warp.cuh).

The lane functions of the CUDA kernel are plain C++ as well; tests/host_emul/warp_emul.cpp drives
them lane by lane on the host with the kernel's own phase order.  Here that emulation is compared
with the C oracle on the same seeded batches the -m gpu parity tests use: decisions and gate bits
bit-exact, metric records bit-exact (required bar: 1e-6 relative).  Every case runs twice: with
each rare path entered only when the lane's own flag asks for it, and with every warp vote forced
to "yes", which proves the rare paths are neutral when entered for nothing.

This checks the arithmetic, not the CUDA plumbing (ring, mbarriers, votes): the -m gpu tests do that.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import params as P, synth
from oracle import c_oracle as co

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host_emul", "warp_emul.cpp")
OUT = os.path.join(HERE, "host_emul", "_build", "libwarp_emul.so")
CSRC = os.path.join(os.path.dirname(HERE), "research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200", "csrc")


def _lib():
    deps = [SRC, os.path.join(CSRC, "cg_warp.cuh"), os.path.join(CSRC, "cg_math.cuh")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC",
                               "-Wno-unknown-pragmas", "-o", OUT, SRC])
    lib = C.CDLL(OUT)
    lib.emul_eval_fused.argtypes = [C.c_void_p] * 6 + [C.c_int64] + [C.c_void_p] * 5 + [C.c_int, C.c_void_p]
    return lib


def emulate(b, p, force_rare=False, has_previous=None, start_tick=None):
    lib = _lib()
    n = b.phase.shape[0]
    dec = np.zeros((n, 6), np.int32)
    bits = np.zeros((n, p.n_words), np.uint32)
    m = np.full((n, 2, p.n_ticks, p.metric_width), -12345.0)
    v = np.full((n, 2, p.n_ticks), 255, np.uint8)
    cnt = np.zeros(10, np.int64)
    ptr = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib.emul_eval_fused(ptr(b.scalars), ptr(b.buckets) if p.n_buckets else None, ptr(b.le) if p.n_buckets else None,
                             ptr(b.phase), ptr(has_previous), ptr(start_tick), n, C.byref(p), ptr(dec), ptr(bits),
                             ptr(m), ptr(v), int(force_rare), ptr(cnt))
    assert rc == 0
    return dec, bits, m, v, cnt


def same_bits(a, b):
    """bit equality with NaN == NaN (and -0.0 == +0.0, as the GPU tests compare)."""
    return np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.where(np.isnan(a), 0.0, a),
                                                                       np.where(np.isnan(b), 0.0, b))


CONFIGS = {
    "cfg2_error_rate_only": dict(make=dict(n_scalar=2, n_buckets=0, checks=P.G_ERR_OK), n=200, gen={}),
    "cfg3_quantiles": dict(make=dict(quantiles=(0.5, 0.95, 0.99), gate_q=1), n=48, gen=dict(resets=False)),
    "cfg4_full_resets": dict(make=dict(), n=96, gen=dict(adversarial=0.1)),
    "k5_feedback_b12": dict(make=dict(n_scalar=5, n_buckets=12, n_samples=100), n=64, gen=dict(adversarial=0.2)),
    "left_open_coincident": dict(make=dict(left_open=True, n_samples=120, n_buckets=16), n=64,
                                 gen=dict(coincident_phase=True, adversarial=0.1)),
    "closed_coincident": dict(make=dict(n_samples=120, n_buckets=16), n=64, gen=dict(coincident_phase=True)),
    "ragged_pitch": dict(make=dict(n_samples=101, row_pitch=104, n_buckets=10), n=33, gen=dict(adversarial=0.1)),
    "single_bucket": dict(make=dict(n_samples=60, n_buckets=1), n=20, gen={}),
    "two_buckets": dict(make=dict(n_samples=60, n_buckets=2), n=20, gen={}),
    "k16_b64": dict(make=dict(n_samples=48, n_scalar=16, n_buckets=64,
                              roles=[1, 3, 4, 8] + [1, 3, 17, 19] * 3), n=12, gen=dict(adversarial=0.2)),
    "odd_tick0": dict(make=dict(n_samples=100, n_buckets=8, tick0_ms=75000), n=24, gen={}),
    "late_tick0_many_ticks": dict(make=dict(n_samples=240, n_buckets=6, tick0_ms=45000, n_ticks=400), n=10, gen={}),
    "two_passes_800_ticks": dict(make=dict(n_samples=600, n_buckets=5, n_ticks=800), n=6, gen=dict(adversarial=0.2)),
    "q_edges": dict(make=dict(n_samples=60, n_buckets=8, quantiles=(0.0, 1.0, -0.5, 1.5), gate_q=1), n=16, gen={}),
    "twelve_samples": dict(make=dict(n_samples=12, n_buckets=4, n_ticks=14), n=17, gen={}),
    "fp32_storage": dict(make=dict(dtype=P.CG_F32, n_samples=120, n_buckets=16), n=48, gen=dict(adversarial=0.1)),
}


@pytest.mark.parametrize("force_rare", [False, True])
@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_lane_arithmetic_matches_oracle(name, force_rare):
    cfg = CONFIGS[name]
    p = P.make_params(**cfg["make"])
    b = synth.generate(p, cfg["n"], seed=0xC0FFEE + sorted(CONFIGS).index(name), **cfg["gen"])
    want_m, want_v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    dec, bits, m, v, cnt = emulate(b, p, force_rare)
    assert np.array_equal(v, want_v), f"{name}: validity bits differ at {np.argwhere(v != want_v)[:5]}"
    assert same_bits(m, want_m), f"{name}: metric records differ at {np.argwhere(~((m == want_m) | (np.isnan(m) & np.isnan(want_m))))[:5]}"
    assert np.array_equal(bits, want_bits), f"{name}: gate bits differ"
    assert np.array_equal(dec, want_dec), f"{name}: decisions differ"


def test_fast_path_carries_the_benchmark_workload():
    """On the benchmark's layout nearly every tick must be decided by the fast path: the exact
    fallback is for edges and pathologies, a fraction of a tick per pair."""
    p = P.make_params()
    b = synth.generate(p, 64, seed=5, adversarial=0.01)
    dec, bits, m, v, cnt = emulate(b, p)
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    assert np.array_equal(dec, want_dec) and np.array_equal(bits, want_bits)
    assert cnt[0] <= 64 * 2, f"{cnt[0]} fallback ticks for 64 pairs"


def test_start_ticks_and_missing_previous_version():
    p = P.make_params(n_samples=120, n_buckets=8)
    b = synth.generate(p, 50, seed=21)
    rng = np.random.default_rng(3)
    st = rng.integers(0, p.n_ticks + 5, size=50).astype(np.int32)
    hp = (rng.random(50) < 0.8).astype(np.uint8)
    want, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, has_previous=hp, start_tick=st)
    dec, *_ = emulate(b, p, has_previous=hp, start_tick=st)
    assert np.array_equal(dec, want)


def test_non_finite_samples_go_to_the_exact_path():
    """NaN and Inf samples (a Prometheus value can be either) must give what the oracle gives."""
    p = P.make_params(n_samples=60, n_buckets=6)
    b = synth.generate(p, 12, seed=77)
    rng = np.random.default_rng(9)
    for i in range(12):
        ver, row, t = rng.integers(0, 2), rng.integers(0, p.n_buckets), rng.integers(0, 60)
        b.buckets[i, ver, row, t] = [np.nan, np.inf, -np.inf][i % 3]
        if i % 2:
            b.scalars[i, ver, rng.integers(0, p.n_scalar), rng.integers(0, 60)] = [np.nan, np.inf][i % 2]
    want_m, want_v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=4)
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=4)
    for force in (False, True):
        dec, bits, m, v, _ = emulate(b, p, force)
        assert np.array_equal(v, want_v) and same_bits(m, want_m)
        assert np.array_equal(bits, want_bits) and np.array_equal(dec, want_dec)
