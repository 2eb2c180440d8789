"""-m gpu: cg_eval_irregular (per-series time stamps) through the C ABI against
the reference-made golden fixtures, the C oracle, and the grid kernels on grid data."""
import numpy as np
import pytest
import torch

import canary_b200  # noqa: F401
from canary_b200 import engine, params as P, synth
from golden_util import golden_record, load_irregular_batches, records_equal
from gpu_util import assert_metrics_match, gpu_fused_metrics
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def gpu_irregular(ib, p, start_tick=None, has_previous=None, path=0):
    t = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    dec, bits, m, v = engine.eval_irregular(t(ib.scalars), t(ib.buckets), t(ib.ts_scalars), t(ib.ts_buckets),
                                            t(ib.cnt_scalars), t(ib.cnt_buckets), t(ib.le), p,
                                            has_previous=t(has_previous), start_tick=t(start_tick),
                                            want_bits=True, want_metrics=True, path=path)
    torch.cuda.synchronize()
    return dec.cpu().numpy(), bits.cpu().numpy().view(np.uint32), m.cpu().numpy(), v.cpu().numpy()


def test_golden_irregular_metrics_and_decisions():
    batches, cases = load_irregular_batches()
    for g, ib in batches.items():
        p = ib.params
        runs = {0: gpu_irregular(ib, p)}
        dec, bits, m, v = runs[0]
        for case in (c for c in cases if c["geometry"] == g):
            i = case["pair"]
            for t in case["ticks"]:
                j = t["tick"]
                for ver, key in enumerate(("new", "old")):
                    rec, valid = golden_record(t[key], p.n_q)
                    assert valid == v[i, ver, j], (g, i, j, key)
                    assert records_equal(rec, m[i, ver, j]), (g, i, j, key, rec, m[i, ver, j])
                assert bool((bits[i, j >> 5] >> (j & 31)) & 1) == t["promote"]
            for run in case["runs"]:
                st = run["start_tick"]
                if st not in runs:
                    runs[st] = gpu_irregular(ib, p, start_tick=np.full(ib.cnt_scalars.shape[0], st, np.int32))
                want = (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
                assert tuple(runs[st][0][i, :5]) == want, (g, i, st)


CONFIGS = {
    "full_layout": dict(make=dict(), n=24, irr=dict(seed=3)),
    "many_gaps": dict(make=dict(n_samples=120, n_buckets=16), n=48,
                      irr=dict(seed=4, miss_prob=0.15, gap_prob=0.08, truncate_prob=0.2, dead_series_prob=0.05)),
    "left_open_edges": dict(make=dict(n_samples=100, n_buckets=8, left_open=True), n=32,
                            irr=dict(seed=5, jitter_ms=0, miss_prob=0.1), gen=dict(coincident_phase=True)),
    "closed_edges": dict(make=dict(n_samples=100, n_buckets=8), n=32,
                         irr=dict(seed=6, jitter_ms=0, miss_prob=0.1), gen=dict(coincident_phase=True)),
    "three_quantiles_k5": dict(make=dict(n_samples=80, n_scalar=5, n_buckets=12, quantiles=(0.5, 0.95, 0.99), gate_q=2),
                               n=32, irr=dict(seed=7)),
    "fp32": dict(make=dict(dtype=P.CG_F32, n_samples=120, n_buckets=16), n=32, irr=dict(seed=8)),
    # per-series misses: the +Inf bucket often has a shorter window than the buckets below it, so the
    # monotonic fix raises the observations (the kernel's one-pass speculation fails and it re-scans)
    "series_level_misses": dict(make=dict(n_samples=80, n_buckets=6), n=32,
                                irr=dict(seed=10, series_miss_prob=0.25, miss_prob=0.0, gap_prob=0.0)),
    "long_window": dict(make=dict(period_ms=30000, elapsed_time=300, n_samples=120, n_buckets=8), n=24,
                        irr=dict(seed=9, jitter_ms=5000)),
}


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_irregular_batch_matches_oracle(name):
    cfg = CONFIGS[name]
    p = P.make_params(**cfg["make"])
    grid = synth.generate(p, cfg["n"], seed=77, adversarial=0.2, **cfg.get("gen", {}))
    ib = synth.make_irregular(grid, **cfg["irr"])
    rng = np.random.default_rng(1)
    st = rng.integers(0, 20, size=cfg["n"]).astype(np.int32)
    hp = (rng.random(cfg["n"]) < 0.9).astype(np.uint8)
    want_dec, want_bits, want_m, want_v = co.eval_irregular(
        ib.scalars, ib.buckets, ib.ts_scalars, ib.ts_buckets, ib.cnt_scalars, ib.cnt_buckets, ib.le, p,
        has_previous=hp, start_tick=st, n_threads=8, want_metrics=True)
    for path in (0, 1, 3):                 # automatic, resident tile, rows through the ring
        dec, bits, m, v = gpu_irregular(ib, p, start_tick=st, has_previous=hp, path=path)
        assert_metrics_match(m, want_m, v, want_v, rtol=1e-6, exact=True, what=f"{name} path {path}")
        assert np.array_equal(bits, want_bits), (name, path)
        assert np.array_equal(dec, want_dec), (name, path)


def test_grid_series_through_the_irregular_kernel_equal_the_grid_kernels():
    p = P.make_params(n_samples=120, n_buckets=16)
    b = synth.generate(p, 40, seed=12, adversarial=0.2)
    want_dec, want_bits, want_m, want_v = gpu_fused_metrics(b)
    for path in (1, 3):
        dec, bits, m, v = gpu_irregular(synth.as_irregular(b), p, path=path)
        assert_metrics_match(m, want_m, v, want_v, rtol=1e-6, exact=True, what=f"grid via irregular path {path}")
        assert np.array_equal(bits, want_bits) and np.array_equal(dec, want_dec)


def test_irregular_rejects_what_it_cannot_run():
    p = P.make_params(n_samples=6, row_pitch=6, n_buckets=2, n_ticks=4)         # pitch not a multiple of 4
    z = lambda *s, dt=torch.float64: torch.zeros(s, dtype=dt, device="cuda")    # noqa: E731
    with pytest.raises(Exception, match="multiple of 4"):
        engine.eval_irregular(z(1, 2, 4, 6), z(1, 2, 2, 6), z(1, 2, 4, 6, dt=torch.int32), z(1, 2, 2, 6, dt=torch.int32),
                              z(1, 2, 4, dt=torch.int32), z(1, 2, 2, dt=torch.int32),
                              torch.tensor([1.0, float("inf")], dtype=torch.float64, device="cuda"), p)
    big = P.make_params(n_samples=480, n_buckets=64)           # values + stamps of a version exceed shared memory
    args = (z(1, 2, 4, 480), z(1, 2, 64, 480), z(1, 2, 4, 480, dt=torch.int32),
            z(1, 2, 64, 480, dt=torch.int32), z(1, 2, 4, dt=torch.int32), z(1, 2, 64, dt=torch.int32),
            torch.cat([torch.arange(63, dtype=torch.float64), torch.tensor([float("inf")], dtype=torch.float64)]).cuda(), big)
    with pytest.raises(Exception, match="shared"):
        engine.eval_irregular(*args, path=1)                   # the resident tile cannot hold it ...
    dec = engine.eval_irregular(*args)                         # ... the rows through the ring can
    assert dec.shape == (1, 6)


def test_layout_too_large_for_a_resident_tile_matches_the_oracle():
    """2 hours of scrapes x 68 series: 392 KB of values and stamps per (pair, version); only the
    row-streaming kernel runs it."""
    p = P.make_params(n_samples=480, n_buckets=64)
    grid = synth.generate(p, 6, seed=5, adversarial=0.2)
    ib = synth.make_irregular(grid, seed=3, jitter_ms=2000, miss_prob=0.02)
    want_dec, want_bits, want_m, want_v = co.eval_irregular(
        ib.scalars, ib.buckets, ib.ts_scalars, ib.ts_buckets, ib.cnt_scalars, ib.cnt_buckets, ib.le, p,
        n_threads=8, want_metrics=True)
    dec, bits, m, v = gpu_irregular(ib, p)
    assert_metrics_match(m, want_m, v, want_v, rtol=1e-6, exact=True, what="large layout")
    assert np.array_equal(bits, want_bits) and np.array_equal(dec, want_dec)


def test_ingested_irregular_answers_give_the_oracle_decisions():
    from canary_b200 import ingest
    from test_ingest import render_irregular
    p = P.make_params(n_samples=96, n_buckets=10)
    ib = synth.make_irregular(synth.generate(p, 12, seed=5, adversarial=0.3), seed=2, miss_prob=0.1, gap_prob=0.05,
                              dead_series_prob=0.05)
    wire, versions, epoch = render_irregular(ib)
    host = ingest.build_irregular_batch(wire, versions, epoch_ms=epoch)
    dec = ingest.evaluate_irregular(host).cpu().numpy()
    want = co.eval_irregular(ib.scalars, ib.buckets, ib.ts_scalars, ib.ts_buckets, ib.cnt_scalars, ib.cnt_buckets,
                             ib.le, p)[0]
    assert np.array_equal(dec, want)
