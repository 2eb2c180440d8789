"""-m gpu: the CUDA path (through the C ABI) against the oracle and the golden fixtures
produced by the unmodified reference.  Bars: decisions / weights / gate bits bit-exact;
rate, quantile and ratio values within 1e-6 relative in fp64 (we also demand bit
equality, which the kernels are designed for); fp32 storage: fp64 arithmetic on the
fp32-rounded inputs, same bars against the oracle run on the same rounded inputs."""
import numpy as np
import pytest
import torch

import canary_b200  # noqa: F401
from canary_b200 import engine, params as P, synth
from golden_util import fdec, golden_record, load_canary_batches, load_json, records_equal
from gpu_util import assert_metrics_match, gpu_fused, gpu_fused_metrics, gpu_grid
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def test_native_library_is_loaded():
    from canary_b200 import _lib
    assert _lib.lib.cg_abi_version() == 1
    maps = open("/proc/self/maps").read()
    assert "libcanarygate.so" in maps


# ---------------------------------------------------------------- golden fixtures
BATCHES, CASES = load_canary_batches()


@pytest.mark.parametrize("geom", sorted(BATCHES))
def test_golden_metrics_and_decisions(geom):
    b = BATCHES[geom]
    p = b.params
    m, v = gpu_grid(b)
    for case in CASES:
        if case["geometry"] != geom:
            continue
        i = case["pair"]
        for t in case["ticks"]:
            j = t["tick"]
            for ver, key in enumerate(("new", "old")):
                rec, valid = golden_record(t[key], p.n_q)
                assert valid == v[i, ver, j], (geom, i, j, key)
                assert records_equal(rec, m[i, ver, j]), (geom, i, j, key, rec, m[i, ver, j])
        for run in case["runs"]:
            st = np.full(b.phase.shape[0], run["start_tick"], np.int32)
            dec, bits = gpu_fused(b, start_tick=st)
            want = (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
            assert tuple(dec[i, :5]) == want, (geom, i, run["start_tick"], dec[i])
            for t in case["ticks"]:
                j = t["tick"]
                assert bool((bits[i, j >> 5] >> (j & 31)) & 1) == t["promote"]


def test_golden_gate_probes():
    doc = load_json("ref_gate_probes.json")
    p = P.make_params(thresholds=doc["thresholds"])
    nm, nv, om, ov, want = [], [], [], [], []
    for c in doc["cases"]:
        for src, mm, vv in ((c["new"], nm, nv), (c["old"], om, ov)):
            rec, valid = np.full(6, np.nan), 0
            for key, slot, bit in (("latency_95th", 0, 1), ("error_rate", 2, 2), ("latency_avg", 3, 4)):
                x = fdec(src[key])
                if x is not None:
                    rec[slot], valid = x, valid | bit
            mm.append(rec)
            vv.append(valid)
        want.append(c["promote"])
    t = lambda a, dt: torch.tensor(np.array(a), dtype=dt, device="cuda")  # noqa: E731
    d = engine.gate(t(nm, torch.float64), t(nv, torch.uint8), t(om, torch.float64), t(ov, torch.uint8), p)
    got = [(int(x) & P.G_PROMOTE) != 0 for x in d.cpu()]
    assert got == want
    # and identical detail bytes to the C oracle
    dc = co.gate(np.array(nm), np.array(nv, np.uint8), np.array(om), np.array(ov, np.uint8), p)
    assert np.array_equal(d.cpu().numpy(), dc)


def test_golden_schedule_probes():
    doc = load_json("ref_schedule_probes.json")
    for c in doc["cases"]:
        bits = c["gate_bits"]
        p = P.make_params(n_ticks=len(bits))
        words = np.zeros((1, p.n_words), np.uint32)
        for j, bit in enumerate(bits):
            if bit:
                words[0, j >> 5] |= np.uint32(1 << (j & 31))
        dec = engine.replay(torch.from_numpy(words.view(np.int32)).cuda(), p).cpu().numpy()[0]
        last = c["weights"][-1]
        if len(last) == 1:
            cur, prev, state = last[0], 0, P.STATE_COMPLETE
        else:
            (prev, cur), state = last, P.STATE_ABORTED
        n_prom = sum(1 for w in c["weights"][1:] if len(w) == 2)
        assert tuple(dec) == (cur, prev, state, n_prom, len(c["eval_ticks"]), c["eval_ticks"][-1]), c["name"]


# ---------------------------------------------------------------- seeded batches vs the oracle
CONFIGS = {
    # BASELINE.json configs[1]: error-rate gate only, two server counters, no histogram
    "cfg2_error_rate_only": dict(make=dict(n_scalar=2, n_buckets=0, checks=P.G_ERR_OK), n=1000, gen={}),
    # configs[2]: 64-bucket histograms, p50/p95/p99
    "cfg3_quantiles": dict(make=dict(quantiles=(0.5, 0.95, 0.99), gate_q=1), n=96, gen=dict(resets=False)),
    # configs[3]: full layout with counter resets and adversarial pairs
    "cfg4_full_resets": dict(make=dict(), n=160, gen=dict(adversarial=0.1)),
    "k5_feedback_b12": dict(make=dict(n_scalar=5, n_buckets=12, n_samples=100), n=64, gen=dict(adversarial=0.2)),
    "left_open_coincident": dict(make=dict(left_open=True, n_samples=120, n_buckets=16), n=64,
                                 gen=dict(coincident_phase=True, adversarial=0.1)),
    "closed_coincident": dict(make=dict(n_samples=120, n_buckets=16), n=64, gen=dict(coincident_phase=True)),
    "ragged_pitch": dict(make=dict(n_samples=101, row_pitch=104, n_buckets=10), n=33, gen=dict(adversarial=0.1)),
    "long_window_30s_scrape": dict(make=dict(period_ms=30000, elapsed_time=300, n_samples=120, n_buckets=8,
                                             tick_step_ms=10000), n=40, gen={}),
    # other scrape periods / windows / tick steps: the row-streaming kernel (path 3) is what runs them
    "scrape_10s": dict(make=dict(period_ms=10000, n_samples=200, n_buckets=20), n=48, gen=dict(adversarial=0.1)),
    "scrape_5s_window_30s_tick_20s": dict(make=dict(period_ms=5000, elapsed_time=30, n_samples=300, n_buckets=12,
                                                    tick_step_ms=20000, attempt_delay_s=20), n=32, gen=dict(adversarial=0.1)),
    "scrape_30s_three_quantiles_500_ticks": dict(make=dict(period_ms=30000, elapsed_time=120, n_samples=180, n_buckets=24,
                                                           quantiles=(0.5, 0.9, 0.99), gate_q=2, n_ticks=500), n=24,
                                                 gen=dict(adversarial=0.1)),
    "short_series": dict(make=dict(n_samples=6, n_buckets=4, n_ticks=40), n=17, gen=dict(resets=False)),
    "one_sample": dict(make=dict(n_samples=2, row_pitch=2, n_buckets=2, n_ticks=12), n=5, gen=dict(resets=False)),
    # streaming-kernel corner cases
    "single_bucket": dict(make=dict(n_samples=60, n_buckets=1), n=20, gen={}),
    "two_buckets": dict(make=dict(n_samples=60, n_buckets=2), n=20, gen={}),
    "k16_b64": dict(make=dict(n_samples=48, n_scalar=16, n_buckets=64,
                              roles=[1, 3, 4, 8] + [1, 3, 17, 19] * 3), n=12, gen=dict(adversarial=0.2)),
    "odd_tick0": dict(make=dict(n_samples=100, n_buckets=8, tick0_ms=75000), n=24, gen={}),
    "late_tick0_many_ticks": dict(make=dict(n_samples=240, n_buckets=6, tick0_ms=45000, n_ticks=400), n=10, gen={}),
    "q_edges": dict(make=dict(n_samples=60, n_buckets=8, quantiles=(0.0, 1.0, -0.5, 1.5), gate_q=1), n=16, gen={}),
    "two_passes_800_ticks": dict(make=dict(n_samples=600, n_buckets=5, n_ticks=800), n=6, gen=dict(adversarial=0.2)),
    "twelve_samples": dict(make=dict(n_samples=12, n_buckets=4, n_ticks=14), n=17, gen={}),
    "k1_b3_tiny_grid": dict(make=dict(n_samples=40, n_scalar=1, n_buckets=3, roles=[3], n_ticks=5), n=300, gen={}),
}


# geometries the streaming kernel covers (window = 4 periods, tick step = 2/3 period, >= 12 samples)
STREAMING = {"cfg2_error_rate_only", "cfg3_quantiles", "cfg4_full_resets", "k5_feedback_b12",
             "left_open_coincident", "closed_coincident", "ragged_pitch",
             "single_bucket", "two_buckets", "k16_b64", "odd_tick0", "late_tick0_many_ticks", "q_edges",
             "two_passes_800_ticks", "twelve_samples", "k1_b3_tiny_grid"}


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_seeded_batch_matches_oracle(name):
    cfg = CONFIGS[name]
    p = P.make_params(**cfg["make"])
    b = synth.generate(p, cfg["n"], seed=0xC0FFEE + sorted(CONFIGS).index(name), **cfg["gen"])
    want_m, want_v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    got_m, got_v = gpu_grid(b)
    assert_metrics_match(got_m, want_m, got_v, want_v, rtol=1e-6, exact=True, what=name)
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    paths = [1, 2, 3] if engine.stream_eligible(p) else [1, 3]
    if (p.row_pitch * p.elem_bytes) % 16:
        paths.remove(3)                      # the row-streaming kernel copies whole rows: 16-byte multiples
    assert engine.stream_eligible(p) == (name in STREAMING), name
    for path in paths:
        got_dec, got_bits, m2, v2 = gpu_fused_metrics(b, path=path)
        assert np.array_equal(got_bits, want_bits), f"{name} path {path}: gate bits differ"
        assert np.array_equal(got_dec, want_dec), f"{name} path {path}: decisions differ"
        assert_metrics_match(m2, want_m, v2, want_v, rtol=1e-6, exact=True, what=f"{name} path {path}")


def test_fp32_storage_matches_oracle_on_rounded_inputs():
    p = P.make_params(dtype=P.CG_F32, n_samples=120, n_buckets=16)
    b = synth.generate(p, 64, seed=11, adversarial=0.1)
    assert b.scalars.dtype == np.float32
    want_m, want_v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    got_m, got_v = gpu_grid(b)
    assert_metrics_match(got_m, want_m, got_v, want_v, rtol=1e-6, exact=True, what="fp32")
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=8)
    for path in (1, 2, 3):
        got_dec, got_bits, m2, v2 = gpu_fused_metrics(b, path=path)
        assert np.array_equal(got_bits, want_bits) and np.array_equal(got_dec, want_dec)
        assert_metrics_match(m2, want_m, v2, want_v, rtol=1e-6, exact=True, what=f"fp32 path {path}")


def test_start_ticks_has_previous_and_running_out_of_data():
    p = P.make_params(n_samples=120, n_buckets=8)
    b = synth.generate(p, 50, seed=21)
    rng = np.random.default_rng(3)
    st = rng.integers(0, p.n_ticks + 5, size=50).astype(np.int32)
    hp = (rng.random(50) < 0.8).astype(np.uint8)
    want, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, has_previous=hp, start_tick=st)
    for path in (1, 2, 3):
        got, _ = gpu_fused(b, has_previous=hp, start_tick=st, path=path)
        assert np.array_equal(got, want), path
    assert (got[hp == 0, 2] == P.STATE_NO_CANARY).all()
    assert (got[:, 2] == P.STATE_IN_PROGRESS).any()     # some start too late to finish


def test_empty_batch_is_a_no_op():
    p = P.make_params(n_samples=16, n_buckets=4)
    z = lambda *s, dt=torch.float64: torch.zeros(s, dtype=dt, device="cuda")  # noqa: E731
    dec = engine.eval_fused(z(0, 2, 4, 16), z(0, 2, 4, 16), z(4), z(0, 2, dt=torch.int32), p)
    assert dec.shape == (0, 6)


def test_oversized_geometry_is_rejected_cleanly():
    p = P.make_params(n_samples=16384, n_buckets=64)     # one row = 128 KB: no ring, no resident tile
    z = lambda *s, dt=torch.float64: torch.zeros(s, dtype=dt, device="cuda")  # noqa: E731
    with pytest.raises(RuntimeError, match="shared memory"):
        engine.eval_fused(z(1, 2, 4, 16384), z(1, 2, 64, 16384), z(64), z(1, 2, dt=torch.int32), p)


def test_cpu_tensors_are_refused():
    p = P.make_params(n_samples=16, n_buckets=4)
    with pytest.raises(ValueError, match="CUDA tensor"):
        engine.eval_grid(torch.zeros(1, 2, 4, 16, dtype=torch.float64), torch.zeros(1, 2, 4, 16, dtype=torch.float64),
                         torch.zeros(4, dtype=torch.float64), torch.zeros(1, 2, dtype=torch.int32), p)


# ---------------------------------------------------------------- size-independent properties
def test_properties_at_scale():
    """Larger batch than the oracle is asked to chew: (1) fused decisions == replay of the
    fused gate bits == replay of gate(metrics); (2) permuting pairs permutes decisions;
    (3) an oracle spot check on a random subsample."""
    p = P.make_params()
    n = 4096
    b = synth.generate(p, 256, seed=5, adversarial=0.05)
    reps = n // 256
    tile = lambda a: np.ascontiguousarray(np.concatenate([a] * reps, axis=0))  # noqa: E731
    big = synth.SynthBatch(tile(b.scalars), tile(b.buckets), b.le, tile(b.phase), p)
    # perturb the copies so they are distinct pairs: shift phases
    big.phase[:] = (big.phase + (np.arange(n)[:, None] // 256) * 37) % p.period_ms
    big.phase[big.phase % 5000 == 0] += 1
    dec, bits = gpu_fused(big)
    d2 = engine.replay(torch.from_numpy(bits.view(np.int32)).cuda(), p).cpu().numpy()
    assert np.array_equal(dec, d2)
    perm = np.random.default_rng(0).permutation(n)
    shuf = synth.SynthBatch(np.ascontiguousarray(big.scalars[perm]), np.ascontiguousarray(big.buckets[perm]),
                            b.le, np.ascontiguousarray(big.phase[perm]), p)
    dec_s, _ = gpu_fused(shuf)
    assert np.array_equal(dec_s, dec[perm])
    idx = np.sort(np.random.default_rng(1).choice(n, 64, replace=False))
    sub = synth.SynthBatch(np.ascontiguousarray(big.scalars[idx]), np.ascontiguousarray(big.buckets[idx]),
                           b.le, np.ascontiguousarray(big.phase[idx]), p)
    want, wbits = co.eval_fused(sub.scalars, sub.buckets, sub.le, sub.phase, p, n_threads=8)
    assert np.array_equal(dec[idx], want) and np.array_equal(bits[idx], wbits)


def test_device_generated_series_match_oracle():
    """The benchmark's on-device generator (cg_synth_fill): whatever it writes, the CUDA
    path and the oracle must agree on it; and shards must be consistent."""
    p = P.make_params()
    le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
    sc, bk, ph = engine.synth_fill(192, p, le, seed=77, adversarial=0.1)
    sc2, bk2, ph2 = engine.synth_fill(64, p, le, seed=77, pair_offset=128, adversarial=0.1)
    assert torch.equal(sc[128:], sc2) and torch.equal(bk[128:], bk2) and torch.equal(ph[128:], ph2)
    dec, bits = engine.eval_fused(sc, bk, le, ph, p, want_bits=True)
    m, v = engine.eval_grid(sc, bk, le, ph, p)
    h = [t.cpu().numpy() for t in (sc, bk, ph)]
    want_m, want_v = co.eval_grid(h[0], h[1], le.cpu().numpy(), h[2], p, n_threads=8)
    assert_metrics_match(m.cpu().numpy(), want_m, v.cpu().numpy(), want_v, what="synth")
    want, wbits = co.eval_fused(h[0], h[1], le.cpu().numpy(), h[2], p, n_threads=8)
    assert np.array_equal(dec.cpu().numpy(), want)
    assert np.array_equal(bits.cpu().numpy().view(np.uint32), wbits)
    states = np.bincount(want[:, 2], minlength=4)
    assert states[P.STATE_COMPLETE] > 0 and states[P.STATE_ABORTED] > 0   # both outcomes occur
    # counters really do reset in this workload
    assert (np.diff(h[1][:, :, :, :p.n_samples], axis=-1) < 0).any()


def test_host_pipeline_matches_device_path():
    p = P.make_params(n_samples=120, n_buckets=16)
    b = synth.generate(p, 300, seed=31)
    want, _ = gpu_fused(b)
    pipe = engine.HostPipeline(p, chunk_pairs=64)
    le = torch.from_numpy(b.le).cuda()
    pin = lambda a: torch.from_numpy(a).pin_memory()  # noqa: E731
    for _ in range(2):
        out = pipe.run(pin(b.scalars), pin(b.buckets), pin(b.phase), le)
        assert np.array_equal(out.numpy(), want)
    assert pipe.h2d_bytes == b.scalars.nbytes + b.buckets.nbytes + b.phase.nbytes
    assert pipe.d2h_bytes == 300 * 24


def test_metric_values_at_scale_8192_pairs():
    """Metric VALUES (not only decisions) on 8,192 pairs of BASELINE.json configs[3], from the device
    generator: every rate / quantile / ratio of the streaming kernel's dense grid against the C oracle,
    bit for bit (required bar 1e-6 relative), plus validity bytes, gate bits and decisions."""
    p = P.make_params()
    n = 8192
    le = torch.from_numpy(synth.default_le(p.n_buckets)).cuda()
    sc, bk, ph = engine.synth_fill(n, p, le, seed=0xC0FFEE + 3, adversarial=0.01)
    dec, bits, m, v = engine.eval_fused(sc, bk, le, ph, p, want_bits=True, want_metrics=True, path=2)
    torch.cuda.synchronize()
    h = [t.cpu().numpy() for t in (sc, bk, ph)]
    le_h = le.cpu().numpy()
    want_m, want_v = co.eval_grid(h[0], h[1], le_h, h[2], p, n_threads=co.max_threads())
    assert_metrics_match(m.cpu().numpy(), want_m, v.cpu().numpy(), want_v, rtol=1e-6, exact=True, what="8192 pairs")
    want_dec, want_bits = co.eval_fused(h[0], h[1], le_h, h[2], p, n_threads=co.max_threads())
    assert np.array_equal(bits.cpu().numpy().view(np.uint32), want_bits)
    assert np.array_equal(dec.cpu().numpy(), want_dec)
