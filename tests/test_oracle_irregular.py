"""CPU: the C oracle for irregular series against (a) the grid oracle on
grid data spelled with explicit time stamps, (b) the Appendix B known answers, and - where
/root/reference exists - (c) the UNMODIFIED reference running over oracle/pyoracle.py."""
import math

import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import params as P, synth
from golden_util import golden_record, load_irregular_batches, records_equal
from oracle import c_oracle as co, ref_loader as rl, soa_bridge as sb


def _irr(ib, p, **kw):
    return co.eval_irregular(ib.scalars, ib.buckets, ib.ts_scalars, ib.ts_buckets, ib.cnt_scalars, ib.cnt_buckets,
                             ib.le, p, **kw)


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("make", [dict(n_samples=120, n_buckets=16), dict(n_samples=100, n_buckets=8, left_open=True),
                                  dict(n_samples=60, n_scalar=5, n_buckets=4, quantiles=(0.5, 0.99), gate_q=1)])
def test_grid_series_with_explicit_time_stamps_change_nothing(make):
    p = P.make_params(**make)
    b = synth.generate(p, 24, seed=5, adversarial=0.2, coincident_phase=bool(make.get("left_open")))
    want_dec, want_bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, n_threads=4)
    want_m, want_v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=4)
    dec, bits, m, v = _irr(synth.as_irregular(b), p, n_threads=4, want_metrics=True)
    assert _same(m, want_m) and np.array_equal(v, want_v)
    assert np.array_equal(bits, want_bits) and np.array_equal(dec, want_dec)


def test_known_answers_appendix_b():
    # the series is both "total" and read back through request_count-like slots: use the raw record
    p = P.make_params(n_samples=8, row_pitch=8, n_scalar=1, n_buckets=0, roles=[P.ROLE_CLIENT_COUNT],
                      tick0_ms=60000, n_ticks=1, checks=0)
    cases = [([7, 22, 37, 52], [100, 110, 120, 130], 40.0),                      # K1
             ([7, 22, 37, 52], [0, 10, 20, 30], 35.333333333333336),           # K2 zero-point clamp
             ([7, 22, 37, 52], [100, 110, 5, 15], 33.33333333333333),          # K3 one reset
             ([20, 35, 50], [100, 110, 120], 31.666666666666664)]              # K4 dStart >= thr -> avg/2
    for ts_s, vals, want in cases:
        sc = np.zeros((1, 2, 1, 8)); ts = np.zeros((1, 2, 1, 8), np.int32); cnt = np.zeros((1, 2, 1), np.int32)
        n = len(vals)
        sc[0, :, 0, :n] = vals
        ts[0, :, 0, :n] = [x * 1000 for x in ts_s]
        cnt[:] = n
        _, _, m, v = co.eval_irregular(sc, None, ts, None, cnt, None, None, p, want_metrics=True)
        assert m[0, 0, 0, p.n_q + 3] == want, (ts_s, vals, m[0, 0, 0])         # request_count = sum(increase(count))
    # K5: one sample in the window -> no output -> `or on() vector(0)` -> 0.0
    sc = np.zeros((1, 2, 1, 8)); ts = np.zeros((1, 2, 1, 8), np.int32); cnt = np.ones((1, 2, 1), np.int32)
    sc[0, :, 0, 0] = 5.0; ts[0, :, 0, 0] = 30000
    _, _, m, v = co.eval_irregular(sc, None, ts, None, cnt, None, None, p, want_metrics=True)
    assert m[0, 0, 0, p.n_q + 3] == 0.0 and not (v[0, 0, 0] & 4)


def test_missing_buckets_drop_out_of_the_quantile_vector():
    p = P.make_params(n_samples=8, row_pitch=8, n_scalar=1, n_buckets=4, roles=[P.ROLE_CLIENT_COUNT],
                      tick0_ms=60000, n_ticks=1, quantiles=(0.5,), checks=0)
    le = np.array([0.1, 0.5, 1.0, np.inf])
    ts_row = np.array([7000, 22000, 37000, 52000], np.int32)
    sc = np.zeros((1, 2, 1, 8)); tsc = np.zeros((1, 2, 1, 8), np.int32); csc = np.full((1, 2, 1), 4, np.int32)
    sc[0, :, 0, :4] = [0, 10, 20, 30]; tsc[0, :, 0, :4] = ts_row
    bk = np.zeros((1, 2, 4, 8)); tbk = np.zeros((1, 2, 4, 8), np.int32); cbk = np.full((1, 2, 4), 4, np.int32)
    for b, slope in enumerate([2.0, 6.0, 9.0, 10.0]):
        bk[0, :, b, :4] = 100 + slope * np.arange(4)
        tbk[0, :, b, :4] = ts_row
    full = co.eval_irregular(sc, bk, tsc, tbk, csc, cbk, le, p, want_metrics=True)[2][0, 0, 0, 0]
    # bucket le=0.5 has one sample in the window: histogram_quantile sees [0.1, 1.0, +Inf] only
    cbk2 = cbk.copy(); cbk2[0, :, 1] = 1
    less = co.eval_irregular(sc, bk, tsc, tbk, csc, cbk2, le, p, want_metrics=True)[2][0, 0, 0, 0]
    le3 = np.array([0.1, 1.0, np.inf]); bk3 = np.ascontiguousarray(bk[:, :, [0, 2, 3]])
    p3 = P.make_params(n_samples=8, row_pitch=8, n_scalar=1, n_buckets=3, roles=[P.ROLE_CLIENT_COUNT],
                       tick0_ms=60000, n_ticks=1, quantiles=(0.5,), checks=0)
    want = co.eval_irregular(sc, bk3, tsc, np.ascontiguousarray(tbk[:, :, [0, 2, 3]]), csc,
                             np.ascontiguousarray(cbk[:, :, [0, 2, 3]]), le3, p3, want_metrics=True)[2][0, 0, 0, 0]
    assert less == want and less != full
    # the +Inf bucket missing -> NaN (highest bound is not +Inf); every bucket missing -> empty vector -> None
    cbk3 = cbk.copy(); cbk3[0, :, 3] = 0
    _, _, m, v = co.eval_irregular(sc, bk, tsc, tbk, csc, cbk3, le, p, want_metrics=True)
    assert math.isnan(m[0, 0, 0, 0]) and (v[0, 0, 0] & 1)
    _, _, m, v = co.eval_irregular(sc, bk, tsc, tbk, csc, np.zeros_like(cbk), le, p, want_metrics=True)
    assert math.isnan(m[0, 0, 0, 0]) and not (v[0, 0, 0] & 1)


@pytest.mark.skipif(not rl.reference_available(), reason="no /root/reference here")
@pytest.mark.parametrize("left_open", [False, True])
def test_c_oracle_equals_reference_on_irregular_pairs(left_open):
    p = P.make_params(n_samples=120, n_buckets=8, left_open=left_open)
    b = synth.generate(p, 8, seed=41, adversarial=0.3, coincident_phase=left_open)
    ib = synth.make_irregular(b, seed=7, miss_prob=0.08, gap_prob=0.03, truncate_prob=0.1, dead_series_prob=0.03)
    dec, _, m, v = _irr(ib, p, want_metrics=True)
    for i in range(8):
        r = sb.run_reference_on_pair(ib, i, left_open=left_open)
        assert (r["cur"], r["prev"], r["state"], r["n_promotions"], r["n_evals"]) == tuple(dec[i, :5]), i
    for i in range(3):
        db = sb.pair_to_db(ib, i, left_open=left_open)
        for j in range(0, p.n_ticks, 7):
            fp = rl.FakePrometheus(db, rl.VirtualClock(p.tick0_ms + j * p.tick_step_ms))
            for ver, pred in enumerate(("v2", "v1")):
                rec, valid = sb.metrics_dict_to_record(rl.reference_get_model_metrics(fp, "m", pred, "ns"))
                assert valid == v[i, ver, j], (i, ver, j)
                for a, c in zip(rec, m[i, ver, j]):
                    assert (math.isnan(a) and math.isnan(c)) or a == c, (i, ver, j, rec, m[i, ver, j])


def test_c_oracle_equals_reference_golden_on_irregular_pairs():
    """tests/golden/ref_irregular_cases.*: answers of the unmodified reference, made in the build container."""
    batches, cases = load_irregular_batches()
    for g, ib in batches.items():
        p = ib.params
        dec0, bits, m, v = _irr(ib, p, want_metrics=True)
        by_start = {0: dec0}
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
                if st not in by_start:
                    by_start[st] = _irr(ib, p, start_tick=np.full(ib.cnt_scalars.shape[0], st, np.int32))[0]
                want = (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
                assert tuple(by_start[st][i, :5]) == want, (g, i, st)
