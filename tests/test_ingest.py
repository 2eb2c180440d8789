"""Ingest of Prometheus range-query answers into the SoA layout: a synthetic batch is rendered
the way Prometheus would send it (values as strings, times as float seconds, labels the
reference's matchers select on), parsed back, and must reproduce the arrays bit for bit."""
import json
import math

import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import ingest, params as P, synth
from oracle import c_oracle as co


def render(batch, epoch_ms=1_700_000_000_000):
    """SynthBatch -> {metric name: query_range answer}, {(deployment, ns): (new, old)}."""
    p = batch.params
    answers = {ingest.SERVER_COUNT: [], ingest.CLIENT_SUM: [], ingest.CLIENT_COUNT: [], ingest.CLIENT_BUCKET: []}
    versions = {}
    T = p.n_samples
    for i in range(batch.phase.shape[0]):
        key = (f"model-{i:03d}", "prod")
        versions[key] = ("v2", "v1")
        for v, pred in enumerate(("v2", "v1")):
            base = {"deployment_name": key[0], "namespace": key[1], "predictor_name": pred}
            ts = [(epoch_ms + int(batch.phase[i, v]) + k * p.period_ms) / 1000.0 for k in range(T)]

            def series(labels, row):
                return {"metric": labels, "values": [[t, repr(float(x))] for t, x in zip(ts, row[:T])]}
            answers[ingest.SERVER_COUNT].append(series(dict(base, code="200", service="predictions"), batch.scalars[i, v, 0]))
            answers[ingest.SERVER_COUNT].append(series(dict(base, code="500", service="predictions"), batch.scalars[i, v, 1]))
            answers[ingest.CLIENT_SUM].append(series(dict(base), batch.scalars[i, v, 2]))
            answers[ingest.CLIENT_COUNT].append(series(dict(base), batch.scalars[i, v, 3]))
            for b in range(p.n_buckets):
                le = "+Inf" if math.isinf(batch.le[b]) else repr(float(batch.le[b]))
                answers[ingest.CLIENT_BUCKET].append(series(dict(base, le=le), batch.buckets[i, v, b]))
    wire = {name: json.dumps({"status": "success", "data": {"resultType": "matrix", "result": res}})
            for name, res in answers.items()}
    return wire, versions, epoch_ms


def test_round_trip_is_bit_exact_and_decisions_agree():
    p = P.make_params(n_samples=96, n_buckets=10)
    b = synth.generate(p, 6, seed=5, adversarial=0.3)
    wire, versions, epoch = render(b)
    host = ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)
    assert host.params.n_scalar == 4 and host.params.n_buckets == 10 and host.params.n_samples == 96
    assert tuple(host.params.scalar_role[:4]) == tuple(p.scalar_role[:4])     # roles from code / service / name
    assert np.array_equal(host.le, b.le)
    assert np.array_equal(host.phase, b.phase)
    assert np.array_equal(host.scalars[..., :96], b.scalars[..., :96])
    assert np.array_equal(host.buckets[..., :96], b.buckets[..., :96])
    assert host.index[("model-003", "prod")] == (3, {"v2": 0, "v1": 1})
    # same decisions from the ingested arrays as from the original ones
    want, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, want_bits=False)
    got, _ = co.eval_fused(host.scalars, host.buckets, host.le, host.phase, host.params, want_bits=False)
    assert np.array_equal(got, want)


def test_special_values_and_errors():
    ans = {"status": "success", "data": {"resultType": "matrix", "result": [
        {"metric": {"le": "+Inf"}, "values": [[1.0, "NaN"], [16.0, "+Inf"], [31.0, "3"]]}]}}
    (labels, ts, v), = ingest.parse_matrix(ans)
    assert ts.tolist() == [1000, 16000, 31000] and math.isnan(v[0]) and math.isinf(v[1]) and v[2] == 3.0
    with pytest.raises(ValueError, match="Prometheus error"):
        ingest.parse_matrix({"status": "error", "error": "boom", "data": {}})
    p = P.make_params(n_samples=8, n_buckets=3)
    b = synth.generate(p, 1, seed=1)
    wire, versions, epoch = render(b)
    doc = json.loads(wire[ingest.CLIENT_COUNT])
    doc["data"]["result"][0]["values"][3][0] += 1.0          # one late scrape
    wire[ingest.CLIENT_COUNT] = json.dumps(doc)
    with pytest.raises(ValueError):          # the grid ingest refuses (IrregularSeries, or a series is missing)
        ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)


# ---- irregular series ---------------------------------------------
def render_irregular(ib, epoch_ms=1_700_000_000_000):
    p = ib.params
    answers = {ingest.SERVER_COUNT: [], ingest.CLIENT_SUM: [], ingest.CLIENT_COUNT: [], ingest.CLIENT_BUCKET: []}
    versions = {}
    for i in range(ib.cnt_scalars.shape[0]):
        key = (f"model-{i:03d}", "prod")
        versions[key] = ("v2", "v1")
        for v, pred in enumerate(("v2", "v1")):
            base = {"deployment_name": key[0], "namespace": key[1], "predictor_name": pred}

            def series(labels, vals, ts, n):
                return {"metric": labels, "values": [[(epoch_ms + int(t)) / 1000.0, repr(float(x))]
                                                      for t, x in zip(ts[:n], vals[:n])]}
            rows = [(ingest.SERVER_COUNT, dict(base, code="200", service="predictions")),
                    (ingest.SERVER_COUNT, dict(base, code="500", service="predictions")),
                    (ingest.CLIENT_SUM, dict(base)), (ingest.CLIENT_COUNT, dict(base))]
            for k, (name, labels) in enumerate(rows):
                n = int(ib.cnt_scalars[i, v, k])
                if n:       # Prometheus does not return a series without samples in the range
                    answers[name].append(series(labels, ib.scalars[i, v, k], ib.ts_scalars[i, v, k], n))
            for b in range(p.n_buckets):
                n = int(ib.cnt_buckets[i, v, b])
                le = "+Inf" if math.isinf(ib.le[b]) else repr(float(ib.le[b]))
                if n:
                    answers[ingest.CLIENT_BUCKET].append(series(dict(base, le=le), ib.buckets[i, v, b], ib.ts_buckets[i, v, b], n))
    wire = {name: json.dumps({"status": "success", "data": {"resultType": "matrix", "result": res}})
            for name, res in answers.items()}
    return wire, versions, epoch_ms


def test_irregular_round_trip_and_decisions():
    p = P.make_params(n_samples=96, n_buckets=10)
    ib = synth.make_irregular(synth.generate(p, 6, seed=5, adversarial=0.3), seed=2, miss_prob=0.1, gap_prob=0.05,
                              dead_series_prob=0.05)
    wire, versions, epoch = render_irregular(ib)
    with pytest.raises(ValueError):          # the grid ingest refuses (IrregularSeries, or a series is missing)
        ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)
    host = ingest.build_irregular_batch(wire, versions, epoch_ms=epoch)
    hp = host.params
    assert hp.n_scalar == 4 and hp.n_buckets == 10 and hp.row_pitch % 4 == 0
    assert tuple(hp.scalar_role[:4]) == tuple(p.scalar_role[:4]) and np.array_equal(host.le, ib.le)
    assert np.array_equal(host.cnt_scalars, ib.cnt_scalars) and np.array_equal(host.cnt_buckets, ib.cnt_buckets)
    for got_v, got_t, cnt, want_v, want_t in ((host.scalars, host.ts_scalars, ib.cnt_scalars, ib.scalars, ib.ts_scalars),
                                              (host.buckets, host.ts_buckets, ib.cnt_buckets, ib.buckets, ib.ts_buckets)):
        for idx in np.ndindex(cnt.shape):
            n = cnt[idx]
            assert np.array_equal(got_v[idx][:n], want_v[idx][:n]) and np.array_equal(got_t[idx][:n], want_t[idx][:n])
    want = co.eval_irregular(ib.scalars, ib.buckets, ib.ts_scalars, ib.ts_buckets, ib.cnt_scalars, ib.cnt_buckets,
                             ib.le, p)[0]
    got = co.eval_irregular(host.scalars, host.buckets, host.ts_scalars, host.ts_buckets, host.cnt_scalars,
                            host.cnt_buckets, host.le, hp)[0]
    assert np.array_equal(got, want)


def test_stale_markers_are_not_samples():
    stale = np.array([ingest.STALE_NAN_BITS], dtype=np.uint64).view(np.float64)[0]
    ts = np.array([0, 15000, 30000, 45000, 60000])
    v = np.array([1.0, 2.0, stale, float("nan"), 5.0])
    t2, v2 = ingest.drop_stale(ts, v)
    assert list(t2) == [0, 15000, 45000, 60000] and math.isnan(v2[2]) and v2[3] == 5.0   # an ordinary NaN stays


def test_two_label_sets_per_bucket_are_refused_or_summed():
    """Real Seldon executors expose several series per `le` (replicas, extra labels); the reference's
    query adds their rates (`sum(rate(..)) by (le)`, mlflow_operator.py:367).  Silently keeping one
    of them would gate on one replica: the default is a clear error, the opt-in adds the counters."""
    p = P.make_params(n_samples=48, n_buckets=6)
    b = synth.generate(p, 3, seed=9, resets=False)
    wire, versions, epoch = render(b)
    doc = json.loads(wire[ingest.CLIENT_BUCKET])
    extra = []
    for item in doc["data"]["result"]:                       # a second replica with the same counts
        extra.append({"metric": dict(item["metric"], replica="b"), "values": item["values"]})
    doc["data"]["result"] += extra
    wire2 = dict(wire)
    wire2[ingest.CLIENT_BUCKET] = json.dumps(doc)
    with pytest.raises(ingest.DuplicateBucketSeries, match="share le="):
        ingest.build_host_batch(wire2, versions, period_ms=p.period_ms, epoch_ms=epoch)
    host = ingest.build_host_batch(wire2, versions, period_ms=p.period_ms, epoch_ms=epoch, aggregate_buckets=True)
    assert host.params.n_buckets == 6
    assert np.array_equal(host.buckets[..., :48], 2.0 * b.buckets[..., :48])
    with pytest.raises(ingest.DuplicateBucketSeries):
        ingest.build_irregular_batch(wire2, versions, epoch_ms=epoch)
    # a reset in one of the replicas cannot be folded into a summed counter
    doc["data"]["result"][-1]["values"][10][1] = "0"
    wire2[ingest.CLIENT_BUCKET] = json.dumps(doc)
    with pytest.raises(ingest.DuplicateBucketSeries, match="reset"):
        ingest.build_host_batch(wire2, versions, period_ms=p.period_ms, epoch_ms=epoch, aggregate_buckets=True)


def test_raw_series_path_needs_no_strings():
    """Remote-read shaped input (int64 ms + fp64 arrays) gives the same arrays as the JSON answers."""
    p = P.make_params(n_samples=64, n_buckets=5)
    b = synth.generate(p, 4, seed=13)
    wire, versions, epoch = render(b)
    raw = {name: [ingest.raw_series(lbl, ts, v) for lbl, ts, v in ingest.parse_matrix(ans)] for name, ans in wire.items()}
    h1 = ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)
    h2 = ingest.build_host_batch(raw, versions, period_ms=p.period_ms, epoch_ms=epoch)
    assert np.array_equal(h1.scalars, h2.scalars) and np.array_equal(h1.buckets, h2.buckets)
    assert np.array_equal(h1.phase, h2.phase) and np.array_equal(h1.le, h2.le)
