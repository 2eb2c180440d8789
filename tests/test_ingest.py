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
    with pytest.raises(ingest.IrregularSeries):
        ingest.build_host_batch(wire, versions, period_ms=p.period_ms, epoch_ms=epoch)
