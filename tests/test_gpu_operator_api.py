"""-m gpu: the reference-facing Python signatures (get_model_metrics /
should_promote_model) driven by a restatement of the reference's canary loop
(mlflow_operator.py:296-352), checked against the goldens the real loop produced."""
import logging

import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import operator_api as api, params as P
from golden_util import fdec, load_canary_batches, same_float

pytestmark = pytest.mark.gpu
BATCHES, CASES = load_canary_batches()
LOG = logging.getLogger("test.canary")


def run_loop_like_the_reference(store, name, namespace, cur_v, prev_v):
    """The operator's loop (:290-352) with its sleeps applied to the store's clock."""
    thresholds = dict(P.REFERENCE_THRESHOLDS)                       # :175-179
    traffic_prev, traffic_current = 90, 10                          # :186-187
    max_traffic, step, interval, max_attempts, attempt_delay = 100, 10, 60, 10, 10   # :290-294
    n_evals = 0
    while traffic_current < max_traffic:
        attempt, ok = 0, False
        while attempt < max_attempts:
            new = api.get_model_metrics(store, name, f"v{cur_v}", namespace)
            old = api.get_model_metrics(store, name, f"v{prev_v}", namespace)
            n_evals += 1
            if api.should_promote_model(new, old, thresholds, LOG):
                traffic_current = min(traffic_current + step, max_traffic)
                traffic_prev = max(traffic_prev - step, 0)
                ok = True
                break
            attempt += 1
            if attempt < max_attempts:
                store.advance(attempt_delay)
        if not ok:
            break
        store.advance(interval)
    return traffic_current, traffic_prev, n_evals, store.now_ms


def test_drop_in_signatures_reproduce_reference_runs():
    b = BATCHES["small"]
    store = api.SeriesStore.from_host(b)
    for case in CASES:
        if case["geometry"] != "small":
            continue
        i = case["pair"]
        store.register("m", "ns", i, new_version="2", old_version="1")
        # per-tick dicts identical to the reference's
        for t in case["ticks"][:6]:
            store.set_time(b.params.tick0_ms + t["tick"] * b.params.tick_step_ms)
            for key, pred in (("new", "v2"), ("old", "v1")):
                got = api.get_model_metrics(store, "m", pred, "ns")
                assert set(got) == set(t[key])
                for k, v in t[key].items():
                    assert same_float(got[k], fdec(v)), (i, t["tick"], key, k, got[k], fdec(v))
        run = case["runs"][0]
        store.set_time(b.params.tick0_ms)
        cur, prev, n_evals, end_ms = run_loop_like_the_reference(store, "m", "ns", "2", "1")
        if run["state"] == P.STATE_COMPLETE:
            assert (cur, prev) == (100, 0)
        else:
            assert (cur, prev) == (run["cur"], run["prev"])
        assert n_evals == run["n_evals"] and end_ms == run["end_ms"]


def test_unknown_deployment_behaves_like_an_empty_prometheus_answer():
    store = api.SeriesStore.from_host(BATCHES["small"])
    m = api.get_model_metrics(store, "nope", "v1", "ns")
    assert m == {"latency_95th": None, "error_responses": 0.0, "error_rate": None,
                 "latency_avg": None, "request_count": 0.0, "feedback_request_count": 0.0}
    assert api.should_promote_model(m, m, dict(P.REFERENCE_THRESHOLDS), LOG) is False


def test_real_prometheus_client_is_refused():
    with pytest.raises(TypeError, match="SeriesStore"):
        api.get_model_metrics(object(), "m", "v1", "ns")


def test_batched_api_and_patch_batch():
    b = BATCHES["small"]
    store = api.SeriesStore.from_host(b)
    dec = api.evaluate_canaries(store)
    runs = {c["pair"]: c["runs"][0] for c in CASES if c["geometry"] == "small"}
    d = dec.cpu().numpy()
    for i, run in runs.items():
        assert tuple(d[i, :5]) == (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
    patches = api.decisions_to_predictor_traffic(dec, ["2"] * len(d), ["1"] * len(d))
    for i, run in runs.items():
        assert [p["traffic"] for p in patches[i]["predictors"]] == run["weights"][-1]


def test_ingested_prometheus_answers_drive_the_same_decisions():
    """Wire format -> ingest -> SeriesStore -> the drop-in signatures and the batched API."""
    from canary_b200 import ingest
    from test_ingest import render
    b = BATCHES["small"]
    wire, versions, epoch = render(b)
    host = ingest.build_host_batch(wire, versions, period_ms=b.params.period_ms, epoch_ms=epoch)
    store = ingest.to_store(host)
    dec = api.evaluate_canaries(store).cpu().numpy()
    for case in CASES:
        if case["geometry"] != "small":
            continue
        i, run = case["pair"], case["runs"][0]
        assert tuple(dec[i, :5]) == (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
        t = case["ticks"][1]
        store.set_time(host.params.tick0_ms + t["tick"] * host.params.tick_step_ms)
        got = api.get_model_metrics(store, f"model-{i:03d}", "v2", "prod")
        for k, v in t["new"].items():
            assert same_float(got[k], fdec(v)), (i, k)


def test_drop_in_signatures_on_irregular_series_reproduce_reference_runs():
    """Same loop, on scrapes that are not a grid: goldens from the unmodified reference
    (tests/golden/ref_irregular_cases.*)."""
    from golden_util import load_irregular_batches
    batches, cases = load_irregular_batches()
    ib = batches["irr"]
    p = ib.params
    store = api.IrregularSeriesStore.from_host(ib)
    for case in (c for c in cases if c["geometry"] == "irr"):
        i = case["pair"]
        store.register("m", "ns", i, new_version="2", old_version="1")
        for t in case["ticks"][:8]:
            store.set_time(p.tick0_ms + t["tick"] * p.tick_step_ms)
            dicts = {}
            for key, pred in (("new", "v2"), ("old", "v1")):
                got = dicts[key] = api.get_model_metrics(store, "m", pred, "ns")
                for k, v in t[key].items():
                    assert same_float(got[k], fdec(v)), (i, t["tick"], key, k, got[k], fdec(v))
            assert api.should_promote_model(dicts["new"], dicts["old"], dict(P.REFERENCE_THRESHOLDS), LOG) == t["promote"]
        run = case["runs"][0]
        store.set_time(p.tick0_ms)
        cur, prev, n_evals, end_ms = run_loop_like_the_reference(store, "m", "ns", "2", "1")
        if run["state"] == P.STATE_COMPLETE:
            assert (cur, prev) == (100, 0)
        else:
            assert (cur, prev) == (run["cur"], run["prev"])
        assert n_evals == run["n_evals"] and end_ms == run["end_ms"]
    # and the batched call gives the same records as the loop
    dec = api.evaluate_canaries(store).cpu().numpy()
    for case in (c for c in cases if c["geometry"] == "irr"):
        run = case["runs"][0]
        assert tuple(dec[case["pair"], :5]) == (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
