"""Bridge between the library's SoA layout and the label-indexed series view the
reference's PromQL strings select from.  TEST INFRASTRUCTURE ONLY.

Given one pair of a SynthBatch it builds an oracle.pyoracle.SeriesDB whose metric
names and labels are the ones mlflow_operator.py:367-410 query, so the UNMODIFIED
reference can be run over exactly the numbers the CUDA kernels see.
"""
from __future__ import annotations

import math

import numpy as np

from . import pyoracle

SERVER_COUNT = "seldon_api_executor_server_requests_seconds_count"
CLIENT_SUM = "seldon_api_executor_client_requests_seconds_sum"
CLIENT_COUNT = "seldon_api_executor_client_requests_seconds_count"
CLIENT_BUCKET = "seldon_api_executor_client_requests_seconds_bucket"

# role bits (include/canary_gate.h)
_R_TOTAL, _R_ERR, _R_SUM, _R_CNT, _R_FB = 1, 2, 4, 8, 16


def _le_label(x: float) -> str:
    return "+Inf" if math.isinf(x) else repr(float(x))


def pair_to_db(batch, pair: int, name: str = "m", namespace: str = "ns",
               versions=("2", "1"), left_open: bool = False) -> pyoracle.SeriesDB:
    """versions[0] is the NEW model (array index 0), versions[1] the OLD one."""
    p = batch.params
    db = pyoracle.SeriesDB(left_open=left_open)
    T = p.n_samples
    for ver in range(2):
        base = {"deployment_name": name, "predictor_name": f"v{versions[ver]}",
                "namespace": namespace}
        irregular = hasattr(batch, "ts_scalars")      # synth.IrregularBatch: per-series time stamps and counts
        ts = None if irregular else [int(batch.phase[pair, ver]) + i * p.period_ms for i in range(T)]

        def series(values, ts_arr, cnt_arr, r):
            if not irregular:
                return list(zip(ts, [float(x) for x in values[pair, ver, r, :T]]))
            n = int(cnt_arr[pair, ver, r])
            return list(zip([int(x) for x in ts_arr[pair, ver, r, :n]], [float(x) for x in values[pair, ver, r, :n]]))

        for k in range(p.n_scalar):
            role = p.scalar_role[k]
            samples = series(batch.scalars, getattr(batch, "ts_scalars", None), getattr(batch, "cnt_scalars", None), k)
            if role & _R_TOTAL:
                labels = dict(base, series=str(k), service="predictions",
                              code="500" if role & _R_ERR else "200")
                if role & _R_FB:
                    labels["service"] = "feedback"
                db.add(SERVER_COUNT, labels, samples)
            elif role & _R_SUM:
                db.add(CLIENT_SUM, dict(base, series=str(k)), samples)
            elif role & _R_CNT:
                db.add(CLIENT_COUNT, dict(base, series=str(k)), samples)
            else:
                raise ValueError(f"scalar series {k} has no role the reference queries")
        for b in range(p.n_buckets):
            db.add(CLIENT_BUCKET, dict(base, le=_le_label(float(batch.le[b]))),
                   series(batch.buckets, getattr(batch, "ts_buckets", None), getattr(batch, "cnt_buckets", None), b))
    return db


def metrics_dict_to_record(m: dict, n_q: int = 1):
    """Reference metrics dict (mlflow_operator.py:364-417) -> (record, valid)."""
    rec = np.full(n_q + 5, np.nan)
    valid = 0
    if m["latency_95th"] is not None:
        rec[0] = m["latency_95th"]
        valid |= 1
    rec[n_q + 0] = m["error_responses"]
    if m["error_rate"] is not None:
        rec[n_q + 1] = m["error_rate"]
        valid |= 2
    if m["latency_avg"] is not None:
        rec[n_q + 2] = m["latency_avg"]
        valid |= 4
    rec[n_q + 3] = m["request_count"]
    rec[n_q + 4] = m["feedback_request_count"]
    return rec, valid


def run_reference_on_pair(batch, pair: int, start_tick: int = 0, left_open: bool = False):
    """Run the unmodified reference canary loop on one pair.  Returns the decision
    record the library must reproduce plus the raw trace."""
    from . import ref_loader
    p = batch.params
    db = pair_to_db(batch, pair, left_open=left_open)
    clock = ref_loader.VirtualClock(p.tick0_ms + start_tick * p.tick_step_ms)
    proms = []

    def factory():
        fp = ref_loader.FakePrometheus(db, clock)
        proms.append(fp)
        return fp
    out = ref_loader.run_reference_canary(factory, clock)
    weights = out["weights"]
    n_evals = proms[0].n_queries // 12            # 6 queries x 2 versions (:302-303)
    last = weights[-1]
    if len(last) == 1:                            # old predictor dropped (:355-358)
        cur, prev, state = last[0], 0, 1
    else:
        prev, cur = last
        state = 2                                 # PromotionFailed, weights frozen (:341-349)
    n_prom = sum(1 for w in weights[1:] if len(w) == 2)
    return dict(cur=cur, prev=prev, state=state, n_promotions=n_prom, n_evals=n_evals,
                weights=weights, end_ms=out["end_ms"],
                events=[e["reason"] for e in out["events"]])
