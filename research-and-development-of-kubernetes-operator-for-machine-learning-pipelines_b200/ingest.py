"""Ingest: Prometheus wire format -> the library's SoA layout.

The reference never sees raw samples: it sends six instant queries per (version, attempt) and
Prometheus does the arithmetic (mlflow_operator.py:367-411).  The B200 path needs the samples
themselves, so the step in front of it is one `/api/v1/query_range`-shaped matrix per metric
name; this module turns those answers into `scalars / buckets / le / phase` arrays plus the
label index `SeriesStore` uses in place of Prometheus' own.

Label rules are the reference's matchers:
  seldon_api_executor_server_requests_seconds_count   every series counts toward Q3 (:383);
      code != "200" also toward Q2 (:375); service == "feedback" also toward Q6 (:410)
  seldon_api_executor_client_requests_seconds_sum / _count   Q4 / Q5 (:393, :397)
  seldon_api_executor_client_requests_seconds_bucket  one series per `le`            (:367)
all selected by (deployment_name, predictor_name, namespace).

`build_host_batch` accepts regular scrape grids only (sample i at phase + i*period) and raises
`IrregularSeries` otherwise; `build_irregular_batch` takes any series: it keeps
each series' own time stamps, drops stale markers the way Prometheus drops them from range vectors,
and leaves absent series as rows with zero samples.  The device side of that is cg_eval_irregular.
"""
from __future__ import annotations

import json
import math
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .params import (CgParams, ROLE_CLIENT_COUNT, ROLE_CLIENT_SUM, ROLE_FEEDBACK, ROLE_SERVER_ERROR,
                     ROLE_SERVER_TOTAL, make_params)

SERVER_COUNT = "seldon_api_executor_server_requests_seconds_count"
CLIENT_SUM = "seldon_api_executor_client_requests_seconds_sum"
CLIENT_COUNT = "seldon_api_executor_client_requests_seconds_count"
CLIENT_BUCKET = "seldon_api_executor_client_requests_seconds_bucket"


class IrregularSeries(ValueError):
    pass


def parse_matrix(answer) -> List[Tuple[Dict[str, str], np.ndarray, np.ndarray]]:
    """One query_range answer (dict or JSON text) -> [(labels, ts_ms int64[], values fp64[])].
    Prometheus sends sample values as strings ("NaN", "+Inf" included) and times as float seconds;
    both columns are converted with one vectorised numpy pass per series.  A list of
    (labels, ts_ms, values) tuples -- `raw_series` output, what a remote-read client holds --
    passes through untouched."""
    if isinstance(answer, (str, bytes)):
        answer = json.loads(answer)
    if isinstance(answer, dict) and "data" in answer:
        if answer.get("status", "success") != "success":
            raise ValueError(f"Prometheus error: {answer.get('error')}")
        answer = answer["data"]
    if isinstance(answer, dict):
        if answer.get("resultType", "matrix") != "matrix":
            raise ValueError("expected a range-query matrix")
        answer = answer["result"]
    out = []
    for item in answer:
        if isinstance(item, tuple):
            labels, ts, v = item
            out.append((dict(labels), np.asarray(ts, dtype=np.int64), np.asarray(v, dtype=np.float64)))
            continue
        vals = item["values"]
        if len(vals) == 0:
            out.append((dict(item["metric"]), np.zeros(0, np.int64), np.zeros(0, np.float64)))
            continue
        cols = np.asarray(vals, dtype=object)
        ts = np.rint(cols[:, 0].astype(np.float64) * 1000.0).astype(np.int64)
        v = cols[:, 1].astype(str).astype(np.float64)          # numpy parses "NaN", "+Inf", "-Inf" as Prometheus writes them
        out.append((dict(item["metric"]), ts, v))
    return out


def raw_series(labels: Dict[str, str], ts_ms, values) -> Tuple[Dict[str, str], np.ndarray, np.ndarray]:
    """One series as a remote-read client holds it: int64 ms time stamps and fp64 values, no strings.
    Stale markers keep their bit pattern (drop_stale removes them in build_irregular_batch); put such
    tuples in a list where a query_range answer would go."""
    return dict(labels), np.ascontiguousarray(ts_ms, dtype=np.int64), np.ascontiguousarray(values, dtype=np.float64)


def _scalar_role(name: str, labels: Dict[str, str]) -> int:
    if name == SERVER_COUNT:
        role = ROLE_SERVER_TOTAL
        if labels.get("code", "") != "200":
            role |= ROLE_SERVER_ERROR
        if labels.get("service", "") == "feedback":
            role |= ROLE_FEEDBACK
        return role
    if name == CLIENT_SUM:
        return ROLE_CLIENT_SUM
    if name == CLIENT_COUNT:
        return ROLE_CLIENT_COUNT
    raise KeyError(name)


def _le_value(text: str) -> float:
    return math.inf if text in ("+Inf", "Inf", "inf") else float(text)


class HostSeries:
    """What `build_host_batch` returns: the numpy SoA arrays, params and the label index."""

    def __init__(self, scalars, buckets, le, phase, params, index, epoch_ms):
        self.scalars, self.buckets, self.le, self.phase = scalars, buckets, le, phase
        self.params, self.index, self.epoch_ms = params, index, epoch_ms


class DuplicateBucketSeries(ValueError):
    pass


def _merge_bucket_series(group, what: str, aggregate: bool):
    """`sum(rate(..._bucket[r])) by (le)` (mlflow_operator.py:367) adds the RATES of every series
    that shares an `le` (replicas, extra labels).  The kernels take one counter row per `le`, so
    several series with one `le` are refused by default; with aggregate=True they are added as
    counters, which gives the same sum of rates only if they are scraped together and none of them
    resets (then increase(a) + increase(b) = increase(a + b) up to one rounding); anything else is
    refused."""
    by_le: Dict[float, List[Tuple[np.ndarray, np.ndarray]]] = {}
    for le, ts, v in group:
        by_le.setdefault(le, []).append((ts, v))
    out = []
    for le in sorted(by_le):
        series = by_le[le]
        if len(series) > 1:
            if not aggregate:
                raise DuplicateBucketSeries(
                    f"{what}: {len(series)} bucket series share le={le}; Prometheus would add their rates "
                    "(sum by (le)).  Pass aggregate_buckets=True to add them as counters (same scrape times, "
                    "no resets), or pre-aggregate.")
            ts0 = series[0][0]
            total = np.zeros(len(ts0), dtype=np.float64)
            for ts, v in series:
                if len(ts) != len(ts0) or not np.array_equal(ts, ts0):
                    raise DuplicateBucketSeries(f"{what}: series sharing le={le} are scraped at different times")
                if len(v) > 1 and np.any(np.diff(v) < 0):
                    raise DuplicateBucketSeries(f"{what}: a series sharing le={le} has a counter reset; "
                                                "its rate cannot be folded into a summed counter")
                total = total + v
            out.append((le, ts0, total))
        else:
            out.append((le, series[0][0], series[0][1]))
    return out


def build_host_batch(answers: Dict[str, Iterable], versions: Dict[Tuple[str, str], Tuple[str, str]],
                     period_ms: int = 15000, epoch_ms: Optional[int] = None, dtype: int = 0,
                     aggregate_buckets: bool = False, **param_overrides) -> HostSeries:
    """answers: {metric name: query_range answer}.  versions: {(deployment_name, namespace):
    (new predictor_name, old predictor_name)}.  Every (pair, version) must expose the same
    scalar label sets (same K, same roles in the same order after sorting by labels) and the
    same `le` set, sampled on a regular grid of `period_ms`."""
    pairs = sorted(versions)
    row_of = {key: i for i, key in enumerate(pairs)}
    ver_of = {}
    for key, (new, old) in versions.items():
        ver_of[(key, new)] = 0
        ver_of[(key, old)] = 1
    scal: Dict[Tuple[int, int], List[Tuple[tuple, int, np.ndarray, np.ndarray]]] = {}
    buck: Dict[Tuple[int, int], List[Tuple[float, np.ndarray, np.ndarray]]] = {}
    for name, answer in answers.items():
        for labels, ts, v in parse_matrix(answer):
            key = (labels.get("deployment_name", ""), labels.get("namespace", ""))
            vk = ver_of.get((key, labels.get("predictor_name", "")))
            if vk is None:
                continue
            slot = (row_of[key], vk)
            if name == CLIENT_BUCKET:
                buck.setdefault(slot, []).append((_le_value(labels["le"]), ts, v))
            else:
                # canonical row order: server counters, client sum, client count; then by labels
                order = ((SERVER_COUNT, CLIENT_SUM, CLIENT_COUNT).index(name), name,
                         tuple(sorted((k, x) for k, x in labels.items()
                                      if k not in ("deployment_name", "namespace", "predictor_name"))))
                scal.setdefault(slot, []).append((order, _scalar_role(name, labels), ts, v))
    n = len(pairs)
    if n == 0:
        raise ValueError("no model pairs")
    all_ts = [ts for group in list(scal.values()) + list(buck.values()) for item in group for ts in [item[-2]]]
    if not all_ts:
        raise ValueError("no series matched")
    if epoch_ms is None:
        epoch_ms = int(min(ts[0] for ts in all_ts if len(ts)))
    T = max(len(ts) for ts in all_ts)
    roles_ref, le_ref = None, None
    K = max((len(g) for g in scal.values()), default=0)
    B = max((len({it[0] for it in g}) for g in buck.values()), default=0)
    dt = np.float32 if dtype == 1 else np.float64
    scalars = np.zeros((n, 2, K, T), dtype=dt)
    buckets = np.zeros((n, 2, B, T), dtype=dt)
    phase = np.zeros((n, 2), dtype=np.int32)

    def grid_phase(ts, what):
        if len(ts) != T:
            raise IrregularSeries(f"{what}: {len(ts)} samples, expected {T}")
        if T > 1 and not np.all(np.diff(ts) == period_ms):
            raise IrregularSeries(f"{what}: samples are not {period_ms} ms apart")
        return int(ts[0] - epoch_ms)

    for i in range(n):
        for v in range(2):
            ph = None
            group = sorted(scal.get((i, v), []), key=lambda it: it[0])
            if len(group) != K:
                raise ValueError(f"pair {pairs[i]} version {v}: {len(group)} scalar series, expected {K}")
            roles = tuple(g[1] for g in group)
            roles_ref = roles if roles_ref is None else roles_ref
            if roles != roles_ref:
                raise ValueError(f"pair {pairs[i]} version {v}: scalar label sets differ from the first pair's")
            for k, (order, role, ts, vals) in enumerate(group):
                p0 = grid_phase(ts, f"{pairs[i]} v{v} {order[1]}")
                ph = p0 if ph is None else ph
                if p0 != ph:
                    raise IrregularSeries(f"{pairs[i]} v{v}: series of one target are scraped at different times")
                scalars[i, v, k] = vals
            hist = _merge_bucket_series(buck.get((i, v), []), f"pair {pairs[i]} version {v}", aggregate_buckets)
            if len(hist) != B:
                raise ValueError(f"pair {pairs[i]} version {v}: {len(hist)} buckets, expected {B}")
            les = np.array([h[0] for h in hist], dtype=np.float64)
            le_ref = les if le_ref is None else le_ref
            if B and not np.array_equal(les, le_ref):
                raise ValueError(f"pair {pairs[i]} version {v}: `le` set differs from the first pair's")
            for b, (_, ts, vals) in enumerate(hist):
                p0 = grid_phase(ts, f"{pairs[i]} v{v} le={les[b]}")
                ph = p0 if ph is None else ph
                if p0 != ph:
                    raise IrregularSeries(f"{pairs[i]} v{v}: series of one target are scraped at different times")
                buckets[i, v, b] = vals
            phase[i, v] = 0 if ph is None else ph
    params = make_params(n_samples=T, n_scalar=K, n_buckets=B, roles=roles_ref, period_ms=period_ms,
                         dtype=dtype, **param_overrides)
    if params.row_pitch * params.elem_bytes % 16:
        pitch = -(-T * params.elem_bytes // 16) * 16 // params.elem_bytes
        pad = lambda a: np.concatenate([a, np.zeros(a.shape[:-1] + (pitch - T,), a.dtype)], axis=-1)  # noqa: E731
        scalars, buckets = np.ascontiguousarray(pad(scalars)), np.ascontiguousarray(pad(buckets))
        params = make_params(n_samples=T, n_scalar=K, n_buckets=B, roles=roles_ref, period_ms=period_ms,
                             dtype=dtype, row_pitch=pitch, **param_overrides)
    index = {key: (row_of[key], {versions[key][0]: 0, versions[key][1]: 1}) for key in pairs}
    return HostSeries(scalars, buckets, le_ref if le_ref is not None else np.zeros(0), phase, params, index, epoch_ms)


def to_store(host: HostSeries, device="cuda", now_ms: Optional[int] = None):
    """Upload an ingested batch to the GPU as an operator_api.SeriesStore."""
    from .operator_api import SeriesStore
    return SeriesStore.from_host(host, device=device, index=host.index, now_ms=now_ms)


# ---- irregular series ---------------------------------------------
STALE_NAN_BITS = 0x7FF0000000000002          # Prometheus value.StaleNaN: "this series ended here"


def drop_stale(ts: np.ndarray, values: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Remove stale markers from one series (raw TSDB / remote-read samples carry them; the JSON
    query API never shows them).  They are not samples: range vectors skip them."""
    v = np.ascontiguousarray(values, dtype=np.float64)
    keep = v.view(np.uint64) != np.uint64(STALE_NAN_BITS)
    return np.asarray(ts)[keep], v[keep]


class HostIrregular:
    """What `build_irregular_batch` returns: values, per-series int32 ms time stamps (relative to
    epoch_ms) and sample counts, in the layout cg_eval_irregular reads, plus the label index."""

    def __init__(self, scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le, params, index, epoch_ms):
        self.scalars, self.buckets, self.ts_scalars, self.ts_buckets = scalars, buckets, ts_scalars, ts_buckets
        self.cnt_scalars, self.cnt_buckets, self.le = cnt_scalars, cnt_buckets, le
        self.params, self.index, self.epoch_ms = params, index, epoch_ms


def build_irregular_batch(answers: Dict[str, Iterable], versions: Dict[Tuple[str, str], Tuple[str, str]],
                          epoch_ms: Optional[int] = None, dtype: int = 0, **param_overrides) -> HostIrregular:
    """Like build_host_batch, for series that are not on a grid.  The scalar rows are the union of
    the label sets seen (a (pair, version) that lacks one gets a row with zero samples), the bucket
    rows the union of the `le` values; samples keep their own time stamps; stale markers are dropped;
    time stamps must be strictly increasing within a series and fit int32 ms from epoch_ms."""
    pairs = sorted(versions)
    if not pairs:
        raise ValueError("no model pairs")
    row_of = {key: i for i, key in enumerate(pairs)}
    ver_of = {}
    for key, (new, old) in versions.items():
        ver_of[(key, new)] = 0
        ver_of[(key, old)] = 1
    scal: Dict[tuple, Dict[Tuple[int, int], Tuple[np.ndarray, np.ndarray]]] = {}
    buck: Dict[float, Dict[Tuple[int, int], Tuple[np.ndarray, np.ndarray]]] = {}
    roles: Dict[tuple, int] = {}
    for name, answer in answers.items():
        for labels, ts, v in parse_matrix(answer):
            key = (labels.get("deployment_name", ""), labels.get("namespace", ""))
            vk = ver_of.get((key, labels.get("predictor_name", "")))
            if vk is None:
                continue
            ts, v = drop_stale(ts, v)
            if len(ts) > 1 and not np.all(np.diff(ts) > 0):
                raise ValueError(f"{name}{labels}: time stamps are not strictly increasing")
            slot = (row_of[key], vk)
            if name == CLIENT_BUCKET:
                per_le = buck.setdefault(_le_value(labels["le"]), {})
                if slot in per_le:
                    raise DuplicateBucketSeries(
                        f"{key} {labels.get('predictor_name')}: several bucket series share le={labels['le']}; Prometheus "
                        "would add their rates (sum by (le)); pre-aggregate them")
                per_le[slot] = (ts, v)
            else:
                order = ((SERVER_COUNT, CLIENT_SUM, CLIENT_COUNT).index(name), name,
                         tuple(sorted((k, x) for k, x in labels.items()
                                      if k not in ("deployment_name", "namespace", "predictor_name"))))
                roles[order] = _scalar_role(name, labels)
                scal.setdefault(order, {})[slot] = (ts, v)
    series = [s for d in list(scal.values()) + list(buck.values()) for s in d.values()]
    if not series:
        raise ValueError("no series matched")
    firsts = [int(ts[0]) for ts, _ in series if len(ts)]
    if epoch_ms is None:
        epoch_ms = min(firsts) if firsts else 0
    T = max(4, max(len(ts) for ts, _ in series))
    dt = np.float32 if dtype == 1 else np.float64
    pitch = -(-T // 4) * 4                                   # multiple of 4 samples: 16-byte rows in both arrays
    n, orders, les = len(pairs), sorted(scal), sorted(buck)
    K, B = len(orders), len(les)

    def fill(groups, R):
        vals = np.zeros((n, 2, R, pitch), dtype=dt)
        tss = np.zeros((n, 2, R, pitch), dtype=np.int32)
        cnt = np.zeros((n, 2, R), dtype=np.int32)
        for r, g in enumerate(groups):
            for (i, vk), (ts, v) in g.items():
                rel = np.asarray(ts, dtype=np.int64) - epoch_ms
                if len(rel) and (rel.min() < -(1 << 31) or rel.max() >= (1 << 31)):
                    raise ValueError("time stamps do not fit int32 milliseconds from epoch_ms")
                m = len(rel)
                vals[i, vk, r, :m], tss[i, vk, r, :m], cnt[i, vk, r] = v, rel, m
        return vals, tss, cnt

    scalars, ts_s, cnt_s = fill([scal[o] for o in orders], K)
    buckets, ts_b, cnt_b = fill([buck[x] for x in les], B)
    params = make_params(n_samples=T, row_pitch=pitch, n_scalar=K, n_buckets=B, roles=[roles[o] for o in orders],
                         dtype=dtype, **param_overrides)
    index = {key: (row_of[key], {versions[key][0]: 0, versions[key][1]: 1}) for key in pairs}
    return HostIrregular(scalars, buckets, ts_s, ts_b, cnt_s, cnt_b, np.array(les, dtype=np.float64), params,
                         index, epoch_ms)


def evaluate_irregular(host: HostIrregular, device="cuda", has_previous=None, start_tick=None, want_bits: bool = False):
    """Upload an irregular batch and run the whole canary on it (engine.eval_irregular)."""
    import torch

    from . import engine
    up = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a)).to(device)   # noqa: E731
    return engine.eval_irregular(up(host.scalars), up(host.buckets), up(host.ts_scalars), up(host.ts_buckets),
                                 up(host.cnt_scalars), up(host.cnt_buckets), up(host.le), host.params,
                                 has_previous=up(has_previous), start_tick=up(start_tick), want_bits=want_bits)
