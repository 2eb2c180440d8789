"""Host-side mirror of the reference operator's interface for the canary path.

Same names, argument meaning, return types and error behaviour as
/root/reference/mlflow_operator.py:

    get_model_metrics(prom, deployment_name, predictor_name, namespace, elapsed_time=60) -> dict   (:363)
    should_promote_model(new_metrics, old_metrics, thresholds, logger) -> bool                     (:419)

so the operator's canary loop (:296-352) keeps calling them unchanged.  `prom` is a
`SeriesStore` (device-resident scrape history) instead of a PrometheusConnect: the six
PromQL queries (:367-410) are evaluated by the CUDA library, not by an HTTP round trip.
A real PrometheusConnect is rejected: this package has no CPU path.

`evaluate_canaries` is the batched form (every MlflowModel resource at once) and
`canary_schedule` the pure function the reference never factored out of its inline loop
(SURVEY.md F3).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Sequence, Tuple

import torch

from . import engine
from .params import (CG_F32, CgParams, G_AVAILABLE, G_AVG_OK, G_ERR_OK, G_P95_OK, G_PROMOTE,
                     M_ERROR_RATE, M_ERROR_RESPONSES, M_FEEDBACK_COUNT, M_LATENCY_AVG,
                     M_REQUEST_COUNT, STATE_NAMES, V_ERROR_RATE, V_LATENCY_AVG, V_QUANTILE,
                     make_params)


class SeriesStore:
    """Device-resident scrape history for many MlflowModel resources.

    Rows are model pairs; `index[(deployment_name, namespace)] -> (row, {predictor_name:
    version})` plays the part of Prometheus' label index for the three labels every query
    of the reference fixes (:367-410).  `now_ms` is the evaluation time of instant queries (the
    reference uses the server's wall clock) in the STORE's time: milliseconds since `epoch_ms`,
    the Unix time of the store's origin (0 for synthetic batches, the first time stamp for
    ingested ones).  Device time is int32 milliseconds: `set_wall_time` converts a Unix time and
    every evaluation refuses a time outside +-2^30 ms of the origin instead of wrapping."""

    TIME_LIMIT_MS = 1 << 30

    def __init__(self, scalars: torch.Tensor, buckets: torch.Tensor, le: torch.Tensor,
                 phase: torch.Tensor, params: CgParams, index: Optional[Dict] = None,
                 now_ms: Optional[int] = None, epoch_ms: int = 0):
        self.scalars, self.buckets, self.le, self.phase = scalars, buckets, le, phase
        self.params = params
        self.index: Dict[Tuple[str, str], Tuple[int, Dict[str, int]]] = index or {}
        self.epoch_ms = int(epoch_ms)
        self.now_ms = params.tick0_ms if now_ms is None else now_ms
        self._cache: Dict[Tuple[int, int, int], tuple] = {}

    @property
    def n_pairs(self) -> int:
        return int(self.phase.shape[0])

    @classmethod
    def from_host(cls, batch, device="cuda", index=None, now_ms=None) -> "SeriesStore":
        """Upload a synth.SynthBatch-like object (numpy arrays) to the GPU."""
        dev = torch.device(device)
        t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
        return cls(t(batch.scalars), t(batch.buckets), t(batch.le), t(batch.phase),
                   batch.params, index=index, now_ms=now_ms, epoch_ms=getattr(batch, "epoch_ms", 0) or 0)

    def register(self, deployment_name: str, namespace: str, row: int, new_version, old_version):
        self.index[(deployment_name, namespace)] = (row, {f"v{new_version}": 0, f"v{old_version}": 1})

    def set_time(self, now_ms: int):
        """Evaluation time in the store's time (ms since `epoch_ms`)."""
        self.now_ms = int(now_ms)

    def set_wall_time(self, unix_ms: int):
        """Evaluation time as Unix milliseconds (what a Prometheus server would call now)."""
        self.now_ms = int(unix_ms) - self.epoch_ms

    def _device_now(self, elapsed_time) -> int:
        """`now_ms` checked against the int32 range the kernels compute in."""
        now, lim = int(self.now_ms), self.TIME_LIMIT_MS
        if not (-lim < now - int(elapsed_time * 1000) and now < lim):
            raise ValueError(f"evaluation time {now} ms is outside +-2^30 ms of the store's origin (epoch_ms = {self.epoch_ms}); "
                             "pass store time, or use set_wall_time() for Unix time")
        return now

    def advance(self, seconds: float):
        self.now_ms += int(round(seconds * 1000))

    # -- one instant evaluation of both versions of one row, cached per (row, now, window)
    def _instant(self, row: int, elapsed_time: int):
        key = (row, self.now_ms, int(elapsed_time))
        hit = self._cache.get(key)
        if hit is not None:
            return hit
        p = self.params.copy()
        p.window_ms = int(elapsed_time * 1000)
        p.tick0_ms = self._device_now(elapsed_time)
        p.n_ticks = 1
        metrics, valid = engine.eval_grid(self.scalars[row:row + 1], self.buckets[row:row + 1],
                                          self.le, self.phase[row:row + 1], p)
        out = (metrics[0, :, 0, :].cpu(), valid[0, :, 0].cpu(), p)
        self._cache.clear()
        self._cache[key] = out
        return out


class IrregularSeriesStore(SeriesStore):
    """The same store for scrapes that are not a perfect grid: every series
    has its own int32 ms time stamps and sample count (synth.IrregularBatch / ingest.HostIrregular
    layout).  get_model_metrics / should_promote_model / evaluate_canaries work on it unchanged."""

    def __init__(self, scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le, params: CgParams,
                 index: Optional[Dict] = None, now_ms: Optional[int] = None, epoch_ms: int = 0):
        self.scalars, self.buckets, self.le = scalars, buckets, le
        self.ts_scalars, self.ts_buckets = ts_scalars, ts_buckets
        self.cnt_scalars, self.cnt_buckets = cnt_scalars, cnt_buckets
        self.phase = None
        self.params = params
        self.index = index or {}
        self.epoch_ms = int(epoch_ms)
        self.now_ms = params.tick0_ms if now_ms is None else now_ms
        self._cache = {}

    @property
    def n_pairs(self) -> int:
        return int(self.cnt_scalars.shape[0])

    @classmethod
    def from_host(cls, batch, device="cuda", index=None, now_ms=None) -> "IrregularSeriesStore":
        dev = torch.device(device)
        t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
        return cls(t(batch.scalars), t(batch.buckets), t(batch.ts_scalars), t(batch.ts_buckets),
                   t(batch.cnt_scalars), t(batch.cnt_buckets), t(batch.le), batch.params,
                   index=index if index is not None else getattr(batch, "index", None), now_ms=now_ms,
                   epoch_ms=getattr(batch, "epoch_ms", 0) or 0)

    def _series(self, rows=slice(None)):
        return (self.scalars[rows], self.buckets[rows], self.ts_scalars[rows], self.ts_buckets[rows],
                self.cnt_scalars[rows], self.cnt_buckets[rows], self.le)

    def _instant(self, row: int, elapsed_time: int):
        key = (row, self.now_ms, int(elapsed_time))
        hit = self._cache.get(key)
        if hit is not None:
            return hit
        p = self.params.copy()
        p.window_ms = int(elapsed_time * 1000)
        p.tick0_ms = self._device_now(elapsed_time)
        p.n_ticks = 1
        _, metrics, valid = engine.eval_irregular(*self._series(slice(row, row + 1)), p, want_metrics=True)
        out = (metrics[0, :, 0, :].cpu(), valid[0, :, 0].cpu(), p)
        self._cache.clear()
        self._cache[key] = out
        return out


def _record_to_dict(rec, valid: int, p: CgParams) -> dict:
    """Metric record -> the dict get_model_metrics builds (:364-417)."""
    nq = p.n_q
    m = {}
    m["latency_95th"] = float(rec[p.gate_q]) if (nq > 0 and valid & V_QUANTILE) else None   # :369-372
    m["error_responses"] = float(rec[nq + M_ERROR_RESPONSES])                                # :377-380
    m["error_rate"] = float(rec[nq + M_ERROR_RATE]) if valid & V_ERROR_RATE else None        # :387-390
    m["latency_avg"] = float(rec[nq + M_LATENCY_AVG]) if valid & V_LATENCY_AVG else None     # :401-404
    m["request_count"] = float(rec[nq + M_REQUEST_COUNT])                                    # :407
    m["feedback_request_count"] = float(rec[nq + M_FEEDBACK_COUNT])                          # :412-415
    return m


def get_model_metrics(prom, deployment_name, predictor_name, namespace, elapsed_time=60):
    """Drop-in for mlflow_operator.py:363.  Unknown deployments / predictors behave like
    Prometheus selecting no series: latency_95th None, sums 0.0, ratios None (:369-415)."""
    if not isinstance(prom, SeriesStore):
        raise TypeError("get_model_metrics needs a SeriesStore: the B200 path evaluates PromQL on "
                        "device-resident series and has no HTTP/CPU fallback")
    hit = prom.index.get((deployment_name, namespace))
    ver = None if hit is None else hit[1].get(predictor_name)
    if hit is None or ver is None:
        return {"latency_95th": None, "error_responses": 0.0, "error_rate": None,
                "latency_avg": None, "request_count": 0.0, "feedback_request_count": 0.0}
    rec, valid, p = prom._instant(hit[0], elapsed_time)
    return _record_to_dict(rec[ver].tolist(), int(valid[ver]), p)


def _dict_to_record(m: dict, p: CgParams):
    nq = max(p.n_q, 1)
    rec = [math.nan] * (nq + 5)
    valid = 0

    def put(key, slot, bit):
        nonlocal valid
        v = m.get(key)
        if v is not None:
            rec[slot] = float(v)
            valid |= bit
    put("latency_95th", p.gate_q if p.n_q else 0, V_QUANTILE)
    put("error_rate", nq + M_ERROR_RATE, V_ERROR_RATE)
    put("latency_avg", nq + M_LATENCY_AVG, V_LATENCY_AVG)
    return rec, valid


def should_promote_model(new_metrics, old_metrics, thresholds, logger):
    """Drop-in for mlflow_operator.py:419.  The three comparisons run on the GPU (cg_gate);
    the log lines are the reference's own (:433, :441-457)."""
    for metric in ("latency_95th", "error_rate", "latency_avg"):                 # :430-434
        if new_metrics.get(metric) is None or old_metrics.get(metric) is None:
            logger.warning(f"Metric {metric} is not available for one of the models.")
            return False
    p = make_params(n_buckets=1, quantiles=(0.95,), thresholds=thresholds)
    nr, nv = _dict_to_record(new_metrics, p)
    orr, ov = _dict_to_record(old_metrics, p)
    dev = torch.device("cuda")
    detail = int(engine.gate(torch.tensor([nr], dtype=torch.float64, device=dev),
                             torch.tensor([nv], dtype=torch.uint8, device=dev),
                             torch.tensor([orr], dtype=torch.float64, device=dev),
                             torch.tensor([ov], dtype=torch.uint8, device=dev), p).cpu()[0])
    if detail & G_P95_OK:                                                        # :440-444
        logger.info("95th percentile latency of the new model is acceptable.")
    else:
        logger.warning("95th percentile latency of the new model exceeds the allowed threshold.")
    if detail & G_ERR_OK:                                                        # :447-451
        logger.info("Error rate of the new model is acceptable.")
    else:
        logger.warning("Error rate of the new model exceeds the allowed threshold.")
    if detail & G_AVG_OK:                                                        # :454-458
        logger.info("Average latency of the new model is acceptable.")
    else:
        logger.warning("Average latency of the new model exceeds the allowed threshold.")
    return bool(detail & G_PROMOTE)                                              # :460


# ---- batched forms (new API; the reference evaluates one resource at a time) -------
def evaluate_canaries(store: SeriesStore, has_previous: Optional[torch.Tensor] = None,
                      start_tick: Optional[torch.Tensor] = None, want_bits: bool = False,
                      out: Optional[torch.Tensor] = None):
    """Run the whole canary (metrics on the dense tick grid -> gate -> step loop) for every
    row of the store.  -> decisions [N, 6] int32: traffic_current, traffic_prev, state,
    n_promotions, n_evals, last_tick."""
    if isinstance(store, IrregularSeriesStore):
        res = engine.eval_irregular(*store._series(), store.params, has_previous=has_previous,
                                    start_tick=start_tick, want_bits=want_bits)
        if out is not None:
            out.copy_(res[0] if want_bits else res)
            return (out, res[1]) if want_bits else out
        return res
    return engine.eval_fused(store.scalars, store.buckets, store.le, store.phase, store.params,
                             has_previous=has_previous, start_tick=start_tick,
                             want_bits=want_bits, out=out)


def canary_schedule(gate_bits: torch.Tensor, params: CgParams, has_previous=None, start_tick=None):
    """The step loop of mlflow_operator.py:296-361 as a pure function of gate outcomes."""
    return engine.replay(gate_bits, params, has_previous=has_previous, start_tick=start_tick)


def decisions_to_predictor_traffic(decisions: torch.Tensor, current_versions: Sequence,
                                   previous_versions: Sequence):
    """Decision records -> the `predictors[*].traffic` values the operator patches into each
    SeldonDeployment (:320-328), with the old predictor dropped at 100 % (:355-358).  With the
    opt-in rollback policy (params.make_params(rollback=True)) a rolled-back canary collapses to the
    OLD predictor alone, the mirror image of completion."""
    rows = decisions.cpu().tolist()
    out = []
    for (cur, prev, state, *_), cv, pv in zip(rows, current_versions, previous_versions):
        if pv is not None and state == 4:
            out.append({"state": STATE_NAMES[state], "predictors": [{"name": f"v{pv}", "traffic": 100}]})
        elif pv is None or state == 1 or state == 3:
            out.append({"state": STATE_NAMES[state], "predictors": [{"name": f"v{cv}", "traffic": 100}]})
        else:
            out.append({"state": STATE_NAMES[state],
                        "predictors": [{"name": f"v{pv}", "traffic": prev},
                                       {"name": f"v{cv}", "traffic": cur}]})
    return out
