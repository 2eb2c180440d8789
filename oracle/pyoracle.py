"""Pure-Python CPU restatement of the PromQL arithmetic the canary path relies on.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import this file;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs use it, and only as the checker.

The reference (mlflow_operator.py) does none of this arithmetic itself: it sends
six PromQL strings to an external Prometheus server (mlflow_operator.py:367,
:375, :383, :393, :397, :410) and parses one float out of each answer.  The
server (github.com/prometheus/prometheus, package promql; functions.go
`extrapolatedRate`, quantile.go `bucketQuantile`) is NOT vendored under
/root/reference and no version is pinned anywhere in the reference, so the
values computed here are "parity unpinned" at that boundary: they follow
SURVEY.md Appendix A (Prometheus 2.x semantics) and are pinned only by the
hand-verifiable known-answer vectors of SURVEY.md Appendix B
(tests/golden/kat_promql.json).  Decisions and traffic weights ARE pinned: they
come from running the reference's own get_model_metrics / should_promote_model /
step loop (oracle/ref_loader.py) on top of these values.

Every function works on plain Python floats (IEEE fp64) and ints; no numpy.
"""
from __future__ import annotations

import math
import re
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

Sample = Tuple[int, float]  # (timestamp in ms, value)


# --------------------------------------------------------------------------
# range selection  (SURVEY.md Appendix A "Range selection")
# --------------------------------------------------------------------------
def select_range(samples: Sequence[Sample], t_ms: int, range_ms: int,
                 left_open: bool = False) -> List[Sample]:
    """Samples with t-r <= ts <= t (Prometheus 2.x) or t-r < ts <= t (3.x)."""
    lo = t_ms - range_ms
    if left_open:
        return [s for s in samples if lo < s[0] <= t_ms]
    return [s for s in samples if lo <= s[0] <= t_ms]


# --------------------------------------------------------------------------
# extrapolated rate / increase for one counter series
# (upstream promql/functions.go extrapolatedRate)
# --------------------------------------------------------------------------
def extrapolated_increase(window: Sequence[Sample], t_ms: int,
                          range_ms: int) -> Optional[float]:
    n = len(window)
    if n < 2:                                   # step 1: no output
        return None
    first_t, first_v = window[0]
    last_t, last_v = window[-1]
    result = last_v - first_v                   # step 2
    prev = first_v
    for _, v in window[1:]:
        if v < prev:                            # counter reset
            result += prev
        prev = v
    range_start = t_ms - range_ms
    d_start = (first_t - range_start) / 1000.0  # step 3
    d_end = (t_ms - last_t) / 1000.0
    sampled = (last_t - first_t) / 1000.0
    avg = sampled / (n - 1)
    if result > 0 and first_v >= 0:             # step 4: zero-point clamp
        d_zero = sampled * (first_v / result)
        if d_zero < d_start:
            d_start = d_zero
    thr = avg * 1.1                             # step 5
    ext = sampled
    ext += d_start if d_start < thr else avg / 2
    ext += d_end if d_end < thr else avg / 2
    return result * (ext / sampled)             # step 6 (increase)


def extrapolated_rate(window: Sequence[Sample], t_ms: int,
                      range_ms: int) -> Optional[float]:
    inc = extrapolated_increase(window, t_ms, range_ms)
    if inc is None:
        return None
    return inc / (range_ms / 1000.0)            # step 6 (rate)


# --------------------------------------------------------------------------
# histogram_quantile  (upstream promql/quantile.go)
# --------------------------------------------------------------------------
def bucket_quantile(q: float, buckets: Sequence[Tuple[float, float]]) -> float:
    """buckets = [(upper bound `le`, count)], any order.  Returns a float
    (possibly NaN / +-Inf).  The empty-input case (no output at all) is the
    caller's business (`histogram_quantile_vector`)."""
    if math.isnan(q):
        return math.nan
    if q < 0:
        return -math.inf
    if q > 1:
        return math.inf
    bs = sorted(buckets, key=lambda b: b[0])
    if not bs or not (math.isinf(bs[-1][0]) and bs[-1][0] > 0):
        return math.nan
    if len(bs) < 2:
        return math.nan
    les = [b[0] for b in bs]
    cnt = [b[1] for b in bs]
    mx = cnt[0]                                 # step 3: force monotonic
    for i in range(1, len(cnt)):
        if cnt[i] > mx:
            mx = cnt[i]
        elif cnt[i] < mx:
            cnt[i] = mx
    obs = cnt[-1]
    if obs == 0:                                # step 4
        return math.nan
    rank = q * obs                              # step 5
    n = len(cnt)
    b = n - 1
    for i in range(n - 1):
        if cnt[i] >= rank:
            b = i
            break
    if b == n - 1:                              # step 6
        return les[n - 2]
    if b == 0 and les[0] <= 0:
        return les[0]
    start, end, c = 0.0, les[b], cnt[b]         # step 7
    if b > 0:
        start = les[b - 1]
        c -= cnt[b - 1]
        rank -= cnt[b - 1]
    return start + (end - start) * (rank / c)


# --------------------------------------------------------------------------
# A tiny label-indexed TSDB and an evaluator for the six query shapes the
# reference issues.  Only what mlflow_operator.py:367-410 needs.
# --------------------------------------------------------------------------
class SeriesDB:
    """series: list of (metric_name, labels dict, [(ts_ms, value), ...])."""

    def __init__(self, left_open: bool = False):
        self.series: List[Tuple[str, Dict[str, str], List[Sample]]] = []
        self.left_open = left_open

    def add(self, name: str, labels: Dict[str, str], samples: Iterable[Sample]):
        self.series.append((name, dict(labels), list(samples)))

    def select(self, name: str, matchers: List[Tuple[str, str, str]]):
        out = []
        for n, labels, samples in self.series:
            if n != name:
                continue
            ok = True
            for key, op, val in matchers:
                have = labels.get(key, "")
                if op == "=" and have != val:
                    ok = False
                elif op == "!=" and have == val:
                    ok = False
            if ok:
                out.append((labels, samples))
        return out


_SEL = r'(?P<name>[a-zA-Z_:][a-zA-Z0-9_:]*)\{(?P<matchers>[^}]*)\}\[(?P<range>\d+)s\]'
_RE_QUANTILE = re.compile(
    r'^histogram_quantile\((?P<q>[-+0-9.eE]+), sum\(rate\(' + _SEL + r'\)\) by \(le\)\)$')
_RE_SUMINC = re.compile(
    r'^sum\(increase\(' + _SEL + r'\)\) or on\(\) vector\(0\)$')
_RE_MATCHER = re.compile(r'\s*([a-zA-Z_][a-zA-Z0-9_]*)\s*(!=|=)\s*"([^"]*)"\s*')


def _parse_matchers(text: str) -> List[Tuple[str, str, str]]:
    out = []
    for part in text.split(","):
        if not part.strip():
            continue
        m = _RE_MATCHER.fullmatch(part)
        if not m:
            raise ValueError(f"unsupported matcher: {part!r}")
        out.append((m.group(1), m.group(2), m.group(3)))
    return out


def _fmt(v: float) -> str:
    """Prometheus sends sample values as strings."""
    if math.isnan(v):
        return "NaN"
    if math.isinf(v):
        return "+Inf" if v > 0 else "-Inf"
    return repr(float(v))


def eval_query(db: SeriesDB, query: str, now_ms: int) -> list:
    """Instant-query evaluator for the two query shapes of the reference.
    Returns the `result` list of a Prometheus /api/v1/query vector answer,
    i.e. what PrometheusConnect.custom_query returns (mlflow_operator.py:368)."""
    query = query.strip()
    m = _RE_QUANTILE.match(query)
    if m:
        q = float(m.group("q"))
        range_ms = int(m.group("range")) * 1000
        sel = db.select(m.group("name"), _parse_matchers(m.group("matchers")))
        by_le: Dict[float, float] = {}
        for labels, samples in sel:
            w = select_range(samples, now_ms, range_ms, db.left_open)
            r = extrapolated_rate(w, now_ms, range_ms)
            if r is None:
                continue
            le = float(labels["le"].replace("+Inf", "inf"))
            by_le[le] = by_le.get(le, 0.0) + r if le in by_le else r
        if not by_le:
            return []                           # empty vector -> None at :369-372
        val = bucket_quantile(q, list(by_le.items()))
        return [{"metric": {}, "value": [now_ms / 1000.0, _fmt(val)]}]
    m = _RE_SUMINC.match(query)
    if m:
        range_ms = int(m.group("range")) * 1000
        sel = db.select(m.group("name"), _parse_matchers(m.group("matchers")))
        total = None
        for labels, samples in sel:
            w = select_range(samples, now_ms, range_ms, db.left_open)
            inc = extrapolated_increase(w, now_ms, range_ms)
            if inc is None:
                continue
            total = inc if total is None else total + inc
        if total is None:                       # `or on() vector(0)`
            total = 0.0
        return [{"metric": {}, "value": [now_ms / 1000.0, _fmt(total)]}]
    raise ValueError(f"query shape not issued by the reference: {query!r}")


# --------------------------------------------------------------------------
# Restatement of the reference's own Python logic, used where the reference
# itself cannot be imported (the GPU box has no /root/reference).  Checked
# against the real thing in tests/test_oracle_vs_reference.py.
# --------------------------------------------------------------------------
def derive_metrics(p95: Optional[float], err: float, total: float,
                   lat_sum: float, lat_count: float, feedback: float) -> dict:
    """mlflow_operator.py:369-415 (defaults and the two divisions)."""
    m = {"latency_95th": p95, "error_responses": err}
    m["error_rate"] = err / total if total > 0 else None          # :387-390
    m["latency_avg"] = lat_sum / lat_count if lat_count > 0 else None  # :401-404
    m["request_count"] = lat_count                                # :407
    m["feedback_request_count"] = feedback                        # :410-415
    return m


def should_promote(new: dict, old: dict, thresholds: dict) -> bool:
    """mlflow_operator.py:419-460 without the logging."""
    for k in ("latency_95th", "error_rate", "latency_avg"):       # :430-434
        if new.get(k) is None or old.get(k) is None:
            return False
    ok = True
    if not (new["latency_95th"] <= old["latency_95th"] * (1 + thresholds["latency_95th"])):
        ok = False                                                # :440
    if not (new["error_rate"] <= old["error_rate"] * (1 + thresholds["error_rate"])):
        ok = False                                                # :447
    if not (new["latency_avg"] <= old["latency_avg"] * (1 + thresholds["latency_avg"])):
        ok = False                                                # :454
    return ok


STATE_IN_PROGRESS, STATE_COMPLETE, STATE_ABORTED, STATE_NO_CANARY, STATE_ROLLED_BACK = 0, 1, 2, 3, 4


def replay_schedule(gate_bits: Sequence[bool], start_tick: int = 0, init_current: int = 10,
                    max_traffic: int = 100, step: int = 10, interval_ticks: int = 6,
                    max_attempts: int = 10, attempt_delay_ticks: int = 1,
                    has_previous: bool = True, rollback: bool = False) -> dict:
    """mlflow_operator.py:184-191 and :290-361 as a pure function over gate
    bits sampled on the tick grid.  Returns the decision record.
    rollback=True is the opt-in extension for the placeholder the reference leaves at :345:
    after the 10th failed attempt all traffic goes back to the old version."""
    if not has_previous:                                          # :188-191, :289
        return dict(cur=100, prev=0, state=STATE_NO_CANARY, n_promotions=0,
                    n_evals=0, last_tick=-1)
    cur, prev = init_current, max_traffic - init_current          # :186-187
    tick, n_prom, n_eval, last = start_tick, 0, 0, -1
    state = STATE_IN_PROGRESS
    while cur < max_traffic:                                      # :296
        attempt, success = 0, False
        while attempt < max_attempts:                             # :300
            if tick >= len(gate_bits):                            # ran out of data
                return dict(cur=cur, prev=prev, state=STATE_IN_PROGRESS,
                            n_promotions=n_prom, n_evals=n_eval, last_tick=last)
            n_eval += 1
            last = tick
            if gate_bits[tick]:                                   # :310
                cur = min(cur + step, max_traffic)                # :312-317
                prev = max(prev - step, 0)
                n_prom += 1
                success = True
                break
            attempt += 1                                          # :337
            if attempt < max_attempts:
                tick += attempt_delay_ticks                       # :340
        if not success:                                           # :347-349
            state = STATE_ABORTED
            break
        tick += interval_ticks                                    # :352
    if cur == max_traffic:                                        # :355
        state = STATE_COMPLETE
    if state == STATE_ABORTED and rollback:                       # not in the reference (:345 is a comment)
        cur, prev, state = 0, max_traffic, STATE_ROLLED_BACK
    return dict(cur=cur, prev=prev, state=state, n_promotions=n_prom,
                n_evals=n_eval, last_tick=last)
