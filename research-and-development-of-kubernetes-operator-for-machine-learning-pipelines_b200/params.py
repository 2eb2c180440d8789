"""ctypes mirror of `cg_params` / `cg_decision` (include/canary_gate.h) and the
reference's constants.

Every default below is a constant hard-coded in the reference operator
(/root/reference/mlflow_operator.py): window `elapsed_time=60` (:363), q = 0.95
(:367), thresholds 0.05 / 0.02 / 0.05 (:175-179), initial split 90/10 (:186-187),
max_traffic 100, step 10, interval 60 s, max_attempts 10, attempt_delay 10 s
(:290-294).
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Optional, Sequence

CG_F64, CG_F32 = 0, 1
ROLE_SERVER_TOTAL, ROLE_SERVER_ERROR, ROLE_CLIENT_SUM, ROLE_CLIENT_COUNT, ROLE_FEEDBACK = 1, 2, 4, 8, 16
MAX_SCALAR, MAX_Q = 16, 4
M_ERROR_RESPONSES, M_ERROR_RATE, M_LATENCY_AVG, M_REQUEST_COUNT, M_FEEDBACK_COUNT, M_EXTRA = 0, 1, 2, 3, 4, 5
V_QUANTILE, V_ERROR_RATE, V_LATENCY_AVG = 1, 2, 4
G_P95_OK, G_ERR_OK, G_AVG_OK, G_AVAILABLE, G_PROMOTE = 1, 2, 4, 8, 0x80
STATE_IN_PROGRESS, STATE_COMPLETE, STATE_ABORTED, STATE_NO_CANARY, STATE_ROLLED_BACK = 0, 1, 2, 3, 4
STATE_NAMES = {0: "in_progress", 1: "complete", 2: "aborted", 3: "no_canary", 4: "rolled_back"}
ABORT_FREEZE, ABORT_ROLLBACK = 0, 1

# mlflow_operator.py:175-179
REFERENCE_THRESHOLDS = {"latency_95th": 0.05, "error_rate": 0.02, "latency_avg": 0.05}


class CgParams(C.Structure):
    _fields_ = [
        ("struct_bytes", C.c_int32), ("dtype", C.c_int32),
        ("n_samples", C.c_int32), ("row_pitch", C.c_int32),
        ("n_scalar", C.c_int32), ("n_buckets", C.c_int32),
        ("scalar_role", C.c_uint8 * MAX_SCALAR),
        ("period_ms", C.c_int32), ("window_ms", C.c_int32),
        ("tick0_ms", C.c_int32), ("tick_step_ms", C.c_int32),
        ("n_ticks", C.c_int32), ("left_open", C.c_int32),
        ("n_q", C.c_int32), ("gate_q", C.c_int32),
        ("q", C.c_double * MAX_Q),
        ("one_plus_thr", C.c_double * 3),
        ("checks", C.c_int32),
        ("init_current", C.c_int32), ("max_traffic", C.c_int32), ("step", C.c_int32),
        ("interval_ticks", C.c_int32), ("max_attempts", C.c_int32),
        ("attempt_delay_ticks", C.c_int32), ("start_tick", C.c_int32),
        ("reserved", C.c_int32), ("on_abort", C.c_int32),
    ]

    @property
    def metric_width(self) -> int:
        return self.n_q + M_EXTRA

    @property
    def n_words(self) -> int:
        return (self.n_ticks + 31) // 32

    @property
    def elem_bytes(self) -> int:
        return 4 if self.dtype == CG_F32 else 8

    def bytes_per_pair(self) -> int:
        """Algorithmic HBM bytes per pair."""
        return 2 * (self.n_scalar + self.n_buckets) * self.n_samples * self.elem_bytes + 24 + 8

    def copy(self) -> "CgParams":
        out = CgParams()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(CgParams))
        return out


class CgDecision(C.Structure):
    _fields_ = [("traffic_current", C.c_int32), ("traffic_prev", C.c_int32),
                ("state", C.c_int32), ("n_promotions", C.c_int32),
                ("n_evals", C.c_int32), ("last_tick", C.c_int32)]


DEFAULT_ROLES = (ROLE_SERVER_TOTAL,                      # server count, code="200"
                 ROLE_SERVER_TOTAL | ROLE_SERVER_ERROR,  # server count, code="500"
                 ROLE_CLIENT_SUM, ROLE_CLIENT_COUNT)


def make_params(n_samples: int = 240, n_scalar: int = 4, n_buckets: int = 64,
                roles: Optional[Sequence[int]] = None, dtype: int = CG_F64,
                row_pitch: Optional[int] = None, period_ms: int = 15000,
                elapsed_time: int = 60, tick0_ms: Optional[int] = None,
                tick_step_ms: int = 10000, n_ticks: Optional[int] = None,
                left_open: bool = False, quantiles: Iterable[float] = (0.95,),
                gate_q: int = 0, thresholds: Optional[dict] = None, checks: int = 7,
                init_current: int = 10, max_traffic: int = 100, step: int = 10,
                interval_s: int = 60, max_attempts: int = 10, attempt_delay_s: int = 10,
                start_tick: int = 0, rollback: bool = False) -> CgParams:
    """Build a parameter block.  With no arguments: the reference's constants on
    BASELINE.json's full layout (1 h of 15 s scrapes, ticks every 10 s).
    rollback=True opts into sending all traffic back to the old version after max_attempts
    failures (state ROLLED_BACK); the reference freezes the weights there (:341-349)."""
    p = CgParams()
    p.struct_bytes = C.sizeof(CgParams)
    p.dtype = dtype
    p.n_samples = n_samples
    p.row_pitch = n_samples if row_pitch is None else row_pitch
    p.n_scalar, p.n_buckets = n_scalar, n_buckets
    if roles is None:
        roles = DEFAULT_ROLES[:n_scalar] if n_scalar <= 4 else tuple(DEFAULT_ROLES) + (
            ROLE_SERVER_TOTAL | ROLE_FEEDBACK,) * (n_scalar - 4)
    if len(roles) != n_scalar:
        raise ValueError("len(roles) must equal n_scalar")
    for i, r in enumerate(roles):
        p.scalar_role[i] = r
    p.period_ms = period_ms
    p.window_ms = int(elapsed_time * 1000)
    p.tick0_ms = p.window_ms if tick0_ms is None else tick0_ms
    p.tick_step_ms = tick_step_ms
    if n_ticks is None:
        span = n_samples * period_ms               # 240 * 15 s = 1 h
        n_ticks = max(0, (span - p.tick0_ms) // tick_step_ms + 1)
    p.n_ticks = n_ticks
    p.left_open = 1 if left_open else 0
    qs = list(quantiles)
    if n_buckets == 0:
        qs = []
    if len(qs) > MAX_Q:
        raise ValueError("too many quantiles")
    p.n_q, p.gate_q = len(qs), gate_q
    for i, q in enumerate(qs):
        p.q[i] = q
    thr = dict(REFERENCE_THRESHOLDS if thresholds is None else thresholds)
    # the reference evaluates (1 + thresholds[k]) in Python floats at run time (:440,447,454)
    p.one_plus_thr[0] = 1 + thr["latency_95th"]
    p.one_plus_thr[1] = 1 + thr["error_rate"]
    p.one_plus_thr[2] = 1 + thr["latency_avg"]
    p.checks = checks
    p.init_current, p.max_traffic, p.step = init_current, max_traffic, step
    if (interval_s * 1000) % tick_step_ms or (attempt_delay_s * 1000) % tick_step_ms:
        raise ValueError("interval and attempt_delay must be multiples of the tick step")
    p.interval_ticks = interval_s * 1000 // tick_step_ms
    p.max_attempts = max_attempts
    p.attempt_delay_ticks = attempt_delay_s * 1000 // tick_step_ms
    p.start_tick = start_tick
    p.on_abort = ABORT_ROLLBACK if rollback else ABORT_FREEZE
    return p
