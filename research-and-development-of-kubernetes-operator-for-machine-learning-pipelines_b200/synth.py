"""Seeded synthetic Prometheus-style series in the library's SoA layout (host side).

Follows SURVEY.md 8(d): per-scrape request counts ~ Poisson(lambda), errors ~
Binomial(requests, p_err), log-normal request latencies binned into cumulative
`le` buckets, counters starting at a random offset, optional counter resets, and a
small share of adversarial pairs (zero traffic, empty histogram, non-monotonic
buckets, old error rate 0 with new > 0).  The series names are the ones the
reference queries (mlflow_operator.py:367-410):

  scalar row 0  seldon_api_executor_server_requests_seconds_count{code="200"}
  scalar row 1  seldon_api_executor_server_requests_seconds_count{code="500"}
  scalar row 2  seldon_api_executor_client_requests_seconds_sum
  scalar row 3  seldon_api_executor_client_requests_seconds_count
  (row 4        seldon_api_executor_server_requests_seconds_count{service="feedback"})
  bucket row b  seldon_api_executor_client_requests_seconds_bucket{le=le[b]}

Large benchmark inputs are produced on the GPU by the library's own generator
kernel (`cg_synth_fill`); this numpy version feeds the CPU tests and small cases.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .params import CgParams, CG_F32


def default_le(n_buckets: int = 64, lo: float = 1e-3, hi: float = 30.0) -> np.ndarray:
    """Geometric upper bounds from `lo` to ~`hi` seconds, last = +Inf."""
    if n_buckets == 0:
        return np.zeros(0, dtype=np.float64)
    if n_buckets == 1:
        return np.array([np.inf])
    le = np.empty(n_buckets, dtype=np.float64)
    if n_buckets > 2:
        le[:-1] = np.geomspace(lo, hi, n_buckets - 1)
    else:
        le[0] = lo
    le[-1] = np.inf
    return le


@dataclass
class SynthBatch:
    scalars: np.ndarray      # [N, 2, K, pitch]
    buckets: np.ndarray      # [N, 2, B, pitch]
    le: np.ndarray           # [B]
    phase: np.ndarray        # [N, 2] int32
    params: CgParams


def _norm_cdf(x):
    return 0.5 * (1.0 + np.vectorize(math.erf)(x / math.sqrt(2.0)))


def generate(p: CgParams, n_pairs: int, seed: int = 0xC0FFEE, resets: bool = True,
             adversarial: float = 0.01, shared_phase: bool = False,
             coincident_phase: bool = False, le: Optional[np.ndarray] = None) -> SynthBatch:
    """Make `n_pairs` model pairs.  Deterministic for a given (params, seed)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    N, K, B, T, pitch = n_pairs, p.n_scalar, p.n_buckets, p.n_samples, p.row_pitch
    le = default_le(B) if le is None else np.asarray(le, dtype=np.float64)

    lam = np.exp(rng.uniform(math.log(5.0), math.log(500.0), size=N))
    p_err_old = np.where(rng.random(N) < 0.3, 0.0,
                         np.exp(rng.uniform(math.log(1e-4), math.log(5e-2), size=N)))
    p_err_new = np.minimum(p_err_old * rng.lognormal(0.0, 0.15, size=N), 1.0)
    mu_old = rng.uniform(math.log(0.01), math.log(0.5), size=N)
    mu_new = mu_old + rng.normal(0.0, 0.04, size=N)
    sigma = 0.5

    adv = rng.random(N) < adversarial
    adv_kind = rng.integers(0, 4, size=N)

    scal_inc = np.zeros((N, 2, K, T), dtype=np.float64)
    buck_inc = np.zeros((N, 2, B, T), dtype=np.float64)
    for ver, (p_err, mu) in enumerate(((p_err_new, mu_new), (p_err_old, mu_old))):
        req = rng.poisson(lam[:, None], size=(N, T)).astype(np.float64)
        zero = adv & (adv_kind == 0)
        req[zero] = 0.0                                    # zero traffic
        pe = p_err.copy()
        if ver == 0:
            pe = np.where(adv & (adv_kind == 3), np.maximum(pe, 0.02), pe)
        else:
            pe = np.where(adv & (adv_kind == 3), 0.0, pe)  # old er = 0, new er > 0
        err = rng.binomial(req.astype(np.int64), pe[:, None]).astype(np.float64)
        if K >= 1:
            scal_inc[:, ver, 0] = req - err                # code="200"
        if K >= 2:
            scal_inc[:, ver, 1] = err                      # code="500"
        mu_t = mu[:, None] + rng.normal(0.0, 0.02, size=(N, T))
        if K >= 3:
            mean_lat = np.exp(mu_t + 0.5 * sigma * sigma)
            scal_inc[:, ver, 2] = req * mean_lat * rng.uniform(0.95, 1.05, size=(N, T))
        if K >= 4:
            scal_inc[:, ver, 3] = req                      # client count
        if K >= 5:
            scal_inc[:, ver, 4:] = rng.poisson(0.05 * lam[:, None, None], size=(N, K - 4, T))
        if B > 0:
            with np.errstate(divide="ignore"):
                z = (np.log(le[None, :, None]) - mu_t[:, None, :]) / sigma   # [N, B, T]
            cdf = _norm_cdf(z)
            cdf[:, -1, :] = 1.0
            cum = np.floor(req[:, None, :] * cdf + 0.5)
            cum = np.maximum.accumulate(cum, axis=1)
            cum[:, -1, :] = req
            empty = adv & (adv_kind == 1)
            cum[empty] = 0.0                               # all-zero histogram
            buck_inc[:, ver] = cum

    def to_counters(inc):
        n_series = inc.shape[2]
        start = np.floor(rng.uniform(0.0, 1e6, size=(N, 2, n_series, 1)))
        if not resets:
            return start + np.cumsum(inc, axis=3)
        reset = rng.random(inc.shape) < (1.0 / 240.0)
        reset[..., 0] = False
        out = np.empty_like(inc)
        run = start[..., 0] + inc[..., 0]
        out[..., 0] = run
        for i in range(1, T):
            run = np.where(reset[..., i], 0.0, run) + inc[..., i]
            out[..., i] = run
        return out

    scal = to_counters(scal_inc)
    buck = to_counters(buck_inc) if B > 0 else buck_inc
    if B > 2:
        nonmono = adv & (adv_kind == 2)                    # raw non-monotonic buckets
        idx = np.nonzero(nonmono)[0]
        for i in idx:
            b = int(rng.integers(1, B - 1))
            buck[i, :, b, :] *= 0.5

    dt = np.float32 if p.dtype == CG_F32 else np.float64
    scalars = np.zeros((N, 2, K, pitch), dtype=dt)
    buckets = np.zeros((N, 2, B, pitch), dtype=dt)
    scalars[..., :T] = scal
    buckets[..., :T] = buck

    if coincident_phase:
        ph = rng.choice(np.arange(0, p.period_ms, math.gcd(p.period_ms, p.tick_step_ms)), size=(N, 2))
    else:
        # never on a tick boundary, so 2.x and 3.x window semantics agree
        g = math.gcd(p.period_ms, p.tick_step_ms)
        ph = rng.integers(1, p.period_ms, size=(N, 2))
        ph = np.where(ph % g == 0, ph + 1, ph)
    if shared_phase:
        ph[:, 1] = ph[:, 0]
    return SynthBatch(scalars=scalars, buckets=buckets, le=le,
                      phase=ph.astype(np.int32), params=p)


@dataclass
class IrregularBatch:
    """Series with their own time stamps.  Only the first cnt samples of a
    row are meaningful; the rest is padding."""
    scalars: np.ndarray      # [N, 2, K, pitch]
    buckets: np.ndarray      # [N, 2, B, pitch]
    ts_scalars: np.ndarray   # [N, 2, K, pitch] int32 ms, strictly increasing within a row
    ts_buckets: np.ndarray   # [N, 2, B, pitch]
    cnt_scalars: np.ndarray  # [N, 2, K] int32
    cnt_buckets: np.ndarray  # [N, 2, B] int32
    le: np.ndarray           # [B]
    params: CgParams


def as_irregular(batch: SynthBatch) -> IrregularBatch:
    """The same grid series, spelled with explicit time stamps (results must not change)."""
    p = batch.params
    N, pitch, T = batch.phase.shape[0], p.row_pitch, p.n_samples
    grid = batch.phase[:, :, None, None].astype(np.int64) + np.arange(pitch, dtype=np.int64) * p.period_ms
    ts_s = np.broadcast_to(grid, (N, 2, p.n_scalar, pitch)).astype(np.int32)
    ts_b = np.broadcast_to(grid, (N, 2, p.n_buckets, pitch)).astype(np.int32)
    return IrregularBatch(scalars=batch.scalars, buckets=batch.buckets,
                          ts_scalars=np.ascontiguousarray(ts_s), ts_buckets=np.ascontiguousarray(ts_b),
                          cnt_scalars=np.full((N, 2, p.n_scalar), T, np.int32),
                          cnt_buckets=np.full((N, 2, p.n_buckets), T, np.int32), le=batch.le, params=p)


def make_irregular(batch: SynthBatch, seed: int = 1, jitter_ms: int = 2000, miss_prob: float = 0.05,
                   series_miss_prob: float = 0.01, gap_prob: float = 0.02, truncate_prob: float = 0.05,
                   dead_series_prob: float = 0.01) -> IrregularBatch:
    """Perturb a grid batch the way real scrapes deviate from a grid:
      * every scrape of a target is late or early by up to jitter_ms (all its series alike);
      * scrapes fail (miss_prob: the whole target; series_miss_prob: one series);
      * targets go away for a few scrapes (gap_prob starts a gap of 2..6 scrapes: windows with 2-3
        samples, and the avg*1.1 extrapolation threshold of Appendix A step 5 is crossed);
      * series start late or end early (truncate_prob) or have no samples at all (dead_series_prob)."""
    p = batch.params
    rng = np.random.default_rng(seed)
    N, T, pitch, K, B = batch.phase.shape[0], p.n_samples, p.row_pitch, p.n_scalar, p.n_buckets
    out = as_irregular(batch)
    vals = [np.zeros_like(batch.scalars), np.zeros_like(batch.buckets)]
    tss = [np.zeros_like(out.ts_scalars), np.zeros_like(out.ts_buckets)]
    cnts = [np.zeros_like(out.cnt_scalars), np.zeros_like(out.cnt_buckets)]
    half = max(1, min(jitter_ms, p.period_ms // 2 - 1))
    for i in range(N):
        for v in range(2):
            t_grid = int(batch.phase[i, v]) + np.arange(T, dtype=np.int64) * p.period_ms
            t_tgt = t_grid + rng.integers(-half, half + 1, size=T)
            keep_tgt = rng.random(T) >= miss_prob
            j = 0
            while j < T:
                if rng.random() < gap_prob:
                    g = int(rng.integers(2, 7))
                    keep_tgt[j:j + g] = False
                    j += g
                j += 1
            for which, (src, R) in enumerate(((batch.scalars, K), (batch.buckets, B))):
                for r in range(R):
                    keep = keep_tgt & (rng.random(T) >= series_miss_prob)
                    if rng.random() < truncate_prob:
                        a, b = sorted(rng.integers(0, T + 1, size=2))
                        keep[:a] = False
                        keep[b:] = False
                    if rng.random() < dead_series_prob:
                        keep[:] = False
                    idx = np.nonzero(keep)[0]
                    n = len(idx)
                    vals[which][i, v, r, :n] = src[i, v, r, idx]
                    tss[which][i, v, r, :n] = t_tgt[idx]
                    cnts[which][i, v, r] = n
    return IrregularBatch(scalars=vals[0], buckets=vals[1], ts_scalars=tss[0], ts_buckets=tss[1],
                          cnt_scalars=cnts[0], cnt_buckets=cnts[1], le=batch.le, params=p)
