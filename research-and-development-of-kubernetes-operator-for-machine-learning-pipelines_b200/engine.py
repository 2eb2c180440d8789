"""Batched device API: torch tensors in, torch tensors out, every FLOP in
libcanarygate.so.  torch is plumbing here (device memory, streams), not compute."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib
from .params import CG_F32, CgParams


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _want(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (no CPU fallback in this library)")
    if t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


def _elem_dtype(p: CgParams):
    return torch.float32 if p.dtype == CG_F32 else torch.float64


def _check_series(scalars, buckets, le, phase, p: CgParams) -> int:
    n = phase.shape[0]
    _want(phase, torch.int32, "phase")
    if tuple(phase.shape) != (n, 2):
        raise ValueError("phase must be [n_pairs, 2]")
    dt = _elem_dtype(p)
    if p.n_scalar:
        _want(scalars, dt, "scalars")
        if tuple(scalars.shape) != (n, 2, p.n_scalar, p.row_pitch):
            raise ValueError(f"scalars must be [{n}, 2, {p.n_scalar}, {p.row_pitch}], got {tuple(scalars.shape)}")
    if p.n_buckets:
        _want(buckets, dt, "buckets")
        if tuple(buckets.shape) != (n, 2, p.n_buckets, p.row_pitch):
            raise ValueError(f"buckets must be [{n}, 2, {p.n_buckets}, {p.row_pitch}], got {tuple(buckets.shape)}")
        _want(le, torch.float64, "le")
        if le.numel() != p.n_buckets:
            raise ValueError("le must have n_buckets entries")
    return n


def eval_grid(scalars, buckets, le, phase, p: CgParams) -> Tuple[torch.Tensor, torch.Tensor]:
    """Batched get_model_metrics (mlflow_operator.py:363-417) on the dense tick grid.
    -> metrics [N, 2, n_ticks, n_q+5] fp64, valid [N, 2, n_ticks] uint8."""
    n = _check_series(scalars, buckets, le, phase, p)
    metrics = torch.empty((n, 2, p.n_ticks, p.metric_width), dtype=torch.float64, device=phase.device)
    valid = torch.empty((n, 2, p.n_ticks), dtype=torch.uint8, device=phase.device)
    _lib.check(_lib.lib.cg_eval_grid(_ptr(scalars if p.n_scalar else None), _ptr(buckets if p.n_buckets else None),
                                     _ptr(le if p.n_buckets else None), _ptr(phase), n, C.byref(p),
                                     _ptr(metrics), _ptr(valid), _stream_ptr()), "cg_eval_grid")
    return metrics, valid


def gate(new_m, new_v, old_m, old_v, p: CgParams) -> torch.Tensor:
    """Batched should_promote_model (mlflow_operator.py:419-460) -> detail bytes."""
    n = new_v.numel()
    for t, dt, nm in ((new_m, torch.float64, "new_m"), (old_m, torch.float64, "old_m"),
                      (new_v, torch.uint8, "new_v"), (old_v, torch.uint8, "old_v")):
        _want(t, dt, nm)
    if new_m.numel() != n * p.metric_width or old_m.numel() != n * p.metric_width:
        raise ValueError("metric records must be [n, n_q+5]")
    detail = torch.empty(n, dtype=torch.uint8, device=new_m.device)
    _lib.check(_lib.lib.cg_gate(_ptr(new_m), _ptr(new_v), _ptr(old_m), _ptr(old_v), n, C.byref(p),
                                _ptr(detail), _stream_ptr()), "cg_gate")
    return detail


def replay(gate_bits, p: CgParams, has_previous=None, start_tick=None, out=None) -> torch.Tensor:
    """The traffic step loop (mlflow_operator.py:296-361) over gate bit-vectors
    -> decisions [N, 6] int32 (cur, prev, state, n_promotions, n_evals, last_tick)."""
    _want(gate_bits, torch.int32, "gate_bits")
    n = gate_bits.shape[0]
    if gate_bits.shape[1] != p.n_words:
        raise ValueError("gate_bits must be [n_pairs, ceil(n_ticks/32)] int32")
    if has_previous is not None:
        _want(has_previous, torch.uint8, "has_previous")
    if start_tick is not None:
        _want(start_tick, torch.int32, "start_tick")
    dec = out if out is not None else torch.empty((n, 6), dtype=torch.int32, device=gate_bits.device)
    _want(dec, torch.int32, "decisions")
    _lib.check(_lib.lib.cg_replay(_ptr(gate_bits), _ptr(has_previous), _ptr(start_tick), n, C.byref(p),
                                  _ptr(dec), _stream_ptr()), "cg_replay")
    return dec


def eval_fused(scalars, buckets, le, phase, p: CgParams, has_previous=None, start_tick=None,
               want_bits: bool = False, want_metrics: bool = False, out=None, path: int = 0):
    """metrics -> gate -> step loop in one pass over the series (mlflow_operator.py:302-317
    for every MlflowModel at once).  -> decisions [N, 6] int32 (+ gate_bits, metrics, valid).
    path: 0 = let the library choose, 1 = resident-tile kernel, 2 = streaming kernel, 3 = row-streaming kernel."""
    n = _check_series(scalars, buckets, le, phase, p)
    dev = phase.device
    if has_previous is not None:
        _want(has_previous, torch.uint8, "has_previous")
    if start_tick is not None:
        _want(start_tick, torch.int32, "start_tick")
    dec = out if out is not None else torch.empty((n, 6), dtype=torch.int32, device=dev)
    _want(dec, torch.int32, "decisions")
    if dec.numel() < n * 6:
        raise ValueError("decisions buffer too small")
    bits = torch.empty((n, p.n_words), dtype=torch.int32, device=dev) if want_bits else None
    metrics = valid = None
    if want_metrics:
        metrics = torch.empty((n, 2, p.n_ticks, p.metric_width), dtype=torch.float64, device=dev)
        valid = torch.empty((n, 2, p.n_ticks), dtype=torch.uint8, device=dev)
    _lib.check(_lib.lib.cg_eval_fused_path(path, _ptr(scalars if p.n_scalar else None),
                                           _ptr(buckets if p.n_buckets else None),
                                           _ptr(le if p.n_buckets else None), _ptr(phase), _ptr(has_previous),
                                           _ptr(start_tick), n, C.byref(p), _ptr(dec), _ptr(bits), _ptr(metrics),
                                           _ptr(valid), _stream_ptr()), "cg_eval_fused")
    res = [dec]
    if want_bits:
        res.append(bits)
    if want_metrics:
        res += [metrics, valid]
    return res[0] if len(res) == 1 else tuple(res)


def eval_irregular(scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le, p: CgParams,
                   has_previous=None, start_tick=None, want_bits: bool = False, want_metrics: bool = False,
                   path: int = 0):
    """The fused step for irregular series: every series has its own int32 ms
    time stamps `ts_*[N, 2, K|B, pitch]` and sample count `cnt_*[N, 2, K|B]`; stale markers already
    dropped (ingest.build_irregular_batch does it).  -> decisions [N, 6] (+ gate_bits, metrics, valid).
    path: 0 = let the library choose, 1 = resident-tile kernel, 3 = row-streaming kernel."""
    n = int(cnt_scalars.shape[0])
    dev = cnt_scalars.device
    if dev.type != "cuda":
        raise ValueError("series must live on the GPU; there is no CPU fallback")
    vdt = torch.float32 if p.dtype == 1 else torch.float64
    for t, dt, shape, what in ((scalars, vdt, (n, 2, p.n_scalar, p.row_pitch), "scalars"),
                               (ts_scalars, torch.int32, (n, 2, p.n_scalar, p.row_pitch), "ts_scalars"),
                               (cnt_scalars, torch.int32, (n, 2, p.n_scalar), "cnt_scalars")) + \
            (((buckets, vdt, (n, 2, p.n_buckets, p.row_pitch), "buckets"),
              (ts_buckets, torch.int32, (n, 2, p.n_buckets, p.row_pitch), "ts_buckets"),
              (cnt_buckets, torch.int32, (n, 2, p.n_buckets), "cnt_buckets")) if p.n_buckets else ()):
        _want(t, dt, what)
        if tuple(t.shape) != shape:
            raise ValueError(f"{what}: shape {tuple(t.shape)}, expected {shape}")
    if p.n_buckets:
        _want(le, torch.float64, "le")
    if has_previous is not None:
        _want(has_previous, torch.uint8, "has_previous")
    if start_tick is not None:
        _want(start_tick, torch.int32, "start_tick")
    dec = torch.empty((n, 6), dtype=torch.int32, device=dev)
    bits = torch.empty((n, p.n_words), dtype=torch.int32, device=dev) if want_bits else None
    metrics = valid = None
    if want_metrics:
        metrics = torch.empty((n, 2, p.n_ticks, p.metric_width), dtype=torch.float64, device=dev)
        valid = torch.empty((n, 2, p.n_ticks), dtype=torch.uint8, device=dev)
    nb = bool(p.n_buckets)
    _lib.check(_lib.lib.cg_eval_irregular_path(int(path), _ptr(scalars), _ptr(buckets if nb else None), _ptr(ts_scalars),
                                               _ptr(ts_buckets if nb else None), _ptr(cnt_scalars),
                                               _ptr(cnt_buckets if nb else None), _ptr(le if nb else None),
                                               _ptr(has_previous), _ptr(start_tick), n, C.byref(p), _ptr(dec), _ptr(bits),
                                               _ptr(metrics), _ptr(valid), _stream_ptr()), "cg_eval_irregular")
    res = [dec]
    if want_bits:
        res.append(bits)
    if want_metrics:
        res += [metrics, valid]
    return res[0] if len(res) == 1 else tuple(res)


def stream_eligible(p: CgParams) -> bool:
    """True if cg_eval_fused takes the streaming kernel for these params."""
    return bool(_lib.lib.cg_stream_eligible(C.byref(p)))


def synth_fill(n_pairs: int, p: CgParams, le: torch.Tensor, seed: int = 0xC0FFEE, pair_offset: int = 0,
               reset_prob: float = 1.0 / 240.0, adversarial: float = 0.01, device=None):
    """Allocate and fill device series with the library's seeded generator kernel
    (benchmark / test utility).  -> scalars, buckets, phase."""
    dev = le.device if device is None else torch.device(device)
    dt = _elem_dtype(p)
    scalars = torch.empty((n_pairs, 2, p.n_scalar, p.row_pitch), dtype=dt, device=dev)
    buckets = torch.empty((n_pairs, 2, p.n_buckets, p.row_pitch), dtype=dt, device=dev)
    phase = torch.empty((n_pairs, 2), dtype=torch.int32, device=dev)
    _lib.check(_lib.lib.cg_synth_fill(_ptr(scalars), _ptr(buckets), _ptr(le), _ptr(phase), n_pairs, pair_offset,
                                      C.byref(p), seed, reset_prob, adversarial, _stream_ptr()), "cg_synth_fill")
    return scalars, buckets, phase


class HostPipeline:
    """End-to-end path for HOST-resident series (what an ingest layer hands over): pinned
    host arrays are streamed to the GPU in chunks of pairs on a copy stream while the
    previous chunk is evaluated on a compute stream; decision records come back to pinned
    host memory.  Staging buffers are allocated once here and reused by every `run`."""

    def __init__(self, p: CgParams, chunk_pairs: int = 4096, device=None):
        self.p = p
        self.chunk = int(chunk_pairs)
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        dt = _elem_dtype(p)
        mk = lambda *s, d=dt: torch.empty(s, dtype=d, device=self.dev)  # noqa: E731
        self.slots = [dict(scalars=mk(self.chunk, 2, p.n_scalar, p.row_pitch),
                           buckets=mk(self.chunk, 2, p.n_buckets, p.row_pitch),
                           phase=mk(self.chunk, 2, d=torch.int32),
                           dec=mk(self.chunk, 6, d=torch.int32),
                           copied=torch.cuda.Event(), done=torch.cuda.Event()) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self.compute_stream = torch.cuda.Stream(device=self.dev)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def run(self, scalars_h, buckets_h, phase_h, le_dev, out_h=None):
        """scalars_h/buckets_h/phase_h: pinned CPU tensors in the library layout.
        -> decisions [N, 6] int32 in pinned host memory (synchronised on return)."""
        p = self.p
        n = phase_h.shape[0]
        for t, nm in ((scalars_h, "scalars_h"), (buckets_h, "buckets_h"), (phase_h, "phase_h")):
            if t.is_cuda or not t.is_contiguous():
                raise ValueError(f"{nm} must be a contiguous CPU tensor")
        if out_h is None:
            out_h = torch.empty((n, 6), dtype=torch.int32, pin_memory=True)
        self.h2d_bytes = self.d2h_bytes = 0
        cur = torch.cuda.current_stream(self.dev)
        self.copy_stream.wait_stream(cur)
        self.compute_stream.wait_stream(cur)
        for c, lo in enumerate(range(0, n, self.chunk)):
            hi = min(lo + self.chunk, n)
            m = hi - lo
            s = self.slots[c & 1]
            with torch.cuda.stream(self.copy_stream):
                if c >= 2:
                    self.copy_stream.wait_event(s["done"])      # slot free again
                s["scalars"][:m].copy_(scalars_h[lo:hi], non_blocking=True)
                s["buckets"][:m].copy_(buckets_h[lo:hi], non_blocking=True)
                s["phase"][:m].copy_(phase_h[lo:hi], non_blocking=True)
                s["copied"].record(self.copy_stream)
            self.h2d_bytes += sum(x[lo:hi].numel() * x.element_size() for x in (scalars_h, buckets_h, phase_h))
            with torch.cuda.stream(self.compute_stream):
                self.compute_stream.wait_event(s["copied"])
                eval_fused(s["scalars"][:m], s["buckets"][:m], le_dev, s["phase"][:m], p, out=s["dec"][:m])
                out_h[lo:hi].copy_(s["dec"][:m], non_blocking=True)
                s["done"].record(self.compute_stream)
            self.d2h_bytes += m * 24
        cur.wait_stream(self.compute_stream)
        self.compute_stream.synchronize()
        return out_h


class RollingStore:
    """Device-resident scrape history that grows one scrape at a time.

    The reference re-asks Prometheus for "now" on every attempt (mlflow_operator.py:302-303) and
    sleeps 10 s or 60 s in between (:340, :352); what changes between two attempts is one scrape
    per series, not the hour.  Here the hour stays in HBM: `append_scrape` moves one value per
    series over PCIe and `evaluate` runs the canary over the newest `p.n_samples` samples.

    Rows have `slack` spare columns after the newest sample; when they are used up every row is
    moved left once (cg_store_shift), so the amortised cost of an append is one column write.
    Times are relative to column 0: `phase[n, 2]` is the scrape offset of each target within a
    period and does not change; the tick grid follows the newest samples.
    """

    def __init__(self, p: CgParams, n_pairs: int, le: torch.Tensor, slack: int = 64, device=None):
        self.dev = le.device if device is None else torch.device(device)
        self.p, self.n, self.le = p.copy(), int(n_pairs), le
        self.T = p.n_samples
        esz = p.elem_bytes
        self.align = 16 // esz                                   # samples per 16 bytes: bulk copies start on 16-byte boundaries
        cap = self.T + max(int(slack), self.align)
        self.cap = -(-cap // self.align) * self.align
        dt = _elem_dtype(p)
        K, B = p.n_scalar, p.n_buckets
        # one spare row at the end: a view that starts `offset` samples into the rows is read row by row
        self._sc = torch.zeros(self.n * 2 * K * self.cap + self.cap, dtype=dt, device=self.dev)
        self._bk = torch.zeros(self.n * 2 * B * self.cap + self.cap, dtype=dt, device=self.dev)
        self.scalars = self._sc[: self.n * 2 * K * self.cap].view(self.n, 2, K, self.cap)
        self.buckets = self._bk[: self.n * 2 * B * self.cap].view(self.n, 2, B, self.cap)
        self.phase = torch.zeros((self.n, 2), dtype=torch.int32, device=self.dev)
        # two staging buffers and a copy stream: the upload of scrape j+1 overlaps the evaluation of scrape j
        self._new_sc = [torch.empty((self.n, 2, K), dtype=dt, device=self.dev) for _ in range(2)]
        self._new_bk = [torch.empty((self.n, 2, B), dtype=dt, device=self.dev) for _ in range(2)]
        self._copy_stream = torch.cuda.Stream(device=self.dev)
        self._ready = [torch.cuda.Event() for _ in range(2)]     # staging buffer filled (copy stream)
        self._used = [None, None]                                # staging buffer consumed (caller's stream)
        self._stage = 0
        self._phase_eff = torch.empty_like(self.phase)
        self.count = 0
        self.h2d_bytes = 0
        self._pp = p.copy()
        self._pp.row_pitch = self.cap

    def load(self, scalars: torch.Tensor, buckets: torch.Tensor, phase: torch.Tensor):
        """Bulk start: device tensors [n, 2, K|B, m] with m <= capacity samples, and the phases."""
        m = scalars.shape[-1] if self.p.n_scalar else buckets.shape[-1]
        if m > self.cap:
            raise ValueError("more samples than the store holds")
        if self.p.n_scalar:
            self.scalars[..., :m].copy_(scalars)
        if self.p.n_buckets:
            self.buckets[..., :m].copy_(buckets)
        self.phase.copy_(phase)
        self.count = m

    def append_scrape(self, new_scalars: torch.Tensor, new_buckets: torch.Tensor):
        """One new scrape of every series: [n, 2, K] and [n, 2, B], pinned host or device memory.
        Host tensors are uploaded on the store's copy stream (they must stay untouched until the
        caller's stream has passed this call); everything else is enqueued on the current stream."""
        p = self._pp
        if self.count == self.cap:                               # slack used up: the oldest samples fall off
            keep = self.T - 1
            _lib.check(_lib.lib.cg_store_shift(_ptr(self.scalars if p.n_scalar else None),
                                               _ptr(self.buckets if p.n_buckets else None), self.n, C.byref(p),
                                               self.count - keep, keep, _stream_ptr()), "cg_store_shift")
            self.count = keep
        self.h2d_bytes = 0
        st = self._stage
        self._stage ^= 1
        cur = torch.cuda.current_stream(self.dev)
        stage_sc, stage_bk = self._new_sc[st], self._new_bk[st]
        from_host = any(t.numel() and not src.is_cuda for src, t in ((new_scalars, stage_sc), (new_buckets, stage_bk)))
        side = self._copy_stream if from_host else cur
        if from_host and self._used[st] is not None:
            side.wait_event(self._used[st])                      # the append that last read this staging buffer
        with torch.cuda.stream(side):
            for src, dst in ((new_scalars, stage_sc), (new_buckets, stage_bk)):
                if dst.numel():
                    dst.copy_(src, non_blocking=True)
                    if not src.is_cuda:
                        self.h2d_bytes += src.numel() * src.element_size()
            if from_host:
                self._ready[st].record(side)
        if from_host:
            cur.wait_event(self._ready[st])
        _lib.check(_lib.lib.cg_store_append(_ptr(self.scalars if p.n_scalar else None),
                                            _ptr(self.buckets if p.n_buckets else None),
                                            _ptr(stage_sc if p.n_scalar else None),
                                            _ptr(stage_bk if p.n_buckets else None), self.n, C.byref(p),
                                            self.count, _stream_ptr()), "cg_store_append")
        if from_host:
            ev = torch.cuda.Event()
            ev.record(cur)
            self._used[st] = ev
        self.count += 1

    def view(self):
        """What `evaluate` hands to the kernels: (first column, samples in view, params, tick0 shift in
        periods).  The view starts on a 16-byte boundary at or before the newest T samples."""
        first = max(0, self.count - self.T)
        off = first - first % self.align
        pp = self._pp.copy()
        pp.n_samples = self.count - off
        pp.tick0_ms = self.p.tick0_ms + first * self.p.period_ms
        return off, self.count - off, pp, first

    def evaluate(self, out: Optional[torch.Tensor] = None, has_previous=None, start_tick=None, want_bits: bool = False):
        """The whole canary (metrics on the dense tick grid -> gate -> step loop) over the newest
        samples of every pair.  -> decisions [n, 6] int32 (and the gate bits)."""
        if self.count < 2:
            raise ValueError("fewer than two scrapes in the store")
        off, _, pp, _ = self.view()
        esz = pp.elem_bytes
        torch.add(self.phase, off * pp.period_ms, out=self._phase_eff)
        dec = out if out is not None else torch.empty((self.n, 6), dtype=torch.int32, device=self.dev)
        _want(dec, torch.int32, "decisions")
        bits = torch.empty((self.n, pp.n_words), dtype=torch.int32, device=self.dev) if want_bits else None
        sc = C.c_void_p(self.scalars.data_ptr() + off * esz) if pp.n_scalar else None
        bk = C.c_void_p(self.buckets.data_ptr() + off * esz) if pp.n_buckets else None
        _lib.check(_lib.lib.cg_eval_fused(sc, bk, _ptr(self.le if pp.n_buckets else None), _ptr(self._phase_eff),
                                          _ptr(has_previous), _ptr(start_tick), self.n, C.byref(pp), _ptr(dec),
                                          _ptr(bits), None, None, _stream_ptr()), "cg_eval_fused")
        return (dec, bits) if want_bits else dec
