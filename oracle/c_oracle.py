"""ctypes wrapper around oracle/libpromql_oracle.so (the C restatement).

TEST INFRASTRUCTURE ONLY (see promql_oracle.c header).  Operates on numpy host
arrays laid out exactly like the device arrays of the product library.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libpromql_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "promql_oracle.c")
    hdr = os.path.join(_HERE, "..", "include", "canary_gate.h")
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(f) and os.path.getmtime(f) > os.path.getmtime(_SO) for f in (src, hdr))
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libpromql_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.cgo_bucket_quantile.restype = C.c_double
        _lib.cgo_bucket_quantile.argtypes = [C.c_double, C.c_void_p, C.c_void_p, C.c_int]
        _lib.cgo_extrapolated_increase.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64,
                                                   C.c_int64, C.c_int, C.c_void_p]
        _lib.cgo_eval_grid.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_void_p,
                                                          C.c_void_p, C.c_int]
        _lib.cgo_gate.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_void_p]
        _lib.cgo_replay.argtypes = [C.c_void_p] * 3 + [C.c_int64, C.c_void_p, C.c_void_p]
        _lib.cgo_eval_fused.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_void_p, C.c_void_p,
                                                           C.c_void_p, C.c_int]
        _lib.cgo_eval_lazy.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_void_p, C.c_void_p, C.c_int]
        _lib.cgo_eval_irregular.argtypes = [C.c_void_p] * 9 + [C.c_int64] + [C.c_void_p] * 5 + [C.c_int]
    return _lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _chk(rc, what):
    if rc != 0:
        raise RuntimeError(f"oracle {what} failed: code {rc}")


def max_threads() -> int:
    return lib().cgo_max_threads()


def extrapolated_increase(ts_ms, values, t_ms, range_ms, left_open=False):
    ts = np.ascontiguousarray(ts_ms, dtype=np.int64)
    v = np.ascontiguousarray(values, dtype=np.float64)
    out = C.c_double(0.0)
    ok = lib().cgo_extrapolated_increase(_ptr(ts), _ptr(v), len(ts), int(t_ms), int(range_ms),
                                         int(left_open), C.byref(out))
    return out.value if ok else None


def bucket_quantile(q, le, counts):
    le = np.ascontiguousarray(le, dtype=np.float64)
    cnt = np.array(counts, dtype=np.float64)
    return lib().cgo_bucket_quantile(float(q), _ptr(le), _ptr(cnt), len(le))


def _np_dtype(p):
    return np.float32 if p.dtype == 1 else np.float64


def eval_grid(scalars, buckets, le, phase, p, n_threads=1):
    n_pairs = phase.shape[0]
    W = p.n_q + 5
    metrics = np.empty((n_pairs, 2, p.n_ticks, W), dtype=np.float64)
    valid = np.empty((n_pairs, 2, p.n_ticks), dtype=np.uint8)
    _chk(lib().cgo_eval_grid(_ptr(scalars), _ptr(buckets), _ptr(le), _ptr(phase), n_pairs,
                             C.byref(p), _ptr(metrics), _ptr(valid), n_threads), "eval_grid")
    return metrics, valid


def gate(new_m, new_v, old_m, old_v, p):
    n = new_v.shape[0]
    detail = np.empty(n, dtype=np.uint8)
    _chk(lib().cgo_gate(_ptr(np.ascontiguousarray(new_m)), _ptr(np.ascontiguousarray(new_v)),
                        _ptr(np.ascontiguousarray(old_m)), _ptr(np.ascontiguousarray(old_v)),
                        n, C.byref(p), _ptr(detail)), "gate")
    return detail


def replay(gate_bits, p, has_previous=None, start_tick=None):
    n_pairs = gate_bits.shape[0]
    dec = np.empty((n_pairs, 6), dtype=np.int32)
    _chk(lib().cgo_replay(_ptr(np.ascontiguousarray(gate_bits, dtype=np.uint32)), _ptr(has_previous),
                          _ptr(start_tick), n_pairs, C.byref(p), _ptr(dec)), "replay")
    return dec


def eval_fused(scalars, buckets, le, phase, p, has_previous=None, start_tick=None,
               n_threads=1, want_bits=True):
    n_pairs = phase.shape[0]
    dec = np.empty((n_pairs, 6), dtype=np.int32)
    bits = np.zeros((n_pairs, (p.n_ticks + 31) // 32), dtype=np.uint32) if want_bits else None
    _chk(lib().cgo_eval_fused(_ptr(scalars), _ptr(buckets), _ptr(le), _ptr(phase),
                              _ptr(has_previous), _ptr(start_tick), n_pairs, C.byref(p),
                              _ptr(dec), _ptr(bits), n_threads), "eval_fused")
    return dec, bits


def eval_lazy(scalars, buckets, le, phase, p, has_previous=None, start_tick=None, n_threads=1):
    n_pairs = phase.shape[0]
    dec = np.empty((n_pairs, 6), dtype=np.int32)
    _chk(lib().cgo_eval_lazy(_ptr(scalars), _ptr(buckets), _ptr(le), _ptr(phase),
                             _ptr(has_previous), _ptr(start_tick), n_pairs, C.byref(p),
                             _ptr(dec), n_threads), "eval_lazy")
    return dec


def eval_irregular(scalars, buckets, ts_scalars, ts_buckets, cnt_scalars, cnt_buckets, le, p,
                   has_previous=None, start_tick=None, n_threads=1, want_metrics=False):
    """Irregular series (per-series int32 ms time stamps + sample counts): decisions, gate bits and,
    on request, the per-tick metric records."""
    n_pairs = cnt_scalars.shape[0]
    dec = np.empty((n_pairs, 6), dtype=np.int32)
    bits = np.zeros((n_pairs, (p.n_ticks + 31) // 32), dtype=np.uint32)
    W = p.n_q + 5
    m = np.empty((n_pairs, 2, p.n_ticks, W), dtype=np.float64) if want_metrics else None
    v = np.empty((n_pairs, 2, p.n_ticks), dtype=np.uint8) if want_metrics else None
    _chk(lib().cgo_eval_irregular(_ptr(scalars), _ptr(buckets), _ptr(ts_scalars), _ptr(ts_buckets),
                                  _ptr(cnt_scalars), _ptr(cnt_buckets), _ptr(le), _ptr(has_previous),
                                  _ptr(start_tick), n_pairs, C.byref(p), _ptr(dec), _ptr(bits), _ptr(m), _ptr(v),
                                  n_threads), "eval_irregular")
    return (dec, bits, m, v) if want_metrics else (dec, bits)
