"""ctypes binding of libcanarygate.so.  There is NO fallback: if the CUDA library is
missing and cannot be built, importing this module raises."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build
from .params import CgParams

_vp, _i64 = C.c_void_p, C.c_int64


def _load():
    path = os.environ.get("CG_LIB_PATH") or _build.LIB    # CG_LIB_PATH: developer aid, a differently tuned build
    if not os.path.exists(path):
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            raise ImportError(
                f"libcanarygate.so is not built ({path}) and building it failed: {e}. "
                "This package has no CPU or PyTorch fallback; run __graft_entry__.build().") from e
    lib = C.CDLL(path)
    lib.cg_abi_version.restype = C.c_int
    lib.cg_last_error.restype = C.c_char_p
    lib.cg_launch_count.restype = _i64
    lib.cg_bytes_per_pair.restype = _i64
    lib.cg_bytes_per_pair.argtypes = [_vp]
    lib.cg_params_default.argtypes = [_vp]
    lib.cg_params_check.argtypes = [_vp]
    lib.cg_metric_width.argtypes = [_vp]
    lib.cg_eval_grid.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]
    lib.cg_gate.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]
    lib.cg_replay.argtypes = [_vp, _vp, _vp, _i64, _vp, _vp, _vp]
    lib.cg_eval_fused.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp]
    lib.cg_synth_fill.argtypes = [_vp, _vp, _vp, _vp, _i64, _i64, _vp, C.c_uint64, C.c_double, C.c_double, _vp]
    lib.cg_eval_fused_path.argtypes = [C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp]
    lib.cg_stream_eligible.argtypes = [_vp]
    lib.cg_eval_irregular.argtypes = [_vp] * 9 + [_i64] + [_vp] * 6
    lib.cg_eval_irregular_path.argtypes = [C.c_int] + [_vp] * 9 + [_i64] + [_vp] * 6
    lib.cg_launch_info.argtypes = [_vp, _i64, _vp, _vp, _vp]
    lib.cg_store_append.argtypes = [_vp, _vp, _vp, _vp, _i64, _vp, C.c_int32, _vp]
    lib.cg_store_shift.argtypes = [_vp, _vp, _i64, _vp, C.c_int32, C.c_int32, _vp]
    if lib.cg_abi_version() != 1:
        raise ImportError("libcanarygate.so ABI version mismatch; rebuild")
    probe = CgParams()
    lib.cg_params_default(C.byref(probe))
    if probe.struct_bytes != C.sizeof(CgParams):
        raise ImportError("cg_params layout differs between params.py and the built library; rebuild")
    return lib


LIB_PATH = _build.LIB
lib = _load()


class CanaryGateError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        msg = lib.cg_last_error().decode("utf-8", "replace")
        raise CanaryGateError(f"{what} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(lib.cg_launch_count())
