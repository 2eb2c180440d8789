"""CPU-only checks of the drop-in boundary: the C-ABI library builds, loads, exports every
symbol include/canary_gate.h declares, and validates parameter blocks without a GPU."""
import ctypes as C
import os
import re

import pytest

import canary_b200  # noqa: F401
from canary_b200 import _lib, params as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "canary_gate.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cg_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    names = declared_functions()
    assert "cg_eval_fused" in names and "cg_eval_grid" in names and len(names) >= 10
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in include/canary_gate.h but not exported"


def test_params_struct_matches_header_defaults():
    p = P.CgParams()
    assert _lib.lib.cg_params_default(C.byref(p)) == 0
    q = P.make_params()
    assert bytes(p) == bytes(q), "params.make_params() and cg_params_default() disagree"
    assert p.struct_bytes == C.sizeof(P.CgParams)
    assert _lib.lib.cg_metric_width(C.byref(p)) == 6
    assert _lib.lib.cg_bytes_per_pair(C.byref(p)) == 261152 == p.bytes_per_pair()
    assert (p.one_plus_thr[0], p.one_plus_thr[1], p.one_plus_thr[2]) == (1.05, 1.02, 1.05)


@pytest.mark.parametrize("mut, frag", [
    (lambda p: setattr(p, "struct_bytes", 8), "ABI mismatch"),
    (lambda p: setattr(p, "row_pitch", 239), "row_pitch < n_samples"),
    (lambda p: setattr(p, "row_pitch", 241), "multiple of 16 bytes"),
    (lambda p: setattr(p, "n_scalar", 17), "n_scalar"),
    (lambda p: setattr(p, "period_ms", 0), "positive"),
    (lambda p: setattr(p, "n_q", 5), "n_q"),
    (lambda p: setattr(p, "gate_q", 1), "gate_q"),
    (lambda p: setattr(p, "checks", 8), "checks"),
    (lambda p: setattr(p, "max_attempts", 0), "schedule"),
    (lambda p: setattr(p, "reserved", 1), "reserved must be 0"),
])
def test_bad_params_are_rejected_with_a_message(mut, frag):
    p = P.make_params()
    mut(p)
    rc = _lib.lib.cg_params_check(C.byref(p))
    assert rc == 1
    assert frag in _lib.lib.cg_last_error().decode()


def test_null_params():
    assert _lib.lib.cg_params_check(None) == 1


def test_launch_info_reports_geometry_without_gpu():
    p = P.make_params()
    g, b, s = C.c_int32(), C.c_int32(), C.c_int32()
    assert _lib.lib.cg_launch_info(C.byref(p), 100000, C.byref(g), C.byref(b), C.byref(s)) == 0
    assert b.value % 32 == 0 and 0 < s.value <= 232448 and g.value >= 1


def test_library_reads_no_environment_and_has_no_debug_switches():
    """The shipped build has no tuning knobs: the ones the profiling tools use exist only in a
    CG_ABLATE developer build (build.py), and cg_params.reserved is rejected unless 0."""
    csrc = os.path.join(ROOT, "research-and-development-of-kubernetes-operator-for-machine-learning-pipelines_b200", "csrc")
    for name in os.listdir(csrc):
        text = open(os.path.join(csrc, name)).read()
        out, depth = [], 0
        for line in text.split("\n"):                       # drop #ifdef CG_ABLATE ... #endif blocks
            if line.strip().startswith("#ifdef CG_ABLATE"):
                depth += 1
            elif depth and line.strip().startswith("#endif"):
                depth -= 1
            elif not depth:
                out.append(line)
        body = "\n".join(out)
        assert "getenv" not in body, f"{name} reads the environment outside CG_ABLATE"
        assert "p.reserved" not in body and "p->reserved &" not in body, f"{name} branches on cg_params.reserved"
