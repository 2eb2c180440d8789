"""Helpers to read tests/golden/*.json (see tests/golden/make_golden.py)."""
import json
import math
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_json(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def fdec(x):
    if x is None:
        return None
    if isinstance(x, (int, float)):
        return float(x)
    if x == "nan":
        return math.nan
    if x == "inf":
        return math.inf
    if x == "-inf":
        return -math.inf
    return float.fromhex(x)


def same_float(a, b):
    """Bit-level equality, NaN == NaN."""
    if a is None or b is None:
        return a is None and b is None
    if math.isnan(a) or math.isnan(b):
        return math.isnan(a) and math.isnan(b)
    return a == b


def load_canary_batches():
    """-> {geometry: SynthBatch-like}, list of cases."""
    import canary_b200  # noqa: F401
    from canary_b200 import params as P, synth
    arrays = np.load(os.path.join(GOLDEN, "ref_canary_cases.npz"))
    doc = load_json("ref_canary_cases.json")
    batches = {}
    for case in doc["cases"]:
        g = case["geometry"]
        if g in batches:
            continue
        p = P.make_params(**case["make_params"])
        batches[g] = synth.SynthBatch(
            scalars=np.ascontiguousarray(arrays[f"{g}_scalars"]),
            buckets=np.ascontiguousarray(arrays[f"{g}_buckets"]),
            le=np.ascontiguousarray(arrays[f"{g}_le"]),
            phase=np.ascontiguousarray(arrays[f"{g}_phase"]), params=p)
    return batches, doc["cases"]


def load_irregular_batches():
    """-> {geometry: synth.IrregularBatch}, list of cases (tests/golden/ref_irregular_cases.*)."""
    import canary_b200  # noqa: F401
    from canary_b200 import params as P, synth
    arrays = np.load(os.path.join(GOLDEN, "ref_irregular_cases.npz"))
    doc = load_json("ref_irregular_cases.json")
    batches = {}
    for case in doc["cases"]:
        g = case["geometry"]
        if g not in batches:
            fields = {f: np.ascontiguousarray(arrays[f"{g}_{f}"]) for f in
                      ("scalars", "buckets", "ts_scalars", "ts_buckets", "cnt_scalars", "cnt_buckets", "le")}
            batches[g] = synth.IrregularBatch(params=P.make_params(**case["make_params"]), **fields)
    return batches, doc["cases"]


def golden_record(md, n_q=1):
    """Golden metrics dict (hex floats / None) -> (record[n_q+5], valid)."""
    rec = np.full(n_q + 5, np.nan)
    valid = 0
    v = fdec(md["latency_95th"])
    if v is not None:
        rec[0] = v
        valid |= 1
    rec[n_q + 0] = fdec(md["error_responses"])
    v = fdec(md["error_rate"])
    if v is not None:
        rec[n_q + 1] = v
        valid |= 2
    v = fdec(md["latency_avg"])
    if v is not None:
        rec[n_q + 2] = v
        valid |= 4
    rec[n_q + 3] = fdec(md["request_count"])
    rec[n_q + 4] = fdec(md["feedback_request_count"])
    return rec, valid


def records_equal(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))
