"""The oracles (Python and C restatements) against SURVEY.md Appendix B KATs."""
import math

import numpy as np
import pytest

from golden_util import load_json
from oracle import c_oracle as co, pyoracle as po

KAT = load_json("kat_promql.json")


@pytest.mark.parametrize("k", KAT["increase"], ids=lambda k: k["id"])
def test_increase_kat(k):
    samples = [(int(t), float(v)) for t, v in k["samples"]]
    windows = [False, True] if k.get("both_windows") else [False]
    for left_open in windows:
        w = po.select_range(samples, k["t_ms"], k["range_ms"], left_open)
        got_py = po.extrapolated_increase(w, k["t_ms"], k["range_ms"])
        got_c = co.extrapolated_increase([s[0] for s in samples], [s[1] for s in samples],
                                         k["t_ms"], k["range_ms"], left_open)
        assert got_py == k["increase"]
        assert got_c == k["increase"]
        if "rate" in k:
            assert po.extrapolated_rate(w, k["t_ms"], k["range_ms"]) == k["rate"]


@pytest.mark.parametrize("k", KAT["quantile"], ids=lambda k: k["id"])
def test_quantile_kat(k):
    bs = [(float(a), float(b)) for a, b in k["buckets"]]
    le = [b[0] for b in bs]
    cnt = [b[1] for b in bs]
    for q, e in zip(k["q"], k["expect"]):
        e = float(e)
        for got in (po.bucket_quantile(q, bs), co.bucket_quantile(q, le, cnt)):
            if math.isnan(e):
                assert math.isnan(got)
            elif math.isinf(e):
                assert got == e
            else:
                assert got == pytest.approx(e, rel=1e-12, abs=1e-15)
        # the two restatements agree to the bit
        a, b = po.bucket_quantile(q, bs), co.bucket_quantile(q, le, cnt)
        assert (math.isnan(a) and math.isnan(b)) or a == b


def test_quantile_edge_rules():
    inf = math.inf
    # highest bound not +Inf -> NaN; fewer than 2 buckets -> NaN (Appendix A step 2)
    assert math.isnan(po.bucket_quantile(0.5, [(0.1, 1.0), (0.2, 2.0)]))
    assert math.isnan(co.bucket_quantile(0.5, [0.1, 0.2], [1.0, 2.0]))
    assert math.isnan(po.bucket_quantile(0.5, [(inf, 2.0)]))
    assert math.isnan(co.bucket_quantile(0.5, [inf], [2.0]))
    # non-monotonic counts are forced monotonic (step 3)
    bs = [(0.1, 10.0), (0.2, 5.0), (0.4, 20.0), (inf, 20.0)]
    a = po.bucket_quantile(0.6, bs)
    b = co.bucket_quantile(0.6, [x[0] for x in bs], [x[1] for x in bs])
    assert a == b == 0.2 + (0.4 - 0.2) * ((0.6 * 20 - 10) / 10)


def test_random_series_py_vs_c():
    rng = np.random.default_rng(5)
    for _ in range(300):
        n = int(rng.integers(1, 9))
        ts = np.sort(rng.choice(np.arange(0, 60001, 250), size=n, replace=False)).astype(np.int64)
        v = np.cumsum(rng.integers(0, 50, size=n)).astype(np.float64) + rng.integers(0, 1000)
        if n > 2 and rng.random() < 0.4:
            k = int(rng.integers(1, n))
            v[k:] -= v[k - 1]          # a reset
        for left_open in (False, True):
            w = po.select_range(list(zip(ts.tolist(), v.tolist())), 60000, 60000, left_open)
            a = po.extrapolated_increase(w, 60000, 60000)
            b = co.extrapolated_increase(ts, v, 60000, 60000, left_open)
            assert a == b
