"""Live check: the oracles under the UNMODIFIED reference.
Runs wherever the reference file is: /root/reference in the build container, or its untouched copy in oracle/_ref/ (made by `make -C oracle _ref`, git-ignored, shipped to the GPU box with the snapshot)."""
import math

import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import params as P, synth
from oracle import c_oracle as co, ref_loader as rl, soa_bridge as sb

pytestmark = pytest.mark.skipif(not rl.reference_available(), reason="no reference file (run make -C oracle _ref where /root/reference is mounted)")


def test_reference_imports_unmodified():
    mod = rl.load_reference()
    assert hasattr(mod, "get_model_metrics") and hasattr(mod, "should_promote_model")
    assert repr(1 + 0.05) == "1.05" and repr(1 + 0.02) == "1.02"


@pytest.mark.parametrize("left_open", [False, True])
def test_c_oracle_equals_reference_on_fresh_pairs(left_open):
    p = P.make_params(n_samples=160, n_buckets=16, left_open=left_open)
    b = synth.generate(p, 10, seed=99, adversarial=0.3, coincident_phase=left_open)
    dec, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p)
    m, v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p)
    for i in range(10):
        r = sb.run_reference_on_pair(b, i, left_open=left_open)
        assert (r["cur"], r["prev"], r["state"], r["n_promotions"], r["n_evals"]) == tuple(dec[i, :5])
    for i in range(3):
        db = sb.pair_to_db(b, i, left_open=left_open)
        for j in range(0, p.n_ticks, 9):
            clock = rl.VirtualClock(p.tick0_ms + j * p.tick_step_ms)
            fp = rl.FakePrometheus(db, clock)
            for ver, pred in enumerate(("v2", "v1")):
                md = rl.reference_get_model_metrics(fp, "m", pred, "ns")
                rec, valid = sb.metrics_dict_to_record(md)
                assert valid == v[i, ver, j]
                for a, c in zip(rec, m[i, ver, j]):
                    assert (math.isnan(a) and math.isnan(c)) or a == c


def test_no_previous_version_means_no_canary():
    clock = rl.VirtualClock(0)
    out = rl.run_reference_canary(lambda: None, clock, previous_version=None)
    assert out["weights"] == [[100]] and out["end_ms"] == 0       # :188-191, :289
    p = P.make_params()
    dec = co.replay(np.zeros((1, p.n_words), np.uint32), p, has_previous=np.zeros(1, np.uint8))[0]
    assert tuple(dec) == (100, 0, P.STATE_NO_CANARY, 0, 0, -1)
