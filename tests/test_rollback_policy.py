"""The opt-in rollback policy.  The reference has no rollback: after the 10th
failed attempt it logs, emits PromotionFailed and leaves the weights where they are
(mlflow_operator.py:341-349; :345 is a placeholder comment).  Default OFF must therefore stay
bit-identical to the reference; ON sends all traffic back to the old version."""
import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import params as P
from oracle import c_oracle as co, pyoracle as po


def _words(bits, n_words):
    w = np.zeros((1, n_words), np.uint32)
    for j, b in enumerate(bits):
        if b:
            w[0, j >> 5] |= np.uint32(1 << (j & 31))
    return w


SCRIPTS = {
    "all_fail": [False] * 60,
    "three_steps_then_fail": ([True] + [False] * 5) * 3 + [False] * 40,
    "all_pass": [True] * 60,
    "fail_nine_then_pass": [False] * 9 + [True] * 60,
    "runs_out_of_data_mid_retry": [True] + [False] * 8,
}


@pytest.mark.parametrize("name", sorted(SCRIPTS))
def test_oracles_agree_and_default_is_the_reference(name):
    bits = SCRIPTS[name]
    for rollback in (False, True):
        p = P.make_params(n_ticks=len(bits), rollback=rollback)
        c = co.replay(_words(bits, p.n_words), p)[0]
        py = po.replay_schedule(bits, rollback=rollback)
        assert tuple(c) == (py["cur"], py["prev"], py["state"], py["n_promotions"], py["n_evals"], py["last_tick"])
    off = co.replay(_words(bits, p.n_words), P.make_params(n_ticks=len(bits)))[0]
    on = co.replay(_words(bits, p.n_words), P.make_params(n_ticks=len(bits), rollback=True))[0]
    if off[2] == P.STATE_ABORTED:
        assert tuple(on[:3]) == (0, 100, P.STATE_ROLLED_BACK) and tuple(on[3:]) == tuple(off[3:])
    else:
        assert tuple(on) == tuple(off)            # the policy only changes what happens after an abort


def test_bad_policy_value_is_refused_by_the_library():
    import ctypes as C

    from canary_b200 import _lib
    p = P.make_params()
    p.on_abort = 7
    assert _lib.lib.cg_params_check(C.byref(p)) != 0 and b"on_abort" in _lib.lib.cg_last_error()


def test_patch_batch_for_a_rolled_back_canary():
    torch = pytest.importorskip("torch")
    from canary_b200 import operator_api as api
    dec = torch.tensor([[0, 100, 4, 2, 16, 40], [30, 70, 2, 2, 16, 40], [100, 0, 1, 9, 9, 48]], dtype=torch.int32)
    out = api.decisions_to_predictor_traffic(dec, [2, 2, 2], [1, 1, 1])
    assert out[0] == {"state": "rolled_back", "predictors": [{"name": "v1", "traffic": 100}]}
    assert out[1]["predictors"] == [{"name": "v1", "traffic": 70}, {"name": "v2", "traffic": 30}]
    assert out[2] == {"state": "complete", "predictors": [{"name": "v2", "traffic": 100}]}


@pytest.mark.gpu
def test_gpu_replay_and_fused_follow_the_policy():
    import torch

    from canary_b200 import engine, synth
    from gpu_util import upload
    p_on = P.make_params(n_samples=120, n_buckets=8, rollback=True)
    p_off = P.make_params(n_samples=120, n_buckets=8)
    b = synth.generate(p_off, 64, seed=9, adversarial=0.5)
    st = upload(b)
    for p in (p_off, p_on):
        want, bits = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p)
        for path in (1, 2):
            got = engine.eval_fused(st.scalars, st.buckets, st.le, st.phase, p, path=path).cpu().numpy()
            assert np.array_equal(got, want), (p.on_abort, path)
        got = engine.replay(torch.from_numpy(bits.view(np.int32)).cuda(), p).cpu().numpy()
        assert np.array_equal(got, want)
    assert (want[:, 2] == P.STATE_ROLLED_BACK).any() and not (want[:, 2] == P.STATE_ABORTED).any()
