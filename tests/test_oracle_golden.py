"""The C oracle (full-scale checker) against the committed golden fixtures, which were
produced by the UNMODIFIED reference (tests/golden/make_golden.py).  Runs anywhere."""
import numpy as np
import pytest

import canary_b200  # noqa: F401
from canary_b200 import params as P
from golden_util import (fdec, golden_record, load_canary_batches, load_json, records_equal)
from oracle import c_oracle as co, pyoracle as po

BATCHES, CASES = load_canary_batches()


def test_gate_probes_c_and_py():
    doc = load_json("ref_gate_probes.json")
    p = P.make_params()
    for c in doc["cases"]:
        new = {k: fdec(v) for k, v in c["new"].items()}
        old = {k: fdec(v) for k, v in c["old"].items()}
        assert po.should_promote(new, old, doc["thresholds"]) == c["promote"], c["name"]

        def rec(m):
            r = np.full(6, np.nan)
            v = 0
            if m["latency_95th"] is not None:
                r[0], v = m["latency_95th"], v | 1
            if m["error_rate"] is not None:
                r[2], v = m["error_rate"], v | 2
            if m["latency_avg"] is not None:
                r[3], v = m["latency_avg"], v | 4
            return r, v
        (nr, nv), (orr, ov) = rec(new), rec(old)
        d = co.gate(nr[None], np.array([nv], np.uint8), orr[None], np.array([ov], np.uint8), p)
        assert bool(d[0] & P.G_PROMOTE) == c["promote"], c["name"]


def test_schedule_probes_c_and_py():
    doc = load_json("ref_schedule_probes.json")
    for c in doc["cases"]:
        bits = c["gate_bits"]
        p = P.make_params(n_ticks=len(bits))
        words = np.zeros((1, p.n_words), np.uint32)
        for j, b in enumerate(bits):
            if b:
                words[0, j >> 5] |= np.uint32(1 << (j & 31))
        dec = co.replay(words, p)[0]
        pyd = po.replay_schedule(bits)
        last = c["weights"][-1]
        if len(last) == 1:
            cur, prev, state = last[0], 0, P.STATE_COMPLETE
        else:
            (prev, cur), state = last, P.STATE_ABORTED
        n_prom = sum(1 for w in c["weights"][1:] if len(w) == 2)
        want = (cur, prev, state, n_prom, len(c["eval_ticks"]), c["eval_ticks"][-1])
        assert tuple(dec) == want, c["name"]
        assert (pyd["cur"], pyd["prev"], pyd["state"], pyd["n_promotions"], pyd["n_evals"],
                pyd["last_tick"]) == want, c["name"]


@pytest.mark.parametrize("geom", sorted(BATCHES))
def test_metrics_and_decisions_match_reference_golden(geom):
    b = BATCHES[geom]
    p = b.params
    m, v = co.eval_grid(b.scalars, b.buckets, b.le, b.phase, p, n_threads=2)
    n_checked = 0
    for case in CASES:
        if case["geometry"] != geom:
            continue
        i = case["pair"]
        for t in case["ticks"]:
            j = t["tick"]
            for ver, key in enumerate(("new", "old")):
                rec, valid = golden_record(t[key], p.n_q)
                assert valid == v[i, ver, j], (geom, i, j, key)
                assert records_equal(rec, m[i, ver, j]), (geom, i, j, key, rec, m[i, ver, j])
            d = co.gate(m[i, 0, j][None], v[i, 0, j:j + 1], m[i, 1, j][None], v[i, 1, j:j + 1], p)
            assert bool(d[0] & P.G_PROMOTE) == t["promote"]
            n_checked += 1
        for run in case["runs"]:
            st = np.full(b.phase.shape[0], run["start_tick"], np.int32)
            dec, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, start_tick=st)
            lazy = co.eval_lazy(b.scalars, b.buckets, b.le, b.phase, p, start_tick=st)
            want = (run["cur"], run["prev"], run["state"], run["n_promotions"], run["n_evals"])
            assert tuple(dec[i, :5]) == want, (geom, i, run["start_tick"])
            assert tuple(lazy[i]) == tuple(dec[i])
            # the virtual clock of the reference ends where the last tick says it should
            if run["state"] == P.STATE_COMPLETE:
                end = p.tick0_ms + (dec[i, 5] + p.interval_ticks) * p.tick_step_ms
            else:
                end = p.tick0_ms + dec[i, 5] * p.tick_step_ms
            assert end == run["end_ms"]
    assert n_checked > 0
