"""Regenerate the golden fixtures in this directory.

Run in the BUILD container (needs /root/reference):  python tests/golden/make_golden.py

  kat_promql.json        SURVEY.md Appendix B known-answer vectors (hand-verifiable;
                         written here literally, then checked against oracle/pyoracle.py
                         before being saved)
  ref_gate_probes.json   should_promote_model (mlflow_operator.py:419-460) quirk probes,
                         answers produced by the UNMODIFIED reference function
  ref_schedule_probes.json  step-loop traces (:296-361) for scripted gate outcomes,
                         produced by the unmodified reference coroutine under a virtual clock
  ref_canary_cases.npz/.json  seeded synthetic pairs (inputs in the .npz) with, from the
                         unmodified reference run on oracle/pyoracle.py's fake Prometheus:
                         per-tick get_model_metrics dicts (:363-417) on a tick subsample,
                         the final weights / state / evaluation count of the canary loop
  ref_irregular_cases.npz/.json  the same for IRREGULAR series (jittered and
                         missed scrapes, gaps, series that start late, end early or are absent)

`python tests/golden/make_golden.py irregular` regenerates only the last pair of files.
"""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import canary_b200  # noqa: E402,F401
from canary_b200 import params as P, synth  # noqa: E402
from oracle import pyoracle as po, ref_loader as rl, soa_bridge as sb  # noqa: E402


def fenc(x):
    if x is None:
        return None
    if math.isnan(x):
        return "nan"
    if math.isinf(x):
        return "inf" if x > 0 else "-inf"
    return float(x).hex()


def kat():
    S = lambda ts, vs: [[int(t * 1000), float(v)] for t, v in zip(ts, vs)]
    inc = [
        dict(id="K1", samples=S([7, 22, 37, 52], [100, 110, 120, 130]), t_ms=60000, range_ms=60000,
             increase=40.0, rate=0.6666666666666666),
        dict(id="K2", samples=S([7, 22, 37, 52], [0, 10, 20, 30]), t_ms=60000, range_ms=60000,
             increase=35.333333333333336),
        dict(id="K3", samples=S([7, 22, 37, 52], [100, 110, 5, 15]), t_ms=60000, range_ms=60000,
             increase=33.33333333333333),
        dict(id="K4", samples=S([20, 35, 50], [100, 110, 120]), t_ms=60000, range_ms=60000,
             increase=31.666666666666664),
        dict(id="K5", samples=S([20], [100]), t_ms=60000, range_ms=60000, increase=None),
        dict(id="U1", samples=[[i * 300000, 10.0 * i] for i in range(11)], t_ms=3000000,
             range_ms=3000000, increase=100.0, both_windows=True),
        dict(id="U2", samples=[[i * 300000, float(v)] for i, v in
                               enumerate([0, 10, 20, 30, 40, 50, 0, 10, 20, 30, 40, 50])],
             t_ms=3000000, range_ms=3000000, increase=90.0),
    ]
    INF = "inf"
    qs = [
        dict(id="Qp", buckets=[[0.1, 50], [0.2, 70], [1, 110], [INF, 120]],
             q=[0, 0.2, 0.5, 0.8, 0.95, 1, -0.1, 1.01],
             expect=[0.0, 0.048, 0.15, 0.72, 1.0, 1.0, "-inf", "inf"]),
        dict(id="Qn", buckets=[[-0.2, 10], [-0.1, 20], [0.3, 20], [INF, 30]],
             q=[0, 0.2, 0.5, 0.8, 1], expect=[-0.2, -0.2, -0.15, 0.3, 0.3]),
        dict(id="Q0", buckets=[[0.1, 0], [INF, 0]], q=[0.95], expect=["nan"]),
    ]
    # sanity: the literals above must be what the restatement gives
    for k in inc:
        w = po.select_range([tuple(s) for s in k["samples"]], k["t_ms"], k["range_ms"])
        got = po.extrapolated_increase(w, k["t_ms"], k["range_ms"])
        assert got == k["increase"], (k["id"], got)
    for k in qs:
        bs = [(float(a), float(b)) for a, b in k["buckets"]]
        for q, e in zip(k["q"], k["expect"]):
            got = po.bucket_quantile(q, bs)
            e = float(e)
            assert (math.isnan(e) and math.isnan(got)) or abs(got - e) <= 1e-12 * max(1, abs(e)) or got == e, (k["id"], q, got)
    return dict(source="SURVEY.md Appendix B", increase=inc, quantile=qs)


def gate_probes():
    thr = dict(P.REFERENCE_THRESHOLDS)
    base = dict(latency_95th=0.2, error_rate=0.01, latency_avg=0.1)
    cases = []

    def add(name, new, old):
        n, o = dict(base, **new), dict(base, **old)
        cases.append(dict(name=name, new={k: fenc(v) for k, v in n.items()},
                          old={k: fenc(v) for k, v in o.items()},
                          promote=bool(rl.reference_should_promote(n, o, thr))))
    add("equal", {}, {})
    add("old_er0_new_er0", dict(error_rate=0.0), dict(error_rate=0.0))
    add("old_er0_new_er_tiny", dict(error_rate=1e-9), dict(error_rate=0.0))
    add("nan_new_p95", dict(latency_95th=math.nan), {})
    add("nan_old_p95", {}, dict(latency_95th=math.nan))
    add("none_new_er", dict(error_rate=None), {})
    add("none_old_avg", {}, dict(latency_avg=None))
    add("none_new_p95", dict(latency_95th=None), {})
    add("tie_p95_exact", dict(latency_95th=0.2 * (1 + 0.05)), {})
    add("just_above_p95", dict(latency_95th=math.nextafter(0.2 * (1 + 0.05), 1.0)), {})
    add("tie_er_exact", dict(error_rate=0.01 * (1 + 0.02)), {})
    add("just_above_er", dict(error_rate=math.nextafter(0.01 * (1 + 0.02), 1.0)), {})
    add("tie_avg_exact", dict(latency_avg=0.1 * (1 + 0.05)), {})
    add("just_above_avg", dict(latency_avg=math.nextafter(0.1 * (1 + 0.05), 1.0)), {})
    add("inf_both", dict(latency_95th=math.inf), dict(latency_95th=math.inf))
    add("only_avg_fails", dict(latency_avg=0.2), {})
    add("all_better", dict(latency_95th=0.1, error_rate=0.0, latency_avg=0.05), {})
    return dict(source="mlflow_operator.py:419-460 run unmodified", thresholds=thr, cases=cases)


def schedule_probes():
    """Scripted gate outcomes as a function of virtual time -> reference trace."""
    scripts = {
        "all_pass": lambda now: True,
        "holds_at_0_10": lambda now: now not in (0, 10000),
        "all_fail": lambda now: False,
        "ten_fail_at_step2": lambda now: now < 60000,
        "nine_fail_then_pass": lambda now: (now // 10000) % 10 == 9,
        "alternating": lambda now: (now // 10000) % 2 == 1,
        "pass_every_7th": lambda now: (now // 10000) % 7 == 3,
        "fail_late": lambda now: now < 400000,
    }
    out = []
    for name, fn in scripts.items():
        clock = rl.VirtualClock(0)
        proms = []

        def canned(q, now, fn=fn):
            ok = fn(now)
            is_new = 'predictor_name="v2"' in q
            if q.startswith("histogram_quantile"):
                return [{"metric": {}, "value": [now / 1000, "0.1" if (ok or not is_new) else "0.5"]}]
            if 'code!="200"' in q:
                return [{"metric": {}, "value": [0, "1"]}]
            if "_sum" in q:
                return [{"metric": {}, "value": [0, "5"]}]
            return [{"metric": {}, "value": [0, "100"]}]

        def fac():
            pr = rl.CannedPrometheus(canned, clock)
            proms.append(pr)
            return pr
        res = rl.run_reference_canary(fac, clock)
        n_ticks = 200
        bits = [bool(fn(j * 10000)) for j in range(n_ticks)]
        out.append(dict(name=name, gate_bits=bits, tick_step_ms=10000,
                        eval_ticks=[t // 10000 for t in proms[0].eval_times_ms],
                        weights=res["weights"], end_ms=res["end_ms"],
                        last_event=res["events"][-1]["reason"] if res["events"] else None))
    return dict(source="mlflow_operator.py:156-361 run unmodified under a virtual clock", cases=out)


def canary_cases():
    cases = []
    arrays = {}
    geoms = [
        ("small", dict(n_samples=120, n_scalar=4, n_buckets=12), 8, 0xC0FFEE + 100, 0.4),
        ("small_k5", dict(n_samples=120, n_scalar=5, n_buckets=12), 3, 0xC0FFEE + 101, 0.0),
        ("full", dict(n_samples=240, n_scalar=4, n_buckets=64), 2, 0xC0FFEE + 102, 0.0),
    ]
    for gname, kw, n_pairs, seed, adv in geoms:
        p = P.make_params(**kw)
        batch = synth.generate(p, n_pairs, seed=seed, adversarial=adv)
        arrays[f"{gname}_scalars"] = batch.scalars
        arrays[f"{gname}_buckets"] = batch.buckets
        arrays[f"{gname}_le"] = batch.le
        arrays[f"{gname}_phase"] = batch.phase
        for pair in range(n_pairs):
            db = sb.pair_to_db(batch, pair)
            ticks = sorted(set(list(range(0, p.n_ticks, 11)) + [0, 1, 2, p.n_ticks - 1]))
            per_tick = []
            for j in ticks:
                clock = rl.VirtualClock(p.tick0_ms + j * p.tick_step_ms)
                fp = rl.FakePrometheus(db, clock)
                rec = {"tick": j}
                for ver, pred in enumerate(("v2", "v1")):
                    md = rl.reference_get_model_metrics(fp, "m", pred, "ns")
                    rec["new" if ver == 0 else "old"] = {k: fenc(v) for k, v in md.items()}
                rec["promote"] = bool(rl.reference_should_promote(
                    rl.reference_get_model_metrics(fp, "m", "v2", "ns"),
                    rl.reference_get_model_metrics(fp, "m", "v1", "ns"),
                    dict(P.REFERENCE_THRESHOLDS)))
                per_tick.append(rec)
            runs = []
            for start in (0, 5):
                r = sb.run_reference_on_pair(batch, pair, start_tick=start)
                runs.append(dict(start_tick=start, cur=r["cur"], prev=r["prev"], state=r["state"],
                                 n_promotions=r["n_promotions"], n_evals=r["n_evals"],
                                 weights=r["weights"], end_ms=r["end_ms"], events=r["events"]))
            cases.append(dict(geometry=gname, make_params=kw, pair=pair, ticks=per_tick, runs=runs))
    np.savez_compressed(os.path.join(HERE, "ref_canary_cases.npz"), **arrays)
    return dict(source="mlflow_operator.py get_model_metrics/should_promote_model/step loop run "
                       "unmodified over oracle/pyoracle.py FakePrometheus", cases=cases)


def irregular_cases():
    cases = []
    arrays = {}
    geoms = [
        ("irr", dict(n_samples=120, n_scalar=4, n_buckets=8), 6, 0xC0FFEE + 200, 0.3,
         dict(seed=7, miss_prob=0.08, gap_prob=0.03, truncate_prob=0.1, dead_series_prob=0.03)),
        ("irr_open", dict(n_samples=80, n_scalar=4, n_buckets=6, left_open=True), 3, 0xC0FFEE + 201, 0.0,
         dict(seed=8, jitter_ms=0, miss_prob=0.1, gap_prob=0.05)),     # no jitter: samples on window edges
    ]
    for gname, kw, n_pairs, seed, adv, irr in geoms:
        p = P.make_params(**kw)
        lo = bool(kw.get("left_open"))
        grid = synth.generate(p, n_pairs, seed=seed, adversarial=adv, coincident_phase=lo)
        batch = synth.make_irregular(grid, **irr)
        for field in ("scalars", "buckets", "ts_scalars", "ts_buckets", "cnt_scalars", "cnt_buckets", "le"):
            arrays[f"{gname}_{field}"] = getattr(batch, field)
        for pair in range(n_pairs):
            db = sb.pair_to_db(batch, pair, left_open=lo)
            ticks = sorted(set(list(range(0, p.n_ticks, 7)) + [0, 1, 2, p.n_ticks - 1]))
            per_tick = []
            for j in ticks:
                fp = rl.FakePrometheus(db, rl.VirtualClock(p.tick0_ms + j * p.tick_step_ms))
                rec = {"tick": j}
                mds = []
                for ver, pred in enumerate(("v2", "v1")):
                    md = rl.reference_get_model_metrics(fp, "m", pred, "ns")
                    mds.append(md)
                    rec["new" if ver == 0 else "old"] = {k: fenc(v) for k, v in md.items()}
                rec["promote"] = bool(rl.reference_should_promote(mds[0], mds[1], dict(P.REFERENCE_THRESHOLDS)))
                per_tick.append(rec)
            runs = []
            for start in (0, 4):
                r = sb.run_reference_on_pair(batch, pair, start_tick=start, left_open=lo)
                runs.append(dict(start_tick=start, cur=r["cur"], prev=r["prev"], state=r["state"],
                                 n_promotions=r["n_promotions"], n_evals=r["n_evals"],
                                 weights=r["weights"], end_ms=r["end_ms"], events=r["events"]))
            cases.append(dict(geometry=gname, make_params=kw, pair=pair, ticks=per_tick, runs=runs))
    np.savez_compressed(os.path.join(HERE, "ref_irregular_cases.npz"), **arrays)
    return dict(source="mlflow_operator.py get_model_metrics/should_promote_model/step loop run unmodified over "
                       "oracle/pyoracle.py FakePrometheus; series with per-series time stamps", cases=cases)


def main():
    assert rl.reference_available(), "needs /root/reference"
    todo = (("kat_promql.json", kat), ("ref_gate_probes.json", gate_probes),
            ("ref_schedule_probes.json", schedule_probes), ("ref_canary_cases.json", canary_cases),
            ("ref_irregular_cases.json", irregular_cases))
    if sys.argv[1:] == ["irregular"]:
        todo = todo[-1:]
    for name, fn in todo:
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(fn(), f, indent=1)
        print("wrote", name)


if __name__ == "__main__":
    main()
