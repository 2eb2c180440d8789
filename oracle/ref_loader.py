"""Import the UNMODIFIED reference operator under stub packages and drive its
canary loop with a fake Prometheus and a virtual clock.

TEST INFRASTRUCTURE ONLY (see oracle/pyoracle.py header).  Needs the reference's one
Python file: /root/reference in the build container, or the untouched copy that
`make -C oracle _ref` puts into oracle/_ref/ (git-ignored) so that the GPU box has it too.
It is what pins decisions and traffic weights to the reference's own code:
`get_model_metrics` (mlflow_operator.py:363-417), `should_promote_model`
(:419-460) and the inline step loop of `create_or_update_seldon_deployment`
(:288-361) run exactly as written; only their collaborators are faked.

Stubs: kopf (decorators + event), kubernetes (config, client,
ApiException), mlflow.tracking.MlflowClient, prometheus_api_client
.PrometheusConnect.  `asyncio.sleep` inside the module is replaced by a
virtual clock so a 540 s canary replays in microseconds.
"""
from __future__ import annotations

import asyncio
import importlib.util
import logging
import os
import sys
import types
from typing import Callable, List, Optional

from . import pyoracle

def _find_reference() -> str:
    """The unmodified reference file: $CG_REFERENCE_FILE, else the read-only mount of the build
    container, else the copy `make -C oracle _ref` (or __graft_entry__.build()) left in oracle/_ref/
    (git-ignored; it travels to the GPU box with the snapshot, like the built .so files)."""
    env = os.environ.get("CG_REFERENCE_FILE")
    if env:
        return env
    mounted = "/root/reference/mlflow_operator.py"
    if os.path.isfile(mounted):
        return mounted
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "mlflow_operator.py")


REFERENCE_FILE = _find_reference()


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_FILE)


class ApiException(Exception):
    def __init__(self, status=0, reason=""):
        super().__init__(f"({status}) {reason}")
        self.status = status


class VirtualClock:
    def __init__(self, start_ms: int = 0):
        self.now_ms = start_ms
        self.sleeps: List[float] = []

    async def sleep(self, seconds):
        self.sleeps.append(seconds)
        self.now_ms += int(round(seconds * 1000))


class Recorder:
    """Fake CustomObjectsApi: every apply is a 404 on GET followed by a
    recorded CREATE, so the weight trace can be read back."""

    def __init__(self):
        self.applied: List[dict] = []
        self.events: List[dict] = []

    def get_namespaced_custom_object(self, **kw):
        raise ApiException(404, "not found")

    def create_namespaced_custom_object(self, group, version, namespace, plural, body):
        import copy
        self.applied.append(copy.deepcopy(body))

    def replace_namespaced_custom_object(self, **kw):  # never reached (GET 404s)
        import copy
        self.applied.append(copy.deepcopy(kw["body"]))

    def delete_namespaced_custom_object(self, **kw):
        pass

    def patch_namespaced_custom_object_status(self, **kw):
        pass

    def weight_trace(self) -> List[List[int]]:
        return [[p["traffic"] for p in sd["spec"]["predictors"]] for sd in self.applied]


class _State:
    """What the stub modules read at call time."""
    recorder: Optional[Recorder] = None
    prom_factory: Optional[Callable[[], object]] = None


def _install_stubs():
    kopf = types.ModuleType("kopf")

    def _deco(*a, **k):
        return lambda fn: fn
    kopf.on = types.SimpleNamespace(create=_deco, update=_deco)

    def _event(body, type=None, reason=None, message=None, **kw):
        if _State.recorder is not None:
            _State.recorder.events.append(dict(type=type, reason=reason, message=message))
    kopf.event = _event

    k8s = types.ModuleType("kubernetes")
    k8s_config = types.ModuleType("kubernetes.config")
    k8s_config.load_incluster_config = lambda: None
    k8s_client = types.ModuleType("kubernetes.client")
    k8s_client.CustomObjectsApi = lambda: _State.recorder
    k8s_client.V1DeleteOptions = lambda: None
    k8s_rest = types.ModuleType("kubernetes.client.rest")
    k8s_rest.ApiException = ApiException
    k8s_client.rest = k8s_rest
    k8s.config = k8s_config
    k8s.client = k8s_client

    mlflow = types.ModuleType("mlflow")
    mlflow_tracking = types.ModuleType("mlflow.tracking")
    mlflow_tracking.MlflowClient = lambda *a, **k: None
    mlflow.tracking = mlflow_tracking

    pac = types.ModuleType("prometheus_api_client")

    def _connect(url=None, disable_ssl=False, **kw):
        return _State.prom_factory()
    pac.PrometheusConnect = _connect

    return {
        "kopf": kopf, "kubernetes": k8s, "kubernetes.config": k8s_config,
        "kubernetes.client": k8s_client, "kubernetes.client.rest": k8s_rest,
        "mlflow": mlflow, "mlflow.tracking": mlflow_tracking,
        "prometheus_api_client": pac,
    }


_ref_module = None


def load_reference():
    """Return the reference module object, importing it once."""
    global _ref_module
    if _ref_module is not None:
        return _ref_module
    if not reference_available():
        raise FileNotFoundError(REFERENCE_FILE)
    stubs = _install_stubs()
    saved = {k: sys.modules.get(k) for k in stubs}
    sys.modules.update(stubs)
    root_level = logging.getLogger().level
    try:
        spec = importlib.util.spec_from_file_location("_cg_reference_mlflow_operator", REFERENCE_FILE)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)   # runs config.load_incluster_config() (:13) against the stub
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        logging.getLogger().setLevel(root_level)
    _ref_module = mod
    return mod


class FakePrometheus:
    """`custom_query(str)` evaluated by oracle/pyoracle.py at the virtual now."""

    def __init__(self, db: pyoracle.SeriesDB, clock: VirtualClock):
        self.db, self.clock = db, clock
        self.n_queries = 0

    def custom_query(self, query: str):
        self.n_queries += 1
        return pyoracle.eval_query(self.db, query, self.clock.now_ms)


class CannedPrometheus:
    """Answers from a function (query, now_ms) -> result list; for gate probes."""

    def __init__(self, fn, clock: VirtualClock):
        self.fn, self.clock = fn, clock
        self.eval_times_ms: List[int] = []

    def custom_query(self, query: str):
        if not self.eval_times_ms or self.eval_times_ms[-1] != self.clock.now_ms:
            self.eval_times_ms.append(self.clock.now_ms)
        return self.fn(query, self.clock.now_ms)


_QUIET = logging.getLogger("cg.reference.quiet")
_QUIET.addHandler(logging.NullHandler())
_QUIET.propagate = False
_QUIET.setLevel(logging.CRITICAL)


def run_reference_canary(prom_factory, clock: VirtualClock, name="m", namespace="ns",
                         current_version="2", previous_version="1", logger=None) -> dict:
    """Run create_or_update_seldon_deployment (mlflow_operator.py:156-361)."""
    mod = load_reference()
    rec = Recorder()
    _State.recorder = rec
    _State.prom_factory = prom_factory
    fake_asyncio = types.SimpleNamespace(sleep=clock.sleep)
    real_asyncio = mod.asyncio
    mod.asyncio = fake_asyncio
    try:
        coro = mod.create_or_update_seldon_deployment(
            name=name, namespace=namespace, minio_secret_name="s",
            body={"metadata": {"uid": "u"}}, current_model_version=current_version,
            previous_model_version=previous_version, new_model_uri="s3://new",
            old_model_uri="s3://old" if previous_version is not None else None,
            prometheus_url="http://fake", logger=logger or _QUIET)
        asyncio.run(coro)
    finally:
        mod.asyncio = real_asyncio
        _State.recorder = None
        _State.prom_factory = None
    return dict(weights=rec.weight_trace(), events=rec.events,
                end_ms=clock.now_ms, sleeps=list(clock.sleeps))


def reference_get_model_metrics(prom, deployment, predictor, namespace, elapsed_time=60):
    return load_reference().get_model_metrics(prom, deployment, predictor, namespace, elapsed_time)


def reference_should_promote(new, old, thresholds, logger=None):
    return load_reference().should_promote_model(new, old, thresholds, logger or _QUIET)
