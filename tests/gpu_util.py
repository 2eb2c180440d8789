"""Shared helpers for the -m gpu tests: upload a host batch, run the CUDA library through
its C ABI (canary_b200.engine -> ctypes -> libcanarygate.so), fetch results."""
import numpy as np
import torch

import canary_b200  # noqa: F401
from canary_b200 import engine, operator_api


def upload(batch):
    return operator_api.SeriesStore.from_host(batch, device="cuda")


def gpu_grid(batch):
    st = upload(batch)
    m, v = engine.eval_grid(st.scalars, st.buckets, st.le, st.phase, batch.params)
    torch.cuda.synchronize()
    return m.cpu().numpy(), v.cpu().numpy()


def gpu_fused_metrics(batch, path=0):
    """decisions, bits, metrics, valid from ONE fused launch (path 2 = streaming kernel)."""
    st = upload(batch)
    dec, bits, m, v = engine.eval_fused(st.scalars, st.buckets, st.le, st.phase, batch.params,
                                        want_bits=True, want_metrics=True, path=path)
    torch.cuda.synchronize()
    return dec.cpu().numpy(), bits.cpu().numpy().view(np.uint32), m.cpu().numpy(), v.cpu().numpy()


def gpu_fused(batch, has_previous=None, start_tick=None, want_bits=True, path=0):
    st = upload(batch)
    hp = None if has_previous is None else torch.from_numpy(has_previous).cuda()
    stt = None if start_tick is None else torch.from_numpy(start_tick).cuda()
    out = engine.eval_fused(st.scalars, st.buckets, st.le, st.phase, batch.params,
                            has_previous=hp, start_tick=stt, want_bits=want_bits, path=path)
    torch.cuda.synchronize()
    if want_bits:
        return out[0].cpu().numpy(), out[1].cpu().numpy().view(np.uint32)
    return out.cpu().numpy()


def assert_metrics_match(got, want, got_v, want_v, rtol=1e-6, exact=True, what=""):
    """Required bar (BASELINE.json north_star): 1e-6 relative in fp64.  The kernels are
    written to reproduce the oracle's IEEE operation order, so `exact` additionally
    demands bit equality (NaN == NaN)."""
    assert np.array_equal(got_v, want_v), f"{what}: validity bits differ"
    nan_g, nan_w = np.isnan(got), np.isnan(want)
    assert np.array_equal(nan_g, nan_w), f"{what}: NaN pattern differs"
    g, w = np.where(nan_g, 0.0, got), np.where(nan_w, 0.0, want)
    inf = np.isinf(w)
    assert np.array_equal(g[inf], w[inf]), f"{what}: infinities differ"
    g, w = np.where(inf, 0.0, g), np.where(inf, 0.0, w)
    err = np.abs(g - w)
    tol = rtol * np.abs(w)
    bad = err > tol
    assert not bad.any(), f"{what}: {bad.sum()} values off by more than {rtol} rel, max abs err {err.max()}"
    if exact:
        neq = g != w
        assert not neq.any(), f"{what}: {neq.sum()} values not bit-identical (max rel {np.max(err / np.maximum(np.abs(w), 1e-300))})"
