"""-m gpu: the device-resident rolling store (engine.RollingStore over cg_store_append / cg_store_shift).
In the operator's life an evaluation adds one scrape, not one hour (mlflow_operator.py:302-303, :340,
:352): feeding the store scrape by scrape must give bit-identical decisions to uploading the same
history in one piece, also after the rows have been compacted and when the view starts on an odd column."""
import numpy as np
import pytest
import torch

import canary_b200  # noqa: F401
from canary_b200 import engine, params as P, synth
from oracle import c_oracle as co

pytestmark = pytest.mark.gpu


def oracle_on_view(store, sc_np, bk_np, le, phase_np):
    off, n_s, pp, _ = store.view()
    q = pp.copy()
    q.row_pitch = n_s
    sc = np.ascontiguousarray(sc_np[..., off:off + n_s])
    bk = np.ascontiguousarray(bk_np[..., off:off + n_s])
    ph = (phase_np + off * pp.period_ms).astype(np.int32)
    return co.eval_fused(sc, bk, le, ph, q, n_threads=8)


@pytest.mark.parametrize("dtype", [P.CG_F64, P.CG_F32])
def test_appends_equal_bulk_upload_also_after_compaction(dtype):
    T, extra, n = 48, 21, 40
    long_p = P.make_params(n_samples=T + extra, n_buckets=8, dtype=dtype)
    b = synth.generate(long_p, n, seed=17, adversarial=0.1)          # the "true" history, T + extra scrapes
    p = P.make_params(n_samples=T, n_buckets=8, dtype=dtype)
    le = torch.from_numpy(b.le).cuda()
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    store = engine.RollingStore(p, n, le, slack=8)
    store.phase.copy_(dev(b.phase))
    held_sc = np.zeros((n, 2, p.n_scalar, 0), b.scalars.dtype)
    held_bk = np.zeros((n, 2, p.n_buckets, 0), b.buckets.dtype)
    base = 0                                                          # true index of the store's column 0
    for i in range(T + extra):
        if store.count == store.cap:
            base += store.count - (T - 1)
        store.append_scrape(dev(b.scalars[..., i]), dev(b.buckets[..., i]))
        assert store.count == i + 1 - base
        if i + 1 < T or (i + 1 - T) % 3 not in (0, 1):                # check even and odd view offsets, not every step
            continue
        got, gbits = store.evaluate(want_bits=True)
        torch.cuda.synchronize()
        # the store's columns are the true samples base .. i
        want, wbits = oracle_on_view(store, b.scalars[..., base:i + 1], b.buckets[..., base:i + 1], b.le, b.phase)
        assert np.array_equal(gbits.cpu().numpy().view(np.uint32), wbits), f"gate bits differ after {i + 1} scrapes"
        assert np.array_equal(got.cpu().numpy(), want), f"decisions differ after {i + 1} scrapes"
        if i + 1 == T:                                                # exactly the bulk upload of the first T scrapes
            bulk = engine.eval_fused(dev(b.scalars[..., :T]), dev(b.buckets[..., :T]), le, dev(b.phase), p)
            assert torch.equal(bulk, got)
    assert base > 0, "the test never compacted the rows"


def test_bulk_load_then_append_from_pinned_host():
    p = P.make_params(n_samples=60, n_buckets=6)
    n = 33
    long_p = P.make_params(n_samples=64, n_buckets=6)
    b = synth.generate(long_p, n, seed=23)
    le = torch.from_numpy(b.le).cuda()
    store = engine.RollingStore(p, n, le, slack=16)
    store.load(torch.from_numpy(np.ascontiguousarray(b.scalars[..., :60])).cuda(),
               torch.from_numpy(np.ascontiguousarray(b.buckets[..., :60])).cuda(), torch.from_numpy(b.phase).cuda())
    for i in range(60, 64):
        store.append_scrape(torch.from_numpy(np.ascontiguousarray(b.scalars[..., i])).pin_memory(),
                            torch.from_numpy(np.ascontiguousarray(b.buckets[..., i])).pin_memory())
        assert store.h2d_bytes == n * 2 * (p.n_scalar + p.n_buckets) * 8
    got = store.evaluate()
    torch.cuda.synchronize()
    want, _ = oracle_on_view(store, b.scalars, b.buckets, b.le, b.phase)
    assert np.array_equal(got.cpu().numpy(), want)
