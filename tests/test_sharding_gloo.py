"""N > 1 host logic on CPU: two gloo ranks shard the pairs, fill their slice of the in-place
gather buffer and all-gather the decision records.  The per-rank evaluator here is the oracle
(no GPU in this test); on GPUs it is engine.eval_fused writing into the same slice."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import canary_b200  # noqa: F401
from canary_b200 import params as P, sharding, synth
from oracle import c_oracle as co


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_pairs, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p = P.make_params(n_samples=80, n_buckets=8)
        b = synth.generate(p, n_pairs, seed=17, adversarial=0.1)     # same seed on every rank
        lo, hi = sharding.shard_range(n_pairs, rank, world)

        def evaluate_local(out):
            dec, _ = co.eval_fused(b.scalars[lo:hi], b.buckets[lo:hi], b.le, b.phase[lo:hi], p, want_bits=False)
            out.copy_(torch.from_numpy(dec))
        got = sharding.evaluate_sharded(evaluate_local, n_pairs, "cpu")
        q.put((rank, got.numpy().copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_pairs", [10, 7, 1])
def test_two_rank_gather_equals_single_process(n_pairs):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_pairs, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    results = dict(q.get(timeout=120) for _ in range(world))
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    p = P.make_params(n_samples=80, n_buckets=8)
    b = synth.generate(p, n_pairs, seed=17, adversarial=0.1)
    want, _ = co.eval_fused(b.scalars, b.buckets, b.le, b.phase, p, want_bits=False)
    for r in range(world):
        assert np.array_equal(results[r], want), f"rank {r}"


def test_shard_ranges_partition_the_pairs():
    for n in (0, 1, 5, 8, 100_000, 1_000_003):
        for world in (1, 2, 4, 8):
            spans = [sharding.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            per = sharding.shard_size(n, world)
            assert all(hi - lo <= per for lo, hi in spans)
            buf = sharding.alloc_gather_buffer(n, world, "cpu")
            assert buf.shape == (per * world, 6)
            assert sum(sharding.my_slice(buf, n, r, world).shape[0] for r in range(world)) == n
