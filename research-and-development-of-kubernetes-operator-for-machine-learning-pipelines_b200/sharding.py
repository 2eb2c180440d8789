"""Multi-GPU form of the canary step: model pairs are independent (each query of the
reference selects one deployment's series, mlflow_operator.py:302-303), so rank r owns the
contiguous pair range [r*ceil(N/G), ...) and no input ever crosses NVLink.  The only exchange
is one all-gather of the 24-byte decision records, done IN PLACE: every rank's kernel writes
its records straight into its slice of the gather buffer (no staging copy), then
`all_gather_into_tensor` (NCCL over NVLink/NVSwitch on GPUs; gloo in the CPU tests) fills in
the other slices.  One process per GPU, launched with torchrun.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist

RECORD_INTS = 6   # cg_decision: traffic_current, traffic_prev, state, n_promotions, n_evals, last_tick


def shard_size(n_pairs: int, world: int) -> int:
    """Pairs per rank (the last rank may own fewer; its tail is padding)."""
    return (n_pairs + world - 1) // world


def shard_range(n_pairs: int, rank: int, world: int) -> Tuple[int, int]:
    per = shard_size(n_pairs, world)
    lo = min(rank * per, n_pairs)
    return lo, min(lo + per, n_pairs)


def alloc_gather_buffer(n_pairs: int, world: int, device) -> torch.Tensor:
    """[world * shard_size, 6] int32; rank r's records live at rows [r*per, (r+1)*per)."""
    return torch.zeros((world * shard_size(n_pairs, world), RECORD_INTS), dtype=torch.int32, device=device)


def my_slice(buf: torch.Tensor, n_pairs: int, rank: int, world: int) -> torch.Tensor:
    """The rows of the gather buffer this rank's kernel must write (its real pairs only)."""
    per = shard_size(n_pairs, world)
    lo, hi = shard_range(n_pairs, rank, world)
    return buf[rank * per: rank * per + (hi - lo)]


def gather_decisions(buf: torch.Tensor, n_pairs: int, group=None) -> torch.Tensor:
    """All-gather in place and return the [n_pairs, 6] view in global pair order."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    per = shard_size(n_pairs, world)
    if world > 1:
        dist.all_gather_into_tensor(buf, buf[rank * per:(rank + 1) * per], group=group)
    return buf[:n_pairs]


def evaluate_sharded(evaluate_local: Callable[[torch.Tensor], None], n_pairs: int, device,
                     group=None, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`evaluate_local(out)` must fill `out` ([local_pairs, 6] int32, a slice of the gather
    buffer) with this rank's decision records, e.g. `lambda out: engine.eval_fused(..., out=out)`.
    Returns all n_pairs records on every rank."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if buf is None:
        buf = alloc_gather_buffer(n_pairs, world, device)
    mine = my_slice(buf, n_pairs, rank, world)
    if mine.shape[0]:
        evaluate_local(mine)
    return gather_decisions(buf, n_pairs, group)
