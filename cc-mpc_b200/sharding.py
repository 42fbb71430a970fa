"""Multi-GPU sharding of the path: one process per GPU, data parallel only.

The reference has no distributed code.  The path shards naturally:
  * Monte Carlo samples are i.i.d.: every rank rolls / reads its own slice of the samples of
    every (obstacle, mode) cell.  The Philox counter is the GLOBAL sample index
    (``sample_offset``), so the union of the shards is the same sample set for any world size.
    The only exchange is ONE all-reduce(sum) of the packed per-cell sums vector
    (collision counts, cost sums, shifted power sums; ``ccmpc_sample_sums`` layout).  Counts
    are integers < 2^53 carried in fp64, so the reduced counts are exact and order independent.
  * MVOE scenario sweeps are independent per scenario: contiguous scenario ranges per rank,
    results stay sharded, no collective.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of `total` items for `rank`: sizes differ by at most one, earlier ranks larger."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad rank / world")
    base, rem = divmod(int(total), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def sample_shard(n_samples: int, rank: Optional[int] = None, world: Optional[int] = None) -> Tuple[int, int]:
    """(sample_offset, n_local) of this rank for a rollout of `n_samples` per cell."""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_range(n_samples, rank, world)
    return lo, hi - lo


def scenario_shard(S: int, rank: Optional[int] = None, world: Optional[int] = None) -> Tuple[int, int]:
    """[lo, hi) scenario range of this rank for a batched MVOE sweep (no exchange needed)."""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    return shard_range(S, rank, world)


def allreduce_sums(sums: torch.Tensor, group=None, async_op: bool = False):
    """The one collective of the path: in-place sum of the packed sums over the sample shards.
    Every rank must have used the same pivot (the GMM / moment means, known everywhere)."""
    if sums.dtype != torch.float64:
        raise TypeError("packed sums are float64")
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    if sums.is_cuda and dist.get_backend(group) == "gloo":
        # host-side test rigs (several ranks on one GPU): gloo reduces a host copy
        host = sums.cpu()
        dist.all_reduce(host, op=dist.ReduceOp.SUM, group=group)
        sums.copy_(host)
        return None
    return dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def counts_from_sums(sums: torch.Tensor, T: int) -> torch.Tensor:
    """int64 [C, T+1] (per-timestep and any-time collision counts) from reduced packed sums."""
    D = 2 * T
    o_coll = 1 + D + D * (D + 1) // 2
    return sums[:, o_coll:o_coll + T + 1].round().to(torch.int64)
