"""Env-batch sharding across the GPUs of one box.

The reference has no distributed code at all: every term of the step (examples/test_viz.py:25-44)
is per-environment, so the batch axis shards with NO data-path collective.  Rank r of W owns the
contiguous env range `shard_range(B, r, W)`, its own PRNG keys (key of GLOBAL env i is the same
whatever W is, so results do not depend on the sharding), its own device state and its own CUDA
graph.  The one optional exchange is the episode-return statistic: a SUM all-reduce of three
float64 numbers (return sum, finished episodes, env-steps) every reporting interval — NCCL on
GPUs, gloo in the CPU tests.

One process per GPU (torchrun): RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR come from the env.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np


def dist_env():
    """(rank, world_size, local_rank) from the torchrun environment (1 process = 1 GPU)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) of the envs rank `rank` owns: contiguous, sizes differ by at most one, the
    first `total % world` ranks get the extra env.  Empty when total < world for high ranks."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    if total < 0:
        raise ValueError("negative batch")
    q, r = divmod(total, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def shard_keys(total: int, rank: int, world: int, seed_hi: int = 0) -> np.ndarray:
    """PRNG keys [n, 2] uint32 of this rank's envs: global env i gets key (seed_hi, i) =
    jax.random.PRNGKey(i) words for seed_hi = 0 — independent of the world size."""
    lo, hi = shard_range(total, rank, world)
    k = np.empty((hi - lo, 2), np.uint32)
    k[:, 0] = np.uint32(seed_hi)
    k[:, 1] = np.arange(lo, hi, dtype=np.uint64).astype(np.uint32)
    return k


@dataclass
class ReturnStats:
    """Whole-job episode statistics after the reduction."""

    return_sum: float
    episodes: float
    env_steps: float

    @property
    def mean_return(self) -> float:
        return self.return_sum / self.episodes if self.episodes > 0 else float("nan")


class ReturnReducer:
    """Accumulates (sum of finished-episode returns, finished episodes, env-steps) locally and
    all-reduces them on demand.  The only collective of the engine; off the step's critical
    path (a few bytes; latency-bound)."""

    def __init__(self, device=None):
        import torch

        self.device = device if device is not None else torch.device("cpu")
        self.acc = torch.zeros(3, dtype=torch.float64, device=self.device)

    def update(self, episode_return, done) -> None:
        """episode_return[B]: return of the episode that JUST finished where done != 0."""
        import torch

        fin = done != 0
        ret = torch.where(fin, episode_return.to(torch.float64), torch.zeros((), dtype=torch.float64,
                                                                             device=episode_return.device))
        self.acc[0] += ret.sum().to(self.device)
        self.acc[1] += fin.sum().to(self.device, torch.float64)
        self.acc[2] += float(done.numel())

    def reduce(self) -> ReturnStats:
        import torch.distributed as dist

        t = self.acc.clone()
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        v = t.cpu().tolist()
        return ReturnStats(*v)


def max_over_ranks(value: float, device=None) -> float:
    """max over ranks of a host float (timing rule: multi-GPU numbers are the max over ranks)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
