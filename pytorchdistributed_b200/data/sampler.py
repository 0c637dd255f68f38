"""Rank-sharded index sampler.

Behavioural contract (what the reference relies on at
``R/02 DDP基本概念/ddp_gpus.py:71-76`` and ``ddp_gpus_torchrun.py:72-77``, i.e.
``torch.utils.data.DistributedSampler``): every rank draws the same seeded
permutation of ``range(len(dataset))`` (seed + epoch), the list is padded by
wrap-around (or truncated with ``drop_last``) to a multiple of the world size,
and rank ``r`` keeps positions ``r, r+W, r+2W, ...``.  ``set_epoch`` reseeds
(``ddp_gpus.py:47``).  ``tests/test_cpu_data.py`` checks index-for-index parity
against torch's sampler over a grid.
"""
from __future__ import annotations

import math
from typing import Iterator, Optional, Sized

import torch
import torch.distributed as dist
from torch.utils.data import Sampler


def _infer(world_size: Optional[int], rank: Optional[int]) -> tuple[int, int]:
    if world_size is None or rank is None:
        if dist.is_available() and dist.is_initialized():
            ws, rk = dist.get_world_size(), dist.get_rank()
        else:
            raise RuntimeError("ShardedSampler needs num_replicas/rank or an initialised process group")
        world_size = ws if world_size is None else world_size
        rank = rk if rank is None else rank
    if not 0 <= rank < world_size:
        raise ValueError(f"rank {rank} outside [0, {world_size})")
    return world_size, rank


class ShardedSampler(Sampler[int]):
    """Drop-in for ``DistributedSampler`` (same ctor keywords, same indices)."""

    def __init__(self, dataset: Sized, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        self.num_replicas, self.rank = _infer(num_replicas, rank)
        self.dataset = dataset
        self.shuffle = shuffle
        self.seed = seed
        self.drop_last = drop_last
        self.epoch = 0
        n = len(dataset)
        if drop_last and n % self.num_replicas:
            self.num_samples = n // self.num_replicas
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas

    def global_order(self) -> list[int]:
        """The padded/truncated order shared by all ranks for the current epoch."""
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=g).tolist()
        else:
            order = list(range(n))
        if self.drop_last:
            return order[: self.total_size]
        short = self.total_size - len(order)
        if short > 0:
            reps = math.ceil(short / max(len(order), 1))
            order = order + (order * reps)[:short]
        return order

    def __iter__(self) -> Iterator[int]:
        order = self.global_order()
        mine = order[self.rank: self.total_size: self.num_replicas]
        assert len(mine) == self.num_samples
        return iter(mine)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)


# Name the reference scripts import.
DistributedSampler = ShardedSampler
