"""Rank-sharded sampling.

Parity: ``get_distributed_sampler`` returns a ``DistributedSampler``-equivalent when the
world is larger than one process and ``None`` otherwise (``R/src/training/utils.py:110-118``);
trainers call ``set_epoch(epoch)`` every epoch (``R/src/training/train_sft.py:141-142``).
:class:`RankSampler` additionally exposes ``state_dict``/``load_state_dict`` so a resumed
run continues mid-epoch (the reference cannot resume at all).
"""
from __future__ import annotations

import math
from typing import Any, Dict, Iterator, Optional, Sized

import torch
from torch.utils.data import Sampler


class RankSampler(Sampler[int]):
    """Deterministic per-rank partition with epoch reshuffle.

    Indices are a seeded permutation of ``range(len(dataset))``, padded by wrap-around to a
    multiple of ``num_replicas`` and strided ``rank::num_replicas`` — every rank gets the
    same count, the union covers the dataset, and shards are disjoint up to the padding.
    """

    def __init__(self, dataset: Sized, num_replicas: int = 1, rank: int = 0, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False) -> None:
        if not 0 <= rank < num_replicas:
            raise ValueError(f"rank {rank} not in [0, {num_replicas})")
        self.n = len(dataset)
        self.num_replicas, self.rank, self.shuffle, self.seed, self.drop_last = num_replicas, rank, shuffle, seed, drop_last
        self.epoch = 0
        self.start = 0  # samples of this rank skipped at the beginning of the current epoch (resume point)
        self.consumed = 0  # samples of this rank handed to the training loop so far in the current epoch
        if drop_last and self.n % num_replicas:
            self.num_samples = self.n // num_replicas
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        if epoch != self.epoch:
            self.start = self.consumed = 0
        self.epoch = epoch

    def _order(self) -> list[int]:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.n, generator=g).tolist()
        else:
            order = list(range(self.n))
        if self.drop_last:
            order = order[: self.total_size]
        elif self.total_size > len(order) and order:
            reps = math.ceil((self.total_size - len(order)) / len(order))
            order = (order + order * reps)[: self.total_size]
        return order

    def __iter__(self) -> Iterator[int]:
        mine = self._order()[self.rank : self.total_size : self.num_replicas]
        return iter(mine[self.start :])

    def __len__(self) -> int:
        return self.num_samples - self.start

    def state_dict(self) -> Dict[str, Any]:
        """``start`` is the number of this rank's samples the loop has already trained on in ``epoch``
        (``infinite_batches`` advances ``consumed`` as it yields), so a resumed run skips exactly those."""
        return {"epoch": self.epoch, "start": min(max(self.consumed, self.start), self.num_samples), "seed": self.seed}

    def load_state_dict(self, state: Dict[str, Any]) -> None:
        self.epoch, self.start = int(state.get("epoch", 0)), int(state.get("start", 0))
        self.seed = int(state.get("seed", self.seed))
        if self.start >= self.num_samples:      # saved exactly at an epoch boundary
            self.epoch, self.start = self.epoch + 1, 0
        self.consumed = self.start


def get_distributed_sampler(dataset: Sized, world_size: int, rank: int, shuffle: bool = True,
                            seed: int = 0, always: bool = False) -> Optional[RankSampler]:
    """``None`` for a single process, else a rank partition (``R/src/training/utils.py:110-118``).  ``always=True``
    also gives a single process a (one-replica) sampler so that its position in the epoch can be saved and resumed."""
    if world_size > 1 or always:
        return RankSampler(dataset, world_size, rank, shuffle=shuffle, seed=seed)
    return None


def split_between_processes(items: list, world_size: int, rank: int) -> list:
    """Rank-local contiguous slice of an identically ordered list — what the reference means
    at ``R/src/training/train_rlhf.py:114`` (defect D7: it uses a context manager as a
    function).  Earlier ranks take the remainder, like accelerate's splitter."""
    n = len(items)
    base, rem = divmod(n, world_size)
    start = rank * base + min(rank, rem)
    return items[start : start + base + (1 if rank < rem else 0)]
