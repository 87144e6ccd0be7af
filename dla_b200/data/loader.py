"""DataLoader construction and pinned host->device staging.

Parity: the reference builds ``DataLoader(batch_size=micro, sampler=…, shuffle=sampler is
None, collate_fn=dataset.collate, num_workers, pin_memory=True)`` in every trainer
(e.g. ``R/src/training/train_dpo.py:62-71``).  :class:`DevicePrefetcher` adds the B200 part:
batches land in pinned memory and are copied on a side stream one step ahead so the H2D copy
overlaps the previous step's compute.
"""
from __future__ import annotations

from typing import Any, Callable, Iterable, Iterator, Optional

import torch
from torch.utils.data import DataLoader, Dataset

from dla_b200.data.sampler import RankSampler, get_distributed_sampler


def to_device(batch: Any, device: torch.device | str, non_blocking: bool = True) -> Any:
    if isinstance(batch, torch.Tensor):
        return batch.to(device, non_blocking=non_blocking)
    if isinstance(batch, dict):
        return {k: to_device(v, device, non_blocking) for k, v in batch.items()}
    if isinstance(batch, (list, tuple)):
        return type(batch)(to_device(v, device, non_blocking) for v in batch)
    return batch


def batch_nbytes(batch: Any) -> int:
    if isinstance(batch, torch.Tensor):
        return batch.numel() * batch.element_size()
    if isinstance(batch, dict):
        return sum(batch_nbytes(v) for v in batch.values())
    if isinstance(batch, (list, tuple)):
        return sum(batch_nbytes(v) for v in batch)
    return 0


def build_loader(dataset: Dataset, micro_batch_size: int, world_size: int, rank: int,
                 shuffle: bool = True, num_workers: int = 0, seed: int = 0,
                 collate_fn: Optional[Callable] = None, drop_last: bool = False):
    # a single process gets a one-replica sampler too (the reference passes ``shuffle=True`` there): same seeded
    # permutation every run, and the position inside the epoch becomes part of the checkpoint
    sampler = get_distributed_sampler(dataset, world_size, rank, shuffle=shuffle, seed=seed, always=shuffle)  # type: ignore[arg-type]
    loader = DataLoader(
        dataset,
        batch_size=micro_batch_size,
        sampler=sampler,
        shuffle=(sampler is None and shuffle),
        collate_fn=collate_fn or getattr(dataset, "collate", None),
        num_workers=num_workers,
        pin_memory=torch.cuda.is_available(),
        drop_last=drop_last,
        persistent_workers=num_workers > 0,
    )
    return loader, sampler


class DevicePrefetcher:
    """Wraps an iterable of (pinned) host batches; yields device batches, copying batch
    ``i+1`` on a dedicated stream while batch ``i`` is being consumed."""

    def __init__(self, loader: Iterable, device: torch.device | str) -> None:
        self.loader, self.device = loader, torch.device(device)
        self.stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def __iter__(self) -> Iterator[Any]:
        it = iter(self.loader)
        if self.stream is None:
            for b in it:
                yield to_device(b, self.device)
            return
        nxt = self._stage(it)
        while nxt is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            cur = nxt
            _record(cur, torch.cuda.current_stream(self.device))
            nxt = self._stage(it)
            yield cur

    def _stage(self, it: Iterator[Any]) -> Any:
        try:
            host = next(it)
        except StopIteration:
            return None
        with torch.cuda.stream(self.stream):
            return to_device(host, self.device, non_blocking=True)


def _record(batch: Any, stream: "torch.cuda.Stream") -> None:
    if isinstance(batch, torch.Tensor):
        batch.record_stream(stream)
    elif isinstance(batch, dict):
        for v in batch.values():
            _record(v, stream)
    elif isinstance(batch, (list, tuple)):
        for v in batch:
            _record(v, stream)


def infinite_batches(loader: DataLoader, sampler: Optional[RankSampler]) -> Iterator[Any]:
    """Epoch loop of the reference (``for epoch in range(10_000): sampler.set_epoch(epoch)``,
    ``R/src/training/train_dpo.py:105-109``) as a generator of (epoch, batch)."""
    epoch = sampler.epoch if sampler is not None else 0
    per_batch = int(getattr(getattr(loader, "loader", loader), "batch_size", None) or 1)
    while True:
        if sampler is not None:
            sampler.set_epoch(epoch)
        n = 0
        base = sampler.start if sampler is not None else 0
        for batch in loader:
            n += 1
            if sampler is not None:     # counted when handed out (a prefetcher may have staged one more batch already)
                sampler.consumed = base + n * per_batch
            yield epoch, batch
        if n == 0:
            raise RuntimeError("dataset produced no batches")
        epoch += 1
        if sampler is not None:
            sampler.start = sampler.consumed = 0
