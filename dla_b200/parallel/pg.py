"""Process-group plumbing: one process per GPU, env rendezvous, NCCL on CUDA / gloo on CPU.

The reference never touches ``torch.distributed`` itself — everything goes through
``accelerate.Accelerator`` (``R/src/training/utils.py:55-75``) with a static single-node rendezvous on port
29500 (``R/config/accelerate_config.yaml:7-13``).  Here the launcher (``dla_b200.cli.launch`` or torchrun)
sets ``RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT`` and :func:`init_distributed` does the
rest.  The helpers mirror the accelerate calls the trainers used: ``gather`` (C4/C5), ``wait_for_everyone``
(C8), ``broadcast`` of objects (C10), rank-0 printing.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Any, Optional

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    backend: str = "none"
    group: Optional[Any] = None
    fused_comm: Optional[Any] = None       # set by parallel.fused.FusedComm: peer-memory path for tiny collectives

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1

    # ---- collectives on small tensors / objects ---------------------------------------
    def barrier(self) -> None:
        """``accelerator.wait_for_everyone()`` (``R/src/training/train_rlhf.py:164``)."""
        if self.distributed:
            if self.backend == "nccl":
                dist.barrier(group=self.group, device_ids=[self.local_rank])
            else:
                dist.barrier(group=self.group)

    def all_reduce(self, t: torch.Tensor, op: str = "sum") -> torch.Tensor:
        if self.distributed:
            ops = {"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}
            if op == "mean":
                dist.all_reduce(t, dist.ReduceOp.SUM, group=self.group)
                t.div_(self.world_size)
            else:
                dist.all_reduce(t, ops[op], group=self.group)
        return t

    def gather(self, t: torch.Tensor) -> torch.Tensor:
        """``accelerator.gather``: concatenate along dim 0 over ranks (``R/src/training/train_sft.py:39``)."""
        if not self.distributed:
            return t
        t = t.contiguous()
        fc = self.fused_comm
        if fc is not None and t.is_cuda and t.is_floating_point() and 0 < t.numel() <= 2048 and t.dim() >= 1:
            # eval losses / logged statistics: one single-block peer-memory kernel instead of an NCCL launch
            return fc.all_gather_small(t).to(t.dtype).view((self.world_size * t.shape[0],) + tuple(t.shape[1:]))
        out = torch.empty((self.world_size * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out, t, group=self.group)
        return out

    def broadcast_object(self, obj: Any, src: int = 0) -> Any:
        """``accelerator.broadcast(config)`` (``R/src/training/utils.py:94-96``)."""
        if not self.distributed:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src, group=self.group)
        return box[0]

    def max_over_ranks(self, value: float) -> float:
        if not self.distributed:
            return value
        t = torch.tensor([value], dtype=torch.float64, device=self.device if self.backend == "nccl" else "cpu")
        dist.all_reduce(t, dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def print0(self, *a: Any, **k: Any) -> None:
        """Rank-0 print (``log_rank_zero``, ``R/src/training/utils.py:105-107``)."""
        if self.is_main:
            print(*a, **k, flush=True)


_CTX: Optional[DistContext] = None


def init_distributed(backend: Optional[str] = None, device: Optional[str] = None, timeout_s: int = 1800) -> DistContext:
    """Initialise (idempotently) from the torchrun-style environment; single process if WORLD_SIZE is unset."""
    global _CTX
    if _CTX is not None:
        return _CTX
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    if use_cuda:
        dev = torch.device("cuda", local_rank % max(torch.cuda.device_count(), 1))
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    ctx = DistContext(rank=rank, world_size=world, local_rank=local_rank, device=dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        ctx.backend = backend or ("nccl" if use_cuda else "gloo")
        if not dist.is_initialized():
            kw = {"device_id": dev} if ctx.backend == "nccl" else {}
            dist.init_process_group(ctx.backend, rank=rank, world_size=world,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kw)
        ctx.group = dist.group.WORLD
    if use_cuda:
        from dla_b200.ops import native
        if native.available():
            native.bind_device(dev.index)
    _CTX = ctx
    return ctx


def get_context() -> DistContext:
    return _CTX if _CTX is not None else init_distributed()


def shutdown() -> None:
    global _CTX
    if dist.is_available() and dist.is_initialized():
        try:
            dist.destroy_process_group()
        except Exception:  # pragma: no cover
            pass
    _CTX = None


def log_rank_zero(ctx: DistContext, message: str) -> None:
    ctx.print0(message)
