"""Name-for-name counterpart of the reference's shared helpers (``R/src/training/utils.py``), for code that was
written against that module.  Each function keeps the reference's name and intent and maps it onto this
framework's runtime (process group + parameter engine instead of ``accelerate.Accelerator``)."""
from __future__ import annotations

from pathlib import Path
from typing import Any, Dict, Iterable, Optional, Tuple

from dla_b200.ckpt import CheckpointManager
from dla_b200.config import HardwarePlan, load_config, resolve_hardware  # noqa: F401  (load_config: utils.py:18-21)
from dla_b200.data.sampler import get_distributed_sampler as _sampler
from dla_b200.parallel.engine import Engine
from dla_b200.parallel.pg import DistContext, init_distributed
from dla_b200.utils import (RunningLoss, flatten_dict, prepare_output_dirs, save_json, seed_everything,  # noqa: F401
                            yield_batch as _yield_batch)


def get_accelerator(config: Dict[str, Any]) -> Tuple[DistContext, HardwarePlan]:
    """``R/src/training/utils.py:55-75``: reads ``hardware.{deepspeed_config, fsdp, gradient_accumulation_steps,
    mixed_precision}`` and returns the distributed context plus the resolved hardware plan (the two things the
    reference's ``Accelerator`` object bundles)."""
    ctx = init_distributed(device=(config.get("hardware") or {}).get("device"))
    return ctx, resolve_hardware(config)


def broadcast_config(ctx: DistContext, config: Dict[str, Any]) -> Dict[str, Any]:
    """``R/src/training/utils.py:94-96``: every rank ends up with rank 0's config."""
    return ctx.broadcast_object(config, src=0)


def save_accelerator_state(ctx: DistContext, engines: Dict[str, Engine], output_dir: str, tag: str,
                           trainer_state: Optional[Dict[str, Any]] = None, **kw: Any) -> Path:
    """``R/src/training/utils.py:99-102``: checkpoint under ``output_dir/tag`` (+ ``latest`` link)."""
    return CheckpointManager(output_dir, ctx).save(tag, engines, trainer_state or {}, **kw)


def log_rank_zero(ctx: DistContext, message: str) -> None:
    """``R/src/training/utils.py:105-107``."""
    ctx.print0(message)


def get_distributed_sampler(dataset, ctx: DistContext, shuffle: bool = True, seed: int = 0):
    """``R/src/training/utils.py:110-118``: ``None`` for a single process, else the rank's partition."""
    return _sampler(dataset, ctx.world_size, ctx.rank, shuffle=shuffle, seed=seed)


def maybe_clip_gradients(engine: Engine, max_norm: Optional[float]) -> Optional[float]:
    """``R/src/training/utils.py:121-123``.  Clipping is part of ``Engine.optimizer_step`` (global norm over
    all shards, coefficient applied inside the fused AdamW kernel); this returns the value to put in
    ``AdamWConfig.max_grad_norm`` (``None``/0 disables it)."""
    return max_norm if max_norm and max_norm > 0 else 0.0


def yield_batch(iterable: Iterable[Dict[str, Any]], ctx: DistContext):
    """``R/src/training/utils.py:126-129``: move each batch to the rank's device."""
    return _yield_batch(iterable, ctx.device)
