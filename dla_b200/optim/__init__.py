"""Learning-rate schedules and optimizer config.

Parity: the SFT trainer builds ``transformers.get_scheduler(name, optimizer, num_warmup_steps,
num_training_steps)`` with ``lr_scheduler`` default ``"cosine"`` and steps it once per micro-batch
(``R/src/training/train_sft.py:105-110,150``, defect D2); the other trainers declare
``lr_scheduler``/``warmup_steps`` but run a constant LR.  :func:`lr_at` restates the HF
lambda schedules; trainers other than SFT use it only when ``optimization.honor_scheduler: true``.
AdamW itself is the fused kernel driven by :class:`dla_b200.parallel.engine.Engine`.
"""
from __future__ import annotations

import math

from dla_b200.parallel.engine import AdamWConfig  # noqa: F401


def lr_multiplier(name: str, step: int, warmup: int, total: int) -> float:
    """Multiplier applied to the base LR at (0-based) scheduler step ``step``."""
    name = (name or "constant").lower()
    if name == "constant":
        return 1.0
    if step < warmup:
        return step / max(1, warmup)
    if name == "constant_with_warmup":
        return 1.0
    if name == "linear":
        return max(0.0, (total - step) / max(1, total - warmup))
    if name == "cosine":
        progress = (step - warmup) / max(1, total - warmup)
        return max(0.0, 0.5 * (1.0 + math.cos(math.pi * progress)))
    raise ValueError(f"unknown lr_scheduler {name!r} (constant, constant_with_warmup, linear, cosine)")


def lr_at(base_lr: float, name: str, step: int, warmup: int, total: int) -> float:
    return base_lr * lr_multiplier(name, step, warmup, total)
