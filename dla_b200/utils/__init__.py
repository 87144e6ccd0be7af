"""Small shared helpers (parity: ``R/src/training/utils.py``)."""
from __future__ import annotations

import json
import os
import random
from dataclasses import dataclass
from pathlib import Path
from typing import Any, Dict, Iterable, Iterator

import numpy as np
import torch


def seed_everything(seed: int) -> None:
    """Seed python / numpy / torch (CPU and every CUDA device) with one integer.

    Parity: ``R/src/training/utils.py:24-29``.  Every rank gets the same seed, like the
    reference; per-rank data order comes from the sampler, not from the seed.
    """
    random.seed(seed)
    np.random.seed(seed % (2**32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def save_json(obj: Any, path: str | os.PathLike) -> None:
    """mkdir -p + indent-2 JSON dump (``R/src/training/utils.py:32-36``), written atomically."""
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    tmp = path.with_suffix(path.suffix + ".tmp")
    with tmp.open("w", encoding="utf-8") as fh:
        json.dump(obj, fh, indent=2)
    os.replace(tmp, path)


@dataclass
class RunningLoss:
    """Count-weighted running mean, reset by the caller after each log line.

    Parity: ``R/src/training/utils.py:39-52``.
    """

    total: float = 0.0
    count: int = 0

    def update(self, value: float, n: int = 1) -> None:
        self.total += float(value) * n
        self.count += n

    @property
    def average(self) -> float:
        return self.total / max(self.count, 1)

    def reset(self) -> None:
        self.total, self.count = 0.0, 0


def prepare_output_dirs(*paths: str | os.PathLike) -> None:
    """``mkdir -p`` each path (``R/src/training/utils.py:78-80``)."""
    for p in paths:
        if p:
            Path(p).mkdir(parents=True, exist_ok=True)


def flatten_dict(cfg: Dict[str, Any], parent: str = "", sep: str = ".") -> Dict[str, Any]:
    """Nested dict -> dotted keys (``R/src/training/utils.py:83-91``)."""
    out: Dict[str, Any] = {}
    for k, v in cfg.items():
        key = f"{parent}{sep}{k}" if parent else str(k)
        if isinstance(v, dict):
            out.update(flatten_dict(v, key, sep))
        else:
            out[key] = v
    return out


def unflatten_dict(flat: Dict[str, Any], sep: str = ".") -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for k, v in flat.items():
        cur = out
        parts = k.split(sep)
        for p in parts[:-1]:
            cur = cur.setdefault(p, {})
        cur[parts[-1]] = v
    return out


def chunk_list(items: list, size: int) -> list[list]:
    """Split a list into consecutive chunks (``R/src/training/generate_teacher_data.py:30-31``)."""
    return [items[i : i + size] for i in range(0, len(items), size)]


def yield_batch(iterable: Iterable[Dict[str, Any]], device: torch.device | str) -> Iterator[Dict[str, Any]]:
    """Move each batch of an iterable to ``device`` (the useful reading of
    ``R/src/training/utils.py:126-129``, which calls ``accelerator.prepare`` per item)."""
    from dla_b200.data.loader import to_device

    for batch in iterable:
        yield to_device(batch, device)


def human_bytes(n: float) -> str:
    for unit in ("B", "KiB", "MiB", "GiB", "TiB"):
        if abs(n) < 1024:
            return f"{n:.1f}{unit}"
        n /= 1024
    return f"{n:.1f}PiB"
