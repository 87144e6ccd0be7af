"""Synthetic data: there is no network on the build / benchmark boxes, so every phase can
run on generated records (``data.source: synthetic``) or on fixed-shape token tensors for
benchmarks (``BASELINE.json``: "synthetic preference data of the named shape")."""
from __future__ import annotations

import random
from typing import Any, Dict, List, Sequence

import torch

_WORDS = ("align model reward policy prefer human answer question safe helpful honest explain why how "
          "the a of to and in is that it for with as on be this by are or from at an which can").split()


def _sentence(rng: random.Random, approx_chars: int) -> str:
    out: List[str] = []
    n = 0
    while n < approx_chars:
        w = rng.choice(_WORDS)
        out.append(w)
        n += len(w) + 1
    return " ".join(out)


def synthetic_records(fields: Sequence[str], num_samples: int, seed: int = 0, approx_chars: int = 256) -> List[Dict[str, Any]]:
    rng = random.Random(seed)
    rows = []
    for i in range(num_samples):
        row: Dict[str, Any] = {}
        for f in fields:
            if f == "reward":
                row[f] = rng.uniform(-1.0, 1.0)
            elif f == "prompt":
                row[f] = f"Q{i}: " + _sentence(rng, approx_chars // 4)
            else:
                row[f] = _sentence(rng, approx_chars)
        rows.append(row)
    return rows


def synthetic_preference_batch(batch: int, seq_len: int, vocab_size: int, seed: int = 0,
                               pin: bool = False, ragged: bool = False) -> Dict[str, Dict[str, torch.Tensor]]:
    """Fixed-length token-id pairs in the nested collate layout of PreferenceDataset."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, Dict[str, torch.Tensor]] = {}
    for side in ("chosen", "rejected"):
        ids = torch.randint(4, vocab_size, (batch, seq_len), generator=g, dtype=torch.long)
        mask = torch.ones(batch, seq_len, dtype=torch.long)
        if ragged:
            lens = torch.randint(max(2, seq_len // 2), seq_len + 1, (batch,), generator=g)
            mask = (torch.arange(seq_len)[None, :] < lens[:, None]).long()
            ids = ids * mask
        if pin and torch.cuda.is_available():
            ids, mask = ids.pin_memory(), mask.pin_memory()
        out[side] = {"input_ids": ids, "attention_mask": mask}
    return out


def synthetic_lm_batch(batch: int, seq_len: int, vocab_size: int, seed: int = 0, pin: bool = False) -> Dict[str, torch.Tensor]:
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(4, vocab_size, (batch, seq_len), generator=g, dtype=torch.long)
    mask = torch.ones(batch, seq_len, dtype=torch.long)
    labels = ids.clone()
    if pin and torch.cuda.is_available():
        ids, mask, labels = ids.pin_memory(), mask.pin_memory(), labels.pin_memory()
    return {"input_ids": ids, "attention_mask": mask, "labels": labels}
