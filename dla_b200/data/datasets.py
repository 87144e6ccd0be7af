"""Datasets and collators for SFT, preference learning and distillation.

Behavioural parity (cited per class) with ``R/src/data/datasets.py``; the implementation is
organised differently: one shared text template + encoder, collators that produce
right-padded ``[B, T]`` tensors (optionally padded to a multiple so the GPU kernels see
TMA-friendly shapes) and an optional sequence-packing collator that honours the
``data.packing`` flag the reference declares but never reads (``R/config/sft_config.yaml:16``).
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence

import torch
from torch.utils.data import Dataset

IGNORE_INDEX = -100
SEP = "\n\n"


@dataclass
class Sample:
    """One record of any phase (parity: ``R/src/data/datasets.py:15-21``)."""

    prompt: str
    response: Optional[str] = None
    chosen: Optional[str] = None
    rejected: Optional[str] = None
    reward: Optional[float] = None


def read_jsonl(path: str | Path) -> List[Dict[str, Any]]:
    """Whole JSONL file -> list of dicts, blank lines skipped (``R/src/data/datasets.py:24-27``)."""
    rows: List[Dict[str, Any]] = []
    with Path(path).open("r", encoding="utf-8") as fh:
        for line in fh:
            if line.strip():
                rows.append(json.loads(line))
    return rows


def write_jsonl(path: str | Path, rows: Sequence[Dict[str, Any]]) -> None:
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    with path.open("w", encoding="utf-8") as fh:
        for r in rows:
            fh.write(json.dumps(r) + "\n")


def _eos_of(tokenizer: Any) -> str:
    return getattr(tokenizer, "eos_token", None) or getattr(tokenizer, "pad_token", None) or "</s>"


def _encode(tokenizer: Any, text: str, max_length: int, specials: bool = True) -> Dict[str, torch.Tensor]:
    enc = tokenizer(text, truncation=True, max_length=max_length, padding=False,
                    add_special_tokens=specials, return_tensors=None)
    ids = torch.as_tensor(enc["input_ids"], dtype=torch.long).reshape(-1)
    mask = torch.as_tensor(enc["attention_mask"], dtype=torch.long).reshape(-1)
    return {"input_ids": ids, "attention_mask": mask}


def pad_batch(batch: List[Dict[str, torch.Tensor]], pad_token_id: int,
              pad_to_multiple_of: int = 1) -> Dict[str, torch.Tensor]:
    """Right-pad every key to the longest item of the batch.

    Fill values follow ``R/src/data/datasets.py:212-229``: ``input_ids`` -> pad id,
    ``attention_mask`` -> 0, ``labels`` -> -100, anything else -> 0.  0-d tensors (e.g.
    ``reward``) are stacked instead of padded — the reference crashes on them (defect D10).
    """
    out: Dict[str, torch.Tensor] = {}
    for key in batch[0].keys():
        items = [ex[key] for ex in batch]
        if items[0].dim() == 0:
            out[key] = torch.stack(items, 0)
            continue
        fill = pad_token_id if key == "input_ids" else IGNORE_INDEX if key == "labels" else 0
        width = max(int(t.shape[0]) for t in items)
        if pad_to_multiple_of > 1:
            width = -(-width // pad_to_multiple_of) * pad_to_multiple_of
        buf = torch.full((len(items), width) + tuple(items[0].shape[1:]), fill, dtype=items[0].dtype)
        for i, t in enumerate(items):
            buf[i, : t.shape[0]] = t
        out[key] = buf
    return out


class _RecordDataset(Dataset):
    def __init__(self, tokenizer: Any, max_length: int, path: str | Path | None,
                 records: Optional[List[Dict[str, Any]]], what: str) -> None:
        if records is None and path is None:
            raise ValueError(f"{what}: give either records= or path=")
        self.records = list(records) if records is not None else read_jsonl(path)  # type: ignore[arg-type]
        self.tokenizer = tokenizer
        self.max_length = int(max_length)
        self.eos = _eos_of(tokenizer)
        self.pad_to_multiple_of = 1

    def __len__(self) -> int:
        return len(self.records)

    @property
    def pad_id(self) -> int:
        return getattr(self.tokenizer, "pad_token_id", None) or 0

    def _join(self, prompt: str, completion: str) -> str:
        return f"{prompt.strip()}{SEP}{completion.strip()}{self.eos}"


class InstructionDataset(_RecordDataset):
    """Prompt/response pairs for SFT.

    Text is ``prompt\\n\\nresponse<eos>`` tokenised on the fly with truncation; labels are
    the input ids with the prompt part set to -100 when ``mask_prompt``
    (``R/src/data/datasets.py:52-80``).  The prompt length is the length of
    ``tokenize(prompt + "\\n\\n", add_special_tokens=False)`` like the reference, plus the
    number of leading special tokens the full encoding added (fixes the BOS off-by-one,
    defect D11; identical when the tokenizer adds none).
    """

    def __init__(self, tokenizer: Any, max_length: int, mask_prompt: bool = True,
                 path: str | Path | None = None, records: Optional[List[Dict[str, Any]]] = None) -> None:
        super().__init__(tokenizer, max_length, path, records, "InstructionDataset")
        self.mask_prompt = mask_prompt

    def __getitem__(self, idx: int) -> Dict[str, torch.Tensor]:
        rec = self.records[idx]
        prompt = rec["prompt"].strip()
        enc = _encode(self.tokenizer, self._join(prompt, rec["response"]), self.max_length)
        labels = enc["input_ids"].clone()
        if self.mask_prompt:
            p = _encode(self.tokenizer, f"{prompt}{SEP}", self.max_length, specials=False)["input_ids"]
            n_prompt = int(p.shape[0])
            bos = getattr(self.tokenizer, "bos_token_id", None)
            if bos is not None and enc["input_ids"].numel() > 0 and int(enc["input_ids"][0]) == bos \
                    and (p.numel() == 0 or int(p[0]) != bos):
                n_prompt += 1
            labels[:n_prompt] = IGNORE_INDEX
        return {"input_ids": enc["input_ids"], "attention_mask": enc["attention_mask"], "labels": labels}

    def collate(self, batch: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        return pad_batch(batch, self.pad_id, self.pad_to_multiple_of)


class PreferenceDataset(_RecordDataset):
    """(prompt, chosen, rejected) triples for reward modelling and DPO.

    Both completions are encoded independently as ``prompt\\n\\ncompletion<eos>``; ``collate``
    returns the nested ``{"chosen": {...}, "rejected": {...}}`` layout, each side padded to
    its own longest sequence (``R/src/data/datasets.py:106-152``).
    """

    def __init__(self, tokenizer: Any, max_length: int, path: str | Path | None = None,
                 records: Optional[List[Dict[str, Any]]] = None) -> None:
        super().__init__(tokenizer, max_length, path, records, "PreferenceDataset")

    def __getitem__(self, idx: int) -> Dict[str, torch.Tensor]:
        rec = self.records[idx]
        item: Dict[str, torch.Tensor] = {}
        for side in ("chosen", "rejected"):
            enc = _encode(self.tokenizer, self._join(rec["prompt"], rec[side]), self.max_length)
            item[f"{side}_input_ids"] = enc["input_ids"]
            item[f"{side}_attention_mask"] = enc["attention_mask"]
        return item

    def collate(self, batch: List[Dict[str, torch.Tensor]]) -> Dict[str, Dict[str, torch.Tensor]]:
        out: Dict[str, Dict[str, torch.Tensor]] = {}
        for side in ("chosen", "rejected"):
            rows = [{"input_ids": ex[f"{side}_input_ids"], "attention_mask": ex[f"{side}_attention_mask"]}
                    for ex in batch]
            out[side] = pad_batch(rows, self.pad_id, self.pad_to_multiple_of)
        return out


class TeacherRolloutDataset(_RecordDataset):
    """Teacher responses for distillation: ``prompt\\n\\nteacher_response<eos>``, labels =
    input ids (no prompt masking), per-sample ``reward`` (default 1.0)
    (``R/src/data/datasets.py:172-190``).  ``collate`` stacks rewards separately, which is
    what the reference intends but crashes on (defect D10)."""

    def __init__(self, path: str | Path | None, tokenizer: Any, max_length: int,
                 records: Optional[List[Dict[str, Any]]] = None) -> None:
        super().__init__(tokenizer, max_length, path, records, "TeacherRolloutDataset")

    def __getitem__(self, idx: int) -> Dict[str, torch.Tensor]:
        rec = self.records[idx]
        enc = _encode(self.tokenizer, self._join(rec["prompt"], rec["teacher_response"]), self.max_length)
        return {"input_ids": enc["input_ids"], "attention_mask": enc["attention_mask"],
                "labels": enc["input_ids"].clone(),
                "reward": torch.tensor(float(rec.get("reward", 1.0)), dtype=torch.float32)}

    def collate(self, batch: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        return pad_batch(batch, self.pad_id, self.pad_to_multiple_of)


class EvalPromptDataset(Dataset):
    """Plain list of prompt records (``R/src/data/datasets.py:199-209``)."""

    def __init__(self, path: str | Path | None = None, records: Optional[List[Dict[str, Any]]] = None) -> None:
        self.records = list(records) if records is not None else read_jsonl(path)  # type: ignore[arg-type]

    def __len__(self) -> int:
        return len(self.records)

    def __getitem__(self, idx: int) -> Dict[str, Any]:
        return self.records[idx]


class PackedCollator:
    """Sequence packing for SFT (``data.packing: true``): concatenates the examples of a batch
    into rows of ``max_length`` tokens with ``cu_seqlens`` boundaries so attention never crosses
    documents.  Returns ``input_ids``/``labels`` ``[R, max_length]``, ``attention_mask`` and
    ``segment_ids`` (0 = padding)."""

    def __init__(self, max_length: int, pad_id: int) -> None:
        self.max_length, self.pad_id = max_length, pad_id

    def __call__(self, batch: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        rows: List[List[Dict[str, torch.Tensor]]] = [[]]
        used = 0
        for ex in sorted(batch, key=lambda e: -int(e["input_ids"].shape[0])):
            n = int(ex["input_ids"].shape[0])
            if used + n > self.max_length and rows[-1]:
                rows.append([])
                used = 0
            rows[-1].append(ex)
            used += n
        R = len(rows)
        ids = torch.full((R, self.max_length), self.pad_id, dtype=torch.long)
        lab = torch.full((R, self.max_length), IGNORE_INDEX, dtype=torch.long)
        seg = torch.zeros((R, self.max_length), dtype=torch.long)
        pos = torch.zeros((R, self.max_length), dtype=torch.long)
        for r, row in enumerate(rows):
            o = 0
            for s, ex in enumerate(row, 1):
                n = min(int(ex["input_ids"].shape[0]), self.max_length - o)
                ids[r, o:o + n] = ex["input_ids"][:n]
                l = ex["labels"][:n].clone()
                l[0] = IGNORE_INDEX  # never predict a document's first token from the previous one
                lab[r, o:o + n] = l
                seg[r, o:o + n] = s
                pos[r, o:o + n] = torch.arange(n)
                o += n
        return {"input_ids": ids, "labels": lab, "attention_mask": (seg > 0).long(),
                "segment_ids": seg, "position_ids": pos}


# ---------------------------------------------------------------------------------------
# record loaders (local JSONL or HF hub) — parity: R/src/data/datasets.py:232-296
# ---------------------------------------------------------------------------------------

def _hf_rows(cfg: Dict[str, Any], split: str):
    from datasets import load_dataset

    name = cfg.get(f"{split}_split") or cfg.get("split", split)
    return load_dataset(cfg["hf_path"], cfg.get("hf_name"), split=name, streaming=False)


def _load_records(cfg: Dict[str, Any], split: str, fields: Sequence[str], extra_paths: Sequence[str]) -> List[Dict[str, Any]]:
    source = cfg.get("source", "local")
    if source == "hf":
        cols = cfg.get("columns", {}) or {}
        template = cfg.get("template")
        records = []
        for row in _hf_rows(cfg, split):
            rec = {f: row[cols.get(f, f)] for f in fields}
            if template:
                rec["prompt"] = template.format(**row)
            records.append(rec)
    elif source == "synthetic":
        from dla_b200.data.synthetic import synthetic_records

        records = synthetic_records(fields, int(cfg.get("num_samples", 256)), int(cfg.get("seed", 0)),
                                    int(cfg.get("approx_chars", 256)))
    else:
        path = cfg.get(f"{split}_path") or cfg.get("path")
        for key in extra_paths:
            path = path or cfg.get(key)
        if not path:
            raise KeyError(f"data config has no path for split {split!r}")
        records = read_jsonl(path)
    limit = cfg.get("limit")
    if limit:
        records = records[: int(limit)]
    return records


def load_instruction_records(cfg: Dict[str, Any], split: str = "train") -> List[Dict[str, Any]]:
    """``source: local|hf|synthetic``; path keys ``{split}_path`` -> ``path``; split names
    ``{split}_split`` -> ``split``; column remap + ``template``; ``limit``
    (``R/src/data/datasets.py:232-260``)."""
    return _load_records(cfg, split, ("prompt", "response"), ())


def load_preference_records(cfg: Dict[str, Any], split: str = "train") -> List[Dict[str, Any]]:
    """As above for (prompt, chosen, rejected); local path also falls back to
    ``preference_path`` (``R/src/data/datasets.py:263-296``)."""
    return _load_records(cfg, split, ("prompt", "chosen", "rejected"), ("preference_path",))


def _max_len(cfg: Dict[str, Any], default: int) -> int:
    return int(cfg.get("max_length") or cfg.get("max_seq_length") or default)


def build_instruction_dataset(cfg: Dict[str, Any], tokenizer: Any, split: str = "train") -> InstructionDataset:
    """``max_length := max_length or max_seq_length or 2048``; ``mask_prompt`` default True
    (``R/src/data/datasets.py:299-306``)."""
    ds = InstructionDataset(tokenizer, _max_len(cfg, 2048), cfg.get("mask_prompt", True),
                            records=load_instruction_records(cfg, split))
    ds.pad_to_multiple_of = int(cfg.get("pad_to_multiple_of", 1))
    return ds


def build_preference_dataset(cfg: Dict[str, Any], tokenizer: Any, split: str = "train") -> PreferenceDataset:
    """Default ``max_length`` 1024 (``R/src/data/datasets.py:309-315``)."""
    ds = PreferenceDataset(tokenizer, _max_len(cfg, 1024), records=load_preference_records(cfg, split))
    ds.pad_to_multiple_of = int(cfg.get("pad_to_multiple_of", 1))
    return ds
