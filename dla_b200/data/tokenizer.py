"""Tokenizers.

The reference uses ``AutoTokenizer.from_pretrained(path, use_fast=True)`` and sets
``pad_token = eos_token`` when missing (``R/src/models/base_model.py:23-25``).  This box has
no hub access, so the default is :class:`ByteTokenizer` — a byte-level tokenizer with the
subset of the HF calling convention the framework uses (``__call__`` with truncation /
padding / ``return_tensors="pt"``, ``decode``, ``pad_token_id`` ...).  If a directory holds
real HF tokenizer files, :func:`load_tokenizer` loads them through transformers instead.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Any, List, Optional, Sequence, Union

import torch


class BatchEncoding(dict):
    """dict with attribute access and ``.to(device)`` like HF's BatchEncoding."""

    def __getattr__(self, item: str) -> Any:
        try:
            return self[item]
        except KeyError as e:  # pragma: no cover
            raise AttributeError(item) from e

    def to(self, device: Union[str, torch.device]) -> "BatchEncoding":
        for k, v in list(self.items()):
            if isinstance(v, torch.Tensor):
                self[k] = v.to(device, non_blocking=True)
        return self


class ByteTokenizer:
    """UTF-8 byte tokenizer: ids 0..3 are specials, byte b -> id b + 4.

    ``vocab_size`` may be larger than 260 (model vocab, e.g. 128256); ids above 259 are
    never produced by encoding and decode to nothing.
    """

    PAD, BOS, EOS, UNK = 0, 1, 2, 3
    OFFSET = 4

    def __init__(self, vocab_size: int = 260, add_bos: bool = False, model_max_length: int = 1 << 20,
                 pad_equals_eos: bool = False, padding_side: str = "right") -> None:
        if vocab_size < 260:
            raise ValueError("ByteTokenizer needs vocab_size >= 260")
        self.vocab_size = vocab_size
        self.add_bos = add_bos
        self.model_max_length = model_max_length
        self.padding_side = padding_side
        self.bos_token, self.eos_token, self.unk_token = "<s>", "</s>", "<unk>"
        self.bos_token_id, self.eos_token_id, self.unk_token_id = self.BOS, self.EOS, self.UNK
        if pad_equals_eos:  # the reference's convention when the tokenizer has no pad token
            self.pad_token, self.pad_token_id = self.eos_token, self.EOS
        else:
            self.pad_token, self.pad_token_id = "<pad>", self.PAD

    # -- encoding ---------------------------------------------------------------------
    def encode(self, text: str, add_special_tokens: bool = True) -> List[int]:
        ids: List[int] = []
        if add_special_tokens and self.add_bos:
            ids.append(self.BOS)
        # the literal eos string maps to the EOS id so "…{eos}" templates work
        parts = text.split(self.eos_token)
        for i, part in enumerate(parts):
            ids.extend(b + self.OFFSET for b in part.encode("utf-8"))
            if i < len(parts) - 1:
                ids.append(self.EOS)
        return ids

    def __call__(self, text: Union[str, Sequence[str]], truncation: bool = False,
                 max_length: Optional[int] = None, padding: Union[bool, str] = False,
                 return_tensors: Optional[str] = None, add_special_tokens: bool = True,
                 **_: Any) -> BatchEncoding:
        single = isinstance(text, str)
        texts = [text] if single else list(text)
        seqs = [self.encode(t, add_special_tokens) for t in texts]
        if truncation and max_length is not None:
            seqs = [s[:max_length] for s in seqs]
        masks = [[1] * len(s) for s in seqs]
        if padding and len(seqs) > 0:
            width = max(len(s) for s in seqs)
            if padding == "max_length" and max_length is not None:
                width = max_length
            for s, m in zip(seqs, masks):
                gap = width - len(s)
                if self.padding_side == "left":
                    s[:0] = [self.pad_token_id] * gap
                    m[:0] = [0] * gap
                else:
                    s.extend([self.pad_token_id] * gap)
                    m.extend([0] * gap)
        if return_tensors == "pt":
            if not padding and len({len(s) for s in seqs}) > 1:
                raise ValueError("return_tensors='pt' needs padding=True for ragged batches")
            return BatchEncoding(input_ids=torch.tensor(seqs, dtype=torch.long),
                                 attention_mask=torch.tensor(masks, dtype=torch.long))
        if single:
            return BatchEncoding(input_ids=seqs[0], attention_mask=masks[0])
        return BatchEncoding(input_ids=seqs, attention_mask=masks)

    # -- decoding ---------------------------------------------------------------------
    def decode(self, ids: Union[Sequence[int], torch.Tensor], skip_special_tokens: bool = True) -> str:
        if isinstance(ids, torch.Tensor):
            ids = ids.tolist()
        out = bytearray()
        pieces: List[str] = []
        for i in ids:
            if self.OFFSET <= i < self.OFFSET + 256:
                out.append(i - self.OFFSET)
            elif not skip_special_tokens and i in (self.PAD, self.BOS, self.EOS, self.UNK):
                pieces.append(out.decode("utf-8", errors="replace"))
                out = bytearray()
                pieces.append({self.PAD: "<pad>", self.BOS: "<s>", self.EOS: "</s>", self.UNK: "<unk>"}[i])
        pieces.append(out.decode("utf-8", errors="replace"))
        return "".join(pieces)

    def batch_decode(self, batch: Any, skip_special_tokens: bool = True) -> List[str]:
        return [self.decode(x, skip_special_tokens) for x in batch]

    # -- persistence ------------------------------------------------------------------
    def save_pretrained(self, path: str | Path) -> None:
        path = Path(path)
        path.mkdir(parents=True, exist_ok=True)
        with (path / "dla_tokenizer.json").open("w") as fh:
            json.dump({"type": "byte", "vocab_size": self.vocab_size, "add_bos": self.add_bos,
                       "pad_equals_eos": self.pad_token_id == self.EOS,
                       "padding_side": self.padding_side}, fh, indent=2)

    @classmethod
    def from_pretrained(cls, path: str | Path) -> "ByteTokenizer":
        with (Path(path) / "dla_tokenizer.json").open() as fh:
            meta = json.load(fh)
        return cls(vocab_size=meta["vocab_size"], add_bos=meta.get("add_bos", False),
                   pad_equals_eos=meta.get("pad_equals_eos", False),
                   padding_side=meta.get("padding_side", "right"))

    def __len__(self) -> int:
        return self.vocab_size


def load_tokenizer(path: Optional[str | Path], vocab_size: int = 260) -> Any:
    """Tokenizer for a model path.

    dla tokenizer file -> ByteTokenizer; HF tokenizer files -> transformers (offline);
    otherwise a fresh ByteTokenizer sized to the model vocab.  ``pad_token`` falls back to
    ``eos_token`` exactly like ``R/src/models/base_model.py:24-25``.
    """
    if path is not None:
        p = Path(path)
        if (p / "dla_tokenizer.json").exists():
            return ByteTokenizer.from_pretrained(p)
        if p.is_dir() and any((p / f).exists() for f in ("tokenizer.json", "tokenizer.model", "vocab.json")):
            from transformers import AutoTokenizer

            tok = AutoTokenizer.from_pretrained(str(p), use_fast=True)
            if tok.pad_token is None:
                tok.pad_token = tok.eos_token
            return tok
    return ByteTokenizer(vocab_size=max(vocab_size, 260))
