"""Checkpoint interop: units <-> HF-named safetensors.

Downstream phases of the reference load ``checkpoints/<phase>/latest`` with ``from_pretrained``
(``R/config/dpo_config.yaml:6-7``, ``R/config/rlhf_config.yaml:5-6``) and the PPO script looks for
``<reward path>/pytorch_model.bin`` (``R/src/training/train_rlhf.py:79-84``) — yet nothing in the reference
writes those (defect D9).  This module writes/reads exactly that layout: ``config.json``, tokenizer files
and ``model-0000k-of-0000n.safetensors`` with HF parameter names (fused QKV / gate-up are split on
export and re-fused on import).
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Callable, Dict, List, Optional, Tuple

import torch

from dla_b200.models.config import ModelConfig
from dla_b200.models.transformer import UnitSpec, Weights


def export_unit(cfg: ModelConfig, unit: str, w: Weights) -> Dict[str, torch.Tensor]:
    """One unit's parameters under HF names."""
    out: Dict[str, torch.Tensor] = {}
    if cfg.arch == "gpt2":
        if unit == "embed":
            out["transformer.wte.weight"] = w["tok_emb"]
            out["transformer.wpe.weight"] = w["pos_emb"]
        elif unit == "head":
            out["transformer.ln_f.weight"], out["transformer.ln_f.bias"] = w["lnf_w"], w["lnf_b"]
            if "lm_head" in w:
                out["lm_head.weight"] = w["lm_head"]
        elif unit.startswith("layer."):
            p = f"transformer.h.{unit.split('.')[1]}."
            out[p + "ln_1.weight"], out[p + "ln_1.bias"] = w["ln1_w"], w["ln1_b"]
            out[p + "attn.c_attn.weight"], out[p + "attn.c_attn.bias"] = w["wqkv"].t(), w["bqkv"]   # HF Conv1D: [in, out]
            out[p + "attn.c_proj.weight"], out[p + "attn.c_proj.bias"] = w["wo"].t(), w["bo"]
            out[p + "ln_2.weight"], out[p + "ln_2.bias"] = w["ln2_w"], w["ln2_b"]
            out[p + "mlp.c_fc.weight"], out[p + "mlp.c_fc.bias"] = w["w_fc"].t(), w["b_fc"]
            out[p + "mlp.c_proj.weight"], out[p + "mlp.c_proj.bias"] = w["w_proj"].t(), w["b_proj"]
    else:
        if unit == "embed":
            out["model.embed_tokens.weight"] = w["tok_emb"]
        elif unit == "head":
            out["model.norm.weight"] = w["final_norm"]
            if "lm_head" in w:
                out["lm_head.weight"] = w["lm_head"]
        elif unit.startswith("layer."):
            p = f"model.layers.{unit.split('.')[1]}."
            q, k, v = w["wqkv"].split([cfg.q_dim, cfg.kv_dim, cfg.kv_dim], dim=0)
            g, u = w["w_gu"].split([cfg.intermediate_size, cfg.intermediate_size], dim=0)
            out[p + "input_layernorm.weight"] = w["attn_norm"]
            out[p + "self_attn.q_proj.weight"], out[p + "self_attn.k_proj.weight"], out[p + "self_attn.v_proj.weight"] = q, k, v
            out[p + "self_attn.o_proj.weight"] = w["wo"]
            out[p + "post_attention_layernorm.weight"] = w["mlp_norm"]
            out[p + "mlp.gate_proj.weight"], out[p + "mlp.up_proj.weight"] = g, u
            out[p + "mlp.down_proj.weight"] = w["w_down"]
    if unit == "reward_head":
        out["scorer.1.weight"], out["scorer.1.bias"] = w["score_w"], w["score_b"]
    if unit == "value_head":
        out["value_head.weight"], out["value_head.bias"] = w["value_w"], w["value_b"]
    return {k: t.detach().contiguous() for k, t in out.items()}


def import_unit(cfg: ModelConfig, unit: str, spec: UnitSpec, get: Callable[[str], Optional[torch.Tensor]]) -> Dict[str, torch.Tensor]:
    """Inverse of :func:`export_unit`: returns ``param name -> tensor`` for the params found."""
    names = {p.name for p in spec.params}
    out: Dict[str, torch.Tensor] = {}

    def put(pname: str, t: Optional[torch.Tensor]) -> None:
        if t is not None and pname in names:
            out[pname] = t

    if cfg.arch == "gpt2":
        if unit == "embed":
            put("tok_emb", get("transformer.wte.weight"))
            put("pos_emb", get("transformer.wpe.weight"))
        elif unit == "head":
            put("lnf_w", get("transformer.ln_f.weight"))
            put("lnf_b", get("transformer.ln_f.bias"))
            put("lm_head", get("lm_head.weight"))
        elif unit.startswith("layer."):
            p = f"transformer.h.{unit.split('.')[1]}."
            tr = lambda t: None if t is None else t.t()
            for a, b in (("ln1_w", "ln_1.weight"), ("ln1_b", "ln_1.bias"), ("bqkv", "attn.c_attn.bias"), ("bo", "attn.c_proj.bias"),
                         ("ln2_w", "ln_2.weight"), ("ln2_b", "ln_2.bias"), ("b_fc", "mlp.c_fc.bias"), ("b_proj", "mlp.c_proj.bias")):
                put(a, get(p + b))
            for a, b in (("wqkv", "attn.c_attn.weight"), ("wo", "attn.c_proj.weight"), ("w_fc", "mlp.c_fc.weight"), ("w_proj", "mlp.c_proj.weight")):
                put(a, tr(get(p + b)))
    else:
        if unit == "embed":
            put("tok_emb", get("model.embed_tokens.weight"))
        elif unit == "head":
            put("final_norm", get("model.norm.weight"))
            put("lm_head", get("lm_head.weight"))
        elif unit.startswith("layer."):
            p = f"model.layers.{unit.split('.')[1]}."
            qkv = [get(p + f"self_attn.{n}_proj.weight") for n in "qkv"]
            if all(t is not None for t in qkv):
                put("wqkv", torch.cat(qkv, 0))
            gu = [get(p + f"mlp.{n}_proj.weight") for n in ("gate", "up")]
            if all(t is not None for t in gu):
                put("w_gu", torch.cat(gu, 0))
            put("attn_norm", get(p + "input_layernorm.weight"))
            put("wo", get(p + "self_attn.o_proj.weight"))
            put("mlp_norm", get(p + "post_attention_layernorm.weight"))
            put("w_down", get(p + "mlp.down_proj.weight"))
    if unit == "reward_head":
        put("score_w", get("scorer.1.weight"))
        put("score_b", get("scorer.1.bias"))
    if unit == "value_head":
        put("value_w", get("value_head.weight"))
        put("value_b", get("value_head.bias"))
    return out


class ShardedSafetensorsWriter:
    """Streams tensors into ``model-XXXXX-of-YYYYY.safetensors`` files of bounded size and writes the
    HF index; a single shard is renamed to ``model.safetensors``."""

    def __init__(self, out_dir: str | Path, max_shard_bytes: int = 5 << 30) -> None:
        self.dir = Path(out_dir)
        self.dir.mkdir(parents=True, exist_ok=True)
        self.max = max_shard_bytes
        self.cur: Dict[str, torch.Tensor] = {}
        self.cur_bytes = 0
        self.files: List[Tuple[Path, List[str]]] = []

    def add(self, tensors: Dict[str, torch.Tensor]) -> None:
        for k, t in tensors.items():
            t = t.detach().to("cpu").contiguous()
            n = t.numel() * t.element_size()
            if self.cur and self.cur_bytes + n > self.max:
                self._flush()
            self.cur[k] = t
            self.cur_bytes += n

    def _flush(self) -> None:
        from safetensors.torch import save_file

        if not self.cur:
            return
        path = self.dir / f"model-tmp-{len(self.files):05d}.safetensors"
        save_file(self.cur, str(path), metadata={"format": "pt"})
        self.files.append((path, list(self.cur)))
        self.cur, self.cur_bytes = {}, 0

    def close(self) -> None:
        self._flush()
        n = len(self.files)
        if n == 1:
            self.files[0][0].rename(self.dir / "model.safetensors")
            return
        weight_map = {}
        for i, (path, keys) in enumerate(self.files):
            name = f"model-{i + 1:05d}-of-{n:05d}.safetensors"
            path.rename(self.dir / name)
            for k in keys:
                weight_map[k] = name
        with (self.dir / "model.safetensors.index.json").open("w") as fh:
            json.dump({"metadata": {}, "weight_map": weight_map}, fh, indent=2)


class SafetensorsReader:
    """Random access over ``model.safetensors`` or an indexed shard set (and ``pytorch_model.bin``)."""

    def __init__(self, path: str | Path) -> None:
        self.dir = Path(path)
        self.map: Dict[str, Path] = {}
        self._bin: Optional[Dict[str, torch.Tensor]] = None
        idx = self.dir / "model.safetensors.index.json"
        if idx.exists():
            with idx.open() as fh:
                for k, f in json.load(fh)["weight_map"].items():
                    self.map[k] = self.dir / f
        elif (self.dir / "model.safetensors").exists():
            from safetensors import safe_open

            with safe_open(str(self.dir / "model.safetensors"), "pt") as f:
                for k in f.keys():
                    self.map[k] = self.dir / "model.safetensors"
        elif (self.dir / "pytorch_model.bin").exists():
            self._bin = torch.load(self.dir / "pytorch_model.bin", map_location="cpu", weights_only=True)
        self._open: Dict[Path, object] = {}

    def has_weights(self) -> bool:
        return bool(self.map) or self._bin is not None

    def get(self, name: str) -> Optional[torch.Tensor]:
        if self._bin is not None:
            for cand in (name, "backbone." + name, name.replace("model.", "backbone.", 1)):
                if cand in self._bin:
                    return self._bin[cand]
            return None
        f = self.map.get(name)
        if f is None:
            return None
        from safetensors import safe_open

        if f not in self._open:
            self._open[f] = safe_open(str(f), "pt")
        return self._open[f].get_tensor(name)  # type: ignore[attr-defined]


def has_pretrained_weights(path: str | Path) -> bool:
    p = Path(path)
    return p.is_dir() and any((p / f).exists() for f in ("model.safetensors", "model.safetensors.index.json", "pytorch_model.bin"))
