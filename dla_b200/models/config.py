"""Model configurations.

The reference loads any HF id through ``AutoModelForCausalLM`` (``R/src/models/base_model.py:30-34``)
and names Mistral-7B (``R/config/sft_config.yaml:6``), phi-2 (``R/config/distill_config.yaml:7``) and,
via ``BASELINE.json``, Llama-3 8B/70B and GPT-2 small.  There is no hub access here, so architectures
are described by :class:`ModelConfig` — built from a preset name, or read from an HF-style
``config.json`` in a checkpoint directory — and weights are random-initialised unless the directory
also holds safetensors.
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from pathlib import Path
from typing import Any, Dict, Optional


@dataclass
class ModelConfig:
    arch: str = "llama"                 # llama (Llama/Mistral: RMSNorm, RoPE, GQA, SwiGLU) | gpt2 (LayerNorm, GELU, learned pos)
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_layers: int = 32
    num_heads: int = 32
    num_kv_heads: int = 8
    head_dim: Optional[int] = None
    max_position_embeddings: int = 8192
    rope_theta: float = 500000.0
    norm_eps: float = 1e-5
    tie_word_embeddings: bool = False
    rotary_pct: float = 1.0             # partial rotary (phi-2 style) when < 1
    bos_token_id: int = 1
    eos_token_id: int = 2
    pad_token_id: Optional[int] = None
    name: str = "custom"
    initializer_range: float = 0.02

    def __post_init__(self) -> None:
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_heads
        if self.arch == "gpt2":
            self.num_kv_heads = self.num_heads

    @property
    def q_dim(self) -> int:
        return self.num_heads * self.head_dim

    @property
    def kv_dim(self) -> int:
        return self.num_kv_heads * self.head_dim

    @property
    def rot_dim(self) -> int:
        rd = int(self.head_dim * self.rotary_pct)
        return rd - rd % 16 if rd >= 16 else 0

    def num_params(self, include_head: bool = True) -> int:
        H, F, V, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_layers
        if self.arch == "gpt2":
            per = (3 * H * H + 3 * H) + (H * H + H) + (H * F + F) + (F * H + H) + 4 * H
            tot = V * H + self.max_position_embeddings * H + L * per + 2 * H
        else:
            per = (self.q_dim + 2 * self.kv_dim) * H + H * self.q_dim + 2 * F * H + H * F + 2 * H
            tot = V * H + L * per + H
        if include_head and not self.tie_word_embeddings:
            tot += V * H
        return tot

    def flops_per_token(self, seq_len: int, backward: bool = False) -> float:
        """Dense matmul + attention FLOPs per token for one forward (x3 with backward)."""
        H, F, L = self.hidden_size, self.intermediate_size, self.num_layers
        lin = 2 * ((self.q_dim + 2 * self.kv_dim) * H + H * self.q_dim + (2 if self.arch == "llama" else 1) * F * H + H * F)
        attn = 2 * 2 * self.q_dim * seq_len / 2  # causal
        head = 2 * self.vocab_size * H
        f = L * (lin + attn) + head
        return f * (3 if backward else 1)

    # ---- HF config.json interop ----------------------------------------------------
    def to_hf_dict(self) -> Dict[str, Any]:
        if self.arch == "gpt2":
            return {"model_type": "gpt2", "architectures": ["GPT2LMHeadModel"], "vocab_size": self.vocab_size,
                    "n_embd": self.hidden_size, "n_inner": self.intermediate_size, "n_layer": self.num_layers,
                    "n_head": self.num_heads, "n_positions": self.max_position_embeddings,
                    "layer_norm_epsilon": self.norm_eps, "tie_word_embeddings": self.tie_word_embeddings,
                    "bos_token_id": self.bos_token_id, "eos_token_id": self.eos_token_id,
                    "activation_function": "gelu_new", "dla_name": self.name}
        return {"model_type": "llama", "architectures": ["LlamaForCausalLM"], "vocab_size": self.vocab_size,
                "hidden_size": self.hidden_size, "intermediate_size": self.intermediate_size,
                "num_hidden_layers": self.num_layers, "num_attention_heads": self.num_heads,
                "num_key_value_heads": self.num_kv_heads, "head_dim": self.head_dim,
                "max_position_embeddings": self.max_position_embeddings, "rope_theta": self.rope_theta,
                "rms_norm_eps": self.norm_eps, "tie_word_embeddings": self.tie_word_embeddings,
                "bos_token_id": self.bos_token_id, "eos_token_id": self.eos_token_id,
                "hidden_act": "silu", "torch_dtype": "bfloat16", "dla_name": self.name}

    @classmethod
    def from_hf_dict(cls, d: Dict[str, Any]) -> "ModelConfig":
        mt = d.get("model_type", "llama")
        if mt == "gpt2":
            H = d["n_embd"]
            return cls(arch="gpt2", vocab_size=d["vocab_size"], hidden_size=H,
                       intermediate_size=d.get("n_inner") or 4 * H, num_layers=d["n_layer"], num_heads=d["n_head"],
                       num_kv_heads=d["n_head"], max_position_embeddings=d.get("n_positions", 1024),
                       norm_eps=d.get("layer_norm_epsilon", 1e-5), tie_word_embeddings=d.get("tie_word_embeddings", True),
                       bos_token_id=d.get("bos_token_id", 50256), eos_token_id=d.get("eos_token_id", 50256),
                       name=d.get("dla_name", "gpt2"))
        if mt in ("llama", "mistral", "qwen2"):
            return cls(arch="llama", vocab_size=d["vocab_size"], hidden_size=d["hidden_size"],
                       intermediate_size=d["intermediate_size"], num_layers=d["num_hidden_layers"],
                       num_heads=d["num_attention_heads"],
                       num_kv_heads=d.get("num_key_value_heads", d["num_attention_heads"]),
                       head_dim=d.get("head_dim"), max_position_embeddings=d.get("max_position_embeddings", 8192),
                       rope_theta=d.get("rope_theta", 10000.0), norm_eps=d.get("rms_norm_eps", 1e-5),
                       tie_word_embeddings=d.get("tie_word_embeddings", False),
                       bos_token_id=d.get("bos_token_id", 1), eos_token_id=d.get("eos_token_id", 2),
                       name=d.get("dla_name", mt))
        raise ValueError(f"unsupported model_type {mt!r} (supported: llama, mistral, qwen2-style, gpt2)")

    def save(self, path: str | Path) -> None:
        path = Path(path)
        path.mkdir(parents=True, exist_ok=True)
        with (path / "config.json").open("w") as fh:
            json.dump(self.to_hf_dict(), fh, indent=2)
        with (path / "generation_config.json").open("w") as fh:
            json.dump({"bos_token_id": self.bos_token_id, "eos_token_id": self.eos_token_id}, fh, indent=2)


PRESETS: Dict[str, ModelConfig] = {
    "llama3-8b": ModelConfig("llama", 128256, 4096, 14336, 32, 32, 8, 128, 8192, 500000.0, 1e-5, False,
                             bos_token_id=128000, eos_token_id=128001, name="llama3-8b"),
    "llama3-70b": ModelConfig("llama", 128256, 8192, 28672, 80, 64, 8, 128, 8192, 500000.0, 1e-5, False,
                              bos_token_id=128000, eos_token_id=128001, name="llama3-70b"),
    "mistral-7b": ModelConfig("llama", 32000, 4096, 14336, 32, 32, 8, 128, 32768, 10000.0, 1e-5, False,
                              name="mistral-7b"),
    # phi-2-sized student on the llama block (the reference's distill student is microsoft/phi-2)
    "phi2-like": ModelConfig("llama", 51200, 2560, 10240, 32, 32, 32, 80, 2048, 10000.0, 1e-5, False,
                             rotary_pct=0.4, name="phi2-like"),
    "gpt2-small": ModelConfig("gpt2", 50257, 768, 3072, 12, 12, 12, 64, 1024, 0.0, 1e-5, True,
                              bos_token_id=50256, eos_token_id=50256, name="gpt2-small"),
    "llama-1b": ModelConfig("llama", 32000, 2048, 5632, 16, 16, 4, 128, 4096, 500000.0, 1e-5, False, name="llama-1b"),
    "tiny-llama": ModelConfig("llama", 512, 256, 512, 2, 2, 1, 128, 512, 10000.0, 1e-5, False, name="tiny-llama"),
    "tiny-gpt2": ModelConfig("gpt2", 512, 128, 512, 2, 2, 2, 64, 256, 0.0, 1e-5, True, name="tiny-gpt2"),
}

_ALIASES = {
    "meta-llama/Meta-Llama-3-8B": "llama3-8b", "meta-llama/Llama-3-8B": "llama3-8b",
    "meta-llama/Meta-Llama-3-70B": "llama3-70b", "mistralai/Mistral-7B-v0.1": "mistral-7b",
    "microsoft/phi-2": "phi2-like", "gpt2": "gpt2-small", "openai-community/gpt2": "gpt2-small",
}


def resolve_model_config(name_or_path: str) -> ModelConfig:
    """Preset name / known hub id -> preset; directory with ``config.json`` -> that config."""
    p = Path(name_or_path)
    if p.is_dir() and (p / "config.json").exists():
        with (p / "config.json").open() as fh:
            return ModelConfig.from_hf_dict(json.load(fh))
    if p.is_symlink() or p.exists():
        q = p.resolve()
        if q.is_dir() and (q / "config.json").exists():
            with (q / "config.json").open() as fh:
                return ModelConfig.from_hf_dict(json.load(fh))
    key = _ALIASES.get(name_or_path, name_or_path)
    if key in PRESETS:
        import copy
        return copy.deepcopy(PRESETS[key])
    raise FileNotFoundError(
        f"model {name_or_path!r}: not a checkpoint directory with config.json and not a preset "
        f"({', '.join(sorted(PRESETS))}); hub downloads are unavailable offline")
