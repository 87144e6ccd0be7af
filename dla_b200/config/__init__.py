"""Config system: YAML -> dict with overlays, dotted overrides, validation and
hardware-block resolution.

Parity with the reference:
  * ``load_config`` reads one YAML file into a plain dict (``R/src/training/utils.py:18-21``).
  * Trainers read ``hardware.{deepspeed_config,fsdp,gradient_accumulation_steps,
    mixed_precision}`` and ``logging.{use_wandb,log_dir}`` (``R/src/training/utils.py:55-75``).
Beyond the reference: ``--overlay`` deep-merges the ablation fragments in
``config/ablations/`` (the reference says "merge manually", ``R/README.md:126``),
``--override a.b=c`` patches single keys, and unknown top-level keys only warn.
"""
from __future__ import annotations

import argparse
import copy
import json
import warnings
from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, Iterable, List, Optional

import yaml

KNOWN_TOP_LEVEL = {
    "experiment_name", "seed", "backend", "model", "data", "optimization", "logging",
    "hardware", "reward_model", "ppo", "sampling", "distill", "models", "benchmarks",
    "latency", "generation", "synthetic",
}


def load_config(path: str | Path) -> Dict[str, Any]:
    """Load a YAML (or JSON) file into a dict. Parity: ``R/src/training/utils.py:18-21``."""
    path = Path(path)
    with path.open("r", encoding="utf-8") as fh:
        if path.suffix == ".json":
            return json.load(fh)
        cfg = yaml.safe_load(fh)
    return cfg or {}


def deep_merge(base: Dict[str, Any], overlay: Dict[str, Any]) -> Dict[str, Any]:
    """Recursively merge ``overlay`` into a copy of ``base`` (overlay wins)."""
    out = copy.deepcopy(base)
    for k, v in (overlay or {}).items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = deep_merge(out[k], v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def _parse_scalar(text: str) -> Any:
    try:
        return yaml.safe_load(text)
    except yaml.YAMLError:
        return text


def apply_overrides(cfg: Dict[str, Any], overrides: Iterable[str]) -> Dict[str, Any]:
    """Apply ``a.b.c=value`` overrides (value parsed as YAML)."""
    out = copy.deepcopy(cfg)
    for item in overrides or []:
        if "=" not in item:
            raise ValueError(f"override must look like key.path=value, got {item!r}")
        key, value = item.split("=", 1)
        cur = out
        parts = key.strip().split(".")
        for p in parts[:-1]:
            nxt = cur.get(p)
            if not isinstance(nxt, dict):
                nxt = {}
                cur[p] = nxt
            cur = nxt
        cur[parts[-1]] = _parse_scalar(value)
    return out


def validate(cfg: Dict[str, Any], phase: Optional[str] = None) -> List[str]:
    """Light validation: returns a list of warnings, raises on hard errors."""
    notes: List[str] = []
    for k in cfg:
        if k not in KNOWN_TOP_LEVEL:
            notes.append(f"unknown top-level config key {k!r} (ignored)")
    required = {
        "sft": [("model", "model_name_or_path"), ("optimization", "micro_batch_size"),
                ("optimization", "learning_rate"), ("optimization", "max_train_steps"),
                ("logging", "output_dir")],
        "reward": [("model", "base_model_name_or_path"), ("optimization", "micro_batch_size"),
                   ("optimization", "learning_rate"), ("optimization", "max_train_steps"),
                   ("logging", "output_dir")],
        "dpo": [("model", "policy_model_name_or_path"), ("model", "reference_model_name_or_path"),
                ("optimization", "micro_batch_size"), ("optimization", "learning_rate"),
                ("optimization", "max_train_steps"), ("logging", "output_dir")],
        "rlhf": [("model", "policy_model_name_or_path"), ("model", "reference_model_name_or_path"),
                 ("ppo",), ("sampling",), ("logging", "output_dir")],
        "distill": [("model", "student_model_name_or_path"), ("data", "teacher_samples_path"),
                    ("optimization", "micro_batch_size"), ("optimization", "learning_rate"),
                    ("optimization", "max_train_steps"), ("logging", "output_dir")],
        "eval": [("models",), ("logging",)],
    }.get(phase or "", [])
    for path in required:
        cur: Any = cfg
        for p in path:
            if not isinstance(cur, dict) or p not in cur:
                raise KeyError(f"config for phase {phase!r} is missing required key {'.'.join(path)}")
            cur = cur[p]
    hw = cfg.get("hardware", {}) or {}
    mp = hw.get("mixed_precision")
    if mp not in (None, "no", "bf16", "fp16", "fp32"):
        raise ValueError(f"hardware.mixed_precision={mp!r} not understood")
    for w in notes:
        warnings.warn(w, stacklevel=2)
    return notes


@dataclass
class HardwarePlan:
    """Resolved view of the ``hardware`` block.

    ``sharding``: ``ddp`` (replicated params, grad all-reduce — the reference's default,
    ``R/config/accelerate_config.yaml:3``), or ``zero3`` / ``fsdp`` (params + grads +
    optimizer state sharded).  ``hardware.deepspeed_config`` and ``hardware.fsdp`` are
    accepted as aliases for the one native sharded engine (``R/src/training/utils.py:58-65``).
    """

    sharding: str = "ddp"
    gradient_accumulation_steps: int = 1
    mixed_precision: Optional[str] = None
    comm: str = "auto"            # auto | nccl | fused
    resident_params: Optional[bool] = None   # keep gathered bf16 params resident (None = by HBM size)
    gradient_clipping: Optional[float] = None
    overlap_comm: bool = True
    min_num_params: int = 0
    num_processes: Optional[int] = None
    optimizer_state_dtype: str = "fp32"
    # spend free HBM on NOT recomputing: "auto" (after the first update, keep the activations of as many trailing layers —
    # and the lm_head logits — as fit with `keep_reserve_gb` to spare), "off" (recompute everything, the reference's
    # behaviour), or an integer number of layers
    keep_activations: Any = "auto"
    keep_reserve_gb: float = 16.0
    raw: Dict[str, Any] = field(default_factory=dict)


def resolve_hardware(cfg: Dict[str, Any], config_dir: Optional[Path] = None) -> HardwarePlan:
    hw = dict(cfg.get("hardware") or {})
    plan = HardwarePlan(raw=hw)
    plan.gradient_accumulation_steps = int(hw.get("gradient_accumulation_steps", 1) or 1)
    plan.mixed_precision = hw.get("mixed_precision")
    plan.comm = hw.get("comm", "auto")
    plan.resident_params = hw.get("resident_params")
    plan.num_processes = hw.get("num_processes")
    plan.optimizer_state_dtype = hw.get("optimizer_state_dtype", "fp32")
    plan.keep_activations = hw.get("keep_activations", "auto")
    plan.keep_reserve_gb = float(hw.get("keep_reserve_gb", 16.0))
    sharding = hw.get("sharding")
    ds = hw.get("deepspeed_config")
    if ds:
        sharding = sharding or "zero3"
        ds_path = Path(ds)
        if not ds_path.exists() and config_dir is not None:
            alt = config_dir.parent / ds if (config_dir.parent / ds).exists() else config_dir / Path(ds).name
            ds_path = alt
        if ds_path.exists():
            with ds_path.open() as fh:
                ds_cfg = json.load(fh)
            zo = ds_cfg.get("zero_optimization", {})
            stage = int(zo.get("stage", 3))
            sharding = hw.get("sharding") or ("zero3" if stage >= 3 else "zero2" if stage == 2 else "zero1" if stage == 1 else "ddp")
            plan.overlap_comm = bool(zo.get("overlap_comm", True))
            gc = ds_cfg.get("gradient_clipping")
            if gc is not None:
                plan.gradient_clipping = float(gc)
            if ds_cfg.get("bf16", {}).get("enabled"):
                plan.mixed_precision = plan.mixed_precision or "bf16"
    fs = hw.get("fsdp")
    if fs:
        strat = fs.get("sharding_strategy", 1) if isinstance(fs, dict) else 1
        # 1 / FULL_SHARD -> zero3 ; 2 / SHARD_GRAD_OP -> zero2 ; 3 / NO_SHARD -> ddp
        table = {1: "zero3", "FULL_SHARD": "zero3", 2: "zero2", "SHARD_GRAD_OP": "zero2", 3: "ddp", "NO_SHARD": "ddp"}
        sharding = hw.get("sharding") or table.get(strat, "zero3")
        if isinstance(fs, dict):
            plan.min_num_params = int(fs.get("min_num_params", 0) or 0)
            if fs.get("offload_params"):
                warnings.warn("hardware.fsdp.offload_params=true ignored: 180 GB HBM per B200, no CPU offload path")
    plan.sharding = {"fsdp": "zero3", None: "ddp"}.get(sharding, sharding)
    if plan.sharding not in ("ddp", "zero1", "zero2", "zero3"):
        raise ValueError(f"hardware.sharding={plan.sharding!r} not understood")
    return plan


def add_config_args(parser: argparse.ArgumentParser) -> argparse.ArgumentParser:
    parser.add_argument("--config", required=True, type=str, help="Path to YAML config")
    parser.add_argument("--overlay", action="append", default=[], help="YAML fragment deep-merged over --config (repeatable)")
    parser.add_argument("--override", action="append", default=[], help="dotted.key=value (repeatable)")
    parser.add_argument("--resume", default=None, help="checkpoint dir or 'latest'")
    return parser


def load_from_args(args: argparse.Namespace, phase: Optional[str] = None) -> Dict[str, Any]:
    cfg = load_config(args.config)
    for ov in getattr(args, "overlay", []) or []:
        cfg = deep_merge(cfg, load_config(ov))
    cfg = apply_overrides(cfg, getattr(args, "override", []) or [])
    validate(cfg, phase)
    cfg["_config_path"] = str(args.config)
    return cfg
