"""Observability: always-on ``metrics.jsonl``, optional wandb, CUDA-event phase timers, NVTX ranges,
throughput / MFU accounting.

Parity: rank 0 logs ``train/loss`` (running mean since the last log), ``eval/loss``, ``eval/acc``,
``train/preference_rate``, ``train/kl``, ``train/reward_mean`` through ``accelerator.log``
(``R/src/training/train_sft.py:156-162`` etc.).  The reference never calls ``init_trackers`` so those
metrics vanish (defect D8); here they always reach ``<log_dir>/metrics.jsonl`` and wandb is optional
(``logging.use_wandb``, ``R/src/training/utils.py:70``).  Beyond the reference: device-timed tokens/s,
MFU against ``MEASURED_PEAKS.json``, grad-norm, lr, peak HBM.
"""
from __future__ import annotations

import contextlib
import json
import os
import time
from pathlib import Path
from typing import Any, Dict, Iterator, List, Optional

import torch


class MetricsLogger:
    def __init__(self, log_dir: Optional[str], is_main: bool = True, use_wandb: bool = False,
                 run_name: Optional[str] = None, config: Optional[Dict[str, Any]] = None) -> None:
        self.is_main = is_main
        self.path: Optional[Path] = None
        self._wandb = None
        if not is_main:
            return
        if log_dir:
            Path(log_dir).mkdir(parents=True, exist_ok=True)
            self.path = Path(log_dir) / "metrics.jsonl"
        if use_wandb:
            try:
                import wandb

                os.environ.setdefault("WANDB_MODE", "offline")
                self._wandb = wandb.init(project=run_name or "dla_b200", dir=log_dir, config=config, reinit=True)
            except Exception as e:  # pragma: no cover - wandb is optional
                print(f"[obs] wandb unavailable ({e}); metrics.jsonl only")

    def log(self, values: Dict[str, Any], step: int) -> None:
        if not self.is_main:
            return
        rec = {"step": int(step), "time": time.time()}
        for k, v in values.items():
            if isinstance(v, torch.Tensor):
                v = v.item()
            rec[k] = v
        if self.path is not None:
            with self.path.open("a", encoding="utf-8") as fh:
                fh.write(json.dumps(rec) + "\n")
        if self._wandb is not None:  # pragma: no cover
            self._wandb.log({k: v for k, v in rec.items() if k not in ("step", "time")}, step=step)

    def close(self) -> None:
        if self._wandb is not None:  # pragma: no cover
            self._wandb.finish()


class PhaseTimer:
    """CUDA-event (or perf_counter on CPU) timers per named phase: ``with timer("fwd"): ...``.
    ``summary()`` synchronises once and returns milliseconds per phase."""

    def __init__(self, device: torch.device | str, enabled: bool = True, nvtx: bool = False) -> None:
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.enabled, self.nvtx = enabled, nvtx and self.cuda
        self._events: Dict[str, List[Any]] = {}
        self._cpu: Dict[str, float] = {}

    @contextlib.contextmanager
    def __call__(self, name: str) -> Iterator[None]:
        if not self.enabled:
            yield
            return
        if self.nvtx:
            torch.cuda.nvtx.range_push(name)
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                self._events.setdefault(name, []).append((a, b))
                if self.nvtx:
                    torch.cuda.nvtx.range_pop()
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self._cpu[name] = self._cpu.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def summary(self, reset: bool = True) -> Dict[str, float]:
        out = dict(self._cpu)
        if self.cuda and self._events:
            torch.cuda.synchronize(self.device)
            for k, evs in self._events.items():
                out[k] = out.get(k, 0.0) + sum(a.elapsed_time(b) for a, b in evs)
        if reset:
            self._events, self._cpu = {}, {}
        return out


def measured_peaks(repo_root: Optional[Path] = None) -> Dict[str, float]:
    """Driver-measured box peaks (roofline denominators); falls back to the profiling guide's numbers."""
    root = repo_root or Path(__file__).resolve().parents[2]
    p = root / "MEASURED_PEAKS.json"
    if p.exists():
        with p.open() as fh:
            d = json.load(fh)
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_tflops": d.get("bf16_tflops", 1590.0),
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", 1400.0), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def mfu(tokens_per_s: float, flops_per_token: float, n_gpus: int = 1, sustained: bool = True) -> float:
    pk = measured_peaks()
    peak = (pk["bf16_tflops_sustained"] if sustained else pk["bf16_tflops"]) * 1e12 * n_gpus
    return tokens_per_s * flops_per_token / peak


def peak_memory_gb(device: torch.device | str) -> float:
    device = torch.device(device)
    if device.type != "cuda":
        return 0.0
    return torch.cuda.max_memory_allocated(device) / (1 << 30)
