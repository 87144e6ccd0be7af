"""Shared training-loop runtime for the phase entrypoints.

One object replaces what every reference trainer re-implements inline around ``Accelerator``
(``R/src/training/train_sft.py:113-174`` and its four siblings): output dirs, effective-batch report,
``accumulate`` semantics, clip + step + zero_grad, running-loss logging every ``log_every_steps``,
checkpoints every ``save_every_steps`` plus ``final``, and the micro-step counting of defect D2
(``max_train_steps`` counts micro-batches; ``optimizer_steps`` is reported too).  Additions: resume,
keep-last-N, ``latest`` link, device-timed tokens/s + MFU, grad-norm / lr metrics.
"""
from __future__ import annotations

import argparse
import time
from pathlib import Path
from typing import Any, Callable, Dict, Optional

import torch

from dla_b200 import config as C
from dla_b200.ckpt import CheckpointManager
from dla_b200.obs import MetricsLogger, PhaseTimer, mfu, peak_memory_gb
from dla_b200.optim import AdamWConfig, lr_at
from dla_b200.parallel.engine import Engine
from dla_b200.parallel.pg import DistContext, init_distributed
from dla_b200.utils import RunningLoss, prepare_output_dirs, seed_everything


def parse_args(description: str, extra: Optional[Callable[[argparse.ArgumentParser], None]] = None) -> argparse.Namespace:
    parser = argparse.ArgumentParser(description=description)
    C.add_config_args(parser)
    if extra:
        extra(parser)
    return parser.parse_args()


def setup(args: argparse.Namespace, phase: str, default_seed: int = 0):
    cfg = C.load_from_args(args, phase)
    seed_everything(int(cfg.get("seed", default_seed)))
    ctx = init_distributed(device=(cfg.get("hardware") or {}).get("device"))
    plan = C.resolve_hardware(cfg, Path(args.config).resolve().parent)
    ctx.print0(f"Loaded config from {args.config} | world={ctx.world_size} sharding={plan.sharding} "
               f"accum={plan.gradient_accumulation_steps} device={ctx.device}")
    return cfg, ctx, plan


class TrainLoop:
    def __init__(self, cfg: Dict[str, Any], phase: str, ctx: DistContext, plan: C.HardwarePlan, engine: Engine,
                 engines: Dict[str, Engine], tokenizer: Any, opt: AdamWConfig, total_steps: int,
                 log_every: int, save_every: int, default_log_dir: str, scheduler: str = "constant",
                 warmup_steps: int = 0, resume: Optional[str] = None, reward_bin: bool = False,
                 tokens_fn: Optional[Callable[[Any], int]] = None, flops_per_token: float = 0.0) -> None:
        self.cfg, self.phase, self.ctx, self.plan = cfg, phase, ctx, plan
        self.engine, self.engines, self.tokenizer, self.opt = engine, engines, tokenizer, opt
        self.total_steps, self.log_every, self.save_every = int(total_steps), max(1, int(log_every)), int(save_every)
        self.scheduler, self.warmup = scheduler, int(warmup_steps)
        log_cfg = cfg.get("logging", {}) or {}
        self.output_dir = log_cfg.get("output_dir", f"checkpoints/{phase}")
        self.log_dir = log_cfg.get("log_dir", default_log_dir)
        if ctx.is_main:
            prepare_output_dirs(self.output_dir, self.log_dir)
        self.logger = MetricsLogger(self.log_dir, ctx.is_main, bool(log_cfg.get("use_wandb")), run_name=cfg.get("experiment_name"),
                                    config=cfg)
        self.ckpt = CheckpointManager(self.output_dir, ctx, keep_last_n=log_cfg.get("keep_last_n"),
                                      async_write=bool(log_cfg.get("async_checkpoint", False)))
        self.accum = plan.gradient_accumulation_steps
        self.global_step = 0          # micro-batches consumed (the reference's "step", defect D2)
        self.micro_steps = 0          # micro_step() calls; accumulation boundaries count these (PPO runs several per step)
        self.optimizer_steps = 0
        self.epoch = 0
        self.running = RunningLoss()
        self.timer = PhaseTimer(ctx.device, enabled=bool(log_cfg.get("profile_phases", False)),
                                nvtx=bool(log_cfg.get("nvtx", False)))
        self.reward_bin = reward_bin
        self.tokens_fn, self.flops_per_token = tokens_fn, flops_per_token
        # optional torch.profiler window: logging.torch_profiler: {start_step: N, steps: K, dir: path}
        # records K micro-steps starting at global step N on every rank, writes a Chrome trace per rank and, on rank 0,
        # the per-kernel table; outside the window nothing is hooked
        tp = log_cfg.get("torch_profiler") or {}
        self._prof_cfg = {"start": int(tp.get("start_step", 0)), "steps": int(tp.get("steps", 0)),
                          "dir": str(tp.get("dir") or Path(self.log_dir) / "profiler")} if tp else None
        self._prof = None
        self._tok_since, self._t_since = 0, time.perf_counter()
        self._ev0 = None
        self.sampler_state: Dict[str, Any] = {}
        eff = int((cfg.get("optimization", {}) or {}).get("micro_batch_size", 1)) * ctx.world_size * self.accum
        target = (cfg.get("optimization", {}) or {}).get("total_batch_size", eff)
        ctx.print0(f"Effective global batch size: {eff} (target {target})")
        if resume:
            path = self.ckpt.resolve(resume)
            if path is not None:
                st = self.ckpt.load(path, {k: e for k, e in engines.items() if e.trainable})
                self.global_step = int(st.get("global_step", 0))
                self.micro_steps = int(st.get("micro_steps", self.global_step))
                self.optimizer_steps = int(st.get("optimizer_steps", 0))
                self.epoch = int(st.get("epoch", 0))
                self.sampler_state = st.get("sampler", {}) or {}
                ctx.print0(f"Resumed from {path} at step {self.global_step}")

    # ------------------------------------------------------------------
    def current_lr(self) -> float:
        return lr_at(self.opt.lr, self.scheduler, self.global_step, self.warmup, self.total_steps)

    def micro_step(self, loss_fn: Callable[[], torch.Tensor], batch: Any = None) -> torch.Tensor:
        """forward + backward of one micro-batch; optimizer update on accumulation boundaries."""
        self._profiler_window()
        boundary = (self.micro_steps + 1) % self.accum == 0
        self.engine.sync_grads = boundary
        with self.timer("fwd"):
            loss = loss_fn()
        with self.timer("bwd"):
            loss.backward()
            self.engine.finish_backward()
        if boundary:
            with self.timer("optim"):
                lr = self.current_lr()
                self.engine.optimizer_step(self.opt, lr=lr, grad_divisor=float(self.ctx.world_size * self.accum)
                                           if self.engine.sharding != "none" else 1.0)
            self.optimizer_steps += 1
            if self.optimizer_steps == 1:
                self._size_kept_activations()
        self.global_step += 1
        self.micro_steps += 1
        if self.tokens_fn is not None and batch is not None:
            self._tok_since += int(self.tokens_fn(batch))
        return loss.detach()

    def _profiler_window(self) -> None:
        pc = self._prof_cfg
        if not pc or pc["steps"] <= 0:
            return
        if self._prof is None and self.global_step == pc["start"]:
            acts = [torch.profiler.ProfilerActivity.CPU]
            if self.ctx.device.type == "cuda":
                acts.append(torch.profiler.ProfilerActivity.CUDA)
            self._prof = torch.profiler.profile(activities=acts)
            self._prof.__enter__()
        elif self._prof is not None and self.global_step >= pc["start"] + pc["steps"]:
            if self.ctx.device.type == "cuda":
                torch.cuda.synchronize(self.ctx.device)
            self._prof.__exit__(None, None, None)
            out = Path(pc["dir"])
            out.mkdir(parents=True, exist_ok=True)
            self._prof.export_chrome_trace(str(out / f"trace_rank{self.ctx.rank}_step{pc['start']}.json"))
            if self.ctx.is_main:
                key = "cuda_time_total" if self.ctx.device.type == "cuda" else "cpu_time_total"
                (out / f"kernels_step{pc['start']}.txt").write_text(self._prof.key_averages().table(sort_by=key, row_limit=40))
                self.ctx.print0(f"[{self.phase}] torch.profiler window of {pc['steps']} step(s) written to {out}")
            self._prof, self._prof_cfg = None, None

    def _size_kept_activations(self) -> None:
        """``hardware.keep_activations``: once the first update has shown the allocator's steady state, spend the free
        HBM on keeping activations (and the lm_head logits) instead of recomputing them in backward.  Sized for the
        LONGEST batch the config allows (micro_batch x max_seq_length, x2 for preference pairs), so shorter batches only
        leave more slack; ``off`` keeps the reference's recompute-everything behaviour."""
        mode = self.plan.keep_activations
        eng = self.engine
        if mode in ("off", False, None, 0, "0") or self.ctx.device.type != "cuda" or not hasattr(eng, "auto_no_recompute"):
            return
        mc, oc = self.cfg.get("model", {}) or {}, self.cfg.get("optimization", {}) or self.cfg.get("ppo", {}) or {}
        max_len = int(mc.get("max_seq_length") or (self.cfg.get("data", {}) or {}).get("max_seq_length") or 0)
        mb = int(oc.get("micro_batch_size") or oc.get("mini_batch_size") or oc.get("batch_size") or 0)
        if not max_len or not mb:
            return
        tokens = mb * max_len * (2 if self.phase in ("dpo", "reward") else 1)
        if isinstance(mode, int) or (isinstance(mode, str) and mode.isdigit()):
            k = eng.set_no_recompute_layers(int(mode))
        else:
            k = eng.auto_no_recompute(tokens, reserve_gb=self.plan.keep_reserve_gb)
        self.ctx.print0(f"[{self.phase}] keeping the activations of the last {k} layer(s) instead of recomputing them "
                        f"(hardware.keep_activations={mode})")

    def after_step(self, loss: torch.Tensor, extra: Optional[Dict[str, Any]] = None, sampler: Any = None) -> None:
        """Logging / checkpoint cadence of the reference loops."""
        log_now = self.global_step % self.log_every == 0
        # the running mean is accumulated on the device: one host sync per log line instead of the
        # reference's ``loss.item()`` every micro-step (``R/src/training/train_sft.py:153``)
        acc = getattr(self, "_loss_acc", None)
        self._loss_acc = loss.float() if acc is None else acc + loss.float()
        self._loss_n = getattr(self, "_loss_n", 0) + 1
        if log_now:
            self.running.update(float(self._loss_acc.item()) / self._loss_n, self._loss_n)
            self._loss_acc, self._loss_n = None, 0
            vals: Dict[str, Any] = {"train/loss": self.running.average, "train/lr": self.current_lr(),
                                    "train/optimizer_steps": self.optimizer_steps}
            if self.engine.last_grad_norm is not None:
                vals["train/grad_norm"] = float(self.engine.last_grad_norm.float().item())
            if self.tokens_fn is not None:
                if self.ctx.device.type == "cuda":
                    torch.cuda.synchronize(self.ctx.device)
                dt = self.ctx.max_over_ranks(time.perf_counter() - self._t_since)
                tps = self._tok_since * self.ctx.world_size / max(dt, 1e-9)
                vals["perf/tokens_per_s"] = tps
                if self.flops_per_token:
                    vals["perf/mfu"] = mfu(tps, self.flops_per_token, self.ctx.world_size)
                vals["perf/peak_hbm_gb"] = peak_memory_gb(self.ctx.device)
                self._tok_since, self._t_since = 0, time.perf_counter()
            if extra:
                vals.update(extra)
            phases = self.timer.summary()
            vals.update({f"time/{k}_ms": v for k, v in phases.items()})
            self.logger.log(vals, self.global_step)
            self.ctx.print0(f"[{self.phase}] step {self.global_step}/{self.total_steps} loss {vals['train/loss']:.4f} "
                            f"lr {vals['train/lr']:.3g}" + (f" tok/s {vals['perf/tokens_per_s']:.0f}" if 'perf/tokens_per_s' in vals else ""))
            self.running.reset()
        if self.save_every and self.global_step % self.save_every == 0:
            self.save(f"step_{self.global_step}", sampler)

    def trainer_state(self, sampler: Any = None) -> Dict[str, Any]:
        st = {"phase": self.phase, "global_step": self.global_step, "micro_steps": self.micro_steps,
              "optimizer_steps": self.optimizer_steps,
              "epoch": self.epoch, "lr": self.current_lr(), "world_size": self.ctx.world_size}
        if sampler is not None and hasattr(sampler, "state_dict"):
            st["sampler"] = sampler.state_dict()
        return st

    def save(self, tag: str, sampler: Any = None) -> None:
        self.ckpt.save(tag, self.engines, self.trainer_state(sampler), tokenizer=self.tokenizer,
                       primary=self.engine.name, reward_bin=self.reward_bin)

    def done(self) -> bool:
        return self.global_step >= self.total_steps

    def finish(self, message: str, sampler: Any = None) -> None:
        self._profiler_window()      # close a window that reaches the end of training
        if self._prof is not None:
            self._prof_cfg["steps"] = 0
            self._prof.__exit__(None, None, None)
            self._prof = None
        self.save("final", sampler)
        self.ckpt.wait()            # a background writer (logging.async_checkpoint) must finish before the process exits
        self.logger.close()
        self.ctx.print0(message)
