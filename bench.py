#!/usr/bin/env python
"""Headline benchmark: DPO tokens/s at Llama-3-8B (BASELINE.json), own arm and reference arm.

    python bench.py --gpus 1 --steps 5 --warmup 3                # this framework
    python bench.py --impl reference --gpus 1 --steps 5 --warmup 3
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one full DPO update on ``micro_batch`` preference pairs per GPU of ``seq_len`` tokens each
side: policy forward on chosen+rejected, frozen-reference forward, DPO loss, backward (activation
checkpointing), global-norm clip and the AdamW update — nothing skipped.  Tokens per step = 2 * seq_len *
micro_batch * n_gpus (weak scaling).  Two measurements:
  * ``value``: CUDA-event time of exactly K steps with the batch resident on the device, max over ranks;
  * ``e2e``  : the same K steps through the public API including, per step, the pinned-host -> device copy
    of that step's batch and a device -> host read of the loss; wall clock bracketed by barrier+synchronize.
The reference arm (``--impl reference``) imports nothing of ``dla_b200``: its rendezvous, barrier, max-over-ranks and
synthetic batches are plain torch inside ``baseline/reference_arm.py`` and its JSON line records that this repo's
kernel library is absent from the process.  ``--recompute on`` runs the own arm with every layer recomputed in
backward (what the reference hard-codes) so the record separates recompute elimination from kernel speed.
Weights are random-init, data synthetic; L2 is not flushed explicitly: every step streams >100 GB of
weights/optimizer state through HBM, far beyond the 126 MB L2 (stated in ``config.l2``).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
os.environ.setdefault("TOKENIZERS_PARALLELISM", "false")


def parse() -> argparse.Namespace:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--model", default=os.environ.get("DLA_BENCH_MODEL", "llama3-8b"))
    ap.add_argument("--seq-len", type=int, default=int(os.environ.get("DLA_BENCH_SEQ", 1024)))
    ap.add_argument("--micro-batch", type=int, default=int(os.environ.get("DLA_BENCH_MB", 4)), help="preference pairs per GPU per step")
    ap.add_argument("--sharding", default=os.environ.get("DLA_BENCH_SHARDING", "zero3"))
    ap.add_argument("--comm", default=os.environ.get("DLA_BENCH_COMM", "auto"))
    ap.add_argument("--state-dtype", default=os.environ.get("DLA_BENCH_STATE", "fp32"))
    ap.add_argument("--resident", default=os.environ.get("DLA_BENCH_RESIDENT", "auto"), choices=["auto", "0", "1"],
                    help="keep gathered parameters resident (auto = by HBM size; 0 = stream units, the 70B mode)")
    ap.add_argument("--recompute", default=os.environ.get("DLA_BENCH_RECOMPUTE", "auto"), choices=["auto", "on"],
                    help="own arm: 'on' recomputes every layer in backward like the reference does (A/B for how much of the "
                         "ratio is recompute elimination); 'auto' keeps activations where free HBM allows")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons every 200 ms while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self) -> None:
        assert self.proc is not None and self.proc.stdout is not None
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


class OwnArm:
    """This framework through its public API (``load_causal_lm`` / ``dpo_step`` / ``Engine``)."""

    impl = "own"

    def __init__(self, args: argparse.Namespace) -> None:
        import torch
        from dla_b200.algo import dpo_step
        from dla_b200.config import HardwarePlan
        from dla_b200.data.synthetic import synthetic_preference_batch
        from dla_b200.models.factory import load_causal_lm
        from dla_b200.ops import native
        from dla_b200.optim import AdamWConfig
        from dla_b200.parallel.pg import init_distributed

        self.torch, self.args = torch, args
        self.ctx = ctx = init_distributed()
        self.world, self.rank, self.device, self.is_main = ctx.world_size, ctx.rank, ctx.device, ctx.is_main
        plan = HardwarePlan(sharding=args.sharding if self.world > 1 else "ddp", comm=args.comm,
                            optimizer_state_dtype=args.state_dtype,
                            resident_params=None if args.resident == "auto" else args.resident == "1")
        self.policy = load_causal_lm(args.model, ctx, plan, gradient_checkpointing=True, seed=0, name="policy")
        self.ref = load_causal_lm(args.model, ctx, plan, gradient_checkpointing=False, trainable=False, seed=0, name="reference")
        self.opt = AdamWConfig(lr=1e-6, weight_decay=0.01, max_grad_norm=1.0)
        self.vocab = self.policy.config.vocab_size
        self.eng = self.policy.engine
        self._dpo_step, self._native, self._synth = dpo_step, native, synthetic_preference_batch
        self.kept = 0
        self.no_recompute = "0" if args.recompute == "on" else os.environ.get("DLA_BENCH_NO_RECOMPUTE", "auto")
        self.note = None

    def barrier(self) -> None:
        self.ctx.barrier()

    def max_over_ranks(self, v: float) -> float:
        return self.ctx.max_over_ranks(v)

    def make_batch(self, mb: int, T: int, seed: int):
        return self._synth(mb, T, self.vocab, seed=seed, pin=True)

    def step(self, batch):
        loss, _ = self._dpo_step(self.policy.model, self.ref.model, batch, beta=0.1, concat=True)
        loss.backward()
        self.eng.finish_backward()
        self.eng.optimizer_step(self.opt, grad_divisor=float(self.world))
        return loss.detach()

    def retry_after_oom(self) -> bool:
        # only reachable on a single GPU when the kept activations leave too little slack: drop a few layers
        # back to recomputation and repeat this warm-up step (parameters were not touched yet)
        if self.world > 1 or self.kept == 0:
            return False
        self.kept = self.eng.set_no_recompute_layers(max(0, self.kept - 4))
        self.eng.zero_grad()
        self.torch.cuda.empty_cache()
        return True

    def after_warmup_step(self, i: int, tokens: int) -> None:
        torch, eng = self.torch, self.eng
        if self.no_recompute == "0":
            return
        if i == 0:
            # after the first step the allocator's steady state is known: spend the free HBM on keeping
            # activations of the last k layers (no recomputation there); nothing is skipped, only not redone
            self.kept = eng.auto_no_recompute(tokens, reserve_gb=float(os.environ.get("DLA_BENCH_RESERVE_GB", 7.5))) \
                if self.no_recompute == "auto" else eng.set_no_recompute_layers(int(self.no_recompute))
            torch.cuda.reset_peak_memory_stats(self.device)
        elif i == 1 and self.kept > 0 and self.world == 1:
            # one step ran with the kept activations: if the measured peak leaves under 3 GiB, give some back
            total = torch.cuda.get_device_properties(self.device).total_memory
            if total - torch.cuda.max_memory_reserved(self.device) < (3 << 30):
                self.kept = eng.set_no_recompute_layers(max(0, self.kept - 2))

    def before_stop(self) -> None:
        self.eng.sync_optimizer()      # the last AdamW runs on a side stream: order the stop event after it

    def launches(self) -> int:
        return int(self._native.launch_count())

    def describe(self) -> dict:
        return {"optimizer": "AdamW fp32 master+moments" if self.args.state_dtype == "fp32" else "AdamW bf16 moments",
                "activation_checkpointing": f"per layer; last {len(self.eng.no_recompute)} of {self.policy.config.num_layers} "
                                            f"layers keep activations (free HBM)",
                "comm": ("fused peer-memory" if getattr(self.eng, "fused", False) or getattr(self.eng, "fused_ddp", False)
                         or getattr(self.eng, "fused_stream", False) else ("nccl" if self.world > 1 else "none")),
                "comm_detail": self._comm_detail()}

    def _comm_detail(self) -> dict:
        """What moved the gradients / parameters between GPUs (the engine's own byte counters over the whole run: a claim
        like ``gpu_launches``, here so the record says more than "N ranks")."""
        try:
            eng = self.eng
            fused = bool(getattr(eng, "fused", False) or getattr(eng, "fused_ddp", False) or getattr(eng, "fused_stream", False))
            arena = getattr(eng, "grad_arena", None)
            return {"gradient_exchange": ("per-unit reduce-scatter kernel inside backward (dla_unit_reduce)" if getattr(eng, "rs_in_backward", False)
                                          else "fused kernels after backward" if fused else "nccl" if self.world > 1 else "none"),
                    "nvswitch_multicast": bool(getattr(arena, "mc_ptr", 0)) if fused else False,
                    "optimizer_steps": int(eng.step_count),
                    "reduce_bytes_per_rank_total": int(eng.stats.get("reduce_bytes", 0)),
                    "param_publish_bytes_per_rank_total": int(eng.stats.get("allgather_bytes", 0))}
        except Exception as e:       # never let a diagnostic break the record
            return {"error": f"{type(e).__name__}: {e}"[:120]}


def main() -> None:
    args = parse()
    import torch

    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device visible to bench.py"}))
        return
    if args.impl == "reference":
        # nothing of dla_b200 is imported on this branch: rendezvous, data and reductions are plain torch
        try:
            from baseline.reference_arm import ReferenceArm

            arm = ReferenceArm(args)
        except Exception as e:  # the install/import genuinely failed on this box
            if int(os.environ.get("RANK", "0")) == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {e}"[:300]}))
            return
    else:
        arm = OwnArm(args)
    world, dev = arm.world, arm.device
    if world != args.gpus and arm.is_main:
        print(f"[bench] note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    K, Wm, T, MB = args.steps, max(args.warmup, 3), args.seq_len, args.micro_batch      # never fewer than 3 warm-up steps
    step_fn = arm.step

    # ---- data: distinct synthetic batches, pinned on the host -----------------------------------
    n_batches = Wm + K
    host = [arm.make_batch(MB, T, 1000 * arm.rank + i) for i in range(max(n_batches, 1))]
    to_dev = lambda b: {s: {k: v.to(dev, non_blocking=True) for k, v in b[s].items()} for s in b}
    h2d_bytes = sum(v.numel() * v.element_size() for s in host[0].values() for v in s.values())

    def sync_all():
        arm.barrier()
        torch.cuda.synchronize(dev)

    # ---- (1) device-timed ------------------------------------------------------------------------
    dev_batches = [to_dev(b) for b in host]
    torch.cuda.synchronize(dev)
    for i in range(Wm):
        try:
            step_fn(dev_batches[i])
        except torch.OutOfMemoryError:
            if not arm.retry_after_oom():
                raise
            step_fn(dev_batches[i])
        arm.after_warmup_step(i, 2 * MB * T)
    sampler = ClockSampler(dev.index or 0)
    sync_all()
    sampler.start()
    l0 = arm.launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        loss = step_fn(dev_batches[Wm + i])
    arm.before_stop()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1)
    l1 = arm.launches()
    clocks = sampler.stop()
    ms = arm.max_over_ranks(ms)
    tokens_per_step = 2 * T * MB * world
    value = tokens_per_step * K / (ms / 1e3)
    last_loss = float(loss.item())

    # ---- (2) end to end through the public API: H2D of each batch + D2H of each loss --------------
    e2e = None
    if not args.no_e2e:
        for i in range(min(Wm, 1)):
            float(step_fn(to_dev(host[i])).item())
        sync_all()
        t0 = time.perf_counter()
        for i in range(K):
            b = to_dev(host[Wm + i])
            float(step_fn(b).item())
        arm.before_stop()
        sync_all()
        dt = arm.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": tokens_per_step * K / dt, "unit": "tokens/s", "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
               "ms_per_step": dt * 1e3 / K}

    if arm.is_main:
        cfg = {"model": args.model, "global_batch": MB * world, "pairs_per_gpu": MB, "seq_len": T,
               "tokens_per_step": tokens_per_step,
               "parallelism": (f"{args.sharding}x{world}" if world > 1 else "single"),
               "l2": "not flushed: each step streams >100 GB of weights/optimizer state (L2 is 126 MB)"}
        cfg.update(arm.describe())
        out = {
            "metric": "DPO tokens/sec at Llama-3-8B (device-timed, max over ranks)" if args.model == "llama3-8b"
                      else f"DPO tokens/sec at {args.model} (device-timed, max over ranks)",
            "value": value, "unit": "tokens/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": arm.impl, "config": cfg,
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(l1 - l0), "loss": last_loss,
            "peak_hbm_gb": torch.cuda.max_memory_allocated(dev) / (1 << 30),
        }
        if arm.impl == "reference":
            out["own_native_so_in_process"] = arm.native_so_in_process()
            out["dla_b200_imported"] = any(m == "dla_b200" or m.startswith("dla_b200.") for m in sys.modules)
        if arm.note:
            out["note"] = arm.note
        print(json.dumps(out))
    arm.barrier()


if __name__ == "__main__":
    main()
