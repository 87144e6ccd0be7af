"""Parameter engine: replicated (DDP) and sharded (ZeRO-1/2/3 = FSDP FULL_SHARD) execution of a
unit-structured model, with the sharded AdamW, gradient accumulation and global-norm clipping.

What it replaces in the reference (all delegated there to accelerate / DeepSpeed / FSDP):
  * ``accelerator.prepare(model, optimizer)`` — DDP wrap, FSDP shard or ``deepspeed.initialize``
    (``R/src/training/train_dpo.py:79-82``; selected by ``R/src/training/utils.py:55-75``)
  * ``accelerator.backward`` — bucketed grad all-reduce / per-unit reduce-scatter
  * parameter all-gather before each unit's forward and backward
  * ``accelerator.clip_grad_norm_`` + ``optimizer.step`` + ``zero_grad``
  * ``accelerator.accumulate`` — collectives and the update only on accumulation boundaries

B200-first choices: one flat bf16 buffer per unit (a decoder layer = 436 MB for Llama-3-8B) is the
communication granule; with 180 GB of HBM the gathered bf16 parameters of an 8B model stay RESIDENT
(16 GB) so only gradients (reduce-scatter) and updated shards (all-gather) cross NVLink once per
optimizer step; models that do not fit (70B) stream units with one-unit-ahead prefetch.  fp32 master
weights and Adam moments exist only for the rank's shard.  On CUDA the update runs the fused
``dla_adamw`` kernel with the clip coefficient read from device memory (no host sync); with
``comm="fused"`` the reduce-scatter itself happens inside the kernel over NVLink peer memory
(:mod:`dla_b200.parallel.fused`).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

from dla_b200.models.config import ModelConfig
from dla_b200.models.transformer import (ParamProvider, UnitSpec, Weights, collect_autograd_grads, init_unit_,
                                         unit_views)
from dla_b200.ops import native as N
from dla_b200.ops import reference as R
from dla_b200.parallel.pg import DistContext

SHARDED = ("zero1", "zero2", "zero3")


@dataclass
class AdamWConfig:
    lr: float = 1e-5
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.0
    max_grad_norm: float = 1.0


class Engine(ParamProvider):
    def __init__(self, cfg: ModelConfig, units: Sequence[UnitSpec], ctx: DistContext, sharding: str = "ddp",
                 trainable: bool = True, checkpoint: bool = True, dtype: Optional[torch.dtype] = None, seed: int = 0,
                 init: bool = True, resident: Optional[bool] = None, grad_dtype: Optional[torch.dtype] = None,
                 state_dtype: torch.dtype = torch.float32, comm: str = "auto", name: str = "model",
                 hbm_budget_bytes: Optional[int] = None) -> None:
        self.cfg, self.units, self.ctx, self.name = cfg, list(units), ctx, name
        self.sharding = sharding if ctx.world_size > 1 else "ddp"
        self.trainable, self.checkpoint = trainable, checkpoint
        self.device = ctx.device
        self.dtype = dtype or (torch.bfloat16 if self.device.type == "cuda" else torch.float32)
        self.grad_dtype = grad_dtype or self.dtype
        self.state_dtype = state_dtype
        if state_dtype not in (torch.float32, torch.bfloat16):
            raise ValueError(f"optimizer_state_dtype must be fp32 or bf16, got {state_dtype}")
        self._moment_flag = 2 if state_dtype == torch.bfloat16 else 0      # DLA_ADAMW_MOMENTS_BF16
        self.W = ctx.world_size if self.sharding in SHARDED else 1     # shard group size
        self.r = ctx.rank if self.W > 1 else 0
        self.comm = comm
        self._by_name = {u.name: u for u in self.units}
        self._order = [u.name for u in self.units]
        self.P = {u.name: u.padded_numel(8 * self.W) for u in self.units}
        self.S = {k: v // self.W for k, v in self.P.items()}
        total = sum(self.P.values())
        if resident is None:
            # keep gathered params (+ full grad buckets when trainable) resident if they fit comfortably
            need = total * _bytes(self.dtype) * (2 if trainable else 1)
            budget = hbm_budget_bytes if hbm_budget_bytes is not None else _default_budget(self.device)
            resident = self.sharding != "zero3" or need <= budget
        self.resident = bool(resident) or self.W == 1
        # fused peer-memory collectives: sharded + resident + trainable on NVLink-connected CUDA ranks
        self.fused = False
        if comm in ("auto", "fused") and trainable and self.W > 1 and self.resident:
            from dla_b200.parallel.fused import FusedComm, fused_available

            if fused_available(ctx):
                self.fused = True
                self.fc = FusedComm.get(ctx)
            elif comm == "fused":
                raise RuntimeError("hardware.comm=fused needs >= 2 CUDA ranks on NCCL with the native library built")
        # replicated (DDP) mode over peer memory: gradient buckets live in a symmetric arena and are all-reduced
        # by one reduce-and-push kernel instead of bucketed NCCL all_reduce calls
        self.fused_ddp = False
        if comm in ("auto", "fused") and trainable and self.sharding == "ddp" and ctx.world_size > 1:
            from dla_b200.parallel.fused import FusedComm, fused_available

            if fused_available(ctx):
                self.fused_ddp = True
                self.fc = FusedComm.get(ctx)
        # streamed units (ZeRO-3 that does not fit resident): parameter all-gather fused into the consuming GEMMs
        self.fused_stream = False
        if comm in ("auto", "fused") and self.W > 1 and not self.resident:
            from dla_b200.parallel.fused import FusedComm, fused_available

            if fused_available(ctx):
                self.fused_stream = True
                self.fc = FusedComm.get(ctx)
        self.gred: Dict[str, torch.Tensor] = {}
        self.base: Dict[str, int] = {}
        self.full: Dict[str, Optional[torch.Tensor]] = {}
        self.shard: Dict[str, torch.Tensor] = {}
        self.gradbuf: Dict[str, Optional[torch.Tensor]] = {}
        self.gshard: Dict[str, torch.Tensor] = {}
        self.master: Dict[str, torch.Tensor] = {}
        self.m: Dict[str, torch.Tensor] = {}
        self.v: Dict[str, torch.Tensor] = {}
        self._views: Dict[str, Weights] = {}
        self._gather_work: Dict[str, Any] = {}
        self._reduce_work: List[Any] = []
        self._passes_expected = 1
        self._passes_seen: Dict[str, int] = {}
        self._plain_used: set = set()
        # optimizer / next-forward overlap: the update runs on a side stream; a unit's first acquire waits for its event
        import os as _os
        self.overlap_optimizer = _os.environ.get("DLA_OVERLAP_OPT", "1") != "0"
        self._opt_stream: Optional["torch.cuda.Stream"] = None
        self._unit_ready: Dict[str, "torch.cuda.Event"] = {}
        self._slots: Dict[str, Weights] = {}
        self.lazy_zero = _os.environ.get("DLA_LAZY_ZERO", "1") != "0"
        self._all_ready: Optional["torch.cuda.Event"] = None
        self.sync_grads = True          # False on non-boundary accumulation micro-steps (``no_sync``)
        # fused path (a) v2: each unit's reduce-scatter is launched on the side stream DURING backward, as soon as the
        # unit's bucket is final, gated by per-unit flags instead of a global barrier (DLA_RS_IN_BACKWARD=0: legacy)
        self.rs_in_backward = (self.fused and self.overlap_optimizer and _os.environ.get("DLA_RS_IN_BACKWARD", "1") != "0")
        self._rs_epoch = 0
        self._rs_launched: set = set()
        # one GPU: a unit's share of the squared gradient norm is summed on the side stream as soon as its bucket is
        # final, under the backward of the layers below it, instead of one 16 GB read after backward (DLA_NORM_IN_BACKWARD=0)
        self.norm_in_backward = (self.ctx.world_size == 1 and self.device.type == "cuda" and self.overlap_optimizer
                                 and _os.environ.get("DLA_NORM_IN_BACKWARD", "1") != "0")
        self._sq_launched: set = set()
        self._sq_stale = False
        self._train_ranges: Optional[Dict[str, List[tuple]]] = None     # set by freeze_except
        self.step_count = 0
        self._sumsq = None
        self._coef = None
        self._last_grad_norm: Optional[torch.Tensor] = None
        self.stats = {"allgather_bytes": 0, "reduce_bytes": 0}
        self._build(seed, init)

    # ------------------------------------------------------------------ construction
    def _build(self, seed: int, init: bool) -> None:
        if self.fused:
            # parameters and gradient buckets of ALL units live in two symmetric arenas so that peers can
            # read gradient shards and write updated parameter shards directly over NVLink
            from dla_b200.parallel.fused import SymmetricArena

            off = 0
            for u in self.units:
                self.base[u.name] = off
                off += self.P[u.name]
            self.param_arena = SymmetricArena(self.ctx, off, self.dtype)
            self.grad_arena = SymmetricArena(self.ctx, off, self.grad_dtype)
            self._flag_base = self.fc.claim_unit_rows(len(self.units))
            self._unit_index = {u.name: i for i, u in enumerate(self.units)}
        if self.fused_ddp:
            from dla_b200.parallel.fused import SymmetricArena

            off = 0
            for u in self.units:
                self.base[u.name] = off
                off += self.P[u.name]
            self._ddp_total = -(-off // (8 * self.ctx.world_size)) * (8 * self.ctx.world_size)
            self.grad_arena = SymmetricArena(self.ctx, self._ddp_total, self.grad_dtype)
        if self.fused_stream:
            from dla_b200.parallel.fused import SymmetricArena

            off, self.sbase = 0, {}
            for u in self.units:
                self.sbase[u.name] = off
                off += self.S[u.name]
            self.shard_arena = SymmetricArena(self.ctx, off, self.dtype)
            self._ag_flags = torch.zeros(8192, dtype=torch.int32, device=self.device)
            self._ag_epoch = 0
            esz = _bytes(self.dtype)
            self._unit_ptrs = {u.name: torch.tensor([ptr + self.sbase[u.name] * esz for ptr in self.shard_arena.ptrs],
                                                    dtype=torch.int64, device=self.device) for u in self.units}
            import os as _os2
            self.stream_rs = self.trainable and _os2.environ.get("DLA_STREAM_RS", "1") != "0"
            if self.stream_rs:
                # gradient buckets of streamed units live in a small symmetric RING (peers sum them in place, in the
                # switch): a bucket exists from the unit's backward until every peer has read its shard of it
                self._ring_k = max(2, int(_os2.environ.get("DLA_STREAM_RING", "3")))
                self._ring_p = max(self.P.values())
                self.grad_ring = SymmetricArena(self.ctx, self._ring_k * self._ring_p, self.grad_dtype, zero=False)
                self._ring_seq = 0
                self._ring_prev: List[Optional[tuple]] = [None] * self._ring_k
                self._ring_of: Dict[str, int] = {}
                self._sflag_base = self.fc.claim_unit_rows(2 * len(self.units))     # rows: 2u = "bucket ready", 2u+1 = "shard read"
                self._unit_index = {u.name: i for i, u in enumerate(self.units)}
                self._srs_epoch = {u.name: 0 for u in self.units}
                self._rs_stream = torch.cuda.Stream(device=self.device)
        for ui, u in enumerate(self.units):
            P, S = self.P[u.name], self.S[u.name]
            if self.fused:
                b = self.base[u.name]
                full = self.param_arena.tensor[b:b + P]
            else:
                full = torch.zeros(P, dtype=self.dtype, device=self.device)
            if init:
                gen = torch.Generator(device=self.device).manual_seed(seed * 1000003 + ui)
                init_unit_(full, u, self.cfg, gen)
            if self.W > 1:
                if self.fused_stream:
                    sb = self.sbase[u.name]
                    self.shard[u.name] = self.shard_arena.tensor[sb:sb + S]
                    self.shard[u.name].copy_(full[self.r * S:(self.r + 1) * S])
                elif self.fused:
                    # the rank's shard IS its slice of the resident copy: the multicast store of the update reaches it too
                    self.shard[u.name] = full[self.r * S:(self.r + 1) * S]
                else:
                    self.shard[u.name] = full[self.r * S:(self.r + 1) * S].clone()
                self.full[u.name] = full if self.resident else None
            else:
                self.shard[u.name] = full            # the rank owns the whole unit
                self.full[u.name] = full
            if self.trainable:
                if self.fused_ddp:
                    self.gradbuf[u.name] = self.grad_arena.tensor[self.base[u.name]:self.base[u.name] + P]
                elif self.fused:
                    self.gradbuf[u.name] = self.grad_arena.tensor[self.base[u.name]:self.base[u.name] + P]
                    self.gred[u.name] = torch.zeros(S, dtype=torch.float32, device=self.device)
                else:
                    self.gradbuf[u.name] = torch.zeros(P, dtype=self.grad_dtype, device=self.device) if self.resident else None
                if self.W > 1 and not self.resident:
                    self.gshard[u.name] = torch.zeros(S, dtype=torch.float32, device=self.device)
            del full
        if self.trainable:
            self._init_optimizer_state()
        if self.fused_stream:
            torch.cuda.synchronize(self.device)
            self.ctx.barrier()          # every rank's shards are in place before anyone pulls them

    def _init_optimizer_state(self) -> None:
        for u in self.units:
            sh = self.shard[u.name]
            self.master[u.name] = sh.float().clone()
            self.m[u.name] = torch.zeros(sh.numel(), dtype=self.state_dtype, device=self.device)
            self.v[u.name] = torch.zeros(sh.numel(), dtype=self.state_dtype, device=self.device)
        if self.device.type == "cuda":
            self._sumsq = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._coef = torch.ones(1, dtype=torch.float32, device=self.device)
            self._norm = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._sq_scratch = torch.empty(N.sumsq_blocks(), dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------ ParamProvider
    def note_plain(self, unit: str) -> None:
        self._plain_used.add(unit)

    def set_no_recompute_layers(self, k: int) -> int:
        """Keep the activations of the LAST ``k`` decoder layers instead of recomputing them in backward
        (they are the first to be consumed, so their memory is released earliest).  Returns k."""
        layers = [u.name for u in self.units if u.name.startswith("layer.")]
        k = max(0, min(int(k), len(layers)))
        self.no_recompute = frozenset(layers[len(layers) - k:]) if k else frozenset()
        return k

    def auto_no_recompute(self, tokens: int, reserve_gb: float = 10.0) -> int:
        """Spend free HBM on skipping recomputation: per decoder layer a plain backward keeps roughly
        ``tokens * (8 H + 2 qkv + 6 F)`` bytes of bf16 activations alive.  Call after a warm-up step."""
        if self.device.type != "cuda" or not self.resident:
            return 0
        cfg = self.cfg
        per_layer = tokens * (8 * cfg.hidden_size + 2 * (cfg.q_dim + 2 * cfg.kv_dim) + 2 * cfg.q_dim + 6 * cfg.intermediate_size)
        torch.cuda.synchronize(self.device)
        free, _ = torch.cuda.mem_get_info(self.device)
        free += torch.cuda.memory_reserved(self.device) - torch.cuda.memory_allocated(self.device)
        budget = max(0, free - reserve_gb * (1 << 30))
        k = int(budget // max(per_layer, 1))
        # the lm_head's bf16 logits (2 bytes x tokens x vocab) are kept only from what is left once EVERY layer keeps its
        # activations: measured at N=1 (Llama-3-8B), giving up a layer (2.6 GB, 2.7 ms of recomputation) for the logits
        # (2.1 GB, one GEMM of the backward) was a net loss of 3 ms; with sharded state (N >= 2) there is room for both
        from dla_b200.ops import functional as _F

        n_layers = sum(1 for u in self.units if u.name.startswith("layer."))
        logits_bytes = tokens * (-(-cfg.vocab_size // 8) * 8) * 2
        keep_logits = self.trainable and k >= n_layers and budget - n_layers * per_layer >= logits_bytes
        if self.ctx.world_size > 1:
            t = torch.tensor([int(keep_logits)], device=self.device)
            dist.all_reduce(t, dist.ReduceOp.MIN, group=self.ctx.group)
            keep_logits = bool(t.item())
        if "DLA_KEEP_LOGITS" not in __import__("os").environ:
            _F.KEEP_LOGITS[0] = keep_logits
        if self.ctx.world_size > 1:     # every rank must take the same decision
            t = torch.tensor([k], device=self.device)
            dist.all_reduce(t, dist.ReduceOp.MIN, group=self.ctx.group)
            k = int(t.item())
        return self.set_no_recompute_layers(k)

    def expect_passes(self, k: int) -> None:
        """How many backward passes traverse each unit per micro-step (2 when chosen/rejected run as
        separate forwards like ``R/src/training/train_dpo.py:113-114``)."""
        self._passes_expected = max(1, int(k))

    def _gather_async(self, unit: str) -> None:
        if self.full.get(unit) is not None or unit in self._gather_work:
            return
        full = torch.empty(self.P[unit], dtype=self.dtype, device=self.device)
        work = dist.all_gather_into_tensor(full, self.shard[unit], group=self.ctx.group, async_op=True)
        self.stats["allgather_bytes"] += full.numel() * full.element_size()
        self._gather_work[unit] = (work, full)

    def _neighbour(self, unit: str, backward: bool) -> Optional[str]:
        i = self._order.index(unit) + (-1 if backward else 1)
        return self._order[i] if 0 <= i < len(self._order) else None

    @property
    def last_grad_norm(self) -> Optional[torch.Tensor]:
        """Global gradient norm of the last optimizer step (device scalar).  With the fused path the value is produced
        on the optimizer side stream, so the reader's stream is ordered after it first."""
        if self.device.type == "cuda":
            self.sync_optimizer()
        return self._last_grad_norm

    def sync_optimizer(self) -> None:
        """Order the current stream after every pending side-stream optimizer update (no host wait)."""
        if not self._unit_ready and self._all_ready is None:
            return
        cur = torch.cuda.current_stream(self.device)
        for e in self._unit_ready.values():
            cur.wait_event(e)
        self._unit_ready.clear()
        if self._all_ready is not None:
            cur.wait_event(self._all_ready)
            self._all_ready = None

    def _wait_unit(self, unit: str) -> None:
        cur = torch.cuda.current_stream(self.device)
        e = self._unit_ready.pop(unit, None)
        if e is not None:
            cur.wait_event(e)
        if self._all_ready is not None:
            cur.wait_event(self._all_ready)
            self._all_ready = None

    def acquire(self, unit: str, backward: bool = False) -> Weights:
        if self._unit_ready or self._all_ready is not None:
            self._wait_unit(unit)          # the side-stream AdamW of the previous step may still be writing this unit
        if unit in self._views:
            return self._views[unit]
        if self.full.get(unit) is None and self.fused_stream:
            return self._acquire_fused_stream(unit, backward)
        if self.full.get(unit) is None:
            self._gather_async(unit)
            work, full = self._gather_work.pop(unit)
            work.wait()
            self.full[unit] = full
            nxt = self._neighbour(unit, backward)
            if nxt is not None:                      # one-unit-ahead prefetch overlaps this unit's math
                self._gather_async(nxt)
        spec = self._by_name[unit]
        w = unit_views(self.full[unit], spec)
        self._attach_grad_slots(unit, w, backward)
        self._views[unit] = w
        return w

    def _attach_grad_slots(self, unit: str, w: Weights, backward: bool) -> None:
        """Point every parameter view at its slice of the unit's gradient bucket.  Streamed units get a bucket only
        when a backward can follow (a checkpointed forward runs under no_grad and needs none); the bucket is
        freed again once its reduce-scatter has been folded into the fp32 shard (``_drain``), so at most a few
        unit-sized buckets are alive — not one per layer (141 GB at 70B)."""
        if not self.trainable:
            return
        gb = self.gradbuf.get(unit)
        if gb is None:
            if not (backward or torch.is_grad_enabled()):
                return
            if self.fused_stream and getattr(self, "stream_rs", False) and backward:
                gb = self._ring_bucket(unit)
            else:
                gb = torch.zeros(self.P[unit], dtype=self.grad_dtype, device=self.device)
            self.gradbuf[unit] = gb
        g = unit_views(gb, self._by_name[unit])
        if self.resident:
            self._slots[unit] = g          # persistent bucket: zero_grad may mark GEMM-fed slots fresh instead of clearing
        for k, t in w.items():
            t.requires_grad_(True)
            t._grad_slot = g[k]

    def _ring_bucket(self, unit: str) -> torch.Tensor:
        """Next slot of the symmetric gradient-bucket ring for ``unit`` (same sequence on every rank).  The slot's
        previous occupant was reduced by kernels that read it from every peer: wait (on this stream) until all ranks
        have published "shard read" for it, then clear the slot."""
        slot = self._ring_seq % self._ring_k
        self._ring_seq += 1
        prev = self._ring_prev[slot]
        if prev is not None:
            self.fc.flag_wait(self._sflag_base + 16 * (2 * prev[0] + 1), prev[1])
            self._ring_prev[slot] = None
        gb = self.grad_ring.tensor[slot * self._ring_p: slot * self._ring_p + self.P[unit]]
        gb.zero_()
        self._ring_of[unit] = slot
        return gb

    def _launch_ring_reduce(self, unit: str) -> None:
        """Streamed unit, fused path (a): the unit's bucket (a ring slot) is final on this rank.  On the side stream,
        behind that point: one kernel announces it, waits for the peers' announcements, sums this rank's shard out of
        all ranks' slots (in the switch when the ring has a multicast address) INTO the fp32 gradient shard, then a
        second tiny kernel tells the peers that this rank is done reading their slots."""
        slot = self._ring_of.pop(unit)
        idx, S = self._unit_index[unit], self.S[unit]
        self._srs_epoch[unit] += 1
        ep = self._srs_epoch[unit]
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._rs_stream.wait_event(ev)
        with torch.cuda.stream(self._rs_stream):
            self.fc.reduce_unit(self.grad_ring, self._sflag_base + 16 * (2 * idx), ep, slot * self._ring_p + self.r * S, S, 1.0,
                                self.gshard[unit], None, accumulate=True)
            self.fc.flag_publish(self._sflag_base + 16 * (2 * idx + 1), ep)
        self._ring_prev[slot] = (idx, ep)
        self.stats["reduce_bytes"] += S * _bytes(self.grad_dtype) * (1 if self.grad_ring.mc_ptr else self.W - 1)
        self.gradbuf[unit] = None
        self._views.pop(unit, None)

    def _acquire_fused_stream(self, unit: str, backward: bool = False) -> Weights:
        """Streamed unit, fused path (b): big 2-D weights are NOT gathered here — each carries an ``_ag`` ticket
        and the first GEMM that consumes it pulls its rows from the owners inside the kernel; only the small
        parameters (norms, biases) and non-GEMM tables are gathered up front by a plain peer-load kernel."""
        spec = self._by_name[unit]
        P, S = self.P[unit], self.S[unit]
        full = torch.empty(P, dtype=self.dtype, device=self.device)
        offs = spec.offsets()
        ptrs = self._unit_ptrs[unit]
        w = unit_views(full, spec)
        order = [p.name for p in spec.params]
        for i, prm in enumerate(spec.params):
            o, n = offs[prm.name]
            end = offs[order[i + 1]][0] if i + 1 < len(order) else P
            gemm_weight = (unit.startswith("layer.") and len(prm.shape) == 2 and prm.shape[1] % 8 == 0
                           and prm.numel >= 256 * 64)
            if gemm_weight:
                w[prm.name]._ag = _AGTicket(self, unit, full, ptrs, S, o, prm.shape[0])
            else:
                self.fc.gather_range(ptrs, full, S, o, end - o)
                self.stats["allgather_bytes"] += (end - o) * _bytes(self.dtype)
        self.full[unit] = full
        self._attach_grad_slots(unit, w, backward)
        self._views[unit] = w
        return w

    def release(self, unit: str, backward: bool = False) -> None:
        if self.resident:
            return
        if self.trainable and not backward and torch.is_grad_enabled():
            return                                  # non-checkpointed training keeps the unit for backward
        self._views.pop(unit, None)
        self.full[unit] = None

    def grads_done(self, unit: str, weights: Weights) -> None:
        collect_autograd_grads(weights)
        if not self.trainable:
            return
        seen = self._passes_seen.get(unit, 0) + 1
        self._passes_seen[unit] = seen
        if seen < self._passes_expected:
            return
        self._passes_seen[unit] = 0
        if self.ctx.world_size == 1:
            if self.norm_in_backward and self.sync_grads:
                self._launch_unit_sumsq(unit)
            return
        self._materialize_fresh(unit)
        if self.resident:
            if self.sync_grads:
                self._launch_reduce(unit)
        else:
            self._launch_reduce(unit)               # streamed units reduce-scatter every micro-step

    def unit_backward_done(self, unit: str) -> None:
        self.grads_done(unit, self._views.get(unit, {}))

    def _launch_reduce(self, unit: str) -> None:
        if self.fused and self.rs_in_backward:
            return self._launch_unit_reduce(unit)
        if self.fused or self.fused_ddp:
            return          # peers pull this bucket themselves inside the fused kernels
        if self.fused_stream and unit in getattr(self, "_ring_of", {}):
            return self._launch_ring_reduce(unit)
        gb = self.gradbuf[unit]
        nbytes = gb.numel() * gb.element_size()
        self.stats["reduce_bytes"] += nbytes
        if self.sharding == "ddp":
            work = dist.all_reduce(gb, group=self.ctx.group, async_op=True)
            self._reduce_work.append((work, unit, None))
        else:
            out = torch.empty(self.S[unit], dtype=gb.dtype, device=self.device)
            work = dist.reduce_scatter_tensor(out, gb, group=self.ctx.group, async_op=True)
            self._reduce_work.append((work, unit, out))
            if not self.resident:
                self._drain(keep=2)                 # bound live full-size grad buckets when streaming units

    def _launch_unit_reduce(self, unit: str) -> None:
        """Fused path (a) v2: this rank's bucket for ``unit`` is final (everything enqueued so far on the current
        stream).  Queue the unit's reduce-scatter kernel on the side stream behind that point; the kernel itself
        announces the bucket to the peers and waits for theirs, so earlier layers' backward keeps running meanwhile."""
        if unit in self._rs_launched:
            return
        self._zero_frozen(unit)
        if self._opt_stream is None:
            self._opt_stream = torch.cuda.Stream(device=self.device)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._opt_stream.wait_event(ev)
        S = self.S[unit]
        with torch.cuda.stream(self._opt_stream):
            self.fc.reduce_unit(self.grad_arena, self._flag_base + 16 * self._unit_index[unit], self._rs_epoch + 1,
                                self.base[unit] + self.r * S, S, 1.0, self.gred[unit], self._sumsq)
        self.stats["reduce_bytes"] += S * _bytes(self.grad_dtype) * (1 if self.grad_arena.mc_ptr else self.W - 1)
        self._rs_launched.add(unit)

    def _launch_unit_sumsq(self, unit: str) -> None:
        """One GPU: this unit's bucket is final (everything enqueued so far on the current stream), so its sum of squares
        can be read now, on the side stream, while the layers below still run their backward."""
        if self._sumsq is None:
            return
        if unit in self._sq_launched:
            self._sq_stale = True       # a second backward added to a bucket already summed: recompute at the step
            return
        self._materialize_fresh(unit)
        self._zero_frozen(unit, local=True)
        if self._opt_stream is None:
            self._opt_stream = torch.cuda.Stream(device=self.device)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._opt_stream.wait_event(ev)
        g = self._local_grad(unit)
        with torch.cuda.stream(self._opt_stream):
            N.call("dla_sumsq", N.ptr(g), g.numel(), int(g.dtype == torch.float32), N.ptr(self._sq_scratch),
                   N.ptr(self._sumsq), 1, N.stream_ptr(self.device))
        self._sq_launched.add(unit)

    def _drain(self, keep: int = 0) -> None:
        while len(self._reduce_work) > keep:
            work, unit, out = self._reduce_work.pop(0)
            work.wait()
            if out is None:
                continue
            if self.resident:
                S = self.S[unit]
                self.gradbuf[unit][self.r * S:(self.r + 1) * S].copy_(out)
            else:
                self.gshard[unit].add_(out.float())
                self.gradbuf[unit] = None           # streamed: the full-size bucket lives only until it is reduced
                self._views.pop(unit, None)

    def finish_backward(self) -> None:
        """Call after ``loss.backward()``: sweeps stray autograd grads, waits for the reductions of this
        micro-step and folds reduce-scatter outputs into the shard accumulators."""
        for unit, w in list(self._views.items()):
            collect_autograd_grads(w)
        for unit in sorted(self._plain_used):          # units that ran without recompute have no per-unit hook
            self._passes_seen[unit] = self._passes_expected - 1
            self.grads_done(unit, self._views.get(unit, {}))
        self._plain_used.clear()
        self._materialize_fresh()       # slices nobody wrote this step become real zeros before peers / reductions read them
        self._drain(0)
        if self.fused and self.sync_grads and not self.rs_in_backward:
            self.fc.barrier(0)      # every rank's gradient buckets are complete before anyone reads them
        if self.fused_ddp and self.sync_grads:
            self.fc.all_reduce_(self.grad_arena, 0, self._ddp_total)
            self.stats["reduce_bytes"] += self._ddp_total * _bytes(self.grad_dtype)

    # ------------------------------------------------------------------ optimizer
    def _local_grad(self, unit: str) -> torch.Tensor:
        """The gradient values this rank's optimizer shard consumes (summed over ranks, not averaged)."""
        if self.W == 1:
            return self.gradbuf[unit]
        if self.resident:
            S = self.S[unit]
            return self.gradbuf[unit][self.r * S:(self.r + 1) * S]
        return self.gshard[unit]

    def optimizer_step(self, opt: AdamWConfig, lr: Optional[float] = None, grad_divisor: float = 1.0) -> None:
        """Global-norm clip + AdamW on the rank's shards, then refresh the bf16 parameters.

        Gradients in the buckets are SUMS over ranks and micro-steps; ``grad_divisor`` (world x
        accumulation steps) turns them into the mean the reference's DDP/accumulate produces."""
        assert self.trainable
        if getattr(self, "fc", None) is not None:
            self.fc.check()         # a flag wait timed out since the last step: stop before applying / saving anything
        lr = opt.lr if lr is None else lr
        self.step_count += 1
        t = self.step_count
        b1, b2 = opt.betas
        bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
        prescale = 1.0 / grad_divisor
        cuda = self.device.type == "cuda"
        self.sync_optimizer()
        if self.fused_stream and getattr(self, "stream_rs", False):
            torch.cuda.current_stream(self.device).wait_stream(self._rs_stream)     # the in-backward reduce-scatters
        self._materialize_fresh()
        overlap = cuda and self.overlap_optimizer and (self.W == 1 or self.fused)
        if overlap and self._opt_stream is None:
            self._opt_stream = torch.cuda.Stream(device=self.device)
        if self.fused:
            if not overlap:
                return self._fused_optimizer_step(opt, lr, prescale, bc1, bc2)
            # the whole reduce / AdamW / push sequence moves to the side stream: it is NVLink- and HBM-bound and
            # overlaps the next step's frozen-reference forward (tensor-bound); the first acquire of any policy
            # unit orders the main stream after it
            main = torch.cuda.current_stream(self.device)
            self._opt_stream.wait_stream(main)
            with torch.cuda.stream(self._opt_stream):
                self._fused_optimizer_step(opt, lr, prescale, bc1, bc2)
                self._all_ready = torch.cuda.Event()
                self._all_ready.record(self._opt_stream)
            return
        # ---- global grad norm -------------------------------------------------------
        if self._train_ranges is not None:
            for u in self.units:
                self._zero_frozen(u.name, local=True)
        if cuda:
            s = N.stream_ptr(self.device)
            if self._sq_launched and self._sq_stale:
                torch.cuda.current_stream(self.device).wait_stream(self._opt_stream)
                self._sq_launched.clear()
            self._sq_stale = False
            if self._sq_launched:
                # partial sums of the units summed during backward are already in _sumsq (it was cleared after the last
                # step's clip); the shared scratch is free once the side stream has drained
                torch.cuda.current_stream(self.device).wait_stream(self._opt_stream)
            else:
                self._sumsq.zero_()
            for u in self.units:
                if u.name in self._sq_launched:
                    continue
                g = self._local_grad(u.name)
                N.call("dla_sumsq", N.ptr(g), g.numel(), int(g.dtype == torch.float32), N.ptr(self._sq_scratch),
                       N.ptr(self._sumsq), 1, s)
            if self.W > 1 and self.fused_stream:
                # norm exchange + clip coefficient over peer memory: no NCCL launch in the optimizer phase
                self.fc.norm_clip(self._sumsq, prescale, float(opt.max_grad_norm or 0.0), self._coef, self._norm)
            else:
                if self.W > 1:
                    dist.all_reduce(self._sumsq, group=self.ctx.group)
                self._sumsq.mul_(prescale * prescale)
                N.call("dla_clip_coef", N.ptr(self._sumsq), float(opt.max_grad_norm or 0.0), N.ptr(self._coef), N.ptr(self._norm), s)
            if self.norm_in_backward:
                self._sumsq.zero_()         # ready for the next backward's side-stream partial sums
                self._sq_launched.clear()
            self._last_grad_norm = self._norm
        else:
            sq = torch.zeros((), dtype=torch.float64)
            for u in self.units:
                sq += self._local_grad(u.name).double().pow(2).sum()
            if self.W > 1:
                sq = sq.reshape(1)
                dist.all_reduce(sq, group=self.ctx.group)
                sq = sq.reshape(())
            norm = float(sq.sqrt()) * prescale
            coef = 1.0
            if opt.max_grad_norm and opt.max_grad_norm > 0:
                coef = min(1.0, opt.max_grad_norm / (norm + 1e-6))
            if not math.isfinite(norm):
                coef = 0.0
            self._last_grad_norm = torch.tensor(norm)
        # ---- update -----------------------------------------------------------------
        if overlap:
            # AdamW is a pure HBM stream (28 B per parameter): run it unit by unit, in forward order, on the side
            # stream while the main stream already runs the next step's reference forward; each unit's first
            # acquire waits for that unit's event only
            main = torch.cuda.current_stream(self.device)
            self._opt_stream.wait_stream(main)
            with torch.cuda.stream(self._opt_stream):
                sp = N.stream_ptr(self.device)
                for u in self.units:
                    self._adamw_cuda(u.name, self._local_grad(u.name), opt, lr, bc1, bc2, prescale, sp)
                    self._zero_unit(u.name)
                    e = torch.cuda.Event()
                    e.record(self._opt_stream)
                    self._unit_ready[u.name] = e
            return
        for u in self.units:
            g = self._local_grad(u.name)
            sh = self.shard[u.name]
            if cuda:
                self._adamw_cuda(u.name, g, opt, lr, bc1, bc2, prescale, N.stream_ptr(self.device))
            elif coef != 0.0:       # coef 0 = non-finite gradient norm: skip the update, weight decay included
                for lo, hi in self._ranges(u.name):
                    R.adamw_step(self.master[u.name][lo:hi], self.m[u.name][lo:hi], self.v[u.name][lo:hi], g[lo:hi], lr, b1, b2,
                                 opt.eps, opt.weight_decay, t, grad_scale=prescale * coef)
                    sh[lo:hi].copy_(self.master[u.name][lo:hi].to(sh.dtype))
        # ---- refresh resident gathered params ----------------------------------------
        if self.W > 1 and self.resident:
            works = []
            for u in self.units:
                works.append(dist.all_gather_into_tensor(self.full[u.name], self.shard[u.name], group=self.ctx.group, async_op=True))
                self.stats["allgather_bytes"] += self.P[u.name] * _bytes(self.dtype)
            for w in works:
                w.wait()
        if self.fused_stream:
            self.fc.barrier(2)      # peers pull my shards inside their GEMMs: publish the updated ones first
        self.zero_grad()

    def _fused_optimizer_step(self, opt: AdamWConfig, lr: float, prescale: float, bc1: float, bc2: float) -> None:
        """Fused path (a): per unit ONE kernel pulls this rank's gradient shard from every peer over NVLink,
        sums in fp32, scales and accumulates the grad-norm partial; after the 4-byte norm all-reduce the
        AdamW kernel consumes the fp32 shard and a peer-store kernel writes the updated bf16 shard into
        every rank's resident parameters.  Without clipping the three collapse into one kernel."""
        fc, s = self.fc, N.stream_ptr(self.device)
        b1, b2 = opt.betas
        clip_on = bool(opt.max_grad_norm and opt.max_grad_norm > 0)
        if self.rs_in_backward:
            # v2: the per-unit reduce-scatters were queued on this stream during backward (units that never signalled
            # are reduced now); what remains is the norm exchange, AdamW and the push of the updated shards
            for u in self.units:
                self._launch_unit_reduce(u.name)
            self._rs_launched.clear()
            self._rs_epoch += 1
            fc.norm_clip(self._sumsq, prescale, float(opt.max_grad_norm or 0.0), self._coef, self._norm)
            self._last_grad_norm = self._norm
            mc = self.param_arena.mc_ptr
            esz = _bytes(self.dtype)
            for u in self.units:
                sh, off = self.shard[u.name], self.base[u.name] + self.r * self.S[u.name]
                for lo, hi in self._ranges(u.name):
                    args = (N.ptr(self.master[u.name]) + 4 * lo, N.ptr(self.m[u.name]) + self.m[u.name].element_size() * lo,
                            N.ptr(self.v[u.name]) + self.v[u.name].element_size() * lo, N.ptr(self.gred[u.name]) + 4 * lo)
                    tail = (hi - lo, lr, b1, b2, opt.eps, opt.weight_decay, bc1, bc2, N.ptr(self._coef), prescale,
                            1 | self._moment_flag, s)
                    if mc:      # one multicast store per 16 bytes updates every rank's resident copy (mine included)
                        N.call("dla_adamw_mc", *args, mc + (off + lo) * esz, *tail)
                    else:
                        N.call("dla_adamw", *args, N.ptr(sh) + esz * lo, *tail)
                        fc.push_shard(sh[lo:hi], self.param_arena, off + lo)
                    self.stats["allgather_bytes"] += (hi - lo) * esz * (1 if mc else self.W - 1)
        elif self._train_ranges is not None:
            raise NotImplementedError("freeze_except is implemented on the in-backward reduce-scatter path (the default); "
                                      "unset DLA_RS_IN_BACKWARD=0 / DLA_OVERLAP_OPT=0 to use it with the fused engine")
        elif clip_on or self._moment_flag:        # the single-kernel reduce+AdamW keeps fp32 moments only
            self._sumsq.zero_()
            for u in self.units:
                off = self.base[u.name] + self.r * self.S[u.name]
                fc.reduce_shard(self.grad_arena, off, self.S[u.name], prescale, self.gred[u.name], self._sumsq)
                self.stats["reduce_bytes"] += self.S[u.name] * _bytes(self.grad_dtype) * (self.W - 1)
            dist.all_reduce(self._sumsq, group=self.ctx.group)
            N.call("dla_clip_coef", N.ptr(self._sumsq), float(opt.max_grad_norm or 0.0), N.ptr(self._coef), N.ptr(self._norm), s)
            self._last_grad_norm = self._norm
            for u in self.units:
                sh, off = self.shard[u.name], self.base[u.name] + self.r * self.S[u.name]
                N.call("dla_adamw", N.ptr(self.master[u.name]), N.ptr(self.m[u.name]), N.ptr(self.v[u.name]),
                       N.ptr(self.gred[u.name]), N.ptr(sh), sh.numel(), lr, b1, b2, opt.eps, opt.weight_decay, bc1, bc2,
                       N.ptr(self._coef), 1.0, 1 | self._moment_flag, s)
                fc.push_shard(sh, self.param_arena, off)
                self.stats["allgather_bytes"] += sh.numel() * _bytes(self.dtype) * (self.W - 1)
        else:
            for u in self.units:
                off = self.base[u.name] + self.r * self.S[u.name]
                fc.reduce_adamw_push(self.grad_arena, self.param_arena, off, self.S[u.name], prescale, self.master[u.name],
                                     self.m[u.name], self.v[u.name], lr, (b1, b2), opt.eps, opt.weight_decay, bc1, bc2, None)
            self._last_grad_norm = None
        fc.barrier(0)               # all peers finished reading my gradients and writing my parameters
        self.zero_grad()

    def _ranges(self, unit: str) -> List[tuple]:
        """Shard-local element ranges the optimizer updates: everything, unless :meth:`freeze_except` narrowed it."""
        if self._train_ranges is None:
            return [(0, self.shard[unit].numel())]
        return self._train_ranges[unit]

    def _adamw_cuda(self, unit: str, g: torch.Tensor, opt: AdamWConfig, lr: float, bc1: float, bc2: float, prescale: float,
                    stream: int) -> None:
        sh, b1, b2 = self.shard[unit], opt.betas[0], opt.betas[1]
        gsz, msz, esz = g.element_size(), self.m[unit].element_size(), sh.element_size()
        for lo, hi in self._ranges(unit):
            N.call("dla_adamw", N.ptr(self.master[unit]) + 4 * lo, N.ptr(self.m[unit]) + msz * lo, N.ptr(self.v[unit]) + msz * lo,
                   N.ptr(g) + gsz * lo, N.ptr(sh) + esz * lo, hi - lo, lr, b1, b2, opt.eps, opt.weight_decay, bc1, bc2,
                   N.ptr(self._coef), prescale, int(g.dtype == torch.float32) | self._moment_flag, stream)

    def freeze_except(self, substring: Optional[str]) -> Dict[str, int]:
        """Only parameters whose qualified name ``<unit>.<param>`` contains ``substring`` stay trainable; ``None`` or
        no match freezes everything (``freeze_except_lora``, ``R/src/models/base_model.py:45-49``).  Frozen parameters
        keep their values exactly: their gradients are cleared before the norm and the reduction, and AdamW (weight
        decay included) runs only over the trainable ranges of each shard.  ``freeze_except("")`` undoes it."""
        if substring == "":
            self._train_ranges = None
            return {"trainable_params": self.num_params(), "total_params": self.num_params()}
        ranges: Dict[str, List[tuple]] = {}
        frozen: Dict[str, List[tuple]] = {}
        n_train = 0
        for u in self.units:
            S = self.shard[u.name].numel()
            lo_s = self.r * S if self.W > 1 else 0
            keep: List[tuple] = []
            drop: List[tuple] = []
            for prm in u.params:
                o, n = u.offsets()[prm.name]
                e = o + -(-n // 8) * 8
                hit = bool(substring) and substring in f"{u.name}.{prm.name}"
                n_train += n if hit else 0
                (keep if hit else drop).append((o, e))
            frozen[u.name] = drop
            loc: List[tuple] = []
            for o, e in keep:
                a, b = max(o, lo_s) - lo_s, min(e, lo_s + S) - lo_s
                if a < b:
                    if loc and loc[-1][1] == a:
                        loc[-1] = (loc[-1][0], b)
                    else:
                        loc.append((a, b))
            ranges[u.name] = loc
        self._train_ranges, self._frozen_full = ranges, frozen
        return {"trainable_params": n_train, "total_params": self.num_params()}

    def trainable_numel(self) -> int:
        if not self.trainable:
            return 0
        if self._train_ranges is None:
            return self.num_params()
        tot = 0
        for u in self.units:
            offs = u.offsets()
            fro = {o for o, _ in self._frozen_full[u.name]}
            tot += sum(p.numel for p in u.params if offs[p.name][0] not in fro)
        return tot

    def _zero_frozen(self, unit: str, local: bool = False) -> None:
        """Clear the gradients of frozen parameters: in the full-size bucket before it is reduced, or (``local``) in
        whatever tensor this rank's optimizer shard reads."""
        if self._train_ranges is None:
            return
        if local:
            g = self._local_grad(unit)
            if g is None:
                return
            prev = 0
            for lo, hi in self._train_ranges[unit] + [(g.numel(), g.numel())]:
                if prev < lo:
                    g[prev:lo].zero_()
                prev = hi
            return
        gb = self.gradbuf.get(unit)
        if gb is None:
            return
        for o, e in self._frozen_full[unit]:
            gb[o:min(e, gb.numel())].zero_()

    def publish_shards(self) -> None:
        """Streamed fused mode: peers pull this rank's parameter shards straight out of the symmetric arena inside their
        GEMMs, so after ANY host-driven overwrite of the shards (pretrained load, resume) every rank's copies must
        have landed before anyone's next forward.  No-op otherwise."""
        if self.fused_stream:
            torch.cuda.synchronize(self.device)
            self.ctx.barrier()

    def zero_grad(self) -> None:
        if torch.cuda.is_available() and self.device.type == "cuda" and torch.cuda.current_stream(self.device) != self._opt_stream:
            self.sync_optimizer()
        if self._sq_launched:
            self._sq_stale = True       # partial norms taken during a backward whose gradients are now discarded
        for k in self.gradbuf:
            self._zero_unit(k)
        for g in self.gshard.values():
            g.zero_()

    def _zero_unit(self, unit: str) -> None:
        """Clear one gradient bucket.  On CUDA, slices that are always (over)written by a wgrad GEMM are only marked
        fresh — the first GEMM of the next step stores instead of accumulating (``ops.functional._slot_gemm``) — so
        a step does not memset the whole 2-bytes-per-parameter bucket (16 GB at Llama-3-8B); everything else
        (norm weights, biases, embedding tables, padding) is zero-filled as before."""
        gb = self.gradbuf.get(unit)
        if gb is None:
            return
        slots = self._slots.get(unit) if (self.lazy_zero and self.device.type == "cuda") else None
        if not slots or not any(getattr(t, "_ow", False) for t in slots.values()):
            gb.zero_()
            return
        for t in slots.values():
            if getattr(t, "_ow", False):
                t._fresh = True
            else:
                t.zero_()

    def _materialize_fresh(self, unit: Optional[str] = None) -> None:
        """Slots still marked fresh received no gradient this step: make their zeros real before anything reads them."""
        for name in ([unit] if unit is not None else list(self._slots)):
            for t in (self._slots.get(name) or {}).values():
                if getattr(t, "_fresh", False):
                    t.zero_()
                    t._fresh = False

    # ------------------------------------------------------------------ state access
    def num_params(self) -> int:
        return sum(u.numel for u in self.units)

    def gathered_unit(self, unit: str) -> Weights:
        """Full (unsharded) views of a unit's parameters for export; collective when sharded."""
        self.sync_optimizer()
        if self.full.get(unit) is not None:
            return unit_views(self.full[unit], self._by_name[unit])
        full = torch.empty(self.P[unit], dtype=self.dtype, device=self.device)
        dist.all_gather_into_tensor(full, self.shard[unit], group=self.ctx.group)
        return unit_views(full, self._by_name[unit])

    def load_unit(self, unit: str, tensors: Dict[str, torch.Tensor]) -> None:
        """Overwrite named parameters of a unit from full-size tensors (every rank passes the same)."""
        self.sync_optimizer()
        spec = self._by_name[unit]
        offs = spec.offsets()
        S = self.S[unit]
        lo, hi = self.r * S, (self.r + 1) * S
        for name, t in tensors.items():
            o, n = offs[name]
            flat = t.reshape(-1).to(self.device, self.dtype)
            if self.full.get(unit) is not None:
                self.full[unit][o:o + n].copy_(flat)
            if self.W > 1:
                a, b = max(o, lo), min(o + n, hi)
                if a < b:
                    self.shard[unit][a - lo:b - lo].copy_(flat[a - o:b - o])
        if self.trainable:
            self.master[unit].copy_(self.shard[unit].float())

    def shard_state(self) -> Dict[str, torch.Tensor]:
        """This rank's exact-resume state: fp32 master + Adam moments per unit."""
        self.sync_optimizer()
        out: Dict[str, torch.Tensor] = {}
        for u in self.units:
            if self.trainable:
                out[f"{u.name}.master"] = self.master[u.name]
                out[f"{u.name}.m"] = self.m[u.name]
                out[f"{u.name}.v"] = self.v[u.name]
            else:
                out[f"{u.name}.param"] = self.shard[u.name]
        return out

    def load_shard_state(self, state: Dict[str, torch.Tensor], step_count: int) -> None:
        self.sync_optimizer()
        for u in self.units:
            if self.trainable and f"{u.name}.master" in state:
                self.master[u.name].copy_(state[f"{u.name}.master"])
                self.m[u.name].copy_(state[f"{u.name}.m"].to(self.state_dtype))
                self.v[u.name].copy_(state[f"{u.name}.v"].to(self.state_dtype))
                self.shard[u.name].copy_(self.master[u.name].to(self.dtype))
            elif f"{u.name}.param" in state:
                self.shard[u.name].copy_(state[f"{u.name}.param"].to(self.dtype))
        self.step_count = step_count
        if self.W > 1 and self.resident:
            for u in self.units:
                src = self.shard[u.name].clone() if self.fused else self.shard[u.name]   # fused: the shard aliases ``full``
                dist.all_gather_into_tensor(self.full[u.name], src, group=self.ctx.group)
        self.publish_shards()

    def memory_report(self) -> Dict[str, float]:
        gb = 1 / (1 << 30)
        tot = lambda d: sum(t.numel() * t.element_size() for t in d.values() if t is not None) * gb
        own = tot(self.shard) if self.W > 1 else 0.0
        return {"params_full_gb": tot(self.full), "params_shard_gb": own, "grads_gb": tot(self.gradbuf) + tot(self.gshard),
                "optimizer_gb": tot(self.master) + tot(self.m) + tot(self.v)}


class _AGTicket:
    """Deferred all-gather of one weight matrix: redeemed by the first GEMM that consumes it."""

    def __init__(self, engine: "Engine", unit: str, full: torch.Tensor, ptrs: torch.Tensor, S: int, offset: int, n_out: int) -> None:
        self.engine, self.unit, self.full, self.ptrs, self.S, self.offset, self.n_out = engine, unit, full, ptrs, S, offset, n_out
        self.done = False

    def run(self, x2: torch.Tensor) -> torch.Tensor:
        eng = self.engine
        eng._ag_epoch += 1
        y = eng.fc.ag_gemm(x2, self.ptrs, self.full, self.S, self.offset, self.n_out, eng._ag_flags, eng._ag_epoch)
        eng.stats["allgather_bytes"] += self.n_out * x2.shape[1] * _bytes(eng.dtype)
        self.done = True
        return y


def _bytes(dt: torch.dtype) -> int:
    return torch.empty((), dtype=dt).element_size()


def _default_budget(device: torch.device) -> int:
    if device.type == "cuda":
        free, total = torch.cuda.mem_get_info(device)
        return int(total * 0.45)
    return 1 << 62
