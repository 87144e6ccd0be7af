"""Peer-memory substrate and the fused compute+collective paths (BASELINE.json (a) and (c)).

Symmetric buffers come from ``torch.distributed._symmetric_memory`` (cuMem + handle exchange done by torch:
that is plumbing); what runs on them is ours: ``csrc/peer.cu`` kernels that load/store peer HBM over
NVLink 5 / NVSwitch directly and synchronise with system-scope flags.

    (a) gradient reduce-scatter + cast/scale + grad-norm partial (+ AdamW + push of the updated shard into
        every peer's resident parameters): :meth:`FusedComm.reduce_shard`, :meth:`FusedComm.push_shard`
    (c) cross-rank gather of sequence log-probs + DPO loss / preference-rate / margin: :meth:`FusedComm.dpo_loss`

NCCL equivalents replaced: ``reduce_scatter_tensor`` + ``all_gather_into_tensor`` per unit per step
 and nothing at all for C12 (the reference logs rank-local statistics).
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

from dla_b200.ops import native as N
from dla_b200.parallel.pg import DistContext

PAD_WORDS = 16384          # uint32 flag words per rank: [16 channels][16 ranks] barrier flags, then per-unit rows
UNIT_FLAG_BASE = 256       # word offset of the per-unit flag rows ([unit][world]) used by the in-backward reduce-scatter
GATHER_NMAX = 2048
DPO_BMAX = 1024
WHITEN_SLOTS = 16


def fused_available(ctx: DistContext) -> bool:
    if ctx.world_size < 2 or ctx.device.type != "cuda" or ctx.backend != "nccl":
        return False
    try:
        import torch.distributed._symmetric_memory as symm  # noqa: F401
        return N.available() and N.has("dla_fused_rs_adamw")
    except Exception:
        return False


class SymmetricArena:
    """One symmetric allocation visible to every rank; ``ptrs_dev`` is the address of a device array of
    ``world`` base pointers (rank order) that kernels index to reach a peer's copy."""

    def __init__(self, ctx: DistContext, numel: int, dtype: torch.dtype, zero: bool = True) -> None:
        import torch.distributed._symmetric_memory as symm

        self.ctx = ctx
        self.tensor = symm.empty(numel, dtype=dtype, device=ctx.device)
        if zero:
            self.tensor.zero_()
        self.handle = symm.rendezvous(self.tensor, ctx.group)
        self.ptrs_dev = int(self.handle.buffer_ptrs_dev)
        self.ptrs = list(self.handle.buffer_ptrs)
        # NVSwitch multicast address of the same allocation (0 when the fabric / driver offers none): multimem.ld_reduce
        # and multimem.st on it reach every rank's copy in one instruction.  Only ld/st-class multimem instructions ever
        # use it; bulk-tensor (TMA) loads in this code base read this GPU's own memory only.
        mc = 0
        if os.environ.get("DLA_NO_MULTICAST", "0") != "1":
            try:
                mc = int(self.handle.multicast_ptr or 0)
            except Exception:
                mc = 0
        self.mc_ptr = mc
        torch.cuda.synchronize(ctx.device)
        ctx.barrier()


class FusedComm:
    _instances: Dict[int, "FusedComm"] = {}

    def __init__(self, ctx: DistContext) -> None:
        self.ctx = ctx
        self.W, self.r = ctx.world_size, ctx.rank
        self.pads = SymmetricArena(ctx, PAD_WORDS, torch.int32)
        self.dpo_buf = SymmetricArena(ctx, 2 * self.W * 4 * DPO_BMAX, torch.float32)
        self.gather_buf = SymmetricArena(ctx, 2 * self.W * GATHER_NMAX, torch.float32)
        self.norm_buf = SymmetricArena(ctx, 2 * self.W, torch.float32)
        self.rl_buf = SymmetricArena(ctx, 2 * self.W * 3 * DPO_BMAX, torch.float32)
        self.whiten_buf = SymmetricArena(ctx, 2 * self.W * WHITEN_SLOTS, torch.float32)
        self._epoch = [0] * 16
        self._unit_rows = 0             # per-unit flag rows handed out so far (see claim_unit_rows)
        # watchdog: kernels whose flag wait times out set a word in pinned host memory; check() reads it without
        # touching the device, the training loop calls it once per optimizer step
        self._status = torch.zeros(1, dtype=torch.int32).pin_memory()
        N.call("dla_peer_watch", self._status.data_ptr())
        ctx.fused_comm = self           # DistContext.gather routes small CUDA tensors through all_gather_small

    @classmethod
    def get(cls, ctx: DistContext) -> "FusedComm":
        key = id(ctx)
        if key not in cls._instances:
            cls._instances[key] = FusedComm(ctx)
        return cls._instances[key]

    # ---------------------------------------------------------------- barrier
    def barrier(self, channel: int = 0, timeout_ms: int = 30000) -> None:
        """All ranks' work enqueued before this point (on the current stream) is complete and visible to
        peers before any rank's work enqueued after it starts."""
        self._epoch[channel] += 1
        N.call("dla_peer_barrier", self.pads.ptrs_dev, self.W, self.r, channel, self._epoch[channel], timeout_ms,
               N.stream_ptr(self.ctx.device))

    def check(self, sync: bool = False) -> None:
        """Raise if any peer-flag wait of this process has timed out (a rank is dead or wedged): the kernels that
        follow such a wait ran on incomplete data, so the job must stop before it checkpoints.  The default reads the
        pinned host mirror only (no device synchronisation); ``sync=True`` also queries the device-side flag."""
        bad = int(self._status[0]) != 0
        if sync and not bad:
            import ctypes

            out = ctypes.c_int(0)
            rc = N.load().dla_peer_status(ctypes.byref(out))
            bad = rc != 0 or out.value != 0
        if bad:
            raise RuntimeError("a peer-flag wait timed out: a rank is dead or wedged; this step's collectives are "
                               "incomplete (see dla_peer_status)")

    def claim_unit_rows(self, n_units: int) -> int:
        """Reserve ``n_units`` rows of per-unit flags in the signal pad; returns the first row's word offset."""
        first = UNIT_FLAG_BASE + self._unit_rows * 16
        if first + n_units * 16 > PAD_WORDS:
            raise RuntimeError(f"signal pad exhausted: {self._unit_rows + n_units} units x 16 flag words > {PAD_WORDS - UNIT_FLAG_BASE}")
        self._unit_rows += n_units
        return first

    # ---------------------------------------------------------------- (a)
    def reduce_shard(self, grads: SymmetricArena, offset: int, n: int, scale: float, out_fp32: torch.Tensor,
                     sumsq: Optional[torch.Tensor]) -> None:
        """out[0:n) = scale * sum_r grads_r[offset : offset+n) read over NVLink; sumsq += sum(out^2)."""
        N.call("dla_fused_rs_adamw", grads.ptrs_dev, None, self.W, self.r, n, offset, scale, N.ptr(out_fp32),
               N.ptr(sumsq), None, None, None, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0, 1.0, None, 1 if sumsq is not None else 0,
               N.stream_ptr(self.ctx.device))

    def reduce_adamw_push(self, grads: SymmetricArena, params: SymmetricArena, offset: int, n: int, scale: float,
                          master: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, betas: Tuple[float, float],
                          eps: float, wd: float, bc1: float, bc2: float, clip: Optional[torch.Tensor]) -> None:
        """One kernel: reduce the shard over NVLink, AdamW it, store the new bf16 values into every rank's
        resident parameters (no separate all-gather)."""
        N.call("dla_fused_rs_adamw", grads.ptrs_dev, params.ptrs_dev, self.W, self.r, n, offset, scale, None, None,
               N.ptr(master), N.ptr(m), N.ptr(v), lr, betas[0], betas[1], eps, wd, bc1, bc2, N.ptr(clip), 2,
               N.stream_ptr(self.ctx.device))

    def push_shard(self, src: torch.Tensor, params: SymmetricArena, offset: int) -> None:
        if params.mc_ptr:
            N.call("dla_peer_push_mc", N.ptr(src), params.mc_ptr, offset, src.numel(), N.stream_ptr(self.ctx.device))
        else:
            N.call("dla_peer_push", N.ptr(src), params.ptrs_dev, self.W, self.r, offset, src.numel(), N.stream_ptr(self.ctx.device))

    def all_reduce_(self, arena: SymmetricArena, offset: int, n: int, scale: float = 1.0) -> None:
        """In-place sum over ranks of ``arena[offset : offset+n)`` on every rank (DDP gradients): barrier,
        reduce-my-slice-and-push kernel (inside the switch when the arena has a multicast address), barrier."""
        self.barrier(0)
        if arena.mc_ptr:
            N.call("dla_peer_allreduce_mc", arena.mc_ptr, self.W, self.r, offset, n, scale, N.stream_ptr(self.ctx.device))
        else:
            N.call("dla_peer_allreduce", arena.ptrs_dev, self.W, self.r, offset, n, scale, N.stream_ptr(self.ctx.device))
        self.barrier(0)

    # ---------------------------------------------------------------- (a) v2: in-backward, per unit
    def reduce_unit(self, grads: SymmetricArena, flag_word: int, epoch: int, offset: int, n: int, scale: float,
                    out_fp32: torch.Tensor, sumsq: Optional[torch.Tensor], timeout_ms: int = 30000, max_blocks: int = 0,
                    accumulate: bool = False) -> None:
        """Reduce-scatter of one unit on the CURRENT stream: waits (inside the kernel) until every rank has flagged
        this unit's bucket complete for ``epoch``, then sums this rank's shard — in NVSwitch when possible.
        ``accumulate``: add to ``out_fp32`` instead of overwriting it."""
        N.call("dla_unit_reduce_acc", self.pads.ptrs_dev, flag_word, grads.mc_ptr or None, grads.ptrs_dev, self.W, self.r, n, offset,
               scale, N.ptr(out_fp32), N.ptr(sumsq), epoch, timeout_ms, max_blocks, int(accumulate), N.stream_ptr(self.ctx.device))

    def flag_publish(self, flag_word: int, epoch: int) -> None:
        """Tell every peer (on the current stream, after everything enqueued so far) that this rank reached ``epoch``."""
        N.call("dla_peer_flag_row", self.pads.ptrs_dev, self.W, self.r, flag_word, epoch, 1, 0, N.stream_ptr(self.ctx.device))

    def flag_wait(self, flag_word: int, epoch: int, timeout_ms: int = 30000) -> None:
        """Hold the current stream until every rank has published ``epoch`` on this row."""
        N.call("dla_peer_flag_row", self.pads.ptrs_dev, self.W, self.r, flag_word, epoch, 2, timeout_ms, N.stream_ptr(self.ctx.device))

    def norm_clip(self, sumsq_local: torch.Tensor, prescale: float, max_norm: float, coef: torch.Tensor, norm: torch.Tensor) -> None:
        """Global grad norm + clip coefficient from every rank's partial ``sum(g^2)`` (reset to 0 afterwards)."""
        self._epoch[4] += 1
        N.call("dla_peer_norm_clip", self.norm_buf.ptrs_dev, self.pads.ptrs_dev, self.W, self.r, N.ptr(sumsq_local), prescale,
               max_norm, self._epoch[4], 4, N.ptr(coef), N.ptr(norm), N.stream_ptr(self.ctx.device))

    # ---------------------------------------------------------------- (b)
    def gather_range(self, unit_ptrs: torch.Tensor, full: torch.Tensor, shard_elems: int, offset: int, n: int) -> None:
        """full[offset : offset+n) <- the owners' shards (plain peer loads; small parameters of a streamed unit)."""
        N.call("dla_peer_gather", N.ptr(unit_ptrs), N.ptr(full), shard_elems, offset, n, self.W, N.stream_ptr(self.ctx.device))

    def ag_gemm(self, x2: torch.Tensor, unit_ptrs: torch.Tensor, full: torch.Tensor, shard_elems: int, b_offset: int,
                n_out: int, flags: torch.Tensor, epoch: int, n_copy: int = 8) -> torch.Tensor:
        """y = x2 @ W^T where W = full[b_offset : b_offset + n_out*K] is gathered from its owners INSIDE the GEMM."""
        M, K = x2.shape
        y = torch.empty(M, n_out, dtype=x2.dtype, device=x2.device)
        N.call("dla_ag_gemm_bf16", N.ptr(x2), N.ptr(unit_ptrs), N.ptr(full), shard_elems, b_offset, N.ptr(y), M, n_out, K,
               x2.stride(0), y.stride(0), self.W, self.r, N.ptr(flags), epoch, n_copy, 0, N.stream_ptr(x2.device))
        return y

    def all_gather_small(self, x: torch.Tensor) -> torch.Tensor:
        """``[world, n]`` fp32 gather of a small fp32 vector (n <= GATHER_NMAX) in one single-block kernel over peer
        memory: stores into every peer's slot, flag exchange, copy-out (eval losses / logged statistics)."""
        x = x.detach().float().contiguous().view(-1)
        n = x.numel()
        assert 0 < n <= GATHER_NMAX
        out = torch.empty(self.W, n, dtype=torch.float32, device=x.device)
        self._epoch[3] += 1
        N.call("dla_peer_allgather_small", self.gather_buf.ptrs_dev, self.pads.ptrs_dev, self.W, self.r, N.ptr(x), n, GATHER_NMAX,
               self._epoch[3], 3, N.ptr(out), N.stream_ptr(x.device))
        return out

    # ---------------------------------------------------------------- (c)
    def dpo_loss(self, pol_pos, pol_neg, ref_pos, ref_neg, beta: float, label_smoothing: float):
        """Returns (global stats [loss, pref_rate, margin], dpos, dneg) — gradients are of the local-batch mean."""
        B = pol_pos.numel()
        assert B <= DPO_BMAX
        dev = pol_pos.device
        out = torch.empty(3, dtype=torch.float32, device=dev)
        dpp, dpn = torch.empty(B, dtype=torch.float32, device=dev), torch.empty(B, dtype=torch.float32, device=dev)
        self._epoch[1] += 1
        N.call("dla_fused_gather_dpo", self.dpo_buf.ptrs_dev, self.pads.ptrs_dev, self.W, self.r, N.ptr(pol_pos), N.ptr(pol_neg),
               N.ptr(ref_pos), N.ptr(ref_neg), B, beta, label_smoothing, self._epoch[1], N.ptr(out), N.ptr(dpp), N.ptr(dpn),
               N.stream_ptr(dev))
        return out, dpp, dpn


    def reinforce_loss(self, pol: torch.Tensor, ref: torch.Tensor, score: torch.Tensor, kl_coef: float):
        """Returns (out [loss, mean kl, mean reward, mean score, baseline] over the GLOBAL batch, dpol [B]).
        Every rank must pass the same number of rows (the kernel reads ``world x B`` entries; ``cli.train_rlhf`` trims
        the global batch to a multiple of the world size for that reason)."""
        B = pol.numel()
        assert B <= DPO_BMAX
        out = torch.empty(5, dtype=torch.float32, device=pol.device)
        dpol = torch.empty(B, dtype=torch.float32, device=pol.device)
        self._epoch[5] += 1
        N.call("dla_fused_gather_reinforce", self.rl_buf.ptrs_dev, self.pads.ptrs_dev, self.W, self.r, N.ptr(pol), N.ptr(ref),
               N.ptr(score), B, kl_coef, self._epoch[5], N.ptr(out), N.ptr(dpol), N.stream_ptr(pol.device))
        return out, dpol

    def whiten_(self, adv: torch.Tensor, mask: torch.Tensor, stats: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Whitens ``adv`` in place with the mean / variance over ALL ranks' masked tokens; ``stats`` (<= 13 per-rank
        sums) come back added up over ranks."""
        assert adv.dtype == torch.float32 and adv.is_contiguous() and mask.dtype == torch.float32 and mask.is_contiguous()
        k = 0 if stats is None else stats.numel()
        out = torch.empty(max(k, 1), dtype=torch.float32, device=adv.device)
        self._epoch[6] += 1
        N.call("dla_fused_gather_whiten", self.whiten_buf.ptrs_dev, self.pads.ptrs_dev, self.W, self.r, N.ptr(adv), N.ptr(mask),
               adv.numel(), N.ptr(stats) if k else None, k, self._epoch[6], N.ptr(out), N.stream_ptr(adv.device))
        return out[:k] if k else None


class _FusedReinforce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, comm, pol, ref, score, kl_coef):
        pol_c, ref_c, sc_c = (t.detach().float().contiguous() for t in (pol, ref, score))
        out, dpol = comm.reinforce_loss(pol_c, ref_c, sc_c, kl_coef)
        ctx.save_for_backward(dpol)
        return out[0], out[1:].detach()

    @staticmethod
    def backward(ctx, g, _gs):
        (dpol,) = ctx.saved_tensors
        return None, dpol * g, None, None, None


def fused_reinforce_loss(comm: "FusedComm", pol, ref, score, kl_coef: float):
    """``(loss, stats)``: loss value and ``stats = [mean kl, mean reward, mean score, baseline]`` are over the global
    rollout batch; the gradient is that of the local-batch mean (the engine divides by world)."""
    return _FusedReinforce.apply(comm, pol, ref, score, kl_coef)


class _FusedDPOLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, comm, pp, pn, rp, rn, beta, ls):
        pp, pn, rp, rn = (t.detach().float().contiguous() for t in (pp, pn, rp, rn))
        out, dpp, dpn = comm.dpo_loss(pp, pn, rp, rn, beta, ls)
        ctx.save_for_backward(dpp, dpn)
        # the loss VALUE is the global mean; its gradient (below) is that of the local mean, which the
        # engine's divide-by-world turns into the global-mean gradient
        return out[0], out[1:].detach()

    @staticmethod
    def backward(ctx, g, _gs):
        dpp, dpn = ctx.saved_tensors
        return None, dpp * g, dpn * g, None, None, None, None


def fused_dpo_loss(comm: FusedComm, pol_pos, pol_neg, ref_pos, ref_neg, beta: float, label_smoothing: float = 0.0):
    return _FusedDPOLoss.apply(comm, pol_pos, pol_neg, ref_pos, ref_neg, beta, label_smoothing)
