"""Alignment objectives: SFT cross-entropy, Bradley-Terry reward loss, DPO, REINFORCE/PPO, distillation.

Each function is the math of one reference hot loop, restated on the native model API:
  * :func:`sft_loss`            — ``outputs.loss`` (``R/src/training/train_sft.py:145-146``)
  * :func:`sequence_logprobs`   — ``compute_logprobs`` / ``sequence_logprob``
                                  (``R/src/training/train_dpo.py:31-39``, ``R/src/training/train_rlhf.py:50-58``)
  * :func:`dpo_step`            — four log-prob passes + ``dpo_loss`` (``R/src/training/train_dpo.py:113-117``)
  * :func:`reward_step`         — ``pairwise_loss`` on chosen/rejected scores (``R/src/training/train_reward.py:140-148``)
  * :func:`reinforce_step`      — the reference's "PPO" (``R/src/training/train_rlhf.py:149-153``)
  * :func:`ppo_losses`          — clipped PPO with value head + GAE (declared keys, ``R/config/rlhf_config.yaml:14,17``)
  * :func:`distill_step`        — CE or ensemble forward-KL (``R/src/training/train_distill.py:127-147``)
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as tF

from dla_b200.models.transformer import CausalLM
from dla_b200.ops import functional as F
from dla_b200.ops import reference as R


def sft_loss(model: CausalLM, batch: Dict[str, torch.Tensor]) -> torch.Tensor:
    return model.loss(batch["input_ids"], batch.get("attention_mask"), batch["labels"],
                      **{k: batch[k] for k in ("position_ids", "segment_ids") if k in batch})


def sequence_logprobs(model: CausalLM, input_ids: torch.Tensor, attention_mask: torch.Tensor,
                      reduction: str = "mean", loss_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-sequence log-prob ``[B]``.  Default = the reference's length-normalised mean over all non-pad
    tokens, prompt included (defect D4 kept for parity); ``reduction="sum"`` and ``loss_mask`` (e.g.
    response-only) give the canonical DPO quantity."""
    tok = model.token_logprobs(input_ids, attention_mask)            # [B, T-1]
    mask = (loss_mask if loss_mask is not None else attention_mask)[:, 1:]
    return F.sequence_logprob(tok, mask, reduction)


def _cat_pairs(chosen: Dict[str, torch.Tensor], rejected: Dict[str, torch.Tensor], pad_id: int = 0):
    Tc, Tr = chosen["input_ids"].shape[1], rejected["input_ids"].shape[1]
    T = max(Tc, Tr)

    def pad(t: torch.Tensor, fill: int) -> torch.Tensor:
        return t if t.shape[1] == T else tF.pad(t, (0, T - t.shape[1]), value=fill)

    ids = torch.cat([pad(chosen["input_ids"], pad_id), pad(rejected["input_ids"], pad_id)], 0)
    mask = torch.cat([pad(chosen["attention_mask"], 0), pad(rejected["attention_mask"], 0)], 0)
    return ids, mask


def dpo_step(policy: CausalLM, ref: CausalLM, batch: Dict[str, Dict[str, torch.Tensor]], beta: float = 0.1,
             label_smoothing: float = 0.0, concat: bool = True, reduction: str = "mean", pad_id: int = 0,
             loss_scale: Optional[float] = None):
    """Returns ``(loss, info)``; ``info`` holds detached per-sequence log-probs and logged statistics.

    ``concat=True`` runs chosen and rejected as one 2B batch (one parameter gather per layer instead of
    two, same math); ``concat=False`` reproduces the reference's separate forwards."""
    ch, rj = batch["chosen"], batch["rejected"]
    B = ch["input_ids"].shape[0]
    if concat:
        ids, mask = _cat_pairs(ch, rj, pad_id)
        # frozen reference first: it does not depend on the optimizer update of the previous step, so it overlaps
        # the side-stream AdamW / gradient exchange (Engine.optimizer_step) that the policy forward must wait for
        with torch.no_grad():
            rf = sequence_logprobs(ref, ids, mask, reduction)
        pol = sequence_logprobs(policy, ids, mask, reduction)
        pol_pos, pol_neg, ref_pos, ref_neg = pol[:B], pol[B:], rf[:B], rf[B:]
    else:
        pol_pos = sequence_logprobs(policy, ch["input_ids"], ch["attention_mask"], reduction)
        pol_neg = sequence_logprobs(policy, rj["input_ids"], rj["attention_mask"], reduction)
        with torch.no_grad():   # the reference forgets no_grad here (defect D16); the math is identical
            ref_pos = sequence_logprobs(ref, ch["input_ids"], ch["attention_mask"], reduction)
            ref_neg = sequence_logprobs(ref, rj["input_ids"], rj["attention_mask"], reduction)
    eng = policy.p
    if (getattr(eng, "fused", False) or getattr(eng, "fused_ddp", False)) and pol_pos.is_cuda and torch.is_grad_enabled():
        # fused path (c): the cross-rank gather of the four log-prob vectors happens inside the loss kernel
        # over NVLink peer memory; loss / preference rate / margin are GLOBAL-batch statistics
        from dla_b200.parallel.fused import fused_dpo_loss

        loss, stats = fused_dpo_loss(eng.fc, pol_pos, pol_neg, ref_pos, ref_neg, beta, label_smoothing)
    else:
        loss, stats = F.dpo_loss(pol_pos, pol_neg, ref_pos, ref_neg, beta, label_smoothing, loss_scale)
    info = {"pol_pos": pol_pos.detach(), "pol_neg": pol_neg.detach(), "ref_pos": ref_pos.detach(),
            "ref_neg": ref_neg.detach(), "preference_rate": stats[0], "margin": stats[1]}
    return loss, info


def reward_step(rm: Any, batch: Dict[str, Dict[str, torch.Tensor]], concat: bool = True, pad_id: int = 0):
    ch, rj = batch["chosen"], batch["rejected"]
    B = ch["input_ids"].shape[0]
    if concat:
        ids, mask = _cat_pairs(ch, rj, pad_id)
        s = rm(ids, mask)
        cs, rs = s[:B], s[B:]
    else:
        cs, rs = rm(ch["input_ids"], ch["attention_mask"]), rm(rj["input_ids"], rj["attention_mask"])
    loss, acc = F.pairwise_loss(cs, rs)
    return loss, {"accuracy": acc, "chosen_scores": cs.detach(), "rejected_scores": rs.detach()}


def reinforce_step(policy_logp: torch.Tensor, ref_logp: torch.Tensor, scores: torch.Tensor, kl_coef: float,
                   baseline: Optional[torch.Tensor] = None):
    """Sequence-level REINFORCE with batch-mean baseline and KL penalty — exactly the reference's update; one
    kernel on CUDA (``csrc/rl.cu``)."""
    return F.reinforce_loss(policy_logp, ref_logp, scores, kl_coef, baseline)


def ppo_losses(new_logp: torch.Tensor, old_logp: torch.Tensor, advantages: torch.Tensor, values: torch.Tensor,
               old_values: torch.Tensor, returns: torch.Tensor, mask: torch.Tensor, clip: float = 0.2,
               vclip: float = 0.2, vf_coef: float = 0.5):
    """Token-level clipped PPO objective + clipped value loss; returns (loss, stats).  One kernel on CUDA."""
    loss, st = F.ppo_loss(new_logp, old_logp, advantages, values, old_values, returns, mask, clip, vclip, vf_coef)
    return loss, {"pg_loss": st[0], "vf_loss": st[1], "approx_kl": st[2], "clipfrac": st[3]}


class _ChunkedDistillKL(torch.autograd.Function):
    """Forward KL(teacher-ensemble-mean || student) over the vocab without holding (K+2) [B,T,V] tensors:
    rows are processed in chunks; backward recomputes the student chunk and emits
    ``(softmax(s) - t) * mask / n`` straight into the dgrad / wgrad GEMMs."""

    @staticmethod
    def forward(ctx, hs, ws, mask, temperature, chunk, *teachers):
        M = hs.shape[0]
        ht, wt = list(teachers[0::2]), list(teachers[1::2])
        m = mask.reshape(-1).float()
        denom = m.sum().clamp(min=1)
        total = torch.zeros((), dtype=torch.float32, device=hs.device)
        for s in range(0, M, chunk):
            e = min(s + chunk, M)
            if hs.is_cuda:
                sl = F.gemm(hs[s:e], ws, out_dtype=torch.float32)
                tls = [F.gemm(h_t[s:e], w_t, out_dtype=torch.float32) for h_t, w_t in zip(ht, wt)]
                kl, _ = F.distill_kl_rows(sl, tls, None, temperature, want_kl=True, want_grad=False)
                total = total + (kl * m[s:e]).sum()
                continue
            sl = hs[s:e].float() @ ws.float().t()
            slp = torch.log_softmax(sl / temperature, -1)
            tp = None
            for h_t, w_t in zip(ht, wt):
                tl = h_t[s:e].float() @ w_t.float().t()
                pr = torch.softmax(tl / temperature, -1)
                tp = pr if tp is None else tp + pr
            tp = tp / len(ht)
            kl = (torch.xlogy(tp, tp) - tp * slp).sum(-1)
            total = total + (kl * m[s:e]).sum()
        ctx.save_for_backward(hs, ws, m, denom, *teachers)
        ctx.meta = (temperature, chunk)
        ctx.wref = ws
        return total / denom

    @staticmethod
    def backward(ctx, g):
        hs, ws, m, denom, *teachers = ctx.saved_tensors
        temperature, chunk = ctx.meta
        ht, wt = list(teachers[0::2]), list(teachers[1::2])
        M = hs.shape[0]
        dh = torch.empty_like(hs)
        slot, acc = F._slot_gemm(ctx.wref)
        dw = None if slot is not None else torch.zeros_like(ws)
        if slot is None:
            acc = True
        wtgt = slot if slot is not None else dw
        for s in range(0, M, chunk):
            e = min(s + chunk, M)
            if hs.is_cuda:
                sl = F.gemm(hs[s:e], ws, out_dtype=torch.float32)
                tls = [F.gemm(h_t[s:e], w_t, out_dtype=torch.float32) for h_t, w_t in zip(ht, wt)]
                _, dl = F.distill_kl_rows(sl, tls, m[s:e] / denom * g, temperature, want_kl=False, want_grad=True)
                F.gemm(dl, ws, b_mn=True, out=dh[s:e])
                F.gemm(dl, hs[s:e], a_mn=True, b_mn=True, out=wtgt, accumulate=acc)
                acc = True
                continue
            sl = hs[s:e].float() @ ws.float().t()
            sp = torch.softmax(sl / temperature, -1)
            tp = None
            for h_t, w_t in zip(ht, wt):
                tl = h_t[s:e].float() @ w_t.float().t()
                pr = torch.softmax(tl / temperature, -1)
                tp = pr if tp is None else tp + pr
            tp = tp / len(ht)
            dl = ((sp - tp) * (m[s:e] / denom * g / temperature)[:, None]).to(hs.dtype)
            dh[s:e] = dl @ ws
            if not acc:
                wtgt.zero_()
                acc = True
            wtgt.add_((dl.t() @ hs[s:e]).to(wtgt.dtype))
        return (dh, dw, None, None, None) + (None,) * len(teachers)


def distill_step(student: CausalLM, teachers: Sequence[CausalLM], batch: Dict[str, torch.Tensor], use_kl: bool,
                 temperature: float = 1.0, chunk: int = 1024) -> torch.Tensor:
    """KL mode: forward KL vs the mean teacher distribution at every non-pad position (no label shift,
    like the reference); CE mode: the SFT loss on teacher tokens."""
    if not (use_kl and teachers):
        return sft_loss(student, batch)
    ids, mask = batch["input_ids"], batch["attention_mask"]
    for t in teachers:
        if t.cfg.vocab_size != student.cfg.vocab_size:
            raise ValueError(f"KL distillation needs equal vocab sizes (student {student.cfg.vocab_size}, "
                             f"teacher {t.cfg.vocab_size}); defect D12 in the reference's default config")
    hs = student.hidden(ids, mask)
    B, T, H = hs.shape
    ws, _ = student._lm_head_weight()
    tw: List[torch.Tensor] = []
    with torch.no_grad():
        for t in teachers:
            ht = t.hidden(ids, mask)
            tw += [ht.reshape(B * T, -1), t._lm_head_weight()[0]]
    return _ChunkedDistillKL.apply(hs.reshape(B * T, H), ws, mask, float(temperature), int(chunk), *tw)
