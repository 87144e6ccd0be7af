"""Plain PyTorch restatements of every op.

Two jobs: (1) the CPU / gloo path of the framework (differentiable through autograd), and
(2) the fp32 oracle every sm_100a kernel is tested against.  Formulas
follow what the reference executes through transformers / ATen.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """K3: ``x * rsqrt(mean(x^2) + eps) * w`` computed in fp32, cast back."""
    xf = x.float()
    y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (y * w.float()).to(x.dtype)


def layernorm(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float) -> torch.Tensor:
    return F.layer_norm(x.float(), (x.shape[-1],), w.float(), None if b is None else b.float(), eps).to(x.dtype)


def rope_tables(positions: torch.Tensor, rot_dim: int, theta: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """cos/sin ``[..., rot_dim/2]`` in fp32 for integer positions."""
    inv = 1.0 / (theta ** (torch.arange(0, rot_dim, 2, device=positions.device, dtype=torch.float32) / rot_dim))
    ang = positions.float()[..., None] * inv
    return ang.cos(), ang.sin()


def rope(x: torch.Tensor, positions: torch.Tensor, theta: float, rot_dim: Optional[int] = None) -> torch.Tensor:
    """K4: rotate-half RoPE (HF Llama convention) on ``x [B, T, H, D]``; ``positions [B, T]``.
    ``rot_dim < D`` gives partial rotary (phi-2 style)."""
    D = x.shape[-1]
    rd = rot_dim or D
    cos, sin = rope_tables(positions, rd, theta)           # [B, T, rd/2]
    cos, sin = cos[:, :, None, :], sin[:, :, None, :]
    xf = x.float()
    x1, x2, rest = xf[..., : rd // 2], xf[..., rd // 2 : rd], xf[..., rd:]
    out = torch.cat([x1 * cos - x2 * sin, x2 * cos + x1 * sin, rest], dim=-1)
    return out.to(x.dtype)


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    """K5 on a fused ``[.., 2F]`` gate|up tensor: ``silu(gate) * up``."""
    g, u = gu.float().chunk(2, dim=-1)
    return (F.silu(g) * u).to(gu.dtype)


def gelu_tanh(x: torch.Tensor) -> torch.Tensor:
    return F.gelu(x.float(), approximate="tanh").to(x.dtype)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = True,
              seqlens: Optional[torch.Tensor] = None, scale: Optional[float] = None,
              segment_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K2: softmax(QK^T * scale + mask) V.  ``q [B, T, Hq, D]``, ``k/v [B, S, Hkv, D]`` (GQA by
    head repetition).  ``seqlens [B]`` masks right padding on the key side; padded query rows
    produce zeros.  ``segment_ids [B, T]`` restricts attention within packed documents."""
    B, T, Hq, D = q.shape
    S, Hk = k.shape[1], k.shape[2]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    qf = q.float().permute(0, 2, 1, 3)
    kf = k.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    vf = v.float().permute(0, 2, 1, 3).repeat_interleave(Hq // Hk, dim=1)
    s = torch.matmul(qf, kf.transpose(-1, -2)) * scale      # [B, Hq, T, S]
    neg = torch.finfo(torch.float32).min
    mask = torch.ones(B, 1, T, S, dtype=torch.bool, device=q.device)
    if causal:
        mask = mask & torch.ones(T, S, dtype=torch.bool, device=q.device).tril(S - T)[None, None]
    if seqlens is not None:
        mask = mask & (torch.arange(S, device=q.device)[None, :] < seqlens[:, None])[:, None, None, :]
    if segment_ids is not None:
        mask = mask & (segment_ids[:, None, :, None] == segment_ids[:, None, None, :])
    s = s.masked_fill(~mask, neg)
    p = torch.softmax(s, dim=-1)
    p = torch.where(mask.any(-1, keepdim=True), p, torch.zeros_like(p))
    o = torch.matmul(p, vf).permute(0, 2, 1, 3)
    if seqlens is not None and T == S:
        o = o * (torch.arange(T, device=q.device)[None, :] < seqlens[:, None])[:, :, None, None]
    return o.to(q.dtype)


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    return F.linear(x, w, b)


def token_logprobs(hidden: torch.Tensor, w: torch.Tensor, labels: torch.Tensor,
                   bias: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """K6/K7 core: per-token ``log_softmax(hidden @ w^T)[label]`` and the log-sum-exp, both fp32
    ``[M]``.  Labels < 0 (ignore) give logp 0."""
    logits = F.linear(hidden.float(), w.float(), None if bias is None else bias.float())
    lse = torch.logsumexp(logits, dim=-1)
    safe = labels.clamp(min=0)
    picked = logits.gather(-1, safe[:, None]).squeeze(-1)
    logp = torch.where(labels >= 0, picked - lse, torch.zeros_like(lse))
    return logp, lse


def sequence_mean_logprob(token_logp: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
    """``sum(logp * mask) / clamp(sum(mask), 1)`` per sequence — the reference's length-normalised
    log-prob (``R/src/training/train_dpo.py:38``; defect D4 kept for parity)."""
    m = mask.to(token_logp.dtype)
    return (token_logp * m).sum(1) / m.sum(1).clamp(min=1)


def dpo_loss(pol_pos, pol_neg, ref_pos, ref_neg, beta: float, label_smoothing: float = 0.0) -> torch.Tensor:
    """K8: ``-logsigmoid(beta * ((pol+ - pol-) - (ref+ - ref-)))`` mean
    (``R/src/training/train_dpo.py:42-44``); optional conservative-DPO label smoothing for the
    declared-but-unread ``model.label_smoothing`` key (``R/config/dpo_config.yaml:9``)."""
    z = beta * ((pol_pos - pol_neg) - (ref_pos - ref_neg))
    loss = -F.logsigmoid(z)
    if label_smoothing:
        loss = (1 - label_smoothing) * loss - label_smoothing * F.logsigmoid(-z)
    return loss.mean()


def pairwise_loss(chosen: torch.Tensor, rejected: torch.Tensor) -> torch.Tensor:
    """K9 Bradley-Terry: ``-logsigmoid(r+ - r-)`` mean (``R/src/models/reward_model.py:67-68``)."""
    return -F.logsigmoid(chosen - rejected).mean()


def reinforce_loss(policy_logp, ref_logp, scores, kl_coef: float, baseline: Optional[torch.Tensor] = None):
    """K11 (``R/src/training/train_rlhf.py:149-153``): kl = pi - ref; r = score - c*kl;
    A = r - mean(r); L = -mean(A.detach() * pi).  ``baseline`` overrides the batch mean (global
    cross-rank baseline, fused path (c))."""
    kl = policy_logp - ref_logp
    rewards = scores - kl_coef * kl
    base = rewards.mean() if baseline is None else baseline
    adv = rewards - base
    return -(adv.detach() * policy_logp).mean(), kl, rewards


def distill_kl(student_logits: torch.Tensor, teacher_logits: list, mask: Optional[torch.Tensor],
               temperature: float = 1.0) -> torch.Tensor:
    """K12 (``R/src/training/train_distill.py:132-144``): forward KL of the ensemble-mean teacher
    distribution against the student, summed over the vocab, masked mean over tokens."""
    slp = F.log_softmax(student_logits.float() / temperature, dim=-1)
    tp = torch.stack([F.softmax(t.float() / temperature, dim=-1) for t in teacher_logits], 0).mean(0)
    kl = F.kl_div(slp, tp, reduction="none").sum(-1)
    if mask is not None:
        m = mask.to(kl.dtype)
        return (kl * m).sum() / m.sum()
    return kl.mean()


def adamw_step(master, m, v, grad, lr, beta1, beta2, eps, wd, step, grad_scale: float = 1.0) -> None:
    """K13 decoupled-weight-decay Adam on fp32 state, in place (torch.optim.AdamW semantics)."""
    g = grad.float() * grad_scale
    master.mul_(1 - lr * wd)
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    master.addcdiv_(m, denom, value=-lr / bc1)


def top_p_sample(logits: torch.Tensor, temperature: float, top_p: float, generator=None) -> torch.Tensor:
    """K16 sampler: temperature, nucleus filter (keep the smallest prefix of the sorted
    distribution with cumulative mass >= top_p), multinomial."""
    if temperature <= 0:
        return logits.argmax(-1)
    probs = torch.softmax(logits.float() / temperature, dim=-1)
    sp, si = probs.sort(-1, descending=True)
    cum = sp.cumsum(-1)
    drop = (cum - sp) >= top_p
    sp = sp.masked_fill(drop, 0.0)
    sp = sp / sp.sum(-1, keepdim=True)
    pick = torch.multinomial(sp, 1, generator=generator)
    return si.gather(-1, pick).squeeze(-1)


def gae(rewards: torch.Tensor, values: torch.Tensor, mask: torch.Tensor, gamma: float, lam: float, whiten: bool = False):
    """Generalised advantage estimation over ``[B, T]`` token rewards/values (worklist #20); ``whiten`` normalises
    the advantages over the masked tokens (returns keep the raw advantages)."""
    B, T = rewards.shape
    adv = torch.zeros_like(rewards)
    last = torch.zeros(B, dtype=rewards.dtype, device=rewards.device)
    for t in reversed(range(T)):
        nxt = values[:, t + 1] if t + 1 < T else torch.zeros_like(last)
        nm = mask[:, t + 1].to(rewards.dtype) if t + 1 < T else torch.zeros_like(last)
        delta = rewards[:, t] + gamma * nxt * nm - values[:, t]
        last = delta + gamma * lam * nm * last
        adv[:, t] = last
    m = mask.to(adv.dtype)
    adv = adv * m
    ret = adv + values
    if whiten:
        denom = m.sum().clamp(min=1)
        mean = (adv * m).sum() / denom
        std = (((adv - mean) ** 2 * m).sum() / denom).sqrt().clamp(min=1e-6)
        adv = (adv - mean) / std * m
    return adv, ret


def ppo_loss(new_logp, old_logp, advantages, values, old_values, returns, mask, clip: float = 0.2, vclip: float = 0.2,
             vf_coef: float = 0.5):
    """Token-level clipped PPO objective + clipped value loss (masked mean); returns ``(loss, stats)`` with
    ``stats = [pg_loss, vf_loss, approx_kl, clipfrac]``.  ``values`` may be None (policy term only)."""
    m = mask.to(new_logp.dtype)
    denom = m.sum().clamp(min=1)
    ratio = torch.exp(new_logp - old_logp)
    pg = torch.max(-advantages * ratio, -advantages * ratio.clamp(1 - clip, 1 + clip))
    pg_loss = (pg * m).sum() / denom
    if values is not None:
        vclipped = old_values + (values - old_values).clamp(-vclip, vclip)
        vf = 0.5 * torch.max((values - returns) ** 2, (vclipped - returns) ** 2)
        vf_loss = (vf * m).sum() / denom
    else:
        vf_loss = torch.zeros((), dtype=new_logp.dtype, device=new_logp.device)
    approx_kl = (0.5 * ((new_logp - old_logp) ** 2) * m).sum() / denom
    clipfrac = (((ratio - 1).abs() > clip).to(m.dtype) * m).sum() / denom
    stats = torch.stack([pg_loss.detach(), vf_loss.detach(), approx_kl.detach(), clipfrac.detach()])
    return pg_loss + vf_coef * vf_loss, stats


def scalar_head(hidden: torch.Tensor, mask: Optional[torch.Tensor], w: torch.Tensor, bias: Optional[torch.Tensor],
                keep: Optional[torch.Tensor] = None, mode: str = "last_token") -> torch.Tensor:
    """K10: pool (``last_token`` = index ``mask.sum(1) - 1``, ``mean`` = masked mean, ``token`` = none) -> dropout
    mask -> ``Linear(H, 1)``; fp32 result ``[B]`` (``[B, T]`` for ``token``).  ``R/src/models/reward_model.py:54-64``."""
    wv = w.float().reshape(-1)
    b = bias.float().reshape(()) if bias is not None else 0.0
    if mode == "token":
        return hidden.float() @ wv + b
    if mode == "last_token":
        idx = (mask.sum(dim=1) - 1).clamp(min=0)
        pooled = hidden[torch.arange(hidden.shape[0], device=hidden.device), idx].float()
    else:
        m = mask.unsqueeze(-1).float()
        pooled = (hidden.float() * m).sum(1) / mask.sum(1, keepdim=True).clamp(min=1).float()
    if keep is not None:
        pooled = pooled * keep
    return pooled @ wv + b
