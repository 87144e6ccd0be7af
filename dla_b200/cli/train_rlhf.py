"""RLHF entrypoint: rollout -> score -> policy update (``ppo.algo: reinforce | ppo``).

Parity with ``R/src/training/train_rlhf.py`` (``ppo.algo: reinforce``, the default): policy (checkpointing on) +
frozen reference + reward model built on ``reward_model.base_model_name_or_path`` (default: the policy path,
``:70-76``) with optional weights from ``reward_model.path`` (``:77-84``); prompts from ``sampling`` (hf column or local
JSONL, ``:34-47``); AdamW(lr=ppo.learning_rate|1e-6, betas=(0.9,0.95)) (``:94-98``); per step: sample
``ppo.batch_size`` prompts identically on every rank and take the rank slice (``:113-114``), generate with
``ppo.generation_params`` (``:123-124``), sequence log-probs of policy (grad) and reference (``:127-129``),
reward score of prompt+response (``:131-147``), ``kl = pi - ref; r = score - kl_coef*kl; A = r - mean(r);
loss = -mean(A.detach()*pi)`` (``:149-153``), clip, step; log loss and kl every 10 steps (``:160-162``); one
save at the end (``:164-165``).  Design fixes: explicit lengths instead of ``!= pad``
masks, reward model in eval mode under no_grad, token ids stay on the device when policy and reward
share a tokenizer, the advantage baseline is the GLOBAL batch mean (cross-rank gather fused with the loss).

``ppo.algo: ppo`` honours the declared-but-unread keys (``mini_batch_size``, ``target_kl``,
``R/config/rlhf_config.yaml:14,17``): value head on the policy trunk, GAE, clipped ratio / value losses,
``ppo.epochs`` passes over mini-batches, early stop at ``target_kl``.
"""
from __future__ import annotations

import random
from typing import Any, Dict, List

import torch

from dla_b200.algo import ppo_losses, reinforce_step, sequence_logprobs
from dla_b200.cli.common import TrainLoop, parse_args, setup
from dla_b200.data import read_jsonl, split_between_processes
from dla_b200.models.factory import build_reward_model, load_causal_lm
from dla_b200.models.reward import ValueHead, value_head_unit
from dla_b200.models.config import resolve_model_config
from dla_b200.ops import functional as F
from dla_b200.optim import AdamWConfig
from dla_b200.rollout import Generator, decode_responses, splice_separator


def load_prompts(cfg: Dict[str, Any]) -> List[str]:
    """``sampling.{source, hf_path, hf_name, split, prompt_key, prompt_path|path}`` (``R/src/training/train_rlhf.py:34-47``)."""
    source = cfg.get("source", "local")
    if source == "hf":
        try:
            from datasets import load_dataset

            ds = load_dataset(cfg["hf_path"], cfg.get("hf_name"), split=cfg.get("split", "train"), streaming=False)
            key = cfg.get("prompt_key", "prompt")
            rows = [row[key] for row in ds if row.get(key)]
            return rows[: int(cfg["limit"])] if cfg.get("limit") else rows      # ``limit`` is inert in the reference preset
        except Exception as e:
            fallback = cfg.get("prompt_path") or cfg.get("path")
            if not fallback:
                raise
            print(f"[rlhf] hf prompts unavailable ({type(e).__name__}); using local fallback {fallback}")
    if source == "synthetic":
        from dla_b200.data.synthetic import synthetic_records

        return [r["prompt"] for r in synthetic_records(("prompt",), int(cfg.get("num_samples", 64)), int(cfg.get("seed", 0)),
                                                       int(cfg.get("approx_chars", 64)))]
    path = cfg.get("prompt_path") or cfg.get("path")
    return [rec["prompt"] for rec in read_jsonl(path) if rec.get("prompt")]


def main() -> None:
    args = parse_args("RLHF PPO loop")
    cfg, ctx, plan = setup(args, "rlhf", default_seed=0)
    mc: Dict[str, Any] = cfg["model"]
    pc: Dict[str, Any] = cfg["ppo"]
    rc: Dict[str, Any] = cfg.get("reward_model", {}) or {}
    seed = int(cfg.get("seed", 0))
    algo = pc.get("algo", "reinforce")
    extra = [value_head_unit(resolve_model_config(mc["policy_model_name_or_path"]))] if algo == "ppo" else None
    policy = load_causal_lm(mc["policy_model_name_or_path"], ctx, plan, gradient_checkpointing=True, seed=seed, name="policy",
                            extra_units=extra)
    ref = load_causal_lm(mc["reference_model_name_or_path"], ctx, plan, gradient_checkpointing=False, trainable=False,
                         seed=seed, name="reference")
    reward_base = rc.get("base_model_name_or_path", mc["policy_model_name_or_path"])
    reward_model, reward_tok = build_reward_model(reward_base, ctx, plan, pooling="last_token", dropout=0.1, trainable=False,
                                                  gradient_checkpointing=False, seed=seed + 1, name="reward",
                                                  checkpoint_path=rc.get("path"))
    reward_model.eval()
    same_tok = type(reward_tok) is type(policy.tokenizer) and getattr(reward_tok, "vocab_size", None) == getattr(policy.tokenizer, "vocab_size", None)
    prompts = load_prompts(cfg["sampling"])
    gen_params = dict(pc.get("generation_params", {"max_new_tokens": 256}))
    gen_params.setdefault("do_sample", bool(pc.get("do_sample", False)))   # HF default without do_sample: greedy (D6)
    opt = AdamWConfig(lr=float(pc.get("learning_rate", 1e-6)), betas=(0.9, 0.95), max_grad_norm=float(pc.get("max_grad_norm", 1.0)))
    steps = int(pc.get("steps", 1024))
    batch_size = int(pc.get("batch_size", 64))
    kl_coef = float(pc.get("kl_coef", 0.1))
    max_len = int(mc.get("max_seq_length", 1024))
    lc = cfg.get("logging", {}) or {}
    loop = TrainLoop(cfg, "rlhf", ctx, plan, policy.engine, {"policy": policy.engine}, policy.tokenizer, opt, steps,
                     int(lc.get("log_every_steps", 10)), int(lc.get("save_every_steps", 0)), "logs/rlhf", resume=args.resume)
    generator = Generator(policy.model)
    vhead = ValueHead(policy.engine) if algo == "ppo" else None
    tok = policy.tokenizer
    pad_id = getattr(tok, "pad_token_id", 0) or 0
    sep_ids = tok("\n\n", add_special_tokens=False)["input_ids"] if same_tok else []
    sep_ids = list(sep_ids[0]) if sep_ids and isinstance(sep_ids[0], (list, tuple)) else list(sep_ids)
    fc = getattr(policy.engine, "fc", None) if (getattr(policy.engine, "fused", False) or getattr(policy.engine, "fused_ddp", False)) else None
    rng = random.Random(seed)

    while not loop.done():
        k = min(batch_size, len(prompts))
        if ctx.world_size > 1 and k >= ctx.world_size:
            k -= k % ctx.world_size        # equal shares per rank: the cross-rank gathers (fused kernel or NCCL) carry ONE row count
        batch_prompts = rng.sample(prompts, k=k)                                        # identical on all ranks
        local_prompts = split_between_processes(batch_prompts, ctx.world_size, ctx.rank)
        if not local_prompts:
            local_prompts = batch_prompts[:1]
        enc = tok(local_prompts, return_tensors="pt", padding=True, truncation=True, max_length=max_len).to(ctx.device)
        # ---- rollout (KV-cache decode, no grad) ------------------------------------------------
        out = generator.generate(enc["input_ids"], enc["attention_mask"], eos_token_id=getattr(tok, "eos_token_id", None),
                                 pad_token_id=pad_id, seed=seed + loop.global_step * 9973 + ctx.rank, **gen_params)
        seqs, amask = out.sequences, out.attention_mask
        # ---- reward ---------------------------------------------------------------------------
        with torch.no_grad():
            if same_tok:
                # the reward model scores "prompt\n\nresponse": splice the separator's ids in on the device
                r_ids, r_mask = splice_separator(out, sep_ids, pad_id)
                scores = reward_model(r_ids, r_mask).float()
            else:                                          # heterogeneous tokenizers: text round trip like the reference
                texts = [f"{p}\n\n{r}" for p, r in zip(local_prompts, decode_responses(tok, out))]
                renc = reward_tok(texts, return_tensors="pt", padding=True, truncation=True, max_length=max_len).to(ctx.device)
                scores = reward_model(renc["input_ids"], renc["attention_mask"]).float()
            ref_tok_logp = ref.model.token_logprobs(seqs, amask)
        info: Dict[str, Any] = {}
        if algo != "ppo":
            def loss_fn():
                pol = sequence_logprobs(policy.model, seqs, amask)
                with torch.no_grad():
                    refl = F.sequence_logprob(ref_tok_logp, amask[:, 1:])
                if fc is not None and pol.is_cuda:
                    # fused path (c): ONE kernel gathers (pol, ref, score) of every rank over peer memory and produces
                    # the global baseline, the global kl / reward / score means and the local gradients
                    from dla_b200.parallel.fused import fused_reinforce_loss

                    loss, st = fused_reinforce_loss(fc, pol, refl, scores, kl_coef)
                    info.update(kl=st[0], reward=st[1], score=st[2])
                    return loss
                with torch.no_grad():
                    rewards_local = scores - kl_coef * (pol.detach() - refl)
                    n_local = torch.tensor(float(rewards_local.numel()), device=rewards_local.device)
                    tot = ctx.all_reduce(torch.stack([rewards_local.sum().float(), n_local]), "sum")
                    baseline = tot[0] / tot[1]                                       # global baseline, any row counts
                loss, kl, rewards = reinforce_step(pol, refl, scores, kl_coef, baseline)
                glob = ctx.all_reduce(torch.stack([kl.detach().sum(), rewards.detach().sum(), scores.sum(), n_local]).float(), "sum")
                info.update(kl=glob[0] / glob[3], reward=glob[1] / glob[3], score=glob[2] / glob[3])
                return loss
            loss = loop.micro_step(loss_fn)
        else:
            loss = _ppo_update(loop, policy, vhead, seqs, amask, out.response_mask, ref_tok_logp, scores, pc, kl_coef, info, ctx, fc)
        extra_log = {"train/kl": float(info["kl"].item()), "train/reward_mean": float(info["reward"].item()),
                     "train/score_mean": float(info["score"].item())} if loop.global_step % loop.log_every == 0 else None
        loop.after_step(loss, extra_log)
    ctx.barrier()
    loop.finish("RLHF PPO loop complete")


def _ppo_update(loop, policy, vhead, seqs, amask, rmask, ref_tok_logp, scores, pc, kl_coef, info, ctx, fc=None):
    """Clipped PPO on the rollout batch: token rewards = -kl_coef * (logp - ref) with the sequence score on
    the last response token, GAE, ``ppo.epochs`` x mini-batches, stop when approx-KL exceeds ``target_kl``.

    Everything that decides control flow is GLOBAL: the advantages are whitened with the mean / variance over all
    ranks' tokens, kl / reward / score are logged as global means, every rank runs the same number of mini-batches
    (rows beyond the smallest local batch are dropped) and the ``target_kl`` stop uses the maximum approx-KL over ranks —
    each ``micro_step`` issues collectives, so a rank-local decision would desynchronise the job."""
    gamma, lam = float(pc.get("gamma", 1.0)), float(pc.get("lam", 0.95))
    clip, vclip, vf = float(pc.get("cliprange", 0.2)), float(pc.get("cliprange_value", 0.2)), float(pc.get("vf_coef", 0.5))
    epochs = int(pc.get("epochs", 1))
    target_kl = pc.get("target_kl")
    B = seqs.shape[0]
    if ctx.distributed:
        B = int(ctx.all_reduce(torch.tensor([B], device=seqs.device), "min").item())
        seqs, amask, rmask, ref_tok_logp, scores = seqs[:B], amask[:B], rmask[:B], ref_tok_logp[:B], scores[:B]
    mbs = max(1, min(int(pc.get("mini_batch_size", B)), B))
    m = rmask[:, 1:].float().contiguous()
    with torch.no_grad():
        h = policy.model.hidden(seqs, amask)
        old_values = vhead(h)[:, :-1]
        old_logp = policy.model.token_logprobs(seqs, amask)
        kl_tok = (old_logp - ref_tok_logp) * m
        rewards = -kl_coef * kl_tok
        last = (rmask.cumsum(1) * rmask).argmax(1).clamp(min=1) - 1
        rewards[torch.arange(B, device=seqs.device), last] += scores
        local = torch.stack([(kl_tok.sum(1) / m.sum(1).clamp(min=1)).sum(), rewards.sum(1).sum(), scores.sum(),
                             torch.tensor(float(B), device=seqs.device)]).float()
        if fc is not None and seqs.is_cuda:
            # fused path (c): GAE locally, then ONE kernel exchanges the whitening sums and the logged sums over peer
            # memory and normalises the local advantages with the global statistics
            adv, ret = F.gae(rewards, old_values * m, m, gamma, lam, whiten=False)
            adv = adv.contiguous()
            tot = fc.whiten_(adv, m, local)
        elif ctx.distributed:
            adv, ret = F.gae(rewards, old_values * m, m, gamma, lam, whiten=False)
            sums = ctx.all_reduce(torch.cat([torch.stack([(adv * m).sum(), (adv * adv * m).sum(), m.sum()]).float(), local]), "sum")
            mean = sums[0] / sums[2].clamp(min=1)
            var = (sums[1] / sums[2].clamp(min=1) - mean * mean).clamp(min=0)
            adv = (adv - mean) / var.sqrt().clamp(min=1e-6) * m
            tot = sums[3:]
        else:
            adv, ret = F.gae(rewards, old_values * m, m, gamma, lam, whiten=True)      # one launch: scan + whitening
            tot = local
    info.update(kl=tot[0] / tot[3], reward=tot[1] / tot[3], score=tot[2] / tot[3])
    loss = torch.zeros((), device=seqs.device)
    stop = False
    step0 = loop.global_step                     # one reference "step" per rollout batch
    for _ in range(epochs):
        perm = torch.randperm(B, device=seqs.device)
        for s in range(0, B, mbs):
            idx = perm[s:s + mbs]
            stats: Dict[str, Any] = {}

            def loss_fn():
                hh = policy.model.hidden(seqs[idx], amask[idx])
                values = vhead(hh)[:, :-1]
                Bm, T, H = hh.shape
                w, _ = policy.model._lm_head_weight()
                tgt = torch.cat([seqs[idx][:, 1:], torch.full_like(seqs[idx][:, :1], -100)], 1)
                new_logp = F.linear_logprob(hh.reshape(Bm * T, H), w, tgt.reshape(-1))[0].view(Bm, T)[:, :-1]
                l, st = ppo_losses(new_logp, old_logp[idx], adv[idx], values, old_values[idx], ret[idx], m[idx], clip, vclip, vf)
                stats.update(st)
                return l
            loss = loop.micro_step(loss_fn)
            if target_kl is not None:
                akl = ctx.all_reduce(stats["approx_kl"].detach().float().reshape(1).clone(), "max")
                if float(akl.item()) > float(target_kl):
                    stop = True
                    break
        if stop:
            break
    loop.global_step = step0 + 1       # the reference counts rollout batches; accumulation boundaries use loop.micro_steps
    return loss


if __name__ == "__main__":
    main()
