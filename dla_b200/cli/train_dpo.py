"""Direct Preference Optimization entrypoint (``python -m dla_b200.cli.train_dpo --config …``).

Parity with ``R/src/training/train_dpo.py``: seed default 0 (``:50``); policy with gradient checkpointing,
frozen reference without (``:54-58``); preference dataset with ``max_seq_length`` default 1024 (``:60-61``);
AdamW(lr, weight_decay=0.01) and no scheduler (``:73-77``); per micro-batch: four sequence log-probs ->
``dpo_loss`` -> backward -> clip -> step (``:110-121``); ``train/loss`` every ``log_every_steps`` (10),
``train/preference_rate`` every ``eval_every_steps`` (100), checkpoints every ``save_every_steps`` (200)
plus ``final`` (``:85-142``).  Honoured extras (declared but unread in the reference):
``model.label_smoothing``, ``optimization.lr_scheduler`` (with ``honor_scheduler``), keep-last-N, resume.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from dla_b200.algo import dpo_step
from dla_b200.cli.common import TrainLoop, parse_args, setup
from dla_b200.data import build_preference_dataset
from dla_b200.data.loader import DevicePrefetcher, build_loader, infinite_batches
from dla_b200.models.factory import load_causal_lm
from dla_b200.optim import AdamWConfig


def pair_tokens(batch: Dict[str, Dict[str, torch.Tensor]]) -> int:
    return int(batch["chosen"]["attention_mask"].sum() + batch["rejected"]["attention_mask"].sum())


def main() -> None:
    args = parse_args("DPO training")
    cfg, ctx, plan = setup(args, "dpo", default_seed=0)
    mc: Dict[str, Any] = cfg["model"]
    oc: Dict[str, Any] = cfg["optimization"]
    lc: Dict[str, Any] = cfg.get("logging", {}) or {}
    seed = int(cfg.get("seed", 0))

    policy = load_causal_lm(mc["policy_model_name_or_path"], ctx, plan, gradient_checkpointing=True, seed=seed, name="policy")
    ref = load_causal_lm(mc["reference_model_name_or_path"], ctx, plan, gradient_checkpointing=False, trainable=False,
                         seed=seed, name="reference")

    data_cfg = dict(cfg["data"])
    data_cfg.setdefault("max_seq_length", mc.get("max_seq_length", 1024))
    dataset = build_preference_dataset(data_cfg, policy.tokenizer, split="train")
    loader, sampler = build_loader(dataset, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=True,
                                   num_workers=int(data_cfg.get("num_workers", 4)), seed=seed)

    honor = bool(oc.get("honor_scheduler", False))
    opt = AdamWConfig(lr=float(oc["learning_rate"]), betas=tuple(oc.get("betas", (0.9, 0.999))),
                      weight_decay=float(oc.get("weight_decay", 0.01)), max_grad_norm=float(oc.get("max_grad_norm", 1.0)))
    concat = bool(data_cfg.get("concat_pairs", True))
    policy.engine.expect_passes(1 if concat else 2)
    loop = TrainLoop(cfg, "dpo", ctx, plan, policy.engine, {"policy": policy.engine, "reference": ref.engine},
                     policy.tokenizer, opt, int(oc["max_train_steps"]), int(lc.get("log_every_steps", 10)),
                     int(lc.get("save_every_steps", 200)), "logs/dpo",
                     scheduler=oc.get("lr_scheduler", "constant") if honor else "constant",
                     warmup_steps=int(oc.get("warmup_steps", 0)) if honor else 0, resume=args.resume,
                     tokens_fn=pair_tokens, flops_per_token=policy.config.flops_per_token(int(data_cfg["max_seq_length"]), True))
    if sampler is not None and loop.sampler_state:
        sampler.load_state_dict(loop.sampler_state)
    beta = float(mc.get("beta", 0.1))
    ls = float(mc.get("label_smoothing", 0.0) or 0.0)
    eval_every = int(lc.get("eval_every_steps", 100))
    reduction = mc.get("logprob_reduction", "mean")
    pad_id = getattr(policy.tokenizer, "pad_token_id", 0) or 0

    for epoch, batch in infinite_batches(DevicePrefetcher(loader, ctx.device), sampler):
        loop.epoch = epoch
        info: Dict[str, Any] = {}

        def loss_fn():
            loss, inf = dpo_step(policy.model, ref.model, batch, beta, ls, concat=concat, reduction=reduction, pad_id=pad_id)
            info.update(inf)
            return loss

        loss = loop.micro_step(loss_fn, batch)
        extra = None
        if eval_every > 0 and loop.global_step % eval_every == 0:
            extra = {"train/preference_rate": float(info["preference_rate"].item()), "train/margin": float(info["margin"].item())}
            loop.logger.log(extra, loop.global_step)
        loop.after_step(loss, sampler=sampler)
        if loop.done():
            break
    loop.finish("DPO training complete", sampler)


if __name__ == "__main__":
    main()
