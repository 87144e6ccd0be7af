"""Reward-model training entrypoint.

Parity with ``R/src/training/train_reward.py``: ``model.{base_model_name_or_path, pooling(last_token),
dropout(0.1), max_seq_length(1024)}`` (``:64-71``); preference data; AdamW(lr, weight_decay(0)) (``:97-101``);
per micro-batch chosen/rejected scores -> ``pairwise_loss`` -> backward -> clip -> step (``:140-152``);
``evaluate`` = mean gathered loss + accuracy over the eval set (``:31-54``) every ``eval_every_steps`` (100);
log every 10, save every 200 + ``final`` (``:112-117,165-173``).  Also writes ``pytorch_model.bin`` so the PPO
phase and the rollout generator find the reward weights where they look (``R/src/training/train_rlhf.py:79-84``).
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from dla_b200.algo import reward_step
from dla_b200.cli.common import TrainLoop, parse_args, setup
from dla_b200.data import build_preference_dataset
from dla_b200.data.loader import DevicePrefetcher, build_loader, infinite_batches, to_device
from dla_b200.models.factory import build_reward_model
from dla_b200.optim import AdamWConfig


@torch.no_grad()
def evaluate(model, loader, ctx) -> Dict[str, float]:
    model.eval()
    correct = total = 0
    losses = []
    for batch in loader:
        batch = to_device(batch, ctx.device)
        loss, info = reward_step(model, batch)
        losses.append(float(ctx.gather(loss.float().reshape(1)).mean().item()))
        correct += int((info["chosen_scores"] > info["rejected_scores"]).sum().item())
        total += int(info["chosen_scores"].numel())
    model.train()
    # accuracy is summed over ranks (the reference reports the rank-local value)
    tot = torch.tensor([correct, total], dtype=torch.float64, device=ctx.device if ctx.backend == "nccl" else "cpu")
    ctx.all_reduce(tot)
    return {"loss": sum(losses) / max(len(losses), 1), "accuracy": float(tot[0] / tot[1].clamp(min=1))}


def main() -> None:
    args = parse_args("Train reward model")
    cfg, ctx, plan = setup(args, "reward", default_seed=0)
    mc: Dict[str, Any] = cfg["model"]
    oc: Dict[str, Any] = cfg["optimization"]
    lc: Dict[str, Any] = cfg.get("logging", {}) or {}
    seed = int(cfg.get("seed", 0))
    model, tokenizer = build_reward_model(mc["base_model_name_or_path"], ctx, plan, pooling=mc.get("pooling", "last_token"),
                                          dropout=float(mc.get("dropout", 0.1)), seed=seed, name="reward")
    data_cfg = dict(cfg["data"])
    data_cfg.setdefault("max_seq_length", mc.get("max_seq_length", 1024))
    train_ds = build_preference_dataset(data_cfg, tokenizer, split="train")
    loader, sampler = build_loader(train_ds, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=True,
                                   num_workers=int(data_cfg.get("num_workers", 4)), seed=seed)
    eval_loader = None
    if data_cfg.get("eval_path") or data_cfg.get("eval_split"):
        eval_ds = build_preference_dataset(data_cfg, tokenizer, split="eval")
        eval_loader, _ = build_loader(eval_ds, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=False,
                                      num_workers=int(data_cfg.get("num_workers", 4)), seed=seed)
    opt = AdamWConfig(lr=float(oc["learning_rate"]), weight_decay=float(oc.get("weight_decay", 0.0)),
                      max_grad_norm=float(oc.get("max_grad_norm", 1.0)))
    honor = bool(oc.get("honor_scheduler", False))
    eng = model.p
    loop = TrainLoop(cfg, "reward", ctx, plan, eng, {"reward": eng}, tokenizer, opt, int(oc["max_train_steps"]),
                     int(lc.get("log_every_steps", 10)), int(lc.get("save_every_steps", 200)), "logs/reward",
                     scheduler=oc.get("lr_scheduler", "constant") if honor else "constant",
                     warmup_steps=int(oc.get("warmup_steps", 0)) if honor else 0, resume=args.resume, reward_bin=True,
                     tokens_fn=lambda b: int(b["chosen"]["attention_mask"].sum() + b["rejected"]["attention_mask"].sum()))
    if sampler is not None and loop.sampler_state:
        sampler.load_state_dict(loop.sampler_state)
    eval_every = int(lc.get("eval_every_steps", 100))
    pad_id = getattr(tokenizer, "pad_token_id", 0) or 0
    concat = bool(data_cfg.get("concat_pairs", True))
    eng.expect_passes(1 if concat else 2)
    for epoch, batch in infinite_batches(DevicePrefetcher(loader, ctx.device), sampler):
        loop.epoch = epoch
        loss = loop.micro_step(lambda: reward_step(model, batch, concat=concat, pad_id=pad_id)[0], batch)
        loop.after_step(loss, sampler=sampler)
        if eval_loader is not None and eval_every > 0 and loop.global_step % eval_every == 0:
            m = evaluate(model, eval_loader, ctx)
            loop.logger.log({"eval/loss": m["loss"], "eval/acc": m["accuracy"]}, loop.global_step)
            ctx.print0(f"[reward] step {loop.global_step} eval/loss {m['loss']:.4f} eval/acc {m['accuracy']:.3f}")
        if loop.done():
            break
    loop.finish("Reward model training complete", sampler)


if __name__ == "__main__":
    main()
