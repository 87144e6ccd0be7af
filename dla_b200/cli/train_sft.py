"""Supervised fine-tuning entrypoint.

Parity with ``R/src/training/train_sft.py``: seed default 42 (``:48``); ``model.{model_name_or_path,
gradient_checkpointing(True), use_flash_attention, max_seq_length(2048)}`` (``:53-60``); instruction dataset with
prompt masking; AdamW(lr, weight_decay(0), betas=(0.9, 0.95)) (``:89-94``); scheduler ``lr_scheduler``
(default cosine) with ``warmup_steps`` stepped once per micro-batch (``:105-110,150``); eval loss over
``eval_path``/``eval_split`` every ``eval_every_steps`` with a cross-rank gather (``:33-42,160-162``);
trainable-parameter report (``:120-123``); checkpoints every ``save_every_steps`` + ``final``.
``data.packing: true`` (declared, never read by the reference, ``R/config/sft_config.yaml:16``) packs sequences.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from dla_b200.algo import sft_loss
from dla_b200.cli.common import TrainLoop, parse_args, setup
from dla_b200.data import PackedCollator, build_instruction_dataset
from dla_b200.data.loader import DevicePrefetcher, build_loader, infinite_batches, to_device
from dla_b200.models.factory import count_trainable_params, load_causal_lm
from dla_b200.optim import AdamWConfig


@torch.no_grad()
def evaluate_loss(model, loader, ctx) -> float:
    """Mean over batches of the rank-gathered mean batch loss (``R/src/training/train_sft.py:33-42``)."""
    losses = []
    for batch in loader:
        batch = to_device(batch, ctx.device)
        loss = sft_loss(model, batch)
        gathered = ctx.gather(loss.float().reshape(1).repeat(batch["input_ids"].shape[0]))
        losses.append(float(gathered.mean().item()))
    return float(sum(losses) / max(len(losses), 1))


def main() -> None:
    args = parse_args("Distributed SFT training")
    cfg, ctx, plan = setup(args, "sft", default_seed=42)
    mc: Dict[str, Any] = cfg["model"]
    oc: Dict[str, Any] = cfg["optimization"]
    lc: Dict[str, Any] = cfg.get("logging", {}) or {}
    seed = int(cfg.get("seed", 42))
    bundle = load_causal_lm(mc["model_name_or_path"], ctx, plan, gradient_checkpointing=mc.get("gradient_checkpointing", True),
                            use_flash_attention=mc.get("use_flash_attention", False), seed=seed, name="model")
    data_cfg = dict(cfg["data"])
    data_cfg.setdefault("max_seq_length", mc.get("max_seq_length", 2048))
    max_len = int(data_cfg.get("max_length") or data_cfg["max_seq_length"])
    train_ds = build_instruction_dataset(data_cfg, bundle.tokenizer, split="train")
    # ``data.packing: true`` (R/config/sft_config.yaml:16): several documents per row; the flash kernels mask (and skip
    # tiles) by document through ``segment_ids``, positions restart per document (``position_ids``)
    collate = PackedCollator(max_len, train_ds.pad_id) if data_cfg.get("packing") and data_cfg.get("honor_packing", True) else None
    loader, sampler = build_loader(train_ds, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=True,
                                   num_workers=int(data_cfg.get("num_workers", 4)), seed=seed, collate_fn=collate)
    eval_loader = None
    if data_cfg.get("eval_path") or data_cfg.get("eval_split"):
        eval_ds = build_instruction_dataset(data_cfg, bundle.tokenizer, split="eval")
        eval_loader, _ = build_loader(eval_ds, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=False,
                                      num_workers=int(data_cfg.get("num_workers", 4)), seed=seed)
    opt = AdamWConfig(lr=float(oc["learning_rate"]), betas=tuple(oc.get("betas", (0.9, 0.95))),
                      weight_decay=float(oc.get("weight_decay", 0.0)), max_grad_norm=float(oc.get("max_grad_norm", 1.0)))
    loop = TrainLoop(cfg, "sft", ctx, plan, bundle.engine, {"model": bundle.engine}, bundle.tokenizer, opt,
                     int(oc["max_train_steps"]), int(lc.get("log_every_steps", 20)), int(lc.get("save_every_steps", 500)),
                     "logs/sft", scheduler=oc.get("lr_scheduler", "cosine"), warmup_steps=int(oc.get("warmup_steps", 0)),
                     resume=args.resume, tokens_fn=lambda b: int(b["attention_mask"].sum()),
                     flops_per_token=bundle.config.flops_per_token(max_len, True))
    if sampler is not None and loop.sampler_state:
        sampler.load_state_dict(loop.sampler_state)
    ctx.print0(f"Trainable parameters: {count_trainable_params(bundle.engine)}")
    eval_every = int(lc.get("eval_every_steps", 200))
    for epoch, batch in infinite_batches(DevicePrefetcher(loader, ctx.device), sampler):
        loop.epoch = epoch
        loss = loop.micro_step(lambda: sft_loss(bundle.model, batch), batch)
        loop.after_step(loss, sampler=sampler)
        if eval_loader is not None and eval_every > 0 and loop.global_step % eval_every == 0:
            ev = evaluate_loss(bundle.model, eval_loader, ctx)
            loop.logger.log({"eval/loss": ev}, loop.global_step)
            ctx.print0(f"[sft] step {loop.global_step} eval/loss {ev:.4f}")
        if loop.done():
            break
    loop.finish("Training complete", sampler)


if __name__ == "__main__":
    main()
