"""Distillation entrypoint (teacher rollouts -> student).

Parity with ``R/src/training/train_distill.py``: student with gradient checkpointing (``:40-44``); when
``distill.use_kl and distill.on_policy`` load the teacher(s) from ``distill.teacher_model_names_or_paths`` or the
single ``distill.teacher_model_name_or_path`` / ``model.teacher_path`` (``:46-61``), frozen; dataset =
``TeacherRolloutDataset(data.teacher_samples_path)`` with ``max_seq_length`` 2048 (``:62-66``);
AdamW(lr, weight_decay(0)) (``:78-82``); KL mode = forward KL against the mean teacher distribution, masked
mean over tokens (``:127-144``), otherwise CE on the teacher tokens (``:145-147``); log ``train/loss`` and
``train/reward_mean`` every 20 steps, save every 400 + ``final`` (``:99-101,156-167``).  ``optimization.temperature``
(declared, unread in the reference, ``R/config/distill_config.yaml:34``) scales both distributions.
"""
from __future__ import annotations

from typing import Any, Dict, List

from dla_b200.algo import distill_step
from dla_b200.cli.common import TrainLoop, parse_args, setup
from dla_b200.data import TeacherRolloutDataset
from dla_b200.data.loader import DevicePrefetcher, build_loader, infinite_batches
from dla_b200.models.factory import load_causal_lm
from dla_b200.optim import AdamWConfig


def main() -> None:
    args = parse_args("Distill aligned teacher into student")
    cfg, ctx, plan = setup(args, "distill", default_seed=0)
    mc: Dict[str, Any] = cfg["model"]
    dc: Dict[str, Any] = cfg.get("distill", {}) or {}
    oc: Dict[str, Any] = cfg["optimization"]
    lc: Dict[str, Any] = cfg.get("logging", {}) or {}
    seed = int(cfg.get("seed", 0))
    student = load_causal_lm(mc["student_model_name_or_path"], ctx, plan, gradient_checkpointing=True,
                             use_flash_attention=mc.get("use_flash_attention", False), seed=seed, name="student")
    use_kl, on_policy = bool(dc.get("use_kl", False)), bool(dc.get("on_policy", False))
    teachers: List[Any] = []
    if use_kl and on_policy:
        paths = list(dc.get("teacher_model_names_or_paths") or [])
        if not paths:
            single = dc.get("teacher_model_name_or_path") or mc.get("teacher_path")
            if single:
                paths = [single]
        if not paths:
            raise ValueError("KL distillation requested but no teacher model path provided")
        for i, p in enumerate(paths):
            teachers.append(load_causal_lm(p, ctx, plan, gradient_checkpointing=False, trainable=False, seed=seed + 17 + i,
                                           name=f"teacher{i}"))
    data_cfg = cfg["data"]
    dataset = TeacherRolloutDataset(data_cfg.get("teacher_samples_path"), student.tokenizer,
                                    max_length=int(mc.get("max_seq_length", 2048)),
                                    records=_synthetic(data_cfg) if data_cfg.get("source") == "synthetic" else None)
    dataset.pad_to_multiple_of = int(data_cfg.get("pad_to_multiple_of", 1))
    loader, sampler = build_loader(dataset, int(oc["micro_batch_size"]), ctx.world_size, ctx.rank, shuffle=True,
                                   num_workers=int(data_cfg.get("num_workers", 4)), seed=seed)
    opt = AdamWConfig(lr=float(oc["learning_rate"]), weight_decay=float(oc.get("weight_decay", 0.0)),
                      max_grad_norm=float(oc.get("max_grad_norm", 1.0)))
    honor = bool(oc.get("honor_scheduler", False))
    engines = {"student": student.engine, **{t.engine.name: t.engine for t in teachers}}
    loop = TrainLoop(cfg, "distill", ctx, plan, student.engine, engines, student.tokenizer, opt, int(oc["max_train_steps"]),
                     int(lc.get("log_every_steps", 20)), int(lc.get("save_every_steps", 400)), "logs/distill",
                     scheduler=oc.get("lr_scheduler", "constant") if honor else "constant",
                     warmup_steps=int(oc.get("warmup_steps", 0)) if honor else 0, resume=args.resume,
                     tokens_fn=lambda b: int(b["attention_mask"].sum()))
    if sampler is not None and loop.sampler_state:
        sampler.load_state_dict(loop.sampler_state)
    temperature = float(oc.get("temperature", 1.0) or 1.0)
    tmodels = [t.model for t in teachers]
    for epoch, batch in infinite_batches(DevicePrefetcher(loader, ctx.device), sampler):
        loop.epoch = epoch
        reward_values = batch.pop("reward")
        loss = loop.micro_step(lambda: distill_step(student.model, tmodels, batch, use_kl and on_policy, temperature), batch)
        extra = {"train/reward_mean": float(reward_values.float().mean().item())} if loop.global_step % loop.log_every == 0 else None
        loop.after_step(loss, extra, sampler=sampler)
        if loop.done():
            break
    loop.finish("Distillation complete", sampler)


def _synthetic(data_cfg: Dict[str, Any]):
    from dla_b200.data.synthetic import synthetic_records

    rows = synthetic_records(("prompt", "teacher_response", "reward"), int(data_cfg.get("num_samples", 64)),
                             int(data_cfg.get("seed", 0)), int(data_cfg.get("approx_chars", 64)))
    return rows


if __name__ == "__main__":
    main()
