#!/usr/bin/env python
"""Secondary BASELINE.json configs (device-timed, max over ranks; synthetic data, random-init weights):

  sft      Llama-3-8B SFT, 4096-token sequences (config 2): tokens/s of fwd + bwd + AdamW
  latency  prefill latency grid of the reference's eval (config 6): B in {1,4,8} x S in {256,512,1024}
  decode   KV-cache generation (PPO rollout / teacher sampling / eval): ms per token, own Generator (eager and
           CUDA-graph) vs transformers generate()
  ppo      Llama-3-8B RLHF step (config 4): KV-cache rollout tokens/s, scoring + update time

    python tools/bench_configs.py sft|latency|ppo [--model llama3-8b] [--impl own|reference]
    torchrun --nproc-per-node N tools/bench_configs.py sft
Appends JSON lines to gpurun_out/configs.jsonl (rank 0)."""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")

from dla_b200.config import HardwarePlan  # noqa: E402
from dla_b200.parallel.pg import init_distributed  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("what", choices=["sft", "latency", "ppo", "decode"])
ap.add_argument("--model", default="llama3-8b")
ap.add_argument("--impl", default="own", choices=["own", "reference"])
ap.add_argument("--steps", type=int, default=4)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--seq", type=int, default=4096)
ap.add_argument("--mb", type=int, default=2)
ap.add_argument("--new-tokens", type=int, default=64)
ap.add_argument("--algo", default="reinforce", choices=["reinforce", "ppo"],
                help="ppo config: the reference's sequence-level REINFORCE update, or clipped PPO with value head + GAE")
args = ap.parse_args()
ctx = init_distributed()
dev, W = ctx.device, ctx.world_size
plan = HardwarePlan(sharding="zero3" if W > 1 else "ddp")


def emit(rec):
    rec.update({"world": W, "model": args.model, "impl": args.impl})
    if ctx.is_main:
        print(json.dumps(rec), flush=True)
        (ROOT / "gpurun_out").mkdir(exist_ok=True)
        with open(ROOT / "gpurun_out" / "configs.jsonl", "a") as fh:
            fh.write(json.dumps(rec) + "\n")


def timed(fn, steps, warm):
    for _ in range(warm):
        fn()
    ctx.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    ctx.barrier()
    torch.cuda.synchronize()
    return ctx.max_over_ranks(e0.elapsed_time(e1) / steps)


def hf_model(train: bool):
    sys.path.insert(0, str(ROOT / "baseline"))
    from reference_arm import llama_hf_config
    from transformers import LlamaForCausalLM

    with torch.device(dev):
        m = LlamaForCausalLM(llama_hf_config(args.model)).to(torch.bfloat16)
    if train:
        m.gradient_checkpointing_enable()
        m.train()
    else:
        m.eval()
    return m


if args.what == "sft":
    from dla_b200.data.synthetic import synthetic_lm_batch

    T, MB = args.seq, args.mb
    if args.impl == "own":
        from dla_b200.algo import sft_loss
        from dla_b200.models.factory import load_causal_lm
        from dla_b200.optim import AdamWConfig

        b = load_causal_lm(args.model, ctx, plan, name="sft")
        V = b.config.vocab_size
        batch = {k: v.to(dev) for k, v in synthetic_lm_batch(MB, T, V, seed=ctx.rank).items()}
        opt = AdamWConfig(lr=2e-5, betas=(0.9, 0.95), weight_decay=0.01)
        first = [True]

        def step():
            loss = sft_loss(b.model, batch)
            loss.backward()
            b.engine.finish_backward()
            b.engine.optimizer_step(opt, grad_divisor=float(W))
            if first[0]:
                first[0] = False
                b.engine.auto_no_recompute(MB * T)
    else:
        m = hf_model(True)
        if W > 1:
            from torch.distributed.fsdp import fully_shard

            for layer in m.model.layers:
                fully_shard(layer)
            fully_shard(m)
        V = m.config.vocab_size
        batch = {k: v.to(dev) for k, v in synthetic_lm_batch(MB, T, V, seed=ctx.rank).items()}
        o = torch.optim.AdamW(m.parameters(), lr=2e-5, betas=(0.9, 0.95), weight_decay=0.01)

        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = m(**batch).loss            # R/src/training/train_sft.py:145-151
            loss.backward()
            torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
            o.step()
            o.zero_grad()

    ms = timed(step, args.steps, args.warmup)
    emit({"config": "sft_4k", "seq_len": T, "micro_batch": MB, "ms_per_step": ms, "tokens_per_s": MB * T * W / ms * 1e3,
          "peak_hbm_gb": torch.cuda.max_memory_allocated(dev) / (1 << 30)})

elif args.what == "latency":
    if args.impl == "own":
        from dla_b200.models.factory import load_causal_lm

        from dla_b200.rollout import GraphedForward

        b = load_causal_lm(args.model, ctx, trainable=False, gradient_checkpointing=False, name="lat")
        V = b.config.vocab_size
        graphs = {}

        def fwd(ids, mask):          # one captured graph per [B, S], as dla_b200.cli.eval_latency does
            key = tuple(ids.shape)
            if key not in graphs:
                graphs[key] = GraphedForward(b.model, *key)
            return graphs[key](ids, mask)
    else:
        m = hf_model(False)
        V = m.config.vocab_size
        fwd = lambda ids, mask: m(input_ids=ids, attention_mask=mask).logits
    with torch.no_grad():
        for B in (1, 4, 8):
            for S in (256, 512, 1024):
                ids = torch.randint(0, V - 1, (B, S), device=dev)
                mask = torch.ones_like(ids)
                ms = timed(lambda: fwd(ids, mask), 20, 5)          # R/config/eval_config.yaml:28-29
                emit({"config": "prefill_latency", "batch_size": B, "seq_length": S, "latency_ms": ms,
                      "tokens_per_second": B * S / ms * 1e3})

elif args.what == "decode":
    B, Tp, Tn = 8, 256, args.new_tokens
    if args.impl == "own":
        from dla_b200.models.factory import load_causal_lm
        from dla_b200.rollout import Generator

        b = load_causal_lm(args.model, ctx, trainable=False, gradient_checkpointing=False, name="gen")
        V = b.config.vocab_size
        ids = torch.randint(4, V, (B, Tp), device=dev)
        variants = {"graph": Generator(b.model), "eager": Generator(b.model, use_cuda_graph=False)}
        runs = {k: (lambda g=g, n=Tn: g.generate(ids, torch.ones_like(ids), max_new_tokens=n, do_sample=True, temperature=0.7,
                                                 top_p=0.9, seed=1, eos_token_id=-1, pad_token_id=0)) for k, g in variants.items()}
        pre = lambda: variants["eager"].generate(ids, torch.ones_like(ids), max_new_tokens=1, do_sample=True, temperature=0.7,
                                                 top_p=0.9, seed=1, eos_token_id=-1, pad_token_id=0)
    else:
        m = hf_model(False)
        V = m.config.vocab_size
        ids = torch.randint(4, V, (B, Tp), device=dev)
        kw = dict(do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=0, eos_token_id=None)   # R/config/rlhf_config.yaml:19-24
        runs = {"hf_generate": lambda: m.generate(input_ids=ids, attention_mask=torch.ones_like(ids), max_new_tokens=Tn,
                                                  min_new_tokens=Tn, **kw)}
        pre = lambda: m.generate(input_ids=ids, attention_mask=torch.ones_like(ids), max_new_tokens=1, min_new_tokens=1, **kw)
    with torch.no_grad():
        t_pre = timed(pre, 3, 2)
        for name, fn in runs.items():
            ms = timed(fn, 3, 2)
            emit({"config": "decode", "variant": name, "batch": B, "prompt_len": Tp, "new_tokens": Tn, "total_ms": ms, "prefill_ms": t_pre,
                  "ms_per_token": (ms - t_pre) / max(Tn - 1, 1), "tokens_per_s": B * Tn * W / ms * 1e3})

elif args.what == "ppo":
    from dla_b200.algo import reinforce_step, sequence_logprobs
    from dla_b200.models.factory import build_reward_model, load_causal_lm
    from dla_b200.ops import functional as F
    from dla_b200.optim import AdamWConfig
    from dla_b200.rollout import Generator

    from dla_b200.models.config import resolve_model_config

    extra = None
    if args.algo == "ppo":
        from dla_b200.models.reward import value_head_unit

        extra = [value_head_unit(resolve_model_config(args.model))]
    pol = load_causal_lm(args.model, ctx, plan, name="policy", extra_units=extra)
    ref = load_causal_lm(args.model, ctx, plan, trainable=False, gradient_checkpointing=False, name="reference")
    rm, _ = build_reward_model(args.model, ctx, plan, trainable=False, gradient_checkpointing=False, name="reward")
    rm.eval()
    V = pol.config.vocab_size
    B, Tp, Tn = 8, 256, args.new_tokens                      # prompts per rank (ppo.batch_size 64 / 8 ranks)
    ids = torch.randint(4, V, (B, Tp), device=dev)
    gen = Generator(pol.model)
    opt = AdamWConfig(lr=1e-6, betas=(0.9, 0.95))
    t = {}
    eng = pol.engine
    fc = getattr(eng, "fc", None) if (getattr(eng, "fused", False) or getattr(eng, "fused_ddp", False)) else None

    class _Loop:            # what cli.train_rlhf._ppo_update needs from TrainLoop
        global_step = 0

        def micro_step(self, fn):
            loss = fn()
            loss.backward()
            eng.finish_backward()
            eng.optimizer_step(opt, grad_divisor=float(W))
            self.global_step += 1
            return loss.detach()

    loop = _Loop()
    if args.algo == "ppo":
        from dla_b200.cli.train_rlhf import _ppo_update
        from dla_b200.models.reward import ValueHead

        vhead = ValueHead(eng)
        pc = {"epochs": 1, "mini_batch_size": 4, "target_kl": 0.1, "gamma": 1.0, "lam": 0.95, "cliprange": 0.2,
              "cliprange_value": 0.2, "vf_coef": 0.5}       # R/config/rlhf_config.yaml:12-22 (mini_batch 32 / 8 ranks)

    def step_ppo():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = gen.generate(ids, torch.ones_like(ids), max_new_tokens=Tn, do_sample=True, temperature=0.7, top_p=0.9, seed=1,
                           eos_token_id=-1, pad_token_id=0)
        torch.cuda.synchronize()
        t["rollout"] = time.perf_counter() - t0
        with torch.no_grad():
            scores = rm(out.sequences, out.attention_mask).float()
            ref_tok = ref.model.token_logprobs(out.sequences, out.attention_mask)
        info = {}
        _ppo_update(loop, pol, vhead, out.sequences, out.attention_mask, out.response_mask, ref_tok, scores, pc, 0.1, info, ctx, fc)
        torch.cuda.synchronize()
        t["total"] = time.perf_counter() - t0

    def step():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = gen.generate(ids, torch.ones_like(ids), max_new_tokens=Tn, do_sample=True, temperature=0.7, top_p=0.9, seed=1,
                           eos_token_id=-1, pad_token_id=0)
        torch.cuda.synchronize()
        t["rollout"] = time.perf_counter() - t0
        with torch.no_grad():
            scores = rm(out.sequences, out.attention_mask).float()
            ref_lp = sequence_logprobs(ref.model, out.sequences, out.attention_mask)
        p_lp = sequence_logprobs(pol.model, out.sequences, out.attention_mask)
        if fc is not None:      # fused path (c): cross-rank gather + global baseline + loss in one peer-memory kernel
            from dla_b200.parallel.fused import fused_reinforce_loss

            loss, _ = fused_reinforce_loss(fc, p_lp, ref_lp, scores, 0.1)
        else:
            loss, kl, rw = reinforce_step(p_lp, ref_lp, scores, 0.1)
        loss.backward()
        pol.engine.finish_backward()
        pol.engine.optimizer_step(opt, grad_divisor=float(W))
        torch.cuda.synchronize()
        t["total"] = time.perf_counter() - t0

    if args.algo == "ppo":
        step = step_ppo
    for _ in range(args.warmup):
        step()
    rs, ts = [], []
    for _ in range(args.steps):
        step()
        rs.append(t["rollout"])
        ts.append(t["total"])
    r, tot = ctx.max_over_ranks(sum(rs) / len(rs)), ctx.max_over_ranks(sum(ts) / len(ts))
    emit({"config": "ppo_clipped_value_gae" if args.algo == "ppo" else "ppo_reinforce", "prompts_per_rank": B,
          "fused_gather": fc is not None, "updates_per_step": (B // 4) if args.algo == "ppo" else 1, "prompt_len": Tp, "new_tokens": Tn, "rollout_s": r,
          "rollout_tokens_per_s": B * Tn * W / r, "decode_ms_per_token": r / Tn * 1e3, "score_and_update_s": tot - r, "step_s": tot,
          "peak_hbm_gb": torch.cuda.max_memory_allocated(dev) / (1 << 30)})
ctx.barrier()
