"""Offline teacher rollouts (+ optional reward scoring) -> JSONL for the distillation phase.

Parity with ``R/src/training/generate_teacher_data.py``: flags ``--teacher --prompts --output --reward_model
--batch_size(4) --max_new_tokens(256) --temperature(0.7) --top_p(0.9)`` (``:19-26``); prompts batched
(``chunk_list``), sampled with ``do_sample=True`` (``:73-79``), the response is everything after the prompt
(``:80,84``), optional reward of ``"{prompt}\\n\\n{response}"`` (``:87-100``), records ``{prompt, teacher_response
[, reward]}`` (``:101-107``).  Here rewards are scored per batch, not one sample at a time.
"""
from __future__ import annotations

import argparse
import json
from pathlib import Path

import torch

from dla_b200.data import read_jsonl
from dla_b200.models.factory import build_reward_model, load_causal_lm
from dla_b200.parallel.pg import init_distributed
from dla_b200.rollout import Generator, decode_responses
from dla_b200.utils import chunk_list


def parse_args() -> argparse.Namespace:
    p = argparse.ArgumentParser(description="Sample teacher responses")
    p.add_argument("--teacher", required=True, help="checkpoint dir or preset name of the teacher model")
    p.add_argument("--prompts", required=True, help="JSONL prompts file")
    p.add_argument("--output", required=True, help="Destination JSONL")
    p.add_argument("--reward_model", default=None, help="Optional reward model checkpoint")
    p.add_argument("--batch_size", type=int, default=4)
    p.add_argument("--max_new_tokens", type=int, default=256)
    p.add_argument("--temperature", type=float, default=0.7)
    p.add_argument("--top_p", type=float, default=0.9)
    p.add_argument("--seed", type=int, default=0)
    return p.parse_args()


def main() -> None:
    args = parse_args()
    ctx = init_distributed()
    teacher = load_causal_lm(args.teacher, ctx, gradient_checkpointing=False, trainable=False, name="teacher")
    tok = teacher.tokenizer
    gen = Generator(teacher.model)
    reward_model = reward_tok = None
    if args.reward_model:
        reward_model, reward_tok = build_reward_model(args.reward_model, ctx, trainable=False, gradient_checkpointing=False,
                                                      checkpoint_path=args.reward_model)
        reward_model.eval()
    prompts = [item["prompt"] for item in read_jsonl(args.prompts)]
    Path(args.output).parent.mkdir(parents=True, exist_ok=True)
    try:
        from tqdm import tqdm
        batches = tqdm(chunk_list(prompts, args.batch_size), desc="Generating")
    except Exception:  # pragma: no cover
        batches = chunk_list(prompts, args.batch_size)
    with Path(args.output).open("w", encoding="utf-8") as writer:
        for bi, batch_prompts in enumerate(batches):
            enc = tok(batch_prompts, return_tensors="pt", padding=True, truncation=True,
                      max_length=min(getattr(tok, "model_max_length", 4096), teacher.config.max_position_embeddings)).to(ctx.device)
            out = gen.generate(enc["input_ids"], enc["attention_mask"], max_new_tokens=args.max_new_tokens,
                               temperature=args.temperature, top_p=args.top_p, do_sample=True,
                               eos_token_id=getattr(tok, "eos_token_id", None), pad_token_id=getattr(tok, "pad_token_id", 0) or 0,
                               seed=args.seed + bi)
            responses = decode_responses(tok, out)
            rewards = [None] * len(responses)
            if reward_model is not None:
                fused = [f"{p}\n\n{r}" for p, r in zip(batch_prompts, responses)]
                renc = reward_tok(fused, return_tensors="pt", truncation=True, padding=True,
                                  max_length=min(getattr(reward_tok, "model_max_length", 4096), 4096)).to(ctx.device)
                with torch.no_grad():
                    rewards = reward_model(renc["input_ids"], renc["attention_mask"]).float().tolist()
            for p, r, rw in zip(batch_prompts, responses, rewards):
                rec = {"prompt": p, "teacher_response": r}
                if rw is not None:
                    rec["reward"] = float(rw)
                writer.write(json.dumps(rec) + "\n")


if __name__ == "__main__":
    main()
