"""Prefill latency / throughput grid.

Parity with ``R/src/eval/eval_latency.py``: for every model x batch x seq: random ids ``[B, S]``, ``warmup_steps``
(3) then ``measure_steps`` (10) full forwards under ``no_grad`` (prefill only, no decode — defect D15 kept as
the metric definition), ``tokens_per_second = B*S*steps/dt`` and ``latency_ms = dt/steps*1e3`` (``:38-62``),
written to ``latency.json`` next to ``logging.output_path`` (``:81-84``).  Timing here is CUDA events on the
launching stream (the reference uses perf_counter around synchronize); ``decode_latency_ms`` per generated
token is reported as an extra when ``latency.decode_tokens`` is set.
"""
from __future__ import annotations

import argparse
import json
import time
from pathlib import Path
from typing import Any, Dict, List

import torch

from dla_b200.config import load_config
from dla_b200.models.factory import load_causal_lm
from dla_b200.parallel.pg import init_distributed
from dla_b200.rollout import Generator, GraphedForward


def parse_args() -> argparse.Namespace:
    p = argparse.ArgumentParser(description="Latency eval")
    p.add_argument("--config", required=True)
    return p.parse_args()


@torch.no_grad()
def measure_model(bundle, batch_sizes: List[int], seq_lengths: List[int], warmup_steps: int, measure_steps: int,
                  decode_tokens: int = 0, use_graph: bool = True) -> List[Dict[str, float]]:
    model, dev, V = bundle.model, bundle.engine.device, bundle.config.vocab_size
    rows: List[Dict[str, float]] = []
    for batch in batch_sizes:
        for seq in seq_lengths:
            ids = torch.randint(0, V - 1, (batch, seq), device=dev)
            mask = torch.ones_like(ids)
            fwd = model.logits
            if use_graph and dev.type == "cuda" and getattr(bundle.engine, "resident", True):
                fwd = GraphedForward(model, batch, seq)          # the grid repeats one shape: replay instead of ~330 launches
            for _ in range(warmup_steps):
                fwd(ids, mask)
            if dev.type == "cuda":
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize(dev)
                e0.record()
                for _ in range(measure_steps):
                    fwd(ids, mask)
                e1.record()
                torch.cuda.synchronize(dev)
                dt = e0.elapsed_time(e1) / 1e3
            else:
                t0 = time.perf_counter()
                for _ in range(measure_steps):
                    fwd(ids, mask)
                dt = time.perf_counter() - t0
            row = {"batch_size": batch, "seq_length": seq, "tokens_per_second": batch * seq * measure_steps / dt,
                   "latency_ms": dt / measure_steps * 1e3}
            if decode_tokens > 0:
                gen = Generator(model)
                t0 = time.perf_counter()
                gen.generate(ids, mask, max_new_tokens=decode_tokens, do_sample=False, eos_token_id=-1)
                if dev.type == "cuda":
                    torch.cuda.synchronize(dev)
                row["decode_latency_ms"] = (time.perf_counter() - t0) * 1e3 / decode_tokens
            rows.append(row)
    return rows


def main() -> None:
    args = parse_args()
    config = load_config(args.config)
    ctx = init_distributed()
    lat = config["latency"]
    results: Dict[str, Any] = {}
    for name, path in config["models"].items():
        bundle = load_causal_lm(path, ctx, gradient_checkpointing=False, trainable=False, name=name)
        results[name] = measure_model(bundle, lat["batch_sizes"], lat["seq_lengths"], lat.get("warmup_steps", 3),
                                      lat.get("measure_steps", 10), int(lat.get("decode_tokens", 0)),
                                      bool(lat.get("cuda_graph", True)))
        del bundle
    if ctx.is_main:
        out = Path((config.get("logging", {}) or {}).get("output_path", "logs/eval/results.json")).with_name("latency.json")
        out.parent.mkdir(parents=True, exist_ok=True)
        out.write_text(json.dumps(results, indent=2))


if __name__ == "__main__":
    main()
