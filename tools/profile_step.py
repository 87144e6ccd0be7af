#!/usr/bin/env python
"""Per-kernel time breakdown of ONE Llama-3-8B DPO step (CUPTI via torch.profiler; numbers taken under the
profiler are for SHARES only, never bench values).  Writes gpurun_out/step_breakdown.json."""
import json
import os
import sys
from collections import defaultdict
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")

from dla_b200.algo import dpo_step  # noqa: E402
from dla_b200.data.synthetic import synthetic_preference_batch  # noqa: E402
from dla_b200.models.factory import load_causal_lm  # noqa: E402
from dla_b200.optim import AdamWConfig  # noqa: E402
from dla_b200.parallel.pg import init_distributed  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b"
ctx = init_distributed()
policy = load_causal_lm(model, ctx, gradient_checkpointing=True, name="policy")
ref = load_causal_lm(model, ctx, gradient_checkpointing=False, trainable=False, name="reference")
opt = AdamWConfig(lr=1e-6, weight_decay=0.01)
batch = synthetic_preference_batch(4, 1024, policy.config.vocab_size, seed=0)
batch = {s: {k: v.cuda() for k, v in batch[s].items()} for s in batch}


def step():
    loss, _ = dpo_step(policy.model, ref.model, batch, beta=0.1)
    loss.backward()
    policy.engine.finish_backward()
    policy.engine.optimizer_step(opt)


step()
policy.engine.auto_no_recompute(2 * 4 * 1024)     # same policy as bench.py: keep activations where HBM allows
for _ in range(2):
    step()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
agg = defaultdict(lambda: [0.0, 0])
fills = []
for ev in prof.events():
    if ev.device_type == torch.autograd.DeviceType.CUDA:
        name = ev.name
        for key in ("gemm2_kernel", "gemm_kernel", "attn_fwd", "attn_bwd2_kernel", "attn_bwd_kernel", "attn_bwd_prep", "attn_bwd_dq", "rmsnorm_fwd", "rmsnorm_bwd",
                    "colsum", "rope", "swiglu_fwd", "swiglu_bwd", "adamw", "sumsq", "embedding", "logprob_grad", "logprob_combine",
                    "Memset", "Memcpy"):
            if key in name:
                if key == "gemm_kernel":
                    key = "gemm_logprob" if "ELi1E" in name or ", 1>" in name or "(int)1>" in name else "gemm_store"
                elif key == "gemm2_kernel":
                    fuse = 0
                    for f in (1, 2, 3):
                        if f"ELi{f}E" in name or f", {f}>" in name or f"(int){f}>" in name:
                            fuse = f
                    key = ("gemm_2cta", "gemm_2cta+swiglu_fwd_epilogue", "gemm_2cta_dgrad+swiglu_bwd_epilogue",
                           "gemm_2cta_qkv+rope_epilogue")[fuse]
                name = key
                break
        if "at::native" in name:       # keep the functor name: which PyTorch elementwise ops are left in the step
            import re
            mm = re.search(r"(\w+Functor|\w+_kernel_cuda|FillFunctor|direct_copy_kernel|CatArray\w*|reduce_kernel|index\w*)", name)
            name = "torch:" + (mm.group(1) if mm else name[:60])
        dur = ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
        if name in ("torch:FillFunctor", "Memset", "Memcpy"):
            fills.append((dur, name))
        agg[name][0] += dur
        agg[name][1] += 1
rows = sorted(((k, v[0] / 1e3, v[1]) for k, v in agg.items()), key=lambda r: -r[1])
total = sum(r[1] for r in rows)
print(f"total kernel time {total:.1f} ms")
for k, ms, n in rows[:30]:
    print(f"{ms:9.2f} ms {100 * ms / total:5.1f}%  x{n:5d}  {k[:100]}")
# GPU idle time: gaps between consecutive kernels on the device timeline (all streams merged), attributed to the kernel
# that FOLLOWS the gap (the one the host was late to launch)
evs = sorted(((ev.time_range.start, ev.time_range.end, ev.name) for ev in prof.events()
              if ev.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e[0])
idle = defaultdict(lambda: [0.0, 0])
busy_until, t_first = None, None
for st_, en_, nm in evs:
    if busy_until is None:
        busy_until, t_first = en_, st_
        continue
    gap = st_ - busy_until
    if gap > 2.0:
        short = nm.split("(")[0][-48:]
        idle[short][0] += gap
        idle[short][1] += 1
    busy_until = max(busy_until, en_)
tot_idle = sum(v[0] for v in idle.values())
print(f"device timeline {(busy_until - t_first) / 1e3:.1f} ms, idle {tot_idle / 1e3:.1f} ms in gaps > 2 us; top followers of a gap:")
for k, v in sorted(idle.items(), key=lambda kv: -kv[1][0])[:14]:
    print(f"   {v[0] / 1e3:7.2f} ms  x{v[1]:4d}  avg {v[0] / v[1]:6.1f} us  before {k}")
fills.sort(reverse=True)
print("largest fill / memset / memcpy events (us):", [(round(d, 1), n) for d, n in fills[:12]], "count", len(fills))
Path(ROOT / "gpurun_out").mkdir(exist_ok=True)
json.dump({"total_ms": total, "timeline_ms": (busy_until - t_first) / 1e3, "idle_ms": tot_idle / 1e3, "rows": rows[:40]},
          open(ROOT / "gpurun_out" / "step_breakdown.json", "w"), indent=1)
