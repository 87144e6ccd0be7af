"""Timing probe for the attention backward: main kernel alone vs pre/post passes, with and without the dQ reduce-add."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.ops import native as N  # noqa: E402

dev = "cuda"
for B, T in ((8, 1024), (2, 4096)):
    Hq, Hk, D = 32, 8, 128
    qkv = torch.randn(B, T, Hq + 2 * Hk, D, device=dev, dtype=torch.bfloat16).requires_grad_()
    o = F.attention_qkv(qkv, Hq, Hk)
    g = torch.randn_like(o)
    for variant, tag in ((1, "gen1"), (0, "gen2"), (17, "gen2 without dQ reduce-add")):
        N.call("dla_attn_set_bwd_variant", variant)
        for _ in range(3):
            torch.autograd.grad(o, qkv, g, retain_graph=True)
        torch.cuda.synchronize()
        with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                torch.autograd.grad(o, qkv, g, retain_graph=True)
            torch.cuda.synchronize()
        rows = [(e.key[:60], e.device_time_total / 5) for e in prof.key_averages() if e.device_time_total > 0]
        print(f"B={B} T={T} {tag}: " + " | ".join(f"{k}: {t:.1f} us" for k, t in sorted(rows, key=lambda r: -r[1])[:4]), flush=True)
N.call("dla_attn_set_bwd_variant", 0)
