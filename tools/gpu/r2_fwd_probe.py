"""Attention forward: optimistic single-pass softmax (default) vs the two-pass path (variant 20), device times."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.ops import native as N  # noqa: E402

for B, T in ((8, 1024), (2, 4096), (1, 8192)):
    Hq, Hk, D = 32, 8, 128
    qkv = torch.randn(B, T, Hq + 2 * Hk, D, device="cuda", dtype=torch.bfloat16)
    fl = 4.0 * B * Hq * T * T * D / 2
    for variant, tag in ((0, "single-pass"), (20, "two-pass")):
        N.call("dla_attn_set_fwd_variant", variant)
        for _ in range(3):
            F.attention_qkv(qkv, Hq, Hk)
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            F.attention_qkv(qkv, Hq, Hk)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = sorted(ts)[len(ts) // 2]
        print(f"B={B} T={T} {tag}: {ms * 1e3:.1f} us  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
    qs, ks, vs = (t.transpose(1, 2).contiguous() for t in (qkv[:, :, :Hq], qkv[:, :, Hq:Hq + Hk], qkv[:, :, Hq + Hk:]))
    sd = lambda: torch.nn.functional.scaled_dot_product_attention(qs, ks, vs, is_causal=True, enable_gqa=True)
    for _ in range(3):
        sd()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sd(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = sorted(ts)[len(ts) // 2]
    print(f"B={B} T={T} SDPA: {ms * 1e3:.1f} us  {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
N.call("dla_attn_set_fwd_variant", 0)
