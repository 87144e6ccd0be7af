#!/usr/bin/env python
"""Tiny single-op drivers for ncu (one GPU, a handful of launches)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from dla_b200.ops import functional as F  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "gemm"
dev = "cuda"
r = lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16)
if what == "gemm":
    a, b = r(8192, 4096), r(14336 * 2, 4096)
    for _ in range(3):
        c = F.gemm(a, b)
elif what == "gemm_wgrad":
    dy, x = r(8192, 4096), r(8192, 14336)
    for _ in range(3):
        c = F.gemm(dy, x, a_mn=True, b_mn=True)
elif what in ("attn_fwd", "attn_bwd"):
    B, T, Hq, Hk, D = 4, 1024, 32, 8, 128
    qkv = r(B, T, Hq + 2 * Hk, D).requires_grad_()
    q, k, v = qkv[:, :, :Hq], qkv[:, :, Hq:Hq + Hk], qkv[:, :, Hq + Hk:]
    for _ in range(3):
        o = F.attention(q, k, v)
        if what == "attn_bwd":
            torch.autograd.grad(o, qkv, r(B, T, Hq, D))
elif what == "attn_fwd_4k":
    B, T, Hq, Hk, D = 2, 4096, 32, 8, 128
    qkv = r(B, T, Hq + 2 * Hk, D)
    for _ in range(3):
        o = F.attention_qkv(qkv, Hq, Hk)
elif what == "skinny":
    x, w = r(8, 4096), r(28672, 4096)
    for _ in range(3):
        y = F.gemm(x, w)
elif what == "decode_attn":
    from dla_b200.ops import native as N
    B, S, Hq, Hk, D = 8, 320, 32, 8, 128
    qkv = r(B, (Hq + 2 * Hk) * D)
    kc, vc = r(B, S, Hk, D), r(B, S, Hk, D)
    lens = torch.full((B,), 300, device=dev, dtype=torch.int32)
    cos, sin = F.rope_tables(8192, D, 500000.0, torch.device(dev))
    o = torch.empty(B, Hq, D, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        N.call("dla_decode_attn_fused", N.ptr(qkv), qkv.stride(0), N.ptr(kc), N.ptr(vc), N.ptr(o), N.ptr(lens), N.ptr(cos), N.ptr(sin),
               B, S, Hq, Hk, D, D, D ** -0.5, N.stream_ptr())
elif what == "lmhead":
    h, w = r(4096, 4096), r(128256, 4096) * 0.02
    lab = torch.randint(0, 128256, (4096,), device=dev)
    for _ in range(2):
        F.linear_logprob(h, w, lab)
elif what == "rmsnorm":
    x, d, w = r(8192, 4096).requires_grad_(), r(8192, 4096).requires_grad_(), torch.ones(4096, device=dev, dtype=torch.bfloat16).requires_grad_()
    for _ in range(3):
        xn, h = F.add_rmsnorm(x.view(8, 1024, 4096), d.view(8, 1024, 4096), w, 1e-5)
        torch.autograd.grad([xn, h], [x, d], [torch.ones_like(xn), torch.ones_like(h)])
elif what == "rope":
    qkv = r(8192, 48 * 128)
    pos = torch.arange(1024, device=dev, dtype=torch.int32).repeat(8)
    for _ in range(3):
        F.rope_qkv(qkv, pos, 32, 8, 128, 500000.0, None, max_pos=8192)
elif what == "swiglu":
    gu = r(8192, 28672).requires_grad_()
    for _ in range(3):
        y = F.swiglu(gu)
        torch.autograd.grad(y, gu, torch.ones_like(y))
elif what == "adamw":
    from dla_b200.ops import native as N
    n = 218_112_000
    master, m, v = (torch.zeros(n, device=dev) for _ in range(3))
    g = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    prm = torch.empty(n, device=dev, dtype=torch.bfloat16)
    coef = torch.ones(1, device=dev)
    for _ in range(3):
        N.call("dla_adamw", N.ptr(master), N.ptr(m), N.ptr(v), N.ptr(g), N.ptr(prm), n, 1e-5, 0.9, 0.999, 1e-8, 0.01, 0.1, 0.001,
               N.ptr(coef), 1.0, 0, N.stream_ptr())
elif what == "mlp_epilogues":      # gate/up GEMM with the SwiGLU epilogue (kFuse 1) and the down dgrad with its backward (kFuse 2)
    h = r(8192, 4096).requires_grad_()
    w_gu, w_down = (r(28672, 4096) * 0.02).requires_grad_(), (r(4096, 14336) * 0.02).requires_grad_()
    for _ in range(2):
        y = F.mlp_swiglu(h, w_gu, w_down)
        torch.autograd.grad(y, [h, w_gu, w_down], torch.ones_like(y))
elif what == "qkv_rope":           # QKV GEMM with the RoPE epilogue (kFuse 3)
    h, w = r(8192, 4096), r(6144, 4096) * 0.02
    pos = torch.arange(1024, device=dev).repeat(8)
    for _ in range(3):
        F.linear_rope_qkv(h, w, pos, 32, 8, 128, 500000.0, None, max_pos=8192)
elif what == "ag_gemm":            # all-gather fused GEMM with this GPU as the only "peer": copy CTAs + flag-gated tiles, no NVLink
    from dla_b200.ops import native as N
    M, Nn, K = 8192, 6144, 4096
    a = r(M, K)
    shard = r(Nn * K) * 0.02            # the "unit": one weight matrix, world = 1
    full = torch.empty_like(shard)
    c = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)
    flags = torch.zeros(8192, dtype=torch.int32, device=dev)
    dptrs = torch.tensor([shard.data_ptr()], dtype=torch.int64, device=dev)
    for ep in range(1, 4):
        N.call("dla_ag_gemm_bf16", N.ptr(a), N.ptr(dptrs), N.ptr(full), Nn * K, 0, N.ptr(c), M, Nn, K, K, Nn, 1, 0, N.ptr(flags), ep, 8, 0,
               N.stream_ptr())
    torch.cuda.synchronize()
    ref = a.float() @ shard.view(Nn, K).float().t()
    err = ((c.float() - ref).norm() / ref.norm()).item()
    assert err < 1e-2, err
torch.cuda.synchronize()
print("ok", what)
