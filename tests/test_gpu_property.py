"""Property-based shape coverage of the contraction kernels (hypothesis): random and ragged shapes against a plain
fp32 PyTorch restatement — every operand layout of the GEMM (incl. extents the TMA descriptors cannot address, split-K,
the K-split last wave of the 2-CTA kernel, accumulation and fp32 outputs), attention with ragged lengths and T that is
no multiple of the 128-row tiles, and the fused lm_head log-prob at real vocabulary sizes."""
import math

import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings, strategies as st  # noqa: E402

from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.ops import reference as R  # noqa: E402

pytestmark = pytest.mark.gpu
DEV = "cuda"
BF16 = torch.bfloat16
COMMON = dict(deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


def _rand(*shape, scale=1.0, seed=0):
    g = torch.Generator(device=DEV).manual_seed(seed)
    return (torch.randn(*shape, device=DEV, generator=g) * scale).to(BF16)


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()


@settings(max_examples=40, **COMMON)
@given(M=st.integers(1, 700), N=st.integers(1, 900), K=st.integers(1, 1500), a_mn=st.booleans(), b_mn=st.booleans(),
       fp32_out=st.booleans(), accumulate=st.booleans())
def test_gemm_any_shape_any_layout(native_lib, M, N, K, a_mn, b_mn, fp32_out, accumulate):
    """All four operand layouts, bf16 / fp32 output, store and accumulate, arbitrary (odd) extents: aligned shapes run the
    TMA kernel directly, the others through zero-padded staging buffers — never a library matmul."""
    a = _rand(K, M, seed=1) if a_mn else _rand(M, K, seed=1)
    b = _rand(K, N, seed=2) if b_mn else _rand(N, K, seed=2)
    want = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    out = torch.full((M, N), 0.5, device=DEV, dtype=torch.float32 if fp32_out else BF16)
    if accumulate:
        want = want + 0.5
    F.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=accumulate)
    assert _rel(out, want) < (2e-3 if fp32_out else 8e-3), (M, N, K, a_mn, b_mn, fp32_out, accumulate)


@settings(max_examples=12, **COMMON)
@given(M=st.sampled_from([64, 128, 256, 384]), N=st.sampled_from([256, 1024, 4096]), K=st.sampled_from([1024, 4096, 8192, 14336]))
def test_gemm_split_k_small_m(native_lib, M, N, K):
    """Few output tiles and a long K: the 1-CTA kernel splits K across CTAs and meets in C by reduce-add."""
    a, b = _rand(M, K, scale=0.5, seed=3), _rand(N, K, scale=0.5, seed=4)
    assert _rel(F.gemm(a, b), a.float() @ b.float().t()) < 8e-3


@settings(max_examples=8, **COMMON)
@given(mt=st.integers(9, 13), nt=st.integers(9, 13), K=st.sampled_from([1024, 2048, 4096]), fp32_out=st.booleans())
def test_gemm_two_cta_tail_wave(native_lib, mt, nt, K, fp32_out):
    """2-CTA kernel with a partial last wave (tile counts just above a multiple of 74 clusters): accumulating / fp32
    outputs take the K-split tail, bf16 stores keep their single rounding."""
    M, N = 256 * mt - 64, 256 * nt - 8
    a, b = _rand(K, M, scale=0.5, seed=5), _rand(K, N, scale=0.5, seed=6)
    out = torch.zeros(M, N, device=DEV, dtype=torch.float32 if fp32_out else BF16)
    F.gemm(a, b, a_mn=True, b_mn=True, out=out, accumulate=True)
    assert _rel(out, a.float().t() @ b.float()) < (2e-3 if fp32_out else 8e-3)


@settings(max_examples=16, **COMMON)
@given(B=st.integers(1, 3), T=st.integers(1, 700), heads=st.sampled_from([(4, 2, 128), (2, 1, 128), (4, 4, 64), (8, 2, 64)]),
       ragged=st.booleans(), variant=st.sampled_from([1, 2]), data=st.data())
def test_attention_ragged_shapes(native_lib, B, T, heads, ragged, variant, data):
    """Forward and both backward kernels at sequence lengths that are no multiple of the 128-row tiles, with per-row
    lengths anywhere in [1, T]."""
    from dla_b200.ops import native as N

    Hq, Hk, D = heads
    qkv = _rand(B, T, Hq + 2 * Hk, D, seed=7)
    seqlens = None
    if ragged:
        seqlens = torch.tensor([data.draw(st.integers(1, T)) for _ in range(B)], device=DEV, dtype=torch.int32)
    valid = torch.ones(B, T, 1, 1, device=DEV, dtype=torch.bool) if seqlens is None else \
        (torch.arange(T, device=DEV)[None, :] < seqlens[:, None])[:, :, None, None]
    N.call("dla_attn_set_bwd_variant", variant)
    try:
        x = qkv.clone().requires_grad_()
        o = F.attention_qkv(x, Hq, Hk, seqlens=seqlens, causal=True)
        g = _rand(B, T, Hq, D, seed=8) * valid
        o.backward(g)
        torch.cuda.synchronize()
    finally:
        N.call("dla_attn_set_bwd_variant", 0)
    xr = qkv.float().requires_grad_()
    oref = R.attention(xr[:, :, :Hq], xr[:, :, Hq:Hq + Hk], xr[:, :, Hq + Hk:], causal=True, seqlens=seqlens)
    (oref * g.float()).sum().backward()
    assert _rel(o * valid, oref * valid) < 1.5e-2
    assert _rel(x.grad, xr.grad) < 3e-2, (B, T, heads, ragged, variant)


@settings(max_examples=6, **COMMON)
@given(M=st.sampled_from([8, 96, 520, 1000]), V=st.sampled_from([50257, 32000, 128256, 1000]), K=st.sampled_from([256, 768]))
def test_lm_head_logprob_real_vocabularies(native_lib, M, V, K):
    """Fused lm_head log-prob (GEMM epilogue + combine) and its chunked backward at GPT-2 / Llama vocabulary sizes,
    including V = 50257 (not a multiple of 8) and ignored labels."""
    h = _rand(M, K, scale=1.0, seed=9)
    w = _rand(V, K, scale=K ** -0.5, seed=10)
    g = torch.Generator(device=DEV).manual_seed(11)
    labels = torch.randint(0, V, (M,), device=DEV, generator=g)
    labels[::7] = -100
    ha, wa = h.clone().requires_grad_(), w.clone().requires_grad_()
    logp, lse = F.linear_logprob(ha, wa, labels)
    hr, wr = h.float().requires_grad_(), w.float().requires_grad_()
    logits = hr @ wr.t()
    ref_lse = torch.logsumexp(logits, dim=-1)
    ref = torch.where(labels >= 0, logits.gather(1, labels.clamp(min=0)[:, None])[:, 0] - ref_lse, torch.zeros_like(ref_lse))
    assert torch.allclose(logp, ref, atol=3e-2, rtol=1e-2), (logp - ref).abs().max()
    wgt = torch.randn(M, device=DEV, generator=g)
    (logp * wgt).sum().backward()
    (ref * wgt).sum().backward()
    assert _rel(ha.grad, hr.grad) < 2e-2 and _rel(wa.grad, wr.grad) < 2e-2
