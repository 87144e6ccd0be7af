"""Differentiable ops.  CUDA tensors run the hand-written sm_100a kernels through
:mod:`dla_b200.ops.native`; CPU tensors run the PyTorch restatements in
:mod:`dla_b200.ops.reference` (used by the gloo plumbing tests and as the numerics oracle).
There is no third path: a CUDA tensor without the native library raises.
"""
from __future__ import annotations

import os as _os

import math
from typing import Optional, Tuple

import torch
from torch.autograd import Function

from dla_b200.ops import native as N
from dla_b200.ops import reference as R

BF16 = torch.bfloat16


def _s(t: torch.Tensor) -> int:
    return N.stream_ptr(t.device)


def _c(t: torch.Tensor) -> torch.Tensor:
    return t if t.is_contiguous() else t.contiguous()


def _check_bf16(*ts: Optional[torch.Tensor]) -> None:
    for t in ts:
        if t is not None and t.dtype != BF16:
            raise TypeError(f"native kernels take bf16 tensors, got {t.dtype}")


# ---- gradient slots ---------------------------------------------------------------------
# A parameter view may carry ``_grad_slot``: its slice of the engine's flat gradient bucket.  The engine does not
# memset 16 GB of buckets per step: slots that are always produced by a wgrad GEMM are only MARKED fresh
# (``_fresh``), and the first GEMM of the step overwrites instead of accumulating.  Every other kind of writer
# (adds, atomics, reductions) goes through ``_slot_add`` which zero-fills a fresh slot first.

def _slot_add(param: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """Slot of ``param`` for an accumulating writer (or None)."""
    slot = getattr(param, "_grad_slot", None) if param is not None else None
    if slot is not None and getattr(slot, "_fresh", False):
        slot.zero_()
        slot._fresh = False
    return slot


def _slot_gemm(param: torch.Tensor) -> Tuple[Optional[torch.Tensor], bool]:
    """``(slot, accumulate)`` for a wgrad GEMM that covers the whole slot: a fresh slot is overwritten."""
    slot = getattr(param, "_grad_slot", None)
    if slot is None:
        return None, False
    slot._ow = True                      # tells the engine this slot may be left un-zeroed
    if getattr(slot, "_fresh", False):
        slot._fresh = False
        return slot, False
    return slot, True


# =======================================================================================
# GEMM
# =======================================================================================

def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = BF16, accumulate: bool = False,
         bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``C[M,N] (+)= op(A) @ op(B)^T (+ bias[N])`` on the tcgen05 kernel.

    ``a``: ``[M, K]`` (or ``[K, M]`` when ``a_mn``), ``b``: ``[N, K]`` (or ``[K, N]`` when ``b_mn``); both
    2-D with unit inner stride.  Shapes the TMA path cannot address (K or a pitch not a multiple
    of 8, unaligned base) run the same kernel on zero-padded staging buffers (:func:`_gemm_padded`).
    ``bias`` (bf16, K-major operands, no accumulate) is added to the fp32 accumulators in the GEMM epilogue.
    """
    assert a.dim() == 2 and b.dim() == 2
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    Nn, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    assert K == Kb, f"gemm: inner dims differ ({K} vs {Kb})"
    if a.stride(1) != 1:
        a = a.contiguous()
    if b.stride(1) != 1:
        b = b.contiguous()
    if out is None:
        out = torch.empty(M, Nn, dtype=out_dtype, device=a.device)
        accumulate = False
    ok = (a.is_cuda and a.dtype == BF16 and b.dtype == BF16 and K % 8 == 0 and a.stride(0) % 8 == 0
          and b.stride(0) % 8 == 0 and a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0
          and out.stride(1) == 1 and out.dtype in (BF16, torch.float32)
          and (not a_mn or M % 8 == 0) and (not b_mn or Nn % 8 == 0))
    if bias is not None:
        assert not (a_mn or b_mn or accumulate) and bias.numel() == Nn, "gemm: bias goes with a plain forward product"
    if not ok:
        if a.is_cuda and a.dtype == BF16 and b.dtype == BF16 and out.dtype in (BF16, torch.float32):
            _gemm_padded(a, b, a_mn, b_mn, out, accumulate, M, Nn, K)
            if bias is not None and out.dtype == BF16 and out.is_contiguous():
                N.call("dla_add_bias", N.ptr(out), None, N.ptr(bias), N.ptr(out), M, Nn, _s(out))
            elif bias is not None:
                out.add_(bias.to(out.dtype))
            return out
        am = a.t() if a_mn else a
        bm = b if b_mn else b.t()
        res = torch.matmul(am.to(torch.float32), bm.to(torch.float32))      # CPU tensors / non-bf16 dtypes only
        if bias is not None:
            res = res + bias.to(torch.float32)
        if accumulate:
            out.add_(res.to(out.dtype))
        else:
            out.copy_(res.to(out.dtype))
        return out
    if M <= SKINNY_MAX_ROWS and K % 32 == 0 and not (a_mn or b_mn or accumulate):
        # decode shapes: bound by the HBM stream of the weight, see csrc/skinny.cu
        N.call("dla_gemm_skinny_bf16", N.ptr(a), N.ptr(b), N.ptr(out), N.ptr(bias) if bias is not None else None, None, M, Nn, K,
               a.stride(0), b.stride(0), out.stride(0), 0, int(out.dtype == torch.float32), int(_STATIC_WEIGHTS[0]), _s(a))
        return out
    if bias is not None:
        N.call("dla_gemm_bias_bf16", N.ptr(a), N.ptr(b), N.ptr(out), N.ptr(bias), M, Nn, K, a.stride(0), b.stride(0), out.stride(0),
               int(out.dtype == torch.float32), _s(a))
        return out
    N.call("dla_gemm_bf16", N.ptr(a), N.ptr(b), N.ptr(out), M, Nn, K, a.stride(0), b.stride(0), out.stride(0),
           int(a_mn), int(b_mn), int(out.dtype == torch.float32), int(accumulate), _s(a))
    return out


def _gemm_padded(a, b, a_mn, b_mn, out, accumulate, M, Nn, K):
    """Shapes the TMA descriptors cannot address directly (K, an MN-major extent or a row pitch that is not a multiple
    of 8 elements, a base that is not 16-byte aligned): the operands are copied into zero-padded staging buffers with
    conforming extents and the SAME tcgen05 kernel runs on them; zero padding along K adds nothing, padded rows /
    columns of the result are dropped.  (Round 1 sent these shapes to a library matmul.)"""
    up8 = lambda n: (n + 7) // 8 * 8
    Kp = up8(K)
    Mp = up8(M) if a_mn else M
    Np = up8(Nn) if b_mn else Nn

    def stage(t, mn, rows_mn, rows_mn_p):
        want = (Kp, rows_mn_p) if mn else (t.shape[0], Kp)
        if tuple(t.shape) == want and t.stride(1) == 1 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0:
            return t
        buf = torch.zeros(want, dtype=BF16, device=t.device)
        buf[: t.shape[0], : t.shape[1]].copy_(t)
        return buf

    a2, b2 = stage(a, a_mn, M, Mp), stage(b, b_mn, Nn, Np)
    direct = (Mp == M and Np == Nn and out.stride(1) == 1)
    tgt = out if direct else torch.empty(Mp, Np, dtype=out.dtype, device=out.device)
    N.call("dla_gemm_bf16", N.ptr(a2), N.ptr(b2), N.ptr(tgt), Mp, Np, Kp, a2.stride(0), b2.stride(0), tgt.stride(0),
           int(a_mn), int(b_mn), int(tgt.dtype == torch.float32), int(accumulate and direct), _s(a2))
    if not direct:
        if accumulate:
            out.add_(tgt[:M, :Nn])
        else:
            out.copy_(tgt[:M, :Nn])
    return out


SKINNY_MAX_ROWS = 16                 # above this the 128-row tcgen05 tiles win (profiles/kernels_r1_skinny.jsonl)
_STATIC_WEIGHTS = [False]


class weights_static:
    """Context: no kernel in the stream writes the weights while it is active (generation, evaluation), so
    decode GEMMs may start early and prefetch their weight stream (``weights_static`` of dla_gemm_skinny_bf16)."""

    def __enter__(self):
        self.prev = _STATIC_WEIGHTS[0]
        _STATIC_WEIGHTS[0] = True
        return self

    def __exit__(self, *exc):
        _STATIC_WEIGHTS[0] = self.prev
        return False


def linear_residual(x: torch.Tensor, w: torch.Tensor, resid: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``resid + x @ w^T (+ bias)``; for decode-sized inputs the add rides in the weight-streaming GEMM's epilogue."""
    x2 = x.reshape(-1, x.shape[-1])
    M, K = x2.shape
    r2 = resid.reshape(M, -1)
    if (x2.is_cuda and not torch.is_grad_enabled() and M <= SKINNY_MAX_ROWS and K % 32 == 0 and x2.dtype == BF16 and w.dtype == BF16
            and x2.stride(1) == 1 and x2.stride(0) % 8 == 0 and w.stride(1) == 1 and w.stride(0) % 8 == 0 and r2.stride(1) == 1
            and r2.dtype == BF16 and x2.data_ptr() % 16 == 0 and w.data_ptr() % 16 == 0):
        out = torch.empty(M, w.shape[0], dtype=BF16, device=x.device)
        N.call("dla_gemm_skinny_bf16", N.ptr(x2), N.ptr(w), N.ptr(out), N.ptr(bias) if bias is not None else None, N.ptr(r2),
               M, w.shape[0], K, x2.stride(0), w.stride(0), out.stride(0), r2.stride(0), 0, int(_STATIC_WEIGHTS[0]), _s(x2))
        return out.view(resid.shape)
    return resid + linear(x, w, bias)


class _Linear(Function):
    """``y = x @ w^T (+ b)``; dgrad and wgrad reuse the same tcgen05 kernel with MN-major operands.
    If the weight carries ``_grad_slot`` (a view into the engine's flat gradient bucket) the wgrad
    GEMM accumulates straight into it and autograd sees no weight gradient."""

    @staticmethod
    def forward(ctx, x, w, bias):
        x2 = x.reshape(-1, x.shape[-1])
        ag = getattr(w, "_ag", None)
        if ag is not None and not ag.done:
            y = ag.run(_c(x2))          # weight rows are all-gathered over NVLink inside this GEMM (path (b))
            if bias is not None:
                N.call("dla_add_bias", N.ptr(y), None, N.ptr(bias), N.ptr(y), y.shape[0], y.shape[1], _s(y))
        else:
            y = gemm(x2, w, bias=bias)      # the bias add rides in the GEMM epilogue
        ctx.save_for_backward(x2, w)
        ctx.has_bias = bias is not None
        ctx.bias_ref = bias
        ctx.wref = w
        ctx.xshape = x.shape
        # a 2-D input gets the GEMM output itself (not a view): callers may modify it in place (RoPE)
        return y if x.dim() == 2 else y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w = ctx.saved_tensors
        dy2 = _c(dy.reshape(-1, dy.shape[-1]))
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn=True).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            slot, acc = _slot_gemm(ctx.wref)
            if slot is not None:
                gemm(dy2, x2, a_mn=True, b_mn=True, out=slot, accumulate=acc)
            else:
                dw = gemm(dy2, x2, a_mn=True, b_mn=True)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            b = ctx.bias_ref
            slot = _slot_add(b)
            H = dy2.shape[1]
            scratch = N.workspace("norm_bwd", N.norm_bwd_blocks() * H * 4, dy2.device)
            if slot is not None:
                N.call("dla_bias_grad", N.ptr(dy2), N.ptr(slot), N.ptr(scratch), 1, dy2.shape[0], H, _s(dy2))
            else:
                db = torch.empty(H, dtype=BF16, device=dy2.device)
                N.call("dla_bias_grad", N.ptr(dy2), N.ptr(db), N.ptr(scratch), 0, dy2.shape[0], H, _s(dy2))
        return dx, dw, db


def linear(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not x.is_cuda:
        return R.linear(x, w, bias)
    _check_bf16(x, w, bias)
    return _Linear.apply(x, w, bias)


# =======================================================================================
# norms
# =======================================================================================

class _AddRMSNorm(Function):
    @staticmethod
    def forward(ctx, x, delta, w, eps):
        H = x.shape[-1]
        x2 = _c(x.reshape(-1, H))
        M = x2.shape[0]
        y = torch.empty_like(x2)
        rstd = torch.empty(M, dtype=torch.float32, device=x.device)
        if delta is not None:
            d2 = _c(delta.reshape(-1, H))
            resid = torch.empty_like(x2)
            N.call("dla_rmsnorm_fwd", N.ptr(x2), N.ptr(d2), N.ptr(w), N.ptr(resid), N.ptr(y), N.ptr(rstd), M, H, eps, _s(x))
        else:
            resid = x2
            N.call("dla_rmsnorm_fwd", N.ptr(x2), None, N.ptr(w), None, N.ptr(y), N.ptr(rstd), M, H, eps, _s(x))
        ctx.save_for_backward(resid, w, rstd)
        ctx.has_delta = delta is not None
        ctx.shape = x.shape
        ctx.wref = w
        return resid.view(x.shape), y.view(x.shape)

    @staticmethod
    def backward(ctx, dresid, dy):
        resid, w, rstd = ctx.saved_tensors
        M, H = resid.shape
        dy2 = _c(dy.reshape(M, H)) if dy is not None else torch.zeros_like(resid)
        dr2 = _c(dresid.reshape(M, H)) if dresid is not None else None
        dx = torch.empty_like(resid)
        scratch = N.workspace("norm_bwd", N.norm_bwd_blocks() * H * 4, resid.device)
        slot = _slot_add(ctx.wref)
        dw = None
        need_w = ctx.needs_input_grad[2]
        if need_w and slot is None:
            dw = torch.empty(H, dtype=BF16, device=resid.device)
        tgt = slot if slot is not None else dw
        N.call("dla_rmsnorm_bwd", N.ptr(dy2), N.ptr(resid), N.ptr(w), N.ptr(rstd), N.ptr(dr2), N.ptr(dx),
               N.ptr(tgt) if need_w else None, N.ptr(scratch), int(slot is not None), M, H, _s(resid))
        dxv = dx.view(ctx.shape)
        return dxv, (dxv if ctx.has_delta else None), dw, None


def add_rmsnorm(x: torch.Tensor, delta: Optional[torch.Tensor], w: torch.Tensor, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``resid = x + delta`` (or ``x``), ``y = rmsnorm(resid) * w``; returns ``(resid, y)``."""
    if not x.is_cuda:
        resid = x if delta is None else x + delta
        return resid, R.rmsnorm(resid, w, eps)
    _check_bf16(x, delta, w)
    return _AddRMSNorm.apply(x, delta, w, eps)


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    return add_rmsnorm(x, None, w, eps)[1]


class _LayerNorm(Function):
    @staticmethod
    def forward(ctx, x, w, b, eps):
        H = x.shape[-1]
        x2 = _c(x.reshape(-1, H))
        M = x2.shape[0]
        y = torch.empty_like(x2)
        stats = torch.empty(2 * M, dtype=torch.float32, device=x.device)
        N.call("dla_layernorm_fwd", N.ptr(x2), N.ptr(w), N.ptr(b), N.ptr(y), N.ptr(stats), M, H, eps, _s(x))
        ctx.save_for_backward(x2, w, stats)
        ctx.refs = (w, b)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, w, stats = ctx.saved_tensors
        M, H = x2.shape
        dy2 = _c(dy.reshape(M, H))
        dx = torch.empty_like(x2)
        scratch = N.workspace("norm_bwd", N.norm_bwd_blocks() * 2 * H * 4, x2.device)
        wref, bref = ctx.refs
        wslot = _slot_add(wref)
        bslot = _slot_add(bref)
        use_slots = wslot is not None and (bref is None or bslot is not None)
        dw = wslot if use_slots else torch.empty(H, dtype=BF16, device=x2.device)
        db = None
        if bref is not None:
            db = bslot if use_slots else torch.empty(H, dtype=BF16, device=x2.device)
        N.call("dla_layernorm_bwd", N.ptr(dy2), N.ptr(x2), N.ptr(w), N.ptr(stats), N.ptr(dx), N.ptr(dw), N.ptr(db),
               N.ptr(scratch), int(use_slots), M, H, _s(x2))
        if use_slots:
            return dx.view(ctx.shape), None, None, None
        return dx.view(ctx.shape), dw, db, None


def layernorm(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float) -> torch.Tensor:
    if not x.is_cuda:
        return R.layernorm(x, w, b, eps)
    _check_bf16(x, w, b)
    return _LayerNorm.apply(x, w, b, eps)


# =======================================================================================
# RoPE on packed QKV
# =======================================================================================
_ROPE_TABLES: dict = {}


def rope_tables(max_pos: int, rot_dim: int, theta: float, device: torch.device) -> Tuple[torch.Tensor, torch.Tensor]:
    key = (max_pos, rot_dim, float(theta), str(device))
    if key not in _ROPE_TABLES:
        cos, sin = R.rope_tables(torch.arange(max_pos, device=device), rot_dim, theta)
        _ROPE_TABLES[key] = (cos.contiguous(), sin.contiguous())
    return _ROPE_TABLES[key]


class _RopeQKV(Function):
    """In-place RoPE on the q and k heads of a packed ``[M, (Hq + 2 Hk) * D]`` projection."""

    @staticmethod
    def forward(ctx, qkv, positions, cos, sin, n_rot_heads, head_dim, rot_dim):
        M = qkv.shape[0]
        N.call("dla_rope_inplace", N.ptr(qkv), N.ptr(cos), N.ptr(sin), N.ptr(positions), M, n_rot_heads, head_dim,
               rot_dim, qkv.stride(0), 1.0, _s(qkv))
        ctx.mark_dirty(qkv)
        ctx.save_for_backward(positions, cos, sin)
        ctx.meta = (n_rot_heads, head_dim, rot_dim)
        return qkv

    @staticmethod
    def backward(ctx, g):
        positions, cos, sin = ctx.saved_tensors
        n_rot_heads, head_dim, rot_dim = ctx.meta
        width = g.shape[1]
        if rot_dim == head_dim and head_dim % 16 == 0 and width % head_dim == 0 and g.stride(1) == 1 and g.stride(0) % 8 == 0:
            # the incoming gradient must not be modified: rotate q/k and copy v into a fresh tensor in ONE pass
            out = torch.empty(g.shape, dtype=g.dtype, device=g.device)
            N.call("dla_rope_copy", N.ptr(g), N.ptr(out), N.ptr(cos), N.ptr(sin), N.ptr(positions), g.shape[0], n_rot_heads,
                   width // head_dim - n_rot_heads, head_dim, g.stride(0), out.stride(0), -1.0, _s(g))
            return out, None, None, None, None, None, None
        g = g.clone(memory_format=torch.contiguous_format)
        N.call("dla_rope_inplace", N.ptr(g), N.ptr(cos), N.ptr(sin), N.ptr(positions), g.shape[0], n_rot_heads,
               head_dim, rot_dim, g.stride(0), -1.0, _s(g))
        return g, None, None, None, None, None, None


def rope_qkv(qkv: torch.Tensor, positions: torch.Tensor, n_q: int, n_kv: int, head_dim: int, theta: float,
             rot_dim: Optional[int] = None, max_pos: int = 8192) -> torch.Tensor:
    """Applies RoPE to the q and k parts of packed ``qkv [M, (n_q + 2 n_kv) * D]``; ``positions [M]``."""
    rd = rot_dim or head_dim
    if not qkv.is_cuda:
        M = qkv.shape[0]
        q, k, v = qkv.split([n_q * head_dim, n_kv * head_dim, n_kv * head_dim], dim=-1)
        q = R.rope(q.reshape(1, M, n_q, head_dim), positions.reshape(1, M), theta, rd).reshape(M, -1)
        k = R.rope(k.reshape(1, M, n_kv, head_dim), positions.reshape(1, M), theta, rd).reshape(M, -1)
        return torch.cat([q, k, v], dim=-1)
    cos, sin = rope_tables(max_pos, rd, theta, qkv.device)
    pos32 = positions.to(torch.int32)
    return _RopeQKV.apply(qkv, pos32, cos, sin, n_q + n_kv, head_dim, rd)


FUSE_ROPE = _os.environ.get("DLA_FUSE_ROPE", "1") != "0"


class _LinearRopeQKV(Function):
    """Packed QKV projection with RoPE in the GEMM epilogue (``dla_gemm_rope_bf16``): the q and k heads are rotated on
    the fp32 accumulators and rounded once.  Backward: inverse rotation of the incoming gradient (one out-of-place
    pass), then the usual dgrad / wgrad GEMMs with the gradient-slot protocol of :class:`_Linear`."""

    @staticmethod
    def forward(ctx, x2, w, positions, cos, sin, n_rot_heads, head_dim):
        M, K = x2.shape
        Nn = w.shape[0]
        y = torch.empty(M, Nn, dtype=BF16, device=x2.device)
        N.call("dla_gemm_rope_bf16", N.ptr(x2), N.ptr(w), N.ptr(y), M, Nn, K, x2.stride(0), w.stride(0), y.stride(0),
               N.ptr(cos), N.ptr(sin), N.ptr(positions), n_rot_heads, _s(x2))
        ctx.save_for_backward(x2, w, positions, cos, sin)
        ctx.wref = w
        ctx.meta = (n_rot_heads, head_dim)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, w, positions, cos, sin = ctx.saved_tensors
        n_rot_heads, head_dim = ctx.meta
        dy = dy if (dy.stride(1) == 1 and dy.stride(0) % 8 == 0) else dy.contiguous()
        g = torch.empty(dy.shape, dtype=dy.dtype, device=dy.device)
        N.call("dla_rope_copy", N.ptr(dy), N.ptr(g), N.ptr(cos), N.ptr(sin), N.ptr(positions), dy.shape[0], n_rot_heads,
               dy.shape[1] // head_dim - n_rot_heads, head_dim, dy.stride(0), g.stride(0), -1.0, _s(dy))
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(g, w, b_mn=True)
        if ctx.needs_input_grad[1]:
            slot, acc = _slot_gemm(ctx.wref)
            if slot is not None:
                gemm(g, x2, a_mn=True, b_mn=True, out=slot, accumulate=acc)
            else:
                dw = gemm(g, x2, a_mn=True, b_mn=True)
        return dx, dw, None, None, None, None, None


def linear_rope_qkv(x2: torch.Tensor, w: torch.Tensor, positions: torch.Tensor, n_q: int, n_kv: int, head_dim: int,
                    theta: float, rot_dim: Optional[int] = None, max_pos: int = 8192) -> torch.Tensor:
    """``rope_qkv(linear(x2, w))`` for a packed ``[(n_q + 2 n_kv) * D, K]`` projection weight, as ONE kernel when the
    shape allows it (CUDA, head_dim 128, full rotary, enough rows for the 2-CTA kernel, weight resident)."""
    rd = rot_dim or head_dim
    fused = (FUSE_ROPE and x2.is_cuda and x2.dim() == 2 and head_dim == 128 and rd == 128 and x2.shape[0] >= 256
             and x2.dtype == BF16 and w.dtype == BF16 and x2.shape[1] % 8 == 0 and x2.stride(1) == 1 and w.stride(1) == 1
             and x2.stride(0) % 8 == 0 and w.stride(0) % 8 == 0 and x2.data_ptr() % 16 == 0 and w.data_ptr() % 16 == 0
             and w.shape[0] == (n_q + 2 * n_kv) * head_dim and getattr(w, "_ag", None) is None)
    if not fused:
        return rope_qkv(linear(x2, w), positions, n_q, n_kv, head_dim, theta, rot_dim, max_pos=max_pos)
    cos, sin = rope_tables(max_pos, rd, theta, x2.device)
    return _LinearRopeQKV.apply(x2, w, positions.to(torch.int32), cos, sin, n_q + n_kv, head_dim)


# =======================================================================================
# activations
# =======================================================================================

class _SwiGLU(Function):
    @staticmethod
    def forward(ctx, gu):
        F2 = gu.shape[-1]
        g2 = _c(gu.reshape(-1, F2))
        h = torch.empty(g2.shape[0], F2 // 2, dtype=gu.dtype, device=gu.device)
        N.call("dla_swiglu_fwd", N.ptr(g2), N.ptr(h), g2.shape[0], F2 // 2, _s(gu))
        ctx.save_for_backward(g2)
        ctx.shape = gu.shape
        return h.view(*gu.shape[:-1], F2 // 2)

    @staticmethod
    def backward(ctx, dh):
        (g2,) = ctx.saved_tensors
        d2 = _c(dh.reshape(g2.shape[0], -1))
        dgu = torch.empty_like(g2)
        N.call("dla_swiglu_bwd", N.ptr(d2), N.ptr(g2), N.ptr(dgu), g2.shape[0], g2.shape[1] // 2, _s(g2))
        return dgu.view(ctx.shape)


class _LinearSwiGLU(Function):
    """``silu(x W_gate^T) * (x W_up^T)`` with the activation computed in the epilogue of the gate/up GEMM (2-CTA tcgen05
    kernel, ``csrc/gemm2.cu``): no separate SwiGLU pass over the ``[M, 2F]`` projection.  gate | up are stored only when
    a backward follows; the backward is the usual SwiGLU backward + dgrad + wgrad (into the weight's gradient slot)."""

    @staticmethod
    def forward(ctx, x, w):
        x2 = _c(x.reshape(-1, x.shape[-1]))
        M, K = x2.shape
        F2 = w.shape[0]
        need_bwd = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        act = torch.empty(M, F2 // 2, dtype=BF16, device=x.device)
        gu = torch.empty(M, F2, dtype=BF16, device=x.device) if need_bwd else None
        N.call("dla_gemm_swiglu_bf16", N.ptr(x2), N.ptr(w), N.ptr(gu), N.ptr(act), M, F2 // 2, K, x2.stride(0), w.stride(0),
               F2, F2 // 2, _s(x))
        if need_bwd:
            ctx.save_for_backward(x2, w, gu)
        ctx.wref = w
        ctx.xshape = x.shape
        return act.view(*x.shape[:-1], F2 // 2)

    @staticmethod
    def backward(ctx, dact):
        x2, w, gu = ctx.saved_tensors
        d2 = _c(dact.reshape(gu.shape[0], -1))
        dgu = torch.empty_like(gu)
        N.call("dla_swiglu_bwd", N.ptr(d2), N.ptr(gu), N.ptr(dgu), gu.shape[0], gu.shape[1] // 2, _s(gu))
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dgu, w, b_mn=True).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            slot, acc = _slot_gemm(ctx.wref)
            if slot is not None:
                gemm(dgu, x2, a_mn=True, b_mn=True, out=slot, accumulate=acc)
            else:
                dw = gemm(dgu, x2, a_mn=True, b_mn=True)
        return dx, dw


class _MLPSwiGLU(Function):
    """The whole gated MLP ``(silu(x Wg^T) * (x Wu^T)) Wd^T`` as one autograd node, so that both SwiGLU passes ride in GEMM
    epilogues: forward as in :class:`_LinearSwiGLU`; backward, the dgrad of the down projection emits d(gate) | d(up)
    directly (``dla_gemm_swiglu_bwd_bf16``) — d(act) never reaches memory — followed by the usual dgrad / wgrad GEMMs."""

    @staticmethod
    def forward(ctx, x, w_gu, w_down):
        x2 = _c(x.reshape(-1, x.shape[-1]))
        M, K = x2.shape
        F2 = w_gu.shape[0]
        need_bwd = any(ctx.needs_input_grad)
        act = torch.empty(M, F2 // 2, dtype=BF16, device=x.device)
        gu = torch.empty(M, F2, dtype=BF16, device=x.device) if need_bwd else None
        N.call("dla_gemm_swiglu_bf16", N.ptr(x2), N.ptr(w_gu), N.ptr(gu), N.ptr(act), M, F2 // 2, K, x2.stride(0), w_gu.stride(0),
               F2, F2 // 2, _s(x))
        y = gemm(act, w_down)
        if need_bwd:
            ctx.save_for_backward(x2, w_gu, w_down, gu, act)
        ctx.wrefs = (w_gu, w_down)
        ctx.xshape = x.shape
        return y.view(*x.shape[:-1], w_down.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w_gu, w_down, gu, act = ctx.saved_tensors
        dy2 = _c(dy.reshape(-1, dy.shape[-1]))
        M, H = dy2.shape
        Fh = act.shape[1]
        dx = dwg = dwd = None
        if ctx.needs_input_grad[2]:
            slot, acc = _slot_gemm(ctx.wrefs[1])
            if slot is not None:
                gemm(dy2, act, a_mn=True, b_mn=True, out=slot, accumulate=acc)
            else:
                dwd = gemm(dy2, act, a_mn=True, b_mn=True)
        dgu = torch.empty_like(gu)
        N.call("dla_gemm_swiglu_bwd_bf16", N.ptr(dy2), N.ptr(w_down), N.ptr(gu), N.ptr(dgu), M, Fh, H, dy2.stride(0),
               w_down.stride(0), gu.stride(0), dgu.stride(0), _s(dy2))
        if ctx.needs_input_grad[0]:
            dx = gemm(dgu, w_gu, b_mn=True).view(ctx.xshape)
        if ctx.needs_input_grad[1]:
            slot, acc = _slot_gemm(ctx.wrefs[0])
            if slot is not None:
                gemm(dgu, x2, a_mn=True, b_mn=True, out=slot, accumulate=acc)
            else:
                dwg = gemm(dgu, x2, a_mn=True, b_mn=True)
        return dx, dwg, dwd


def mlp_swiglu(x: torch.Tensor, w_gu: torch.Tensor, w_down: torch.Tensor) -> torch.Tensor:
    """``linear(swiglu(linear(x, w_gu)), w_down)``.  When the shapes suit the 2-CTA kernel both SwiGLU passes are GEMM
    epilogues (:class:`_MLPSwiGLU`); otherwise the composition of the separate ops."""
    F2, K = w_gu.shape
    M = x.numel() // x.shape[-1]
    ok = (x.is_cuda and FUSE_SWIGLU[0] and x.dtype == BF16 and w_gu.dtype == BF16 and w_down.dtype == BF16 and F2 % 256 == 0
          and M >= 256 and K % 8 == 0 and w_down.shape[0] % 8 == 0 and ((M + 255) // 256) * (F2 // 256) >= 74
          and all(w.stride(1) == 1 and w.stride(0) % 8 == 0 and w.data_ptr() % 16 == 0 for w in (w_gu, w_down))
          and all(getattr(w, "_ag", None) is None or w._ag.done for w in (w_gu, w_down)))
    if not ok:
        return linear(linear_swiglu(x, w_gu), w_down)
    return _MLPSwiGLU.apply(x, w_gu, w_down)


FUSE_SWIGLU = [True]        # A/B switch (tests, DLA_FUSE_SWIGLU=0)


def linear_swiglu(x: torch.Tensor, w_gu: torch.Tensor) -> torch.Tensor:
    """``swiglu(linear(x, w_gu))`` for ``w_gu = [gate rows | up rows]``; on CUDA, when the shape suits the 2-CTA kernel
    (F % 128 == 0, at least half a wave of 256 x 128-feature tiles), the activation is fused into the GEMM epilogue."""
    F2, K = w_gu.shape
    M = x.numel() // x.shape[-1]
    ag = getattr(w_gu, "_ag", None)
    fused = (x.is_cuda and FUSE_SWIGLU[0] and x.dtype == BF16 and w_gu.dtype == BF16 and F2 % 256 == 0 and M >= 256 and K % 8 == 0
             and ((M + 255) // 256) * (F2 // 256) >= 74 and w_gu.stride(1) == 1 and w_gu.stride(0) % 8 == 0
             and w_gu.data_ptr() % 16 == 0 and (ag is None or ag.done))
    if not fused:
        return swiglu(linear(x, w_gu))
    return _LinearSwiGLU.apply(x, w_gu)


if _os.environ.get("DLA_FUSE_SWIGLU", "1") == "0":
    FUSE_SWIGLU[0] = False


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    if not gu.is_cuda:
        return R.swiglu(gu)
    _check_bf16(gu)
    return _SwiGLU.apply(gu)


class _GeluBias(Function):
    @staticmethod
    def forward(ctx, x, bias):
        Fd = x.shape[-1]
        x2 = _c(x.reshape(-1, Fd))
        y = torch.empty_like(x2)
        N.call("dla_gelu_fwd", N.ptr(x2), N.ptr(bias), N.ptr(y), x2.shape[0], Fd, _s(x))
        ctx.save_for_backward(x2, bias) if bias is not None else ctx.save_for_backward(x2)
        ctx.has_bias = bias is not None
        ctx.bias_ref = bias
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        saved = ctx.saved_tensors
        x2 = saved[0]
        bias = saved[1] if ctx.has_bias else None
        d2 = _c(dy.reshape(x2.shape))
        dx = torch.empty_like(x2)
        N.call("dla_gelu_bwd", N.ptr(d2), N.ptr(x2), N.ptr(bias), N.ptr(dx), x2.shape[0], x2.shape[1], _s(x2))
        db = None
        if ctx.has_bias and ctx.needs_input_grad[1]:
            H = x2.shape[1]
            scratch = N.workspace("norm_bwd", N.norm_bwd_blocks() * H * 4, x2.device)
            slot = _slot_add(ctx.bias_ref)
            if slot is not None:
                N.call("dla_bias_grad", N.ptr(dx), N.ptr(slot), N.ptr(scratch), 1, x2.shape[0], H, _s(x2))
            else:
                db = torch.empty(H, dtype=BF16, device=x2.device)
                N.call("dla_bias_grad", N.ptr(dx), N.ptr(db), N.ptr(scratch), 0, x2.shape[0], H, _s(x2))
        return dx.view(ctx.shape), db


def gelu_bias(x: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """tanh-GELU of ``x + bias`` (GPT-2's ``gelu_new``)."""
    if not x.is_cuda:
        return R.gelu_tanh(x if bias is None else x + bias)
    _check_bf16(x, bias)
    return _GeluBias.apply(x, bias)


# =======================================================================================
# embedding
# =======================================================================================

class _Embedding(Function):
    @staticmethod
    def forward(ctx, ids, table, pos_table, T):
        flat = _c(ids.reshape(-1))
        M, (V, H) = flat.numel(), table.shape
        out = torch.empty(M, H, dtype=table.dtype, device=table.device)
        N.call("dla_embedding_fwd", N.ptr(flat), N.ptr(table), N.ptr(pos_table), N.ptr(out), M, H, V, T, _s(table))
        ctx.save_for_backward(flat)
        ctx.refs = (table, pos_table)
        ctx.meta = (M, H, V, T)
        return out.view(*ids.shape, H)

    @staticmethod
    def backward(ctx, dout):
        (flat,) = ctx.saved_tensors
        M, H, V, T = ctx.meta
        table, pos_table = ctx.refs
        d2 = _c(dout.reshape(M, H))
        tslot = _slot_add(table)
        pslot = _slot_add(pos_table)
        dt = tslot if tslot is not None else torch.zeros(V, H, dtype=BF16, device=d2.device)
        dp = None
        if pos_table is not None:
            dp = pslot if pslot is not None else torch.zeros_like(pos_table)
        N.call("dla_embedding_bwd", N.ptr(flat), N.ptr(d2), N.ptr(dt), N.ptr(dp), M, H, V, T, _s(d2))
        return None, (None if tslot is not None else dt), (None if (pos_table is None or pslot is not None) else dp), None


def embedding(ids: torch.Tensor, table: torch.Tensor, pos_table: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Token (+ learned position, GPT-2) embedding; ``ids [B, T]``."""
    if not table.is_cuda:
        out = torch.nn.functional.embedding(ids, table)
        if pos_table is not None:
            out = out + pos_table[: ids.shape[-1]]
        return out
    _check_bf16(table, pos_table)
    return _Embedding.apply(ids, table, pos_table, ids.shape[-1])


# =======================================================================================
# attention
# =======================================================================================

def segment_starts(segment_ids: torch.Tensor) -> torch.Tensor:
    """``[B, T]`` int32: for every token the index of the first token of its document (``segment_ids`` as produced by
    ``PackedCollator``: consecutive equal ids form a document, 0 = padding at the end of the row).  This is the form
    the flash kernels take: the keys a query may see are ``[start[q], q]``."""
    B, T = segment_ids.shape
    idx = torch.arange(T, device=segment_ids.device, dtype=torch.int32).expand(B, T)
    first = torch.ones(B, T, dtype=torch.bool, device=segment_ids.device)
    first[:, 1:] = segment_ids[:, 1:] != segment_ids[:, :-1]
    return torch.cummax(torch.where(first, idx, torch.zeros_like(idx)), dim=1).values.to(torch.int32).contiguous()


class _Attention(Function):
    @staticmethod
    def forward(ctx, q, k, v, seqlens, causal, scale, seg_start=None):
        B, T, Hq, D = q.shape
        Hk = k.shape[2]
        o = torch.empty(B, T, Hq, D, dtype=q.dtype, device=q.device)
        lse = torch.empty(B, Hq, T, dtype=torch.float32, device=q.device)
        N.call("dla_attn_fwd_seg", N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(o), N.ptr(lse), B, T, Hq, Hk, D,
               q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
               o.stride(0), o.stride(1), o.stride(2), N.ptr(seqlens), N.ptr(seg_start), scale, int(causal), _s(q))
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.seqlens, ctx.seg_start = seqlens, seg_start
        ctx.meta = (causal, scale)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        seqlens, seg_start = ctx.seqlens, ctx.seg_start
        causal, scale = ctx.meta
        B, T, Hq, D = q.shape
        Hk = k.shape[2]
        do = _c(do)
        # dq/dk/dv take the strides of q/k/v.  When q, k, v are views of one fused projection the three
        # gradients are views of one packed [B, T, Hq + 2 Hk, D] buffer, so autograd's view backward
        # reassembles d(qkv) without extra copies.
        packed = (q.stride() == k.stride() == v.stride() and q.stride(2) == D and q.stride(1) == (Hq + 2 * Hk) * D
                  and k.data_ptr() - q.data_ptr() == Hq * D * 2 and v.data_ptr() - k.data_ptr() == Hk * D * 2)
        if packed:
            dqkv = torch.empty(B, T, Hq + 2 * Hk, D, dtype=q.dtype, device=q.device)
            dq, dk, dv = dqkv[:, :, :Hq], dqkv[:, :, Hq:Hq + Hk], dqkv[:, :, Hq + Hk:]
        else:
            q, k, v = _c(q), _c(k), _c(v)
            dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        dq_acc = torch.empty(B * Hq * (-(-T // 128) * 128) * D, dtype=torch.float32, device=q.device)   # padded to whole query tiles
        delta = torch.empty(B, Hq, T, dtype=torch.float32, device=q.device)
        assert do.stride(2) == o.stride(2) and do.stride(1) == o.stride(1) and do.stride(0) == o.stride(0)
        N.call("dla_attn_bwd_seg", N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(o), N.ptr(do), N.ptr(lse),
               N.ptr(dq), N.ptr(dk), N.ptr(dv), N.ptr(dq_acc), N.ptr(delta), B, T, Hq, Hk, D,
               q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
               o.stride(0), o.stride(1), o.stride(2), N.ptr(seqlens), N.ptr(seg_start), scale, int(causal), _s(q))
        return dq, dk, dv, None, None, None, None


class _AttentionPacked(Function):
    """Attention on a packed ``[B, T, Hq + 2 Hk, D]`` projection as ONE autograd input: the backward returns the
    packed gradient directly.  With three sliced views as inputs autograd's SliceBackward materialises three
    zero-filled full-size tensors, copies each slice in and adds them up (8 launches and ~0.8 GB of traffic per
    Llama-3-8B layer)."""

    @staticmethod
    def forward(ctx, qkv, n_q, n_kv, seqlens, causal, scale, seg_start=None):
        B, T, _, D = qkv.shape
        q, k, v = qkv[:, :, :n_q], qkv[:, :, n_q:n_q + n_kv], qkv[:, :, n_q + n_kv:]
        o = torch.empty(B, T, n_q, D, dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty(B, n_q, T, dtype=torch.float32, device=qkv.device)
        N.call("dla_attn_fwd_seg", N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(o), N.ptr(lse), B, T, n_q, n_kv, D,
               q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
               o.stride(0), o.stride(1), o.stride(2), N.ptr(seqlens), N.ptr(seg_start), scale, int(causal), _s(qkv))
        ctx.save_for_backward(qkv, o, lse)
        ctx.seqlens, ctx.seg_start = seqlens, seg_start
        ctx.meta = (n_q, n_kv, causal, scale)
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse = ctx.saved_tensors
        seqlens, seg_start = ctx.seqlens, ctx.seg_start
        n_q, n_kv, causal, scale = ctx.meta
        B, T, _, D = qkv.shape
        q, k, v = qkv[:, :, :n_q], qkv[:, :, n_q:n_q + n_kv], qkv[:, :, n_q + n_kv:]
        do = _c(do)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[:, :, :n_q], dqkv[:, :, n_q:n_q + n_kv], dqkv[:, :, n_q + n_kv:]
        dq_acc = torch.empty(B * n_q * (-(-T // 128) * 128) * D, dtype=torch.float32, device=qkv.device)  # padded to whole query tiles
        delta = torch.empty(B, n_q, T, dtype=torch.float32, device=qkv.device)
        N.call("dla_attn_bwd_seg", N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(o), N.ptr(do), N.ptr(lse),
               N.ptr(dq), N.ptr(dk), N.ptr(dv), N.ptr(dq_acc), N.ptr(delta), B, T, n_q, n_kv, D,
               q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
               o.stride(0), o.stride(1), o.stride(2), N.ptr(seqlens), N.ptr(seg_start), scale, int(causal), _s(qkv))
        return dqkv, None, None, None, None, None, None


def attention_qkv(qkv: torch.Tensor, n_q: int, n_kv: int, seqlens: Optional[torch.Tensor] = None, causal: bool = True,
                  scale: Optional[float] = None, segment_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Causal GQA attention on a packed, contiguous ``[B, T, n_q + 2 n_kv, D]`` projection; returns ``[B, T, n_q, D]``."""
    if not qkv.is_cuda or not qkv.is_contiguous() or (segment_ids is not None and not causal):
        return attention(qkv[:, :, :n_q], qkv[:, :, n_q:n_q + n_kv], qkv[:, :, n_q + n_kv:], seqlens=seqlens, causal=causal,
                         scale=scale, segment_ids=segment_ids)
    scale = scale if scale is not None else 1.0 / math.sqrt(qkv.shape[-1])
    _check_bf16(qkv)
    if seqlens is not None:
        seqlens = seqlens.to(torch.int32)
    seg_start = _seg_start_arg(segment_ids, seqlens, qkv.shape[1])
    if seg_start is not None and seqlens is None:
        seqlens = (segment_ids > 0).sum(1).to(torch.int32)
    return _AttentionPacked.apply(qkv, n_q, n_kv, seqlens, causal, scale, seg_start)


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, seqlens: Optional[torch.Tensor] = None,
              causal: bool = True, scale: Optional[float] = None,
              segment_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Causal GQA attention on ``[B, T, H, D]`` tensors (views of a packed QKV are fine: strides are
    passed through).  ``seqlens [B]`` int32 masks right padding."""
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if not q.is_cuda or (segment_ids is not None and not causal):
        return R.attention(q, k, v, causal=causal, seqlens=seqlens, scale=scale, segment_ids=segment_ids)
    _check_bf16(q, k, v)
    if seqlens is not None:
        seqlens = seqlens.to(torch.int32)
    seg_start = _seg_start_arg(segment_ids, seqlens, q.shape[1])
    if seg_start is not None and seqlens is None:
        seqlens = (segment_ids > 0).sum(1).to(torch.int32)
    return _Attention.apply(q, k, v, seqlens, causal, scale, seg_start)


_SEG_CACHE: dict = {}


def _seg_start_arg(segment_ids: Optional[torch.Tensor], seqlens: Optional[torch.Tensor], T: int) -> Optional[torch.Tensor]:
    """The kernels' ``seg_start`` for a batch of packed rows; computed once per batch tensor (every layer of a forward
    and of its recomputation passes the same ``segment_ids`` object)."""
    if segment_ids is None:
        return None
    key = (segment_ids.data_ptr(), segment_ids._version, tuple(segment_ids.shape))
    hit = _SEG_CACHE.get("k")
    if hit is not None and hit[0] == key and hit[1] is segment_ids:
        return hit[2]
    ss = segment_starts(segment_ids)
    _SEG_CACHE["k"] = (key, segment_ids, ss)
    return ss


# =======================================================================================
# fused lm_head -> per-token log-prob
# =======================================================================================
LOGPROB_BWD_CHUNK = 2048


class _LinearLogprob(Function):
    """``logp[m] = log_softmax(h[m] @ W^T)[label[m]]`` without materialising ``[M, V]``.

    Forward: tcgen05 GEMM whose epilogue reduces every 128x256 logits tile to (max, sumexp, picked).
    Backward: per row-chunk, recompute bf16 logits with the GEMM, turn them into
    ``g * (onehot - softmax)`` in place, and feed two more GEMMs (dgrad, wgrad)."""

    @staticmethod
    def forward(ctx, h, w, labels):
        M, K = h.shape
        V = w.shape[0]
        nbytes = int(N.load().dla_lmhead_scratch_bytes(M, V))
        scratch = N.workspace("lmhead", nbytes, h.device)
        logp = torch.empty(M, dtype=torch.float32, device=h.device)
        lse = torch.empty(M, dtype=torch.float32, device=h.device)
        Vp = (V + 7) // 8 * 8
        # memory-for-compute: with HBM to spare (KEEP_LOGITS, set by Engine.auto_no_recompute or DLA_KEEP_LOGITS=1) the
        # epilogue also stores the bf16 logits and the backward skips its recomputation GEMM
        kept = None
        if KEEP_LOGITS[0] and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]):   # (grad mode is off inside forward)
            kept = torch.empty(M, Vp, dtype=BF16, device=h.device)
        N.call("dla_lmhead_logprob_keep", N.ptr(h), N.ptr(w), N.ptr(labels), M, V, K, h.stride(0), w.stride(0),
               N.ptr(scratch), N.ptr(logp), N.ptr(lse), N.ptr(kept), Vp, _s(h))
        if kept is not None:
            ctx.save_for_backward(h, w, labels, lse, kept)
        else:
            ctx.save_for_backward(h, w, labels, lse)
        ctx.kept = kept is not None
        ctx.kept_state = [False]        # set once the backward has turned the kept logits into gradients in place
        ctx.wref = w
        return logp, lse

    @staticmethod
    def backward(ctx, g, _g_lse):
        saved = ctx.saved_tensors
        h, w, labels, lse = saved[:4]
        kept = saved[4] if ctx.kept else None
        if kept is not None:
            if ctx.kept_state[0]:
                raise RuntimeError("linear_logprob: the kept logits were turned into gradients by the first backward; "
                                   "run with DLA_KEEP_LOGITS=0 (F.KEEP_LOGITS[0] = False) to backpropagate twice")
            ctx.kept_state[0] = True
        M, K = h.shape
        V = w.shape[0]
        g = _c(g.float())
        dh = torch.empty_like(h) if ctx.needs_input_grad[0] else None
        slot, acc = _slot_gemm(ctx.wref)
        dw = None
        if ctx.needs_input_grad[1] and slot is None:
            dw = torch.zeros_like(w)
            acc = True
        wtgt = slot if slot is not None else dw
        Vp = (V + 7) // 8 * 8
        chunk = min(LOGPROB_BWD_CHUNK, M)
        if kept is None:
            buf = N.workspace("lmhead_bwd", chunk * Vp * 2, h.device).view(BF16)[: chunk * Vp].view(chunk, Vp)
        for s in range(0, M, chunk):
            e = min(s + chunk, M)
            n = e - s
            if kept is None:
                rows = buf[:n]
                dl = rows.narrow(1, 0, V)           # [n, V] view, row pitch Vp
                gemm(h[s:e], w, out=dl)              # recompute bf16 logits for this row chunk
            else:
                rows = kept[s:e]                     # the forward kept them: transformed in place, chunk by chunk
                dl = rows.narrow(1, 0, V)
            N.call("dla_logprob_grad_inplace", N.ptr(rows), N.ptr(lse[s:e]), N.ptr(labels[s:e]), N.ptr(g[s:e]),
                   n, V, Vp, _s(h))
            if dh is not None:
                gemm(dl, w, b_mn=True, out=dh[s:e])
            if wtgt is not None:
                gemm(dl, h[s:e], a_mn=True, b_mn=True, out=wtgt, accumulate=acc)
                acc = True               # later row chunks add to the first one
        return dh, dw, None


KEEP_LOGITS = [_os.environ.get("DLA_KEEP_LOGITS", "0") == "1"]


def linear_logprob(h: torch.Tensor, w: torch.Tensor, labels: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-token label log-prob and log-sum-exp (fp32 ``[M]``) of ``h [M, K] @ w[V, K]^T``.
    Labels < 0 are ignored (log-prob 0, no gradient)."""
    if not h.is_cuda:
        return R.token_logprobs(h, w, labels)
    _check_bf16(h, w)
    return _LinearLogprob.apply(_c(h), w, _c(labels.to(torch.int64)))


# =======================================================================================
# sequence reductions and losses
# =======================================================================================

class _SeqMean(Function):
    @staticmethod
    def forward(ctx, tok, mask, use_sum):
        B, T = tok.shape
        tok, mask = _c(tok.float()), _c(mask.to(torch.int64))
        out = torch.empty(B, dtype=torch.float32, device=tok.device)
        N.call("dla_seq_mean_fwd", N.ptr(tok), N.ptr(mask), N.ptr(out), B, T, int(use_sum), _s(tok))
        ctx.save_for_backward(mask)
        ctx.meta = (B, T, use_sum)
        return out

    @staticmethod
    def backward(ctx, g):
        (mask,) = ctx.saved_tensors
        B, T, use_sum = ctx.meta
        g = _c(g.float())
        dtok = torch.empty(B, T, dtype=torch.float32, device=g.device)
        N.call("dla_seq_mean_bwd", N.ptr(g), N.ptr(mask), N.ptr(dtok), B, T, int(use_sum), _s(g))
        return dtok, None, None


def sequence_logprob(token_logp: torch.Tensor, mask: torch.Tensor, reduction: str = "mean") -> torch.Tensor:
    """Per-sequence masked mean (reference semantics, ``R/src/training/train_dpo.py:38``) or sum."""
    if not token_logp.is_cuda:
        if reduction == "sum":
            return (token_logp * mask.to(token_logp.dtype)).sum(1)
        return R.sequence_mean_logprob(token_logp, mask)
    return _SeqMean.apply(token_logp, mask, reduction == "sum")


class _DPOLoss(Function):
    @staticmethod
    def forward(ctx, pp, pn, rp, rn, beta, ls, scale):
        B = pp.numel()
        pp, pn, rp, rn = (_c(t.float()) for t in (pp, pn, rp, rn))
        out = torch.empty(3, dtype=torch.float32, device=pp.device)
        dpp = torch.empty(B, dtype=torch.float32, device=pp.device)
        dpn = torch.empty(B, dtype=torch.float32, device=pp.device)
        N.call("dla_dpo_loss", N.ptr(pp), N.ptr(pn), N.ptr(rp), N.ptr(rn), B, beta, ls, scale, N.ptr(out),
               N.ptr(dpp), N.ptr(dpn), _s(pp))
        ctx.save_for_backward(dpp, dpn)
        return out[0], out[1:].detach()

    @staticmethod
    def backward(ctx, g, _gs):
        dpp, dpn = ctx.saved_tensors
        return dpp * g, dpn * g, None, None, None, None, None


def dpo_loss(pol_pos, pol_neg, ref_pos, ref_neg, beta: float, label_smoothing: float = 0.0,
             scale: Optional[float] = None):
    """Returns ``(loss, stats)`` with ``stats = [preference_rate, mean_margin]``.  ``scale`` defaults to
    ``1/B`` (batch mean, ``R/src/training/train_dpo.py:44``)."""
    B = pol_pos.numel()
    scale = 1.0 / B if scale is None else scale
    if not pol_pos.is_cuda:
        z = beta * ((pol_pos - pol_neg) - (ref_pos - ref_neg))
        loss = R.dpo_loss(pol_pos, pol_neg, ref_pos, ref_neg, beta, label_smoothing) * (scale * B)
        stats = torch.stack([(pol_pos > pol_neg).float().sum() * scale, z.detach().sum() * scale])
        return loss, stats
    return _DPOLoss.apply(pol_pos, pol_neg, ref_pos.detach(), ref_neg.detach(), beta, label_smoothing, scale)


class _PairwiseLoss(Function):
    @staticmethod
    def forward(ctx, c, r, scale):
        B = c.numel()
        c, r = _c(c.float()), _c(r.float())
        out = torch.empty(2, dtype=torch.float32, device=c.device)
        dc = torch.empty(B, dtype=torch.float32, device=c.device)
        dr = torch.empty(B, dtype=torch.float32, device=c.device)
        N.call("dla_pairwise_loss", N.ptr(c), N.ptr(r), B, scale, N.ptr(out), N.ptr(dc), N.ptr(dr), _s(c))
        ctx.save_for_backward(dc, dr)
        return out[0], out[1].detach()

    @staticmethod
    def backward(ctx, g, _ga):
        dc, dr = ctx.saved_tensors
        return dc * g, dr * g, None


def pairwise_loss(chosen: torch.Tensor, rejected: torch.Tensor, scale: Optional[float] = None):
    """Bradley-Terry loss; returns ``(loss, accuracy)`` (``R/src/models/reward_model.py:67-68``)."""
    B = chosen.numel()
    scale = 1.0 / B if scale is None else scale
    if not chosen.is_cuda:
        return R.pairwise_loss(chosen, rejected) * (scale * B), (chosen > rejected).float().sum() * scale
    return _PairwiseLoss.apply(chosen, rejected, scale)


# =======================================================================================
# scalar heads (reward / value), RL objectives, distillation KL
# =======================================================================================

HEAD_MODES = {"last_token": 0, "mean": 1, "token": 2}


def _grad_to(param: torch.Tensor, g32: torch.Tensor) -> Optional[torch.Tensor]:
    """Route an fp32 parameter gradient: into the engine's bucket view when the parameter has one."""
    slot = _slot_add(param)
    if slot is not None:
        slot.add_(g32.view_as(slot).to(slot.dtype))
        return None
    return g32.view_as(param).to(param.dtype)


class _ScalarHead(Function):
    @staticmethod
    def forward(ctx, hidden, mask, w, bias, keep, mode):
        B, T, H = hidden.shape
        hidden = _c(hidden)
        mask = _c(mask.to(torch.int64)) if mask is not None else None
        keep = _c(keep.float()) if keep is not None else None
        out = torch.empty((B, T) if mode == 2 else (B,), dtype=torch.float32, device=hidden.device)
        N.call("dla_scalar_head_fwd", N.ptr(hidden), N.ptr(mask) if mask is not None else None, N.ptr(w),
               N.ptr(bias) if bias is not None else None, N.ptr(keep) if keep is not None else None, N.ptr(out), B, T, H, mode,
               _s(hidden))
        ctx.save_for_backward(hidden, mask, w, keep)
        ctx.refs = (w, bias)
        ctx.mode = mode
        return out

    @staticmethod
    def backward(ctx, dout):
        hidden, mask, w, keep = ctx.saved_tensors
        wref, bref = ctx.refs
        B, T, H = hidden.shape
        dout = _c(dout.float())
        dh = torch.empty_like(hidden) if ctx.needs_input_grad[0] else None
        need_w = ctx.needs_input_grad[2]
        need_b = bref is not None and ctx.needs_input_grad[3]
        dw32 = torch.zeros(H, dtype=torch.float32, device=hidden.device) if need_w else None
        db32 = torch.zeros(1, dtype=torch.float32, device=hidden.device) if need_b else None
        N.call("dla_scalar_head_bwd", N.ptr(hidden), N.ptr(mask) if mask is not None else None, N.ptr(w),
               N.ptr(keep) if keep is not None else None, N.ptr(dout), N.ptr(dh) if dh is not None else None,
               N.ptr(dw32) if need_w else None, N.ptr(db32) if need_b else None, B, T, H, ctx.mode, _s(hidden))
        dw = _grad_to(wref, dw32) if need_w else None
        db = _grad_to(bref, db32) if need_b else None
        return dh, None, dw, db, None, None


def scalar_head(hidden: torch.Tensor, mask: Optional[torch.Tensor], w: torch.Tensor, bias: Optional[torch.Tensor],
                keep: Optional[torch.Tensor] = None, mode: str = "last_token") -> torch.Tensor:
    """Reward head (``mode`` ``last_token`` / ``mean`` -> ``[B]``) or value head (``token`` -> ``[B, T]``), fp32 out.
    ``w`` is the ``[1, H]`` (or ``[H]``) row of ``Linear(H, 1)``; ``keep`` the scaled dropout mask ``[B, H]``."""
    if not hidden.is_cuda:
        return R.scalar_head(hidden, mask, w, bias, keep, mode)
    _check_bf16(hidden, w)
    return _ScalarHead.apply(hidden, mask, w, bias, keep, HEAD_MODES[mode])


class _Reinforce(Function):
    @staticmethod
    def forward(ctx, pol, ref, score, kl_coef, baseline):
        B = pol.numel()
        pol, ref, score = (_c(t.float()) for t in (pol, ref, score))
        base = _c(baseline.float().reshape(1)) if baseline is not None else None
        out = torch.empty(3, dtype=torch.float32, device=pol.device)
        kl, rew, dpol = (torch.empty(B, dtype=torch.float32, device=pol.device) for _ in range(3))
        N.call("dla_reinforce_loss", N.ptr(pol), N.ptr(ref), N.ptr(score), B, float(kl_coef), N.ptr(base) if base is not None else None,
               N.ptr(out), N.ptr(kl), N.ptr(rew), N.ptr(dpol), _s(pol))
        ctx.save_for_backward(dpol)
        ctx.mark_non_differentiable(kl, rew)
        return out[0], kl, rew

    @staticmethod
    def backward(ctx, g, _gk, _gr):
        (dpol,) = ctx.saved_tensors
        return dpol * g, None, None, None, None


def reinforce_loss(policy_logp: torch.Tensor, ref_logp: torch.Tensor, scores: torch.Tensor, kl_coef: float,
                   baseline: Optional[torch.Tensor] = None):
    """``(loss, kl [B], rewards [B])`` of the reference's sequence-level REINFORCE (``R/src/training/train_rlhf.py:149-153``);
    ``kl`` / ``rewards`` are detached statistics, the gradient flows through the loss only (as in the reference)."""
    if not policy_logp.is_cuda:
        loss, kl, rew = R.reinforce_loss(policy_logp, ref_logp, scores, kl_coef, baseline)
        return loss, kl.detach(), rew.detach()
    return _Reinforce.apply(policy_logp, ref_logp.detach(), scores.detach(), kl_coef, baseline)


class _PPOLoss(Function):
    @staticmethod
    def forward(ctx, new_lp, old_lp, adv, values, old_values, returns, mask, clip, vclip, vf_coef):
        n = new_lp.numel()
        f = lambda t: _c(t.float()) if t is not None else None
        new_c, old_c, adv_c, val_c, oval_c, ret_c, m_c = (f(t) for t in (new_lp, old_lp, adv, values, old_values, returns, mask))
        out = torch.empty(5, dtype=torch.float32, device=new_lp.device)
        dlp = torch.empty_like(new_c)
        dv = torch.empty_like(val_c) if val_c is not None else None
        q = lambda t: N.ptr(t) if t is not None else None
        N.call("dla_ppo_loss", q(new_c), q(old_c), q(adv_c), q(val_c), q(oval_c), q(ret_c), q(m_c), n, float(clip), float(vclip),
               float(vf_coef), N.ptr(out), N.ptr(dlp), q(dv), _s(new_c))
        ctx.save_for_backward(dlp, dv if dv is not None else dlp)
        ctx.has_v = dv is not None
        ctx.shapes = (new_lp.shape, values.shape if values is not None else None)
        stats = out[1:].clone()
        ctx.mark_non_differentiable(stats)
        return out[0], stats

    @staticmethod
    def backward(ctx, g, _gs):
        dlp, dv = ctx.saved_tensors
        return ((dlp * g).view(ctx.shapes[0]), None, None, (dv * g).view(ctx.shapes[1]) if ctx.has_v else None, None, None, None,
                None, None, None)


def ppo_loss(new_logp, old_logp, advantages, values, old_values, returns, mask, clip: float = 0.2, vclip: float = 0.2,
             vf_coef: float = 0.5):
    """Clipped PPO + clipped value loss, masked token mean.  Returns ``(loss, stats)`` with
    ``stats = [pg_loss, vf_loss, approx_kl, clipfrac]`` (detached)."""
    if not new_logp.is_cuda:
        return R.ppo_loss(new_logp, old_logp, advantages, values, old_values, returns, mask, clip, vclip, vf_coef)
    return _PPOLoss.apply(new_logp, old_logp.detach(), advantages.detach(), values, None if old_values is None else old_values.detach(),
                          None if returns is None else returns.detach(), mask, clip, vclip, vf_coef)


def gae(rewards: torch.Tensor, values: torch.Tensor, mask: torch.Tensor, gamma: float, lam: float, whiten: bool = False):
    """``(advantages, returns)`` over ``[B, T]``; ``whiten`` normalises the advantages over the masked tokens."""
    if not rewards.is_cuda:
        return R.gae(rewards, values, mask, gamma, lam, whiten)
    B, T = rewards.shape
    r, v, m = (_c(t.float()) for t in (rewards, values, mask))
    adv, ret = torch.empty_like(r), torch.empty_like(r)
    N.call("dla_gae", N.ptr(r), N.ptr(v), N.ptr(m), N.ptr(adv), N.ptr(ret), B, T, float(gamma), float(lam), int(whiten), _s(r))
    return adv, ret


def distill_kl_rows(student_logits: torch.Tensor, teacher_logits, coef: Optional[torch.Tensor], temperature: float,
                    want_kl: bool = True, want_grad: bool = False):
    """Row-wise ensemble KL on fp32 logits ``[rows, V]`` (unit inner stride): ``(kl [rows] | None, dl bf16 [rows, V] | None)``
    with ``dl = (softmax(s/T) - mean_k softmax(t_k/T)) * coef / T``.  CUDA only; see ``csrc/distill.cu``."""
    import ctypes

    rows, V = student_logits.shape
    K = len(teacher_logits)
    assert student_logits.dtype == torch.float32 and all(t.dtype == torch.float32 and t.shape == student_logits.shape for t in teacher_logits)
    assert student_logits.stride(1) == 1 and all(t.stride(1) == 1 for t in teacher_logits)
    ptrs = (ctypes.c_void_p * K)(*[t.data_ptr() for t in teacher_logits])
    lds = (ctypes.c_int64 * K)(*[t.stride(0) for t in teacher_logits])
    kl = torch.empty(rows, dtype=torch.float32, device=student_logits.device) if want_kl else None
    dl = torch.empty(rows, V, dtype=BF16, device=student_logits.device) if want_grad else None
    cf = _c(coef.float()) if coef is not None else None
    N.call("dla_distill_kl", N.ptr(student_logits), student_logits.stride(0), ctypes.cast(ptrs, ctypes.c_void_p),
           ctypes.cast(lds, ctypes.c_void_p), K, N.ptr(cf) if cf is not None else None, rows, V, float(temperature),
           N.ptr(kl) if kl is not None else None, N.ptr(dl) if dl is not None else None, V, _s(student_logits))
    return kl, dl
