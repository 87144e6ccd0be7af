"""ctypes binding of ``libdla_b200.so`` (C ABI declared in ``include/dla_b200.h``).

The library is built in-tree by :mod:`dla_b200.ops.build`.  On a machine with a CUDA device the
native path is mandatory: a missing or stale library raises instead of silently falling back to
eager PyTorch (the CPU path exists only for tensors that live on the CPU).
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_int64, c_longlong, c_size_t, c_uint32, c_uint64, c_void_p
from pathlib import Path
from typing import Dict, Optional, Tuple

import torch

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent / "libdla_b200.so"

P, I, I64, F = c_void_p, c_int, c_int64, c_float

_SIGNATURES = {
    "dla_rmsnorm_fwd": [P, P, P, P, P, P, I64, I, F, P],
    "dla_rmsnorm_bwd": [P, P, P, P, P, P, P, P, I, I64, I, P],
    "dla_layernorm_fwd": [P, P, P, P, P, I64, I, F, P],
    "dla_layernorm_bwd": [P, P, P, P, P, P, P, P, I, I64, I, P],
    "dla_rope_inplace": [P, P, P, P, I64, I, I, I, I64, F, P],
    "dla_rope_copy": [P, P, P, P, P, I64, I, I, I, I64, I64, F, P],
    "dla_swiglu_fwd": [P, P, I64, I, P],
    "dla_swiglu_bwd": [P, P, P, I64, I, P],
    "dla_gelu_fwd": [P, P, P, I64, I, P],
    "dla_gelu_bwd": [P, P, P, P, I64, I, P],
    "dla_add_bias": [P, P, P, P, I64, I, P],
    "dla_bias_grad": [P, P, P, I, I64, I, P],
    "dla_embedding_fwd": [P, P, P, P, I64, I, I, I, P],
    "dla_embedding_bwd": [P, P, P, P, I64, I, I, I, P],
    "dla_adamw": [P, P, P, P, P, I64, F, F, F, F, F, F, F, P, F, I, P],
    "dla_sumsq": [P, I64, I, P, P, I, P],
    "dla_clip_coef": [P, F, P, P, P],
    "dla_cast": [P, P, I64, I, I, F, P],
    "dla_dpo_loss": [P, P, P, P, I, F, F, F, P, P, P, P],
    "dla_pairwise_loss": [P, P, I, F, P, P, P, P],
    "dla_seq_mean_fwd": [P, P, P, I, I, I, P],
    "dla_seq_mean_bwd": [P, P, P, I, I, I, P],
    "dla_gemm_bf16": [P, P, P, I, I, I, I64, I64, I64, I, I, I, I, P],
    "dla_gemm_bias_bf16": [P, P, P, P, I, I, I, I64, I64, I64, I, P],
    "dla_gemm_rope_bf16": [P, P, P, I, I, I, I64, I64, I64, P, P, P, I, P],
    "dla_gemm_set_variant": [I],
    "dla_gemm_swiglu_bf16": [P, P, P, P, I, I, I, I64, I64, I64, I64, P],
    "dla_gemm_swiglu_bwd_bf16": [P, P, P, P, I, I, I, I64, I64, I64, I64, P],
    "dla_set_device": [I],
    "dla_lmhead_logprob": [P, P, P, I, I, I, I64, I64, P, P, P, P],
    "dla_lmhead_logprob_keep": [P, P, P, I, I, I, I64, I64, P, P, P, P, I64, P],
    "dla_logprob_grad_inplace": [P, P, P, P, I, I, I64, P],
    "dla_attn_fwd": [P, P, P, P, P, I, I, I, I, I] + [I64] * 9 + [P, F, I, P],
    "dla_attn_bwd": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I] + [I64] * 9 + [P, F, I, P],
    "dla_attn_fwd_seg": [P, P, P, P, P, I, I, I, I, I] + [I64] * 9 + [P, P, F, I, P],
    "dla_attn_bwd_seg": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I] + [I64] * 9 + [P, P, F, I, P],
    "dla_attn_decode": [P, P, P, P, P, I, I, I, I, I, F, P],
    "dla_gemm_skinny_bf16": [P, P, P, P, P, I, I, I, I64, I64, I64, I64, I, I, P],
    "dla_decode_attn_fused": [P, I64, P, P, P, P, P, P, I, I, I, I, I, I, F, P],
    "dla_scalar_head_fwd": [P, P, P, P, P, P, I, I, I, I, P],
    "dla_scalar_head_bwd": [P, P, P, P, P, P, P, P, I, I, I, I, P],
    "dla_reinforce_loss": [P, P, P, I, F, P, P, P, P, P, P],
    "dla_ppo_loss": [P, P, P, P, P, P, P, I, F, F, F, P, P, P, P],
    "dla_gae": [P, P, P, P, P, I, I, F, F, I, P],
    "dla_distill_kl": [P, I64, P, P, I, P, I, I, F, P, P, I64, P],
    "dla_attn_set_fwd_variant": [I],
    "dla_attn_set_bwd_variant": [I],
    "dla_peer_allgather_small": [P, P, I, I, P, I, I, c_uint32, I, P, P],
    "dla_sample_top_p": [P, I, I, I, I64, F, F, c_uint64, c_uint64, P, P, P],
    "dla_fused_rs_adamw": [P, P, I, I, I64, I64, F, P, P, P, P, P, F, F, F, F, F, F, F, P, I, P],
    "dla_peer_barrier": [P, I, I, I, c_uint32, I, P],
    "dla_peer_push": [P, P, I, I, I64, I64, P],
    "dla_peer_gather": [P, P, I64, I64, I64, I, P],
    "dla_peer_allreduce": [P, I, I, I64, I64, F, P],
    "dla_ag_gemm_bf16": [P, P, P, I64, I64, P, I, I, I, I64, I64, I, I, P, c_uint32, I, I, P],
    "dla_fused_gather_dpo": [P, P, I, I, P, P, P, P, I, F, F, c_uint32, P, P, P, P],
    "dla_adamw_mc": [P, P, P, P, P, I64, F, F, F, F, F, F, F, P, F, I, P],
    "dla_unit_reduce": [P, I, P, P, I, I, I64, I64, F, P, P, c_uint32, I, I, P],
    "dla_unit_reduce_acc": [P, I, P, P, I, I, I64, I64, F, P, P, c_uint32, I, I, I, P],
    "dla_peer_flag_row": [P, I, I, I, c_uint32, I, I, P],
    "dla_peer_norm_clip": [P, P, I, I, P, F, F, c_uint32, I, P, P, P],
    "dla_peer_push_mc": [P, P, I64, I64, P],
    "dla_peer_allreduce_mc": [P, I, I, I64, I64, F, P],
    "dla_peer_watch": [P],
    "dla_fused_gather_reinforce": [P, P, I, I, P, P, P, I, F, c_uint32, P, P, P],
    "dla_fused_gather_whiten": [P, P, I, I, P, P, I, P, I, c_uint32, P, P],
}


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def load() -> ctypes.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"native library {_LIB_PATH} is missing — run `python -m dla_b200.ops.build` "
            "(CUDA tensors have no eager fallback in this framework)")
    lib = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
    lib.dla_last_error.restype = ctypes.c_char_p
    lib.dla_launch_count.restype = c_longlong
    for name in ("dla_lmhead_scratch_bytes", "dla_sampler_scratch_bytes"):
        if hasattr(lib, name):
            getattr(lib, name).restype = c_size_t
            getattr(lib, name).argtypes = [I, I]
    for name, sig in _SIGNATURES.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = sig
            fn.restype = c_int
    _LIB = lib
    return lib


def has(name: str) -> bool:
    return hasattr(load(), name)


def launch_count() -> int:
    return int(load().dla_launch_count()) if _LIB is not None or available() else 0


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device: Optional[torch.device] = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


_DEVICE_BOUND = False


def bind_device(device: Optional[int] = None) -> None:
    """Tell the library which GPU this process drives (one process per GPU)."""
    global _DEVICE_BOUND
    lib = load()
    dev = torch.cuda.current_device() if device is None else int(device)
    rc = lib.dla_set_device(dev)
    if rc != 0:
        raise RuntimeError(f"dla_set_device({dev}) failed: {lib.dla_last_error().decode()}")
    _DEVICE_BOUND = True


def call(name: str, *args) -> None:
    """Invoke a ``dla_*`` entry point; raises RuntimeError with the library's message on failure."""
    lib = load()
    if not _DEVICE_BOUND:
        bind_device()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.dla_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{name} failed (rc={rc}): {msg}")


# ---------------------------------------------------------------------------------------
# scratch workspaces (cached per device / stream so hot loops do not allocate)
# ---------------------------------------------------------------------------------------
_WORKSPACES: Dict[Tuple[str, int, int], torch.Tensor] = {}


def workspace(tag: str, nbytes: int, device: torch.device) -> torch.Tensor:
    """A cached uint8 device buffer of at least ``nbytes``, keyed by tag + device + stream."""
    key = (tag, device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream)
    buf = _WORKSPACES.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = buf
    return buf


def norm_bwd_blocks() -> int:
    return int(load().dla_norm_bwd_blocks())


def sumsq_blocks() -> int:
    return int(load().dla_sumsq_blocks())


def require_cuda_native(t: torch.Tensor) -> bool:
    """True when ``t`` should go through the native kernels (it lives on a CUDA device)."""
    return t.is_cuda
