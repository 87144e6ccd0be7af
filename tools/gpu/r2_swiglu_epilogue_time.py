"""CUDA-event time of the two SwiGLU-epilogue GEMMs at Llama-3-8B sizes (8192 tokens)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from dla_b200.ops import native as N  # noqa: E402

dev = "cuda"
M, H, Fi = 8192, 4096, 14336
r = lambda *s, sc=1.0: (torch.randn(*s, device=dev) * sc).to(torch.bfloat16)
x, w_gu, w_down = r(M, H), r(2 * Fi, H, sc=0.02), r(H, Fi, sc=0.02)
gu, act, dy, dgu = torch.empty(M, 2 * Fi, dtype=torch.bfloat16, device=dev), torch.empty(M, Fi, dtype=torch.bfloat16, device=dev), r(M, H), torch.empty(M, 2 * Fi, dtype=torch.bfloat16, device=dev)
s = N.stream_ptr(torch.device(dev))


def fwd():
    N.call("dla_gemm_swiglu_bf16", N.ptr(x), N.ptr(w_gu), N.ptr(gu), N.ptr(act), M, Fi, H, x.stride(0), w_gu.stride(0), gu.stride(0), act.stride(0), s)


def bwd():
    N.call("dla_gemm_swiglu_bwd_bf16", N.ptr(dy), N.ptr(w_down), N.ptr(gu), N.ptr(dgu), M, Fi, H, dy.stride(0), w_down.stride(0), gu.stride(0), dgu.stride(0), s)


for name, fn, flop in (("gate/up + swiglu (kFuse 1)", fwd, 2 * M * 2 * Fi * H), ("down dgrad + swiglu bwd (kFuse 2)", bwd, 2 * M * Fi * H)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"{name}: {ms * 1e3:.1f} us, {flop / ms / 1e9:.0f} TFLOP/s")
