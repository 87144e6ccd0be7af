#!/usr/bin/env python
"""Per-kernel timing on one B200 with CUDA events (warm-up, L2 flush between iterations) and roofline
fractions against MEASURED_PEAKS.json.  Writes JSON lines to gpurun_out/kernels.jsonl."""
from __future__ import annotations

import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from dla_b200.obs import measured_peaks  # noqa: E402
from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.ops import native as N  # noqa: E402

DEV = "cuda"
PK = measured_peaks()
FLUSH = None


def timeit(fn, iters=10, warm=3, flush=True):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            FLUSH.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def rnd(*s):
    return torch.randn(*s, device=DEV, dtype=torch.bfloat16)


def main():
    out = []
    which = sys.argv[1:] or ["gemm", "attn", "elem", "logprob"]
    if "gemm" in which:
        shapes = [(8192, 6144, 4096, "qkv"), (8192, 4096, 4096, "wo"), (8192, 28672, 4096, "gate_up"), (8192, 4096, 14336, "down"),
                  (8192, 8192, 8192, "square"), (4096, 128256, 4096, "lm_head"), (2048, 4096, 4096, "small_m")]
        for M, Nn, K, tag in shapes:
            a, b = rnd(M, K), rnd(Nn, K)
            c = torch.empty(M, Nn, device=DEV, dtype=torch.bfloat16)
            flops = 2.0 * M * Nn * K
            t_own = timeit(lambda: F.gemm(a, b, out=c))
            N.call("dla_gemm_set_variant", 2)
            t_2cta = timeit(lambda: F.gemm(a, b, out=c))
            N.call("dla_gemm_set_variant", 5)            # last-wave K split off
            t_2cta_notail = timeit(lambda: F.gemm(a, b, out=c))
            N.call("dla_gemm_set_variant", 6)
            N.call("dla_gemm_set_variant", 0)
            t_lib = timeit(lambda: torch.matmul(a, b.t(), out=c))
            rec = {"kernel": "gemm_nt", "tag": tag, "M": M, "N": Nn, "K": K, "ms": t_own, "tflops": flops / t_own / 1e9, "tflops_2cta": flops / t_2cta / 1e9, "tflops_2cta_no_tail_split": flops / t_2cta_notail / 1e9,
                   "cublas_ms": t_lib, "cublas_tflops": flops / t_lib / 1e9,
                   "frac_of_measured_peak": flops / t_own / 1e9 / PK["bf16_tflops"]}
            # dgrad (B mn-major) and wgrad (both mn-major) of the same Linear
            dy = rnd(M, Nn)
            dx = torch.empty(M, K, device=DEV, dtype=torch.bfloat16)
            dw = torch.empty(Nn, K, device=DEV, dtype=torch.bfloat16)
            rec["dgrad_tflops"] = flops / timeit(lambda: F.gemm(dy, b, b_mn=True, out=dx)) / 1e9
            rec["wgrad_tflops"] = flops / timeit(lambda: F.gemm(dy, a, a_mn=True, b_mn=True, out=dw)) / 1e9
            N.call("dla_gemm_set_variant", 2)
            rec["dgrad_tflops_2cta"] = flops / timeit(lambda: F.gemm(dy, b, b_mn=True, out=dx)) / 1e9
            rec["wgrad_tflops_2cta"] = flops / timeit(lambda: F.gemm(dy, a, a_mn=True, b_mn=True, out=dw)) / 1e9
            N.call("dla_gemm_set_variant", 5)
            rec["dgrad_tflops_2cta_no_tail_split"] = flops / timeit(lambda: F.gemm(dy, b, b_mn=True, out=dx)) / 1e9
            rec["wgrad_tflops_2cta_no_tail_split"] = flops / timeit(lambda: F.gemm(dy, a, a_mn=True, b_mn=True, out=dw)) / 1e9
            N.call("dla_gemm_set_variant", 6)
            N.call("dla_gemm_set_variant", 0)
            out.append(rec)
            print(json.dumps(rec), flush=True)
    if "smallm" in which:
        for M in (128, 256, 512, 1024):
            for Nn, K, tag in [(6144, 4096, "qkv"), (4096, 4096, "wo"), (28672, 4096, "gate_up"), (4096, 14336, "down")]:
                a, b = rnd(M, K), rnd(Nn, K)
                c = torch.empty(M, Nn, device=DEV, dtype=torch.bfloat16)
                flops = 2.0 * M * Nn * K
                N.call("dla_gemm_set_variant", 4)
                t_on = timeit(lambda: F.gemm(a, b, out=c))
                N.call("dla_gemm_set_variant", 3)
                t_off = timeit(lambda: F.gemm(a, b, out=c))
                N.call("dla_gemm_set_variant", 4)
                t_lib = timeit(lambda: torch.matmul(a, b.t(), out=c))
                rec = {"kernel": "gemm_small_m", "tag": tag, "M": M, "N": Nn, "K": K, "ms_splitk": t_on, "ms_no_split": t_off, "cublas_ms": t_lib,
                       "tflops": flops / t_on / 1e9, "weight_gbs": 2.0 * Nn * K / t_on / 1e6}
                out.append(rec)
                print(json.dumps(rec), flush=True)
    if "skinny" in which:
        for M in (1, 8, 16):
            for Nn, K, tag in [(6144, 4096, "qkv"), (4096, 4096, "wo"), (28672, 4096, "gate_up"), (4096, 14336, "down"), (128256, 4096, "lm_head")]:
                a, b = rnd(M, K), rnd(Nn, K)
                c = torch.empty(M, Nn, device=DEV, dtype=torch.bfloat16)
                byt = 2.0 * (Nn * K + M * K + M * Nn)
                t_own = timeit(lambda: F.gemm(a, b, out=c))
                t_tile = timeit(lambda: N.call("dla_gemm_bf16", N.ptr(a), N.ptr(b), N.ptr(c), M, Nn, K, K, K, Nn, 0, 0, 0, 0, N.stream_ptr()))
                t_lib = timeit(lambda: torch.matmul(a, b.t(), out=c))
                rec = {"kernel": "gemm_skinny", "tag": tag, "M": M, "N": Nn, "K": K, "ms": t_own, "gbs": byt / t_own / 1e6,
                       "frac_hbm": byt / t_own / 1e6 / PK["hbm_gbs"], "tile_kernel_ms": t_tile, "cublas_ms": t_lib}
                out.append(rec)
                print(json.dumps(rec), flush=True)
    if "attn" in which:
        for B, T, Hq, Hk, D in [(8, 1024, 32, 8, 128), (2, 4096, 32, 8, 128)]:
            qkv = rnd(B, T, Hq + 2 * Hk, D).requires_grad_()
            q, k, v = qkv[:, :, :Hq], qkv[:, :, Hq:Hq + Hk], qkv[:, :, Hq + Hk:]
            fl = 4.0 * B * Hq * T * T * D / 2
            N.call("dla_attn_set_fwd_variant", 1)
            t_f1 = timeit(lambda: F.attention(q.detach(), k.detach(), v.detach()))
            N.call("dla_attn_set_fwd_variant", 3)
            t_f3 = timeit(lambda: F.attention(q.detach(), k.detach(), v.detach()))
            N.call("dla_attn_set_fwd_variant", 0)
            t_f = timeit(lambda: F.attention(q.detach(), k.detach(), v.detach()))
            o = F.attention(q, k, v)
            g = rnd(B, T, Hq, D)
            N.call("dla_attn_set_bwd_variant", 1)
            t_b1 = timeit(lambda: torch.autograd.grad(o, qkv, g, retain_graph=True))
            N.call("dla_attn_set_bwd_variant", 2)
            t_b2 = timeit(lambda: torch.autograd.grad(o, qkv, g, retain_graph=True))
            N.call("dla_attn_set_bwd_variant", 0)
            t_b = timeit(lambda: torch.autograd.grad(o, qkv, g, retain_graph=True))
            # library arms on the same box: PyTorch SDPA (cuDNN / flash backends pick themselves) and flash-attn 2
            qs, ks, vs = (t.detach().transpose(1, 2).contiguous().requires_grad_() for t in (q, k, v))
            sdpa = lambda: torch.nn.functional.scaled_dot_product_attention(qs, ks, vs, is_causal=True, enable_gqa=True)
            t_sdpa = timeit(lambda: sdpa())
            os_ = sdpa()
            gs = g.transpose(1, 2).contiguous()
            t_sdpa_b = timeit(lambda: torch.autograd.grad(os_, (qs, ks, vs), gs, retain_graph=True))
            t_fa = t_fa_b = None
            try:
                from flash_attn import flash_attn_func

                qf, kf, vf = (t.detach().contiguous().requires_grad_() for t in (q, k, v))
                t_fa = timeit(lambda: flash_attn_func(qf, kf, vf, causal=True))
                of = flash_attn_func(qf, kf, vf, causal=True)
                t_fa_b = timeit(lambda: torch.autograd.grad(of, (qf, kf, vf), g, retain_graph=True))
            except Exception as e:  # not built for this arch / not importable
                print(f"# flash_attn unavailable: {type(e).__name__}: {str(e)[:100]}", flush=True)
            tf = lambda ms, mult=1.0: (mult * fl / ms / 1e9) if ms else None
            rec = {"kernel": "attention", "B": B, "T": T, "Hq": Hq, "Hk": Hk, "D": D, "fwd_ms": t_f, "fwd_tflops": tf(t_f),
                   "fwd_gen1_ms": t_f1, "fwd_gen3_ms": t_f3, "bwd_ms": t_b, "bwd_tflops": tf(t_b, 2.5), "bwd_gen1_ms": t_b1,
                   "bwd_gen1_tflops": tf(t_b1, 2.5), "bwd_gen2_ms": t_b2, "bwd_gen2_tflops": tf(t_b2, 2.5), "sdpa_fwd_ms": t_sdpa, "sdpa_fwd_tflops": tf(t_sdpa), "sdpa_bwd_ms": t_sdpa_b,
                   "sdpa_bwd_tflops": tf(t_sdpa_b, 2.5), "flash_attn2_fwd_ms": t_fa, "flash_attn2_fwd_tflops": tf(t_fa),
                   "flash_attn2_bwd_ms": t_fa_b, "flash_attn2_bwd_tflops": tf(t_fa_b, 2.5),
                   "note": "own bwd includes the delta / zeroing pre-pass and the fp32 -> bf16 dQ cast"}
            out.append(rec)
            print(json.dumps(rec), flush=True)
    if "elem" in which:
        M, H, Fi = 8192, 4096, 14336
        x, d, w = rnd(M, H), rnd(M, H), rnd(H)
        t = timeit(lambda: F.add_rmsnorm(x, d, w, 1e-5))
        rec = {"kernel": "add_rmsnorm_fwd", "M": M, "H": H, "ms": t, "gbs": 4 * M * H * 2 / t / 1e6, "frac_hbm": 4 * M * H * 2 / t / 1e6 / PK["hbm_gbs"]}
        out.append(rec); print(json.dumps(rec), flush=True)
        gu = rnd(M, 2 * Fi)
        t = timeit(lambda: F.swiglu(gu))
        rec = {"kernel": "swiglu_fwd", "M": M, "F": Fi, "ms": t, "gbs": 3 * M * Fi * 2 / t / 1e6, "frac_hbm": 3 * M * Fi * 2 / t / 1e6 / PK["hbm_gbs"]}
        out.append(rec); print(json.dumps(rec), flush=True)
        n = 218_112_000
        master, m_, v_ = (torch.zeros(n, device=DEV) for _ in range(3))
        g, p = torch.zeros(n, device=DEV, dtype=torch.bfloat16), torch.empty(n, device=DEV, dtype=torch.bfloat16)
        t = timeit(lambda: N.call("dla_adamw", N.ptr(master), N.ptr(m_), N.ptr(v_), N.ptr(g), N.ptr(p), n, 1e-5, 0.9, 0.95, 1e-8, 0.01,
                                  0.1, 0.05, None, 1.0, 0, N.stream_ptr()), flush=False)
        rec = {"kernel": "adamw", "n": n, "ms": t, "gbs": n * 28 / t / 1e6, "frac_hbm": n * 28 / t / 1e6 / PK["hbm_gbs"]}
        out.append(rec); print(json.dumps(rec), flush=True)
    if "logprob" in which:
        M, V, K = 8192, 128256, 4096
        h, w = rnd(M, K), rnd(V, K) * 0.02
        lab = torch.randint(0, V, (M,), device=DEV)
        t = timeit(lambda: F.linear_logprob(h, w, lab), iters=5)
        rec = {"kernel": "lmhead_logprob_fwd", "M": M, "V": V, "K": K, "ms": t, "tflops": 2.0 * M * V * K / t / 1e9}

        def ref():
            logits = torch.matmul(h, w.t())
            lp = torch.log_softmax(logits.float(), -1)
            return lp.gather(-1, lab[:, None])
        t2 = timeit(ref, iters=3)
        rec["reference_ms"] = t2
        out.append(rec); print(json.dumps(rec), flush=True)
    Path(ROOT / "gpurun_out").mkdir(exist_ok=True)
    with open(ROOT / "gpurun_out" / "kernels.jsonl", "a") as fh:
        for r in out:
            fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
