#!/usr/bin/env python
"""Roofline numbers for the three fused compute+collective paths, N ranks on one box (torchrun).
Device-timed (CUDA events), max over ranks, L2 irrelevant (buffers are hundreds of MB).

  (a) per-unit gradient reduce-scatter + scale + norm partial (fused peer reads) vs NCCL reduce_scatter_tensor
      + cast + sumsq, and the full optimizer step (fused vs NCCL engine) on Llama-3-8B-sized units
  (b) all-gather fused GEMM vs NCCL all_gather_into_tensor + tcgen05 GEMM, on the QKV and gate-up shapes
  (c) fused gather + DPO loss vs all_gather x4 + loss kernel

Writes gpurun_out/fused_paths_N<world>.json (rank 0)."""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.ops import native as N  # noqa: E402
from dla_b200.parallel.fused import FusedComm, SymmetricArena  # noqa: E402
from dla_b200.parallel.pg import init_distributed  # noqa: E402

NVLINK_GBS = 770.0          # measured peer-copy bandwidth per direction (B200_PROFILING.md)
ctx = init_distributed()
W, r, dev = ctx.world_size, ctx.rank, ctx.device
fc = FusedComm.get(ctx)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ctx.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return ctx.max_over_ranks(e0.elapsed_time(e1) / iters)


out = {"world": W, "nvlink_gbs_per_dir": NVLINK_GBS}

# ---------------------------------------------------------------- (a)
P = 218_112_000                      # one Llama-3-8B decoder layer
P = -(-P // (8 * W)) * (8 * W)
S = P // W
grads = SymmetricArena(ctx, P, torch.bfloat16)
grads.tensor.normal_()
params = SymmetricArena(ctx, P, torch.bfloat16)
gred = torch.empty(S, device=dev)
sumsq = torch.zeros(1, device=dev)
master, m_, v_ = (torch.zeros(S, device=dev) for _ in range(3))


def fused_a():
    fc.barrier(0)
    fc.reduce_shard(grads, r * S, S, 1.0 / W, gred, sumsq)


nccl_out = torch.empty(S, device=dev, dtype=torch.bfloat16)
scratch = torch.empty(N.sumsq_blocks(), device=dev)


def nccl_a():
    dist.reduce_scatter_tensor(nccl_out, grads.tensor, group=ctx.group)
    N.call("dla_sumsq", N.ptr(nccl_out), S, 0, N.ptr(scratch), N.ptr(sumsq), 1, N.stream_ptr())


def fused_a_full():
    fc.barrier(0)
    fc.reduce_adamw_push(grads, params, r * S, S, 1.0 / W, master, m_, v_, 1e-6, (0.9, 0.999), 1e-8, 0.01, 0.1, 0.001, None)
    fc.barrier(0)


shard_bf16 = torch.empty(S, device=dev, dtype=torch.bfloat16)


def nccl_a_full():
    dist.reduce_scatter_tensor(nccl_out, grads.tensor, group=ctx.group)
    N.call("dla_adamw", N.ptr(master), N.ptr(m_), N.ptr(v_), N.ptr(nccl_out), N.ptr(shard_bf16), S, 1e-6, 0.9, 0.999, 1e-8, 0.01,
           0.1, 0.001, None, 1.0 / W, 0, N.stream_ptr())
    dist.all_gather_into_tensor(params.tensor, shard_bf16, group=ctx.group)


link_bytes = (W - 1) * S * 2                      # bytes each rank must pull over NVLink
t_fa, t_na = timed(fused_a), timed(nccl_a)
t_faf, t_naf = timed(fused_a_full), timed(nccl_a_full)
roof_a = link_bytes / (NVLINK_GBS * 1e9) * 1e3
roof_af = max(2 * link_bytes / 2, link_bytes) / (NVLINK_GBS * 1e9) * 1e3      # pull grads (in) and push params (out) use opposite directions
out["a_reduce_scatter_norm"] = {"unit_elems": P, "fused_ms": t_fa, "nccl_ms": t_na, "roofline_ms": roof_a,
                                "fused_frac_of_roofline": roof_a / t_fa, "nccl_frac_of_roofline": roof_a / t_na}
out["a_reduce_adamw_allgather"] = {"fused_ms": t_faf, "nccl_ms": t_naf, "roofline_ms": roof_af,
                                   "fused_frac_of_roofline": roof_af / t_faf, "nccl_frac_of_roofline": roof_af / t_naf}

# ---- (a) v2: flag-gated per-unit reduce-scatter (no barrier), in-switch sum vs P2P pull; AdamW with multicast store
row0 = fc.claim_unit_rows(2)
ep_v2 = [0, 0]
os.environ["DLA_NO_MULTICAST"] = "1"
grads_p2p = SymmetricArena(ctx, P, torch.bfloat16)
os.environ.pop("DLA_NO_MULTICAST")
grads_p2p.tensor.copy_(grads.tensor)
coef, norm = torch.ones(1, device=dev), torch.zeros(1, device=dev)


def v2_rs_switch():
    ep_v2[0] += 1
    fc.reduce_unit(grads, row0, ep_v2[0], r * S, S, 1.0, gred, sumsq)


def v2_rs_p2p():
    ep_v2[1] += 1
    fc.reduce_unit(grads_p2p, row0 + 16, ep_v2[1], r * S, S, 1.0, gred, sumsq)


def v2_full():
    ep_v2[0] += 1
    fc.reduce_unit(grads, row0, ep_v2[0], r * S, S, 1.0, gred, sumsq)
    fc.norm_clip(sumsq, 1.0 / W, 1.0, coef, norm)
    if params.mc_ptr:
        N.call("dla_adamw_mc", N.ptr(master), N.ptr(m_), N.ptr(v_), N.ptr(gred), params.mc_ptr + r * S * 2, S, 1e-6, 0.9, 0.999, 1e-8,
               0.01, 0.1, 0.001, N.ptr(coef), 1.0 / W, 1, N.stream_ptr())
    else:
        N.call("dla_adamw", N.ptr(master), N.ptr(m_), N.ptr(v_), N.ptr(gred), N.ptr(shard_bf16), S, 1e-6, 0.9, 0.999, 1e-8, 0.01,
               0.1, 0.001, N.ptr(coef), 1.0 / W, 1, N.stream_ptr())
        fc.push_shard(shard_bf16, params, r * S)
    fc.barrier(0)


t_v2s = timed(v2_rs_switch) if grads.mc_ptr else None
t_v2p = timed(v2_rs_p2p)
t_v2f = timed(v2_full)
# with the sum done in the switch each rank still SENDS its (W-1) foreign shards (same egress bytes as the pull model) but
# receives one reduced shard instead of W-1; the all-gather by multicast store is the mirror image (sends S, receives
# (W-1) S).  A reduce-scatter and an all-gather in flight together therefore load each direction with W*S elements.
out["a_v2_unit_reduce"] = {"multicast": bool(grads.mc_ptr), "switch_ms": t_v2s, "p2p_ms": t_v2p, "nccl_ms": t_na,
                           "egress_bytes": link_bytes, "ingress_bytes_switch": S * 2, "ingress_bytes_p2p": link_bytes,
                           "roofline_ms": roof_a, "switch_frac_of_roofline": (roof_a / t_v2s) if t_v2s else None,
                           "p2p_frac_of_roofline": roof_a / t_v2p}
out["a_v2_reduce_norm_adamw_push"] = {"ms": t_v2f, "legacy_fused_ms": t_faf, "nccl_ms": t_naf, "roofline_ms": roof_af,
                                      "frac_of_roofline": roof_af / t_v2f,
                                      "bytes_per_direction": {"out": link_bytes + S * 2, "in": S * 2 + link_bytes}}

# ---------------------------------------------------------------- (b)
bf16_tflops = 1632.6
try:
    bf16_tflops = json.load(open(ROOT / "MEASURED_PEAKS.json"))["bf16_tflops"]
except Exception:
    pass
res_b = {}
for tag, Nn, K in (("qkv", 6144, 4096), ("gate_up", 28672, 4096)):
    M = 8192
    unit = -(-Nn * K // (8 * W)) * (8 * W)
    Sb = unit // W
    shards = SymmetricArena(ctx, Sb, torch.bfloat16)
    shards.tensor.normal_(std=0.02)
    ptrs = torch.tensor(shards.ptrs, dtype=torch.int64, device=dev)
    full = torch.empty(unit, device=dev, dtype=torch.bfloat16)
    x = torch.randn(M, K, device=dev).to(torch.bfloat16)
    flags = torch.zeros(256, dtype=torch.int32, device=dev)
    ep = [0]

    ncopy = [12]

    def fused_b():
        ep[0] += 1
        fc.ag_gemm(x, ptrs, full, Sb, 0, Nn, flags, ep[0], n_copy=ncopy[0])

    y = torch.empty(M, Nn, device=dev, dtype=torch.bfloat16)

    def nccl_b():
        dist.all_gather_into_tensor(full, shards.tensor, group=ctx.group)
        F.gemm(x, full[:Nn * K].view(Nn, K), out=y)

    def gemm_only():
        F.gemm(x, full[:Nn * K].view(Nn, K), out=y)

    fc.barrier(0)
    sweep = {}
    for nc in (2, 4, 8, 12, 16):
        ncopy[0] = nc
        sweep[nc] = timed(fused_b)
    best = min(sweep, key=sweep.get)
    ncopy[0] = best
    t_f, t_n, t_g = sweep[best], timed(nccl_b), timed(gemm_only)
    flops = 2.0 * M * Nn * K
    roof = max(flops / (bf16_tflops * 1e12), (W - 1) * Sb * 2 / (NVLINK_GBS * 1e9)) * 1e3
    res_b[tag] = {"M": M, "N": Nn, "K": K, "copy_cta_sweep_ms": sweep, "copy_ctas": best, "fused_ms": t_f, "nccl_ag_plus_gemm_ms": t_n, "gemm_only_ms": t_g, "roofline_ms": roof,
                  "fused_frac_of_roofline": roof / t_f, "nccl_frac_of_roofline": roof / t_n}
out["b_all_gather_gemm"] = res_b

# ---------------------------------------------------------------- (c)
B = 4
pp, pn, rp, rn = (torch.randn(B, device=dev) for _ in range(4))


def fused_c():
    fc.dpo_loss(pp, pn, rp, rn, 0.1, 0.0)


def nccl_c():
    g = [ctx.gather(t) for t in (pp, pn, rp, rn)]
    F.dpo_loss(g[0], g[1], g[2], g[3], 0.1)


out["c_gather_dpo_loss"] = {"pairs_per_rank": B, "fused_us": timed(fused_c, 50) * 1e3, "nccl_us": timed(nccl_c, 50) * 1e3,
                            "note": "latency-bound: O(B) floats; roofline = one NVLink round trip (~2-4 us)"}
fc.check()
if ctx.is_main:
    Path(ROOT / "gpurun_out").mkdir(exist_ok=True)
    json.dump(out, open(ROOT / "gpurun_out" / f"fused_paths_N{W}.json", "w"), indent=1)
    print(json.dumps(out, indent=1))
ctx.barrier()
