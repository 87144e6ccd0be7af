"""Multi-GPU checks (>= 2 B200s on one box): the fused peer-memory kernels against their NCCL
equivalents, and end-to-end loss parity of the fused engine vs the NCCL engine."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _init(rank, world, port):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    sys.path.insert(0, str(ROOT))
    from dla_b200.parallel.pg import init_distributed

    return init_distributed()


def _kernels_worker(rank, world, port):
    import torch.distributed as dist

    ctx = _init(rank, world, port)
    from dla_b200.parallel.fused import FusedComm, SymmetricArena

    dev = ctx.device
    fc = FusedComm.get(ctx)
    n_total = 8 * world * 4096 + 8 * world * 3
    S = n_total // world
    grads = SymmetricArena(ctx, n_total, torch.bfloat16)
    params = SymmetricArena(ctx, n_total, torch.bfloat16)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    grads.tensor.copy_(torch.randn(n_total, device=dev, generator=g).to(torch.bfloat16))
    # ---- (a) reduce-scatter: fused peer reads vs NCCL
    fc.barrier(0)
    out = torch.empty(S, device=dev)
    sumsq = torch.zeros(1, device=dev)
    fc.reduce_shard(grads, rank * S, S, 0.5, out, sumsq)
    ref = torch.empty(S, device=dev)
    dist.reduce_scatter_tensor(ref, grads.tensor.float(), group=ctx.group)
    ref *= 0.5
    torch.cuda.synchronize()
    assert torch.allclose(out, ref, atol=1e-5, rtol=1e-5), (out - ref).abs().max()
    assert torch.allclose(sumsq, ref.pow(2).sum().reshape(1), rtol=1e-3)
    # ---- all-gather by peer stores vs NCCL
    shard = torch.randn(S, device=dev, generator=g).to(torch.bfloat16)
    fc.push_shard(shard, params, rank * S)
    fc.barrier(0)
    want = torch.empty(n_total, device=dev, dtype=torch.bfloat16)
    dist.all_gather_into_tensor(want, shard, group=ctx.group)
    torch.cuda.synchronize()
    assert torch.equal(params.tensor, want)
    # ---- fused reduce + AdamW + push (no clipping) vs reduce_scatter + torch AdamW + all_gather
    master = torch.randn(S, device=dev, generator=g)
    m, v = torch.zeros(S, device=dev), torch.zeros(S, device=dev)
    p_ref = master.clone().requires_grad_()
    opt = torch.optim.AdamW([p_ref], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01)
    p_ref.grad = ref.clone() / 0.5 * 0.25
    opt.step()
    fc.barrier(0)
    fc.reduce_adamw_push(grads, params, rank * S, S, 0.25, master, m, v, 1e-2, (0.9, 0.95), 1e-8, 0.01, 1 - 0.9, 1 - 0.95, None)
    fc.barrier(0)
    torch.cuda.synchronize()
    assert torch.allclose(master, p_ref.detach(), atol=1e-5, rtol=1e-5)
    gathered = torch.empty(n_total, device=dev, dtype=torch.bfloat16)
    dist.all_gather_into_tensor(gathered, master.to(torch.bfloat16), group=ctx.group)
    assert torch.equal(params.tensor, gathered)
    # ---- (c) fused gather + DPO loss vs gathered reference
    from dla_b200.ops import reference as R

    B = 5
    pp, pn, rp, rn = (torch.randn(B, device=dev, generator=g) for _ in range(4))
    stats, dpp, dpn = fc.dpo_loss(pp, pn, rp, rn, 0.3, 0.1)
    allv = [ctx.gather(t) for t in (pp, pn, rp, rn)]
    want_loss = R.dpo_loss(*allv, 0.3, 0.1)
    a = pp.clone().requires_grad_()
    b = pn.clone().requires_grad_()
    R.dpo_loss(a, b, rp, rn, 0.3, 0.1).backward()
    torch.cuda.synchronize()
    assert torch.allclose(stats[0], want_loss, atol=1e-5)
    assert torch.allclose(stats[1], (allv[0] > allv[1]).float().mean(), atol=1e-6)
    assert torch.allclose(dpp, a.grad, atol=1e-6) and torch.allclose(dpn, b.grad, atol=1e-6)
    # ---- tiny all-gather over peer memory (ctx.gather routes through it once a FusedComm exists) vs NCCL,
    # several calls back to back with changing sizes (the buffer halves alternate with the epoch parity)
    assert ctx.fused_comm is fc
    for it, n in enumerate([1, 7, 2048, 3, 64, 1]):
        x = torch.randn(n, device=dev, generator=g) + 10 * rank + it
        got = ctx.gather(x)
        want_g = torch.empty(world * n, device=dev)
        dist.all_gather_into_tensor(want_g, x, group=ctx.group)
        assert torch.equal(got, want_g), (it, n)
    xb = torch.randn(4, 3, device=dev, generator=g).to(torch.bfloat16)
    got = ctx.gather(xb)
    want_b = torch.empty(world * 4, 3, device=dev, dtype=torch.bfloat16)
    dist.all_gather_into_tensor(want_b, xb, group=ctx.group)
    assert got.dtype == torch.bfloat16 and torch.equal(got, want_b)
    fc.check()
    ctx.barrier()


def _e2e_worker(rank, world, port, out):
    ctx = _init(rank, world, port)
    from dla_b200.algo import dpo_step
    from dla_b200.config import HardwarePlan
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig

    res = {}
    for comm in ("nccl", "fused"):
        plan = HardwarePlan(sharding="zero3", comm=comm, resident_params=True)
        pol = load_causal_lm("tiny-llama", ctx, plan, seed=11, name=f"policy_{comm}")
        ref = load_causal_lm("tiny-llama", ctx, plan, trainable=False, gradient_checkpointing=False, seed=11, name=f"ref_{comm}")
        assert pol.engine.fused == (comm == "fused")
        losses = []
        for step in range(4):
            full = synthetic_preference_batch(2 * world, 128, pol.config.vocab_size, seed=50 + step, ragged=True)
            batch = {s: {k: v[rank * 2:(rank + 1) * 2].cuda() for k, v in full[s].items()} for s in full}
            loss, info = dpo_step(pol.model, ref.model, batch, beta=0.5)
            loss.backward()
            pol.engine.finish_backward()
            pol.engine.optimizer_step(AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=1.0), grad_divisor=float(world))
            # NCCL path reports the rank-local loss, the fused path the global one: compare global means
            losses.append(float(ctx.all_reduce(loss.detach().clone(), "mean")) if comm == "nccl" else float(loss))
        flat = torch.cat([pol.engine.gathered_unit(u.name)[p.name].reshape(-1).float() for u in pol.engine.units for p in u.params])
        res[comm] = (losses, flat.cpu(), float(pol.engine.last_grad_norm))
    if rank == 0:
        torch.save(res, out)
    ctx.barrier()


def test_fused_peer_kernels_match_nccl():
    mp.spawn(_kernels_worker, args=(2, _free_port()), nprocs=2, join=True)


def test_fused_engine_matches_nccl_engine(tmp_path):
    out = tmp_path / "e2e.pt"
    mp.spawn(_e2e_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    res = torch.load(out)
    ln, pn, gn = res["nccl"]
    lf, pf, gf = res["fused"]
    assert ln == pytest.approx(lf, abs=2e-3), (ln, lf)
    assert gn == pytest.approx(gf, rel=2e-2)
    rel = float((pn - pf).norm() / pn.norm())
    assert rel < 2e-3, rel


def _ag_worker(rank, world, port, out):
    import torch.distributed as dist

    ctx = _init(rank, world, port)
    from dla_b200.parallel.fused import FusedComm, SymmetricArena

    dev = ctx.device
    fc = FusedComm.get(ctx)
    # a "unit": [pad 512 elems | W (N x K) | tail 1024 elems], sharded contiguously over the ranks
    M, N_, K = 640, 1536, 512
    b_off = 512
    P = b_off + N_ * K + 1024
    P = -(-P // (8 * world)) * (8 * world)
    S = P // world
    g = torch.Generator(device=dev).manual_seed(7)                      # same "full unit" on every rank
    unit = (torch.randn(P, device=dev, generator=g) * 0.05).to(torch.bfloat16)
    shards = SymmetricArena(ctx, S, torch.bfloat16)
    shards.tensor.copy_(unit[rank * S:(rank + 1) * S])
    fc.barrier(0)
    ptrs = torch.tensor(shards.ptrs, dtype=torch.int64, device=dev)
    x = torch.randn(M, K, device=dev, generator=torch.Generator(device=dev).manual_seed(100 + rank)).to(torch.bfloat16)
    flags = torch.zeros(64, dtype=torch.int32, device=dev)
    for epoch in (1, 2):                                               # second call reuses the flags with a new epoch
        full = torch.full((P,), float("nan"), device=dev, dtype=torch.bfloat16)
        y = fc.ag_gemm(x, ptrs, full, S, b_off, N_, flags, epoch, n_copy=4)
        torch.cuda.synchronize()
        w = unit[b_off:b_off + N_ * K].view(N_, K)
        assert torch.equal(full[b_off:b_off + N_ * K].view(N_, K), w), "gathered weight differs"
        ref = x.float() @ w.float().t()
        assert torch.allclose(y.float(), ref, atol=0.05, rtol=2e-2), (y.float() - ref).abs().max()
    # plain peer gather of the tail
    fc.gather_range(ptrs, full, S, b_off + N_ * K, 1024)
    torch.cuda.synchronize()
    assert torch.equal(full[b_off + N_ * K:b_off + N_ * K + 1024], unit[b_off + N_ * K:b_off + N_ * K + 1024])
    fc.check()
    ctx.barrier()


def _e2e_stream_worker(rank, world, port, out):
    ctx = _init(rank, world, port)
    from dla_b200.algo import dpo_step
    from dla_b200.config import HardwarePlan
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.ops import functional as F
    from dla_b200.optim import AdamWConfig

    # this test compares two ways of moving the same bytes: keep the math identical (the all-gather GEMM has no RoPE
    # epilogue, so its arm rounds the projection before rotating it; the NCCL arm would rotate the fp32 accumulators)
    F.FUSE_ROPE = False
    res = {}
    for comm in ("nccl", "fused"):
        plan = HardwarePlan(sharding="zero3", comm=comm, resident_params=False)
        pol = load_causal_lm("tiny-llama", ctx, plan, seed=13, name=f"policy_{comm}")
        ref = load_causal_lm("tiny-llama", ctx, plan, trainable=False, gradient_checkpointing=False, seed=13, name=f"ref_{comm}")
        assert pol.engine.fused_stream == (comm == "fused") and not pol.engine.resident
        losses = []
        for step in range(3):
            full = synthetic_preference_batch(2 * world, 128, pol.config.vocab_size, seed=70 + step, ragged=True)
            batch = {s: {k: v[rank * 2:(rank + 1) * 2].cuda() for k, v in full[s].items()} for s in full}
            loss, info = dpo_step(pol.model, ref.model, batch, beta=0.5)
            loss.backward()
            pol.engine.finish_backward()
            pol.engine.optimizer_step(AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=1.0), grad_divisor=float(world))
            losses.append(float(ctx.all_reduce(loss.detach().clone(), "mean")))
        flat = torch.cat([pol.engine.gathered_unit(u.name)[p.name].reshape(-1).float() for u in pol.engine.units for p in u.params])
        res[comm] = (losses, flat.cpu())
    if rank == 0:
        torch.save(res, out)
    ctx.barrier()


def test_all_gather_fused_gemm_matches_gather_then_matmul(tmp_path):
    mp.spawn(_ag_worker, args=(2, _free_port(), str(tmp_path / "x")), nprocs=2, join=True)


def test_streamed_engine_with_fused_all_gather_matches_nccl(tmp_path):
    out = tmp_path / "e2e_stream.pt"
    mp.spawn(_e2e_stream_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    res = torch.load(out)
    (ln, pn), (lf, pf) = res["nccl"], res["fused"]
    assert ln == pytest.approx(lf, abs=2e-3), (ln, lf)
    assert float((pn - pf).norm() / pn.norm()) < 2e-3


def _e2e_ddp_worker(rank, world, port, out):
    ctx = _init(rank, world, port)
    from dla_b200.algo import dpo_step
    from dla_b200.config import HardwarePlan
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig

    res = {}
    for comm in ("nccl", "fused"):
        plan = HardwarePlan(sharding="ddp", comm=comm)
        pol = load_causal_lm("tiny-llama", ctx, plan, seed=17, name=f"policy_{comm}")
        ref = load_causal_lm("tiny-llama", ctx, plan, trainable=False, gradient_checkpointing=False, seed=17, name=f"ref_{comm}")
        assert pol.engine.fused_ddp == (comm == "fused") and pol.engine.W == 1
        for step in range(3):
            full = synthetic_preference_batch(2 * world, 128, pol.config.vocab_size, seed=90 + step, ragged=True)
            batch = {s: {k: v[rank * 2:(rank + 1) * 2].cuda() for k, v in full[s].items()} for s in full}
            loss, info = dpo_step(pol.model, ref.model, batch, beta=0.5)
            loss.backward()
            pol.engine.finish_backward()
            pol.engine.optimizer_step(AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=1.0), grad_divisor=float(world))
        flat = torch.cat([pol.engine.gathered_unit(u.name)[p.name].reshape(-1).float() for u in pol.engine.units for p in u.params])
        # replicas must stay identical across ranks
        mx = flat.clone()
        ctx.all_reduce(mx, "max")
        assert torch.equal(mx, flat), "DDP replicas diverged"
        res[comm] = flat.cpu()
    if rank == 0:
        torch.save(res, out)
    ctx.barrier()


def test_replicated_engine_with_peer_allreduce_matches_nccl(tmp_path):
    out = tmp_path / "e2e_ddp.pt"
    mp.spawn(_e2e_ddp_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    res = torch.load(out)
    assert float((res["nccl"] - res["fused"]).norm() / res["nccl"].norm()) < 2e-3


# ------------------------------------------------------------------------------------------------------------------
# round 2: in-backward per-unit reduce-scatter, NVSwitch multimem paths, norm exchange, RLHF gathers
# ------------------------------------------------------------------------------------------------------------------

def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [w for w in (2, 4, 8) if w <= n] or [2]


def _v2_kernels_worker(rank, world, port):
    import torch.distributed as dist

    ctx = _init(rank, world, port)
    from dla_b200.ops import native as N
    from dla_b200.ops import reference as R
    from dla_b200.parallel.fused import FusedComm, SymmetricArena, fused_reinforce_loss

    dev = ctx.device
    fc = FusedComm.get(ctx)
    n_total = 8 * world * 4096 + 8 * world * 5
    S = n_total // world
    gen = torch.Generator(device=dev).manual_seed(200 + rank)
    arenas = {"switch": SymmetricArena(ctx, n_total, torch.bfloat16)}
    os.environ["DLA_NO_MULTICAST"] = "1"
    arenas["p2p"] = SymmetricArena(ctx, n_total, torch.bfloat16)
    os.environ.pop("DLA_NO_MULTICAST")
    assert arenas["p2p"].mc_ptr == 0
    has_mc = arenas["switch"].mc_ptr != 0
    if rank == 0:
        print(f"[v2] world={world} multicast={'yes' if has_mc else 'NO (P2P fallback only)'}", flush=True)
    row0 = fc.claim_unit_rows(4)
    data = torch.randn(n_total, device=dev, generator=gen).to(torch.bfloat16)
    ref = torch.empty(S, device=dev)
    dist.reduce_scatter_tensor(ref, data.float(), group=ctx.group)
    # ---- per-unit reduce-scatter gated by the unit's own flags (no barrier before it), twice (epochs 1, 2)
    for name, arena in arenas.items():
        for epoch, unit in ((1, 0), (2, 0), (1, 1)):
            arena.tensor.copy_(data)
            out = torch.full((S,), float("nan"), device=dev)
            sumsq = torch.zeros(1, device=dev)
            row = row0 + 16 * (unit + (2 if name == "p2p" else 0))
            fc.reduce_unit(arena, row, epoch, rank * S, S, 0.5, out, sumsq)
            torch.cuda.synchronize()
            tol = dict(atol=2e-2, rtol=1e-2) if arena.mc_ptr else dict(atol=1e-5, rtol=1e-5)   # the switch returns bf16
            assert torch.allclose(out, ref * 0.5, **tol), (name, (out - ref * 0.5).abs().max())
            assert torch.allclose(sumsq, out.pow(2).sum().reshape(1), rtol=1e-3), name
            fc.barrier(0)          # nobody overwrites the arena while a peer still reads it
    # ---- norm exchange: every rank gets the same total, the local partial is reset
    part = torch.full((1,), float(rank + 1), device=dev)
    coef, norm = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    for it in range(3):
        part.fill_(float(rank + 1 + it))
        fc.norm_clip(part, 0.5, 1.0, coef, norm)
        torch.cuda.synchronize()
        tot = sum(r + 1 + it for r in range(world))
        assert float(norm) == pytest.approx((tot ** 0.5) * 0.5, rel=1e-6) and float(part) == 0.0
        assert float(coef) == pytest.approx(min(1.0, 1.0 / ((tot ** 0.5) * 0.5 + 1e-6)), rel=1e-6)
    # ---- AdamW whose bf16 result is stored into every rank's copy (multicast) / pushed (P2P), all-reduce variants
    for name, arena in arenas.items():
        master = torch.randn(S, device=dev, generator=gen)
        want_master = master.clone().requires_grad_()
        m, v = torch.zeros(S, device=dev), torch.zeros(S, device=dev)
        grad = torch.randn(S, device=dev, generator=gen)
        opt = torch.optim.AdamW([want_master], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01)
        want_master.grad = grad.clone()
        opt.step()
        one = torch.ones(1, device=dev)
        fc.barrier(0)
        if arena.mc_ptr:
            N.call("dla_adamw_mc", N.ptr(master), N.ptr(m), N.ptr(v), N.ptr(grad), arena.mc_ptr + rank * S * 2, S, 1e-2, 0.9, 0.95,
                   1e-8, 0.01, 1 - 0.9, 1 - 0.95, N.ptr(one), 1.0, 1, N.stream_ptr(dev))
        else:
            shard = torch.empty(S, device=dev, dtype=torch.bfloat16)
            N.call("dla_adamw", N.ptr(master), N.ptr(m), N.ptr(v), N.ptr(grad), N.ptr(shard), S, 1e-2, 0.9, 0.95, 1e-8, 0.01,
                   1 - 0.9, 1 - 0.95, N.ptr(one), 1.0, 1, N.stream_ptr(dev))
            fc.push_shard(shard, arena, rank * S)
        fc.barrier(0)
        torch.cuda.synchronize()
        assert torch.allclose(master, want_master.detach(), atol=1e-5, rtol=1e-5)
        gathered = torch.empty(n_total, device=dev, dtype=torch.bfloat16)
        dist.all_gather_into_tensor(gathered, master.to(torch.bfloat16), group=ctx.group)
        assert torch.equal(arena.tensor, gathered), name
        # a clip coefficient of 0 (non-finite norm) must leave everything untouched
        zero = torch.zeros(1, device=dev)
        snap = (master.clone(), m.clone(), v.clone())
        N.call("dla_adamw", N.ptr(master), N.ptr(m), N.ptr(v), N.ptr(torch.full_like(grad, float("inf"))), None, S, 1e-2, 0.9, 0.95,
               1e-8, 0.01, 1 - 0.9, 1 - 0.95, N.ptr(zero), 1.0, 1, N.stream_ptr(dev))
        torch.cuda.synchronize()
        assert torch.equal(master, snap[0]) and torch.equal(m, snap[1]) and torch.equal(v, snap[2])
        # DDP all-reduce in place
        arena.tensor.copy_(data)
        want_ar = data.float().clone()
        dist.all_reduce(want_ar, group=ctx.group)
        fc.all_reduce_(arena, 0, n_total, 0.25)
        torch.cuda.synchronize()
        assert torch.allclose(arena.tensor.float(), want_ar * 0.25, atol=3e-2, rtol=2e-2), name
        mx = arena.tensor.float().clone()
        dist.all_reduce(mx, dist.ReduceOp.MAX, group=ctx.group)
        assert torch.equal(mx, arena.tensor.float()), "replicas differ after the all-reduce"
    # ---- (c) REINFORCE over the global batch, several calls back to back (the buffer halves alternate)
    B = 6
    for it in range(3):
        pol = torch.randn(B, device=dev, generator=gen)
        refl = torch.randn(B, device=dev, generator=gen)
        score = torch.randn(B, device=dev, generator=gen)
        pg = pol.clone().requires_grad_()
        loss, st = fused_reinforce_loss(fc, pg, refl, score, 0.1)
        loss.backward()
        ap, ar, asc = (ctx.gather(t) for t in (pol, refl, score))
        rew = asc - 0.1 * (ap - ar)
        base = rew.mean()
        want = -((rew - base) * ap).mean()
        torch.cuda.synchronize()
        assert torch.allclose(loss.detach(), want, atol=1e-5), (it, loss, want)
        assert torch.allclose(st[0], (ap - ar).mean(), atol=1e-5) and torch.allclose(st[1], rew.mean(), atol=1e-5)
        assert torch.allclose(st[2], asc.mean(), atol=1e-5) and torch.allclose(st[3], base, atol=1e-5)
        mine = (score - 0.1 * (pol - refl)) - base
        assert torch.allclose(pg.grad, -mine / B, atol=1e-6)
    # ---- (c) PPO: global advantage whitening + summed statistics
    Bt, T = 3, 37
    adv = torch.randn(Bt, T, device=dev, generator=gen) * (rank + 1)
    mask = (torch.rand(Bt, T, device=dev, generator=gen) > 0.3).float()
    stats = torch.tensor([1.0 + rank, 2.0, 0.5 * rank, float(Bt)], device=dev)
    all_adv, all_mask = ctx.gather(adv.reshape(1, -1)).reshape(-1), ctx.gather(mask.reshape(1, -1)).reshape(-1)
    mean = (all_adv * all_mask).sum() / all_mask.sum()
    var = (((all_adv - mean) ** 2) * all_mask).sum() / all_mask.sum()
    want_adv = (adv - mean) / var.sqrt() * mask
    got = adv.clone().contiguous()
    tot = fc.whiten_(got, mask.contiguous(), stats)
    torch.cuda.synchronize()
    assert torch.allclose(got, want_adv, atol=1e-4, rtol=1e-4), (got - want_adv).abs().max()
    want_tot = stats.clone()
    dist.all_reduce(want_tot, group=ctx.group)
    assert torch.allclose(tot, want_tot, atol=1e-5)
    # ---- DPO gather back to back with different sizes (double buffered now)
    for it, Bd in enumerate([3, 7, 2]):
        pp, pn, rp, rn = (torch.randn(Bd, device=dev, generator=gen) for _ in range(4))
        st2, _, _ = fc.dpo_loss(pp, pn, rp, rn, 0.3, 0.0)
        allv = [ctx.gather(t) for t in (pp, pn, rp, rn)]
        torch.cuda.synchronize()
        assert torch.allclose(st2[0], R.dpo_loss(*allv, 0.3, 0.0), atol=1e-5), it
    fc.check(sync=True)
    ctx.barrier()


@pytest.mark.parametrize("world", _worlds())
def test_round2_peer_kernels_match_nccl(world):
    mp.spawn(_v2_kernels_worker, args=(world, _free_port()), nprocs=world, join=True)


def _e2e_v2_worker(rank, world, port, out):
    ctx = _init(rank, world, port)
    from dla_b200.algo import dpo_step
    from dla_b200.config import HardwarePlan
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig

    res = {}
    for tag, comm, env, keep in (("nccl", "nccl", {}, 0), ("v2", "fused", {}, 0), ("v2_keep", "fused", {}, 2),
                                 ("v2_p2p", "fused", {"DLA_NO_MULTICAST": "1"}, 1), ("legacy", "fused", {"DLA_RS_IN_BACKWARD": "0"}, 0)):
        os.environ.update(env)
        plan = HardwarePlan(sharding="zero3", comm=comm, resident_params=True)
        pol = load_causal_lm("tiny-llama", ctx, plan, seed=11, name=f"policy_{tag}")
        ref = load_causal_lm("tiny-llama", ctx, plan, trainable=False, gradient_checkpointing=False, seed=11, name=f"ref_{tag}")
        for k in env:
            os.environ.pop(k)
        assert pol.engine.fused == (comm == "fused")
        if comm == "fused":
            assert pol.engine.rs_in_backward == (tag != "legacy")
        pol.engine.set_no_recompute_layers(keep)       # kept layers signal through the unit-boundary node
        losses = []
        for step in range(4):
            full = synthetic_preference_batch(2 * world, 128, pol.config.vocab_size, seed=50 + step, ragged=True)
            batch = {s: {k: v[rank * 2:(rank + 1) * 2].cuda() for k, v in full[s].items()} for s in full}
            loss, info = dpo_step(pol.model, ref.model, batch, beta=0.5)
            loss.backward()
            pol.engine.finish_backward()
            pol.engine.optimizer_step(AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=1.0), grad_divisor=float(world))
            losses.append(float(ctx.all_reduce(loss.detach().clone(), "mean")) if comm == "nccl" else float(loss))
        flat = torch.cat([pol.engine.gathered_unit(u.name)[p.name].reshape(-1).float() for u in pol.engine.units for p in u.params])
        mx = flat.clone()
        ctx.all_reduce(mx, "max")
        assert torch.equal(mx, flat), f"{tag}: resident parameter copies differ across ranks"
        res[tag] = (losses, flat.cpu(), float(pol.engine.last_grad_norm))
    if rank == 0:
        torch.save(res, out)
    ctx.barrier()


@pytest.mark.parametrize("world", _worlds())
def test_in_backward_reduce_scatter_engine_matches_nccl(tmp_path, world):
    out = tmp_path / "e2e_v2.pt"
    mp.spawn(_e2e_v2_worker, args=(world, _free_port(), str(out)), nprocs=world, join=True)
    res = torch.load(out)
    ln, pn, gn = res["nccl"]
    for tag in ("v2", "v2_keep", "v2_p2p", "legacy"):
        lf, pf, gf = res[tag]
        assert ln == pytest.approx(lf, abs=3e-3), (tag, ln, lf)
        assert gn == pytest.approx(gf, rel=3e-2), tag
        assert float((pn - pf).norm() / pn.norm()) < 3e-3, tag


def _watchdog_worker(rank, world, port):
    ctx = _init(rank, world, port)
    from dla_b200.parallel.fused import FusedComm

    fc = FusedComm.get(ctx)
    fc.barrier(0)
    torch.cuda.synchronize()
    fc.check()                                   # healthy so far
    if rank == 0:
        # rank 1 never enters this barrier (a wedged peer): the wait gives up after 300 ms of wall time, the kernel
        # mirrors the fact into pinned host memory, and check() turns it into an exception without a device query
        import time

        t0 = time.perf_counter()
        fc.barrier(7, timeout_ms=300)
        torch.cuda.synchronize()
        waited = time.perf_counter() - t0
        assert 0.25 < waited < 5.0, waited
        with pytest.raises(RuntimeError, match="timed out"):
            fc.check()
        with pytest.raises(RuntimeError, match="timed out"):
            fc.check(sync=True)
    ctx.barrier()                                # NCCL: both ranks leave together


def test_watchdog_turns_a_missing_peer_into_an_exception():
    """a peer that stops taking part must not hang the others forever or let them compute on half-exchanged
    data — the flag wait is bounded by wall time and the training loop's per-step poll raises."""
    mp.spawn(_watchdog_worker, args=(2, _free_port()), nprocs=2, join=True)


def _freeze_worker(rank, world, port, out):
    ctx = _init(rank, world, port)
    from dla_b200.algo import dpo_step
    from dla_b200.config import HardwarePlan
    from dla_b200.data.synthetic import synthetic_preference_batch
    from dla_b200.models.factory import freeze_except_lora, load_causal_lm
    from dla_b200.optim import AdamWConfig

    res = {}
    for comm in ("nccl", "fused"):
        plan = HardwarePlan(sharding="zero3", comm=comm, resident_params=True)
        pol = load_causal_lm("tiny-llama", ctx, plan, seed=21, name=f"policy_{comm}")
        ref = load_causal_lm("tiny-llama", ctx, plan, trainable=False, gradient_checkpointing=False, seed=21, name=f"ref_{comm}")
        info = freeze_except_lora(pol.engine, "w_down")
        assert 0 < info["trainable_params"] < info["total_params"]
        names = [(u.name, p.name) for u in pol.engine.units for p in u.params]
        before = {n: pol.engine.gathered_unit(n[0])[n[1]].clone() for n in names}
        for step in range(3):
            full = synthetic_preference_batch(2 * world, 128, pol.config.vocab_size, seed=80 + step, ragged=True)
            batch = {s: {k: v[rank * 2:(rank + 1) * 2].cuda() for k, v in full[s].items()} for s in full}
            loss, _ = dpo_step(pol.model, ref.model, batch, beta=0.5)
            loss.backward()
            pol.engine.finish_backward()
            pol.engine.optimizer_step(AdamWConfig(lr=1e-3, weight_decay=0.1, max_grad_norm=1.0), grad_divisor=float(world))
        after = {n: pol.engine.gathered_unit(n[0])[n[1]] for n in names}
        for n in names:
            if "w_down" in f"{n[0]}.{n[1]}":
                assert not torch.equal(before[n], after[n]), (comm, n)
            else:
                assert torch.equal(before[n], after[n]), (comm, n, "a frozen parameter moved")
        res[comm] = (torch.cat([after[n].reshape(-1).float() for n in names]).cpu(), float(pol.engine.last_grad_norm))
    if rank == 0:
        torch.save(res, out)
    ctx.barrier()


def test_freeze_except_on_the_sharded_engines(tmp_path):
    """Substring freezing (P15) on ZeRO-3 with both engines: frozen parameters stay bit-identical on every rank (no weight
    decay, no update), the trainable ones move, the clipped norm counts trainable gradients only, fused == NCCL."""
    out = tmp_path / "freeze.pt"
    mp.spawn(_freeze_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    res = torch.load(out)
    (pn, gn), (pf, gf) = res["nccl"], res["fused"]
    assert gn == pytest.approx(gf, rel=3e-2)
    assert float((pn - pf).norm() / pn.norm()) < 3e-3
