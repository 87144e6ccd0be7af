"""End-to-end model numerics on the GPU: native bf16 kernels vs the fp32 PyTorch restatement with the
same weights, and a few DPO updates through the engine."""
import pytest
import torch

from dla_b200.algo import dpo_step, sequence_logprobs
from dla_b200.data.synthetic import synthetic_preference_batch
from dla_b200.models import CausalLM, LocalParams, build_units, resolve_model_config

pytestmark = pytest.mark.gpu


def _pair(name, seed=0):
    cfg = resolve_model_config(name)
    units = build_units(cfg)
    gpu = LocalParams(cfg, units, "cuda", torch.bfloat16, trainable=True, checkpoint=True, seed=seed)
    cpu = LocalParams(cfg, units, "cpu", torch.float32, trainable=True, checkpoint=False, init=False)
    for u in units:
        cpu.flat[u.name].copy_(gpu.flat[u.name].float().cpu())
    return cfg, CausalLM(cfg, gpu), CausalLM(cfg, cpu)


@pytest.mark.parametrize("name,T", [("tiny-llama", 128), ("tiny-llama", 72), ("tiny-gpt2", 64)])
def test_model_forward_backward_matches_reference(native_lib, name, T):
    torch.manual_seed(0)
    cfg, mg, mc = _pair(name)
    B = 3
    ids = torch.randint(4, cfg.vocab_size, (B, T))
    lens = torch.tensor([T, T - 9, T // 2])
    mask = (torch.arange(T)[None] < lens[:, None]).long()
    ids = ids * mask
    lp_g = sequence_logprobs(mg, ids.cuda(), mask.cuda())
    lp_c = sequence_logprobs(mc, ids, mask)
    assert torch.allclose(lp_g.float().cpu(), lp_c, atol=3e-2, rtol=2e-2), (lp_g, lp_c)
    w = torch.tensor([1.0, -2.0, 0.5])
    (lp_g * w.cuda()).sum().backward()
    (lp_c * w).sum().backward()
    from dla_b200.models.transformer import collect_autograd_grads
    for u in mc.p.units:
        collect_autograd_grads(mc.p.acquire(u.name))
    for (n, gg), (_, gc) in zip(mg.p.named_grads(), mc.p.named_grads()):
        gg, gc = gg.float().cpu(), gc.float()
        denom = gc.norm().clamp(min=1e-6)
        rel = ((gg - gc).norm() / denom).item()
        assert rel < 8e-2 or gc.norm() < 1e-4, f"{name}:{n} grad rel err {rel:.3f} (|g|={gc.norm():.3g})"


def test_dpo_updates_reduce_loss(native_lib):
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.pg import init_distributed

    ctx = init_distributed()
    policy = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=True, seed=1, name="policy")
    ref = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=False, trainable=False, seed=1, name="reference")
    batch = synthetic_preference_batch(4, 128, policy.config.vocab_size, seed=3, ragged=True)
    batch = {s: {k: v.cuda() for k, v in batch[s].items()} for s in batch}
    losses = []
    for _ in range(8):
        loss, info = dpo_step(policy.model, ref.model, batch, beta=0.5)
        loss.backward()
        policy.engine.finish_backward()
        policy.engine.optimizer_step(AdamWConfig(lr=2e-3, weight_decay=0.0))
        losses.append(loss.item())
    assert abs(losses[0] - 0.6931) < 1e-3          # identical policy and reference at step 0
    assert losses[-1] < losses[0] - 0.05, losses
    assert float(info["preference_rate"]) >= 0.75


@pytest.mark.parametrize("no_recompute", [0, 2])
def test_engine_fast_paths_are_bit_identical_to_plain(native_lib, no_recompute):
    """The engine's step-level optimisations change scheduling, not numbers: lazy gradient-bucket zeroing (first
    wgrad GEMM overwrites) and the side-stream optimizer overlap must give bit-identical parameters to the plain
    path (memset buckets, AdamW on the main stream) after several DPO steps, with gradient accumulation."""
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.pg import init_distributed

    ctx = init_distributed()
    ref = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=False, trainable=False, seed=1, name="reference")
    # every token id occurs once per micro-batch and T < 128 (one kv tile): the embedding scatter-add and the dQ
    # reduce-add then have a single contribution per address, so the backward is bit-reproducible
    batches = []
    gen = torch.Generator().manual_seed(5)
    for i in range(6):
        ids = (torch.randperm(ref.config.vocab_size - 4, generator=gen)[:4 * 96] + 4).view(4, 96).cuda()
        ones = torch.ones_like(ids)
        batches.append({"chosen": {"input_ids": ids[:2], "attention_mask": ones[:2]},
                        "rejected": {"input_ids": ids[2:], "attention_mask": ones[2:]}})
    results = []
    for fast in (False, True):
        pol = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=True, seed=2, name="policy")
        eng = pol.engine
        eng.lazy_zero = fast
        eng.overlap_optimizer = fast
        eng.set_no_recompute_layers(no_recompute)
        opt = AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
        for step in range(3):
            for micro in range(2):                       # two micro-steps accumulate into the same buckets
                loss, _ = dpo_step(pol.model, ref.model, batches[2 * step + micro], beta=0.3)
                loss.backward()
                eng.finish_backward()
            eng.optimizer_step(opt, grad_divisor=2.0)
        eng.sync_optimizer()
        torch.cuda.synchronize()
        results.append({u.name: eng.shard[u.name].clone() for u in eng.units})
    for name in results[0]:
        assert torch.equal(results[0][name], results[1][name]), f"unit {name} differs between plain and fast paths"


def test_grad_norm_summed_during_backward_matches_the_pass_after_it(native_lib):
    """One GPU: each unit's sum of squares is taken on the side stream as soon as the unit's backward is done.  Same
    norm (to fp32 summation order) and same parameters as the single pass after backward; a second backward before the
    optimizer step falls back to that pass."""
    from dla_b200.models.factory import load_causal_lm
    from dla_b200.optim import AdamWConfig
    from dla_b200.parallel.pg import init_distributed

    ctx = init_distributed()
    ref = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=False, trainable=False, seed=1, name="reference")
    gen = torch.Generator().manual_seed(11)
    batches = []
    for i in range(4):
        ids = (torch.randperm(ref.config.vocab_size - 4, generator=gen)[:4 * 96] + 4).view(4, 96).cuda()
        ones = torch.ones_like(ids)
        batches.append({"chosen": {"input_ids": ids[:2], "attention_mask": ones[:2]},
                        "rejected": {"input_ids": ids[2:], "attention_mask": ones[2:]}})
    out = []
    for early in (False, True):
        pol = load_causal_lm("tiny-llama", ctx, gradient_checkpointing=True, seed=2, name="policy")
        eng = pol.engine
        eng.norm_in_backward = early
        eng.set_no_recompute_layers(1)
        opt = AdamWConfig(lr=1e-3, weight_decay=0.01, max_grad_norm=0.05)       # small enough that the clip engages
        norms = []
        for step in range(3):
            loss, _ = dpo_step(pol.model, ref.model, batches[step], beta=0.3)
            loss.backward()
            eng.finish_backward()
            assert len(eng._sq_launched) == (len(eng.units) if early else 0)
            eng.optimizer_step(opt)
            norms.append(float(eng._last_grad_norm))
        # two backwards into the same buckets, then one step: the early partial sums are stale and must not be used
        for micro in range(2):
            loss, _ = dpo_step(pol.model, ref.model, batches[3], beta=0.3)
            loss.backward()
            eng.finish_backward()
        eng.optimizer_step(opt, grad_divisor=2.0)
        norms.append(float(eng._last_grad_norm))
        # a backward whose gradients are thrown away must not leave its partial sums behind
        loss, _ = dpo_step(pol.model, ref.model, batches[0], beta=0.3)
        loss.backward()
        eng.finish_backward()
        eng.zero_grad()
        loss, _ = dpo_step(pol.model, ref.model, batches[1], beta=0.3)
        loss.backward()
        eng.finish_backward()
        eng.optimizer_step(opt)
        norms.append(float(eng._last_grad_norm))
        eng.sync_optimizer()
        torch.cuda.synchronize()
        out.append((norms, {u.name: eng.shard[u.name].float().clone() for u in eng.units}))
    (n0, p0), (n1, p1) = out
    assert all(n > 0.05 for n in n0), n0
    for a, b in zip(n0, n1):
        assert abs(a - b) <= 1e-5 * abs(a), (n0, n1)
    for name in p0:
        assert torch.allclose(p0[name], p1[name], rtol=1e-2, atol=1e-3), f"unit {name} differs"


def test_packed_rows_equal_the_documents_run_one_by_one(native_lib):
    """``data.packing`` on the GPU: three documents packed into one row (segment ids + per-document positions) give
    each document the same token log-probs and the same weight gradients as running the documents as separate rows."""
    torch.manual_seed(5)
    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    lens = [150, 61, 200]            # documents straddle the 128-token tile boundaries of the flash kernels
    T = 448
    docs = [torch.randint(4, cfg.vocab_size, (n,)) for n in lens]

    def model():
        return CausalLM(cfg, LocalParams(cfg, units, "cuda", torch.bfloat16, trainable=True, checkpoint=True, seed=9))

    # packed: one row
    ids = torch.zeros(1, T, dtype=torch.long)
    seg = torch.zeros(1, T, dtype=torch.long)
    pos = torch.zeros(1, T, dtype=torch.int32)
    o = 0
    for k, d in enumerate(docs):
        ids[0, o:o + len(d)], seg[0, o:o + len(d)], pos[0, o:o + len(d)] = d, k + 1, torch.arange(len(d), dtype=torch.int32)
        o += len(d)
    mp_ = model()
    lp = mp_.token_logprobs(ids.cuda(), (seg > 0).long().cuda(), position_ids=pos.cuda(), segment_ids=seg.cuda())[0]
    # separate rows, right padded
    Tm = 256
    ids2 = torch.zeros(3, Tm, dtype=torch.long)
    mask2 = torch.zeros(3, Tm, dtype=torch.long)
    for k, d in enumerate(docs):
        ids2[k, :len(d)], mask2[k, :len(d)] = d, 1
    ms = model()
    lp2 = ms.token_logprobs(ids2.cuda(), mask2.cuda())
    o = 0
    tot_p, tot_s = 0, 0
    for k, n in enumerate(lens):
        a, b = lp[o:o + n - 1].float(), lp2[k, :n - 1].float()       # the last token of a document predicts nothing of it
        assert torch.allclose(a, b, atol=3e-2, rtol=2e-2), (k, (a - b).abs().max())
        tot_p, tot_s = tot_p + a.sum(), tot_s + b.sum()
        o += n
    tot_p.backward()
    tot_s.backward()
    for (n, gp), (_, gs) in zip(mp_.p.named_grads(), ms.p.named_grads()):
        gp, gs = gp.float(), gs.float()
        rel = ((gp - gs).norm() / gs.norm().clamp(min=1e-6)).item()
        assert rel < 5e-2 or gs.norm() < 1e-4, f"{n}: packed vs separate grad rel err {rel:.3f}"
