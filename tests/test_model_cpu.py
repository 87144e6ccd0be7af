"""CPU checks of the model / checkpointing / engine plumbing (fp32, PyTorch reference ops)."""
import copy

import pytest
import torch

from dla_b200.algo import dpo_step, sequence_logprobs, sft_loss
from dla_b200.data.synthetic import synthetic_lm_batch, synthetic_preference_batch
from dla_b200.models import CausalLM, LocalParams, build_units, resolve_model_config
from dla_b200.models.transformer import collect_autograd_grads


def _grads(model):
    for u in model.p.units:
        collect_autograd_grads(model.p.acquire(u.name))
    return {n: g.clone() for n, g in model.p.named_grads()}


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-gpt2"])
def test_checkpointed_backward_equals_plain_autograd(name):
    cfg = resolve_model_config(name)
    units = build_units(cfg)
    a = LocalParams(cfg, units, "cpu", torch.float32, trainable=True, checkpoint=True, seed=3)
    b = LocalParams(cfg, units, "cpu", torch.float32, trainable=True, checkpoint=False, init=False)
    for u in units:
        b.flat[u.name].copy_(a.flat[u.name])
    ma, mb = CausalLM(cfg, a), CausalLM(cfg, b)
    batch = synthetic_lm_batch(2, 24, cfg.vocab_size, seed=1)
    batch["attention_mask"][1, 17:] = 0
    batch["labels"][1, 17:] = -100
    la, lb = sft_loss(ma, batch), sft_loss(mb, batch)
    assert torch.allclose(la, lb, atol=1e-6)
    la.backward()
    lb.backward()
    ga, gb = _grads(ma), _grads(mb)
    for k in ga:
        assert gb[k].abs().sum() > 0, f"{k} has no gradient"
        assert torch.allclose(ga[k], gb[k], atol=1e-5, rtol=1e-4), k


def test_llama_matches_hf_llama():
    """Our fused-QKV / fused-gate-up Llama block equals transformers' LlamaForCausalLM on the same weights."""
    transformers = pytest.importorskip("transformers")
    from dla_b200.models.io import export_unit

    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    p = LocalParams(cfg, units, "cpu", torch.float32, trainable=False, seed=5)
    m = CausalLM(cfg, p)
    hf_cfg = transformers.LlamaConfig(**{k: v for k, v in cfg.to_hf_dict().items()
                                         if k not in ("model_type", "architectures", "dla_name", "torch_dtype")},
                                      attn_implementation="eager")
    hf = transformers.LlamaForCausalLM(hf_cfg).eval()
    sd = {}
    for u in units:
        sd.update(export_unit(cfg, u.name, p.acquire(u.name)))
    missing, unexpected = hf.load_state_dict(sd, strict=False)
    assert not [k for k in missing if "rotary" not in k], missing
    ids = torch.randint(4, cfg.vocab_size, (2, 20))
    mask = torch.ones_like(ids)
    mask[1, 13:] = 0
    with torch.no_grad():
        ours = m.logits(ids, mask)
        theirs = hf(input_ids=ids, attention_mask=mask).logits
    valid = mask.bool()
    assert torch.allclose(ours[valid], theirs[valid], atol=2e-4, rtol=1e-3)


def test_dpo_concat_equals_separate_and_learns():
    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    pol = LocalParams(cfg, units, "cpu", torch.float32, trainable=True, checkpoint=True, seed=1)
    ref = LocalParams(cfg, units, "cpu", torch.float32, trainable=False, seed=1)
    mp, mr = CausalLM(cfg, pol), CausalLM(cfg, ref)
    batch = synthetic_preference_batch(3, 16, cfg.vocab_size, seed=2, ragged=True)
    l1, i1 = dpo_step(mp, mr, batch, 0.3, concat=True)
    l2, i2 = dpo_step(mp, mr, batch, 0.3, concat=False)
    assert torch.allclose(l1, l2, atol=1e-6)
    assert torch.allclose(i1["pol_pos"], i2["pol_pos"], atol=1e-5)
    assert abs(l1.item() - 0.693147) < 1e-5
    losses = []
    for _ in range(6):
        pol.zero_grad()
        loss, _ = dpo_step(mp, mr, batch, 0.3)
        loss.backward()
        for u in units:
            w = pol.acquire(u.name)
            collect_autograd_grads(w)
            pol.flat[u.name].data.add_(pol.grad[u.name], alpha=-0.5)
        losses.append(loss.item())
    assert losses[-1] < losses[0] - 0.1, losses


def test_no_recompute_layers_give_same_gradients():
    from dla_b200.models.config import ModelConfig
    from dla_b200.parallel.engine import Engine
    from dla_b200.parallel.pg import DistContext

    cfg = resolve_model_config("tiny-llama")
    units = build_units(cfg)
    ctx = DistContext()
    a = Engine(cfg, units, ctx, trainable=True, checkpoint=True, dtype=torch.float32, seed=2)
    b = Engine(cfg, units, ctx, trainable=True, checkpoint=True, dtype=torch.float32, seed=2)
    assert b.set_no_recompute_layers(1) == 1 and b.no_recompute == frozenset({"layer.1"})
    batch = synthetic_lm_batch(2, 16, cfg.vocab_size, seed=4)
    for eng in (a, b):
        loss = sft_loss(CausalLM(cfg, eng), batch)
        loss.backward()
        eng.finish_backward()
    for u in units:
        assert torch.allclose(a.gradbuf[u.name], b.gradbuf[u.name], atol=1e-6), u.name
        assert a.gradbuf[u.name].abs().sum() > 0
