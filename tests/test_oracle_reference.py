"""Oracle tests: our math vs the reference's own functions imported read-only from /root/reference
(SURVEY §4 item 1).  Skipped when the reference tree is absent."""
import sys
import types

import pytest
import torch

from conftest import REFERENCE_ROOT

pytestmark = pytest.mark.skipif(not REFERENCE_ROOT.exists(), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def ref():
    if "accelerate" not in sys.modules:
        import importlib.machinery

        import transformers  # noqa: F401  (resolve its optional-dependency probes before the stand-in exists)
        acc = types.ModuleType("accelerate")
        acc.__spec__ = importlib.machinery.ModuleSpec("accelerate", None)
        for n in ("Accelerator", "DeepSpeedPlugin", "FullyShardedDataParallelPlugin"):
            setattr(acc, n, type(n, (), {}))
        util = types.ModuleType("accelerate.utils")
        util.__spec__ = importlib.machinery.ModuleSpec("accelerate.utils", None)
        util.DistributedDataParallelKwargs = type("DistributedDataParallelKwargs", (), {})
        acc.utils = util
        sys.modules["accelerate"], sys.modules["accelerate.utils"] = acc, util
    sys.path.insert(0, str(REFERENCE_ROOT))
    try:
        from src.data import datasets as rdata
        from src.models import reward_model as rreward
        from src.training import train_dpo as rdpo
        from src.training import train_rlhf as rrlhf
    finally:
        sys.path.pop(0)
    return types.SimpleNamespace(data=rdata, reward=rreward, dpo=rdpo, rlhf=rrlhf)


def test_dpo_loss_matches(ref):
    from dla_b200.ops import functional as F

    torch.manual_seed(0)
    a, b, c, d = (torch.randn(9) for _ in range(4))
    want = ref.dpo.dpo_loss(a, b, c, d, 0.1)
    got, stats = F.dpo_loss(a, b, c, d, 0.1)
    assert torch.allclose(got, want, atol=1e-6)
    assert abs(stats[0].item() - (a > b).float().mean().item()) < 1e-6


def test_pairwise_loss_matches(ref):
    from dla_b200.ops import functional as F

    torch.manual_seed(1)
    c, r = torch.randn(11), torch.randn(11)
    got, acc = F.pairwise_loss(c, r)
    assert torch.allclose(got, ref.reward.pairwise_loss(c, r), atol=1e-6)


class _HFLike(torch.nn.Module):
    """Wraps our CausalLM in the ``model(input_ids=, attention_mask=).logits`` calling convention."""

    def __init__(self, m):
        super().__init__()
        self.m = m

    def forward(self, input_ids, attention_mask):
        return types.SimpleNamespace(logits=self.m.logits(input_ids, attention_mask))


def test_sequence_logprob_matches_reference_compute_logprobs(ref):
    from dla_b200.algo import sequence_logprobs
    from dla_b200.models import build_causal_lm

    m = build_causal_lm("tiny-llama", "cpu", trainable=False, seed=2)
    ids = torch.randint(4, 512, (3, 17))
    mask = torch.ones_like(ids)
    mask[1, 10:] = 0
    mask[2, 5:] = 0
    want = ref.dpo.compute_logprobs(_HFLike(m), ids, mask)
    want2 = ref.rlhf.sequence_logprob(_HFLike(m), ids, mask)
    got = sequence_logprobs(m, ids, mask)
    assert torch.allclose(got, want, atol=1e-5)
    assert torch.allclose(got, want2, atol=1e-5)


class _Backbone(torch.nn.Module):
    def __init__(self, m):
        super().__init__()
        self.m = m
        self.config = types.SimpleNamespace(hidden_size=m.cfg.hidden_size)

    def forward(self, input_ids, attention_mask):
        return types.SimpleNamespace(last_hidden_state=self.m.hidden(input_ids, attention_mask))


@pytest.mark.parametrize("pooling", ["last_token", "mean"])
def test_reward_model_matches(ref, pooling):
    from dla_b200.models import CausalLM, LocalParams, RewardModel, resolve_model_config, reward_units

    cfg = resolve_model_config("tiny-llama")
    prov = LocalParams(cfg, reward_units(cfg), "cpu", torch.float32, trainable=False, seed=4)
    ours = RewardModel(cfg, prov, pooling=pooling, dropout=0.0).eval()
    theirs = ref.reward.RewardModel(_Backbone(CausalLM(cfg, prov)), pooling=pooling, dropout=0.0).eval()
    w = prov.acquire("reward_head")
    with torch.no_grad():
        theirs.scorer[1].weight.copy_(w["score_w"])
        theirs.scorer[1].bias.copy_(w["score_b"])
    ids = torch.randint(4, 512, (4, 12))
    mask = torch.ones_like(ids)
    mask[0, 7:] = 0
    mask[3, 3:] = 0
    with torch.no_grad():
        assert torch.allclose(ours(ids, mask), theirs(ids, mask).float(), atol=1e-5)


def test_pad_batch_and_datasets_match(ref):
    from dla_b200.data import ByteTokenizer, InstructionDataset, PreferenceDataset, pad_batch

    tok = ByteTokenizer(vocab_size=300)
    items = [{"input_ids": torch.arange(5), "attention_mask": torch.ones(5, dtype=torch.long), "labels": torch.arange(5)},
             {"input_ids": torch.arange(3), "attention_mask": torch.ones(3, dtype=torch.long), "labels": torch.arange(3)}]
    want, got = ref.data.pad_batch(items, 7), pad_batch(items, 7)
    assert want.keys() == got.keys()
    for k in want:
        assert torch.equal(want[k], got[k]), k
    sft = [{"prompt": " What is 2+2? ", "response": "4 "}, {"prompt": "Name a colour", "response": "blue is a colour"}]
    a = ref.data.InstructionDataset(tok, 64, True, records=sft)
    b = InstructionDataset(tok, 64, True, records=sft)
    for i in range(2):
        for k in ("input_ids", "attention_mask", "labels"):
            assert torch.equal(a[i][k], b[i][k]), (i, k)
    ca, cb = a.collate([a[0], a[1]]), b.collate([b[0], b[1]])
    for k in ca:
        assert torch.equal(ca[k], cb[k]), k
    prefs = [{"prompt": "p1", "chosen": "good answer", "rejected": "bad"}, {"prompt": "p2 longer", "chosen": "x", "rejected": "yyyy yyy"}]
    pa, pb = ref.data.PreferenceDataset(tok, 32, records=prefs), PreferenceDataset(tok, 32, records=prefs)
    ca, cb = pa.collate([pa[0], pa[1]]), pb.collate([pb[0], pb[1]])
    for side in ("chosen", "rejected"):
        for k in ("input_ids", "attention_mask"):
            assert torch.equal(ca[side][k], cb[side][k]), (side, k)


def test_reinforce_loss_matches_reference_lines(ref):
    """The PPO 5-liner (R/src/training/train_rlhf.py:149-153) restated vs ours."""
    from dla_b200.algo import reinforce_step

    torch.manual_seed(3)
    pol = torch.randn(6, requires_grad=True)
    refl, score = torch.randn(6), torch.randn(6)
    kl = pol - refl
    rewards = score - 0.1 * kl
    adv = rewards - rewards.mean()
    want = -(adv.detach() * pol).mean()
    got, kl2, rw2 = reinforce_step(pol, refl, score, 0.1)
    assert torch.allclose(got, want, atol=1e-6)


def test_distill_kl_matches_reference_expression():
    """R/src/training/train_distill.py:132-144 restated inline vs the chunked implementation."""
    import torch.nn.functional as tF

    from dla_b200.algo import distill_step
    from dla_b200.models import build_causal_lm

    s = build_causal_lm("tiny-llama", "cpu", trainable=True, checkpoint=True, seed=1)
    t1 = build_causal_lm("tiny-llama", "cpu", trainable=False, seed=2)
    t2 = build_causal_lm("tiny-llama", "cpu", trainable=False, seed=3)
    ids = torch.randint(4, 512, (2, 10))
    mask = torch.ones_like(ids)
    mask[1, 6:] = 0
    batch = {"input_ids": ids, "attention_mask": mask, "labels": ids.clone()}
    got = distill_step(s, [t1, t2], batch, use_kl=True, chunk=7)
    with torch.no_grad():
        slp = tF.log_softmax(s.logits(ids, mask), -1)
        tp = torch.stack([tF.softmax(t.logits(ids, mask), -1) for t in (t1, t2)], 0).mean(0)
        kl = tF.kl_div(slp, tp, reduction="none").sum(-1)
        want = (kl * mask).sum() / mask.sum()
    assert torch.allclose(got, want, atol=1e-5)
    got.backward()
    assert sum(float(g.abs().sum()) for _, g in s.p.named_grads()) > 0
