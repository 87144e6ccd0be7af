"""Reference arm of bench.py: the reference's own DPO step on stock PyTorch / transformers.

What runs: ``compute_logprobs`` and ``dpo_loss`` imported UNMODIFIED from the reference tree
(``baseline/_ref/distributed-llm-alignment/src/training/train_dpo.py``, falling back to ``/root/reference``),
driven in the order of its hot loop (``R/src/training/train_dpo.py:110-121``: four log-prob passes,
loss, backward, clip_grad_norm_, AdamW step, zero_grad) on ``transformers.LlamaForCausalLM`` (cuBLAS GEMMs,
PyTorch SDPA, ATen log_softmax) with torch DDP / FSDP2 + NCCL.  None of dla_b200's models, kernels or
engine is on this path.

Why not ``pip install`` + the reference's ``main()``: the tree has no setup.py/pyproject (pip refuses it,
see DESIGN.md) and its trainers import ``accelerate`` — not installed, not installable offline — and
load weights with ``from_pretrained(..., device_map="auto")`` from the hub.  So a stand-in ``accelerate``
module is registered only to let the file import; the model is random-init from the named config, and
the distributed wrapping accelerate would do (``prepare``) is done with the torch primitives it delegates to.
"""
from __future__ import annotations

import os
import sys
import types
from pathlib import Path
from typing import Any, Dict

import torch
import torch.distributed as dist

HERE = Path(__file__).resolve().parent
CANDIDATES = [HERE / "_ref" / "distributed-llm-alignment", Path("/root/reference/distributed-llm-alignment")]


def locate() -> Path | None:
    for c in CANDIDATES:
        if (c / "src" / "training" / "train_dpo.py").exists():
            return c
    return None


def import_reference_fns():
    root = locate()
    if root is None:
        raise FileNotFoundError("reference tree not found under baseline/_ref or /root/reference")
    if "accelerate" not in sys.modules:
        try:
            import accelerate  # noqa: F401
        except Exception:
            import importlib.machinery

            import transformers  # noqa: F401  (resolve its optional-dependency probes before the stand-in exists)
            acc = types.ModuleType("accelerate")
            acc.__spec__ = importlib.machinery.ModuleSpec("accelerate", None)
            for n in ("Accelerator", "DeepSpeedPlugin", "FullyShardedDataParallelPlugin"):
                setattr(acc, n, type(n, (), {}))
            util = types.ModuleType("accelerate.utils")
            util.__spec__ = importlib.machinery.ModuleSpec("accelerate.utils", None)
            util.DistributedDataParallelKwargs = type("DistributedDataParallelKwargs", (), {"__init__": lambda self, **k: None})
            acc.utils = util
            sys.modules["accelerate"], sys.modules["accelerate.utils"] = acc, util
    sys.path.insert(0, str(root))
    try:
        from src.training.train_dpo import compute_logprobs, dpo_loss  # type: ignore
    finally:
        sys.path.pop(0)
    return compute_logprobs, dpo_loss, root


def llama_hf_config(name: str):
    from transformers import LlamaConfig

    presets = {
        "llama3-8b": dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                          num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=8192,
                          rope_theta=500000.0, rms_norm_eps=1e-5),
        "llama3-70b": dict(vocab_size=128256, hidden_size=8192, intermediate_size=28672, num_hidden_layers=80,
                           num_attention_heads=64, num_key_value_heads=8, max_position_embeddings=8192,
                           rope_theta=500000.0, rms_norm_eps=1e-5),
        "llama-1b": dict(vocab_size=32000, hidden_size=2048, intermediate_size=5632, num_hidden_layers=16,
                         num_attention_heads=16, num_key_value_heads=4, max_position_embeddings=4096,
                         rope_theta=500000.0, rms_norm_eps=1e-5),
        "tiny-llama": dict(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                           num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=512,
                           rope_theta=10000.0, rms_norm_eps=1e-5),
    }
    cfg = LlamaConfig(**presets[name], tie_word_embeddings=False, attn_implementation="sdpa")
    cfg.use_cache = False
    return cfg


def param_bytes(cfg) -> int:
    """bf16 bytes of one LlamaForCausalLM of this config (untied head)."""
    h, f, L, v = cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers, cfg.vocab_size
    kv = h // cfg.num_attention_heads * cfg.num_key_value_heads
    per_layer = h * (h + 2 * kv) + h * h + 3 * h * f + 2 * h
    return 2 * (L * per_layer + 2 * v * h + h)


def build_on_meta(cfg):
    """The module tree with no storage behind it: what accelerate/DeepSpeed ZeRO-3 do for a model that does not
    fit one GPU (``zero3_init_flag``), here with torch's meta device."""
    from transformers import LlamaForCausalLM

    with torch.device("meta"):
        return LlamaForCausalLM(cfg).to(torch.bfloat16)


@torch.no_grad()
def materialize_sharded(model, cfg, device: torch.device, seed: int) -> None:
    """Allocate only this rank's shards and fill them with HF's initialisation (normal(0, initializer_range)
    for matrices and embeddings, ones for the norm gains); rebuild the rotary tables the meta device dropped."""
    model.to_empty(device=device)
    gen = torch.Generator(device=device).manual_seed(seed)
    for _, p in model.named_parameters():
        local = p.to_local() if hasattr(p, "to_local") else p
        if p.dim() == 1:
            local.fill_(1.0)
        else:
            local.normal_(0.0, cfg.initializer_range, generator=gen)
    rot = model.model.rotary_emb
    model.model.rotary_emb = type(rot)(config=cfg, device=device)


class ReferenceDPO:
    """Policy + frozen reference as the reference builds them (policy with gradient checkpointing,
    reference without, AdamW(lr, weight_decay=0.01), max_grad_norm 1.0) wrapped for N ranks."""

    def __init__(self, model_name: str, device: torch.device, world: int, lr: float = 1e-6, beta: float = 0.1,
                 sharded: bool = True) -> None:
        from transformers import LlamaForCausalLM

        self.compute_logprobs, self.dpo_loss, self.root = import_reference_fns()
        cfg = llama_hf_config(model_name)
        self.device, self.world, self.beta = device, world, beta
        torch.manual_seed(0)
        # two replicas that would not fit one GPU (70B) are built on the meta device and materialised shard by shard
        hbm = torch.cuda.get_device_properties(device).total_memory if device.type == "cuda" else 1 << 62
        self.meta_init = world > 1 and sharded and (2 * param_bytes(cfg) > 0.5 * hbm or os.environ.get("DLA_REF_META_INIT") == "1")
        if self.meta_init:
            policy, ref = build_on_meta(cfg), build_on_meta(cfg)
        else:
            with torch.device(device):
                policy = LlamaForCausalLM(cfg).to(torch.bfloat16)
                ref = LlamaForCausalLM(cfg).to(torch.bfloat16)
        policy.gradient_checkpointing_enable()
        policy.train()      # HF only checkpoints in train mode; Llama has no dropout so the math is unchanged
        ref.eval()
        for p in ref.parameters():
            p.requires_grad = False
        if world > 1:
            if sharded:
                from torch.distributed.fsdp import fully_shard

                for m in (policy, ref):
                    for layer in m.model.layers:
                        fully_shard(layer)
                    fully_shard(m)
                if self.meta_init:
                    rank = dist.get_rank()
                    materialize_sharded(policy, cfg, device, seed=1000 + rank)
                    materialize_sharded(ref, cfg, device, seed=2000 + rank)
            else:
                policy = torch.nn.parallel.DistributedDataParallel(policy, device_ids=[device.index], find_unused_parameters=False)
        self.policy, self.ref = policy, ref
        self.opt = torch.optim.AdamW(policy.parameters(), lr=lr, weight_decay=0.01)

    def step(self, batch: Dict[str, Dict[str, torch.Tensor]]) -> torch.Tensor:
        ch, rj = batch["chosen"], batch["rejected"]
        with torch.autocast("cuda", dtype=torch.bfloat16):      # accelerate mixed_precision=bf16 wraps forward
            pol_pos = self.compute_logprobs(self.policy, ch["input_ids"], ch["attention_mask"])
            pol_neg = self.compute_logprobs(self.policy, rj["input_ids"], rj["attention_mask"])
            ref_pos = self.compute_logprobs(self.ref, ch["input_ids"], ch["attention_mask"])
            ref_neg = self.compute_logprobs(self.ref, rj["input_ids"], rj["attention_mask"])
            loss = self.dpo_loss(pol_pos, pol_neg, ref_pos, ref_neg, self.beta)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(self.policy.parameters(), 1.0)
        self.opt.step()
        self.opt.zero_grad()
        return loss.detach()


def synthetic_preference_batch(batch: int, seq_len: int, vocab_size: int, seed: int = 0,
                               pin: bool = False) -> Dict[str, Dict[str, torch.Tensor]]:
    """Same shapes, seeds and value range as the own arm's generator (``dla_b200/data/synthetic.py``), restated
    with plain torch so this process never imports the framework under test."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, Dict[str, torch.Tensor]] = {}
    for side in ("chosen", "rejected"):
        ids = torch.randint(4, vocab_size, (batch, seq_len), generator=g, dtype=torch.long)
        mask = torch.ones(batch, seq_len, dtype=torch.long)
        if pin and torch.cuda.is_available():
            ids, mask = ids.pin_memory(), mask.pin_memory()
        out[side] = {"input_ids": ids, "attention_mask": mask}
    return out


class ReferenceArm:
    """Everything ``bench.py --impl reference`` needs, on stock torch only: rendezvous, barrier, max over ranks,
    data, and the step.  ``bench.py`` drives it through the same timing loop as the own arm."""

    impl = "reference"

    def __init__(self, args: Any) -> None:
        import datetime

        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        local = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        self.device = torch.device("cuda", local % max(torch.cuda.device_count(), 1))
        torch.cuda.set_device(self.device)
        if self.world > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            dist.init_process_group("nccl", rank=self.rank, world_size=self.world, device_id=self.device,
                                    timeout=datetime.timedelta(seconds=1800))
        self.is_main = self.rank == 0
        self.runner = ReferenceDPO(args.model, self.device, self.world, sharded=(args.sharding != "ddp"))
        self.vocab = llama_hf_config(args.model).vocab_size
        self.note = ("reference compute_logprobs/dpo_loss unmodified from baseline/_ref, HF LlamaForCausalLM + torch "
                     "AdamW + FSDP2/NCCL; accelerate/deepspeed absent so its main() cannot run (see DESIGN.md)")

    # -- plumbing --------------------------------------------------------------------------------
    def barrier(self) -> None:
        if self.world > 1:
            dist.barrier(device_ids=[self.device.index])

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, dist.ReduceOp.MAX)
        return float(t.item())

    def make_batch(self, mb: int, T: int, seed: int):
        return synthetic_preference_batch(mb, T, self.vocab, seed=seed, pin=True)

    # -- the step --------------------------------------------------------------------------------
    def step(self, batch):
        return self.runner.step(batch)

    def after_warmup_step(self, i: int, tokens: int) -> None:
        pass

    def retry_after_oom(self) -> bool:
        return False

    def before_stop(self) -> None:
        pass

    def launches(self) -> int:
        return 0        # none of this repo's kernels exist in this process; see ``native_so_in_process``

    def describe(self) -> Dict[str, Any]:
        return {"optimizer": "torch AdamW on bf16 params", "activation_checkpointing": True,
                "meta_init": bool(self.runner.meta_init)}

    @staticmethod
    def native_so_in_process() -> bool:
        """True if this repo's kernel library got mapped into the process after all (it must not be)."""
        try:
            return "libdla_b200" in open("/proc/self/maps").read()
        except OSError:
            return False
