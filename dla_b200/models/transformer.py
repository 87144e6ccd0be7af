"""Native decoder-only transformers (Llama/Mistral family and GPT-2) built on :mod:`dla_b200.ops`.

Design (B200-first, not an nn.Module tree):
  * parameters live in one flat bf16 buffer per *unit* (embedding, each decoder layer, head); a unit
    is the granule the sharded engine all-gathers / reduce-scatters and the
    granule of activation checkpointing (``R/src/models/base_model.py:36-37``).
  * QKV and gate/up projections are stored fused so each is ONE tcgen05 GEMM; RoPE runs in place on
    the packed QKV; residual adds are fused into the RMSNorm kernels.
  * a :class:`ParamProvider` hands out views of the unit's parameters for forward and again for the
    recompute+backward pass; the distributed engine implements the same interface.

Reference behaviour replaced: ``transformers.AutoModelForCausalLM`` forward/backward reached from
``R/src/training/train_dpo.py:32`` and friends.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from dla_b200.models.config import ModelConfig
from dla_b200.ops import functional as F

Weights = Dict[str, torch.Tensor]


# =======================================================================================
# parameter specs / units
# =======================================================================================

@dataclass
class ParamSpec:
    name: str
    shape: Tuple[int, ...]
    init: str = "normal"       # normal | ones | zeros | scaled (normal / sqrt(2L))
    decay: bool = True

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass
class UnitSpec:
    name: str
    params: List[ParamSpec]

    @property
    def numel(self) -> int:
        return sum(p.numel for p in self.params)

    def offsets(self, align: int = 8) -> Dict[str, Tuple[int, int]]:
        """name -> (offset, numel) inside the unit's flat buffer; every param starts 16-byte aligned."""
        out, off = {}, 0
        for p in self.params:
            out[p.name] = (off, p.numel)
            off += -(-p.numel // align) * align
        return out

    def padded_numel(self, multiple: int = 8, align: int = 8) -> int:
        off = 0
        for p in self.params:
            off += -(-p.numel // align) * align
        return -(-off // multiple) * multiple


def build_units(cfg: ModelConfig, with_lm_head: bool = True) -> List[UnitSpec]:
    H, Fi, V = cfg.hidden_size, cfg.intermediate_size, cfg.vocab_size
    units: List[UnitSpec] = []
    if cfg.arch == "gpt2":
        units.append(UnitSpec("embed", [ParamSpec("tok_emb", (V, H)), ParamSpec("pos_emb", (cfg.max_position_embeddings, H), decay=False)]))
        for i in range(cfg.num_layers):
            units.append(UnitSpec(f"layer.{i}", [
                ParamSpec("ln1_w", (H,), "ones", False), ParamSpec("ln1_b", (H,), "zeros", False),
                ParamSpec("wqkv", (3 * H, H)), ParamSpec("bqkv", (3 * H,), "zeros", False),
                ParamSpec("wo", (H, H), "scaled"), ParamSpec("bo", (H,), "zeros", False),
                ParamSpec("ln2_w", (H,), "ones", False), ParamSpec("ln2_b", (H,), "zeros", False),
                ParamSpec("w_fc", (Fi, H)), ParamSpec("b_fc", (Fi,), "zeros", False),
                ParamSpec("w_proj", (H, Fi), "scaled"), ParamSpec("b_proj", (H,), "zeros", False),
            ]))
        head = [ParamSpec("lnf_w", (H,), "ones", False), ParamSpec("lnf_b", (H,), "zeros", False)]
    else:
        units.append(UnitSpec("embed", [ParamSpec("tok_emb", (V, H))]))
        qkv = cfg.q_dim + 2 * cfg.kv_dim
        for i in range(cfg.num_layers):
            units.append(UnitSpec(f"layer.{i}", [
                ParamSpec("attn_norm", (H,), "ones", False),
                ParamSpec("wqkv", (qkv, H)),
                ParamSpec("wo", (H, cfg.q_dim), "scaled"),
                ParamSpec("mlp_norm", (H,), "ones", False),
                ParamSpec("w_gu", (2 * Fi, H)),
                ParamSpec("w_down", (H, Fi), "scaled"),
            ]))
        head = [ParamSpec("final_norm", (H,), "ones", False)]
    if with_lm_head and not cfg.tie_word_embeddings:
        head.append(ParamSpec("lm_head", (V, H)))
    units.append(UnitSpec("head", head))
    return units


def init_unit_(flat: torch.Tensor, unit: UnitSpec, cfg: ModelConfig, generator: Optional[torch.Generator] = None) -> None:
    """Random-initialise a unit's flat buffer in place (normal(0, 0.02), residual projections scaled
    by 1/sqrt(2L) like GPT-2/Llama reference inits)."""
    offs = unit.offsets()
    for p in unit.params:
        o, n = offs[p.name]
        view = flat[o:o + n]
        if p.init == "ones":
            view.fill_(1.0)
        elif p.init == "zeros":
            view.zero_()
        else:
            std = cfg.initializer_range
            if p.init == "scaled":
                std = std / math.sqrt(2 * cfg.num_layers)
            # draw in fp32 chunks to bound peak memory for 1 GB embeddings
            step = 1 << 24
            for s in range(0, n, step):
                e = min(s + step, n)
                view[s:e].copy_(torch.randn(e - s, device=flat.device, dtype=torch.float32, generator=generator) * std)


def unit_views(flat: torch.Tensor, unit: UnitSpec) -> Weights:
    offs = unit.offsets()
    return {p.name: flat[offs[p.name][0]: offs[p.name][0] + p.numel].view(p.shape) for p in unit.params}


# =======================================================================================
# parameter providers
# =======================================================================================

class ParamProvider:
    """Interface between the model and whoever owns the parameters."""

    trainable: bool = True
    checkpoint: bool = True
    no_recompute: frozenset = frozenset()     # units that keep their activations instead of recomputing (HBM permitting)

    def note_plain(self, unit: str) -> None:
        pass

    def acquire(self, unit: str, backward: bool = False) -> Weights:  # pragma: no cover - interface
        raise NotImplementedError

    def release(self, unit: str, backward: bool = False) -> None:
        pass

    def grads_done(self, unit: str, weights: Weights) -> None:
        pass

    def unit_backward_done(self, unit: str) -> None:
        """A unit that kept its activations has finished its backward (see ``_UnitBoundary``)."""
        pass


class LocalParams(ParamProvider):
    """All parameters resident on one device, one flat buffer per unit (+ a flat grad buffer when
    trainable).  Used single-process and as the building block of the replicated engine."""

    def __init__(self, cfg: ModelConfig, units: Sequence[UnitSpec], device: torch.device | str,
                 dtype: torch.dtype = torch.bfloat16, trainable: bool = True, checkpoint: bool = True,
                 seed: int = 0, init: bool = True, grad_dtype: Optional[torch.dtype] = None) -> None:
        self.cfg, self.units = cfg, list(units)
        self.device, self.dtype = torch.device(device), dtype
        self.trainable, self.checkpoint = trainable, checkpoint
        self.flat: Dict[str, torch.Tensor] = {}
        self.grad: Dict[str, torch.Tensor] = {}
        gen = torch.Generator(device=self.device).manual_seed(seed) if init else None
        for u in self.units:
            buf = torch.zeros(u.padded_numel(), dtype=dtype, device=self.device)
            if init:
                init_unit_(buf, u, cfg, gen)
            self.flat[u.name] = buf
            if trainable:
                self.grad[u.name] = torch.zeros(u.padded_numel(), dtype=grad_dtype or dtype, device=self.device)
        self._by_name = {u.name: u for u in self.units}
        self._views: Dict[str, Weights] = {}

    def acquire(self, unit: str, backward: bool = False) -> Weights:
        if unit not in self._views:
            spec = self._by_name[unit]
            w = unit_views(self.flat[unit], spec)
            if self.trainable:
                g = unit_views(self.grad[unit], spec)
                for k, t in w.items():
                    t.requires_grad_(True)
                    t._grad_slot = g[k]            # native ops accumulate weight grads here
            self._views[unit] = w
        return self._views[unit]

    def grads_done(self, unit: str, weights: Weights) -> None:
        collect_autograd_grads(weights)

    def unit_backward_done(self, unit: str) -> None:
        collect_autograd_grads(self._views.get(unit, {}))

    # convenience -----------------------------------------------------------------------
    def zero_grad(self) -> None:
        for g in self.grad.values():
            g.zero_()

    def named_parameters(self):
        for u in self.units:
            for k, t in self.acquire(u.name).items():
                yield f"{u.name}.{k}", t

    def named_grads(self):
        for u in self.units:
            g = unit_views(self.grad[u.name], u)
            for k, t in g.items():
                yield f"{u.name}.{k}", t

    def num_params(self) -> int:
        return sum(u.numel for u in self.units)


def collect_autograd_grads(weights: Weights) -> None:
    """CPU / eager path: ops that went through plain autograd leave ``.grad`` on the views; fold it
    into the flat gradient slots so both paths present gradients the same way."""
    for t in weights.values():
        if t.grad is not None:
            slot = getattr(t, "_grad_slot", None)
            if slot is not None:
                if getattr(slot, "_fresh", False):      # lazily zeroed bucket slice (see ops.functional._slot_add)
                    slot.zero_()
                    slot._fresh = False
                slot.add_(t.grad.to(slot.dtype))
                t.grad = None


# =======================================================================================
# unit execution with activation checkpointing
# =======================================================================================

class _CheckpointedUnit(torch.autograd.Function):
    """Runs ``fn(weights, *inputs)`` without saving activations; backward re-acquires the unit's
    parameters, recomputes and backpropagates, then tells the provider the unit's grads are final
    (so a sharded engine can launch that unit's reduce-scatter while earlier layers still run)."""

    @staticmethod
    def forward(ctx, provider, unit, fn, anchor, *inputs):
        # ``anchor`` is a dummy scalar that requires grad: parameters are not autograd inputs of this
        # Function, so without it a unit whose inputs carry no grad (the embedding) would cut the graph.
        ctx.provider, ctx.unit, ctx.fn = provider, unit, fn
        ctx.save_for_backward(*[t for t in inputs if isinstance(t, torch.Tensor)])
        ctx.layout = [isinstance(t, torch.Tensor) for t in inputs]
        ctx.consts = [t for t in inputs if not isinstance(t, torch.Tensor)]
        with torch.no_grad():
            w = provider.acquire(unit, backward=False)
            out = fn(w, *inputs)
            provider.release(unit, backward=False)
        return out

    @staticmethod
    def backward(ctx, *grads):
        saved = list(ctx.saved_tensors)
        consts = list(ctx.consts)
        inputs = []
        for is_t in ctx.layout:
            if is_t:
                t = saved.pop(0)
                if t.is_floating_point():
                    t = t.detach().requires_grad_(True)
                inputs.append(t)
            else:
                inputs.append(consts.pop(0))
        provider, unit = ctx.provider, ctx.unit
        w = provider.acquire(unit, backward=True)
        with torch.enable_grad():
            out = ctx.fn(w, *inputs)
        outs = out if isinstance(out, tuple) else (out,)
        pairs = [(o, g) for o, g in zip(outs, grads) if g is not None and o.requires_grad]
        if pairs:
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        provider.grads_done(unit, w)
        provider.release(unit, backward=True)
        in_grads = tuple(t.grad if isinstance(t, torch.Tensor) and t.is_floating_point() else None for t in inputs)
        return (None, None, None, None) + in_grads


_ANCHORS: Dict[str, torch.Tensor] = {}


def _anchor(device: torch.device) -> torch.Tensor:
    key = str(device)
    if key not in _ANCHORS:
        _ANCHORS[key] = torch.zeros((), device=device, requires_grad=True)
    return _ANCHORS[key]


class _UnitBoundary(torch.autograd.Function):
    """Identity on a unit's input whose backward tells the provider that the unit's weight gradients are final.

    For units that keep their activations (plain autograd, no recompute) there is no per-unit backward call to hang
    that signal on.  The node created here is older than every node of the unit, and every op of the unit sits on a
    path from this input to the unit's outputs, so autograd runs it only after all of them — each of which wrote its
    weight gradient inside its own backward — have run.  A sharded engine can then start that unit's reduce-scatter
    while earlier layers are still in backward, exactly as for recomputed units."""

    @staticmethod
    def forward(ctx, x, provider, unit):
        ctx.provider, ctx.unit = provider, unit
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        ctx.provider.unit_backward_done(ctx.unit)
        return g, None, None


def run_unit(provider: ParamProvider, unit: str, fn: Callable, *inputs):
    grad_mode = torch.is_grad_enabled() and provider.trainable
    if grad_mode and provider.checkpoint and unit not in provider.no_recompute:
        dev = next(t.device for t in inputs if isinstance(t, torch.Tensor))
        return _CheckpointedUnit.apply(provider, unit, fn, _anchor(dev), *inputs)
    w = provider.acquire(unit, backward=False)
    hooked = False
    if grad_mode:
        for i, t in enumerate(inputs):
            if isinstance(t, torch.Tensor) and t.is_floating_point() and t.requires_grad:
                inputs = inputs[:i] + (_UnitBoundary.apply(t, provider, unit),) + inputs[i + 1:]
                hooked = True
                break
    out = fn(w, *inputs)
    if not grad_mode:
        provider.release(unit, backward=False)
    elif not hooked:
        provider.note_plain(unit)       # no differentiable input to hang the signal on: swept in finish_backward
    return out


# =======================================================================================
# layer math
# =======================================================================================

def _llama_layer(cfg: ModelConfig):
    Hq, Hk, D = cfg.num_heads, cfg.num_kv_heads, cfg.head_dim

    def fn(w: Weights, x: torch.Tensor, delta: Optional[torch.Tensor], positions: torch.Tensor,
           seqlens: Optional[torch.Tensor], segment_ids: Optional[torch.Tensor] = None):
        # x: residual stream [B, T, H]; delta: the previous block's MLP output not yet added
        B, T, H = x.shape
        x, h = F.add_rmsnorm(x, delta, w["attn_norm"], cfg.norm_eps)
        # RoPE rides in the QKV GEMM's epilogue where the shape allows (head_dim 128, full rotary), else a pass of its own
        qkv = F.linear_rope_qkv(h.reshape(B * T, H), w["wqkv"], positions.reshape(-1), Hq, Hk, D, cfg.rope_theta,
                                cfg.rot_dim or None, max_pos=cfg.max_position_embeddings)
        qkv4 = qkv.view(B, T, Hq + 2 * Hk, D)
        attn = F.attention_qkv(qkv4, Hq, Hk, seqlens=seqlens, causal=True, segment_ids=segment_ids)
        a = F.linear(attn.reshape(B, T, Hq * D), w["wo"])
        x, h = F.add_rmsnorm(x, a, w["mlp_norm"], cfg.norm_eps)
        m = F.mlp_swiglu(h, w["w_gu"], w["w_down"])     # both SwiGLU passes ride in GEMM epilogues (fwd: gate/up, bwd: down dgrad)
        return x, m

    return fn


def _gpt2_layer(cfg: ModelConfig):
    Hh, D = cfg.num_heads, cfg.head_dim

    def fn(w: Weights, x: torch.Tensor, delta: Optional[torch.Tensor], positions: torch.Tensor,
           seqlens: Optional[torch.Tensor], segment_ids: Optional[torch.Tensor] = None):
        B, T, H = x.shape
        if delta is not None:
            x = x + delta
        h = F.layernorm(x, w["ln1_w"], w["ln1_b"], cfg.norm_eps)
        qkv4 = F.linear(h, w["wqkv"], w["bqkv"]).view(B, T, 3 * Hh, D)
        attn = F.attention_qkv(qkv4, Hh, Hh, seqlens=seqlens, causal=True, segment_ids=segment_ids)
        x = x + F.linear(attn.reshape(B, T, H), w["wo"], w["bo"])
        h = F.layernorm(x, w["ln2_w"], w["ln2_b"], cfg.norm_eps)
        m = F.linear(F.gelu_bias(F.linear(h, w["w_fc"]), w["b_fc"]), w["w_proj"], w["b_proj"])
        return x, m

    return fn


class CausalLM:
    """Decoder-only LM.  ``hidden()`` returns the final-normed hidden states ``[B, T, H]``; the heads
    (token log-probs, logits, CE loss) never materialise ``[B, T, V]`` unless asked for logits."""

    def __init__(self, cfg: ModelConfig, provider: ParamProvider) -> None:
        self.cfg, self.p = cfg, provider
        self._layer_fn = _gpt2_layer(cfg) if cfg.arch == "gpt2" else _llama_layer(cfg)

    # ---- pieces ---------------------------------------------------------------------
    def _embed(self, w: Weights, input_ids: torch.Tensor) -> torch.Tensor:
        return F.embedding(input_ids, w["tok_emb"], w.get("pos_emb"))

    def _final_norm(self, w: Weights, x: torch.Tensor, delta: torch.Tensor) -> torch.Tensor:
        if self.cfg.arch == "gpt2":
            return F.layernorm(x + delta, w["lnf_w"], w["lnf_b"], self.cfg.norm_eps)
        return F.add_rmsnorm(x, delta, w["final_norm"], self.cfg.norm_eps)[1]

    def hidden(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
               position_ids: Optional[torch.Tensor] = None, segment_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
        B, T = input_ids.shape
        dev = input_ids.device
        seqlens = None
        if attention_mask is not None:
            seqlens = attention_mask.sum(1).to(torch.int32)      # right padding (reference collators)
        if position_ids is None:
            position_ids = torch.arange(T, device=dev, dtype=torch.int32).expand(B, T)
        x = run_unit(self.p, "embed", self._embed, input_ids)
        delta = None
        for i in range(self.cfg.num_layers):
            x, delta = run_unit(self.p, f"layer.{i}", self._layer_fn, x, delta, position_ids, seqlens, segment_ids)
        return run_unit(self.p, "head", lambda w, a, b: self._final_norm(w, a, b), x, delta)

    def _lm_head_weight(self) -> Tuple[torch.Tensor, str]:
        if self.cfg.tie_word_embeddings:
            return self.p.acquire("embed")["tok_emb"], "embed"
        return self.p.acquire("head")["lm_head"], "head"

    # ---- heads ----------------------------------------------------------------------
    def token_logprobs(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                       labels: Optional[torch.Tensor] = None, **kw) -> torch.Tensor:
        """``[B, T-1]`` fp32 log-prob of token ``t+1`` given the prefix (the shift of
        ``R/src/training/train_dpo.py:33-36``).  ``labels`` (default ``input_ids``) may carry -100."""
        h = self.hidden(input_ids, attention_mask, **kw)
        B, T, H = h.shape
        src = labels if labels is not None else input_ids
        # all T positions go through the fused head (row count stays a multiple of 8 for the TMA path);
        # the last position has no target and is marked ignored
        tgt = torch.cat([src[:, 1:], torch.full_like(src[:, :1], -100)], dim=1)
        w, unit = self._lm_head_weight()
        logp, _ = F.linear_logprob(h.reshape(B * T, H), w, tgt.reshape(-1))
        return logp.view(B, T)[:, :-1]

    def logits(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
               last_only: bool = False, **kw) -> torch.Tensor:
        h = self.hidden(input_ids, attention_mask, **kw)
        if last_only:
            if attention_mask is not None:
                idx = attention_mask.sum(1) - 1
                h = h[torch.arange(h.shape[0], device=h.device), idx].unsqueeze(1)
            else:
                h = h[:, -1:]
        w, _ = self._lm_head_weight()
        return F.linear(h, w)

    def loss(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor], labels: torch.Tensor, **kw) -> torch.Tensor:
        """Shifted cross-entropy, mean over labels != -100 (HF ``ForCausalLM`` loss that the reference
        reads as ``outputs.loss``, ``R/src/training/train_sft.py:145-146``)."""
        lp = self.token_logprobs(input_ids, attention_mask, labels=labels, **kw)
        valid = (labels[:, 1:] >= 0)
        return -(lp * valid).sum() / valid.sum().clamp(min=1)


def build_causal_lm(name_or_cfg, device: torch.device | str = "cpu", dtype: Optional[torch.dtype] = None,
                    trainable: bool = True, checkpoint: bool = True, seed: int = 0, init: bool = True) -> CausalLM:
    """Single-process convenience constructor (random init)."""
    from dla_b200.models.config import resolve_model_config

    cfg = name_or_cfg if isinstance(name_or_cfg, ModelConfig) else resolve_model_config(name_or_cfg)
    device = torch.device(device)
    dtype = dtype or (torch.bfloat16 if device.type == "cuda" else torch.float32)
    prov = LocalParams(cfg, build_units(cfg), device, dtype, trainable=trainable, checkpoint=checkpoint, seed=seed, init=init)
    return CausalLM(cfg, prov)
