"""KV-cache generation engine (prefill + per-token decode) with on-device sampling.

Replaces ``model.generate(...)`` of transformers as used for PPO rollouts
(``R/src/training/train_rlhf.py:123-124``), teacher sampling (``R/src/training/generate_teacher_data.py:73-79``)
and eval (``R/src/eval/eval_alignment.py:71-77``).  Differences by design: prompts may be
right-padded — per-row lengths are tracked explicitly instead of re-deriving a mask from ``pad == eos``;
rows are decoded as a batch; sampling (temperature, nucleus, multinomial, EOS tracking) never leaves
the device.  Output is HF-shaped: ``sequences [B, T_prompt_max + new]`` = each row's prompt followed
directly by its continuation, right-padded with ``pad_token_id``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch

from dla_b200.models.config import ModelConfig
from dla_b200.models.transformer import CausalLM, Weights
from dla_b200.ops import functional as F
from dla_b200.ops import native as N
from dla_b200.ops import reference as R


@dataclass
class GenerationOutput:
    sequences: torch.Tensor          # [B, L] prompt + response, right padded
    attention_mask: torch.Tensor     # [B, L]
    prompt_lens: torch.Tensor        # [B]
    response_mask: torch.Tensor      # [B, L] 1 on generated tokens (EOS included)
    lengths: torch.Tensor            # [B] total valid tokens per row


class KVCache:
    def __init__(self, cfg: ModelConfig, batch: int, max_len: int, device: torch.device, dtype: torch.dtype) -> None:
        shape = (cfg.num_layers, batch, max_len, cfg.num_kv_heads, cfg.head_dim)
        self.k = torch.zeros(shape, dtype=dtype, device=device)
        self.v = torch.zeros(shape, dtype=dtype, device=device)
        self.max_len = max_len


class Generator:
    def __init__(self, model: CausalLM) -> None:
        self.model, self.cfg, self.p = model, model.cfg, model.p

    # ---- per-layer math with cache ------------------------------------------------------
    def _qkv(self, w: Weights, x: torch.Tensor, positions: torch.Tensor):
        cfg = self.cfg
        B, T, H = x.shape
        if cfg.arch == "gpt2":
            h = F.layernorm(x, w["ln1_w"], w["ln1_b"], cfg.norm_eps)
            qkv = F.linear(h, w["wqkv"], w["bqkv"]).reshape(B * T, -1)
        else:
            h = F.rmsnorm(x, w["attn_norm"], cfg.norm_eps)
            qkv = F.linear(h.reshape(B * T, H), w["wqkv"])
            qkv = F.rope_qkv(qkv, positions.reshape(-1), cfg.num_heads, cfg.num_kv_heads, cfg.head_dim, cfg.rope_theta,
                             cfg.rot_dim or None, max_pos=cfg.max_position_embeddings)
        return qkv.view(B, T, cfg.num_heads + 2 * cfg.num_kv_heads, cfg.head_dim)

    def _post_attn(self, w: Weights, x: torch.Tensor, attn: torch.Tensor) -> torch.Tensor:
        cfg = self.cfg
        B, T, H = x.shape
        if cfg.arch == "gpt2":
            x = x + F.linear(attn.reshape(B, T, H), w["wo"], w["bo"])
            h = F.layernorm(x, w["ln2_w"], w["ln2_b"], cfg.norm_eps)
            return x + F.linear(F.gelu_bias(F.linear(h, w["w_fc"]), w["b_fc"]), w["w_proj"], w["b_proj"])
        a = F.linear(attn.reshape(B, T, cfg.q_dim), w["wo"])
        x, h = F.add_rmsnorm(x, a, w["mlp_norm"], cfg.norm_eps)
        return x + F.linear(F.swiglu(F.linear(h, w["w_gu"])), w["w_down"])

    def _final(self, x: torch.Tensor) -> torch.Tensor:
        w = self.p.acquire("head")
        if self.cfg.arch == "gpt2":
            h = F.layernorm(x, w["lnf_w"], w["lnf_b"], self.cfg.norm_eps)
        else:
            h = F.rmsnorm(x, w["final_norm"], self.cfg.norm_eps)
        lw, _ = self.model._lm_head_weight()
        out = F.linear(h, lw)
        self.p.release("head")
        return out

    def _decode_attention(self, q: torch.Tensor, cache: KVCache, layer: int, lens: torch.Tensor) -> torch.Tensor:
        """q [B, Hq, D] against the first lens[b] cached positions."""
        cfg = self.cfg
        B = q.shape[0]
        if q.is_cuda and cfg.head_dim in (64, 128):
            o = torch.empty_like(q)
            N.call("dla_attn_decode", N.ptr(q), N.ptr(cache.k[layer]), N.ptr(cache.v[layer]), N.ptr(o), N.ptr(lens), B,
                   cache.max_len, cfg.num_heads, cfg.num_kv_heads, cfg.head_dim, cfg.head_dim ** -0.5, N.stream_ptr(q.device))
            return o
        return R.attention(q[:, None], cache.k[layer], cache.v[layer], causal=False, seqlens=lens)[:, 0]

    # ---- public ------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, max_new_tokens: int = 256,
                 temperature: float = 1.0, top_p: float = 1.0, do_sample: bool = False, eos_token_id: Optional[int] = None,
                 pad_token_id: Optional[int] = None, seed: Optional[int] = None, **_: object) -> GenerationOutput:
        cfg = self.cfg
        dev = input_ids.device
        B, Tp = input_ids.shape
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        eos = cfg.eos_token_id if eos_token_id is None else eos_token_id
        pad = eos if pad_token_id is None else pad_token_id
        plen = attention_mask.sum(1).to(torch.int32)
        max_len = Tp + max_new_tokens
        if cfg.arch == "gpt2":
            max_len = min(max_len, cfg.max_position_embeddings)
        dtype = self.p.acquire("embed")["tok_emb"].dtype
        cache = KVCache(cfg, B, max_len, dev, dtype)
        Hq, Hk = cfg.num_heads, cfg.num_kv_heads
        gen = torch.Generator(device=dev)
        seed = seed if seed is not None else int(torch.randint(0, 2**31 - 1, (1,)).item())
        gen.manual_seed(seed)

        # ---- prefill: full causal attention over the (right-padded) prompts, filling the cache
        pos = torch.arange(Tp, device=dev, dtype=torch.int32).expand(B, Tp)
        x = F.embedding(input_ids, self.p.acquire("embed")["tok_emb"], self.p.acquire("embed").get("pos_emb"))
        self.p.release("embed")
        for i in range(cfg.num_layers):
            w = self.p.acquire(f"layer.{i}")
            qkv = self._qkv(w, x, pos)
            q, k, v = qkv[:, :, :Hq], qkv[:, :, Hq:Hq + Hk], qkv[:, :, Hq + Hk:]
            cache.k[i, :, :Tp].copy_(k)
            cache.v[i, :, :Tp].copy_(v)
            attn = F.attention(q, k, v, seqlens=plen, causal=True)
            x = self._post_attn(w, x, attn)
            self.p.release(f"layer.{i}")
        last = x[torch.arange(B, device=dev), (plen - 1).long().clamp(min=0)].unsqueeze(1)      # [B,1,H]
        logits = self._final(last)[:, 0]

        seqs = torch.full((B, max_len), pad, dtype=torch.long, device=dev)
        seqs[:, :Tp] = torch.where(attention_mask.bool(), input_ids, torch.full_like(input_ids, pad))
        lens = plen.clone()
        finished = torch.zeros(B, dtype=torch.bool, device=dev)
        rows = torch.arange(B, device=dev)
        steps = min(max_new_tokens, max_len - int(plen.min().item()))
        for step in range(steps):
            nxt = sample(logits, temperature if do_sample else 0.0, top_p, gen, seed=seed, step=step)
            nxt = torch.where(finished, torch.full_like(nxt, pad), nxt)
            can = (~finished) & (lens < max_len)
            idx = lens.long().clamp(max=max_len - 1)
            seqs[rows, idx] = torch.where(can, nxt, seqs[rows, idx])
            newly = can & (nxt == eos)
            lens = lens + can.to(torch.int32)
            finished = finished | newly | (lens >= max_len)
            if step == steps - 1 or bool(finished.all()):
                break
            # ---- decode one token per row at position lens-1
            tok_pos = (lens - 1).clamp(min=0)
            x = F.embedding(nxt[:, None], self.p.acquire("embed")["tok_emb"], None)
            if cfg.arch == "gpt2":
                x = x + self.p.acquire("embed")["pos_emb"][tok_pos.long()][:, None]
            self.p.release("embed")
            for i in range(cfg.num_layers):
                w = self.p.acquire(f"layer.{i}")
                qkv = self._qkv(w, x, tok_pos[:, None])
                cache.k[i, rows, tok_pos.long()] = qkv[:, 0, Hq:Hq + Hk]
                cache.v[i, rows, tok_pos.long()] = qkv[:, 0, Hq + Hk:]
                attn = self._decode_attention(qkv[:, 0, :Hq].contiguous(), cache, i, lens)
                x = self._post_attn(w, x, attn[:, None])
                self.p.release(f"layer.{i}")
            logits = self._final(x)[:, 0]
        L = int(lens.max().item())
        seqs = seqs[:, :L]
        ar = torch.arange(L, device=dev)[None]
        amask = (ar < lens[:, None]).long()
        rmask = ((ar >= plen[:, None]) & (ar < lens[:, None])).long()
        return GenerationOutput(seqs, amask, plen.long(), rmask, lens.long())


def sample(logits: torch.Tensor, temperature: float, top_p: float, generator: Optional[torch.Generator] = None,
           seed: int = 0, step: int = 0) -> torch.Tensor:
    """Next-token choice on the device: greedy when ``temperature <= 0`` else nucleus sampling.  CUDA runs
    the fused temperature + top-p + multinomial kernel (``csrc/sampler.cu``, counter-based RNG keyed by
    ``seed``/``step``); CPU runs the sort-based PyTorch restatement."""
    if logits.is_cuda:
        lg = logits if logits.is_contiguous() else logits.contiguous()
        if lg.dtype not in (torch.float32, torch.bfloat16):
            lg = lg.float()
        out = torch.empty(lg.shape[0], dtype=torch.int64, device=lg.device)
        N.call("dla_sample_top_p", N.ptr(lg), int(lg.dtype == torch.float32), lg.shape[0], lg.shape[1], lg.stride(0),
               float(temperature), float(min(max(top_p, 1e-6), 1.0)), int(seed) & 0xFFFFFFFFFFFFFFFF, int(step), None,
               N.ptr(out), N.stream_ptr(lg.device))
        return out
    if temperature <= 0:
        return logits.argmax(-1)
    return R.top_p_sample(logits, temperature, top_p, generator)


def decode_responses(tokenizer, out: GenerationOutput) -> List[str]:
    texts = []
    for b in range(out.sequences.shape[0]):
        s, e = int(out.prompt_lens[b]), int(out.lengths[b])
        texts.append(tokenizer.decode(out.sequences[b, s:e], skip_special_tokens=True))
    return texts
