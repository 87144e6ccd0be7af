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
from typing import Sequence, List, Optional

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


@dataclass
class _DecodeState:
    logits: torch.Tensor      # [B, V] next-token logits of every row
    seqs: torch.Tensor        # [B, max_len]
    lens: torch.Tensor        # [B] int32 valid tokens per row
    finished: torch.Tensor    # [B] bool
    nxt: torch.Tensor         # [B] token chosen by the last pick
    counter: torch.Tensor     # [1] int64 RNG counter, advanced on the device


@dataclass
class _DecodeSession:
    cache: KVCache
    st: _DecodeState
    rows: torch.Tensor
    fused: bool
    temp: float
    top_p: float
    eos: int
    pad: int
    max_len: int
    graph: Optional["torch.cuda.CUDAGraph"] = None


_SEED_BASE = 0x5EED5EED            # kernel-side constant; the per-call seed enters through the device counter


class Generator:
    def __init__(self, model: CausalLM, use_cuda_graph: bool = True, finish_check_every: int = 8) -> None:
        self.model, self.cfg, self.p = model, model.cfg, model.p
        self.use_cuda_graph = use_cuda_graph
        self.finish_check_every = max(1, int(finish_check_every))    # host reads `finished` only this often
        self._sessions: dict = {}

    def _graphable(self, dev: torch.device) -> bool:
        """The decode step can be captured when acquiring a unit launches nothing (parameters resident)."""
        return bool(self.use_cuda_graph and dev.type == "cuda" and getattr(self.p, "resident", True))

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
            x = F.linear_residual(attn.reshape(B, T, H), w["wo"], x, w["bo"])
            h = F.layernorm(x, w["ln2_w"], w["ln2_b"], cfg.norm_eps)
            return F.linear_residual(F.gelu_bias(F.linear(h, w["w_fc"]), w["b_fc"]), w["w_proj"], x, w["b_proj"])
        a = F.linear(attn.reshape(B, T, cfg.q_dim), w["wo"])
        x, h = F.add_rmsnorm(x, a, w["mlp_norm"], cfg.norm_eps)
        return F.linear_residual(F.swiglu(F.linear(h, w["w_gu"])), w["w_down"], x)

    def _fused_decode_ok(self, dev: torch.device, max_len: int) -> bool:
        cfg = self.cfg
        g = cfg.num_heads // cfg.num_kv_heads
        stripes = 256 // (cfg.head_dim // 8) if cfg.head_dim in (64, 128) else 0
        smem = 4 * (g * cfg.head_dim + g * max_len + stripes * g * cfg.head_dim)
        return dev.type == "cuda" and cfg.head_dim in (64, 128) and g in (1, 2, 4, 8) and smem <= 220 * 1024

    def _decode_layer_attention(self, w: Weights, x: torch.Tensor, cache: KVCache, layer: int, lens: torch.Tensor) -> torch.Tensor:
        """Norm + QKV projection of one token per row, then RoPE + cache append + attention in one kernel
        (``csrc/decode.cu``); ``lens`` already counts the new token.  Returns ``[B, Hq, D]``."""
        cfg = self.cfg
        B, _, H = x.shape
        if cfg.arch == "gpt2":
            h = F.layernorm(x, w["ln1_w"], w["ln1_b"], cfg.norm_eps)
            qkv = F.linear(h, w["wqkv"], w["bqkv"]).reshape(B, -1)
            cos = sin = None
            rot = 0
        else:
            h = F.rmsnorm(x, w["attn_norm"], cfg.norm_eps)
            qkv = F.linear(h.reshape(B, H), w["wqkv"])
            rot = cfg.rot_dim or cfg.head_dim
            cos, sin = F.rope_tables(cfg.max_position_embeddings, rot, cfg.rope_theta, x.device)
        o = torch.empty(B, cfg.num_heads, cfg.head_dim, dtype=qkv.dtype, device=x.device)
        N.call("dla_decode_attn_fused", N.ptr(qkv), qkv.stride(0), N.ptr(cache.k[layer]), N.ptr(cache.v[layer]), N.ptr(o), N.ptr(lens),
               N.ptr(cos) if cos is not None else None, N.ptr(sin) if sin is not None else None, B, cache.max_len, cfg.num_heads,
               cfg.num_kv_heads, cfg.head_dim, rot, cfg.head_dim ** -0.5, N.stream_ptr(x.device))
        return o

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

    # ---- decode session -----------------------------------------------------------------
    def _session(self, key: tuple, B: int, max_len: int, dev: torch.device, dtype: torch.dtype, temp: float, top_p: float,
                 eos: int, pad: int) -> "_DecodeSession":
        """Fixed device buffers for one (batch, length, sampling) signature.  Sessions are kept (most recent two) so
        that repeated rollouts of the same shape — every PPO step — replay the graph captured the first time.
        A session holds no reference back to the generator: dropping the generator frees its graphs at once
        (a CUDA graph must never be destroyed by the cyclic GC in the middle of another capture)."""
        ses = self._sessions.get(key)
        if ses is not None:
            return ses
        cfg = self.cfg
        st = _DecodeState(logits=torch.zeros(B, cfg.vocab_size, dtype=dtype, device=dev),
                          seqs=torch.full((B, max_len), pad, dtype=torch.long, device=dev),
                          lens=torch.zeros(B, dtype=torch.int32, device=dev), finished=torch.zeros(B, dtype=torch.bool, device=dev),
                          nxt=torch.zeros(B, dtype=torch.long, device=dev), counter=torch.zeros(1, dtype=torch.long, device=dev))
        ses = _DecodeSession(cache=KVCache(cfg, B, max_len, dev, dtype), st=st, rows=torch.arange(B, device=dev),
                             fused=self._fused_decode_ok(dev, max_len), temp=temp, top_p=top_p, eos=eos, pad=pad, max_len=max_len)
        while len(self._sessions) >= 2:
            self._sessions.pop(next(iter(self._sessions)))
        self._sessions[key] = ses
        return ses

    def _pick(self, ses: "_DecodeSession") -> None:
        """Sample one token per row and update sequences / lengths / finished flags (no host sync)."""
        st, max_len = ses.st, ses.max_len
        nxt = sample(st.logits, ses.temp, ses.top_p, None, seed=_SEED_BASE, step=0, step_dev=st.counter)
        nxt = torch.where(st.finished, torch.full_like(nxt, ses.pad), nxt)
        can = (~st.finished) & (st.lens < max_len)
        idx = st.lens.long().clamp(max=max_len - 1)
        st.seqs[ses.rows, idx] = torch.where(can, nxt, st.seqs[ses.rows, idx])
        st.lens.add_(can.to(torch.int32))
        st.finished.logical_or_((can & (nxt == ses.eos)) | (st.lens >= max_len))
        st.nxt.copy_(nxt)
        st.counter.add_(1)

    def _advance(self, ses: "_DecodeSession") -> None:
        """One decode forward for the tokens chosen by ``_pick`` (position lens-1), refreshing ``st.logits``."""
        cfg, st, cache, rows = self.cfg, ses.st, ses.cache, ses.rows
        Hq, Hk = cfg.num_heads, cfg.num_kv_heads
        tok_pos = (st.lens - 1).clamp(min=0)
        x = F.embedding(st.nxt[:, None], self.p.acquire("embed")["tok_emb"], None)
        if cfg.arch == "gpt2":
            x = x + self.p.acquire("embed")["pos_emb"][tok_pos.long()][:, None]
        self.p.release("embed")
        pos64 = None if ses.fused else tok_pos.long()
        for i in range(cfg.num_layers):
            w = self.p.acquire(f"layer.{i}")
            if ses.fused:
                attn = self._decode_layer_attention(w, x, cache, i, st.lens)
            else:
                qkv = self._qkv(w, x, tok_pos[:, None])
                cache.k[i, rows, pos64] = qkv[:, 0, Hq:Hq + Hk]
                cache.v[i, rows, pos64] = qkv[:, 0, Hq + Hk:]
                attn = self._decode_attention(qkv[:, 0, :Hq].contiguous(), cache, i, st.lens)
            x = self._post_attn(w, x, attn[:, None])
            self.p.release(f"layer.{i}")
        st.logits.copy_(self._final(x)[:, 0])

    def _capture(self, ses: "_DecodeSession") -> None:
        """Record ``_advance`` + ``_pick`` as one CUDA graph.  Garbage collection is run first and held off during
        the capture: freeing an unrelated dead CUDA graph while a stream is capturing invalidates the capture."""
        import gc

        gc.collect()
        was_enabled = gc.isenabled()
        gc.disable()
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._advance(ses)
                self._pick(ses)
            ses.graph = graph
        finally:
            if was_enabled:
                gc.enable()

    def reset(self) -> None:
        """Drop cached decode sessions (KV buffers and captured graphs), e.g. after parameters were re-allocated."""
        self._sessions.clear()

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
        if dev.type == "cuda":
            # KV buffers (and the captured decode graph keyed by them) come in 64-position buckets, so rollouts whose
            # prompt lengths differ a little share one session; rows still stop after max_new_tokens
            max_len = (max_len + 63) // 64 * 64
        if cfg.arch == "gpt2":
            max_len = min(max_len, cfg.max_position_embeddings)
        elif Tp + max_new_tokens > cfg.max_position_embeddings:
            raise ValueError(f"prompt ({Tp}) + max_new_tokens ({max_new_tokens}) exceeds the model's {cfg.max_position_embeddings} "
                             f"positions (the RoPE tables end there)")
        else:
            max_len = min(max_len, cfg.max_position_embeddings)
        dtype = self.p.acquire("embed")["tok_emb"].dtype
        temp = float(temperature) if do_sample else 0.0
        seed = seed if seed is not None else int(torch.randint(0, 2**31 - 1, (1,)).item())
        if dev.type != "cuda":
            return self._generate_host(input_ids, attention_mask, plen, max_len, max_new_tokens, temp, top_p, eos, pad, seed)
        ses = self._session((B, max_len, str(dev), dtype, temp, float(top_p), int(eos), int(pad)), B, max_len, dev, dtype, temp,
                            float(top_p), int(eos), int(pad))
        cache, st = ses.cache, ses.st
        Hq, Hk = cfg.num_heads, cfg.num_kv_heads

        # ---- prefill: full causal attention over the (right-padded) prompts, filling the cache
        logits = self._prefill(input_ids, plen, cache)
        st.logits.copy_(logits)
        st.seqs.fill_(pad)
        st.seqs[:, :Tp] = torch.where(attention_mask.bool(), input_ids, torch.full_like(input_ids, pad))
        st.lens.copy_(plen)
        st.finished.zero_()
        st.counter.fill_((int(seed) & 0x7FFFFFFF) << 24)     # the seed selects the stream; one count per decoded token
        steps = min(max_new_tokens, max_len - int(plen.min().item()))

        done = 0
        with F.weights_static():      # nothing in this stream writes the weights until generation returns
            if steps > 0:
                self._pick(ses)
                done = 1
            if ses.graph is None and self._graphable(dev) and steps > 4:
                # one eager token first (work areas and device binding exist before capture), then capture
                # "forward + pick" once: ~12 launches per layer collapse into one graph launch per token.
                # Capture only records; the state is untouched until the first replay.
                self._advance(ses)
                self._pick(ses)
                done = 2
                self._capture(ses)
            use_graph = ses.graph is not None and self.use_cuda_graph
            while done < steps:
                if use_graph:
                    ses.graph.replay()
                else:
                    self._advance(ses)
                    self._pick(ses)
                done += 1
                if done % self.finish_check_every == 0 and done < steps and bool(st.finished.all()):
                    break
        lens = st.lens.clone()
        L = int(lens.max().item())
        seqs = st.seqs[:, :L].clone()
        ar = torch.arange(L, device=dev)[None]
        amask = (ar < lens[:, None]).long()
        rmask = ((ar >= plen[:, None]) & (ar < lens[:, None])).long()
        return GenerationOutput(seqs, amask, plen.long(), rmask, lens.long())

    def _prefill(self, input_ids: torch.Tensor, plen: torch.Tensor, cache: KVCache) -> torch.Tensor:
        """Full causal forward over the right-padded prompts; fills ``cache[:, :, :Tp]`` and returns the logits
        of every row's last prompt token ``[B, V]``."""
        cfg = self.cfg
        dev = input_ids.device
        B, Tp = input_ids.shape
        Hq, Hk = cfg.num_heads, cfg.num_kv_heads
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
        return self._final(last)[:, 0]

    def _generate_host(self, input_ids, attention_mask, plen, max_len, max_new_tokens, temp, top_p, eos, pad, seed) -> GenerationOutput:
        """CPU restatement of the same loop (tests, tiny models): PyTorch ops, torch.Generator sampling."""
        cfg = self.cfg
        dev = input_ids.device
        B, Tp = input_ids.shape
        Hq, Hk = cfg.num_heads, cfg.num_kv_heads
        dtype = self.p.acquire("embed")["tok_emb"].dtype
        cache = KVCache(cfg, B, max_len, dev, dtype)
        gen = torch.Generator(device=dev)
        gen.manual_seed(seed)
        logits = self._prefill(input_ids, plen, cache)
        seqs = torch.full((B, max_len), pad, dtype=torch.long, device=dev)
        seqs[:, :Tp] = torch.where(attention_mask.bool(), input_ids, torch.full_like(input_ids, pad))
        lens = plen.clone()
        finished = torch.zeros(B, dtype=torch.bool, device=dev)
        rows = torch.arange(B, device=dev)
        steps = min(max_new_tokens, max_len - int(plen.min().item()))
        for step in range(steps):
            nxt = sample(logits, temp, top_p, gen, seed=seed, step=step)
            nxt = torch.where(finished, torch.full_like(nxt, pad), nxt)
            can = (~finished) & (lens < max_len)
            idx = lens.long().clamp(max=max_len - 1)
            seqs[rows, idx] = torch.where(can, nxt, seqs[rows, idx])
            newly = can & (nxt == eos)
            lens = lens + can.to(torch.int32)
            finished = finished | newly | (lens >= max_len)
            if step == steps - 1 or bool(finished.all()):
                break
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


class GraphedForward:
    """``model.logits(ids, mask)`` for one fixed ``[B, S]`` shape, captured once and replayed as a single CUDA graph.

    A prefill of a few hundred tokens is launch-bound (≈330 launches of 5-30 us each for Llama-3-8B): serving and the
    latency grid (``R/src/eval/eval_latency.py:42-53`` runs the same shape ``measure_steps`` times) replay the graph
    instead.  Inputs are copied into fixed buffers; the returned logits tensor is the graph's output buffer (valid
    until the next call).  Requires resident parameters (acquiring a unit must not launch collectives)."""

    def __init__(self, model: CausalLM, batch: int, seq: int, last_only: bool = False) -> None:
        import gc

        dev = model.p.device if hasattr(model.p, "device") else torch.device("cuda")
        self.ids = torch.zeros(batch, seq, dtype=torch.long, device=dev)
        self.mask = torch.ones(batch, seq, dtype=torch.long, device=dev)
        with torch.no_grad(), F.weights_static():
            for _ in range(2):                       # work areas, function attributes and device binding exist before capture
                model.logits(self.ids, self.mask, last_only=last_only)
            torch.cuda.synchronize(dev)
            gc.collect()
            was = gc.isenabled()
            gc.disable()                             # a dead CUDA graph freed mid-capture would invalidate this capture
            try:
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):
                    self.out = model.logits(self.ids, self.mask, last_only=last_only)
            finally:
                if was:
                    gc.enable()

    @torch.no_grad()
    def __call__(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        self.ids.copy_(input_ids)
        if attention_mask is not None:
            self.mask.copy_(attention_mask)
        else:
            self.mask.fill_(1)
        self.graph.replay()
        return self.out


def sample(logits: torch.Tensor, temperature: float, top_p: float, generator: Optional[torch.Generator] = None,
           seed: int = 0, step: int = 0, step_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Next-token choice on the device: greedy when ``temperature <= 0`` else nucleus sampling.  CUDA runs
    the fused temperature + top-p + multinomial kernel (``csrc/sampler.cu``, counter-based RNG keyed by
    ``seed``/``step`` plus the device counter ``step_dev`` when given); CPU runs the sort-based PyTorch restatement."""
    if logits.is_cuda:
        lg = logits if logits.is_contiguous() else logits.contiguous()
        if lg.dtype not in (torch.float32, torch.bfloat16):
            lg = lg.float()
        out = torch.empty(lg.shape[0], dtype=torch.int64, device=lg.device)
        N.call("dla_sample_top_p", N.ptr(lg), int(lg.dtype == torch.float32), lg.shape[0], lg.shape[1], lg.stride(0),
               float(temperature), float(min(max(top_p, 1e-6), 1.0)), int(seed) & 0xFFFFFFFFFFFFFFFF, int(step),
               N.ptr(step_dev) if step_dev is not None else None, N.ptr(out), N.stream_ptr(lg.device))
        return out
    if temperature <= 0:
        return logits.argmax(-1)
    return R.top_p_sample(logits, temperature, top_p, generator)


def splice_separator(out: GenerationOutput, sep_ids: Sequence[int], pad_id: int = 0):
    """``(ids, mask)`` of ``prompt + sep + response`` per row, built on the device from a rollout's right-padded
    ``prompt + response`` rows.  The reward model is trained on ``f"{prompt}\n\n{completion}"`` and the reference
    scores rollouts through that text form (``R/src/training/train_rlhf.py:139-147``); when policy and reward model
    share a tokenizer the same token sequence is produced here without the decode / re-tokenise round trip."""
    seqs, plen, lens = out.sequences, out.prompt_lens.long(), out.lengths.long()
    k = len(sep_ids)
    if k == 0:
        return seqs, out.attention_mask
    B, L = seqs.shape
    dev = seqs.device
    sep = torch.tensor(list(sep_ids), dtype=seqs.dtype, device=dev)
    j = torch.arange(L + k, device=dev)[None, :]
    p = plen[:, None]
    src = torch.where(j < p, j, j - k).clamp_(0, L - 1)
    ids = seqs.gather(1, src.expand(B, L + k))
    in_sep = (j >= p) & (j < p + k)
    ids = torch.where(in_sep, sep[(j - p).clamp(0, k - 1)], ids)
    mask = (j < (lens[:, None] + k))
    ids = torch.where(mask, ids, torch.full_like(ids, pad_id))
    return ids, mask.to(out.attention_mask.dtype)


def decode_responses(tokenizer, out: GenerationOutput) -> List[str]:
    texts = []
    for b in range(out.sequences.shape[0]):
        s, e = int(out.prompt_lens[b]), int(out.lengths[b])
        texts.append(tokenizer.decode(out.sequences[b, s:e], skip_special_tokens=True))
    return texts
