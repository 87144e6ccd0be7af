"""Reward model and value head on top of the native backbone.

Parity: ``RewardModel`` = backbone without LM head + ``Sequential(Dropout(p), Linear(hidden, 1))`` with
``last_token`` (index ``mask.sum(1) - 1``, right padding) or masked ``mean`` pooling, returning ``[B]``
scores (``R/src/models/reward_model.py:38-64``).  ``pairwise_loss`` lives in :mod:`dla_b200.ops.functional`.
The value head (per-token scalar on the final hidden states) backs the ``ppo.algo: ppo`` mode that
``BASELINE.json`` asks for and the reference only declares.
"""
from __future__ import annotations

from typing import List, Literal, Optional

import torch

from dla_b200.models.config import ModelConfig
from dla_b200.ops import functional as F
from dla_b200.models.transformer import CausalLM, ParamProvider, ParamSpec, UnitSpec, Weights, build_units, run_unit

Pooling = Literal["last_token", "mean"]


def reward_units(cfg: ModelConfig) -> List[UnitSpec]:
    units = build_units(cfg, with_lm_head=False)
    units.append(UnitSpec("reward_head", [ParamSpec("score_w", (1, cfg.hidden_size)), ParamSpec("score_b", (1,), "zeros", False)]))
    return units


def value_head_unit(cfg: ModelConfig) -> UnitSpec:
    return UnitSpec("value_head", [ParamSpec("value_w", (1, cfg.hidden_size), "zeros"), ParamSpec("value_b", (1,), "zeros", False)])


def pool_last_token(hidden: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
    """``hidden[b, mask[b].sum() - 1]`` — assumes right padding like ``R/src/models/reward_model.py:57-59``."""
    idx = attention_mask.sum(dim=1) - 1
    return hidden[torch.arange(hidden.shape[0], device=hidden.device), idx.clamp(min=0)]


def pool_mean(hidden: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
    m = attention_mask.unsqueeze(-1).to(hidden.dtype)
    return (hidden * m).sum(1) / attention_mask.sum(1, keepdim=True).clamp(min=1).to(hidden.dtype)


class RewardModel:
    def __init__(self, cfg: ModelConfig, provider: ParamProvider, pooling: Pooling = "last_token", dropout: float = 0.1) -> None:
        self.cfg, self.p = cfg, provider
        self.backbone = CausalLM(cfg, provider)
        self.pooling, self.dropout = pooling, dropout
        self.training = True

    def train(self, mode: bool = True) -> "RewardModel":
        self.training = mode
        return self

    def eval(self) -> "RewardModel":
        return self.train(False)

    def _head(self, w: Weights, hidden: torch.Tensor, attention_mask: torch.Tensor, keep: Optional[torch.Tensor]) -> torch.Tensor:
        return F.scalar_head(hidden, attention_mask, w["score_w"], w["score_b"], keep, self.pooling)

    def __call__(self, input_ids: torch.Tensor, attention_mask: torch.Tensor) -> torch.Tensor:
        hidden = self.backbone.hidden(input_ids, attention_mask)
        keep = None
        if self.training and self.dropout > 0 and torch.is_grad_enabled():
            # the mask is drawn here (not inside the checkpointed unit) so recompute sees the same one
            keep = (torch.rand(hidden.shape[0], hidden.shape[-1], device=hidden.device) >= self.dropout).float() / (1 - self.dropout)
        return run_unit(self.p, "reward_head", self._head, hidden, attention_mask, keep)

    forward = __call__


class ValueHead:
    """Per-token value ``[B, T]`` from final hidden states; parameters live in unit ``value_head``."""

    def __init__(self, provider: ParamProvider) -> None:
        self.p = provider

    @staticmethod
    def _fn(w: Weights, hidden: torch.Tensor) -> torch.Tensor:
        return F.scalar_head(hidden, None, w["value_w"], w["value_b"], None, "token")

    def __call__(self, hidden: torch.Tensor) -> torch.Tensor:
        return run_unit(self.p, "value_head", self._fn, hidden)
