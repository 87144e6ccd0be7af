"""Model layer (parity: ``R/src/models/``)."""
from dla_b200.models.config import PRESETS, ModelConfig, resolve_model_config  # noqa: F401
from dla_b200.models.reward import RewardModel, ValueHead, reward_units, value_head_unit  # noqa: F401
from dla_b200.models.transformer import (  # noqa: F401
    CausalLM, LocalParams, ParamProvider, ParamSpec, UnitSpec, build_causal_lm, build_units,
)
