"""Optimizer-side configs (reference: core/optimizer/optimizer_config.py, parameter_group_config.py,
loss_scaler_config.py, learning_rate_scheduler/learning_rate_scheduler_config.py)."""
from __future__ import annotations

from enum import Enum
from typing import Optional

from pydantic import Field

from ..config import BaseConfig


class LearningRateDecayStyle(Enum):
    CONSTANT = "constant"
    LINEAR = "linear"
    COSINE = "cosine"


class LearningRateSchedulerConfig(BaseConfig):
    learning_rate: float = Field(0.0, description="peak learning rate (reached at the end of warm-up)")
    learning_rate_minimum: float = Field(0.0, description="floor of the schedule and value after the decay")
    learning_rate_decay_style: LearningRateDecayStyle = Field(LearningRateDecayStyle.COSINE, description="shape of the decay after warm-up")
    learning_rate_decay_iters: int = Field(0, description="iterations over which the schedule runs, warm-up included")
    learning_rate_warmup_steps: int = Field(0, description="linear warm-up steps")


class LossScalerConfig(BaseConfig):
    """Dynamic loss scaling for fp16: start high, halve (after `hysteresis` overflows) whenever gradients overflow and
    skip that step, double again after `window` clean steps."""

    enable: bool = Field(False, description="enable dynamic loss scaling")
    initial_scale: float = Field(2.0**32, description="initial loss scale")
    window: int = Field(1000, description="clean steps before the scale is raised")
    hysteresis: float = Field(2, description="overflows tolerated before the scale is lowered")
    consecutive_hysteresis: bool = Field(False, description="refill the hysteresis budget on every clean step")
    min_scale: float = Field(1.0, description="lower bound of the scale")
    factor: float = Field(2.0, description="multiplicative step of the scale")


class OptimizerConfig(BaseConfig):
    method: str = Field("adamw", description="optimization method")
    beta1: float = Field(0.9, description="decay of the first-moment running average")
    beta2: float = Field(0.95, description="decay of the second-moment running average")
    eps: float = Field(1e-8, description="term added to the denominator")
    gradient_clipping: float = Field(1.0, description="clip the global l2 gradient norm to this value; 0 disables", ge=0.0)
    allreduce_bucket_size: int = Field(500000000, description="elements per collective bucket", gt=0)
    loss_scaler: LossScalerConfig = Field(LossScalerConfig(), description="loss scaler configuration")
    zero: bool = Field(False, description="ZeRO stage 1: shard optimizer state over the data-parallel group")
    zero_save_static: bool = Field(False, description="checkpoint the per-rank optimizer shards as they are (fast, layout specific)")
    debug_log: bool = Field(False, description="log per-parameter norms each step")


class OptimizerParamGroupConfig(BaseConfig):
    name: Optional[str] = Field(None, description="name used in logged metrics")
    learning_rate_scheduler: LearningRateSchedulerConfig = Field(LearningRateSchedulerConfig(), description="schedule of this group")
    weight_decay: float = Field(1e-2, description="decoupled weight decay of this group")
