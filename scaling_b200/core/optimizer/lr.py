"""Warm-up + decay learning-rate schedule (reference: core/optimizer/learning_rate_scheduler/*.py)."""
from __future__ import annotations

import math

from .configs import LearningRateDecayStyle, LearningRateSchedulerConfig


class LearningRateScheduler:
    def __init__(self, config: LearningRateSchedulerConfig):
        self.config = config

    def get_lr(self, step_index: int) -> float:
        c = self.config
        if c.learning_rate_warmup_steps > 0 and step_index <= c.learning_rate_warmup_steps:
            return c.learning_rate * step_index / c.learning_rate_warmup_steps
        if c.learning_rate_decay_style == LearningRateDecayStyle.CONSTANT:
            return c.learning_rate
        if step_index > c.learning_rate_decay_iters:
            return c.learning_rate_minimum
        progress = (step_index - c.learning_rate_warmup_steps) / (c.learning_rate_decay_iters - c.learning_rate_warmup_steps)
        assert 0.0 <= progress <= 1.0
        if c.learning_rate_decay_style == LearningRateDecayStyle.LINEAR:
            coeff = 1.0 - progress
        else:
            coeff = 0.5 * (math.cos(math.pi * progress) + 1.0)
        return c.learning_rate_minimum + coeff * (c.learning_rate - c.learning_rate_minimum)
