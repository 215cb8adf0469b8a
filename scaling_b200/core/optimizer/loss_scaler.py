"""Dynamic loss scaler (reference: core/optimizer/loss_scaler.py:1-132).

The overflow flag is produced on the device by the fused Σg² kernel (one flag for the whole flat gradient shard) and
MAX-reduced together with the norm; the host reads it once per step — the reference syncs the host once per parameter.
"""
from __future__ import annotations

from typing import Any, Dict, NamedTuple, Optional

import torch

from .configs import LossScalerConfig


class LossScalerOutput(NamedTuple):
    overflow: Optional[bool]
    no_overflow_steps: Optional[int]
    current_loss_scale: Optional[float]


class LossScaler:
    def __init__(self, config: LossScalerConfig, parameter_groups: Any = None) -> None:
        self.config = config
        self._current_scale = config.initial_scale
        self._current_hysteresis = config.hysteresis
        self._no_overflow_steps = 0

    @property
    def scale(self) -> float:
        return self._current_scale if self.config.enable else 1.0

    def step(self, overflow: Optional[bool] = None) -> LossScalerOutput:
        if not self.config.enable:
            return LossScalerOutput(None, None, None)
        overflow = bool(overflow)
        c = self.config
        if overflow:
            if c.hysteresis == 1 or self._current_hysteresis == 1:
                self._current_scale = max(self._current_scale / c.factor, c.min_scale)
            else:
                self._current_hysteresis -= 1
            self._no_overflow_steps = 0
        else:
            if c.consecutive_hysteresis:
                self._current_hysteresis = c.hysteresis
            if self._no_overflow_steps > 0 and self._no_overflow_steps % c.window == 0:
                if not c.consecutive_hysteresis:
                    self._current_hysteresis = c.hysteresis
                self._current_scale *= c.factor
            self._no_overflow_steps += 1
        return LossScalerOutput(overflow, self._no_overflow_steps, self._current_scale)

    def scale_loss(self, loss: torch.Tensor) -> torch.Tensor:
        return loss * self._current_scale if self.config.enable else loss

    def state_dict(self) -> Dict[str, Any]:
        return {"current_scale": self._current_scale, "current_hysteresis": self._current_hysteresis,
                "no_overflow_steps": self._no_overflow_steps}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        self._current_scale = state_dict["current_scale"]
        self._current_hysteresis = state_dict["current_hysteresis"]
        self._no_overflow_steps = state_dict["no_overflow_steps"]
