"""Activation selector (reference: core/nn/activation_function.py:1-17)."""
from enum import Enum
from typing import Callable

import torch

from ...ops import functional as OF


class ActivationFunction(Enum):
    GELU = "gelu"
    SILU = "silu"


def get_activation_function(activation_function: ActivationFunction) -> Callable[[torch.Tensor], torch.Tensor]:
    if activation_function == ActivationFunction.GELU:
        return lambda x: OF.activation(x, "gelu")
    if activation_function == ActivationFunction.SILU:
        return lambda x: OF.activation(x, "silu")
    raise NotImplementedError
