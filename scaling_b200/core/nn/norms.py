"""LayerNorm / RMSNorm modules (reference: core/nn/norm/layernorm.py:1-87, rms_norm.py:1-63, get_norm.py,
layernorm_config.py).  Under sequence parallelism the norm works on the local [b, s/tp, h] shard and all-gathers its
output along the sequence for the column-parallel GEMMs that follow.  ``forward(x, residual=r)`` fuses the residual
add into the same kernel pass and returns ``(y, x + r)``.
"""
from __future__ import annotations

from enum import Enum
from typing import Optional, Union

import torch
from pydantic import Field

from ...ops import functional as OF
from ..config import BaseConfig
from ..topology import Topology
from .parameter_meta import CoreParameterMeta
from .tp_ops import gather_from_sequence_parallel_region


class LayerNormOptimizationType(Enum):
    TORCH = "torch"
    SM100 = "sm100"  # accepted for explicitness; on a B200 the sm_100a kernel is used for either value


class LayerNormConfig(BaseConfig):
    optimization_type: LayerNormOptimizationType = Field(
        LayerNormOptimizationType.TORCH, description="kept for config compatibility; kernels are selected by device")
    layernorm_epsilon: float = Field(1.0e-5, description="value added to the variance for numerical stability")


class NormType(Enum):
    LAYERNORM = "layernorm"
    RMS = "rms"


def _sp(topology: Optional[Topology]) -> bool:
    return topology is not None and topology.config.sequence_parallel and topology.config.model_parallel_size > 1


class LayerNorm(torch.nn.Module):
    def __init__(self, config: LayerNormConfig, normalized_shape: int, device: Optional[torch.device] = None,
                 dtype: torch.dtype = torch.float32, bitfit_bias_name: Optional[str] = None,
                 topology: Optional[Topology] = None):
        super().__init__()
        self.config = config
        self.normalized_shape = normalized_shape
        self.topology = topology
        dev = device if device is not None else (topology.device if topology is not None else None)
        self.weight = torch.nn.Parameter(torch.ones(normalized_shape, device=dev, dtype=dtype))
        CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=False)
        self.bias_name = "bias" if not bitfit_bias_name else f"bias_{bitfit_bias_name}"
        bias = torch.nn.Parameter(torch.zeros(normalized_shape, device=dev, dtype=dtype))
        setattr(self, self.bias_name, bias)
        CoreParameterMeta.register_on_parameter(bias, is_model_parallel=False)

    def get_bias(self) -> torch.Tensor:
        return getattr(self, self.bias_name)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None, gather: bool = True):
        out = OF.layer_norm(x, self.weight, self.get_bias(), self.config.layernorm_epsilon, residual=residual)
        y, s = (out, None) if residual is None else out
        if gather and _sp(self.topology):
            y = gather_from_sequence_parallel_region(y, self.topology)
        return y if residual is None else (y, s)


class RMSNorm(torch.nn.Module):
    def __init__(self, config: LayerNormConfig, dimensions: int, device: Optional[torch.device] = None,
                 dtype: torch.dtype = torch.float32, topology: Optional[Topology] = None):
        super().__init__()
        self.config = config
        self.eps = config.layernorm_epsilon
        self.topology = topology
        dev = device if device is not None else (topology.device if topology is not None else None)
        self.weight = torch.nn.Parameter(torch.ones(dimensions, device=dev, dtype=dtype))
        CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=False)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None, gather: bool = True):
        out = OF.rms_norm(x, self.weight, self.eps, residual=residual)
        y, s = (out, None) if residual is None else out
        if gather and _sp(self.topology):
            y = gather_from_sequence_parallel_region(y, self.topology)
        return y if residual is None else (y, s)


def get_norm(norm_type: NormType, layernorm_config: Optional[LayerNormConfig], dimensions: int,
             device: Optional[torch.device], dtype: torch.dtype, bitfit_bias_name: Optional[str] = None,
             topology: Optional[Topology] = None) -> Union[LayerNorm, RMSNorm]:
    assert layernorm_config is not None
    if norm_type == NormType.LAYERNORM:
        return LayerNorm(layernorm_config, dimensions, device=device, dtype=dtype, bitfit_bias_name=bitfit_bias_name,
                         topology=topology)
    if norm_type == NormType.RMS:
        return RMSNorm(layernorm_config, dimensions, device=device, dtype=dtype, topology=topology)
    raise NotImplementedError(f"{norm_type}")
