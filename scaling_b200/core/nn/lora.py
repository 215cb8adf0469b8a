"""LoRA adapters built from the tensor-parallel linears (reference: core/nn/lora.py:1-166, lora_config.py:1-44).

Parameter names/shapes follow the reference (``dense_in`` = A with the rank dimension TP-split, ``dense_out`` = B) so
checkpoints interchange.  ``get_delta_weights`` returns this rank's slice of (α/r)·B·A for merging into the base weight.
"""
from __future__ import annotations

import math
from enum import Enum
from functools import partial
from typing import List, Optional, Union

import torch
from pydantic import Field

from ..config import BaseConfig
from ..topology import Topology
from .linear import ColumnParallelLinear, RowParallelLinear
from .tp_ops import all_concat


class LoRAModuleType(Enum):
    QUERY = "query"
    KEY = "key"
    VALUE = "value"
    DENSE = "dense"


class LoRaConfig(BaseConfig):
    name: str = Field("lora", description="suffix of the parameter names / separate checkpoint file")
    rank: int = Field(64, description="rank r of the low-rank update (over all heads)")
    parallel_modules: List[LoRAModuleType] = Field(
        [LoRAModuleType.DENSE, LoRAModuleType.KEY, LoRAModuleType.VALUE, LoRAModuleType.QUERY],
        description="attention projections that receive a LoRA adapter")
    dropout: float = Field(0.0, description="dropout on the rank-r activations")
    alpha: int = Field(1, description="scaling numerator: update is scaled by alpha / rank")
    bias: bool = Field(False, description="add biases to the LoRA projections")
    kaiming_a: float = Field(math.sqrt(5), description="`a` of the kaiming-uniform init of A")


class ParallelLoRa(torch.nn.Module):
    def __init__(self, in_features: int, out_features: int, lora_module_type: LoRAModuleType, rank: int,
                 bias: bool = False, alpha: int = 1, dtype: torch.dtype = torch.float32,
                 kaiming_a: Optional[float] = 1.0e-5, device: Optional[torch.device] = None,
                 dropout: Optional[float] = None, topology: Optional[Topology] = None):
        super().__init__()
        assert rank <= in_features, f"LoRa rank {rank} exceeds the input dimensionality {in_features}"
        self.scaling = alpha / rank
        self.topology = topology
        self.lora_module_type = lora_module_type
        self.model_parallel_size = 1 if topology is None else topology.config.model_parallel_size
        if self.model_parallel_size > 1:
            assert rank % self.model_parallel_size == 0, "rank must be divisible by the model parallel size"
            assert out_features % self.model_parallel_size == 0, "out_features must be divisible by the model parallel size"
        self.dropout = torch.nn.Dropout(dropout) if dropout else None
        a_init = partial(torch.nn.init.kaiming_uniform_, a=kaiming_a)
        kw = dict(bias=bias, device=device, dtype=dtype, topology=topology)
        self.dense_out: Union[ColumnParallelLinear, RowParallelLinear]
        if lora_module_type != LoRAModuleType.DENSE:
            self.dense_in = ColumnParallelLinear(in_features, rank, init_method=a_init, parallel_output=False, **kw)
            self.dense_out = ColumnParallelLinear(rank, out_features, init_method=torch.nn.init.zeros_,
                                                  parallel_output=True, **kw)
        else:
            self.dense_in = ColumnParallelLinear(in_features, rank, init_method=a_init, parallel_output=True, **kw)
            self.dense_out = RowParallelLinear(rank, out_features, init_method=torch.nn.init.zeros_,
                                               parallel_input=True, parallel_output=False, **kw)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = self.dense_in(x)
        if self.dropout is not None:
            h = self.dropout(h)
        return self.dense_out(h) * self.scaling

    def get_delta_weights(self) -> torch.Tensor:
        """This rank's slice of the merged update (α/r)·B·A, laid out like the base weight it is added to."""
        tp = self.model_parallel_size > 1
        is_dense = self.lora_module_type == LoRAModuleType.DENSE
        a = self.dense_in.weight
        b = self.dense_out.weight
        if tp:
            a = all_concat(a, dim=-2, topology=self.topology)
            b = all_concat(b, dim=-1 if is_dense else -2, topology=self.topology)
        delta = (b.float() @ a.float()) * self.scaling
        if tp:
            r = self.topology.model_parallel_rank
            if is_dense:
                n = delta.shape[1] // self.model_parallel_size
                delta = delta[:, r * n:(r + 1) * n]
            else:
                n = delta.shape[0] // self.model_parallel_size
                delta = delta[r * n:(r + 1) * n]
        return delta.to(self.dense_in.weight.dtype)
