"""Tensor-parallel MLP blocks (reference: core/nn/mlp.py:1-167).

``ParallelSwiGLUMLP`` keeps the reference's three parameters (``dense_in`` = gate, ``siglu_weight`` = up,
``dense_out``) but runs gate and up as ONE dual GEMM with the silu·mul in its TMEM epilogue, one dgrad and one wgrad
GEMM in backward (one dX all-reduce instead of two under TP).  Row-parallel biases are added once after the reduction.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from ...ops import functional as OF
from ...parallel.symm import get_tp_comm
from ..topology import Topology
from .activation import ActivationFunction, get_activation_function
from .linear import ColumnParallelLinear, RowParallelLinear
from .tp_ops import all_reduce_scatter_to_sequence_parallel, copy_to_tensor_model_parallel_region

InitFn = Callable[[torch.Tensor], torch.Tensor]


def _features(io_features: int, factor: float) -> int:
    f = io_features * factor
    assert float(int(f)) == f, "io_features * intermediate_feature_factor does not result in a natural number"
    return int(f)


def _sp(topology: Optional[Topology]) -> bool:
    return topology is not None and topology.config.sequence_parallel


def _row_parallel_out(layer: RowParallelLinear, h: torch.Tensor, topology: Optional[Topology]) -> torch.Tensor:
    """Row-parallel projection with its reduction (reduce-scatter under SP; otherwise the all-reduce, fused into the
    GEMM launch when the TP group has a peer-memory communicator) and the bias added once."""
    if not _sp(topology):
        return layer(h, reduce=True)
    y = None
    if layer.parallel_input and h.is_cuda:  # fused GEMM + reduce-scatter over NVLink peer memory
        y = OF.linear_reduce_scatter(h, layer.weight, get_tp_comm(topology, sequence_parallel=True))
    if y is None:
        y = all_reduce_scatter_to_sequence_parallel(layer(h), topology)
    b = layer.get_bias()
    return y if b is None else y + b


class ParallelMLP(torch.nn.Module):
    """h -> factor·h -> activation -> h."""

    def __init__(self, io_features: int, intermediate_feature_factor: float, bias: bool = True,
                 device: Optional[torch.device] = None, dtype: torch.dtype = torch.float32,
                 topology: Optional[Topology] = None, init_method: InitFn = torch.nn.init.xavier_normal_,
                 bitfit_bias_name: Optional[str] = None,
                 activation_function: ActivationFunction = ActivationFunction.GELU):
        super().__init__()
        f = _features(io_features, intermediate_feature_factor)
        self.topology = topology
        self.dense_in = ColumnParallelLinear(io_features, f, bias=bias, device=device, dtype=dtype, topology=topology,
                                             init_method=init_method, parallel_output=True,
                                             bitfit_bias_name=bitfit_bias_name)
        self.dense_out = RowParallelLinear(f, io_features, bias=bias, device=device, dtype=dtype, topology=topology,
                                           init_method=init_method, parallel_input=True, parallel_output=True,
                                           bitfit_bias_name=bitfit_bias_name)
        self.activation_function = get_activation_function(activation_function)
        self.activation_name = ActivationFunction(activation_function).value

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = None
        if x.is_cuda:
            # up-projection with bias + activation in the GEMM epilogue (one launch instead of GEMM, bias add, activation)
            comm = get_tp_comm(self.topology)
            xin = x if (comm is not None or _sp(self.topology)) else copy_to_tensor_model_parallel_region(x, self.topology)
            h = OF.linear_act(xin, self.dense_in.weight, self.dense_in.get_bias(), self.activation_name, comm=comm)
        if h is None:
            h = self.activation_function(self.dense_in(x))
        return _row_parallel_out(self.dense_out, h, self.topology)


class ParallelSwiGLUMLP(torch.nn.Module):
    """dense_out( silu(dense_in(x)) * siglu_weight(x) )."""

    def __init__(self, io_features: int, intermediate_feature_factor: float, bias: bool = True,
                 device: Optional[torch.device] = None, dtype: torch.dtype = torch.float32,
                 topology: Optional[Topology] = None, init_method: InitFn = torch.nn.init.xavier_normal_,
                 bitfit_bias_name: Optional[str] = None):
        super().__init__()
        f = _features(io_features, intermediate_feature_factor)
        self.topology = topology
        kw = dict(bias=bias, device=device, dtype=dtype, topology=topology, init_method=init_method,
                  bitfit_bias_name=bitfit_bias_name)
        self.dense_in = ColumnParallelLinear(io_features, f, parallel_output=True, **kw)
        self.siglu_weight = ColumnParallelLinear(io_features, f, parallel_output=True, **kw)
        self.dense_out = RowParallelLinear(f, io_features, parallel_input=True, parallel_output=True, **kw)

    def supports_sequence_parallel_shard(self) -> bool:
        """True when ``forward(x, sequence_parallel_shard=True)`` can take the norm's ungathered [b, s/tp, h] output and
        run the all-gather inside the gate|up dual GEMM."""
        return (_sp(self.topology) and self.dense_in.get_bias() is None and self.siglu_weight.get_bias() is None
                and get_tp_comm(self.topology, sequence_parallel=True) is not None)

    def forward(self, x: torch.Tensor, sequence_parallel_shard: bool = False) -> torch.Tensor:
        if sequence_parallel_shard:
            sp_comm = get_tp_comm(self.topology, sequence_parallel=True)
            if OF.sp_gather_fusable(x, sp_comm, self.dense_in.weight.shape[0]) and \
                    OF.fused_view((self.dense_in.weight.data, self.siglu_weight.weight.data)) is not None:
                h = OF.swiglu_linear(x, self.dense_in.weight, self.siglu_weight.weight, sp=sp_comm)
                return _row_parallel_out(self.dense_out, h, self.topology)
            from .tp_ops import gather_from_sequence_parallel_region

            x = gather_from_sequence_parallel_region(x, self.topology)
        comm = get_tp_comm(self.topology) if x.is_cuda else None
        if comm is None and not _sp(self.topology):
            x = copy_to_tensor_model_parallel_region(x, self.topology)  # one dX all-reduce for gate and up together
        bg, bu = self.dense_in.get_bias(), self.siglu_weight.get_bias()
        if bg is None and bu is None:
            h = OF.swiglu_linear(x, self.dense_in.weight, self.siglu_weight.weight, comm=comm)
        else:
            if comm is not None:
                x = OF._CopyToRegion.apply(x, comm)
            h = torch.nn.functional.silu(OF.linear(x, self.dense_in.weight, bg)) * OF.linear(x, self.siglu_weight.weight, bu)
        return _row_parallel_out(self.dense_out, h, self.topology)
