"""Megatron-style tensor-parallel linears and embedding (reference: core/nn/linear/column_parallel_linear.py:1-159,
row_parallel_linear.py:1-169, vocab_parallel_embedding.py:1-147).

Parameter names, shapes and TP split dimensions are the checkpoint contract and are kept; the math runs through
:mod:`scaling_b200.ops.functional` (tcgen05 GEMM with main-grad accumulation on B200, F.linear elsewhere).
Under sequence parallelism a row-parallel bias is added once, after the reduce-scatter, by the caller
(``skip_bias_add``) — the reference adds it before the reduction and so multiplies it by tp.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch

from ...ops import functional as OF
from ...parallel.symm import get_tp_comm
from ..topology import Topology
from .parameter_meta import CoreParameterMeta
from .tp_ops import (
    all_concat,
    all_reduce,
    all_reduce_scatter_to_sequence_parallel,
    all_shard,
    copy_to_tensor_model_parallel_region,
    get_device,
)

InitFn = Callable[[torch.Tensor], torch.Tensor]


def _bias_name(bitfit_bias_name: Optional[str]) -> str:
    return "bias" if not bitfit_bias_name else f"bias_{bitfit_bias_name}"


class ColumnParallelLinear(torch.nn.Module):
    """Y = X·Aᵀ + b with A split along its output dimension: weight [out/tp, in]."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device: Optional[torch.device] = None,
                 dtype: torch.dtype = torch.float32, topology: Optional[Topology] = None,
                 init_method: InitFn = torch.nn.init.xavier_normal_, parallel_output: bool = False,
                 bitfit_bias_name: Optional[str] = None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self._device = get_device(topology=topology, device=device)
        self.dtype, self.topology, self.init_method = dtype, topology, init_method
        self.parallel_output = parallel_output
        self.model_parallel_size = 1 if topology is None else topology.config.model_parallel_size
        assert out_features % self.model_parallel_size == 0, (
            f"cannot column parallelize, out_features ({out_features}) needs to be divisible by model parallel size "
            f"({self.model_parallel_size})")
        self.output_features_per_partition = out_features // self.model_parallel_size
        self.weight = torch.nn.Parameter(torch.empty(self.output_features_per_partition, in_features,
                                                     device=self._device, dtype=dtype))
        init_method(self.weight)
        CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=True, model_parallel_dimension=0)
        self.bias_name: Optional[str] = None
        if bias:
            self.bias_name = _bias_name(bitfit_bias_name)
            b = torch.nn.Parameter(torch.zeros(self.output_features_per_partition, device=self._device, dtype=dtype))
            setattr(self, self.bias_name, b)
            CoreParameterMeta.register_on_parameter(b, is_model_parallel=True, model_parallel_dimension=0)

    def get_bias(self) -> Optional[torch.Tensor]:
        return None if self.bias_name is None else getattr(self, self.bias_name)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sp = self.topology is not None and self.topology.config.sequence_parallel
        comm = get_tp_comm(self.topology) if x.is_cuda else None
        if comm is not None:  # dX all-reduce runs inside the dgrad GEMM launch
            y = OF.linear(x, self.weight, self.get_bias(), comm=comm, reduce_input_grad=True)
        else:
            if not sp:  # under SP the gather in front already turns the gradient into a reduce-scatter
                x = copy_to_tensor_model_parallel_region(x, self.topology)
            y = OF.linear(x, self.weight, self.get_bias())
        if not self.parallel_output:
            y = all_concat(y, dim=-1, topology=self.topology)
        return y


class RowParallelLinear(torch.nn.Module):
    """Y = X·Aᵀ + b with A split along its input dimension: weight [out, in/tp]; partial sums are reduced."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, device: Optional[torch.device] = None,
                 dtype: torch.dtype = torch.float32, topology: Optional[Topology] = None,
                 init_method: InitFn = torch.nn.init.xavier_normal_, parallel_input: bool = False,
                 parallel_output: bool = False, bitfit_bias_name: Optional[str] = None):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self._device = get_device(topology=topology, device=device)
        self.dtype, self.topology, self.init_method = dtype, topology, init_method
        self.parallel_input, self.parallel_output = parallel_input, parallel_output
        self.model_parallel_size = 1 if topology is None else topology.config.model_parallel_size
        assert in_features % self.model_parallel_size == 0, (
            f"cannot row parallelize, in_features ({in_features}) needs to be divisible by model parallel size "
            f"({self.model_parallel_size})")
        self.input_features_per_partition = in_features // self.model_parallel_size
        self.weight = torch.nn.Parameter(torch.empty(out_features, self.input_features_per_partition,
                                                     device=self._device, dtype=dtype))
        init_method(self.weight)
        CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=True, model_parallel_dimension=1)
        self.bias_name: Optional[str] = None
        if bias:
            self.bias_name = _bias_name(bitfit_bias_name)
            b = torch.nn.Parameter(torch.zeros(out_features, device=self._device, dtype=dtype))
            setattr(self, self.bias_name, b)
            CoreParameterMeta.register_on_parameter(b, is_model_parallel=False)

    def get_bias(self) -> Optional[torch.Tensor]:
        return None if self.bias_name is None else getattr(self, self.bias_name)

    def forward(self, x: torch.Tensor, skip_bias_add: bool = False, reduce: bool = False) -> torch.Tensor:
        """``parallel_output=True`` returns this rank's partial sum (the caller reduces, e.g. reduce-scatter under
        SP) and never adds the bias — unless ``reduce`` asks for the all-reduced, biased result anyway (callers
        that know they are not sequence parallel; the reduction then runs inside the GEMM launch over NVLink peer
        memory when a TPComm exists).  ``skip_bias_add`` lets the caller fuse bias + dropout + residual."""
        if not self.parallel_input:
            x = all_shard(x, dim=-1, topology=self.topology)
        if self.parallel_output and not reduce:
            return OF.linear(x, self.weight, None)
        b = None if skip_bias_add else self.get_bias()
        comm = get_tp_comm(self.topology) if x.is_cuda else None
        if comm is not None:
            return OF.linear(x, self.weight, b, comm=comm, reduce_output=True)
        y = all_reduce(OF.linear(x, self.weight, None), self.topology)
        return y if b is None else y + b


class VocabParallelEmbedding(torch.nn.Module):
    """Embedding table split along the vocabulary: each TP rank owns rows [start, end)."""

    def __init__(self, num_embeddings: int, embedding_dim: int, device: Optional[torch.device] = None,
                 dtype: torch.dtype = torch.float32, topology: Optional[Topology] = None,
                 init_method: InitFn = torch.nn.init.xavier_normal_, finetunable_token_ids: Optional[List[int]] = None):
        super().__init__()
        self.num_embeddings, self.embedding_dim = num_embeddings, embedding_dim
        self._device = get_device(topology=topology, device=device)
        self.dtype, self.topology = dtype, topology
        self.model_parallel_size = 1 if topology is None else topology.config.model_parallel_size
        rank = 0 if topology is None else topology.model_parallel_rank
        assert num_embeddings % self.model_parallel_size == 0, (
            f"cannot parallelize embedding, num_embeddings ({num_embeddings}) needs to be divisible by model parallel "
            f"size ({self.model_parallel_size})")
        self.num_embeddings_per_partition = num_embeddings // self.model_parallel_size
        self.vocab_start_index = rank * self.num_embeddings_per_partition
        self.vocab_end_index = self.vocab_start_index + self.num_embeddings_per_partition
        self.weight = torch.nn.Parameter(torch.empty(self.num_embeddings_per_partition, embedding_dim,
                                                     device=self._device, dtype=dtype))
        init_method(self.weight)
        CoreParameterMeta.register_on_parameter(self.weight, is_model_parallel=True, model_parallel_dimension=0)
        self.finetunable_token_ids = list(finetunable_token_ids or [])
        self._grad_mask: Optional[torch.Tensor] = None
        if self.finetunable_token_ids:
            mask = torch.zeros(self.num_embeddings_per_partition, 1, device=self._device, dtype=dtype)
            for t in self.finetunable_token_ids:
                if self.vocab_start_index <= t < self.vocab_end_index:
                    mask[t - self.vocab_start_index] = 1
            self._grad_mask = mask
            self.weight.register_hook(lambda g: g * self._grad_mask.to(g.dtype))
            # the row mask lives in an autograd hook, so this weight's gradient must travel through autograd instead
            # of being accumulated straight into the optimizer's main-grad arena by the embedding backward kernel
            self.weight._sb_no_main_grad = True

    def grad_mask(self) -> Optional[torch.Tensor]:
        """[V_local, 1] 0/1 mask of trainable rows (``finetunable_token_ids``), or None when all rows train."""
        return self._grad_mask

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        sp = self.topology is not None and bool(self.topology.config.sequence_parallel)
        if self.model_parallel_size > 1 and self.weight.is_cuda:
            # every row lives on one rank: its owner stores it into the peers' outputs over NVLink (no dense all-reduce)
            out = OF.vocab_embedding_reduced(x, self.weight, self.vocab_start_index, self.vocab_end_index,
                                             get_tp_comm(self.topology, sequence_parallel=sp), sp)
            if out is not None:
                return out
        out = OF.vocab_embedding(x, self.weight, self.vocab_start_index, self.vocab_end_index)
        if self.topology is not None and self.topology.config.sequence_parallel:
            return all_reduce_scatter_to_sequence_parallel(out, self.topology)
        return all_reduce(out, self.topology)
