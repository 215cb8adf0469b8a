"""Tensor-/sequence-parallel conjugate operators as autograd Functions (reference: core/nn/linear/utils.py:20-383).

Six (forward, backward) pairs; all of them are the identity when the TP group has one rank:

    copy_to_tensor_model_parallel_region   identity            / all-reduce
    all_reduce                             all-reduce          / identity
    all_concat(dim)                        all-gather + concat / keep own slice
    all_shard(dim)                         keep own slice      / all-gather + concat
    all_reduce_scatter_to_sequence_parallel  reduce-scatter(seq, dim 1) / all-gather(seq)
    gather_from_sequence_parallel_region     all-gather(seq)            / reduce-scatter(seq) (or plain split)

This module is the torch.distributed (NCCL / gloo) implementation — the oracle and the baseline inside the product.
The hot TP collectives are additionally available fused with their GEMMs over NVLink peer memory
(:mod:`scaling_b200.parallel`).
"""
from __future__ import annotations

from typing import Any, Optional

import torch
import torch.distributed as dist

from ..topology import Topology


def get_device(topology: Optional[Topology] = None, device: Optional[torch.device] = None) -> torch.device:
    assert topology is None or device is None, "cannot specify both device and topology"
    if topology is not None:
        return topology.device
    if device is not None:
        return device
    if torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def _tp(topology: Optional[Topology]) -> tuple[int, int, Any]:
    if topology is None or topology.config.model_parallel_size == 1:
        return 1, 0, None
    return topology.config.model_parallel_size, topology.model_parallel_rank, topology.model_parallel_group


# ---- raw collectives --------------------------------------------------------------------------------------------
def _all_reduce_raw(x: torch.Tensor, group: Any) -> torch.Tensor:
    x = x.contiguous()
    dist.all_reduce(x, group=group)
    return x


def _all_gather_dim(x: torch.Tensor, dim: int, size: int, group: Any) -> torch.Tensor:
    x = x.contiguous()
    if x.dim() == 0:
        x = x.reshape(1)
    out = torch.empty((size * x.shape[0], *x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=group)
    dim = dim % x.dim()
    if dim == 0:
        return out
    return torch.cat(list(out.view(size, *x.shape).unbind(0)), dim=dim)


def _shard_dim(x: torch.Tensor, dim: int, size: int, rank: int) -> torch.Tensor:
    n = x.shape[dim] // size
    return x.narrow(dim, rank * n, n).contiguous()


def _peer_comm(x: torch.Tensor, group: Any):
    """The TP group's peer-memory communicator when ``x`` can take the NVLink peer-store path (CUDA, bf16, [b, s, h])."""
    if not x.is_cuda:
        return None
    from ...ops import kernels as K
    from ...parallel.symm import comm_for_group

    comm = comm_for_group(group)
    return comm if K.seq_peer_ok(x, comm) else None


def _reduce_scatter_seq(x: torch.Tensor, size: int, group: Any) -> torch.Tensor:
    """[b, s, h] -> [b, s/size, h], summed over the group. The sequence is dimension 1."""
    assert x.shape[1] % size == 0, "sequence length must be divisible by the tensor parallel size"
    comm = _peer_comm(x, group)
    if comm is not None:  # peer stores + device barrier + local sum (own kernels), no NCCL on the layer path
        from ...ops import kernels as K

        return K.seq_reduce_scatter(x, comm)
    # reduce_scatter_tensor splits dim 0 -> bring the sequence dim to the front
    xt = x.transpose(0, 1).contiguous()
    out = torch.empty((xt.shape[0] // size, *xt.shape[1:]), dtype=x.dtype, device=x.device)
    dist.reduce_scatter_tensor(out, xt, group=group)
    return out.transpose(0, 1).contiguous()


def _all_gather_seq(x: torch.Tensor, size: int, group: Any) -> torch.Tensor:
    comm = _peer_comm(x, group)
    if comm is not None:
        from ...ops import kernels as K

        return K.seq_all_gather(x, comm)
    xt = x.transpose(0, 1).contiguous()
    out = torch.empty((xt.shape[0] * size, *xt.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, xt, group=group)
    return out.transpose(0, 1).contiguous()


# ---- autograd functions -----------------------------------------------------------------------------------------
class _Copy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, g):
        return _all_reduce_raw(g, ctx.group), None


class _AllReduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, group):
        return _all_reduce_raw(x.clone() if x.requires_grad else x, group)

    @staticmethod
    def backward(ctx, g):
        return g, None


class _AllConcat(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, size, rank, group):
        ctx.args = (dim, size, rank)
        return _all_gather_dim(x, dim, size, group)

    @staticmethod
    def backward(ctx, g):
        dim, size, rank = ctx.args
        return _shard_dim(g, dim, size, rank), None, None, None, None


class _AllShard(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dim, size, rank, group):
        ctx.args = (dim, size, group)
        return _shard_dim(x, dim, size, rank)

    @staticmethod
    def backward(ctx, g):
        dim, size, group = ctx.args
        return _all_gather_dim(g, dim, size, group), None, None, None, None


class _ReduceScatterSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, size, group):
        ctx.args = (size, group)
        return _reduce_scatter_seq(x, size, group)

    @staticmethod
    def backward(ctx, g):
        size, group = ctx.args
        return _all_gather_seq(g, size, group), None, None


class _GatherSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, size, rank, group, reduce_scatter_grads):
        ctx.args = (size, rank, group, reduce_scatter_grads)
        return _all_gather_seq(x, size, group)

    @staticmethod
    def backward(ctx, g):
        size, rank, group, rs = ctx.args
        if rs:
            return _reduce_scatter_seq(g, size, group), None, None, None, None
        return _shard_dim(g, 1, size, rank), None, None, None, None


# ---- public API (reference names) -------------------------------------------------------------------------------
def copy_to_tensor_model_parallel_region(input_tensor: torch.Tensor, topology: Optional[Topology]) -> torch.Tensor:
    size, _, group = _tp(topology)
    return input_tensor if size == 1 else _Copy.apply(input_tensor, group)


def all_reduce(input_tensor: torch.Tensor, topology: Optional[Topology]) -> torch.Tensor:
    size, _, group = _tp(topology)
    return input_tensor if size == 1 else _AllReduce.apply(input_tensor, group)


def all_concat(input_tensor: torch.Tensor, dim: int, topology: Optional[Topology]) -> torch.Tensor:
    size, rank, group = _tp(topology)
    return input_tensor if size == 1 else _AllConcat.apply(input_tensor, dim, size, rank, group)


def all_concat_with_group(input_tensor: torch.Tensor, dim: int, group: Any) -> torch.Tensor:
    size = dist.get_world_size(group)
    if size == 1:
        return input_tensor
    return _AllConcat.apply(input_tensor, dim, size, dist.get_rank(group), group)


def all_shard(input_tensor: torch.Tensor, dim: int, topology: Optional[Topology]) -> torch.Tensor:
    size, rank, group = _tp(topology)
    return input_tensor if size == 1 else _AllShard.apply(input_tensor, dim, size, rank, group)


def all_reduce_scatter_to_sequence_parallel(input_tensor: torch.Tensor, topology: Optional[Topology]) -> torch.Tensor:
    size, _, group = _tp(topology)
    return input_tensor if size == 1 else _ReduceScatterSeq.apply(input_tensor, size, group)


def gather_from_sequence_parallel_region(input_tensor: torch.Tensor, topology: Optional[Topology],
                                         tensor_parallel_output_grad: bool = True) -> torch.Tensor:
    """``tensor_parallel_output_grad=True``: the consumer is tensor parallel, so gradients arrive as partial sums and
    are reduce-scattered; ``False``: gradients are replicated and are simply split (the reference's two branches are
    identical by mistake — SURVEY §9.2 #4)."""
    size, rank, group = _tp(topology)
    return input_tensor if size == 1 else _GatherSeq.apply(input_tensor, size, rank, group, tensor_parallel_output_grad)
