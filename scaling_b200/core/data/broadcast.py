"""Batch broadcast from TP rank 0 to its tensor-parallel group (reference: core/data/broadcast_data.py:1-165).

One metadata broadcast (all shapes of all tensors in one int64 vector) and one flat payload broadcast per dtype;
bool travels as int8.  With mp == 1 nothing is communicated and tensors are only moved to the device.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from ..topology import Topology

MAX_DIMS = 8


_MAX_DATA_DIM = MAX_DIMS


def _build_tensor_sizes(tensors: List[Optional[torch.Tensor]], model_parallel_rank: int) -> List[int]:
    """Shape table as one flat int list, ``MAX_DIMS`` entries per tensor: the sizes, padded with -1 (reference
    broadcast_data.py:20-41).  Only TP rank 0 knows the shapes; the others send all -1.  Extension: -2 in the first entry
    marks a tensor that is None."""
    sizes = [-1] * (len(tensors) * _MAX_DATA_DIM)
    if model_parallel_rank == 0:
        for i, t in enumerate(tensors):
            if t is None:
                sizes[i * _MAX_DATA_DIM] = -2
                continue
            assert t.dim() <= _MAX_DATA_DIM, f"tensors with more than {_MAX_DATA_DIM} dimensions cannot be broadcast"
            for d, n in enumerate(t.shape):
                sizes[i * _MAX_DATA_DIM + d] = int(n)
    return sizes


def _unpack_sizes(flat: List[int]) -> tuple:
    """(shapes, element counts) from a table built by :func:`_build_tensor_sizes`."""
    shapes: List[List[int]] = []
    numels: List[int] = []
    for i in range(0, len(flat), _MAX_DATA_DIM):
        shape = [int(n) for n in flat[i:i + _MAX_DATA_DIM] if n >= 0]
        shapes.append(shape)
        k = 1
        for n in shape:
            k *= n
        numels.append(k)
    return shapes, numels


def sync_sizes(tensors: Optional[List[Optional[torch.Tensor]]], n: int, topology: Topology) -> List[Optional[List[int]]]:
    """Broadcast the shapes (or "is None") of ``n`` tensors; -1 padding marks unused dims, -2 a missing tensor."""
    if topology.model_parallel_rank == 0:
        assert tensors is not None and len(tensors) == n
    table = _build_tensor_sizes(tensors if topology.model_parallel_rank == 0 else [None] * n, topology.model_parallel_rank)
    sizes = torch.tensor(table, dtype=torch.long).to(topology.device)
    dist.broadcast(sizes, src=topology.model_parallel_ranks[0], group=topology.model_parallel_group)
    flat = sizes.cpu().tolist()
    shapes, _ = _unpack_sizes(flat)
    return [None if flat[i * _MAX_DATA_DIM] == -2 else shape for i, shape in enumerate(shapes)]


def broadcast_data(tensors: Optional[List[Optional[torch.Tensor]]], dtype: torch.dtype, topology: Topology,
                   n: Optional[int] = None) -> List[Optional[torch.Tensor]]:
    """Returns the list of tensors (dtype ``dtype``) on ``topology.device`` on every rank of the TP group."""
    device = topology.device
    if topology.config.model_parallel_size == 1 or not topology.is_distributed_initialized:
        assert tensors is not None
        return [None if t is None else t.to(device=device, dtype=dtype, non_blocking=True) for t in tensors]
    if n is None:
        cnt = torch.tensor([len(tensors) if tensors is not None else 0], dtype=torch.long, device=device)
        dist.broadcast(cnt, src=topology.model_parallel_ranks[0], group=topology.model_parallel_group)
        n = int(cnt.item())
    shapes = sync_sizes(tensors, n, topology)
    wire = torch.int8 if dtype == torch.bool else dtype
    numels = [0 if s is None else int(torch.Size(s).numel()) for s in shapes]
    total = sum(numels)
    if topology.model_parallel_rank == 0:
        assert tensors is not None
        parts = [t.to(device=device, dtype=wire, non_blocking=True).reshape(-1) for t in tensors if t is not None]
        flat = torch.cat(parts) if parts else torch.empty(0, dtype=wire, device=device)
    else:
        flat = torch.empty(total, dtype=wire, device=device)
    if total > 0:
        dist.broadcast(flat, src=topology.model_parallel_ranks[0], group=topology.model_parallel_group)
    out: List[Optional[torch.Tensor]] = []
    off = 0
    for s, k in zip(shapes, numels):
        if s is None:
            out.append(None)
            continue
        t = flat[off:off + k].view(s)
        off += k
        out.append(t.to(dtype) if wire != dtype else t)
    return out
