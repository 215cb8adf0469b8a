"""Merge / split tensor-parallel parameters for layout-independent checkpoints (reference: core/utils/param_merge.py).

Merge is one ``all_gather`` per tensor over the TP group (the reference broadcasts ``tp`` times through the parameter's
own storage); the merged tensor lands on the CPU.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from ..nn.parameter_meta import CoreParameterMeta
from ..topology import Topology


def merge_parameter(parameter: torch.Tensor, core_parameter_meta: CoreParameterMeta, topology: Optional[Topology]) -> torch.Tensor:
    """Full (TP-merged) tensor on the CPU. Collective over the TP group when the parameter is model parallel."""
    local = parameter.detach()
    if topology is None or topology.config.model_parallel_size == 1 or not core_parameter_meta.is_model_parallel \
            or not topology.is_distributed_initialized:
        return local.cpu().clone()
    tp = topology.config.model_parallel_size
    local = local.contiguous()
    flat = local.reshape(-1)
    gathered = torch.empty(tp * flat.numel(), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, flat, group=topology.model_parallel_group)
    parts = [g.view(local.shape) for g in gathered.view(tp, -1).unbind(0)]
    replicas = getattr(core_parameter_meta, "kv_replicas", 1)
    if replicas > 1:
        parts = parts[::replicas]
    return torch.cat(parts, dim=core_parameter_meta.model_parallel_dimension).cpu()


def split_parameter(parameter: torch.Tensor, core_parameter_meta: CoreParameterMeta, topology: Optional[Topology]) -> torch.Tensor:
    """This rank's TP slice of a merged tensor."""
    if topology is None or topology.config.model_parallel_size == 1 or not core_parameter_meta.is_model_parallel:
        return parameter
    tp = topology.config.model_parallel_size
    rank = topology.model_parallel_rank
    replicas = getattr(core_parameter_meta, "kv_replicas", 1)
    parts = tp // replicas
    dim = core_parameter_meta.model_parallel_dimension
    assert parameter.shape[dim] % parts == 0, f"cannot split dimension {dim} of {tuple(parameter.shape)} into {parts}"
    n = parameter.shape[dim] // parts
    return parameter.narrow(dim, (rank // replicas) * n, n).clone()
