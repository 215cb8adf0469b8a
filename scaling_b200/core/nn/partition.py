"""Assignment of the layer list to pipeline stages (reference: core/nn/parallel_module/pipeline_partitioning.py:1-136).

``balanced`` minimises the heaviest stage (by trainable parameter count) exactly, with a small dynamic program over
contiguous splits, instead of a binary search + probe; layers that expose ``estimate_parameter_count`` are not even
instantiated to be counted.
"""
from __future__ import annotations

from typing import List, NamedTuple, Optional, Sequence

import numpy as np

from .layers import LayerSpec


class PipePartitionCoordinates(NamedTuple):
    start: int
    end: int

    @property
    def length(self) -> int:
        return self.end - self.start


def pipe_partition_from_indices(partition_array: Sequence[int], num_layers: Optional[int] = None) -> List[PipePartitionCoordinates]:
    if num_layers is not None and partition_array[-1] != num_layers:
        raise ValueError(f"Last entry of partition_array needs to match num_layers; got partition_array="
                         f"{list(partition_array)}, num_layers={num_layers}.")
    if len(partition_array) < 2 or partition_array[0] != 0 or any(e <= s for s, e in zip(partition_array[:-1], partition_array[1:])):
        raise ValueError(f"partition_array must start at 0 and be strictly increasing (no empty stage); got {list(partition_array)}")
    return [PipePartitionCoordinates(int(s), int(e)) for s, e in zip(partition_array[:-1], partition_array[1:])]


def pipe_partition_uniform(item_count: int, partition_count: int) -> List[PipePartitionCoordinates]:
    """Equal stage lengths; the remainder goes to the first stages."""
    assert item_count >= partition_count, f"cannot partition {item_count} layers on {partition_count} pipe parallel stages"
    base, rest = divmod(item_count, partition_count)
    bounds = [0]
    for i in range(partition_count):
        bounds.append(bounds[-1] + base + (1 if i < rest else 0))
    return pipe_partition_from_indices(bounds)


def count_layer_params(layer_specs: Sequence[LayerSpec]) -> List[int]:
    counts = []
    for spec in layer_specs:
        n = spec.estimate_trainable_parameters()
        if n is None:
            layer = spec.initialize()
            n = sum(p.numel() for p in layer.parameters() if p.requires_grad)
            del layer
        counts.append(int(n))
    return counts


def balanced_bounds(weights: Sequence[float], partition_count: int) -> List[int]:
    """Contiguous split of ``weights`` into ``partition_count`` non-empty parts minimising the largest part sum."""
    n = len(weights)
    assert n >= partition_count
    prefix = np.concatenate([[0.0], np.cumsum(np.asarray(weights, dtype=np.float64))])
    INF = float("inf")
    best = np.full((partition_count + 1, n + 1), INF)
    cut = np.zeros((partition_count + 1, n + 1), dtype=np.int64)
    best[0, 0] = 0.0
    for k in range(1, partition_count + 1):
        for i in range(k, n - (partition_count - k) + 1):
            for j in range(k - 1, i):
                cost = max(best[k - 1, j], prefix[i] - prefix[j])
                if cost < best[k, i] - 1e-12:
                    best[k, i] = cost
                    cut[k, i] = j
    bounds = [n]
    i = n
    for k in range(partition_count, 0, -1):
        i = int(cut[k, i])
        bounds.append(i)
    return bounds[::-1]


def pipe_partition_balanced(layer_specs: Sequence[LayerSpec], partition_count: int, eps: float = 1e-3) -> List[PipePartitionCoordinates]:
    return pipe_partition_from_indices(balanced_bounds(count_layer_params(layer_specs), partition_count))
