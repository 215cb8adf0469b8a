"""A tensor/pipeline-parallel MLP built from the core building blocks (reference: examples/mlp_example/model.py).

Layers alternate column-parallel (→ ReLU, output stays sharded) and row-parallel (sums the shards), so the network can
be cut across TP ranks; every layer is its own pipeline-partitionable unit.
"""
from __future__ import annotations

from typing import Any, Dict, List, Tuple, Union

import torch

from scaling_b200.core import (
    BaseContext,
    BaseLayer,
    BaseLayerIO,
    ColumnParallelLinear,
    LayerSpec,
    Optimizer,
    OptimizerParamGroup,
    OptimizerParamGroupConfig,
    ParallelModule,
    RowParallelLinear,
    Topology,
)

from .config import MLPConfig
from .data import MNISTDatasetBatch


class MLPContext(BaseContext[MLPConfig]):
    """Run state of the example (iteration / sample counters, RNG, topology) typed with the example's config."""

    config: MLPConfig


class MLPLayerIO(BaseLayerIO):
    def __init__(self, activations: torch.Tensor):
        self.activations = activations


class MLPBaseLayer(BaseLayer[MLPLayerIO, MLPLayerIO, MLPLayerIO]):
    @staticmethod
    def input_to_tuple(input: Union[MLPLayerIO, MNISTDatasetBatch]) -> Tuple[Any, ...]:
        return (input.inputs,) if isinstance(input, MNISTDatasetBatch) else (input.activations,)

    @staticmethod
    def tuple_to_input(d: Tuple[Any, ...]) -> MLPLayerIO:
        return MLPLayerIO(activations=d[0])

    @staticmethod
    def output_to_tuple(output: MLPLayerIO) -> Tuple[Any, ...]:
        return (output.activations,)

    @staticmethod
    def tuple_to_last_stage_activation(d: Tuple[Any, ...]) -> MLPLayerIO:
        return MLPLayerIO(activations=d[0])


def _activations(x: Union[MLPLayerIO, MNISTDatasetBatch]) -> torch.Tensor:
    if isinstance(x, MNISTDatasetBatch):
        return torch.flatten(x.inputs, start_dim=1)
    return x.activations


class MLPLinearColumnParallel(MLPBaseLayer):
    def __init__(self, in_features: int, out_features: int, topology: Topology, relu: bool = True):
        super().__init__()
        self.linear = ColumnParallelLinear(in_features, out_features, topology=topology, parallel_output=True)
        self.relu = relu

    def forward(self, x: Union[MLPLayerIO, MNISTDatasetBatch]) -> MLPLayerIO:
        h = self.linear(_activations(x))
        return MLPLayerIO(activations=torch.relu(h) if self.relu else h)


class MLPLinearRowParallel(MLPBaseLayer):
    def __init__(self, in_features: int, out_features: int, topology: Topology, relu: bool = True):
        super().__init__()
        self.linear = RowParallelLinear(in_features, out_features, topology=topology, parallel_input=True)
        self.relu = relu

    def forward(self, x: Union[MLPLayerIO, MNISTDatasetBatch]) -> MLPLayerIO:
        h = self.linear(_activations(x))
        return MLPLayerIO(activations=torch.relu(h) if self.relu else h)


def loss_function(output: MLPLayerIO, batch: MNISTDatasetBatch) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
    logits = output.activations.float()
    loss = torch.nn.functional.cross_entropy(logits, batch.targets)
    accuracy = (logits.argmax(dim=1) == batch.targets).float().mean()
    return loss, {"accuracy": accuracy}


def metrics_aggregation_fn(topology: Topology, metrics: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    out = {k: torch.stack([m[k] for m in metrics]).mean() for k in metrics[0]}
    if topology.config.data_parallel_size > 1 and topology.is_distributed_initialized:
        for k in out:
            torch.distributed.all_reduce(out[k], group=topology.data_parallel_group)
            out[k] = out[k] / topology.config.data_parallel_size
    return out


def get_layer_specs(context: MLPContext) -> List[LayerSpec]:
    a, topology = context.config.architecture, context.topology
    specs = [LayerSpec(MLPLinearColumnParallel, in_features=784, out_features=a.hidden_dim, topology=topology)]
    for _ in range(a.n_hidden_layers):
        specs.append(LayerSpec(MLPLinearRowParallel, in_features=a.hidden_dim, out_features=a.hidden_dim, topology=topology))
        specs.append(LayerSpec(MLPLinearColumnParallel, in_features=a.hidden_dim, out_features=a.hidden_dim, topology=topology))
    specs.append(LayerSpec(MLPLinearRowParallel, in_features=a.hidden_dim, out_features=10, topology=topology, relu=False))
    return specs


def init_model(context: MLPContext) -> ParallelModule:
    return ParallelModule(layer_specs=get_layer_specs(context), topology=context.topology,
                          profiler_config=context.config.profiler)


def init_optimizer(context: MLPContext, model: ParallelModule) -> Optimizer:
    named = model.named_parameters_with_meta()
    lr = context.config.learning_rate_scheduler
    groups = [
        OptimizerParamGroup([x for x in named if not x[0].endswith("bias")],
                            OptimizerParamGroupConfig(name="weight_decay_params", weight_decay=context.config.architecture.weight_decay,
                                                      learning_rate_scheduler=lr)),
        OptimizerParamGroup([x for x in named if x[0].endswith("bias")],
                            OptimizerParamGroupConfig(name="no_weight_decay_params", weight_decay=0.0, learning_rate_scheduler=lr)),
    ]
    return Optimizer(config=context.config.optimizer, parameter_groups=groups, topology=context.topology)
