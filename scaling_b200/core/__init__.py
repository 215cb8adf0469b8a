"""Model-agnostic core.  The flat namespace mirrors ``scaling.core`` name for name (reference core/__init__.py), but it
is resolved lazily: ``from scaling_b200.core import Topology`` imports ``core.topology`` only."""
from __future__ import annotations

import importlib
from typing import Any, Dict, List

_EXPORTS: Dict[str, str] = {
    "config": "BaseConfig overwrite_recursive",
    "context": "BaseContext BaseContextGeneric DeterminedBaseContext",
    "data": ("BaseBlendedDataset BaseDataset BaseDatasetBatch BaseDatasetItem BaseLayerIO BlendedDatasetConfig DataLoader "
             "FileDataset MemoryMapDataset MemoryMapDatasetBuilder broadcast_data"),
    "logging": "LoggerConfig logger",
    "nn": ("ActivationFunction BaseLayer ColumnParallelLinear CoreParameterMeta LayerNorm LayerNormConfig "
           "LayerNormOptimizationType LayerSpec LoRaConfig LoRAModuleType MaskedSoftmax MaskedSoftmaxConfig "
           "MaskedSoftmaxKernel NormType ParallelLoRa ParallelMLP ParallelModule ParallelSelfAttention ParallelSwiGLUMLP "
           "PipelineScheduleInference PipelineScheduleTrain PipePartitionCoordinates RelativePositionEmbeddingType RMSNorm "
           "RotaryConfig RotaryEmbedding RotaryEmbeddingComplex RowParallelLinear TiedLayerSpec VocabParallelEmbedding "
           "get_activation_function get_norm pipe_partition_uniform"),
    "optimizer": ("BaseOptimizer LearningRateDecayStyle LearningRateScheduler LearningRateSchedulerConfig LossScaler "
                  "LossScalerConfig Optimizer OptimizerConfig OptimizerParamGroup OptimizerParamGroupConfig"),
    "profiler": "Profiler ProfilerConfig SynchronizedTimer",
    "runner": "LaunchConfig RunnerConfig RunnerDockerConfig RunnerType runner_main",
    "topology": "ActivationCheckpointingType PipePartitionMethod Topology TopologyConfig",
    "trainer": "BaseTrainer DeterminedBaseTrainer TrainerConfig",
}
_WHERE: Dict[str, str] = {name: module for module, names in _EXPORTS.items() for name in names.split()}
__all__: List[str] = sorted(_WHERE)


def __getattr__(name: str) -> Any:
    module = _WHERE.get(name)
    if module is None:
        if name in _EXPORTS or name in ("determined", "utils"):  # plain sub-packages / modules
            return importlib.import_module(f"{__name__}.{name}")
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    value = getattr(importlib.import_module(f"{__name__}.{module}"), name)
    globals()[name] = value  # resolve once
    return value


def __dir__() -> List[str]:
    return sorted(set(globals()) | set(__all__))
