"""Model-agnostic core — same flat namespace as ``scaling.core`` (reference core/__init__.py:1-71)."""
from .config import BaseConfig, overwrite_recursive  # noqa: F401
from .context import BaseContext, BaseContextGeneric, DeterminedBaseContext  # noqa: F401
from .data import (  # noqa: F401
    BaseBlendedDataset,
    BaseDataset,
    BaseDatasetBatch,
    BaseDatasetItem,
    BaseLayerIO,
    BlendedDatasetConfig,
    DataLoader,
    FileDataset,
    MemoryMapDataset,
    MemoryMapDatasetBuilder,
    broadcast_data,
)
from .logging import LoggerConfig, logger  # noqa: F401
from .nn import (  # noqa: F401
    ActivationFunction,
    BaseLayer,
    ColumnParallelLinear,
    CoreParameterMeta,
    LayerNorm,
    LayerNormConfig,
    LayerNormOptimizationType,
    LayerSpec,
    LoRaConfig,
    LoRAModuleType,
    MaskedSoftmax,
    MaskedSoftmaxConfig,
    MaskedSoftmaxKernel,
    NormType,
    ParallelLoRa,
    ParallelMLP,
    ParallelModule,
    ParallelSelfAttention,
    ParallelSwiGLUMLP,
    PipelineScheduleInference,
    PipelineScheduleTrain,
    PipePartitionCoordinates,
    RelativePositionEmbeddingType,
    RMSNorm,
    RotaryConfig,
    RotaryEmbedding,
    RotaryEmbeddingComplex,
    RowParallelLinear,
    TiedLayerSpec,
    VocabParallelEmbedding,
    get_activation_function,
    get_norm,
    pipe_partition_uniform,
)
from .optimizer import (  # noqa: F401
    BaseOptimizer,
    LearningRateDecayStyle,
    LearningRateScheduler,
    LearningRateSchedulerConfig,
    LossScaler,
    LossScalerConfig,
    Optimizer,
    OptimizerConfig,
    OptimizerParamGroup,
    OptimizerParamGroupConfig,
)
from .profiler import Profiler, ProfilerConfig, SynchronizedTimer  # noqa: F401
from .runner import LaunchConfig, RunnerConfig, RunnerDockerConfig, RunnerType, runner_main  # noqa: F401
from .topology import ActivationCheckpointingType, PipePartitionMethod, Topology, TopologyConfig  # noqa: F401
from .trainer import BaseTrainer, DeterminedBaseTrainer, TrainerConfig  # noqa: F401
