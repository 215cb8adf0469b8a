from .activation import ActivationFunction, get_activation_function  # noqa: F401
from .attention import (  # noqa: F401
    ParallelSelfAttention,
    RelativePositionEmbeddingType,
    cumulative_seq_lengths_to_dense_attention_mask,
    get_max_seq_length,
    repeat_kv,
)
from .engine import Buffers, BufferType, EvaluationStepOutput, ParallelModule, TrainStepOutput  # noqa: F401
from .inference_module import HiddenStateRecorder, InferenceModule, RecorderSetting  # noqa: F401
from .layers import BaseLayer, LayerSpec, TiedLayerSpec  # noqa: F401
from .linear import ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding  # noqa: F401
from .lora import LoRaConfig, LoRAModuleType, ParallelLoRa  # noqa: F401
from .mlp import ParallelMLP, ParallelSwiGLUMLP  # noqa: F401
from .norms import LayerNorm, LayerNormConfig, LayerNormOptimizationType, NormType, RMSNorm, get_norm  # noqa: F401
from .p2p import CommunicationMetaBase, GradOutput, ModelParallelCommunicator, PipeCommunicator  # noqa: F401
from .parameter_meta import CoreParameterMeta  # noqa: F401
from .partition import (  # noqa: F401
    PipePartitionCoordinates,
    pipe_partition_balanced,
    pipe_partition_from_indices,
    pipe_partition_uniform,
)
from .partitioned_module import PipePartitionedModule, TiedLayerIndex, key_match  # noqa: F401
from .rotary import RotaryConfig, RotaryEmbedding, RotaryEmbeddingComplex  # noqa: F401
from .schedule import (  # noqa: F401
    InstructionBackwardPass,
    InstructionBase,
    InstructionForwardPass,
    InstructionLoadMicroBatch,
    InstructionLoss,
    InstructionOptimizerStep,
    InstructionRecvActivation,
    InstructionRecvGrad,
    InstructionReduceTiedGrads,
    InstructionSendActivation,
    InstructionSendGrad,
    InstructionStoreMicroBatch,
    PipelineScheduleInference,
    PipelineScheduleTrain,
    SimulationEngine,
)
from .softmax import MaskedSoftmax, MaskedSoftmaxConfig, MaskedSoftmaxKernel  # noqa: F401
from .tp_ops import (  # noqa: F401
    all_concat,
    all_reduce,
    all_reduce_scatter_to_sequence_parallel,
    all_shard,
    copy_to_tensor_model_parallel_region,
    gather_from_sequence_parallel_region,
)
