from .configs import (  # noqa: F401
    LearningRateDecayStyle,
    LearningRateSchedulerConfig,
    LossScalerConfig,
    OptimizerConfig,
    OptimizerParamGroupConfig,
)
from .loss_scaler import LossScaler, LossScalerOutput  # noqa: F401
from .lr import LearningRateScheduler  # noqa: F401
from .optimizer import (  # noqa: F401
    BaseOptimizer,
    Optimizer,
    OptimizerStepOutput,
    allreduce_no_retain,
    allreduce_tensor_in_float32,
)
from .param_group import (  # noqa: F401
    NCCL_START_ALIGNMENT_FACTOR,
    OptimizerParamGroup,
    flatten_dense_tensors_aligned,
    get_data_parallel_partitions,
)
