"""Configuration of the MLP example (reference: examples/mlp_example/config.py)."""
from pydantic import Field

from scaling_b200.core import (
    BaseConfig,
    LearningRateSchedulerConfig,
    OptimizerConfig,
    ProfilerConfig,
    RunnerConfig,
    TopologyConfig,
    TrainerConfig,
)
from scaling_b200.core.logging import LoggerConfig


class MLPArchitectureConfig(BaseConfig):
    n_hidden_layers: int = Field(0, ge=0, description="number of hidden column/row-parallel layer pairs between the input and output layers")
    hidden_dim: int = Field(64, gt=0, description="width of every hidden layer")


class TrainingConfig(BaseConfig):
    weight_decay: float = Field(0.0001, description="AdamW weight decay of the weight matrices (biases are not decayed)")


class DataConfig(BaseConfig):
    root: str = Field("./.data", description="directory with MNIST in torchvision's layout; a synthetic digit set is used when it is absent")
    synthetic_size: int = Field(4096, gt=0, description="number of synthetic samples when MNIST is not on disk")


class MLPConfig(BaseConfig):
    runner: RunnerConfig = Field(RunnerConfig(), description="how the job is launched")
    logger: LoggerConfig = Field(LoggerConfig(), description="logging sinks")
    topology: TopologyConfig = Field(
        TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=1, micro_batch_size=2,
                       gradient_accumulation_steps=1),
        description="parallel layout and batch sizes")
    optimizer: OptimizerConfig = Field(OptimizerConfig(), description="AdamW / ZeRO / loss scaling")
    learning_rate_scheduler: LearningRateSchedulerConfig = Field(LearningRateSchedulerConfig(), description="warmup and decay")
    training: TrainingConfig = Field(TrainingConfig(), description="regularisation")
    trainer: TrainerConfig = Field(TrainerConfig(), description="iterations, checkpoints, evaluation")
    profiler: ProfilerConfig = Field(ProfilerConfig(), description="per-instruction timers")
    architecture: MLPArchitectureConfig = Field(MLPArchitectureConfig(), description="the network")
    data: DataConfig = Field(DataConfig(), description="where the digits come from")
