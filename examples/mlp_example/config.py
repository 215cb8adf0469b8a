"""What the MLP example can be told: the core blocks (launch, logging, layout, optimiser, schedule, trainer loop,
profiler) plus two blocks of its own — the network and where the digits come from."""
from pydantic import Field

from scaling_b200 import core as C
from scaling_b200.core.logging import LoggerConfig


class MLPArchitectureConfig(C.BaseConfig):
    hidden_dim: int = Field(64, gt=0, description="width of every hidden layer")
    n_hidden_layers: int = Field(0, ge=0, description="extra row/column-parallel layer pairs between input and output layer")
    weight_decay: float = Field(1e-4, ge=0.0, description="AdamW weight decay of the weight matrices; biases are never decayed")


class DataConfig(C.BaseConfig):
    root: str = Field("./.data", description="MNIST in torchvision's on-disk layout; when absent a synthetic digit set is used (nothing is downloaded)")
    synthetic_size: int = Field(4096, gt=0, description="number of synthetic training samples")


def _small_topology() -> C.TopologyConfig:
    return C.TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=1, micro_batch_size=2,
                            gradient_accumulation_steps=1)


class MLPConfig(C.BaseConfig):
    architecture: MLPArchitectureConfig = Field(MLPArchitectureConfig(), description="the network")
    data: DataConfig = Field(DataConfig(), description="where the digits come from")
    topology: C.TopologyConfig = Field(_small_topology(), description="parallel layout and batch sizes")
    optimizer: C.OptimizerConfig = Field(C.OptimizerConfig(), description="AdamW, ZeRO, loss scaling, clipping")
    learning_rate_scheduler: C.LearningRateSchedulerConfig = Field(C.LearningRateSchedulerConfig(), description="warm-up and decay")
    trainer: C.TrainerConfig = Field(C.TrainerConfig(), description="iterations, checkpoints, evaluation")
    profiler: C.ProfilerConfig = Field(C.ProfilerConfig(), description="per-instruction timers")
    logger: LoggerConfig = Field(LoggerConfig(), description="logging sinks")
    runner: C.RunnerConfig = Field(C.RunnerConfig(), description="how the job is launched")
