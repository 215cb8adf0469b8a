"""A complete tiny training package for engine tests, independent of the transformer suite (modelled on the role of
the reference's tests/core/minimal/): constant synthetic data, embedding → column linear → row linear → norm → head
(optionally tied to the embedding), MSE-free cross-entropy loss."""
from __future__ import annotations

from pathlib import Path
from typing import Any, List, Optional

import torch
from pydantic import Field

from scaling_b200.core import (
    BaseConfig,
    BaseContext,
    BaseDataset,
    BaseDatasetBatch,
    BaseDatasetItem,
    BaseLayer,
    BaseLayerIO,
    BaseTrainer,
    ColumnParallelLinear,
    LayerNorm,
    LayerNormConfig,
    LayerSpec,
    LearningRateSchedulerConfig,
    Optimizer,
    OptimizerConfig,
    OptimizerParamGroup,
    OptimizerParamGroupConfig,
    ParallelModule,
    ProfilerConfig,
    RowParallelLinear,
    TiedLayerSpec,
    Topology,
    TopologyConfig,
    TrainerConfig,
    VocabParallelEmbedding,
    broadcast_data,
)
from scaling_b200.core.nn.tp_ops import all_concat

VOCAB, HIDDEN, SEQ = 32, 16, 8


class MinimalConfig(BaseConfig):
    topology: TopologyConfig
    optimizer: OptimizerConfig = Field(OptimizerConfig())
    learning_rate_scheduler: LearningRateSchedulerConfig = Field(LearningRateSchedulerConfig(learning_rate=0.01, learning_rate_decay_style="constant"))
    trainer: TrainerConfig = Field(TrainerConfig())
    profiler: ProfilerConfig = Field(ProfilerConfig())
    weight_tying: bool = Field(False)
    precision: str = Field("float32")
    n_hidden_layers: int = Field(2)


class MinimalContext(BaseContext[MinimalConfig]):
    pass


class Item(BaseDatasetItem):
    def __init__(self, tokens: torch.Tensor):
        self.tokens = tokens


class Batch(BaseDatasetBatch):
    def __init__(self, inputs: Optional[torch.Tensor] = None, targets: Optional[torch.Tensor] = None):
        self.inputs = inputs
        self.targets = targets

    def only_inputs(self) -> "Batch":
        return Batch(inputs=self.inputs)

    def only_targets(self) -> "Batch":
        return Batch(targets=self.targets)


class MinimalDataset(BaseDataset[Item, Batch, Batch]):
    def __init__(self, seed: int = 0, n: int = 256):
        self.n = n
        super().__init__(seed=seed)

    def ident(self) -> str:
        return "minimal"

    def __len__(self) -> int:
        return self.n

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        self.seed = seed
        g = torch.Generator().manual_seed(seed)
        self.perm = torch.randperm(self.n, generator=g) if shuffle else torch.arange(self.n)

    def __getitem__(self, index: int) -> Item:
        i = int(self.perm[index])
        g = torch.Generator().manual_seed(1000 + i)
        start = int(torch.randint(0, VOCAB, (1,), generator=g))
        return Item((start + 3 * torch.arange(SEQ + 1)) % VOCAB)  # learnable: next = (token + 3) mod V

    def collate(self, batch: List[Item]) -> Batch:
        t = torch.stack([b.tokens for b in batch])
        return Batch(inputs=t[:, :-1].contiguous(), targets=t[:, 1:].contiguous())

    @staticmethod
    def sync_batch_to_model_parallel(topology: Topology, batch: Optional[Batch]) -> Batch:
        tensors = None if batch is None else [batch.inputs, batch.targets]
        inputs, targets = broadcast_data(tensors, torch.long, topology, n=2)
        return Batch(inputs=inputs, targets=targets)


class IO(BaseLayerIO):
    def __init__(self, activations: torch.Tensor):
        self.activations = activations


class _Proto(BaseLayer):
    @staticmethod
    def input_to_tuple(input: Any):
        return (input.activations,)

    @staticmethod
    def tuple_to_input(d):
        return IO(d[0])

    @staticmethod
    def output_to_tuple(output: Any):
        return (output.activations,)

    @staticmethod
    def tuple_to_last_stage_activation(d):
        return IO(d[0])


class EmbeddingLayer(_Proto):
    def __init__(self, topology: Topology, dtype: torch.dtype):
        super().__init__()
        self.embedding = VocabParallelEmbedding(VOCAB, HIDDEN, topology=topology, dtype=dtype)

    def forward(self, x: Any) -> IO:
        ids = x.inputs if isinstance(x, Batch) else x.activations
        return IO(self.embedding(ids))

    @staticmethod
    def input_to_tuple(input: Any):
        return (input.inputs,) if isinstance(input, Batch) else (input.activations,)

    @staticmethod
    def tuple_to_input(d):
        return IO(d[0])


class HiddenLayer(_Proto):
    def __init__(self, topology: Topology, dtype: torch.dtype):
        super().__init__()
        self.up = ColumnParallelLinear(HIDDEN, 4 * HIDDEN, topology=topology, dtype=dtype, parallel_output=True)
        self.down = RowParallelLinear(4 * HIDDEN, HIDDEN, topology=topology, dtype=dtype, parallel_input=True)
        self.norm = LayerNorm(LayerNormConfig(), HIDDEN, dtype=dtype, topology=topology)
        self.dropout = torch.nn.Dropout(0.1)
        self.topology = topology
        self.register_buffer("buffer_for_testing", torch.ones(3, dtype=dtype))

    def forward(self, x: IO) -> IO:
        h = self.down(torch.nn.functional.gelu(self.up(self.norm(x.activations))))
        with self.topology.model_parallel_constant_rng():
            h = self.dropout(h)
        return IO(x.activations + h)


class HeadLayer(_Proto):
    def __init__(self, topology: Topology, dtype: torch.dtype):
        super().__init__()
        self.linear = ColumnParallelLinear(HIDDEN, VOCAB, bias=False, topology=topology, dtype=dtype)

    def forward(self, x: IO) -> IO:
        return IO(self.linear(x.activations))


class TiedHeadLayer(_Proto):
    def __init__(self, topology: Topology, dtype: torch.dtype):
        super().__init__()
        self.embedding = VocabParallelEmbedding(VOCAB, HIDDEN, topology=topology, dtype=dtype)
        self.topology = topology

    def forward(self, x: IO) -> IO:
        from scaling_b200.core.nn.tp_ops import copy_to_tensor_model_parallel_region

        h = copy_to_tensor_model_parallel_region(x.activations, self.topology)
        logits = torch.nn.functional.linear(h, self.embedding.weight)
        return IO(all_concat(logits, dim=-1, topology=self.topology))


def loss_function(output: IO, batch: Batch):
    logits = output.activations.float()
    loss = torch.nn.functional.cross_entropy(logits.reshape(-1, VOCAB), batch.targets.reshape(-1))
    acc = (logits.argmax(-1) == batch.targets).float().mean()
    return loss, {"accuracy": acc}


def metrics_aggregation_fn(topology: Topology, metrics: List[dict]) -> dict:
    import torch.distributed as dist

    keys = metrics[0].keys()
    out = {k: torch.stack([m[k] for m in metrics]).mean() for k in keys}
    if topology.config.data_parallel_size > 1 and topology.is_distributed_initialized:
        for k in out:
            dist.all_reduce(out[k], group=topology.data_parallel_group)
            out[k] = out[k] / topology.config.data_parallel_size
    return out


def layer_specs(config: MinimalConfig, topology: Topology) -> List[LayerSpec]:
    dtype = {"float32": torch.float32, "bfloat16": torch.bfloat16, "float16": torch.float16}[config.precision]
    specs: List[LayerSpec] = []
    if config.weight_tying:
        specs.append(TiedLayerSpec(module_class=EmbeddingLayer, key="emb", tied_weight_attributes=["embedding.weight"],
                                   topology=topology, dtype=dtype))
    else:
        specs.append(LayerSpec(EmbeddingLayer, topology=topology, dtype=dtype))
    for _ in range(config.n_hidden_layers):
        specs.append(LayerSpec(HiddenLayer, topology=topology, dtype=dtype))
    if config.weight_tying:
        specs.append(TiedLayerSpec(module_class=TiedHeadLayer, key="emb", tied_weight_attributes=["embedding.weight"],
                                   topology=topology, dtype=dtype))
    else:
        specs.append(LayerSpec(HeadLayer, topology=topology, dtype=dtype))
    return specs


def run_minimal(rank: int, world_size: int, master_port: int, config_dict: dict, return_state: bool = False):
    """Worker body: build everything from ``config_dict`` and train; returns the list of losses (rank 0 / first stage)."""
    config_dict = dict(config_dict)
    config_dict["topology"] = dict(config_dict["topology"], world_size=world_size, global_rank=rank, local_slot=rank)
    config = MinimalConfig.from_dict(config_dict)
    topology = Topology(config.topology)
    context = MinimalContext(config=config, topology=topology)
    context.initialize(master_addr="127.0.0.1", master_port=str(master_port), seed=config.trainer.seed)
    model = ParallelModule(layer_specs(config, topology), topology=topology, profiler_config=config.profiler)
    named = model.named_parameters_with_meta()
    groups = [
        OptimizerParamGroup([x for x in named if not x[0].endswith("bias")],
                            OptimizerParamGroupConfig(name="wd", weight_decay=0.01, learning_rate_scheduler=config.learning_rate_scheduler)),
        OptimizerParamGroup([x for x in named if x[0].endswith("bias")],
                            OptimizerParamGroupConfig(name="nowd", weight_decay=0.0, learning_rate_scheduler=config.learning_rate_scheduler)),
    ]
    optimizer = Optimizer(config.optimizer, groups, topology)
    dataset = MinimalDataset(seed=config.trainer.seed) if topology.is_io_rank else None
    trainer = BaseTrainer(config=config.trainer, context=context, parallel_module=model, optimizer=optimizer,
                          dataset=dataset, sync_batch_to_model_parallel=MinimalDataset.sync_batch_to_model_parallel,
                          loss_function=loss_function, metrics_aggregation_fn=metrics_aggregation_fn)
    metrics = trainer.run_training(return_metrics=True)
    losses = [m["training/loss"] for m in (metrics or [])]
    if return_state:
        return {"losses": losses, "params_count": (trainer.parameters_total, trainer.parameters_unique)}
    return losses
