"""Transformer configuration tree (reference: transformer/context/config.py:1-459).  Field names are the YAML contract
and are identical to the reference's, so its config files load unchanged."""
from __future__ import annotations

from copy import deepcopy
from enum import Enum
from pathlib import Path
from typing import Any, List, Mapping, Optional

import torch
from pydantic import Field, model_validator

from ..core import (
    BaseConfig,
    BlendedDatasetConfig,
    LayerNormConfig,
    LearningRateSchedulerConfig,
    LoggerConfig,
    LoRaConfig,
    MaskedSoftmaxConfig,
    NormType,
    OptimizerConfig,
    ProfilerConfig,
    RelativePositionEmbeddingType,
    RunnerConfig,
    TopologyConfig,
    TrainerConfig,
)

_VERSION_DOC = ("reported as `model_version`; should increase alpha-numerically with newer training states — the "
                "largest version across base model and adapters wins, hence the leading full stop of the default")


class Precision(Enum):
    FLOAT16 = "float16"
    BFLOAT16 = "bfloat16"
    FLOAT32 = "float32"

    @property
    def dtype(self) -> torch.dtype:
        return {"float16": torch.float16, "bfloat16": torch.bfloat16, "float32": torch.float32}[self.value]


class MLPType(Enum):
    DEFAULT = "default"
    SWIGLU = "swiglu"


class TrainingConfig(BaseConfig, populate_by_name=True):
    weight_decay: float = Field(0.0001, description="decoupled weight decay")
    finetune: bool = Field(False, description="train only `finetunable_parameters`")
    finetunable_parameters: List[str] = Field([], description="regexes of parameter names that are trained when finetuning")
    parameters_exclude: List[str] = Field([], description="regexes of parameter names that are never trained")
    use_separate_lr_on_embeddings: bool = Field(False, description="own learning-rate schedule for embedding.weight",
                                                alias="use_seperate_lr_on_embeddings")
    use_deterministic_torch_algorithms: bool = Field(False, description="torch.use_deterministic_algorithms(True)")

    @model_validator(mode="after")
    def _check_finetune(self) -> "TrainingConfig":
        if self.finetune and not self.finetunable_parameters:
            raise ValueError("Can not set finetune when finetunable_parameters is empty")
        if not self.finetune and self.finetunable_parameters:
            raise ValueError("Can not set finetunable_parameters when finetune is False")
        return self


class BitfitBiasConfig(BaseConfig):
    name: str = Field(description="suffix of the bias parameters (`bias_<name>`)")
    version: str = Field(".unknown.", description=_VERSION_DOC)


class SoftpromptConfig(BaseConfig):
    name: str = Field(description="suffix of the parameter (`softprompt_<name>`)")
    n_tokens: int = Field(description="number of learned prefix positions")
    version: str = Field(".unknown.", description=_VERSION_DOC)


class AdapterConfig(BaseConfig):
    name: str = Field(description="suffix of the adapter modules (`attn_adapter_<name>`, `mlp_adapter_<name>`)")
    attention_downsampling_factor: Optional[float] = Field(None, description="bottleneck width / hidden size after attention")
    mlp_downsampling_factor: Optional[float] = Field(None, description="bottleneck width / hidden size after the mlp")
    init_std: float = Field(1.0e-5, description="std of the normal init")
    version: str = Field(".unknown.", description=_VERSION_DOC)


class EmbeddingHeadConfig(BaseConfig):
    name: str = Field(description="suffix of the projection parameters")
    proj_layers: List[int] = Field(description="output widths of the projection stack")


class TransformerArchitectureConfig(BaseConfig):
    """Architecture of the transformer (constant over a training run)."""

    vocab_size: int = Field(0, description="vocabulary size (matches the tokenizer)")
    vocab_file: Optional[Path] = Field(None, description="tokenizer file, copied into checkpoints as vocab.json")
    hidden_size: int = Field(0, description="model width")
    num_layers: int = Field(0, description="number of transformer layers")
    num_attention_heads: int = Field(0, description="number of query heads")
    num_local_attention_heads: int = Field(0, description="heads that use sliding-window attention")
    local_attention_window_size: Optional[int] = Field(None, description="sliding-window size")
    rotary_embedding_base: int = Field(10000, description="rotary frequency base")
    rotary_percentage: float = Field(1.0, description="fraction of each head's features that is rotated")
    sequence_length: int = Field(2048, description="training sequence length (also the rotary table length)")
    norm_type: NormType = Field(NormType.LAYERNORM, description="'layernorm' or 'rms'")
    relative_position_embedding_type: RelativePositionEmbeddingType = Field(
        RelativePositionEmbeddingType.ROTARY, description="'none', 'rotary' or 'rotary_complex'")
    mlp_type: MLPType = Field(MLPType.DEFAULT, description="'default' or 'swiglu'")
    mlp_factor: float = Field(4.0, description="mlp width / hidden size")
    attention_bias: bool = Field(True, description="biases in the attention projections")
    attention_qkv_in_one: bool = Field(True, description="one fused qkv parameter (requires kv heads == heads)")
    attention_num_kv_heads: Optional[int] = Field(None, description="number of key/value heads (GQA / MQA)")
    attention_use_matmul: bool = Field(False, description="kept for config compatibility")
    mlp_bias: bool = Field(True, description="biases in the mlp")
    key_query_norm: bool = Field(False, description="normalise queries and keys per head")
    weight_tying: bool = Field(True, description="share embedding and LM-head weights")
    masked_softmax: MaskedSoftmaxConfig = Field(MaskedSoftmaxConfig(), description="attention kernel selection")
    layernorm: LayerNormConfig = Field(LayerNormConfig(), description="norm settings")
    precision: Precision = Field(Precision.FLOAT32, description="parameter / activation dtype")
    dropout_embedding: float = Field(0.0, description="dropout after the embedding", ge=0.0, le=1.0)
    dropout_attention_probs: float = Field(0.0, description="dropout on attention probabilities", ge=0.0, le=1.0)
    dropout_after_attention: float = Field(0.0, description="dropout after the attention block", ge=0.0, le=1.0)
    dropout_after_mlp: float = Field(0.0, description="dropout after the mlp block", ge=0.0, le=1.0)
    bitfit_bias_config: Optional[BitfitBiasConfig] = Field(None, description="bitfit finetuning")
    finetunable_token_ids: List[int] = Field([], description="embedding rows that stay trainable when finetuning")
    image_encoder: bool = Field(False, description="add the CLIP image encoder to the input embedding")
    dropout_image_encoder: float = Field(0.0, description="dropout after the image projection", ge=0.0, le=1.0)
    softprompt_config: Optional[SoftpromptConfig] = Field(None, description="soft-prompt finetuning")
    adapter_config: Optional[AdapterConfig] = Field(None, description="adapter finetuning")
    lora_config: Optional[LoRaConfig] = Field(None, description="LoRA finetuning")
    embedding_head_config: Optional[EmbeddingHeadConfig] = Field(None, description="pooled embedding head")
    causal: bool = Field(True, description="causal attention")


class DataConfig(BaseConfig):
    """Dataset selection and packing options."""

    legacy_dataset: bool = Field(False, description="Megatron indexed-dataset reader")
    load_mmap_index_to_memory: bool = Field(False, description="hold the document index in RAM")
    use_mmap: bool = Field(True, description="memory-map the token file instead of explicit reads")
    load_data_item_mmap_index_to_memory: bool = Field(False, description="hold the packed-window index in RAM")
    finetuning_dataset: bool = Field(False, description="prompt/completion finetuning dataset")
    finetuning_chat_dataset: bool = Field(False, description="chat finetuning dataset")
    finetuning_dataset_memory_map: bool = Field(False, description="finetuning data is a memory map")
    data_prefixes: Optional[List[Path]] = Field(None, description="training data prefixes")
    validation_data_prefixes: Optional[List[Path]] = Field(None, description="validation data prefixes")
    blended_dataset: BlendedDatasetConfig = Field(BlendedDatasetConfig(), description="blending of several datasets")
    only_full_sequences: bool = Field(False, description="start every window at a document start")
    allow_incomplete_sequences_every_n: int = Field(0, description="with only_full_sequences: relax every n-th window")


class TransformerConfig(BaseConfig):
    version: str = Field(".unknown.", description=_VERSION_DOC)
    runner: RunnerConfig = Field(RunnerConfig(), description="job launch")
    logger: LoggerConfig = Field(LoggerConfig(), description="logging sinks")
    topology: TopologyConfig = Field(
        TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=1, micro_batch_size=2,
                       gradient_accumulation_steps=1), description="parallel layout and batch sizes")
    optimizer: OptimizerConfig = Field(OptimizerConfig(), description="optimizer")
    learning_rate_scheduler: LearningRateSchedulerConfig = Field(LearningRateSchedulerConfig(), description="schedule")
    embedding_learning_rate_scheduler: LearningRateSchedulerConfig = Field(LearningRateSchedulerConfig(), description="schedule of the embedding group")
    training: TrainingConfig = Field(TrainingConfig(), description="what is trained")
    trainer: TrainerConfig = Field(TrainerConfig(), description="loop and checkpoints")
    profiler: ProfilerConfig = Field(ProfilerConfig(), description="instruction profiler")
    transformer_architecture: TransformerArchitectureConfig = Field(TransformerArchitectureConfig(), description="model")
    data: DataConfig = Field(DataConfig(), description="data")
    determined_experiment_id: Optional[int] = Field(None, description="determined experiment id")
    determined_trial_id: Optional[int] = Field(None, description="determined trial id")

    @classmethod
    def from_dict(cls, d: Mapping[str, Any], overwrite_values: Optional[dict] = None) -> "TransformerConfig":
        """Finetuning parameter families (bitfit biases, adapters, softprompts) automatically get their own
        checkpoint files."""
        d = dict(deepcopy(d))
        arch = d.get("transformer_architecture") or {}
        separate = set()
        for key, prefix in (("bitfit_bias_config", "bias"), ("adapter_config", "adapter"), ("softprompt_config", "softprompt")):
            if arch.get(key) is not None:
                separate.add(f"{prefix}_{arch[key]['name']}")
        if separate:
            d.setdefault("trainer", {})
            d["trainer"] = dict(d["trainer"], separate_file_for_parameters=sorted(separate))
        return super().from_dict(d, overwrite_values=overwrite_values)
