"""Model assembly, loss and optimizer groups of the transformer suite (reference: transformer/model/model.py:1-408).

Layer list (its positions are the checkpoint layer indices): ``0 EmbeddingInput``, ``1..L TransformerLayer``,
``L+1 LayerNormWrapper``, ``L+2 TransformerLMHead | TransformerLMHeadTied`` [, ``L+3 TransformerEmbeddingHead``].
"""
from __future__ import annotations

import math
import re
from typing import Any, Dict, List, Optional, Sequence, Set, Tuple, Union

import torch
import torch.distributed as dist

from ...core import (
    BaseOptimizer,
    CoreParameterMeta,
    LayerSpec,
    LearningRateSchedulerConfig,
    Optimizer,
    OptimizerConfig,
    OptimizerParamGroup,
    OptimizerParamGroupConfig,
    ParallelModule,
    TiedLayerSpec,
    Topology,
)
from ...core.logging import logger
from ...ops import functional as OF
from ..config import TrainingConfig, TransformerArchitectureConfig
from ..context import TransformerContext
from ..data.text_dataset_batch import TextDatasetBatch
from .layers import (
    EmbeddingInput,
    LayerNormWrapper,
    TransformerEmbeddingHead,
    TransformerLayer,
    TransformerLayerIO,
    TransformerLMHead,
    TransformerLMHeadTied,
)

NamedParameterMeta = Tuple[str, torch.Tensor, CoreParameterMeta]


def loss_function(output: TransformerLayerIO, batch: TextDatasetBatch) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
    """Loss-weighted mean token cross-entropy (fp32) and masked accuracy.

    When the LM head left the logits sharded over the vocabulary (training on a B200) the fused vocab-parallel kernel
    is used: per-token (max, Σexp, target logit, arg-max) are combined across the TP group — 16 bytes per token —
    instead of all-gathering [tokens, V] logits and copying them to fp32 (reference model.py:56-67)."""
    assert batch.target_token_ids is not None, "target_token_ids not set in batch; revisit Batch.only_targets()"
    assert batch.loss_weights is not None, "loss_weights not set in batch; revisit Batch.only_targets()"
    weights = batch.loss_weights.float().reshape(-1)
    targets = batch.target_token_ids.reshape(-1)
    logits = output.activations.reshape(-1, output.activations.shape[-1])
    shard = getattr(output, "vocab_parallel_logits", None)
    if shard is not None:
        partial, comm = (shard[2], shard[3]) if len(shard) > 2 else (None, None)
        losses, argmax = OF.vocab_parallel_cross_entropy(logits, targets, vocab_start=shard[0], group=shard[1],
                                                         partial=partial, comm=comm)
    else:
        lf = logits.float()
        losses = torch.nn.functional.cross_entropy(lf, targets, reduction="none")
        argmax = lf.argmax(-1)
    loss = (losses * weights).sum() / weights.sum()
    mask = (weights > 0).float()
    accuracy = ((argmax == targets).float() * mask).sum() / mask.sum()
    return loss, {"accuracy": accuracy.detach()}


def metrics_aggregation_fn(topology: Topology, metrics: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
    """Mean over micro batches, then over the data-parallel group (stays on the device)."""
    out = {k: torch.stack([m[k].detach().float() for m in metrics]).mean() for k in metrics[0]}
    if topology.config.data_parallel_size > 1 and topology.is_distributed_initialized:
        for k in out:
            dist.all_reduce(out[k], group=topology.data_parallel_group)
            out[k] = out[k] / topology.config.data_parallel_size
    return out


class TransformerParallelModule(ParallelModule[TransformerLayerIO, TextDatasetBatch]):
    """The engine, typed for the transformer IO.  The data-dependent ``cumulative_seq_lengths`` never travels between
    stages (only its padded, static-shape copy does) — handled by ``TransformerLayerBaseIO.output_to_tuple`` /
    ``tuple_to_input`` rather than by overriding engine internals."""


def get_transformer_layer_specs(architecture_config: TransformerArchitectureConfig, topology: Optional[Topology] = None) -> List[LayerSpec]:
    a = architecture_config
    std = math.sqrt(2 / (a.hidden_size * 5))

    def init_method(x: torch.Tensor) -> torch.Tensor:
        return torch.nn.init.normal_(x, mean=0.0, std=std)

    common = dict(architecture_config=a, topology=topology)
    specs: List[LayerSpec] = []
    tied = ["embedding.weight"]
    if a.weight_tying:
        specs.append(TiedLayerSpec(module_class=EmbeddingInput, key="embedding_tying", tied_weight_attributes=tied,
                                   init_method=init_method, **common))
    else:
        specs.append(LayerSpec(module_class=EmbeddingInput, init_method=init_method, **common))
    for i in range(a.num_layers):
        specs.append(LayerSpec(TransformerLayer, layer_index=i, init_method=init_method, **common))
    specs.append(LayerSpec(LayerNormWrapper, layer_index=a.num_layers, **common))
    if a.weight_tying:
        specs.append(TiedLayerSpec(module_class=TransformerLMHeadTied, key="embedding_tying", tied_weight_attributes=tied,
                                   init_method=init_method, **common))
    else:
        specs.append(LayerSpec(module_class=TransformerLMHead, init_method=init_method, **common))
    if a.embedding_head_config is not None:
        specs.append(LayerSpec(module_class=TransformerEmbeddingHead, **common))
    return specs


def init_model(context: TransformerContext, use_continuous_recommunication: bool = False) -> TransformerParallelModule:
    specs = get_transformer_layer_specs(context.config.transformer_architecture, topology=context.topology)
    return TransformerParallelModule(layer_specs=specs, topology=context.topology,
                                     profiler_config=context.config.profiler,
                                     use_continuous_recommunication=use_continuous_recommunication)


# ---- optimizer groups ------------------------------------------------------------------------------------------
def _first_match(name: str, patterns: Sequence[str]) -> Optional[str]:
    return next((p for p in patterns if re.search(p, name) is not None), None)


def _find_matching_param(key: NamedParameterMeta, data: Sequence[str]) -> Optional[str]:
    """First pattern of ``data`` that matches the parameter's name (reference model.py:385-386)."""
    return _first_match(key[0], data)


def _filter_by_param(parameters_exclude: Sequence[str], parameter_list: List[NamedParameterMeta]) -> List[NamedParameterMeta]:
    return [p for p in parameter_list if _find_matching_param(p, parameters_exclude) is None]


def _find_global_unmatched_parameters(found: Set[str], finetunable_parameters: Sequence[str]) -> Set[str]:
    """Patterns that matched no parameter on ANY rank (a pipeline stage only sees its own layers)."""
    missing = set(finetunable_parameters) - found
    if dist.is_available() and dist.is_initialized():
        gathered: List[Any] = [None] * dist.get_world_size()
        dist.all_gather_object(gathered, missing)
        missing = set.intersection(*gathered)
    return missing


def _extract_parameters(training: TrainingConfig, named: List[NamedParameterMeta]):
    """Split into (embedding group, no-weight-decay group, weight-decay group) — `.bias` suffix ⇒ no decay (so bitfit
    biases `bias_<n>` do get decay, like the reference model.py:343-345); finetuning keeps only matching names and a
    pattern that matches nothing on any rank is an error (reference model.py:329-364)."""
    wd, no_wd, emb = [], [], []
    found: Set[str] = set()
    for item in named:
        name = item[0]
        if training.finetune:
            m = _find_matching_param(item, training.finetunable_parameters)
            if m is None:
                continue
            found.add(m)
        if name.endswith(".bias"):
            no_wd.append(item)
        elif training.use_separate_lr_on_embeddings and name == "embedding.weight":
            assert not training.finetune, "Can not use separate lr on embeddings with finetuning"
            emb.append(item)
        else:
            wd.append(item)
    if training.finetune:
        missing = _find_global_unmatched_parameters(found, training.finetunable_parameters)
        if missing:
            raise ValueError(f"Unmatched finetunable parameters: {missing}")
    if training.parameters_exclude:
        wd = _filter_by_param(training.parameters_exclude, wd)
        no_wd = _filter_by_param(training.parameters_exclude, no_wd)
        emb = _filter_by_param(training.parameters_exclude, emb)
    return emb, no_wd, wd


def get_parameter_groups(context: TransformerContext, model: TransformerParallelModule,
                         learning_rate_scheduler_config: Optional[LearningRateSchedulerConfig] = None,
                         embedding_learning_rate_scheduler_config: Optional[LearningRateSchedulerConfig] = None) -> List[OptimizerParamGroup]:
    cfg = context.config
    lr = learning_rate_scheduler_config or cfg.learning_rate_scheduler
    emb_lr = embedding_learning_rate_scheduler_config or cfg.embedding_learning_rate_scheduler
    named = model.named_parameters_with_meta()
    emb, no_wd, wd = _extract_parameters(cfg.training, named)
    topo = context.topology
    distributed = topo.is_distributed_initialized
    counts = torch.tensor([len(wd), len(no_wd), len(emb)], dtype=torch.long, device=topo.device)
    if distributed:
        dist.all_reduce(counts, op=dist.ReduceOp.MAX)  # a group must exist everywhere if it exists anywhere
    assert int(counts.sum()) > 0, "did not specify any finetuneable parameters on any rank"
    logger.info(f"training parameters: {sorted({n for n, _, _ in wd + no_wd + emb})}")
    # true freezing: parameters outside every group neither compute nor keep gradients
    trained = {id(p) for _, p, _ in wd + no_wd + emb}
    for _, p, _ in named:
        if id(p) not in trained:
            p.requires_grad_(False)
    if len(trained) < len(named):
        # frozen weights get the trained arenas' layout, so the fused q|k|v / gate|up GEMMs keep applying to them
        from ...core.optimizer.param_group import pack_frozen_parameters

        model._frozen_arenas = pack_frozen_parameters([p for _, p, _ in named])  # type: ignore[attr-defined]
    groups: List[OptimizerParamGroup] = []
    c = counts.tolist()
    if c[0] > 0:
        groups.append(OptimizerParamGroup(wd, OptimizerParamGroupConfig(
            name="weight_decay_params", weight_decay=cfg.training.weight_decay, learning_rate_scheduler=lr)))
    if c[1] > 0:
        groups.append(OptimizerParamGroup(no_wd, OptimizerParamGroupConfig(
            name="no_weight_decay_params", weight_decay=0.0, learning_rate_scheduler=lr)))
    if c[2] > 0:
        groups.append(OptimizerParamGroup(emb, OptimizerParamGroupConfig(
            name="embedding_weight_decay_params", weight_decay=cfg.training.weight_decay, learning_rate_scheduler=emb_lr)))
    return groups


def init_optimizer(context: TransformerContext, model: TransformerParallelModule,
                   optimizer_config: Optional[OptimizerConfig] = None,
                   learning_rate_scheduler_config: Optional[LearningRateSchedulerConfig] = None,
                   embedding_learning_rate_scheduler_config: Optional[LearningRateSchedulerConfig] = None) -> BaseOptimizer:
    groups = get_parameter_groups(context, model, learning_rate_scheduler_config, embedding_learning_rate_scheduler_config)
    return Optimizer(config=optimizer_config or context.config.optimizer, parameter_groups=groups, topology=context.topology)
