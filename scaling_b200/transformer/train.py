"""Training entry point of the transformer suite (reference: transformer/train.py:1-304).

``main(launch_config, overwrite_config, return_metrics)`` is what every per-GPU process runs: config → topology →
logger → context (device + torch.distributed) → model → optimizer → datasets on IO ranks → trainer.run_training().
"""
from __future__ import annotations

import os
import shutil
from pathlib import Path
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from ..core import BaseDataset, DeterminedBaseTrainer, LaunchConfig, Topology
from ..core.logging import logger
from ..core.nn.engine import EvaluationStepOutput, TrainStepOutput
from .config import DataConfig, TransformerConfig
from .context import TransformerContext
from .data.text_dataset import TextBlendedDataset, TextDataset
from .dataset_loader import load_datasets
from .model import TransformerParallelModule, init_model, init_optimizer, loss_function, metrics_aggregation_fn
from .utils.get_tflops import (
    get_model_flop_utilization_palm,
    get_tflops_aleph_alpha,
    get_tflops_bloom,
    get_tflops_electra,
    get_tflops_megatron,
)


class TransformerTrainer(DeterminedBaseTrainer[TransformerContext, TransformerParallelModule]):
    def save_checkpoint(self, save_dir: Optional[Path] = None) -> Path:
        save_dir = super().save_checkpoint(save_dir=save_dir)
        vocab = self.context.config.transformer_architecture.vocab_file
        if vocab is not None and save_dir is not None and self.context.topology.config.global_rank == 0:
            step_dir = Path(save_dir) / f"global_step{self.context.iterations}"
            if step_dir.is_dir():
                shutil.copy(vocab, step_dir / "vocab.json")
        return save_dir

    def _tflops_metrics(self, iter_time_s: float) -> Dict[str, float]:
        a, topo = self.context.config.transformer_architecture, self.context.topology
        kw = dict(iter_time_s=iter_time_s, topology=topo, transformer_architecture=a)
        return {
            "runtime/tflops_megatron": get_tflops_megatron(parameter_count=self.parameters_total, **kw),
            "runtime/tflops_megatron_layout_independent": get_tflops_megatron(parameter_count=self.parameters_unique, **kw),
            "runtime/tflops_bloom": get_tflops_bloom(**kw),
            "runtime/tflops_electra": get_tflops_electra(**kw),
            "runtime/tflops_aleph_alpha": get_tflops_aleph_alpha(**kw),
            "runtime/mfu_palm": get_model_flop_utilization_palm(parameter_count=self.parameters_unique, **kw),
            "runtime/tokens_per_second_per_gpu": topo.config.global_batch_size * a.sequence_length / iter_time_s / topo.config.world_size,
        }

    def log_metrics(self, train_step_output: TrainStepOutput, eval_step_output: Optional[EvaluationStepOutput] = None) -> Dict[str, Any]:
        logger.info(f"completed step {self.context.iterations}")
        t = train_step_output
        m: Dict[str, Any] = {"training/loss": t.loss, "runtime/step_duration": t.step_duration}
        m.update(t.debug_dict or {})
        m.update({f"training/{k}": v for k, v in (t.metrics or {}).items()})
        for key, val in (("global_grad_norm", t.global_grad_norm), ("global_grad_norm_clipped", t.global_grad_norm_clipped),
                         ("no_overflow_steps", t.no_overflow_steps), ("current_loss_scale", t.current_loss_scale)):
            if val is not None:
                m[f"training/{key}"] = val
        if t.overflow is not None:
            m["training/overflow"] = int(t.overflow)
        for name, lr in (t.learning_rates or {}).items():
            m[f"training/learning_rate_{name}"] = lr
        m.update(self._tflops_metrics(t.step_duration))
        if eval_step_output is not None:
            m.update({f"evaluation/{k}": v for k, v in (eval_step_output.metrics or {}).items()})
            m["evaluation/loss"] = eval_step_output.loss
            m["evaluation/step_duration"] = eval_step_output.step_duration
        logger.log_metrics(m, step=self.context.iterations)
        return m


def _get_sync_batch(data_config: DataConfig) -> Callable:
    if data_config.finetuning_dataset:
        from .data.finetuning import FinetuningTextDataset

        return FinetuningTextDataset.sync_batch_to_model_parallel
    if data_config.finetuning_chat_dataset:
        from .data.finetuning import FinetuningChatDataset

        return FinetuningChatDataset.sync_batch_to_model_parallel
    return TextDataset.sync_batch_to_model_parallel


def _get_dataset_type(data_config: DataConfig) -> type:
    if data_config.legacy_dataset:
        from .data.legacy import LegacyBlendedDataset

        return LegacyBlendedDataset
    if data_config.finetuning_dataset:
        from .data.finetuning import FinetuningTextBlendedDataset

        return FinetuningTextBlendedDataset
    if data_config.finetuning_chat_dataset:
        from .data.finetuning import FinetuningChatBlendedDataset

        return FinetuningChatBlendedDataset
    return TextBlendedDataset


def _read_datasets(config: TransformerConfig) -> Tuple[Optional[BaseDataset], Optional[BaseDataset]]:
    logger.info("loading dataset")
    train, val = load_datasets(config.data, config.transformer_architecture, config)
    cls = _get_dataset_type(config.data)
    seed = config.trainer.seed
    blended = cls(seed=seed, config=config.data.blended_dataset, datasets=train)
    blended_val = cls(seed=seed, config=config.data.blended_dataset, datasets=val) if config.data.validation_data_prefixes else None
    return blended, blended_val


def _init_transformer_config(launch_config: LaunchConfig, overwrite_config: Optional[Dict[str, Any]]) -> TransformerConfig:
    config_dict = launch_config.payload or overwrite_config or {}
    config_dict = launch_config.overwrite_config_dict_with_launcher_args(dict(config_dict))
    return TransformerConfig.from_dict(config_dict, overwrite_values=overwrite_config)


def build_trainer(launch_config: LaunchConfig, overwrite_config: Optional[Dict[str, Any]] = None,
                  determined_context: Any = None, determined_profiler: Any = None,
                  dataset: Optional[BaseDataset] = None) -> TransformerTrainer:
    """Everything ``main`` does short of running the loop (used by benchmarks and tests)."""
    config = _init_transformer_config(launch_config, overwrite_config)
    topology = Topology(config=config.topology)
    name = f"RANK {topology.config.global_rank}"
    if config.runner.use_determined:
        logger.configure_determined(config.logger, name=name, global_rank=topology.config.global_rank,
                                    determined_context=determined_context)
    else:
        logger.configure(config.logger, name=name, global_rank=topology.config.global_rank)
    logger.log_config(config)
    if config.training.use_deterministic_torch_algorithms:
        os.environ["CUBLAS_WORKSPACE_CONFIG"] = ":4096:8"
        torch.use_deterministic_algorithms(True)
    context = TransformerContext(config=config, topology=topology)
    if determined_context is not None:
        context.initialize_with_determined(master_addr=launch_config.master_addr, master_port=str(launch_config.master_port),
                                           determined_context=determined_context, determined_profiler=determined_profiler,
                                           seed=config.trainer.seed)
    else:
        context.initialize(master_addr=launch_config.master_addr, master_port=str(launch_config.master_port),
                           seed=config.trainer.seed)
    model = init_model(context=context)
    optimizer = init_optimizer(context=context, model=model)
    train_ds: Optional[BaseDataset] = dataset
    val_ds: Optional[BaseDataset] = None
    if topology.is_io_rank and dataset is None:
        train_ds, val_ds = _read_datasets(config)
    return TransformerTrainer(config=config.trainer, context=context, parallel_module=model, optimizer=optimizer,
                              dataset=train_ds, sync_batch_to_model_parallel=_get_sync_batch(config.data),
                              loss_function=loss_function, metrics_aggregation_fn=metrics_aggregation_fn,
                              dataset_evaluation=val_ds)


def main(launch_config: LaunchConfig, overwrite_config: Optional[Dict[str, Any]] = None, return_metrics: bool = False,
         determined_context: Any = None, determined_profiler: Any = None) -> Optional[List[Dict[str, Any]]]:
    trainer = build_trainer(launch_config, overwrite_config, determined_context, determined_profiler)
    return trainer.run_training(return_metrics=return_metrics)


if __name__ == "__main__":
    main(LaunchConfig.from_launcher_args())
