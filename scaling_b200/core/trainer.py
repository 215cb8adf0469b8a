"""Config-driven training loop with checkpointing (reference: core/trainer/trainer.py:33-558, trainer_config.py).

``BaseTrainer``: optional checkpoint load at construction, data loaders on IO ranks only (seeded with the restored
``consumed_samples``), ``run_training`` = step → save → eval → log.  Checkpoints live in
``save_dir/global_step{N}/`` with a ``latest`` pointer.  ``DeterminedBaseTrainer`` adds Determined storage-manager
checkpoints and preemption handling behind an optional import.
"""
from __future__ import annotations

import shutil
import sys
import uuid
from pathlib import Path
from typing import Any, Callable, Dict, Generic, List, Optional, TypeVar, Union

import torch
from pydantic import Field

from .config import BaseConfig
from .context import BaseContext, BaseContextGeneric, DeterminedBaseContextGeneric
from .data import BaseDataset, DataLoader
from .logging import DeterminedLogger, logger
from .nn.attention import ParallelSelfAttention
from .nn.engine import EvaluationStepOutput, ParallelModule, TrainStepOutput
from .optimizer import BaseOptimizer


class TrainerConfig(BaseConfig):
    save_dir: Optional[Path] = Field(None, description="directory for saving checkpoints")
    save_interval: Optional[int] = Field(None, description="save every N steps (needs save_dir)")
    load_dir: Optional[Path] = Field(None, description="directory to resume from")
    train_iterations: Optional[int] = Field(None, description="train for this number of iterations")
    assert_checkpoint_loaded: bool = Field(True, description="fail if load_dir yields no checkpoint")
    load_optimizer_states: bool = Field(True, description="restore optimizer state on load")
    delete_past_optimizer_states: bool = Field(True, description="keep optimizer state only in the newest checkpoint")
    load_context: bool = Field(True, description="restore iteration / consumed-sample counters and RNG on load")
    allowed_missing_keys_in_checkpoint: Optional[List[str]] = Field(None, description="regexes of parameters that may be absent from the checkpoint (e.g. new adapters)")
    allowed_unexpected_keys_in_checkpoint: Optional[List[str]] = Field(None, description="regexes of checkpoint entries that may be dropped")
    ignore_keys_in_checkpoint: Optional[List[str]] = Field(None, description="regexes of parameters that are re-initialised instead of loaded")
    merge_lora_after_loading_checkpoint: Optional[bool] = Field(False, description="fold LoRA weights into the base weights after loading")
    seed: int = Field(42, description="base seed")
    dataloader_num_workers: int = Field(0, description="worker processes of the data loader")
    dataloader_pin_memory: bool = Field(True, description="pinned host batches")
    dataloader_prefetch_factor: Optional[int] = Field(None, description="batches prefetched per worker")
    eval_iterations: int = Field(1, description="evaluation steps per evaluation")
    eval_interval: Optional[int] = Field(None, description="evaluate every N steps")
    separate_file_for_parameters: Optional[List[str]] = Field(None, description="parameter-name fragments that get their own checkpoint file per layer")


ParallelModuleGeneric = TypeVar("ParallelModuleGeneric", bound=ParallelModule)


class BaseTrainer(Generic[BaseContextGeneric, ParallelModuleGeneric]):
    def __init__(self, config: TrainerConfig, context: BaseContextGeneric, parallel_module: ParallelModuleGeneric,
                 optimizer: BaseOptimizer, dataset: Optional[BaseDataset],
                 sync_batch_to_model_parallel: Callable[..., Any], loss_function: Callable[..., Any],
                 metrics_aggregation_fn: Optional[Callable] = None, dataset_evaluation: Optional[BaseDataset] = None):
        self.config = config
        self.context = context
        self.parallel_module = parallel_module
        self.parameters_total, self.parameters_unique = parallel_module.get_params_count()
        logger.log_config_dict({"parameters_total": self.parameters_total, "parameters_unique": self.parameters_unique})
        self.optimizer = optimizer
        self.dataset = dataset
        self.dataset_evaluation = dataset_evaluation
        loaded = self.load_checkpoint(
            load_dir=config.load_dir, load_optimizer_states=config.load_optimizer_states, load_context=config.load_context,
            allowed_missing_keys_in_checkpoint=config.allowed_missing_keys_in_checkpoint,
            allowed_unexpected_keys_in_checkpoint=config.allowed_unexpected_keys_in_checkpoint,
            ignore_keys_in_checkpoint=config.ignore_keys_in_checkpoint)
        if config.assert_checkpoint_loaded:
            assert loaded, ("checkpoint could not be loaded. if this is intended you may change the parameter "
                            "'assert_checkpoint_loaded' in the TrainerConfig to False.")
        if config.merge_lora_after_loading_checkpoint:
            for module in parallel_module.modules():
                if isinstance(module, ParallelSelfAttention) and module.lora_config is not None:
                    module.merge_lora_weights()
            self.optimizer.refresh_optimizer_after_model_change()
            logger.info("Merged LoRa weights")
        # loaders are created after the checkpoint load so that consumed_samples is the restored value
        self.dataloader: Optional[DataLoader] = None
        self.dataloader_evaluation: Optional[DataLoader] = None
        topo = context.topology
        if topo.is_io_rank:
            assert dataset is not None, "io ranks need a dataset"
            kw = dict(seed=config.seed, topology=topo, num_workers=config.dataloader_num_workers,
                      pin_memory=config.dataloader_pin_memory, prefetch_factor=config.dataloader_prefetch_factor)
            self.dataloader = DataLoader(consumed_samples=context.consumed_samples, dataset=dataset, **kw)
            if dataset_evaluation is not None:
                self.dataloader_evaluation = DataLoader(consumed_samples=context.consumed_samples_evaluation,
                                                        dataset=dataset_evaluation, **kw)
        self.sync_batch_to_model_parallel = sync_batch_to_model_parallel
        self.loss_function = loss_function
        self.metrics_aggregation_fn = metrics_aggregation_fn

    # ---- checkpoints --------------------------------------------------------------------------------------
    def save_checkpoint(self, save_dir: Optional[Path] = None) -> Path:
        save_dir = Path(save_dir or self.config.save_dir)  # type: ignore[arg-type]
        step_dir = save_dir / f"global_step{self.context.iterations}"
        step_dir.mkdir(exist_ok=True, parents=True)
        self.parallel_module.save_checkpoint(step_dir, separate_file_for_parameters=self.config.separate_file_for_parameters)
        self.optimizer.save_checkpoint(step_dir)
        self.context.save_checkpoint(step_dir)
        topo = self.context.topology
        if topo.is_distributed_initialized:
            torch.distributed.barrier()
        if topo.config.global_rank == 0:
            (save_dir / "latest").write_text(f"global_step{self.context.iterations}", encoding="UTF-8")
            if self.config.delete_past_optimizer_states:
                for old in save_dir.glob("global_step*"):
                    if old.is_dir() and old != step_dir:
                        for f in old.glob("optimizer_state_*.pt"):
                            f.unlink()
        logger.info(f"saved checkpoint: {step_dir}")
        return save_dir

    def load_checkpoint(self, load_dir: Optional[Path] = None, load_optimizer_states: bool = True, load_context: bool = True,
                        allowed_missing_keys_in_checkpoint: Optional[List[str]] = None,
                        allowed_unexpected_keys_in_checkpoint: Optional[List[str]] = None,
                        ignore_keys_in_checkpoint: Optional[List[str]] = None) -> bool:
        if load_dir is None:
            return False
        load_dir = Path(load_dir)
        if (load_dir / "latest").is_file():
            step_dir = load_dir / (load_dir / "latest").read_text(encoding="UTF-8").strip()
        elif any(load_dir.glob("*.pt")):
            logger.info(f"no latest file found, using load dir directly instead: {load_dir}")
            step_dir = load_dir
        else:
            logger.error(f"no files found in load dir: {load_dir}")
            return False
        if not step_dir.is_dir():
            logger.error(f"iteration_dir does not exist: {step_dir}")
            return False
        self.parallel_module.load_checkpoint(
            step_dir, allowed_missing_keys_in_checkpoint=allowed_missing_keys_in_checkpoint,
            allowed_unexpected_keys_in_checkpoint=allowed_unexpected_keys_in_checkpoint,
            ignore_keys_in_checkpoint=ignore_keys_in_checkpoint)
        if load_optimizer_states and any(step_dir.glob("optimizer_state_*.pt")):
            self.optimizer.load_checkpoint(step_dir)
        else:
            self.optimizer.refresh_optimizer_after_model_change()
        if load_context:
            self.context.load_checkpoint(step_dir)
        logger.info(f"loaded checkpoint: {step_dir}")
        return True

    # ---- steps -----------------------------------------------------------------------------------------------
    def train_step(self) -> TrainStepOutput:
        out = self.parallel_module.train_step(
            dataloader=self.dataloader, optimizer=self.optimizer,
            sync_batch_to_model_parallel=self.sync_batch_to_model_parallel, loss_function=self.loss_function,
            metrics_aggregation_fn=self.metrics_aggregation_fn)
        self.context.step()
        return out

    def eval_step(self) -> EvaluationStepOutput:
        if self.context.topology.is_io_rank:
            assert self.dataloader_evaluation is not None, (
                "needs an evaluation dataloader on io ranks if evaluation is to be performed. "
                "Remember to give the trainer a dataset_evaluation")
        out = self.parallel_module.evaluation_step(
            dataloader=self.dataloader_evaluation, sync_batch_to_model_parallel=self.sync_batch_to_model_parallel,
            loss_function=self.loss_function, metrics_aggregation_fn=self.metrics_aggregation_fn)
        self.context.consumed_samples_evaluation += self.context.topology.config.global_batch_size
        return out

    def evaluate(self) -> EvaluationStepOutput:
        """``eval_iterations`` evaluation steps (each one global batch) averaged into one result — the reference defines
        the option but runs a single step regardless."""
        outs = [self.eval_step() for _ in range(max(1, int(self.config.eval_iterations or 1)))]
        if len(outs) == 1:
            return outs[0]
        n = float(len(outs))
        metrics = None
        if outs[0].metrics is not None:
            metrics = {k: sum(o.metrics[k] for o in outs) / n for k in outs[0].metrics}
        loss = None if outs[0].loss is None else sum(o.loss for o in outs) / n
        return EvaluationStepOutput(loss, metrics, sum(o.step_duration for o in outs))

    def log_metrics(self, train_step_output: TrainStepOutput, eval_step_output: Optional[EvaluationStepOutput]) -> Dict[str, Any]:
        logger.info(f"completed step {self.context.iterations}")
        t = train_step_output
        metrics: Dict[str, Any] = {f"training/{k}": v for k, v in (t.metrics or {}).items()}
        assert "training/loss" not in metrics
        metrics["training/loss"] = t.loss
        metrics["training/step_duration"] = t.step_duration
        metrics["runtime/step_duration"] = t.step_duration
        if t.global_grad_norm is not None:
            metrics["training/global_grad_norm"] = t.global_grad_norm
        if t.global_grad_norm_clipped is not None:
            metrics["training/global_grad_norm_clipped"] = t.global_grad_norm_clipped
        for name, lr in (t.learning_rates or {}).items():
            metrics[f"training/learning_rate_{name}"] = lr
        if t.overflow is not None:
            metrics["training/overflow"] = int(t.overflow)
        if t.no_overflow_steps is not None:
            metrics["training/no_overflow_steps"] = t.no_overflow_steps
        if t.current_loss_scale is not None:
            metrics["training/current_loss_scale"] = t.current_loss_scale
        if t.debug_dict:
            metrics.update(t.debug_dict)
        if eval_step_output is not None:
            for k, v in (eval_step_output.metrics or {}).items():
                metrics[f"evaluation/{k}"] = v
            metrics["evaluation/loss"] = eval_step_output.loss
            metrics["evaluation/step_duration"] = eval_step_output.step_duration
        logger.log_metrics(metrics, step=self.context.iterations)
        return metrics

    def _after_step(self) -> None:
        """Hook between the step and checkpointing (preemption handling in the Determined trainer)."""

    def run_training(self, return_metrics: bool = False) -> Optional[List[Dict[str, Union[float, int]]]]:
        collected: List[Dict[str, Any]] = []
        while self.context.iterations < (self.config.train_iterations or 0):
            out = self.train_step()
            self._after_step()
            if (self.config.save_interval is not None
                    and (self.config.save_dir is not None or isinstance(logger._impl, DeterminedLogger))
                    and self.context.iterations % self.config.save_interval == 0):
                self.save_checkpoint()
            ev = None
            if self.config.eval_interval is not None and self.context.iterations % self.config.eval_interval == 0:
                ev = self.evaluate()
            if self.context.topology.config.global_rank == 0:
                m = self.log_metrics(out, ev)
                if return_metrics:
                    collected.append(m)
        return collected if return_metrics else None


class DeterminedBaseTrainer(BaseTrainer[DeterminedBaseContextGeneric, ParallelModuleGeneric]):
    """Trainer that, when the context carries a Determined core context, stores checkpoints through Determined's
    storage manager, resumes from ``latest_checkpoint`` and saves + exits on preemption (reference :317-558).
    Without Determined it behaves exactly like :class:`BaseTrainer`."""

    def save_checkpoint(self, save_dir: Optional[Path] = None) -> Path:
        if getattr(self.context, "_use_determined", False):
            return self.determined_save_checkpoint()
        return super().save_checkpoint(save_dir=save_dir)

    def determined_save_checkpoint(self) -> Path:
        det = self.context.determined_context
        assert det is not None, "tried to save a Determined checkpoint but no determined context is given"
        manager = det.checkpoint._storage_manager
        if self.context.topology.config.global_rank == 0:
            storage_id = str(uuid.uuid4())
            with manager.store_path(storage_id) as path:
                det.distributed.broadcast((storage_id, path))
                super().save_checkpoint(save_dir=Path(path))
                from determined.common import storage  # type: ignore

                resources = storage.StorageManager._list_directory(path)
                all_res = [resources] if isinstance(manager, storage.SharedFSStorageManager) else det.distributed.gather(resources)
            merged = {k: v for d in all_res for k, v in d.items()}
            det.checkpoint._report_checkpoint(storage_id, merged, {"steps_completed": self.context.iterations})
            return Path(path)
        storage_id, path = det.distributed.broadcast(None)
        super().save_checkpoint(save_dir=Path(path))
        return Path(path)

    def load_checkpoint(self, load_dir: Optional[Path] = None, load_optimizer_states: bool = True, load_context: bool = True,
                        allowed_missing_keys_in_checkpoint: Optional[List[str]] = None,
                        allowed_unexpected_keys_in_checkpoint: Optional[List[str]] = None,
                        ignore_keys_in_checkpoint: Optional[List[str]] = None) -> bool:
        resumed = False
        if getattr(self.context, "_use_determined", False):
            import determined as det  # type: ignore

            info = det.get_cluster_info()
            if info is not None and info.latest_checkpoint is not None:
                resumed = True
                with self.context.determined_context.checkpoint.restore_path(info.latest_checkpoint) as p:
                    return super().load_checkpoint(Path(p), True, True, allowed_missing_keys_in_checkpoint,
                                                   allowed_unexpected_keys_in_checkpoint, ignore_keys_in_checkpoint)
        return super().load_checkpoint(load_dir, load_optimizer_states or resumed, load_context or resumed,
                                       allowed_missing_keys_in_checkpoint, allowed_unexpected_keys_in_checkpoint,
                                       ignore_keys_in_checkpoint)

    def _after_step(self) -> None:
        if not getattr(self.context, "_use_determined", False):
            return
        if self.context.determined_profiler is not None:
            self.context.determined_profiler.update_batch_idx(self.context.iterations)
        if self.context.determined_context.preempt.should_preempt():
            self.determined_save_checkpoint()
            print("exiting program after preemption.", flush=True)
            sys.exit()
