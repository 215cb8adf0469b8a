"""Determined-AI entry point (reference: transformer/train_determined.py:1-183): torch-distributed launcher env →
``LaunchConfig``, trial hyper-parameters → config overrides, then the ordinary ``train.main``.

The hyper-parameter mapping is a table (``HPARAM_TARGETS``) rather than code: ``hparam name → path in the config``;
token-denominated budgets (``target_train_tokens``, ``warmup_tokens``) are converted to iterations with the trial's
``global_batch_size × sequence_length``.
"""
from __future__ import annotations

import argparse
import os
from typing import Any, Dict, List, Optional, Tuple, Union

from ..core.determined import determined_available, determined_profiler_from_ctx, init, maybe_periodic_stacktraces
from ..core.runner import LaunchConfig
from .config import TransformerConfig
from .train import main as train_main

# hparam -> (config section, *nested keys)
HPARAM_TARGETS: Dict[str, Tuple[str, ...]] = {
    "learning_rate": ("learning_rate_scheduler", "learning_rate"),
    "wandb_project": ("logger", "wandb_project"),
    "model_parallel_size": ("topology", "model_parallel_size"),
    "pipe_parallel_size": ("topology", "pipe_parallel_size"),
    "sequence_parallel": ("topology", "sequence_parallel"),
    "global_batch_size": ("topology", "global_batch_size"),
    "micro_batch_size": ("topology", "micro_batch_size"),
    "activation_checkpointing_type": ("topology", "activation_checkpointing_type"),
    "pipe_partition_method": ("topology", "pipe_partition_method"),
    "pipe_partition_overwrite": ("topology", "pipe_partition_overwrite"),
    "train_iterations": ("trainer", "train_iterations"),
    "kernel": ("transformer_architecture", "masked_softmax", "kernel"),
    "sequence_length": ("transformer_architecture", "sequence_length"),
}
_FALSY_SKIPPED = {"learning_rate"}  # the reference ignores a zero/None learning rate


def from_launcher_args_determined() -> LaunchConfig:
    """Determined starts the job through ``torch.distributed.run``: the device index is ``LOCAL_RANK``."""
    parser = argparse.ArgumentParser(description="process launch")
    parser.add_argument("--config", type=str, default=None, help="path to config file")
    parser.add_argument("remaining_args", nargs=argparse.REMAINDER)
    args = parser.parse_args()
    world_size = int(os.environ["WORLD_SIZE"])
    payload = None
    if args.config is not None:
        payload = TransformerConfig.from_yaml(args.config, overwrite_values={"topology": {"world_size": world_size}}).as_dict()
    return LaunchConfig(master_addr=os.environ["MASTER_ADDR"], master_port=int(os.environ["MASTER_PORT"]),
                        world_size=world_size, global_rank=int(os.environ["RANK"]),
                        local_slot=int(os.environ["LOCAL_RANK"]), payload=payload)


def _set(d: Dict[str, Any], path: Tuple[str, ...], value: Any) -> None:
    for k in path[:-1]:
        d = d.setdefault(k, {})
    d[path[-1]] = value


def hparams_to_overwrite(hparams: Optional[Dict[str, Any]], overwrite_config: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """Config overrides implied by a trial's hyper-parameters.  A ``layout`` entry (written when an individual run is
    restarted) replaces the whole hyper-parameter set."""
    out: Dict[str, Any] = dict(overwrite_config or {})
    if not hparams:
        return out
    if "layout" in hparams:
        hparams = hparams["layout"]
    if not hparams:
        return out
    for section in ("topology", "transformer_architecture", "logger", "trainer", "learning_rate_scheduler"):
        out.setdefault(section, {})
    for budget, target in (("target_train_tokens", ("trainer", "train_iterations")),
                           ("warmup_tokens", ("learning_rate_scheduler", "learning_rate_warmup_steps"))):
        if hparams.get(budget):
            assert "sequence_length" in hparams and "global_batch_size" in hparams, \
                f"{budget} needs sequence_length and global_batch_size to be converted to steps"
            _set(out, target, int(hparams[budget] / (hparams["global_batch_size"] * hparams["sequence_length"])))
    for name, path in HPARAM_TARGETS.items():
        if name not in hparams or (name in _FALSY_SKIPPED and not hparams[name]):
            continue
        _set(out, path, hparams[name])  # an explicit train_iterations wins over target_train_tokens (applied later)
    return out


def main(determined_context: Any, profiler: Any, overwrite_config: Optional[Dict[str, Any]] = None,
         return_metrics: bool = False, det_experiment_id: Optional[int] = None, det_trial_id: Optional[int] = None,
         info: Any = None) -> Optional[List[Dict[str, Union[float, int]]]]:
    launch_config = from_launcher_args_determined()
    overwrite = dict(overwrite_config or {})
    overwrite.setdefault("runner", {})["use_determined"] = True
    overwrite["determined_experiment_id"] = det_experiment_id
    overwrite["determined_trial_id"] = det_trial_id
    hparams = info.trial.hparams if info is not None else None
    overwrite = hparams_to_overwrite(hparams, overwrite)
    print(f"{overwrite=}")
    return train_main(launch_config, overwrite, return_metrics, determined_context, profiler)


if __name__ == "__main__":
    assert determined_available(), "train_determined needs the `determined` package"
    import determined as det  # type: ignore

    info = det.get_cluster_info()
    assert info is not None
    experiment_config = det.ExperimentConfig(info.trial._config)
    distributed = det.core.DistributedContext.from_torch_distributed()
    with maybe_periodic_stacktraces(experiment_config.debug_enabled()):
        with init(distributed=distributed) as ctx:
            with determined_profiler_from_ctx(ctx, experiment_config, info) as prof:
                main(ctx, prof, det_experiment_id=info.trial.experiment_id, det_trial_id=info.trial.trial_id, info=info)
