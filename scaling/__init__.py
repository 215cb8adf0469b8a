"""``scaling`` — the reference's import namespace, served by :mod:`scaling_b200`.

User scripts, examples and tests written against Aleph Alpha's *Scaling* import from paths such as
``scaling.core``, ``scaling.core.nn.parallel_module.inference_module``, ``scaling.core.logging.logging`` or
``scaling.transformer.context.config`` (reference: src/scaling/core/__init__.py:1-71,
src/scaling/transformer/__init__.py:3-14 and the sub-paths listed in SURVEY §9.1).  This package makes every one of those
module paths importable and resolves its attributes in the scaling_b200 modules that implement them — there is no
second implementation behind it and no per-module stub file: one table (``_ROUTES``) says which scaling_b200 modules
stand behind which reference path, and a meta-path finder synthesises the alias modules on demand.

``python -m scaling.core.runner.launch`` and ``python -m scaling.transformer.train`` work too (the alias loader hands
runpy a trampoline into the implementing module).
"""
from __future__ import annotations

import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import sys
import types
from typing import Any, Dict, List, Optional, Sequence

__version__ = "0.1.0"
_ALIAS_MARKER = "scaling_b200 alias"
_B = "scaling_b200."

# reference module path (without the leading "scaling.") -> implementing scaling_b200 modules, searched in order
_ROUTES: Dict[str, Sequence[str]] = {
    # ---- core
    "core": ["core"],
    "core.config": ["core.config"],
    "core.config.base": ["core.config"],
    "core.context": ["core.context"],
    "core.context.context": ["core.context"],
    "core.data": ["core.data"],
    "core.data.base_dataset": ["core.data.base"],
    "core.data.base_layer_io": ["core.data.base"],
    "core.data.blended_dataset": ["core.data.blended"],
    "core.data.blended_dataset_config": ["core.data.blended"],
    "core.data.broadcast_data": ["core.data.broadcast"],
    "core.data.dataloader": ["core.data.loader"],
    "core.data.file_dataset": ["core.data.memmap"],
    "core.data.file_handles": ["core.data.memmap"],
    "core.data.memory_map": ["core.data.memmap"],
    "core.determined": ["core.determined"],
    "core.determined.core": ["core.determined"],
    "core.logging": ["core.logging"],
    "core.logging.logging": ["core.logging"],
    "core.logging.logger_config": ["core.logging"],
    "core.logging.color_formatter": ["core.logging"],
    "core.nn": ["core.nn"],
    "core.nn.activation_function": ["core.nn.activation"],
    "core.nn.attention": ["core.nn.attention"],
    "core.nn.attention.attention": ["core.nn.attention"],
    "core.nn.linear": ["core.nn.linear", "core.nn.tp_ops"],
    "core.nn.linear.column_parallel_linear": ["core.nn.linear"],
    "core.nn.linear.row_parallel_linear": ["core.nn.linear"],
    "core.nn.linear.vocab_parallel_embedding": ["core.nn.linear"],
    "core.nn.linear.utils": ["core.nn.tp_ops"],
    "core.nn.lora": ["core.nn.lora"],
    "core.nn.lora_config": ["core.nn.lora"],
    "core.nn.masked_softmax": ["core.nn.softmax"],
    "core.nn.masked_softmax.masked_softmax": ["core.nn.softmax"],
    "core.nn.masked_softmax.masked_softmax_config": ["core.nn.softmax"],
    "core.nn.mlp": ["core.nn.mlp"],
    "core.nn.norm": ["core.nn.norms"],
    "core.nn.norm.get_norm": ["core.nn.norms"],
    "core.nn.norm.layernorm": ["core.nn.norms"],
    "core.nn.norm.layernorm_config": ["core.nn.norms"],
    "core.nn.norm.rms_norm": ["core.nn.norms"],
    "core.nn.parallel_module": ["core.nn.engine", "core.nn.p2p", "core.nn.layers", "core.nn.partitioned_module",
                                "core.nn.partition", "core.nn.inference_module", "core.nn.act_ckpt"],
    "core.nn.parallel_module.activation_checkpointing": ["core.nn.act_ckpt"],
    "core.nn.parallel_module.base_layer": ["core.nn.layers"],
    "core.nn.parallel_module.layer_spec": ["core.nn.layers"],
    "core.nn.parallel_module.buffers": ["core.nn.engine"],
    "core.nn.parallel_module.communicator": ["core.nn.p2p"],
    "core.nn.parallel_module.inference_module": ["core.nn.inference_module"],
    "core.nn.parallel_module.parallel_module": ["core.nn.engine"],
    "core.nn.parallel_module.partitioned_module": ["core.nn.partitioned_module"],
    "core.nn.parallel_module.pipeline_partitioning": ["core.nn.partition"],
    "core.nn.parallel_module.tied_layer_index": ["core.nn.partitioned_module"],
    "core.nn.parameter_meta": ["core.nn.parameter_meta"],
    "core.nn.pipeline_schedule": ["core.nn.schedule"],
    "core.nn.pipeline_schedule.base": ["core.nn.schedule"],
    "core.nn.pipeline_schedule.inference": ["core.nn.schedule"],
    "core.nn.pipeline_schedule.instructions": ["core.nn.schedule"],
    "core.nn.pipeline_schedule.train": ["core.nn.schedule"],
    "core.nn.rotary": ["core.nn.rotary"],
    "core.nn.rotary_config": ["core.nn.rotary"],
    "core.optimizer": ["core.optimizer"],
    "core.optimizer.allreduce": ["core.optimizer.optimizer"],
    "core.optimizer.base": ["core.optimizer.optimizer"],
    "core.optimizer.optimizer": ["core.optimizer.optimizer"],
    "core.optimizer.optimizer_config": ["core.optimizer.configs"],
    "core.optimizer.parameter_group": ["core.optimizer.param_group"],
    "core.optimizer.parameter_group_config": ["core.optimizer.configs"],
    "core.optimizer.loss_scaler": ["core.optimizer.loss_scaler", "core.optimizer.configs"],
    "core.optimizer.loss_scaler_config": ["core.optimizer.configs"],
    "core.optimizer.learning_rate_scheduler": ["core.optimizer.lr", "core.optimizer.configs"],
    "core.optimizer.learning_rate_scheduler.learning_rate_scheduler": ["core.optimizer.lr", "core.optimizer.configs"],
    "core.optimizer.learning_rate_scheduler.learning_rate_scheduler_config": ["core.optimizer.configs"],
    "core.profiler": ["core.profiler"],
    "core.profiler.profiler": ["core.profiler"],
    "core.profiler.profiler_config": ["core.profiler"],
    "core.profiler.timer": ["core.profiler"],
    "core.runner": ["core.runner"],
    "core.runner.launch": ["core.runner.launch"],
    "core.runner.launch_config": ["core.runner.config"],
    "core.runner.runner": ["core.runner.runner"],
    "core.runner.runner_config": ["core.runner.config"],
    "core.topology": ["core.topology"],
    "core.topology.rng_tracker": ["core.topology"],
    "core.topology.topology": ["core.topology"],
    "core.topology.topology_config": ["core.topology"],
    "core.trainer": ["core.trainer"],
    "core.trainer.trainer": ["core.trainer"],
    "core.trainer.trainer_config": ["core.trainer"],
    "core.utils": ["core.utils"],
    "core.utils.determined_utils": ["core.determined"],
    "core.utils.param_merge": ["core.utils.param_merge"],
    "core.utils.port": ["core.utils.port"],
    # ---- transformer
    "transformer": ["transformer"],
    "transformer.context": ["transformer.config", "transformer.context"],
    "transformer.context.config": ["transformer.config"],
    "transformer.context.context": ["transformer.context"],
    "transformer.data": ["transformer.data"],
    "transformer.data.finetuning_chat_dataset": ["transformer.data.finetuning"],
    "transformer.data.finetuning_text_dataset": ["transformer.data.finetuning"],
    "transformer.data.inference_settings": ["transformer.data.inference_settings"],
    "transformer.data.legacy_blended_dataset": ["transformer.data.legacy"],
    "transformer.data.legacy_dataset": ["transformer.data.legacy"],
    "transformer.data.legacy_dataset.indexed_dataset": ["transformer.data.legacy"],
    "transformer.data.text_dataset": ["transformer.data.text_dataset"],
    "transformer.data.text_dataset_batch": ["transformer.data.text_dataset_batch"],
    "transformer.data.text_dataset_item": ["transformer.data.text_dataset", "transformer.data.text_dataset_batch"],
    "transformer.data.utils": ["transformer.data.utils"],
    "transformer.dataset_loader": ["transformer.dataset_loader"],
    "transformer.inference": ["transformer.inference"],
    "transformer.inference.inference_model": ["transformer.inference.inference_model"],
    "transformer.inference.sample": ["transformer.inference.sample"],
    "transformer.model": ["transformer.model"],
    "transformer.model.model": ["transformer.model.model"],
    "transformer.model.image_encoder": ["transformer.model.image_encoder"],
    "transformer.model.image_encoder.clip": ["transformer.model.image_encoder"],
    "transformer.model.image_encoder.image_encoder": ["transformer.model.image_encoder"],
    "transformer.model.layers": ["transformer.model.layers"],
    "transformer.model.layers.base": ["transformer.model.layers"],
    "transformer.model.layers.embedding": ["transformer.model.layers"],
    "transformer.model.layers.embedding_head": ["transformer.model.layers"],
    "transformer.model.layers.layer": ["transformer.model.layers"],
    "transformer.model.layers.layernorm": ["transformer.model.layers"],
    "transformer.model.layers.lm_head": ["transformer.model.layers"],
    "transformer.model.layers.lm_head_tied": ["transformer.model.layers"],
    "transformer.tokenizer": ["transformer.tokenizer"],
    "transformer.tokenizer.tokenizer": ["transformer.tokenizer"],
    "transformer.train": ["transformer.train"],
    "transformer.train_determined": ["transformer.train_determined"],
    "transformer.utils": ["transformer.utils"],
    "transformer.utils.get_tflops": ["transformer.utils.get_tflops"],
}


class _AliasModule(types.ModuleType):
    """Module whose attributes live in one or more scaling_b200 modules (looked up lazily, cached on first use)."""

    _targets: Sequence[str] = ()

    def _target_modules(self) -> List[types.ModuleType]:
        return [importlib.import_module(_B + t) for t in self._targets]

    def __getattr__(self, name: str) -> Any:
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        for mod in self._target_modules():
            try:
                value = getattr(mod, name)
            except AttributeError:
                continue
            # sub-modules of the implementation must not leak in under a reference package name
            if isinstance(value, types.ModuleType) and value.__name__.startswith(_B) and \
                    f"{self.__name__}.{name}"[len("scaling."):] in _ROUTES:
                continue
            self.__dict__[name] = value
            return value
        raise AttributeError(f"module {self.__name__!r} has no attribute {name!r} "
                             f"(looked in {', '.join(_B + t for t in self._targets)})")

    def __dir__(self) -> List[str]:
        names = set(self.__dict__)
        for mod in self._target_modules():
            names.update(getattr(mod, "__all__", None) or [n for n in dir(mod) if not n.startswith("_")])
        return sorted(names)


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname: str, path: Any = None, target: Any = None) -> Optional[importlib.machinery.ModuleSpec]:
        if not fullname.startswith("scaling."):
            return None
        root = sys.modules.get("scaling")
        if getattr(root, "_ALIAS_MARKER", None) != _ALIAS_MARKER:  # another `scaling` (the reference itself) is loaded
            return None
        key = fullname[len("scaling."):]
        if key not in _ROUTES:
            return None
        is_pkg = any(k.startswith(key + ".") for k in _ROUTES)
        return importlib.machinery.ModuleSpec(fullname, self, is_package=is_pkg)

    def create_module(self, spec: importlib.machinery.ModuleSpec) -> types.ModuleType:
        mod = _AliasModule(spec.name)
        mod._targets = tuple(_ROUTES[spec.name[len("scaling."):]])
        mod.__doc__ = f"alias of {', '.join(_B + t for t in mod._targets)} (reference path {spec.name})"
        if spec.submodule_search_locations is not None:
            mod.__path__ = []  # a package: sub-modules come from this finder, not from disk
        return mod

    def exec_module(self, module: types.ModuleType) -> None:
        all_names: List[str] = []
        for mod in module._target_modules():  # type: ignore[attr-defined]
            all_names.extend(getattr(mod, "__all__", None) or [])
        if all_names:
            module.__all__ = sorted(set(all_names))

    # runpy (`python -m scaling.core.runner.launch`) asks the loader for the code object of the module
    def get_code(self, fullname: str) -> Any:
        # a trampoline rather than the implementing module's own code object: that code uses relative imports, which
        # must keep resolving inside scaling_b200
        target = _B + _ROUTES[fullname[len("scaling."):]][0]
        return compile(f"import runpy as _r\n_r.run_module({target!r}, run_name='__main__', alter_sys=True)\n",
                       f"<scaling alias of {target}>", "exec")

    def get_source(self, fullname: str) -> Optional[str]:
        return None

    def is_package(self, fullname: str) -> bool:
        key = fullname[len("scaling."):]
        return any(k.startswith(key + ".") for k in _ROUTES)


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())


def __getattr__(name: str) -> Any:
    if name in ("core", "transformer"):
        return importlib.import_module(f"scaling.{name}")
    raise AttributeError(f"module 'scaling' has no attribute {name!r}")
