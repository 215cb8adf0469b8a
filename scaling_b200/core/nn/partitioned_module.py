"""Pipeline-partitioned module: only the owning stage instantiates a layer; tied layers share parameters; layer-wise,
layout-independent checkpoints.

Reference: core/nn/parallel_module/partitioned_module.py:1-371, tied_layer_index.py:1-224.  On-disk contract
(SURVEY §5.4): ``model_state_layer_{global index}_{LayerClassName}[_{separate name}].pt`` holding a dict
``parameter name -> TP-merged CPU tensor``; only dp-rank 0 / mp-rank 0 write.
"""
from __future__ import annotations

import re
from collections import defaultdict
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Set, Tuple, Union

import torch
import torch.distributed as dist

from ..logging import logger
from ..topology import PipePartitionMethod, Topology
from ..utils.param_merge import merge_parameter, split_parameter
from .layers import LayerSpec, TiedLayerSpec
from .parameter_meta import CoreParameterMeta
from .partition import (
    PipePartitionCoordinates,
    pipe_partition_balanced,
    pipe_partition_from_indices,
    pipe_partition_uniform,
)


def key_match(key: str, list_of_patterns: Sequence[str]) -> bool:
    """True if any regex pattern is found in ``key``."""
    return any(re.search(p, key) is not None for p in list_of_patterns)


def _resolve(module: torch.nn.Module, dotted: str) -> Tuple[torch.nn.Module, str]:
    parts = dotted.split(".")
    for p in parts[:-1]:
        module = getattr(module, p)
    return module, parts[-1]


class TiedLayerIndex:
    """Bookkeeping for ``TiedLayerSpec`` groups: which stages hold a copy, the process group across those stages (per
    dp/mp coordinate), and parameter sharing between copies that live on the same stage."""

    def __init__(self, pipe_partition_coordinates: List[PipePartitionCoordinates], layer_specs: List[LayerSpec],
                 topology: Topology):
        self.topology = topology
        own = pipe_partition_coordinates[topology.pipe_parallel_rank]
        stage_of = {}
        for stage, c in enumerate(pipe_partition_coordinates):
            for i in range(c.start, c.end):
                stage_of[i] = stage
        self.layers_by_key: Dict[str, List[int]] = defaultdict(list)
        self.attrs_by_key: Dict[str, Set[str]] = defaultdict(set)
        for i, spec in enumerate(layer_specs):
            if isinstance(spec, TiedLayerSpec):
                self.layers_by_key[spec.key].append(i)
                self.attrs_by_key[spec.key].update(spec.tied_weight_attributes)
        self.stages_by_key = {k: sorted({stage_of[i] for i in v}) for k, v in self.layers_by_key.items()}
        # process groups: created on every rank, in sorted key order, for every (dp, mp) coordinate
        self.group_by_key: Dict[str, Any] = {}
        self.ranks_by_key: Dict[str, List[int]] = {}
        c = topology.config
        for key in sorted(self.layers_by_key):
            stages = self.stages_by_key[key]
            for d in range(c.data_parallel_size):
                for m in range(c.model_parallel_size):
                    ranks = sorted(topology.get_global_rank(pipe_parallel_rank=s, data_parallel_rank=d, model_parallel_rank=m)
                                   for s in stages)
                    group = dist.new_group(ranks) if (len(ranks) > 1 and topology.is_distributed_initialized) else None
                    if c.global_rank in ranks:
                        self.group_by_key[key] = group
                        self.ranks_by_key[key] = ranks
        # local bookkeeping
        self.own = own
        self.main_layer_by_key = {k: v[0] for k, v in self.layers_by_key.items()}
        self.local_main_by_key: Dict[str, int] = {}
        for k, v in self.layers_by_key.items():
            local = [i for i in v if own.start <= i < own.end]
            if local:
                self.local_main_by_key[k] = local[0]
        self.key_of_layer = {i: k for k, v in self.layers_by_key.items() for i in v}
        self.local_main_modules: Dict[str, torch.nn.Module] = {}

    def share(self, layer_index: int, module: torch.nn.Module) -> None:
        """Called for each instantiated local layer, in order: later local copies re-use the first copy's parameters."""
        key = self.key_of_layer.get(layer_index)
        if key is None:
            return
        if self.local_main_by_key[key] == layer_index:
            self.local_main_modules[key] = module
            return
        main = self.local_main_modules[key]
        for attr in sorted(self.attrs_by_key[key]):
            src_mod, src_name = _resolve(main, attr)
            dst_mod, dst_name = _resolve(module, attr)
            setattr(dst_mod, dst_name, getattr(src_mod, src_name))

    def is_tied(self, layer_index: int) -> bool:
        return layer_index in self.key_of_layer

    def tied_attributes(self, layer_index: int) -> Set[str]:
        key = self.key_of_layer.get(layer_index)
        return set() if key is None else set(self.attrs_by_key[key])

    def layer_index_is_tied_global_duplicate(self, layer_index: int) -> bool:
        key = self.key_of_layer.get(layer_index)
        return key is not None and self.main_layer_by_key[key] != layer_index

    def layer_index_is_tied_local_duplicate(self, layer_index: int) -> bool:
        key = self.key_of_layer.get(layer_index)
        return key is not None and self.local_main_by_key.get(key) != layer_index

    def layer_index_to_tied_local_duplicate_parameter_names(self, layer_index: int) -> Set[str]:
        return self.tied_attributes(layer_index) if self.layer_index_is_tied_local_duplicate(layer_index) else set()

    def local_parameters_and_process_groups(self) -> List[Tuple[torch.nn.Parameter, Any, List[int]]]:
        """(parameter, group over the stages holding a copy, those stages) for every tied parameter on this stage."""
        out = []
        for key in sorted(self.local_main_modules):
            module = self.local_main_modules[key]
            for attr in sorted(self.attrs_by_key[key]):
                mod, name = _resolve(module, attr)
                out.append((getattr(mod, name), self.group_by_key.get(key), self.stages_by_key[key]))
        return out


class PipePartitionedModule(torch.nn.Module):
    """Either multi-process (``topology``: this process builds its own stage) or single-process (``devices``: every
    stage is built here, one device per stage — inference)."""

    def __init__(self, layer_specs: List[LayerSpec], devices: Optional[Sequence[Union[int, str, torch.device]]] = None,
                 topology: Optional[Topology] = None, pipe_partition_method: Optional[PipePartitionMethod] = None,
                 pipe_partition_overwrite: Optional[List[int]] = None):
        super().__init__()
        assert (devices is None) ^ (topology is None), "Exactly either one of 'devices' or 'topology' must be specified"
        self.topology = topology
        self.devices: Optional[List[torch.device]] = None
        if devices is not None:
            self.devices = [d if isinstance(d, torch.device) else (torch.device("cuda", d) if isinstance(d, int) else torch.device(d))
                            for d in devices]
        self._layer_specs = layer_specs
        if topology is None:
            self.pipe_partition_method = pipe_partition_method or PipePartitionMethod.UNIFORM
            self.pipe_partition_overwrite = pipe_partition_overwrite
        else:
            assert pipe_partition_method is None, "pipe partition method should be provided through the topology"
            assert pipe_partition_overwrite is None, "pipe partition overwrite should be provided through the topology"
            self.pipe_partition_method = topology.config.pipe_partition_method
            self.pipe_partition_overwrite = topology.config.pipe_partition_overwrite
        self._initialize_layers()

    # ---- construction -----------------------------------------------------------------------------
    def _get_pipe_partition_coordinates(self) -> List[PipePartitionCoordinates]:
        n = len(self.devices) if self.devices is not None else self.topology.config.pipe_parallel_size
        if len(self._layer_specs) < n:
            raise RuntimeError(f"Number of layers ({len(self._layer_specs)}) is smaller than number of pipe partitions {n}")
        if self.pipe_partition_overwrite is not None:
            return pipe_partition_from_indices(self.pipe_partition_overwrite, num_layers=len(self._layer_specs))
        if self.pipe_partition_method == PipePartitionMethod.UNIFORM or n == 1:
            return pipe_partition_uniform(len(self._layer_specs), n)
        if self.pipe_partition_method == PipePartitionMethod.BALANCED:
            return pipe_partition_balanced(self._layer_specs, n)
        raise NotImplementedError(f"Pipe partition method not known: {self.pipe_partition_method}")

    def _initialize_layers(self) -> None:
        coords = self._get_pipe_partition_coordinates()
        self._all_pipe_partition_coordinates = coords
        self.tied_layer_index: Optional[TiedLayerIndex] = None
        layers: List[torch.nn.Module] = []
        if self.topology is not None:
            own = coords[self.topology.pipe_parallel_rank]
            self._pipe_partition_coordinates = [own]
            logger.info(f"pipe_parallel_rank {self.topology.pipe_parallel_rank} gets layers {own.start}:{own.end}")
            self.tied_layer_index = TiedLayerIndex(coords, self._layer_specs, self.topology)
            for i in range(own.start, own.end):
                layer = self._layer_specs[i].initialize(self.topology.device)
                self.tied_layer_index.share(i, layer)
                self._register_meta(layer, i)
                layers.append(layer)
            self._layer_devices = [self.topology.device] * len(layers)
        else:
            assert self.devices is not None
            self._pipe_partition_coordinates = coords
            self._layer_devices = []
            for stage, c in enumerate(coords):
                for i in range(c.start, c.end):
                    spec = self._layer_specs[i]
                    if isinstance(spec, TiedLayerSpec):  # single process: no tying across stages needed
                        spec = LayerSpec(spec.module_class, **spec.kwargs)
                    layer = spec.initialize(self.devices[stage])
                    self._register_meta(layer, i)
                    layers.append(layer)
                    self._layer_devices.append(self.devices[stage])
        self._layers = torch.nn.ModuleList(layers)

    def _register_meta(self, layer: torch.nn.Module, layer_index: int) -> None:
        tied_attrs = self.tied_layer_index.tied_attributes(layer_index) if self.tied_layer_index is not None else set()
        for name, tensor in list(layer.named_parameters(remove_duplicate=False)) + list(layer.named_buffers()):
            meta = getattr(tensor, "core_parameter_meta", None)
            if meta is None:
                meta = CoreParameterMeta.register_on_parameter(tensor, is_model_parallel=False)
            is_tied = name in tied_attrs
            if meta.layer_index is not None and is_tied:
                meta.set_is_tied(True)
                meta.set_layer_index(layer_index)  # shared parameter: remember every layer that uses it
                continue
            meta.set(layer_index=layer_index, parameter_name=name, layer_class_name=layer.__class__.__name__,
                     is_tied=is_tied)

    @property
    def layer_index_offset(self) -> int:
        return self._pipe_partition_coordinates[0].start

    def local_layers(self) -> List[Tuple[int, torch.nn.Module]]:
        """(global layer index, module) for every layer held by this process."""
        return [(self.layer_index_offset + i, l) for i, l in enumerate(self._layers)]

    # ---- checkpoint -------------------------------------------------------------------------------
    def _checkpoint_tensors(self, layer: torch.nn.Module) -> List[Tuple[str, torch.Tensor]]:
        persistent = set(layer.state_dict().keys())
        out = [(n, t) for n, t in list(layer.named_parameters(remove_duplicate=False)) + list(layer.named_buffers())
               if n in persistent]
        return out

    def save_checkpoint(self, dir_: Union[Path, str], separate_file_for_parameters: Optional[List[str]] = None) -> None:
        """One file per layer (plus one per ``separate_file_for_parameters`` entry); TP shards are merged."""
        if self.topology is not None and self.topology.data_parallel_rank != 0:
            return
        dir_ = Path(dir_)
        writer = self.topology is None or self.topology.model_parallel_rank == 0
        for layer_index, layer in self.local_layers():
            groups: Dict[str, Dict[str, torch.Tensor]] = {"": {}}
            for sep in separate_file_for_parameters or []:
                groups[sep] = {}
            for name, tensor in self._checkpoint_tensors(layer):
                merged = merge_parameter(tensor, tensor.core_parameter_meta, self.topology)
                target = ""
                for sep in separate_file_for_parameters or []:
                    if sep in name:
                        target = sep
                groups[target][name] = merged
            if not writer:
                continue
            for sep, sd in groups.items():
                if not sd:
                    continue
                fname = f"model_state_layer_{layer_index}_{layer.__class__.__name__}{'' if sep == '' else '_' + sep}.pt"
                torch.save(sd, str(dir_ / fname))

    def load_checkpoint(self, dir_: Union[Path, str, Sequence[Union[Path, str]]],
                        add_bias_names_if_not_exist: Optional[List[str]] = None,
                        add_bias_names_if_not_exist_exceptions: Optional[List[str]] = None,
                        allowed_missing_keys_in_checkpoint: Optional[List[str]] = None,
                        allowed_unexpected_keys_in_checkpoint: Optional[List[str]] = None,
                        ignore_keys_in_checkpoint: Optional[List[str]] = None) -> None:
        paths = [Path(p) for p in (dir_ if isinstance(dir_, (list, tuple)) else [dir_])]
        missing_dirs = [p for p in paths if not p.is_dir()]
        if missing_dirs:
            raise RuntimeError(f"Weight set directories missing: {missing_dirs}")
        allowed_missing = list(allowed_missing_keys_in_checkpoint or [])
        ignore = list(ignore_keys_in_checkpoint or [])
        allowed_unexpected = list(allowed_unexpected_keys_in_checkpoint or []) + ignore
        unexpected: Set[str] = set()
        missing: Set[str] = set()
        for layer_index, layer in self.local_layers():
            state: Dict[str, torch.Tensor] = {}
            for p in paths:
                for f in sorted(p.glob(f"model_state_layer_{layer_index}_{layer.__class__.__name__}*.pt")):
                    state.update(torch.load(f, map_location="cpu", weights_only=False))
            if add_bias_names_if_not_exist:
                for k in list(state.keys()):
                    if add_bias_names_if_not_exist_exceptions and any(
                            part == e for part in k.split(".") for e in add_bias_names_if_not_exist_exceptions):
                        continue
                    if k.endswith(".bias"):
                        for bias_name in add_bias_names_if_not_exist:
                            if bias_name and f"{k}_{bias_name}" not in state:
                                state[f"{k}_{bias_name}"] = state[k].clone()
            for name, tensor in self._checkpoint_tensors(layer):
                if name in state and tensor.core_parameter_meta.is_model_parallel:
                    state[name] = split_parameter(state[name], tensor.core_parameter_meta, self.topology)
            for k in [k for k in state if key_match(k, ignore)]:
                del state[k]
            # copy in place so that views into flat parameter arenas stay valid
            own = dict(self._checkpoint_tensors(layer))
            for k, v in state.items():
                if k in own:
                    if tuple(own[k].shape) != tuple(v.shape):
                        raise RuntimeError(f"size mismatch for {k} in layer {layer_index}: checkpoint {tuple(v.shape)} "
                                           f"vs model {tuple(own[k].shape)}")
                    with torch.no_grad():
                        own[k].copy_(v.to(device=own[k].device, dtype=own[k].dtype))
                else:
                    unexpected.add(k)
            missing.update(k for k in own if k not in state and not key_match(k, ignore))
        bad_unexpected = {k for k in unexpected if not key_match(k, allowed_unexpected)}
        if unexpected - bad_unexpected:
            logger.warning(f"Ignoring unexpected keys in checkpoint: {unexpected - bad_unexpected}")
        if bad_unexpected:
            raise RuntimeError(f"Unexpected keys in checkpoint: {bad_unexpected}. You may add keys to "
                               f"'allowed_unexpected_keys_in_checkpoint'.")
        bad_missing = {k for k in missing if not key_match(k, allowed_missing)}
        if missing - bad_missing:
            logger.warning(f"Ignoring missing keys in checkpoint: {missing - bad_missing}")
        if bad_missing:
            raise RuntimeError(f"Missing keys in checkpoint: {bad_missing}. You may add keys to "
                               f"'allowed_missing_keys_in_checkpoint'.")
