"""Single-process, multi-device pipeline for inference with hidden-state recording (reference:
core/nn/parallel_module/inference_module.py:1-122).

Every pipeline stage is built in this process, one device per stage; the layer IO hops devices at stage boundaries.
Tied layer specs become plain specs (the checkpoint loader fills both copies from the same file).  The recorder
attaches forward hooks to the named sub-modules of chosen layers and returns ``{layer index: {module name: output}}``
(the empty name ``""`` is the layer itself).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import torch

from ..topology import PipePartitionMethod
from .layers import LayerSpec
from .partitioned_module import PipePartitionedModule


@dataclass
class RecorderSetting:
    include_modules: Optional[Sequence[str]] = ("",)
    exclude_modules: Optional[Sequence[str]] = None

    def __post_init__(self) -> None:
        assert self.include_modules is None or self.exclude_modules is None, \
            "Cannot specify both include_modules and exclude_modules"

    def wants(self, name: str) -> bool:
        if self.include_modules is not None:
            return name in self.include_modules
        return self.exclude_modules is not None and name not in self.exclude_modules


class HiddenStateRecorder:
    """Context manager: while active, outputs of the selected sub-modules land in ``current_record``."""

    def __init__(self, module: PipePartitionedModule, recorder_settings_per_layer: Dict[int, RecorderSetting]):
        self._module = module
        self.recorder_settings_per_layer = recorder_settings_per_layer
        self.current_record: Dict[int, Dict[str, Any]] = {k: {} for k in recorder_settings_per_layer}
        self._hooks: List[Any] = []

    def __enter__(self) -> "HiddenStateRecorder":
        self.start_recording()
        return self

    def __exit__(self, *_exc: Any) -> None:
        self.stop_recording()

    def start_recording(self) -> None:
        for layer_index, setting in self.recorder_settings_per_layer.items():
            record = self.current_record[layer_index]
            for name, sub in self._module._layers[layer_index].named_modules():
                if setting.wants(name):
                    self._hooks.append(sub.register_forward_hook(
                        lambda _m, _i, out, record=record, name=name: record.__setitem__(name, out)))

    def stop_recording(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []

    def delete_records(self) -> None:
        self.current_record = {k: {} for k in self.recorder_settings_per_layer}


class InferenceModule(PipePartitionedModule):
    def __init__(self, layer_specs: List[LayerSpec], devices: Optional[Sequence[Union[int, str, torch.device]]] = None,
                 pipe_partition_method: PipePartitionMethod = PipePartitionMethod.UNIFORM,
                 pipe_partition_overwrite: Optional[List[int]] = None):
        if devices is None:
            devices = [0] if torch.cuda.is_available() else ["cpu"]
        super().__init__(layer_specs=layer_specs, devices=devices, pipe_partition_method=pipe_partition_method,
                         pipe_partition_overwrite=pipe_partition_overwrite)
        self.eval()

    @torch.no_grad()
    def forward(self, x: Any) -> Any:
        current = None
        for layer, device in zip(self._layers, self._layer_devices):
            if device != current:  # stage boundary
                if hasattr(x, "to_"):
                    x.to_(device)
                elif torch.is_tensor(x):
                    x = x.to(device)
                current = device
            x = layer(x)
        return x

    @torch.no_grad()
    def forward_with_hidden_state_recorder(self, x: Any, recorder_settings_per_layer: Optional[Dict[int, RecorderSetting]] = None
                                           ) -> Tuple[Any, Dict[int, Dict[str, Any]]]:
        recorder = HiddenStateRecorder(self, recorder_settings_per_layer or {})
        with recorder:
            x = self.forward(x)
        return x, recorder.current_record
