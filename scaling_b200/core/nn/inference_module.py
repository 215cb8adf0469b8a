"""Single-process, multi-device pipeline for inference with hidden-state recording (reference:
core/nn/parallel_module/inference_module.py:1-122)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Sequence, Set, Union

import torch

from .layers import LayerSpec
from .partitioned_module import PipePartitionedModule
from ..topology import PipePartitionMethod


@dataclass
class RecorderSetting:
    """Which layers to record (``None`` = all) and whether to also keep attention scores where a layer exposes them."""
    layer_indices: Optional[Set[int]] = None
    include_attention_scores: bool = False


@dataclass
class HiddenStateRecorder:
    setting: RecorderSetting
    hidden_states: Dict[int, Any] = field(default_factory=dict)

    def wants(self, layer_index: int) -> bool:
        return self.setting.layer_indices is None or layer_index in self.setting.layer_indices

    def record(self, layer_index: int, output: Any) -> None:
        act = getattr(output, "activations", output)
        self.hidden_states[layer_index] = act.detach() if torch.is_tensor(act) else act


class InferenceModule(PipePartitionedModule):
    """All pipeline stages live in this process, one device per stage; activations hop between devices."""

    def __init__(self, layer_specs: List[LayerSpec], devices: Optional[Sequence[Union[int, str, torch.device]]] = None,
                 pipe_partition_method: PipePartitionMethod = PipePartitionMethod.UNIFORM,
                 pipe_partition_overwrite: Optional[List[int]] = None):
        if devices is None:
            devices = [0] if torch.cuda.is_available() else ["cpu"]
        super().__init__(layer_specs=layer_specs, devices=devices, pipe_partition_method=pipe_partition_method,
                         pipe_partition_overwrite=pipe_partition_overwrite)
        self.eval()

    @torch.no_grad()
    def forward(self, x: Any, recorder_setting: Optional[RecorderSetting] = None) -> Any:
        recorder = HiddenStateRecorder(recorder_setting) if recorder_setting is not None else None
        for (layer_index, layer), device in zip(self.local_layers(), self._layer_devices):
            if hasattr(x, "to_"):
                x.to_(device)
            elif torch.is_tensor(x):
                x = x.to(device)
            x = layer(x)
            if recorder is not None and recorder.wants(layer_index):
                recorder.record(layer_index, x)
        if recorder is not None:
            return x, recorder.hidden_states
        return x
