"""Per-parameter metadata attached as ``param.core_parameter_meta`` (reference: core/nn/parameter_meta.py:1-150).

A small dataclass: where the parameter lives in the layer list, how (and whether) it is split over the tensor-parallel
group and whether it is tied across layers.  ``key`` is the layout-independent identifier optimizer checkpoints are
keyed by and ``state_dict`` is stored next to every optimizer entry, so both formats are part of the on-disk contract
 and must not change.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Set, Tuple

import torch

_KEY = "layer_index_{layer}_parameter_name_{name}_is_model_parallel_{mp}_model_parallel_dimension_{dim}"
_STATE_FIELDS = ("local_shape", "is_model_parallel", "model_parallel_dimension", "layer_index", "parameter_name", "is_tied",
                 "tied_layer_indices", "tied_grad_on_model_parallel")


@dataclass(eq=False)
class CoreParameterMeta:
    local_shape: Tuple[int, ...]
    is_model_parallel: bool
    model_parallel_dimension: Optional[int] = None
    layer_index: Optional[int] = None
    parameter_name: Optional[str] = None
    layer_class_name: Optional[str] = None
    is_tied: bool = False
    tied_layer_indices: Set[int] = field(default_factory=set)
    tied_grad_on_model_parallel: bool = False

    def __post_init__(self) -> None:
        self.local_shape = tuple(self.local_shape)
        self.tied_layer_indices = set(self.tied_layer_indices or ())
        if self.layer_index is not None:
            self.set_layer_index(self.layer_index)

    # ---- identity -------------------------------------------------------------------------------------------
    def key_for_layer(self, layer_index: Optional[int]) -> str:
        return _KEY.format(layer=layer_index, name=self.parameter_name, mp=self.is_model_parallel,
                           dim=self.model_parallel_dimension)

    @property
    def key(self) -> str:
        return self.key_for_layer(self.layer_index)

    def possible_keys(self) -> List[str]:
        """A tied parameter is stored under every layer that uses it."""
        layers = sorted(self.tied_layer_indices) if self.is_tied else [self.layer_index]
        return [self.key_for_layer(i) for i in layers]

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, CoreParameterMeta) and other.key == self.key

    def __hash__(self) -> int:
        return hash(self.key)

    def __repr__(self) -> str:
        return (f"CoreParameterMeta [{self.parameter_name}] layer_index [{self.layer_index}] "
                f"layer_class_name [{self.layer_class_name}] is_model_parallel [{self.is_model_parallel}]")

    @property
    def is_model_parallel_duplicate(self) -> bool:
        """Replicated on every TP rank (norm weights, row-parallel biases, …)."""
        return not self.is_model_parallel

    # ---- checkpoint -----------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        return {name: getattr(self, name) for name in _STATE_FIELDS}

    @classmethod
    def from_state_dict(cls, state_dict: Dict[str, Any]) -> "CoreParameterMeta":
        return cls(**state_dict)

    # ---- mutation while the module tree is assembled -----------------------------------------------------------
    def set_layer_index(self, layer_index: int) -> None:
        self.layer_index = layer_index
        if self.is_tied:
            self.tied_layer_indices.add(layer_index)

    def set_parameter_name(self, parameter_name: str) -> None:
        self.parameter_name = parameter_name

    def set_layer_class_name(self, layer_class_name: str) -> None:
        self.layer_class_name = layer_class_name

    def set_is_tied(self, is_tied: bool) -> None:
        self.is_tied = is_tied

    def set(self, layer_index: int, parameter_name: str, layer_class_name: str, is_tied: bool) -> None:
        self.is_tied, self.parameter_name, self.layer_class_name = is_tied, parameter_name, layer_class_name
        self.set_layer_index(layer_index)

    @staticmethod
    def register_on_parameter(parameter: torch.Tensor, is_model_parallel: bool, model_parallel_dimension: Optional[int] = None,
                              **fields: Any) -> "CoreParameterMeta":
        """Create the metadata of ``parameter`` and hang it on the tensor (once)."""
        if hasattr(parameter, "core_parameter_meta"):
            raise AssertionError("core_parameter_meta already registered")
        meta = CoreParameterMeta(tuple(parameter.shape), is_model_parallel, model_parallel_dimension, **fields)
        parameter.core_parameter_meta = meta  # type: ignore[attr-defined]
        return meta
