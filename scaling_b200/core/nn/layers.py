"""Layer contract and lazy layer constructors (reference: core/nn/parallel_module/base_layer.py:1-70,
layer_spec.py:1-32)."""
from __future__ import annotations

from abc import abstractmethod
from typing import Any, Generic, List, Optional, Tuple, Type, TypeVar

import torch

from ..data.base import BaseDatasetBatch, BaseLayerIO

BaseLossInputGeneric = TypeVar("BaseLossInputGeneric")
BaseLayerInputGeneric = TypeVar("BaseLayerInputGeneric")
BaseLayerOutputGeneric = TypeVar("BaseLayerOutputGeneric", bound=BaseLayerIO)
BaseLayerLastLayerOutputGeneric = TypeVar("BaseLayerLastLayerOutputGeneric", bound=BaseLayerIO)
BaseDatasetBatchGeneric = TypeVar("BaseDatasetBatchGeneric", bound=BaseDatasetBatch)


class BaseLayer(torch.nn.Module, Generic[BaseLayerInputGeneric, BaseLayerOutputGeneric, BaseLayerLastLayerOutputGeneric]):
    """A pipeline layer: ``forward(io) -> io`` plus the tuple protocol used for stage-to-stage transfer and
    activation checkpointing (tensors and picklable values only)."""

    @abstractmethod
    def forward(self, x: Any) -> Any:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def input_to_tuple(input: Any) -> Tuple[Any, ...]:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def tuple_to_input(d: Tuple[Any, ...]) -> Any:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def output_to_tuple(output: Any) -> Tuple[Any, ...]:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def tuple_to_last_stage_activation(d: Tuple[Any, ...]) -> Any:
        raise NotImplementedError

    def _forward_tuple_input(self, *args: Any) -> Any:
        return self(self.tuple_to_input(tuple(args)))


class LayerSpec:
    """Deferred constructor: only the owning pipeline stage instantiates the module."""

    def __init__(self, module_class: Type[torch.nn.Module], **kwargs: Any) -> None:
        self.module_class = module_class
        self.kwargs = kwargs

    def initialize(self, device: Optional[torch.device] = None) -> torch.nn.Module:
        module = self.module_class(**self.kwargs)
        if device is None:
            topo = self.kwargs.get("topology")
            device = topo.device if topo is not None else None
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        return module.to(device)

    def estimate_trainable_parameters(self) -> Optional[int]:
        fn = getattr(self.module_class, "estimate_parameter_count", None)
        return None if fn is None else int(fn(**self.kwargs))


class TiedLayerSpec(LayerSpec):
    """Layer whose ``tied_weight_attributes`` are shared with every other spec of the same ``key``."""

    def __init__(self, key: str, tied_weight_attributes: List[str], module_class: Type[torch.nn.Module], **kwargs: Any):
        super().__init__(module_class=module_class, **kwargs)
        self.key = key
        assert len(set(tied_weight_attributes)) == len(tied_weight_attributes), (
            f"duplicates in tied_weight_attributes: {tied_weight_attributes}")
        self.tied_weight_attributes = tied_weight_attributes
