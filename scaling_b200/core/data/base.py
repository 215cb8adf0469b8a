"""Dataset / batch / layer-IO contracts (reference: core/data/base_dataset.py:1-108, base_layer_io.py:1-13)."""
from __future__ import annotations

from abc import abstractmethod
from typing import Any, Generic, List, Optional, TypeVar

import torch


class BaseLayerIO:
    """Bag of tensors passed between layers; helpers act on every tensor attribute in place."""

    def to_(self, device: torch.device, non_blocking: bool = False) -> None:
        for name, attr in self.__dict__.items():
            if isinstance(attr, torch.Tensor):
                setattr(self, name, attr.to(device, non_blocking=non_blocking))

    def contiguous_(self) -> None:
        for name, attr in self.__dict__.items():
            if isinstance(attr, torch.Tensor):
                setattr(self, name, attr.contiguous())

    def pin_memory(self):  # used by torch's DataLoader(pin_memory=True)
        for name, attr in self.__dict__.items():
            if isinstance(attr, torch.Tensor) and not attr.is_cuda:
                setattr(self, name, attr.pin_memory())
        return self


class BaseDatasetItem:
    """One sample."""


TBatch = TypeVar("TBatch", bound="BaseDatasetBatch")


class BaseDatasetBatch(BaseLayerIO):
    """A collated batch; the first pipeline stage keeps ``only_inputs()``, the last ``only_targets()``."""

    @abstractmethod
    def only_inputs(self: TBatch) -> TBatch:
        return self

    @abstractmethod
    def only_targets(self: TBatch) -> TBatch:
        return self


BaseDatasetItemGeneric = TypeVar("BaseDatasetItemGeneric", bound=BaseDatasetItem)
BaseDatasetBatchBeforeSyncGeneric = TypeVar("BaseDatasetBatchBeforeSyncGeneric", bound=BaseDatasetBatch)
BaseDatasetBatchGeneric = TypeVar("BaseDatasetBatchGeneric", bound=BaseDatasetBatch)


class BaseDataset(torch.utils.data.Dataset, Generic[BaseDatasetItemGeneric, BaseDatasetBatchBeforeSyncGeneric, BaseDatasetBatchGeneric]):
    def __init__(self, seed: int, shuffle: bool = True) -> None:
        self.seed: Optional[int] = None
        self.set_seed(seed=seed, shuffle=shuffle)

    @abstractmethod
    def ident(self) -> str:
        raise NotImplementedError

    @abstractmethod
    def __len__(self) -> int:
        raise NotImplementedError

    @abstractmethod
    def __getitem__(self, index: int) -> Any:
        raise NotImplementedError

    @abstractmethod
    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        raise NotImplementedError

    @abstractmethod
    def collate(self, batch: List[Any]) -> Any:
        raise NotImplementedError

    @staticmethod
    @abstractmethod
    def sync_batch_to_model_parallel(topology: Any, batch: Optional[Any]) -> Any:
        raise NotImplementedError

    def __repr__(self) -> str:
        return self.__class__.__name__
