"""Datasets, loaders and the TP batch broadcast (reference: core/data/__init__.py) — lazy flat namespace."""
from __future__ import annotations

import importlib
from typing import Any, Dict, List

_EXPORTS: Dict[str, str] = {
    "base": ("BaseDataset BaseDatasetBatch BaseDatasetBatchBeforeSyncGeneric BaseDatasetBatchGeneric BaseDatasetItem "
             "BaseDatasetItemGeneric BaseLayerIO"),
    "blended": "BaseBlendedDataset BlendedDatasetConfig blend_index weights_by_num_docs weights_examples_proportional",
    "broadcast": "broadcast_data sync_sizes",
    "loader": "DataLoader RandomSampler",
    "memmap": "FileDataset FileHandle MemoryMapDataset MemoryMapDatasetBuilder RetryableException",
}
_WHERE: Dict[str, str] = {name: module for module, names in _EXPORTS.items() for name in names.split()}
__all__: List[str] = sorted(_WHERE)


def __getattr__(name: str) -> Any:
    module = _WHERE.get(name)
    if module is None:
        if name in _EXPORTS:
            return importlib.import_module(f"{__name__}.{name}")
        raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
    value = getattr(importlib.import_module(f"{__name__}.{module}"), name)
    globals()[name] = value
    return value


def __dir__() -> List[str]:
    return sorted(set(globals()) | set(__all__))
