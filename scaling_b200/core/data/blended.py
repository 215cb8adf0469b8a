"""Blend several datasets into one interleaved index (reference: core/data/blended_dataset.py:1-398,
blended_dataset_config.py:1-106).

The interleave index is built by the native C++ ``sb_blend_index`` (csrc/blend_index.cpp), the in-tree equivalent of
the Rust ``blended_dataset_loop.sample`` wheel the reference depends on; cache files use the same three names
(``{stem}.bin``, ``{stem}.meta.json``, ``{stem}.input.json``).  A pure-NumPy fallback keeps the class usable when the
native library is not built.  Global rank 0 builds, the others poll for the files.
"""
from __future__ import annotations

import ctypes
import hashlib
import json
import time
from pathlib import Path
from typing import Any, Generic, List, Optional, Sequence, TypeVar

import numpy as np
import torch
from pydantic import Field

from ..config import BaseConfig
from ..logging import logger
from .base import BaseDataset, BaseDatasetBatchBeforeSyncGeneric, BaseDatasetBatchGeneric, BaseDatasetItemGeneric

# fourth type parameter of BaseBlendedDataset: the type of the blended member datasets (reference blended_dataset.py:128-140)
BaseDatasetGeneric = TypeVar("BaseDatasetGeneric", bound=BaseDataset)


class BlendedDatasetConfig(BaseConfig):
    weight_by_num_documents: bool = Field(True, description="derive weights from document counts (overrides `weights`)")
    weighted_sampler_alpha: float = Field(1.0, description="p_i ∝ n_i^alpha; 1 = proportional, 0 = uniform over datasets")
    weights: Optional[List[float]] = Field(None, description="explicit per-dataset weights, same order as the datasets")
    weight_examples_proportional: bool = Field(False, description="examples-proportional mixing with rate limit and temperature (T5, arXiv:1910.10683 p.31)")
    ep_maximum: Optional[int] = Field(None, description="rate limit K of examples-proportional mixing")
    ep_temperature: float = Field(1.0, description="temperature T of examples-proportional mixing")
    minimum_dataset_size: int = Field(0, description="repeat the blend until it has at least this many items")
    cache_directory: Optional[Path] = Field(None, description="where the blend index is cached (required for >1 dataset)")
    shuffle_dataset_indices: bool = Field(False, description="shuffle the blend index after creation")
    load_dataset_indices_to_memory: bool = Field(False, description="read the blend index into RAM instead of mmap")


def weights_by_num_docs(sizes: Sequence[int], alpha: float = 0.3) -> np.ndarray:
    """Per-dataset multipliers for "documents to draw = weight x size": sampling probability q_i ∝ p_i^alpha with
    p_i = n_i / Σn, expressed relative to the unbiased rate (w_i ∝ q_i / p_i) — alpha < 1 boosts small datasets,
    alpha = 1 keeps every dataset at its natural share (reference :24-59; XLM / XLM-R style)."""
    n = np.asarray(sizes, dtype=np.float64)
    p = n / n.sum()
    q = p ** alpha
    q = q / q.sum()
    w = q / p
    return w / w.sum()


def weights_examples_proportional(sizes: Sequence[int], temperature: float = 1.0, maximum: Optional[float] = None) -> np.ndarray:
    """Examples-proportional mixing (T5, arXiv:1910.10683 p.31): rate r_m = min(e_m, K) / Σ min(e_n, K), optionally
    sharpened by 1/T, again expressed relative to the unbiased rate (reference :62-125)."""
    assert temperature is not None and temperature != 0, "expect a non-zero temperature"
    n = np.asarray(sizes, dtype=np.float64)
    p = n / n.sum()
    limited = n.copy()
    if maximum:
        assert maximum > 0, f"examples-proportional sampling requires maximum limit > 0 (current max = {maximum})"
        limited = np.minimum(limited, float(maximum))
    r = limited / limited.sum()
    if temperature != 1.0:
        r = r ** (1.0 / temperature)
        r = r / r.sum()
    w = r / p
    return w / w.sum()


def blend_index(counts: np.ndarray) -> np.ndarray:
    """[total, 2] int64 rows (dataset index, index within dataset), most-behind-its-target dataset first."""
    counts = np.ascontiguousarray(counts, dtype=np.int64)
    total = int(counts[counts > 0].sum())
    out = np.empty((total, 2), dtype=np.int64)
    try:
        from ...ops._lib import host_lib

        lib = host_lib()
        n = lib.sb_blend_index(counts.ctypes.data_as(ctypes.c_void_p), len(counts), out.ctypes.data_as(ctypes.c_void_p), total)
        assert n == total, f"native blend index wrote {n} of {total} rows"
        return out
    except (RuntimeError, OSError, AttributeError):
        sampled = np.zeros(len(counts), dtype=np.int64)
        for w in range(total):
            ratio = np.where(sampled < counts, sampled / np.maximum(counts, 1), np.inf)
            i = int(ratio.argmin())
            out[w] = (i, sampled[i])
            sampled[i] += 1
        return out


class BaseBlendedDataset(BaseDataset[BaseDatasetItemGeneric, BaseDatasetBatchBeforeSyncGeneric, BaseDatasetBatchGeneric],
                         Generic[BaseDatasetItemGeneric, BaseDatasetBatchBeforeSyncGeneric, BaseDatasetBatchGeneric,
                                 BaseDatasetGeneric]):
    def __init__(self, seed: int, config: BlendedDatasetConfig, datasets: Sequence[BaseDataset], shuffle: bool = True):
        assert len(datasets) > 0, "at least one dataset is needed"
        self.config = config
        self.datasets = list(datasets)
        if len(self.datasets) > 1:
            assert config.cache_directory is not None, "cache_directory must be set to blend more than one dataset"
        self.dataset_indices: Optional[np.ndarray] = None
        self.random_index: Optional[np.ndarray] = None
        self.size = 0
        super().__init__(seed=seed, shuffle=shuffle)

    def ident(self) -> str:
        h = hashlib.md5()
        for d in self.datasets:
            h.update(d.ident().encode("utf-8"))
        h.update(self.config.as_str().encode("utf-8"))
        return h.hexdigest()

    def _cache_stem(self, seed: int) -> str:
        assert self.config.cache_directory is not None
        return str(Path(self.config.cache_directory) / f"blended_dataset_{self.ident()}_seed_{seed}")

    def _weights(self) -> np.ndarray:
        sizes = [len(d) for d in self.datasets]
        c = self.config
        if c.weight_examples_proportional:
            w = weights_examples_proportional(sizes, temperature=c.ep_temperature, maximum=c.ep_maximum)
        elif c.weight_by_num_documents:
            w = weights_by_num_docs(sizes, alpha=c.weighted_sampler_alpha)
        else:
            assert c.weights is not None and len(c.weights) == len(sizes), "weights must match the datasets"
            w = list(c.weights)
        return np.asarray(w, dtype=np.float64)

    def _targets(self) -> np.ndarray:
        sizes = np.asarray([len(d) for d in self.datasets], dtype=np.float64)
        w = self._weights()
        rel = w / w.max()
        if self.config.weight_examples_proportional:
            t = [max(1, int(round(p * n))) for n, p in zip(sizes, rel)]
        else:
            t = [max(1, int(p * n)) for n, p in zip(sizes, rel)]
        t = np.asarray(t, dtype=np.int64)
        if self.config.minimum_dataset_size > 0 and t.sum() < self.config.minimum_dataset_size:
            t = t * int(np.ceil(self.config.minimum_dataset_size / t.sum()))
        return t

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        if self.seed is not None and self.seed == seed:
            return
        self.seed = seed
        for d in self.datasets:
            d.set_seed(seed=seed, shuffle=shuffle)
        if len(self.datasets) == 1:
            self.size = len(self.datasets[0])
            if self.config.minimum_dataset_size > self.size:
                self.size = self.size * int(np.ceil(self.config.minimum_dataset_size / self.size))
            self.dataset_indices = None
            return
        stem = self._cache_stem(seed)
        f_bin, f_meta, f_in = Path(stem + ".bin"), Path(stem + ".meta.json"), Path(stem + ".input.json")
        rank0 = (not torch.distributed.is_initialized()) or torch.distributed.get_rank() == 0
        if not (f_bin.is_file() and f_meta.is_file() and f_in.is_file()) and rank0:
            start = time.time()
            Path(stem).parent.mkdir(parents=True, exist_ok=True)
            targets = self._targets()
            index = blend_index(targets)
            tmp = Path(stem + ".bin.tmp")
            index.tofile(tmp)
            tmp.rename(f_bin)
            f_in.write_text(json.dumps({"counts": targets.tolist()}))
            f_meta.write_text(json.dumps({"dtype": "int64", "shape": list(index.shape)}))
            logger.info(f"{self.__class__.__name__} blend index for seed {seed} built in {time.time() - start:.2f}s")
        waited = 0
        while not (f_bin.is_file() and f_meta.is_file() and f_in.is_file()):
            time.sleep(0.5)
            waited += 1
            if waited % 120 == 0:
                logger.info(f"{self.__class__.__name__} waiting for blend index {stem} ({waited / 120:.0f} min)")
        meta = json.loads(f_meta.read_text())
        shape = tuple(meta["shape"])
        if self.config.load_dataset_indices_to_memory:
            self.dataset_indices = np.fromfile(f_bin, dtype=np.dtype(meta["dtype"])).reshape(shape)
        else:
            self.dataset_indices = np.memmap(f_bin, mode="r", order="C", dtype=np.dtype(meta["dtype"]), shape=shape)
        self.size = shape[0]
        # same observable state as the reference (blended_dataset.py:372-384): an in-memory index is permuted in place
        # (random_index None), a memory-mapped one is read through random_index (identity when not shuffling)
        do_shuffle = self.config.shuffle_dataset_indices and shuffle
        if self.config.load_dataset_indices_to_memory:
            if do_shuffle:
                np.random.RandomState(seed=seed).shuffle(self.dataset_indices)
            self.random_index = None
        else:
            self.random_index = np.arange(self.size)
            if do_shuffle:
                np.random.RandomState(seed=seed).shuffle(self.random_index)

    def __len__(self) -> int:
        return self.size

    def __getitem__(self, index: int) -> Any:
        if self.dataset_indices is None:
            ds = self.datasets[0]
            return ds[index % len(ds)]
        if self.random_index is not None:
            index = int(self.random_index[index])
        d, i = (int(x) for x in self.dataset_indices[index])
        ds = self.datasets[d]
        return ds[i % len(ds)]

    def collate(self, batch: List[Any]) -> Any:
        return self.datasets[0].collate(batch)

    @staticmethod
    def sync_batch_to_model_parallel(topology: Any, batch: Optional[Any]) -> Any:
        raise NotImplementedError

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}_{self.datasets[0].__class__.__name__}"
