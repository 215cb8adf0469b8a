"""Megatron-era dataset formats (reference: transformer/data/legacy_dataset/indexed_dataset.py:1-476,
legacy_blended_dataset.py:1-391).  Byte-for-byte the on-disk layouts, so existing corpora and caches keep working:

``MMIDIDX`` (``impl="mmap"``): ``{prefix}.idx`` = magic ``MMIDIDX\\0\\0`` (9 B) · version u64=1 · dtype code u8 ·
n_items u64 · n_docs u64 · sizes i32[n] · byte pointers i64[n] · doc_idx i64[n_docs]; ``{prefix}.bin`` = the items
back to back.  ``TNTIDX`` (``impl="lazy"``/``"cached"``): magic ``TNTIDX\\0\\0`` (8 B) · version u64=1 · dtype code
u64 · element size u64 · n_items u64 · n_sizes u64 · n_docs u64 · dim_offsets i64[n+1] · data_offsets i64[n+1] ·
sizes i64[n_sizes] · doc_idx i64[n_docs].

:class:`LegacyBlendedDataset` keeps the old blend semantics — the greedy interleave stops as soon as the FIRST dataset
has delivered its quota (the current blend runs until all have) and its cache is an ``MMIDIDX`` pair plus a
``.done`` marker — but builds the index with the native C++ ``sb_blend_index`` instead of a Python loop.
"""
from __future__ import annotations

import hashlib
import struct
import time
from pathlib import Path
from typing import Any, List, Optional, Sequence, Union

import numpy as np
import torch

from ...core import BaseBlendedDataset, BlendedDatasetConfig
from ...core.data.blended import blend_index, weights_by_num_docs, weights_examples_proportional
from .text_dataset import TextDataset
from .text_dataset_batch import TextDatasetBatch, TextDatasetBatchBeforeSync

DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float32, 7: np.float64, 8: np.uint16}
MMAP_MAGIC = b"MMIDIDX\x00\x00"
LAZY_MAGIC = b"TNTIDX\x00\x00"


def code(dtype: Any) -> int:
    for k, v in DTYPES.items():
        if np.dtype(v) == np.dtype(dtype):
            return k
    raise ValueError(dtype)


def index_file_path(prefix_path: str) -> str:
    return prefix_path + ".idx"


def data_file_path(prefix_path: str) -> str:
    return prefix_path + ".bin"


def best_fitting_dtype(vocab_size: Optional[int] = None) -> Any:
    return np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32


def infer_dataset_impl(path: str) -> Optional[str]:
    if not IndexedDataset.exists(path):
        return None
    with open(index_file_path(path), "rb") as f:
        magic = f.read(9)
    if magic[:8] == LAZY_MAGIC:
        return "cached"
    if magic == MMAP_MAGIC:
        return "mmap"
    return None


class _Sizes(np.ndarray):
    """Item sizes, readable both as the reference's ``ds.sizes`` property and as the ``ds.sizes()`` accessor of the
    memory-map datasets."""

    def __call__(self) -> np.ndarray:
        return np.asarray(self)


def _as_sizes(a: np.ndarray) -> _Sizes:
    return np.asarray(a).view(_Sizes)


# ------------------------------------------------------------------------------------------------ MMIDIDX
class Index:
    """Reader / writer of the ``MMIDIDX`` index file."""

    _HDR_MAGIC = MMAP_MAGIC

    @staticmethod
    def write(path: str, dtype: Any, sizes: Sequence[int], doc_idx: Sequence[int]) -> None:
        sizes_a = np.asarray(sizes, dtype=np.int32)
        pointers = np.zeros(len(sizes_a), dtype=np.int64)
        if len(sizes_a) > 1:
            np.cumsum(sizes_a[:-1].astype(np.int64) * np.dtype(dtype).itemsize, out=pointers[1:])
        with open(path, "wb") as f:
            f.write(MMAP_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", code(dtype)))
            f.write(struct.pack("<Q", len(sizes_a)))
            f.write(struct.pack("<Q", len(doc_idx)))
            f.write(sizes_a.tobytes(order="C"))
            f.write(pointers.tobytes(order="C"))
            f.write(np.asarray(doc_idx, dtype=np.int64).tobytes(order="C"))

    def __init__(self, path: str, skip_warmup: bool = False):
        with open(path, "rb") as f:
            assert f.read(9) == MMAP_MAGIC, ("Index file doesn't match expected format. "
                                             "Make sure that --dataset-impl is configured properly.")
            assert struct.unpack("<Q", f.read(8)) == (1,)
            (dtype_code,) = struct.unpack("<B", f.read(1))
            self._dtype = DTYPES[dtype_code]
            (self._len,) = struct.unpack("<Q", f.read(8))
            (self._doc_count,) = struct.unpack("<Q", f.read(8))
            offset = f.tell()
        self._mmap = np.memmap(path, mode="r", order="C")
        buf = memoryview(self._mmap)
        self._sizes = np.frombuffer(buf, dtype=np.int32, count=self._len, offset=offset)
        self._pointers = np.frombuffer(buf, dtype=np.int64, count=self._len, offset=offset + self._sizes.nbytes)
        self._doc_idx = np.frombuffer(buf, dtype=np.int64, count=self._doc_count,
                                      offset=offset + self._sizes.nbytes + self._pointers.nbytes)

    @property
    def dtype(self) -> Any:
        return self._dtype

    @property
    def sizes(self) -> np.ndarray:
        return self._sizes

    @property
    def doc_idx(self) -> np.ndarray:
        return self._doc_idx

    def __getitem__(self, i: int):
        return self._pointers[i], self._sizes[i]

    def __len__(self) -> int:
        return self._len


class MMapIndexedDataset(torch.utils.data.Dataset):
    def __init__(self, path: str, skip_warmup: bool = False):
        self._do_init(str(path))

    def __getstate__(self) -> str:
        return self._path

    def __setstate__(self, state: str) -> None:
        self._do_init(state)

    def _do_init(self, path: str) -> None:
        self._path = path
        self._index = Index(index_file_path(path))
        self._bin = np.memmap(data_file_path(path), mode="r", order="C")

    def __len__(self) -> int:
        return len(self._index)

    def get(self, idx: int, offset: int = 0, length: Optional[int] = None) -> np.ndarray:
        """Tokens [offset, offset + length) of item ``idx`` without touching the rest of it."""
        ptr, size = self._index[idx]
        if length is None:
            length = int(size) - offset
        ptr = int(ptr) + offset * np.dtype(self._index.dtype).itemsize
        return np.frombuffer(memoryview(self._bin), dtype=self._index.dtype, count=length, offset=ptr)

    def __getitem__(self, idx: Union[int, slice]) -> Union[np.ndarray, List[np.ndarray]]:
        if isinstance(idx, (int, np.integer)):
            return self.get(int(idx))
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            return [self.get(i) for i in range(start, stop)]
        raise TypeError(f"unsupported index {idx!r}")

    @property
    def sizes(self) -> _Sizes:
        return _as_sizes(self._index.sizes)

    @property
    def doc_idx(self) -> np.ndarray:
        return self._index.doc_idx

    @staticmethod
    def exists(path: str) -> bool:
        return Path(index_file_path(path)).is_file() and Path(data_file_path(path)).is_file()


class MMapIndexedDatasetBuilder:
    def __init__(self, out_file: str, dtype: Any = np.int64):
        self._data_file = open(out_file, "wb")
        self._dtype = dtype
        self._sizes: List[int] = []
        self._doc_idx: List[int] = [0]

    def add_item(self, tensor: Union[torch.Tensor, np.ndarray]) -> None:
        a = np.asarray(tensor.numpy() if torch.is_tensor(tensor) else tensor, dtype=self._dtype)
        self._data_file.write(a.tobytes(order="C"))
        self._sizes.append(a.size)

    def end_document(self) -> None:
        self._doc_idx.append(len(self._sizes))

    def finalize(self, index_file: str) -> None:
        self._data_file.close()
        Index.write(index_file, self._dtype, self._sizes, self._doc_idx)


# ------------------------------------------------------------------------------------------------ TNTIDX
class IndexedDataset(torch.utils.data.Dataset):
    """Reader of the older ``TNTIDX`` format; items are read from the data file on access."""

    _HDR_MAGIC = LAZY_MAGIC

    def __init__(self, path: str):
        self.path = path
        with open(index_file_path(path), "rb") as f:
            assert f.read(8) == LAZY_MAGIC, ("Index file doesn't match expected format. "
                                             "Make sure that --dataset-impl is configured properly.")
            assert struct.unpack("<Q", f.read(8)) == (1,)
            dtype_code, self.element_size = struct.unpack("<QQ", f.read(16))
            self.dtype = DTYPES[dtype_code]
            self._len, n_sizes = struct.unpack("<QQ", f.read(16))
            (n_docs,) = struct.unpack("<Q", f.read(8))
            self.dim_offsets = np.fromfile(f, dtype=np.int64, count=self._len + 1)
            self.data_offsets = np.fromfile(f, dtype=np.int64, count=self._len + 1)
            self._sizes = np.fromfile(f, dtype=np.int64, count=n_sizes)
            self.doc_idx = np.fromfile(f, dtype=np.int64, count=n_docs)
        self._data = np.memmap(data_file_path(path), mode="r", dtype=self.dtype, order="C")

    @property
    def sizes(self) -> _Sizes:
        return _as_sizes(self._sizes)

    def _item(self, i: int) -> np.ndarray:
        if i < 0 or i >= self._len:
            raise IndexError("index out of range")
        shape = self._sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]]
        flat = np.asarray(self._data[self.data_offsets[i]:self.data_offsets[i + 1]])
        return flat.reshape(tuple(int(s) for s in shape))

    def __getitem__(self, idx: Union[int, slice]) -> Union[np.ndarray, List[np.ndarray]]:
        if isinstance(idx, (int, np.integer)):
            return self._item(int(idx))
        start, stop, step = idx.indices(len(self))
        if step != 1:
            raise ValueError("Slices into indexed_dataset must be contiguous")
        return [self._item(i) for i in range(start, stop)]

    def __len__(self) -> int:
        return self._len

    def size(self, index: int) -> np.ndarray:
        return self._sizes[self.dim_offsets[index]:self.dim_offsets[index + 1]]

    @staticmethod
    def exists(path: str) -> bool:
        return Path(index_file_path(path)).is_file() and Path(data_file_path(path)).is_file()


class IndexedCachedDataset(IndexedDataset):
    """``impl="cached"``: the whole data file is read into RAM once."""

    def __init__(self, path: str):
        super().__init__(path)
        self._data = np.fromfile(data_file_path(path), dtype=self.dtype)


class IndexedDatasetBuilder:
    def __init__(self, out_file: str, dtype: Any = np.int32):
        self.out_file = open(out_file, "wb")
        self.dtype = dtype
        self.element_size = np.dtype(dtype).itemsize
        self.data_offsets, self.dim_offsets, self.sizes, self.doc_idx = [0], [0], [], [0]

    def add_item(self, tensor: Union[torch.Tensor, np.ndarray]) -> None:
        a = np.asarray(tensor.numpy() if torch.is_tensor(tensor) else tensor, dtype=self.dtype)
        self.out_file.write(a.tobytes(order="C"))
        self.data_offsets.append(self.data_offsets[-1] + a.size)
        self.sizes.extend(a.shape)
        self.dim_offsets.append(self.dim_offsets[-1] + a.ndim)

    def end_document(self) -> None:
        self.doc_idx.append(len(self.sizes))

    def finalize(self, index_file: str) -> None:
        self.out_file.close()
        with open(index_file, "wb") as f:
            f.write(LAZY_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<QQ", code(self.dtype), self.element_size))
            f.write(struct.pack("<QQ", len(self.data_offsets) - 1, len(self.sizes)))
            f.write(struct.pack("<Q", len(self.doc_idx)))
            for arr in (self.dim_offsets, self.data_offsets, self.sizes, self.doc_idx):
                f.write(np.asarray(arr, dtype=np.int64).tobytes(order="C"))


def make_builder(out_file: str, impl: str, vocab_size: Optional[int] = None):
    if impl == "mmap":
        return MMapIndexedDatasetBuilder(out_file, dtype=best_fitting_dtype(vocab_size))
    return IndexedDatasetBuilder(out_file)


def make_dataset(path: str, impl: str, skip_warmup: bool = False):
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        print("Path should be a basename that both .idx and .bin can be appended to get full filenames.")
        return None
    if impl == "infer":
        impl = infer_dataset_impl(path) or "mmap"
    if impl == "lazy":
        return IndexedDataset(path)
    if impl == "cached":
        return IndexedCachedDataset(path)
    if impl == "mmap":
        return MMapIndexedDataset(path, skip_warmup)
    print(f"Unknown dataset implementation: {impl}")
    return None


def get_indexed_dataset_(data_prefix: str, data_impl: str, skip_warmup: bool):
    return make_dataset(data_prefix, data_impl, skip_warmup)


# ------------------------------------------------------------------------------------------------ legacy blend
class LegacyBlendedDataset(BaseBlendedDataset[Any, TextDatasetBatchBeforeSync, TextDatasetBatch, Any]):
    def __init__(self, seed: int, config: BlendedDatasetConfig, datasets: Sequence[TextDataset], shuffle: bool = True):
        self.ep_maximum_dict: Optional[dict] = None
        self.weights: Optional[np.ndarray] = (np.asarray(config.weights, dtype=np.float64)
                                              if config.weights is not None else None)
        self.num_datasets = len(datasets)
        self.dataset_indices_: Any = None
        super().__init__(seed=seed, config=config, datasets=datasets, shuffle=shuffle)

    # cache naming kept so indices written by the reference are found again
    def get_data_index_cache_filename_stem(self, seed: int) -> str:
        assert self.config.cache_directory is not None, "cache directory is needed"
        Path(self.config.cache_directory).mkdir(exist_ok=True, parents=True)
        assert self.weights is not None
        prefix_str = "-".join(Path(d.data_prefix).name for d in self.datasets)
        weight_str = "-".join(str(round(w * 100) / 100) for w in self.weights.tolist())
        prefix_hash = hashlib.md5(prefix_str.encode("utf-8")).hexdigest()
        weight_hash = hashlib.md5(weight_str.encode("utf-8")).hexdigest()
        return str(Path(self.config.cache_directory)
                   / f"index_cache_blended_dataset_seed_{seed}_seq_len_{self.datasets[0].sequence_length}"
                     f"_prefix_{prefix_hash}_weights_{weight_hash}")

    def _resolve_weights(self, sizes: List[int]) -> np.ndarray:
        c = self.config
        if c.weight_by_num_documents:
            if c.weight_examples_proportional:
                assert self.ep_maximum_dict is not None or c.ep_maximum is not None, \
                    "ep_maximum should be defined when using examples_proportional sampling"
                if self.ep_maximum_dict is not None:  # per-dataset rate limits keyed by the prefix stem
                    limits = [self.ep_maximum_dict.get(Path(d.data_prefix).stem, c.ep_maximum if c.ep_maximum is not None else n)
                              for d, n in zip(self.datasets, sizes)]
                    capped = np.minimum(np.asarray(sizes, dtype=np.float64), np.asarray(limits, dtype=np.float64))
                    p = np.asarray(sizes, dtype=np.float64) / float(np.sum(sizes))
                    r = capped / capped.sum()
                    if c.ep_temperature != 1.0:
                        r = r ** (1.0 / c.ep_temperature)
                        r = r / r.sum()
                    w = r / p
                    return w / w.sum()
                return weights_examples_proportional(sizes, c.ep_temperature, c.ep_maximum)
            return weights_by_num_docs(sizes, c.weighted_sampler_alpha)
        assert self.weights is not None, "explicit weights are needed when weight_by_num_documents is off"
        w = np.asarray(self.weights, dtype=np.float64)
        assert w.sum() > 0.0
        return w / w.sum()

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        assert shuffle, "Blended datasets should always be shuffled"
        if seed == self.seed:
            return
        self.seed = seed
        for d in self.datasets:
            d.set_seed(seed=seed, shuffle=True)
        if self.num_datasets == 1:
            self.size = len(self.datasets[0])
            return
        sizes = [len(d) for d in self.datasets]
        self.weights = self._resolve_weights(sizes)
        stem = self.get_data_index_cache_filename_stem(seed)
        f_bin, f_idx, f_done = Path(stem + ".bin"), Path(stem + ".idx"), Path(stem + ".done")
        rank0 = (not torch.distributed.is_initialized()) or torch.distributed.get_rank() == 0
        if not f_done.is_file() and rank0:
            start = time.time()
            rel = self.weights / self.weights.max()
            rnd = (lambda v: int(round(v))) if self.config.weight_examples_proportional else int
            targets = np.asarray([max(1, rnd(p * n)) for n, p in zip(sizes, rel)], dtype=np.int64)
            index = blend_index(targets)
            # old semantics: stop at the first dataset that has delivered its quota
            done_at = np.full(len(targets), len(index), dtype=np.int64)
            for d, t in enumerate(targets):
                hits = np.flatnonzero((index[:, 0] == d) & (index[:, 1] == t - 1))
                if len(hits):
                    done_at[d] = hits[0] + 1
            index = index[:int(done_at.min())]
            builder = MMapIndexedDatasetBuilder(str(f_bin), dtype=np.int32)
            builder._data_file.write(index.astype(np.int32).tobytes(order="C"))
            builder._sizes = [2] * len(index)
            builder.finalize(str(f_idx))
            f_done.write_text("True")
            print(f"{self.__class__.__name__} compute_data_index done in {time.time() - start:.2f} seconds for seed {seed}",
                  flush=True)
        waited = 0
        while not (f_bin.is_file() and f_idx.is_file() and f_done.is_file()):
            time.sleep(1)
            waited += 1
            if waited % 60 == 0:
                print(f"BlendedDataset waiting on index for seed {seed}; elapsed {waited / 60:.0f} minutes", flush=True)
        self.dataset_indices_ = MMapIndexedDataset(stem, True)
        self.size = len(self.dataset_indices_)

    def __len__(self) -> int:
        return max(self.size, self.config.minimum_dataset_size)

    def __getitem__(self, idx: int) -> Any:
        if self.size < self.config.minimum_dataset_size:
            idx %= self.size
        if self.num_datasets == 1:
            return self.datasets[0][idx]
        d, i = (int(v) for v in self.dataset_indices_[idx])
        return self.datasets[d][i]

    @staticmethod
    def sync_batch_to_model_parallel(topology, batch):
        return TextDataset.sync_batch_to_model_parallel(topology, batch)
