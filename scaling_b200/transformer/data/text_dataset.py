"""Packed language-modelling dataset (reference: transformer/data/text_dataset.py:1-462).

Shuffled documents are packed back to back into fixed windows of ``sequence_length + 1`` tokens without padding;
documents may straddle windows and consecutive windows of one document overlap by the one token that is the last
target of one and the first input of the next.  The window index is cached next to the data (built by global rank 0,
the others poll), so start-up after the first run is immediate.  ``only_full_sequences`` starts every window at a
document start (optionally relaxed every n-th window).
"""
from __future__ import annotations

import ctypes
import hashlib
import json
import time
from pathlib import Path
from typing import List, Optional, Union

import numpy as np
import torch

from ...core import BaseBlendedDataset, BaseDataset, BaseDatasetItem, FileDataset, MemoryMapDataset, MemoryMapDatasetBuilder, broadcast_data
from ...core.topology import Topology
from .text_dataset_batch import TextDatasetBatch, TextDatasetBatchBeforeSync


class TextDatasetItem(BaseDatasetItem):
    def __init__(self, token_ids: torch.Tensor):
        self.token_ids = token_ids


def build_window_index(sizes: np.ndarray, order: np.ndarray, sequence_length: int, only_full_sequences: bool = False,
                       allow_incomplete_every_n: int = 0) -> List[np.ndarray]:
    """List of windows; each window is a flat int64 array of (document, start, end) triples covering seq+1 tokens."""
    windows: List[np.ndarray] = []
    current: List[int] = []
    filled = 0
    want = sequence_length + 1
    full = relaxed = 0
    relaxing = False
    for doc in order.tolist():
        n = int(sizes[doc])
        pos = 0
        while pos < n - 1:  # a piece needs at least one (input, target) pair
            end = min(n, pos + want - filled)
            if only_full_sequences and not relaxing:
                if end - pos < want:
                    if allow_incomplete_every_n and full / allow_incomplete_every_n - relaxed >= 1:
                        relaxing = True
                    else:
                        break
                else:
                    full += 1
            current += [doc, pos, end]
            filled += end - pos
            if filled == want:
                windows.append(np.asarray(current, dtype=np.int64))
                current, filled = [], 0
                if relaxing:
                    relaxed += 1
                relaxing = False
            pos = end - 1
    return windows


class TextDataset(BaseDataset[TextDatasetItem, TextDatasetBatchBeforeSync, TextDatasetBatch]):
    def __init__(self, data_prefix: Union[str, Path], sequence_length: int, seed: int, legacy_dataset: bool = False,
                 load_mmap_index_to_memory: bool = False, load_data_item_mmap_index_to_memory: bool = False,
                 only_full_sequences: bool = False, allow_incomplete_sequences_every_n: int = 0, shuffle: bool = True,
                 use_mmap: bool = True):
        self.data_prefix = Path(data_prefix)
        self.sequence_length = sequence_length
        self.legacy_dataset = legacy_dataset
        self.only_full_sequences = only_full_sequences
        self.allow_incomplete_sequences_every_n = allow_incomplete_sequences_every_n
        self.load_data_item_mmap_index_to_memory = load_data_item_mmap_index_to_memory
        self.use_mmap = use_mmap
        if legacy_dataset:
            from .legacy import MMapIndexedDataset

            assert not (only_full_sequences or load_mmap_index_to_memory), "not supported for legacy datasets"
            self.memory_map = MMapIndexedDataset(str(data_prefix))
        elif use_mmap:
            self.memory_map = MemoryMapDataset(self.data_prefix, load_index_to_memory=load_mmap_index_to_memory)
        else:
            self.memory_map = FileDataset(self.data_prefix, load_index_to_memory=load_mmap_index_to_memory)
        self.data_item_index: Optional[Union[MemoryMapDataset, FileDataset]] = None
        super().__init__(seed=seed, shuffle=shuffle)

    def ident(self) -> str:
        return f"{hashlib.md5(str(self.data_prefix).encode('utf-8')).hexdigest()}-seq-{self.sequence_length}"

    def _cache_stem(self, seed: int, shuffle: bool) -> str:
        stem = f"{self.data_prefix}_window_index_seed_{seed}_seq_{self.sequence_length}"
        if not shuffle:
            stem += "_ordered"
        if self.only_full_sequences:
            stem += f"_full_every_{self.allow_incomplete_sequences_every_n}"
        return stem

    def set_seed(self, seed: int, shuffle: bool = True) -> None:
        if self.seed is not None and self.seed == seed:
            return
        self.seed = seed
        stem = self._cache_stem(seed, shuffle)
        meta = Path(stem + ".meta.json")
        rank0 = (not torch.distributed.is_initialized()) or torch.distributed.get_rank() == 0
        if not meta.is_file() and rank0:
            for suffix in (".bin", ".idx"):
                Path(stem + suffix).unlink(missing_ok=True)
            order = np.arange(len(self.memory_map))
            if shuffle:
                np.random.RandomState(seed=seed).shuffle(order)
            sizes = np.asarray(self.memory_map.sizes())
            builder = MemoryMapDatasetBuilder(Path(stem), dtype=np.int64, index_dtype=np.int64)
            for w in build_window_index(sizes, order, self.sequence_length, self.only_full_sequences,
                                        self.allow_incomplete_sequences_every_n):
                builder.add(w)
            builder.finalize()
        waited = 0.0
        while not (meta.is_file() and Path(stem + ".bin").is_file() and Path(stem + ".idx").is_file()):
            time.sleep(0.25)
            waited += 0.25
            if waited % 60 == 0:
                print(f"TextDataset waiting for window index {stem} ({waited / 60:.0f} min)", flush=True)
        cls = MemoryMapDataset if self.use_mmap else FileDataset
        self.data_item_index = cls(Path(stem), load_index_to_memory=self.load_data_item_mmap_index_to_memory)

    def __len__(self) -> int:
        assert self.data_item_index is not None
        return len(self.data_item_index)

    def __getitem__(self, index: int) -> TextDatasetItem:
        assert self.data_item_index is not None, "data item index not set"
        pieces = np.asarray(self.data_item_index[index]).reshape(-1, 3)
        native = self._native_gather(pieces)
        if native is not None:
            return TextDatasetItem(token_ids=torch.from_numpy(native.astype(np.int64)))
        out = np.empty(self.sequence_length + 1, dtype=np.int64)
        w = 0
        for doc, start, end in pieces.tolist():
            n = end - start
            out[w:w + n] = self.memory_map[doc][start:end]
            w += n
        return TextDatasetItem(token_ids=torch.from_numpy(out))

    def _native_gather(self, pieces: np.ndarray) -> Optional[np.ndarray]:
        """The window's (document, start, end) pieces copied by the C++ ``sb_gather_window_i32`` straight out of the
        memory-mapped int32 token file (the reference loops over the pieces in Python,
        text_dataset.py:371-385).  None when the corpus is not an int32 memory map or the library is unavailable."""
        state = getattr(self, "_gather_state", None)
        if state is None:
            state = False
            if isinstance(self.memory_map, MemoryMapDataset) and not self.legacy_dataset:
                try:
                    import ctypes

                    from ...ops._lib import host_lib

                    tokens = self.memory_map.tokens()
                    if tokens.dtype == np.int32:
                        offsets = np.ascontiguousarray(self.memory_map.offsets(), dtype=np.int64)
                        state = (host_lib().sb_gather_window_i32, ctypes, tokens, offsets)
                except (RuntimeError, OSError, AttributeError):
                    state = False
            self._gather_state = state
        if state is False:
            return None
        fn, ctypes, tokens, offsets = state
        out = np.empty(self.sequence_length + 1, dtype=np.int32)
        p64 = np.ascontiguousarray(pieces, dtype=np.int64)
        n = fn(tokens.ctypes.data_as(ctypes.c_void_p), offsets.ctypes.data_as(ctypes.c_void_p),
               p64.ctypes.data_as(ctypes.c_void_p), len(p64), out.ctypes.data_as(ctypes.c_void_p), len(out))
        return out if n == len(out) else None

    def collate(self, batch: List[TextDatasetItem]) -> TextDatasetBatchBeforeSync:
        return TextDatasetBatchBeforeSync(token_ids=torch.stack([b.token_ids for b in batch]))

    @staticmethod
    def sync_batch_to_model_parallel(topology: Optional[Topology], batch: Optional[TextDatasetBatchBeforeSync]) -> TextDatasetBatch:
        """TP rank 0 holds the host batch; everyone ends up with inputs/targets on the device."""
        if topology is None:
            assert batch is not None
            tokens = batch.token_ids
        else:
            tensors = [batch.token_ids] if batch is not None else None
            tokens = broadcast_data(tensors, torch.long, topology, n=1)[0]
        return TextDatasetBatch(input_token_ids=tokens[:, :-1].contiguous(), target_token_ids=tokens[:, 1:].contiguous())

    @staticmethod
    def jsonl_to_memory_map(data_file_jsonl: Union[str, Path], prefix_path_memory_map: Union[str, Path], tokenizer=None) -> None:
        """``{"text": ...}`` lines → tokenised memory map (one document per line, EOS appended)."""
        from ..tokenizer import Tokenizer

        tokenizer = tokenizer or Tokenizer.default()
        with MemoryMapDatasetBuilder(Path(prefix_path_memory_map)) as builder, open(data_file_jsonl, encoding="UTF-8") as f:
            for line in f:
                if line.strip():
                    ids = tokenizer.encode(json.loads(line)["text"]) + [tokenizer.eos_token_id]
                    builder.add(np.asarray(ids))


class TextBlendedDataset(BaseBlendedDataset[TextDatasetItem, TextDatasetBatchBeforeSync, TextDatasetBatch, TextDataset]):
    @staticmethod
    def sync_batch_to_model_parallel(topology, batch):
        return TextDataset.sync_batch_to_model_parallel(topology, batch)
