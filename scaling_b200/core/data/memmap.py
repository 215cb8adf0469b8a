"""Token storage: memory-mapped and plain-file readers of the ``{prefix}.bin/.idx/.meta.json`` format.

On-disk contract (reference: core/data/memory_map.py:1-250, file_dataset.py:1-196, file_handles.py:1-70):
``.bin`` = all documents back to back in ``dtype``; ``.idx`` = ``[start, size]`` pairs in ``index_dtype``;
``.meta.json`` = ``{"dtype", "index_dtype", "document_count"}``.  Files written by either framework are readable by
the other.  ``FileDataset`` reads with ``np.fromfile`` behind a retry loop (ESTALE / short reads with exponential
back-off), so network-filesystem hiccups do not SIGBUS the trainer the way an mmap would.
"""
from __future__ import annotations

import json
import time
from errno import ESTALE
from pathlib import Path
from typing import IO, Any, Callable, Iterator, Optional, TypeVar, Union

import numpy as np

V = TypeVar("V")


class _PrefixFiles:
    prefix_path: Path

    @property
    def file_path_data(self) -> Path:
        return Path(str(self.prefix_path) + ".bin")

    @property
    def file_path_index(self) -> Path:
        return Path(str(self.prefix_path) + ".idx")

    @property
    def file_path_meta(self) -> Path:
        return Path(str(self.prefix_path) + ".meta.json")

    def _read_meta(self) -> None:
        for f in (self.file_path_data, self.file_path_index, self.file_path_meta):
            assert f.is_file(), f"cannot initialize dataset, file not found: {f}"
        meta = json.loads(self.file_path_meta.read_text())
        self.dtype = np.dtype(meta["dtype"])
        self.index_dtype = np.dtype(meta["index_dtype"])
        self.document_count = int(meta["document_count"])


class MemoryMapDataset(_PrefixFiles):
    """Random access to documents through ``numpy.memmap`` — small slices of files far larger than RAM."""

    def __init__(self, prefix_path: Union[str, Path], load_index_to_memory: bool = False) -> None:
        self.prefix_path = Path(prefix_path)
        self.load_index_to_memory = load_index_to_memory
        self._read_meta()
        self._data = np.memmap(self.file_path_data, mode="r", order="C", dtype=self.dtype)
        index = np.memmap(self.file_path_index, mode="r", order="C", dtype=self.index_dtype)
        index = index[: 2 * self.document_count].reshape(self.document_count, 2)
        self._index = np.array(index) if load_index_to_memory else index

    def __len__(self) -> int:
        return self.document_count

    def sizes(self, idx: Optional[int] = None) -> np.ndarray:
        if idx is None:
            return np.asarray(self._index[:, 1])
        return np.asarray(self._index[idx, 1])

    def offsets(self) -> np.ndarray:
        return np.asarray(self._index[:, 0])

    def tokens(self) -> np.ndarray:
        """The whole flat token array (memory mapped)."""
        return self._data

    def __getitem__(self, idx: int) -> np.ndarray:
        if not isinstance(idx, (int, np.integer)):
            raise NotImplementedError
        if not 0 <= idx < self.document_count:
            raise IndexError(f"cannot retrieve document idx {idx} from {self.document_count} documents")
        start, size = (int(x) for x in self._index[idx])
        return self._data[start:start + size]

    def __iter__(self) -> Iterator[np.ndarray]:
        for i in range(len(self)):
            yield self[i]

    def close(self) -> None:
        for name in ("_data", "_index"):
            arr = getattr(self, name, None)
            mm = getattr(arr, "_mmap", None)
            if mm is not None:
                try:
                    mm.close()
                except Exception:  # noqa: BLE001
                    pass
            if hasattr(self, name):
                delattr(self, name)


class MemoryMapDatasetBuilder(_PrefixFiles):
    """Appends 1-D arrays as documents; ``finalize`` writes the meta file. Usable as a context manager."""

    def __init__(self, prefix_path: Union[str, Path], dtype: Any = np.int32, index_dtype: Any = np.int64):
        self.prefix_path = Path(prefix_path)
        self.dtype = np.dtype(dtype)
        self.index_dtype = np.dtype(index_dtype)
        assert not self.file_path_data.is_file(), f"data file already exists: {self.file_path_data}"
        assert not self.file_path_index.is_file(), f"index file already exists: {self.file_path_index}"
        self.file_path_data.parent.mkdir(parents=True, exist_ok=True)
        self._data_file: Optional[IO[bytes]] = open(self.file_path_data, "wb")
        self._index_file: Optional[IO[bytes]] = open(self.file_path_index, "wb")
        self.current_index = 0
        self.document_count = 0

    def add(self, np_array: np.ndarray) -> None:
        assert np_array.ndim == 1, "cannot add arrays of more than one dimension"
        assert self._data_file is not None and self._index_file is not None, "builder has been finalized"
        arr = np.ascontiguousarray(np_array.astype(self.dtype))
        self._data_file.write(arr.tobytes(order="C"))
        self._index_file.write(np.array([self.current_index, len(arr)], dtype=self.index_dtype).tobytes(order="C"))
        self.current_index += len(arr)
        self.document_count += 1

    def finalize(self) -> None:
        assert self._data_file is not None and self._index_file is not None, "The Builder has been finalized already"
        self._data_file.close()
        self._index_file.close()
        self._data_file = self._index_file = None
        meta = {"dtype": self.dtype.name, "index_dtype": self.index_dtype.name, "document_count": self.document_count}
        self.file_path_meta.write_text(json.dumps(meta))

    def __enter__(self) -> "MemoryMapDatasetBuilder":
        return self

    def __exit__(self, *exc: Any) -> bool:
        if self._data_file is not None:
            self.finalize()
        return False

    def __del__(self) -> None:
        for f in (getattr(self, "_data_file", None), getattr(self, "_index_file", None)):
            if f is not None:
                f.close()


# --------------------------------------------------------------------------------------------------------------
class RetryableException(Exception):
    pass


def is_retryable(e: Exception) -> bool:
    if isinstance(e, OSError):
        return e.errno == ESTALE
    return isinstance(e, RetryableException)


class FileHandle:
    """File handle whose operations are retried on stale-handle errors (network filesystems)."""

    def __init__(self, path: Union[str, Path], mode: str = "rb"):
        self._path = Path(path)
        self._mode = mode
        self._handle: Optional[IO[Any]] = None

    def retry_operation(self, func: Callable[[IO[Any]], V], max_attempts: int = 5, max_delay: int = 32) -> V:
        for attempt in range(1, max_attempts + 1):
            try:
                if self._handle is None:
                    self._handle = open(self._path, self._mode)
                return func(self._handle)
            except Exception as e:  # noqa: BLE001
                if not is_retryable(e):
                    raise
                try:
                    self.close()
                except Exception as e_close:  # noqa: BLE001
                    print(f"closing {self._path} failed: {e_close}", flush=True)
                print(f"retryable error on {self._path}: {e} (attempt {attempt}/{max_attempts})", flush=True)
                if attempt < max_attempts:
                    time.sleep(min(2 ** attempt, max_delay))
        raise Exception(f"Stale file handle even after {max_attempts} retries for {self._path}.")

    def close(self) -> None:
        if self._handle is not None:
            self._handle.close()
            self._handle = None

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def retry_array_from_file(file_handle: FileHandle, dtype: Any, count: int, offset: int, max_attempts: int = 5,
                          max_delay: int = 32) -> np.ndarray:
    """``count`` items of ``dtype`` at BYTE ``offset`` of the file behind ``file_handle``, re-opened and retried when the
    read comes back empty (``np.fromfile`` does not look at errno: a stale network-filesystem handle shows as a
    zero-length result) — reference file_dataset.py:151-199.  A short but non-empty read is an error, not a retry."""

    def op(f: IO[Any]) -> np.ndarray:
        res = np.fromfile(f, dtype=dtype, count=count, offset=offset)
        f.seek(0)  # fromfile leaves the position behind the data; the handle is reused with absolute offsets
        if len(res) == 0 and count != 0:
            raise RetryableException(f"expected {count} items of {dtype} at offset {offset} of {file_handle._path}, got none")
        if len(res) != count:
            raise Exception(f"expected {count} items of {dtype} at offset {offset} of {file_handle._path}, got {len(res)}")
        return res

    return file_handle.retry_operation(op, max_attempts=max_attempts, max_delay=max_delay)


class FileDataset(_PrefixFiles):
    """Same format as :class:`MemoryMapDataset`, read with explicit file IO and retries."""

    def __init__(self, prefix_path: Union[str, Path], load_index_to_memory: bool = False) -> None:
        self.prefix_path = Path(prefix_path)
        self.load_index_to_memory = load_index_to_memory
        self._read_meta()
        self._data_handle = FileHandle(self.file_path_data)
        self._index_handle = FileHandle(self.file_path_index)
        self._index: Optional[np.ndarray] = None
        if load_index_to_memory:
            self._index = self._read(self._index_handle, self.index_dtype, 2 * self.document_count, 0).reshape(-1, 2)

    @staticmethod
    def _read(handle: FileHandle, dtype: np.dtype, count: int, offset_items: int) -> np.ndarray:
        return retry_array_from_file(handle, dtype, count, offset_items * np.dtype(dtype).itemsize)

    def __len__(self) -> int:
        return self.document_count

    def _index_row(self, idx: int) -> tuple[int, int]:
        if self._index is not None:
            s, n = self._index[idx]
        else:
            s, n = self._read(self._index_handle, self.index_dtype, 2, 2 * idx)
        return int(s), int(n)

    def sizes(self, idx: Optional[int] = None) -> np.ndarray:
        if idx is not None:
            return np.asarray(self._index_row(idx)[1])
        if self._index is not None:
            return self._index[:, 1]
        full = self._read(self._index_handle, self.index_dtype, 2 * self.document_count, 0).reshape(-1, 2)
        return full[:, 1]

    def __getitem__(self, idx: int) -> np.ndarray:
        if not 0 <= idx < self.document_count:
            raise IndexError(f"cannot retrieve document idx {idx} from {self.document_count} documents")
        start, size = self._index_row(idx)
        return self._read(self._data_handle, self.dtype, size, start)

    def __iter__(self) -> Iterator[np.ndarray]:
        for i in range(len(self)):
            yield self[i]

    def close(self) -> None:
        self._data_handle.close()
        self._index_handle.close()
