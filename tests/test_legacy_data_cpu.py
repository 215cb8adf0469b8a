"""Megatron-era dataset formats and the legacy blend (reference: tests/transformer/test_data legacy cases)."""
import numpy as np
import pytest
import torch

from scaling_b200.core import BlendedDatasetConfig
from scaling_b200.transformer.data import legacy as L
from scaling_b200.transformer.data.text_dataset import TextDataset


def _write_mmap(prefix, docs, dtype=np.uint16):
    b = L.MMapIndexedDatasetBuilder(str(prefix) + ".bin", dtype=dtype)
    for d in docs:
        b.add_item(torch.tensor(d))
        b.end_document()
    b.finalize(str(prefix) + ".idx")


def test_mmap_indexed_roundtrip(tmp_path):
    docs = [[5, 6, 7, 0], [9, 0], list(range(1, 40)) + [0]]
    _write_mmap(tmp_path / "corpus", docs)
    raw = (tmp_path / "corpus.idx").read_bytes()
    assert raw[:9] == b"MMIDIDX\x00\x00" and raw[17] == 8  # dtype code of uint16
    ds = L.make_dataset(str(tmp_path / "corpus"), "infer")
    assert isinstance(ds, L.MMapIndexedDataset) and len(ds) == 3
    assert [ds[i].tolist() for i in range(3)] == docs
    assert ds.sizes.tolist() == [4, 2, 40] and ds.sizes().tolist() == [4, 2, 40]
    assert ds.get(2, offset=3, length=4).tolist() == [4, 5, 6, 7]
    assert [a.tolist() for a in ds[0:2]] == docs[:2]
    assert ds.doc_idx.tolist() == [0, 1, 2, 3]
    with pytest.raises(ValueError):
        ds[0:3:2]


def test_lazy_and_cached_roundtrip(tmp_path):
    b = L.make_builder(str(tmp_path / "old.bin"), impl="lazy")
    items = [np.arange(6).reshape(2, 3), np.arange(4)]
    for it in items:
        b.add_item(it)
        b.end_document()
    b.finalize(str(tmp_path / "old.idx"))
    assert L.infer_dataset_impl(str(tmp_path / "old")) == "cached"
    for impl in ("lazy", "cached"):
        ds = L.make_dataset(str(tmp_path / "old"), impl)
        assert len(ds) == 2 and ds[0].tolist() == items[0].tolist() and ds[1].tolist() == items[1].tolist()
        assert ds.size(0).tolist() == [2, 3]
    assert L.make_dataset(str(tmp_path / "missing"), "mmap") is None


def test_text_dataset_on_legacy_corpus_and_legacy_blend(tmp_path):
    rng = np.random.RandomState(0)
    for name, n in (("a", 60), ("b", 20)):
        docs = [rng.randint(1, 100, size=rng.randint(5, 30)).tolist() + [0] for _ in range(n)]
        _write_mmap(tmp_path / name, docs)
    dsets = [TextDataset(tmp_path / n, sequence_length=16, seed=1, legacy_dataset=True) for n in ("a", "b")]
    assert all(len(d) > 0 for d in dsets)
    assert dsets[0][0].token_ids.shape == (17,)
    cfg = BlendedDatasetConfig(cache_directory=tmp_path / "cache", weighted_sampler_alpha=1.0)
    blend = L.LegacyBlendedDataset(seed=1, config=cfg, datasets=dsets)
    # alpha = 1: every dataset's quota is its own length; the old loop stops when the first one is exhausted
    idx = np.stack([np.asarray(blend.dataset_indices_[i]) for i in range(len(blend))])
    counts = np.bincount(idx[:, 0], minlength=2)
    assert (counts <= [len(d) for d in dsets]).all() and (counts == [len(d) for d in dsets]).any()
    assert len(blend) < sum(len(d) for d in dsets) or counts.tolist() == [len(d) for d in dsets]
    # within a dataset, indices are consumed in order; the interleave follows the lowest sampled/target ratio
    for d in range(2):
        assert idx[idx[:, 0] == d][:, 1].tolist() == list(range(counts[d]))
    assert blend[0].token_ids.shape == (17,)
    # cached: a second instance reads the same index
    again = L.LegacyBlendedDataset(seed=1, config=cfg, datasets=dsets)
    assert len(again) == len(blend)
    assert (tmp_path / "cache").is_dir() and any(p.suffix == ".done" for p in (tmp_path / "cache").iterdir())
