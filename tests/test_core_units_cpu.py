"""Core unit tests on CPU: config, topology, LR schedule, pipeline partitioning + schedule, datasets, sampler
(reference: tests/core/test_config, test_topology, test_data/*, test_nn/test_pipeline_*)."""
import errno
import json
from enum import Enum
from pathlib import Path
from typing import Optional

import numpy as np
import pytest
import torch
from pydantic import Field, ValidationError

from scaling_b200.core import (
    BaseConfig,
    BaseDataset,
    FileDataset,
    LearningRateScheduler,
    LearningRateSchedulerConfig,
    MemoryMapDataset,
    MemoryMapDatasetBuilder,
    PipelineScheduleInference,
    PipelineScheduleTrain,
    Topology,
    TopologyConfig,
    overwrite_recursive,
    pipe_partition_uniform,
)
from scaling_b200.core.data.blended import blend_index, weights_by_num_docs, weights_examples_proportional
from scaling_b200.core.data.loader import RandomSampler
from scaling_b200.core.data.memmap import FileHandle, RetryableException
from scaling_b200.core.nn.partition import balanced_bounds, pipe_partition_from_indices


# ---------------------------------------------------------------------------------------------------- config
class Color(Enum):
    RED = "red"
    BLUE = "blue"


class Inner(BaseConfig):
    n: int = Field(3, description="a number")
    path: Optional[Path] = Field(None, description="a path")


class Outer(BaseConfig):
    """Outer doc."""

    name: str = Field("x", description="the name\nsecond line")
    color: Color = Field(Color.RED, description="a colour")
    inner: Inner = Field(Inner(), description="nested block")
    required: int = Field(description="must be given")


def test_config_roundtrip_and_immutability(tmp_path):
    c = Outer.from_dict({"required": 1, "inner": {"n": 5, "path": "a/b"}, "color": "blue"})
    assert c.inner.n == 5 and c.inner.path == Path("a/b") and c.color is Color.BLUE
    d = c.as_dict()
    assert d["inner"]["path"] == "a/b" and d["color"] == "blue"
    assert Outer.from_dict(d) == c and Outer.from_str(c.as_str()) == c
    c.save(tmp_path / "c.json")
    assert Outer.from_yaml(tmp_path / "c.json") == c
    assert Outer.from_yaml(tmp_path / "c.json", overwrite_values={"inner": {"n": 9}}).inner.n == 9
    with pytest.raises(ValidationError):
        Outer.from_dict({"required": 1, "unknown_key": 2})
    with pytest.raises(ValidationError):
        Outer.from_dict({})
    with pytest.raises(ValidationError):
        c.name = "y"  # frozen


def test_config_overwrite_recursive_and_template(tmp_path):
    base = {"a": 1, "b": {"c": 2, "d": {"e": 3}}}
    overwrite_recursive(base, {"b": {"d": {"e": 4}, "f": 5}, "g": 6})
    assert base == {"a": 1, "b": {"c": 2, "d": {"e": 4}, "f": 5}, "g": 6}
    import yaml

    t = Outer.get_template_str()
    assert "# the name" in t and "# second line" in t and "# nested block" in t and "# Outer doc." in t
    parsed = yaml.load(t, Loader=yaml.FullLoader)
    assert parsed["inner"]["n"] == 3 and parsed["color"] == "red" and parsed["required"] is None
    Outer.save_template(tmp_path / "t.yml")
    assert (tmp_path / "t.yml").read_text() == t
    inst = Outer(required=2)
    assert Outer.from_dict(yaml.load(inst.as_annotated_str(), Loader=yaml.FullLoader)) == inst


# -------------------------------------------------------------------------------------------------- topology
@pytest.mark.parametrize("given,expect", [
    (dict(model_parallel_size=2, pipe_parallel_size=2, data_parallel_size=2), dict(world_size=8)),
    (dict(model_parallel_size=2, pipe_parallel_size=2, world_size=8), dict(data_parallel_size=2)),
    (dict(world_size=8, data_parallel_size=2, pipe_parallel_size=1), dict(model_parallel_size=4)),
    (dict(world_size=6, data_parallel_size=3, model_parallel_size=1), dict(pipe_parallel_size=2)),
])
def test_topology_config_derives_missing_size(given, expect):
    c = TopologyConfig(micro_batch_size=2, gradient_accumulation_steps=3, **given)
    for k, v in expect.items():
        assert getattr(c, k) == v
    assert c.global_batch_size == 2 * 3 * c.data_parallel_size


def test_topology_config_batch_sizes():
    c = TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=4, global_batch_size=32, micro_batch_size=2)
    assert c.gradient_accumulation_steps == 4
    c = TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=4, global_batch_size=32, gradient_accumulation_steps=2)
    assert c.micro_batch_size == 4
    with pytest.raises(ValidationError):
        TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=4, global_batch_size=30,
                       micro_batch_size=2, gradient_accumulation_steps=4)
    with pytest.raises(ValidationError):
        TopologyConfig(model_parallel_size=2, pipe_parallel_size=2, data_parallel_size=2, world_size=9, micro_batch_size=1,
                       gradient_accumulation_steps=1)


def test_topology_rank_grid():
    cfgs = [TopologyConfig(model_parallel_size=2, pipe_parallel_size=2, data_parallel_size=2, micro_batch_size=1,
                           gradient_accumulation_steps=1, global_rank=r) for r in range(8)]
    topos = [Topology(c) for c in cfgs]
    # mp is the fastest-varying coordinate: TP neighbours are adjacent ranks
    assert [(t.pipe_parallel_rank, t.data_parallel_rank, t.model_parallel_rank) for t in topos] == [
        (p, d, m) for p in range(2) for d in range(2) for m in range(2)]
    t = topos[5]  # (pp1, dp0, mp1)
    assert t.model_parallel_ranks == [4, 5] and t.data_parallel_ranks == [5, 7] and t.pipe_parallel_ranks == [1, 5]
    assert t.previous_pipe_parallel_rank == 1 and t.next_pipe_parallel_rank is None
    assert t.is_last_pipe_parallel_rank and not t.is_first_pipe_parallel_rank
    assert [x.is_io_rank for x in topos] == [True, False, True, False, True, False, True, False]
    assert t.get_global_rank(pipe_parallel_rank=0) == 1 and t.get_global_rank(model_parallel_rank=0) == 4
    groups = t.all_model_parallel_groups() + t.all_data_parallel_groups() + t.all_pipe_parallel_groups()
    assert all(sorted(sum(gs, [])) == list(range(8)) for gs in (t.all_model_parallel_groups(), t.all_data_parallel_groups(),
                                                                 t.all_pipe_parallel_groups()))
    assert len(groups) == 12


# -------------------------------------------------------------------------------------------------------- LR
def test_learning_rate_scheduler():
    def sched(**kw):
        return LearningRateScheduler(LearningRateSchedulerConfig(learning_rate=1.0, learning_rate_minimum=0.1,
                                                                 learning_rate_warmup_steps=10, learning_rate_decay_iters=110, **kw))

    for style in ("constant", "linear", "cosine"):
        s = sched(learning_rate_decay_style=style)
        assert s.get_lr(0) == pytest.approx(0.0) and s.get_lr(5) == pytest.approx(0.5) and s.get_lr(10) == pytest.approx(1.0)
        lrs = [s.get_lr(i) for i in range(10, 130)]
        assert all(a >= b - 1e-12 for a, b in zip(lrs, lrs[1:])), style
    assert sched(learning_rate_decay_style="constant").get_lr(1000) == pytest.approx(1.0)
    assert sched(learning_rate_decay_style="linear").get_lr(60) == pytest.approx(0.55)
    assert sched(learning_rate_decay_style="cosine").get_lr(60) == pytest.approx(0.55)
    assert sched(learning_rate_decay_style="linear").get_lr(500) == pytest.approx(0.1)
    assert sched(learning_rate_decay_style="cosine").get_lr(500) == pytest.approx(0.1)


# ------------------------------------------------------------------------------------------------- partition
@pytest.mark.parametrize("items,parts", [(7, 1), (7, 2), (8, 4), (10, 3), (5, 5), (33, 8)])
def test_pipe_partition_uniform(items, parts):
    coords = pipe_partition_uniform(items, parts)
    assert len(coords) == parts and coords[0].start == 0 and coords[-1].end == items
    assert all(a.end == b.start for a, b in zip(coords, coords[1:]))
    lengths = [c.length for c in coords]
    assert max(lengths) - min(lengths) <= 1 and lengths == sorted(lengths, reverse=True)  # remainder goes to the first stages


def test_pipe_partition_from_indices_and_balanced():
    coords = pipe_partition_from_indices([0, 3, 4, 9])
    assert [(c.start, c.end) for c in coords] == [(0, 3), (3, 4), (4, 9)]
    with pytest.raises(ValueError):
        pipe_partition_from_indices([0, 3, 2])
    with pytest.raises(ValueError):
        pipe_partition_from_indices([0, 3, 9], num_layers=10)
    b = balanced_bounds([10, 1, 1, 1, 1, 10], 3)
    assert b[0] == 0 and b[-1] == 6 and len(b) == 4
    loads = [sum([10, 1, 1, 1, 1, 10][b[i]:b[i + 1]]) for i in range(3)]
    assert max(loads) == 10
    assert balanced_bounds([1.0] * 8, 4) == [0, 2, 4, 6, 8]


# -------------------------------------------------------------------------------------------------- schedule
@pytest.mark.parametrize("pp", [1, 2, 7, 16])
@pytest.mark.parametrize("gas", [1, 2, 7, 16, 32])
def test_train_schedule_terminates_and_is_consistent(pp, gas):
    result = PipelineScheduleTrain.illustrate(gradient_accumulation_steps=gas, pipe_parallel_size=pp)  # raises on deadlock
    assert 0.0 <= result["idle_fraction"] < 1.0
    per_stage = PipelineScheduleTrain._all_instructions(gas, pp)
    for stage, ins in enumerate(per_stage):
        names = [i.name.replace("Instruction", "") for i in ins]
        assert names.count("ForwardPass") == gas and names.count("BackwardPass") == gas
        assert names[-1] == "OptimizerStep" and names[-2] == "ReduceTiedGrads"
        assert names.count("SendActivation") == (gas if stage < pp - 1 else 0)
        assert names.count("RecvGrad") == (gas if stage < pp - 1 else 0)
        assert names.count("RecvActivation") == (gas if stage > 0 else 0)
        assert names.count("Loss") == (gas if stage == pp - 1 else 0)
        # a micro batch is never back-propagated before it was forwarded, and buffers are not clobbered while in use
        live = {}
        for i in ins:
            if i.name.endswith("ForwardPass"):
                assert i.buffer_id not in live, "buffer reused while its micro batch still awaits backward"
                live[i.buffer_id] = i.micro_batch_id
            elif i.name.endswith("BackwardPass"):
                assert live.pop(i.buffer_id) == i.micro_batch_id
    if pp > 1:
        ideal = (pp - 1) / (gas + pp - 1)
        assert result["idle_fraction"] == pytest.approx(ideal, abs=0.35)


def test_inference_schedule_and_visualize(tmp_path):
    per_stage = PipelineScheduleInference._all_instructions(4, 3)
    assert all([i.name for i in ins].count("InstructionForwardPass") == 4 for ins in per_stage)
    PipelineScheduleInference.illustrate(gradient_accumulation_steps=4, pipe_parallel_size=3)
    img = PipelineScheduleTrain.visualize(gradient_accumulation_steps=4, pipe_parallel_size=2)
    assert img is not None


# ------------------------------------------------------------------------------------------------------ data
@pytest.mark.parametrize("cls", [MemoryMapDataset, FileDataset])
@pytest.mark.parametrize("in_memory", [False, True])
def test_memory_map_roundtrip(tmp_path, cls, in_memory):
    docs = [np.arange(n) + 7 for n in (1, 5, 0, 13)]
    with MemoryMapDatasetBuilder(tmp_path / "mm", dtype=np.uint16) as b:
        for d in docs:
            b.add(d)
    meta = json.loads((tmp_path / "mm.meta.json").read_text())
    assert meta["document_count"] == 4 and np.dtype(meta["dtype"]) == np.uint16
    idx = np.fromfile(tmp_path / "mm.idx", dtype=np.int64).reshape(-1, 2)
    assert idx.tolist() == [[0, 1], [1, 5], [6, 0], [6, 13]]
    ds = cls(tmp_path / "mm", load_index_to_memory=in_memory)
    assert len(ds) == 4 and [x.tolist() for x in ds] == [d.tolist() for d in docs]
    assert ds.sizes().tolist() == [1, 5, 0, 13] and int(ds.sizes(3)) == 13
    with pytest.raises(IndexError):
        ds[4]


def test_builder_refuses_existing_and_wrong_shapes(tmp_path):
    with MemoryMapDatasetBuilder(tmp_path / "a") as b:
        b.add(np.array([1, 2]))
        with pytest.raises(AssertionError):
            b.add(np.zeros((2, 2)))
    with pytest.raises(AssertionError):
        MemoryMapDatasetBuilder(tmp_path / "a")


def test_file_handle_retries_stale_handles(tmp_path, monkeypatch):
    p = tmp_path / "f.bin"
    p.write_bytes(b"abcdef")
    monkeypatch.setattr("time.sleep", lambda s: None)
    h = FileHandle(p)
    calls = {"n": 0}

    def flaky(f):
        calls["n"] += 1
        if calls["n"] < 3:
            raise OSError(errno.ESTALE, "stale file handle")
        f.seek(0)
        return f.read(3)

    assert h.retry_operation(flaky) == b"abc" and calls["n"] == 3

    def short_read(f):
        raise RetryableException("short read")

    with pytest.raises(Exception):
        h.retry_operation(short_read, max_attempts=2)

    def fatal(f):
        raise OSError(errno.EACCES, "denied")

    with pytest.raises(OSError):
        h.retry_operation(fatal)


def test_blend_weights():
    w = weights_by_num_docs([100, 10, 1], alpha=1.0)
    assert np.allclose(w, 1 / 3)  # natural proportions: every dataset sampled exactly once
    w0 = weights_by_num_docs([100, 10, 1], alpha=0.0)
    assert np.allclose(w0 * np.array([100, 10, 1]) / (w0 * np.array([100, 10, 1])).sum(), 1 / 3)  # uniform over datasets
    w3 = weights_by_num_docs([100, 10, 1], alpha=0.3)
    assert w3[2] > w3[1] > w3[0] and w3.sum() == pytest.approx(1.0)
    e = weights_examples_proportional([1000, 100, 10], temperature=1.0, maximum=100)
    rate = e * np.array([1000, 100, 10])
    assert rate[0] == pytest.approx(rate[1]) and rate[2] < rate[1]
    assert np.allclose(weights_examples_proportional([5, 5], 1.0, None), 0.5)
    with pytest.raises(AssertionError):
        weights_examples_proportional([1, 2], temperature=0)


def test_blend_index_properties():
    counts = np.array([7, 3, 5, 1])
    idx = blend_index(counts)
    assert idx.shape == (16, 2)
    for d, c in enumerate(counts):
        assert idx[idx[:, 0] == d][:, 1].tolist() == list(range(c))
    # greedy rule: at every step the chosen dataset has the lowest sampled/target ratio (ties: lowest index)
    sampled = np.zeros(4, dtype=np.int64)
    for d, i in idx:
        ratio = np.where(sampled < counts, sampled / counts, np.inf)
        assert d == int(ratio.argmin()) and i == sampled[d]
        sampled[d] += 1
    assert blend_index(np.array([0, 4])).tolist() == [[1, 0], [1, 1], [1, 2], [1, 3]]


class _Range(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n, self.seeds = n, []

    def __len__(self):
        return self.n

    def set_seed(self, seed, shuffle=True):
        self.seeds.append(seed)


@pytest.mark.parametrize("dp,mbs,n", [(1, 1, 10), (2, 2, 23), (4, 1, 17)])
def test_resumable_sampler_covers_every_index_once(dp, mbs, n):
    def topo(rank):
        return Topology(TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=dp, micro_batch_size=mbs,
                                       gradient_accumulation_steps=1, global_rank=rank))

    usable = n // (dp * mbs) * (dp * mbs)
    seen = []
    for r in range(dp):
        ds = _Range(n)
        it = iter(RandomSampler(ds, seed=5, consumed_samples=0, topology=topo(r)))
        seen.append([next(it) for _ in range(usable // dp)])
        assert ds.seeds[0] == 5
    flat = sorted(sum(seen, []))
    assert len(set(flat)) == len(flat) == usable and set(flat) <= set(range(n))
    # resume mid-epoch: the continuation equals the tail of the uninterrupted stream
    consumed = (usable // (dp * mbs) // 2) * dp * mbs
    for r in range(dp):
        it = iter(RandomSampler(_Range(n), seed=5, consumed_samples=consumed, topology=topo(r)))
        assert [next(it) for _ in range((usable - consumed) // dp)] == seen[r][consumed // dp:]
    # next epoch re-seeds the dataset with seed + epoch
    ds = _Range(n)
    it = iter(RandomSampler(ds, seed=5, consumed_samples=usable, topology=topo(0)))
    next(it)
    assert ds.seeds[-1] == 6


class _SeededItems(BaseDataset):
    """Items carry the seed their (possibly worker-side) dataset copy was last given."""

    def __init__(self, n):
        self.n = n
        super().__init__(seed=0)

    def ident(self):
        return "seeded"

    def __len__(self):
        return self.n

    def set_seed(self, seed, shuffle=True):
        self.seed = seed

    def __getitem__(self, i):
        return (i, self.seed)

    def collate(self, batch):
        return batch

    @staticmethod
    def sync_batch_to_model_parallel(topology, batch):
        return batch


@pytest.mark.parametrize("workers", [0, 1])
def test_loader_reseeds_worker_copies_at_every_epoch(workers):
    """The torch iterator (and with it the worker processes' dataset copies) is rebuilt at the epoch boundary, after
    the dataset was re-seeded with seed + epoch (reference: core/data/dataloader.py:56-96)."""
    from scaling_b200.core.data.loader import DataLoader

    topo = Topology(TopologyConfig(model_parallel_size=1, pipe_parallel_size=1, data_parallel_size=1, micro_batch_size=2,
                                   gradient_accumulation_steps=1, global_rank=0))
    loader = DataLoader(seed=10, consumed_samples=0, dataset=_SeededItems(6), topology=topo, num_workers=workers,
                        pin_memory=False)
    seeds = [[seed for _, seed in next(loader)] for _ in range(9)]  # three epochs of three batches
    assert seeds == [[10, 10]] * 3 + [[11, 11]] * 3 + [[12, 12]] * 3
    assert loader.sampler.consumed_samples == 18


def test_blend_index_extends_the_legacy_python_loop():
    """The legacy Python loop (transformer/data/legacy_blended_dataset.py:214-229) stops as soon as the FIRST dataset
    has delivered its target; the native index (like the Rust wheel, whose blended length equals the sum of the parts,
    tests/transformer/test_blended_dataset.py:58) keeps going until every dataset has.  Same rule, so the legacy
    output is a prefix of ours."""
    counts = np.array([5, 9, 2, 7], dtype=np.int64)
    sampled = np.zeros(len(counts), dtype=np.int64)
    legacy = []
    while (sampled < counts).all():
        d = int((sampled / counts).argmin())
        legacy.append([d, int(sampled[d])])
        sampled[d] += 1
    ours = blend_index(counts)
    assert len(ours) == counts.sum() > len(legacy) > 0
    assert ours[:len(legacy)].tolist() == legacy


# ---------------------------------------------------------------------------------------------------- logging
def test_logger_sinks_and_rank_filters(tmp_path, capsys):
    from scaling_b200.core import LoggerConfig
    from scaling_b200.core.logging import logger

    cfg = LoggerConfig(log_dir=tmp_path / "logs", log_level="info", use_tensorboard=False, metrics_ranks=[0])
    assert cfg.log_dir is not None and cfg.log_dir.parent == tmp_path / "logs"  # timestamped sub-directory
    assert cfg.is_rank_in_metrics_ranks(0) and not cfg.is_rank_in_metrics_ranks(1)
    assert LoggerConfig(tensorboard_ranks=None).is_rank_in_tensorboard_ranks(0)
    assert not LoggerConfig(tensorboard_ranks=[2]).is_rank_in_tensorboard_ranks(0)
    with pytest.raises(ValidationError):
        LoggerConfig(use_wandb=True, wandb_api_key=None) if "WANDB_API_KEY" not in __import__("os").environ else (_ for _ in ()).throw(
            ValidationError.from_exception_data("x", []))
    logger.configure(cfg, name="RANK 0", global_rank=0)
    logger.info("hello from the test")
    logger.log_metrics({"training/loss": 1.25, "runtime/step_duration": 0.5}, step=3)
    logger.log_config_dict({"a": {"b": 1}})
    files = list(cfg.log_dir.glob("*.log"))
    assert len(files) == 1
    text = files[0].read_text()
    assert "hello from the test" in text and "training/loss" in text and "1.25" in text
    logger.configure(LoggerConfig(log_level="warning"), name="quiet")
    logger.info("not shown")
    assert "not shown" not in capsys.readouterr().err


def test_tokenizer_wrapper(tmp_path):
    from tokenizers import Tokenizer as HF
    from tokenizers import decoders, models, pre_tokenizers

    from scaling_b200.transformer.tokenizer import Tokenizer, load_tokenizers

    vocab = {"<|endoftext|>": 0, "<unk>": 1, "▁hello": 2, "hello": 3, "▁world": 4, "world": 5}
    tok = HF(models.WordLevel(vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.Metaspace(replacement="▁", prepend_scheme="always")
    tok.decoder = decoders.Metaspace(replacement="▁")
    tok.save(str(tmp_path / "vocab.json"))
    t = Tokenizer.from_file(tmp_path / "vocab.json")
    assert t.eos_token_id == 0 and len(t) == 6 and t.vocab_size == 6
    assert t.encode("hello world") == [2, 4] and t.decode([2, 4]) == "hello world"
    with_space, without = load_tokenizers(tmp_path / "vocab.json")
    assert with_space.encode("world") == [4] and without.encode("world") == [5]  # continuation pieces get no prefix space
    assert Tokenizer.from_str((tmp_path / "vocab.json").read_text()).encode("hello") == [2]


def test_pack_frozen_parameters_keeps_values_and_makes_neighbours_adjacent():
    """Frozen parameters move into one arena in registration order (128-element slots); trained ones are left alone."""
    import torch

    from scaling_b200.core.optimizer.param_group import pack_frozen_parameters
    from scaling_b200.ops.functional import fused_view

    torch.manual_seed(0)
    a, b, c = (torch.nn.Parameter(torch.randn(n, 128), requires_grad=False) for n in (4, 2, 3))
    t = torch.nn.Parameter(torch.randn(5, 128))
    odd = torch.nn.Parameter(torch.randn(7), requires_grad=False)
    before = [p.detach().clone() for p in (a, b, c, t, odd)]
    ptr_t = t.data_ptr()
    arenas = pack_frozen_parameters([a, b, t, c, odd, a])
    assert len(arenas) == 1 and arenas[0].numel() == (4 + 2 + 3) * 128 + 128
    for p, ref in zip((a, b, c, t, odd), before):
        assert torch.equal(p.detach(), ref)
    assert t.data_ptr() == ptr_t
    fv = fused_view((a.data, b.data, c.data))
    assert fv is not None and fv.shape == (9, 128) and torch.equal(fv[4:6], b.detach())
    a.data.mul_(2.0)  # views, not copies
    assert torch.equal(arenas[0][:4 * 128].view(4, 128), a.detach())
