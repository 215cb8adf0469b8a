"""Benchmark driver for this framework's arm of ``bench.py`` (kept inside the package so it uses only public API).

Two timed phases, both through ``TransformerTrainer.train_step`` (the call a user makes):
  1. *device* — micro batches are staged on the GPU before the clock starts: measures the training step itself;
  2. *end to end* — the stock ``DataLoader`` with pinned host batches: every step copies its inputs host→device and
     reads the loss back (device→host).  Bytes are counted from the tensors that are actually copied.
"""
from __future__ import annotations

import itertools
import json
import os
from pathlib import Path
from typing import Any, Iterator

import torch
import torch.distributed as dist


class _StagedLoader:
    """Cycles over batches that already live on the device (phase 1)."""

    def __init__(self, loader: Any, n: int, device: torch.device):
        self.batches = []
        for _ in range(n):
            b = next(loader)
            b.to_(device)
            self.batches.append(b)
        self._it = itertools.cycle(self.batches)

    def __iter__(self) -> Iterator[Any]:
        return self

    def __next__(self) -> Any:
        return next(self._it)


class _CountingLoader:
    """Stock loader + byte counter for the host→device traffic of phase 2."""

    def __init__(self, loader: Any):
        self.loader = loader
        self.bytes = 0

    def __iter__(self) -> Iterator[Any]:
        return self

    def __next__(self) -> Any:
        b = next(self.loader)
        for v in vars(b).values():
            if torch.is_tensor(v):
                assert not v.is_cuda
                self.bytes += v.numel() * v.element_size()
        return b


def run_bench(args: Any, bench: Any) -> None:
    rank, local, world, addr, port = bench.dist_env()
    n = args.gpus
    assert world == n, f"WORLD_SIZE {world} != --gpus {n}"
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    from ..core import LaunchConfig
    from ..ops import _lib
    from ..ops import functional as OF
    from .train import build_trainer

    OF.set_ops(args.ops)
    gb = args.seqs_per_gpu * n
    prefix = Path("/tmp/scaling_bench_ours") / "data"
    bench.write_synthetic_dataset(prefix, n_docs=gb * (2 * (args.steps + args.warmup) + 4), vocab=32000, rank=rank)
    cfg = bench.model_config(args, n, prefix)
    launch = LaunchConfig(master_port=port, master_addr=addr, world_size=world, global_rank=rank, local_slot=local, payload=cfg)
    trainer = build_trainer(launch)
    topo = trainer.context.topology
    gas = topo.config.gradient_accumulation_steps

    # ---- phase 1: device-timed, inputs staged on the GPU -----------------------------------------------------
    real_loader = trainer.dataloader
    if real_loader is not None:
        trainer.dataloader = _StagedLoader(real_loader, gas * (args.warmup + args.steps), topo.device)  # no batch repeats
    _lib.reset_launch_count()
    launches_before = [0]

    def step():
        return trainer.train_step()

    # count launches of the timed steps only
    for _ in range(args.warmup):
        out = step()
    launches_before[0] = _lib.launch_count()
    ms, clocks, out = bench.timed_steps(step, 0, args.steps, world, local)
    launches = _lib.launch_count() - launches_before[0]

    # ---- phase 2: end to end through the stock pinned-memory loader ---------------------------------------------
    counter = None
    if real_loader is not None:
        counter = _CountingLoader(real_loader)
        trainer.dataloader = counter
    for _ in range(min(args.warmup, 2)):
        step()
    if counter is not None:
        counter.bytes = 0
    ms_e2e, _clocks2, out2 = bench.timed_steps(step, 0, args.steps, world, local)
    h2d = torch.tensor([counter.bytes if counter is not None else 0], dtype=torch.float64, device=topo.device)
    lt = torch.tensor([launches], dtype=torch.float64, device=topo.device)
    if world > 1:
        dist.all_reduce(h2d)
        dist.all_reduce(lt)
    if rank == 0:
        tokens = gb * bench.SEQ * args.steps
        extra = {
            "e2e": {"value": tokens / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
                    "h2d_bytes_per_step": int(h2d.item() / args.steps), "d2h_bytes_per_step": 16},
            "gpu_launches": int(lt.item()),
            "ops": args.ops,
            "last_loss": out.loss,
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9,
            "native_library": str(_lib.lib_path()),
        }
        print(json.dumps(bench.result_line(args, n, "scaling_b200", ms, clocks, extra)))
