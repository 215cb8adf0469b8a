"""Per-instruction profiler (reference: core/profiler/profiler.py:24-139, timer.py:16-23, profiler_config.py).

Timers record CUDA events on the current stream instead of calling ``cuda.synchronize`` on both ends, so profiling
does not serialise the GPU; elapsed times are resolved once, when the window closes.  On CPU they
fall back to ``time.perf_counter``.  Output format (``profile.json``) is the reference's, consumed by the schedule
simulator.
"""
from __future__ import annotations

import json
import time
from contextlib import contextmanager
from pathlib import Path
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
import torch.distributed as dist
from pydantic import Field

from .config import BaseConfig
from .topology import Topology


class ProfilerConfig(BaseConfig):
    profile_steps: int = Field(0, description="number of steps to time; 0 disables profiling")
    profile_start_at_step: int = Field(10, description="first profiled step (lets clocks and caches settle)")
    profiler_output: Optional[Path] = Field(None, description="where rank 0 writes profile.json")


class SynchronizedTimer:
    """Wall-clock timer bracketed by device synchronisation — used once per step for ``step_duration``."""

    def __init__(self) -> None:
        self._start: Optional[float] = None

    @staticmethod
    def _sync() -> None:
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def start(self) -> None:
        self._sync()
        self._start = time.time()

    def stop(self) -> float:
        assert self._start is not None, "timer was not started"
        self._sync()
        d = time.time() - self._start
        self._start = None
        return d


class Profiler:
    def __init__(self, config: ProfilerConfig, topology: Topology):
        self.config = config
        self.topology = topology
        self.current_step = -1
        self._pending: List[Tuple[str, int, int, int, Any, Any]] = []
        self.observations: List[Dict[str, Any]] = []
        self._flushed = False

    @property
    def active(self) -> bool:
        c = self.config
        return c.profile_steps > 0 and c.profile_start_at_step <= self.current_step < c.profile_start_at_step + c.profile_steps

    def step(self) -> None:
        self.current_step += 1

    @contextmanager
    def time(self, timer_name: str, micro_batch_id: int = -1, buffer_id: int = -1) -> Iterator[None]:
        if not self.active:
            yield
            return
        if torch.cuda.is_available():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            yield
            e1.record()
            self._pending.append((timer_name, self.current_step, micro_batch_id, buffer_id, e0, e1))
        else:
            t0 = time.perf_counter()
            yield
            self._pending.append((timer_name, self.current_step, micro_batch_id, buffer_id, t0, time.perf_counter()))

    def _resolve(self) -> None:
        if torch.cuda.is_available() and self._pending:
            torch.cuda.synchronize()
        topo = self.topology
        for name, step, mb, buf, a, b in self._pending:
            dur = a.elapsed_time(b) / 1e3 if hasattr(a, "elapsed_time") else b - a
            self.observations.append({
                "timer_name": name, "step": step, "micro_batch_id": mb, "buffer_id": buf,
                "pipe_parallel_rank": topo.pipe_parallel_rank, "data_parallel_rank": topo.data_parallel_rank,
                "model_parallel_rank": topo.model_parallel_rank, "duration": dur})
        self._pending.clear()

    def flush(self) -> None:
        """Call at the end of every step; when the window has just closed, gather to rank 0 and write the file."""
        c = self.config
        if c.profile_steps == 0 or self._flushed or self.current_step < c.profile_start_at_step + c.profile_steps - 1:
            return
        self._resolve()
        gathered: Optional[List[Any]] = None
        if self.topology.is_distributed_initialized and self.topology.config.world_size > 1:
            gathered = [None] * self.topology.config.world_size if self.topology.config.global_rank == 0 else None
            dist.gather_object(self.observations, gathered, dst=0)
        else:
            gathered = [self.observations]
        if self.topology.config.global_rank in (0, None) and c.profiler_output is not None and gathered is not None:
            tc = self.topology.config
            out = {"pipe_parallel_size": tc.pipe_parallel_size, "data_parallel_size": tc.data_parallel_size,
                   "model_parallel_size": tc.model_parallel_size,
                   "gradient_accumulation_steps": tc.gradient_accumulation_steps,
                   "observations": [o for obs in gathered for o in obs]}
            Path(c.profiler_output).parent.mkdir(parents=True, exist_ok=True)
            Path(c.profiler_output).write_text(json.dumps(out, indent=1))
        self._flushed = True
