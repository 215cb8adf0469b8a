"""Run state: counters, seeds, per-rank context checkpoints (reference: core/context/context.py:31-197)."""
from __future__ import annotations

import random
from pathlib import Path
from typing import Any, Dict, Generic, Optional, TypeVar, Union

import numpy as np
import torch
import torch.distributed as dist

from .config import BaseConfig
from .topology import Topology

ConfigT = TypeVar("ConfigT", bound=BaseConfig)


class BaseContext(Generic[ConfigT]):
    def __init__(self, config: ConfigT, topology: Topology) -> None:
        self.config = config
        self.topology = topology
        self.iterations = 0
        self.consumed_samples = 0
        self.consumed_samples_evaluation = 0
        self._use_determined = False
        self.determined_context: Any = None
        self.determined_profiler: Any = None

    def initialize(self, master_addr: str, master_port: str, torch_distributed_timeout_minutes: int = 20,
                   seed: int = 42, distributed: bool = True) -> None:
        """Bring up the device (and torch.distributed unless ``distributed=False``) and seed every host RNG with
        ``seed + global_rank``."""
        if self.topology is not None:
            if distributed:
                self.topology.initialize_distributed(
                    master_addr=master_addr, master_port=master_port,
                    torch_distributed_timeout_minutes=torch_distributed_timeout_minutes, seed=seed,
                )
            else:
                self.topology.initialize_device()
            seed = seed + (self.topology.config.global_rank or 0)
        random.seed(seed)
        np.random.seed(seed % (2**32))
        torch.manual_seed(seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed(seed)

    def step(self) -> None:
        self.iterations += 1
        self.consumed_samples += self.topology.config.global_batch_size

    def state_dict(self) -> Dict[str, Any]:
        sd = {
            "iterations": self.iterations,
            "consumed_samples": self.consumed_samples,
            "consumed_samples_evaluation": self.consumed_samples_evaluation,
            "random_rng_state": random.getstate(),
            "np_rng_state": np.random.get_state(),
            "torch_rng_state": torch.get_rng_state(),
            "topology": self.topology.state_dict(),
        }
        if torch.cuda.is_available():
            sd["torch_cuda_rng_state"] = torch.cuda.get_rng_state()
        return sd

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        self.iterations = state_dict["iterations"]
        self.consumed_samples = state_dict["consumed_samples"]
        self.consumed_samples_evaluation = state_dict.get("consumed_samples_evaluation", 0)
        random.setstate(state_dict["random_rng_state"])
        np.random.set_state(state_dict["np_rng_state"])
        torch.set_rng_state(state_dict["torch_rng_state"])
        if torch.cuda.is_available() and state_dict.get("torch_cuda_rng_state") is not None:
            torch.cuda.set_rng_state(state_dict["torch_cuda_rng_state"])
        self.topology.load_state_dict(state_dict.get("topology"))

    def save_checkpoint(self, dir: Union[Path, str]) -> None:
        dir = Path(dir)
        rank = self.topology.config.global_rank
        if rank == 0:
            self.config.save(dir / "config.yml")
        torch.save(self.state_dict(), str(dir / f"context_global_rank_{rank}.pt"))

    def load_checkpoint(self, dir: Union[Path, str]) -> None:
        """Own file if present; ranks that did not exist when the checkpoint was written (re-layout) catch up through
        a MAX all-reduce of the counters."""
        dir = Path(dir)
        f = dir / f"context_global_rank_{self.topology.config.global_rank}.pt"
        if f.is_file():
            self.load_state_dict(torch.load(str(f), weights_only=False))
        if self.topology.is_distributed_initialized:
            t = torch.tensor([self.iterations, self.consumed_samples, self.consumed_samples_evaluation],
                             dtype=torch.long, device=self.topology.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            self.iterations, self.consumed_samples, self.consumed_samples_evaluation = (int(x) for x in t.tolist())


class DeterminedBaseContext(BaseContext[ConfigT]):
    """Context that can additionally carry a Determined core context (optional dependency; reference :167-197)."""

    def initialize_with_determined(self, master_addr: str, master_port: str, determined_context: Any,
                                   determined_profiler: Optional[Any], torch_distributed_timeout_minutes: int = 20,
                                   seed: int = 42, distributed: bool = True) -> None:
        self._use_determined = True
        self.determined_context = determined_context
        self.determined_profiler = determined_profiler
        self.initialize(master_addr, master_port, torch_distributed_timeout_minutes, seed, distributed)


BaseContextGeneric = TypeVar("BaseContextGeneric", bound=BaseContext)
DeterminedBaseContextGeneric = TypeVar("DeterminedBaseContextGeneric", bound=DeterminedBaseContext)
