"""3-D process grid, process groups and the tensor-parallel-constant RNG.

Reference behaviour: core/topology/topology_config.py:22-206 (any 3 of {mp, pp, dp, world} and any 2 of
{global, micro, grad-acc} are enough), core/topology/topology.py:20-441 (rank grid ``arange(world).reshape(pp, dp, mp)``
with mp fastest so tensor-parallel neighbours are adjacent GPUs, every group created on every rank in the same order),
core/topology/rng_tracker.py:59-95 (a forkable RNG stream seeded identically across a TP group).

Differences by design: the backend/device are not hard-coded — ``nccl``+``cuda`` when a GPU is present, ``gloo``+``cpu``
otherwise (BASELINE config 1 and all CPU tests run the full 3-D engine on gloo); and a cheap counter-based
``dropout_seed`` is exposed for the Philox kernels, which need no RNG state forking at all.
"""
from __future__ import annotations

import contextlib
import datetime
import os
from enum import Enum
from typing import Any, Dict, Iterator, List, Optional

import numpy as np
import torch
import torch.distributed as dist
from pydantic import Field, model_validator

from .config import BaseConfig


class PipePartitionMethod(Enum):
    UNIFORM = "uniform"
    BALANCED = "balanced"


class ActivationCheckpointingType(Enum):
    EVERY_PIPE_STAGE = "every_pipe_stage"
    EVERY_LAYER = "every_layer"
    DISABLED = "disabled"


class TopologyConfig(BaseConfig):
    global_rank: Optional[int] = Field(None, description="rank of this process in the job", ge=0)
    world_size: int = Field(description="total number of processes (= GPUs)", gt=0)
    local_slot: Optional[int] = Field(None, description="device index on this node", ge=0)
    model_parallel_size: int = Field(description="tensor-parallel degree", gt=0)
    pipe_parallel_size: int = Field(description="pipeline-parallel degree", gt=0)
    data_parallel_size: int = Field(description="data-parallel degree", gt=0)
    global_batch_size: int = Field(description="samples per optimizer step over all data-parallel ranks", gt=0)
    micro_batch_size: int = Field(description="samples per forward/backward pass on one data-parallel rank", gt=0)
    gradient_accumulation_steps: int = Field(description="micro batches per optimizer step", gt=0)
    pipe_partition_method: PipePartitionMethod = Field(PipePartitionMethod.UNIFORM, description="layer → stage assignment")
    pipe_partition_overwrite: Optional[List[int]] = Field(None, description="explicit stage boundaries (layer indices)")
    activation_checkpointing_type: ActivationCheckpointingType = Field(
        ActivationCheckpointingType.DISABLED, description="recompute activations per layer / per stage / never"
    )
    sequence_parallel: bool = Field(False, description="Megatron sequence parallelism over the tensor-parallel group")

    @model_validator(mode="before")
    @classmethod
    def _derive_sizes(cls, values: Dict[Any, Any]) -> Dict[Any, Any]:
        v = dict(values)
        mp, pp, dp, world = (v.get(k) for k in ("model_parallel_size", "pipe_parallel_size", "data_parallel_size", "world_size"))
        if sum(x is not None for x in (mp, pp, dp, world)) < 3:
            raise AssertionError(
                "At least 3 out of 4 parallelization parameters (model_parallel_size, pipe_parallel_size, "
                "data_parallel_size and world_size) need to be set."
            )
        if world is None:
            world = mp * pp * dp
        if mp is None:
            mp = world // (pp * dp)
        if pp is None:
            pp = world // (mp * dp)
        if dp is None:
            dp = world // (mp * pp)
        assert mp * pp * dp == world, f"mp {mp} * pp {pp} * dp {dp} != world_size {world}"
        gb, mb, ga = (v.get(k) for k in ("global_batch_size", "micro_batch_size", "gradient_accumulation_steps"))
        if sum(x is not None for x in (gb, mb, ga)) < 2:
            raise AssertionError(
                "At least 2 out of 3 batch size parameters (global_batch_size, micro_batch_size, "
                "and gradient_accumulation_steps) need to be set."
            )
        if ga is None:
            ga = gb // (mb * dp)
        if mb is None:
            mb = gb // (ga * dp)
        if gb is None:
            gb = mb * ga * dp
        assert gb == mb * ga * dp, (
            f"global_batch_size {gb} does not equal the product of micro_batch_size ({mb}) and "
            f"gradient_accumulation_steps ({ga}) and data_parallel_size ({dp})."
        )
        v.update(model_parallel_size=mp, pipe_parallel_size=pp, data_parallel_size=dp, world_size=world,
                 global_batch_size=gb, micro_batch_size=mb, gradient_accumulation_steps=ga)
        return v


# --------------------------------------------------------------------------------------------------------------
# tensor-parallel-constant RNG
# --------------------------------------------------------------------------------------------------------------
class CudaRNGStateTracker:
    """A second RNG stream, seeded identically on every rank of a TP group.

    ``fork()`` swaps it in for the duration of a block so ATen dropout on TP-replicated activations draws the same
    mask on every TP rank (reference rng_tracker.py:59-95).  Works on CPU generators as well, which the reference does
    not.  The sm_100a dropout kernels do not use it — they take ``Topology.dropout_seed`` + a call counter.
    """

    def __init__(self, seed: int, device: torch.device):
        self.seed = seed
        self.device = device
        self._cuda = device.type == "cuda"
        orig = self._get()
        if self._cuda:
            torch.cuda.manual_seed(seed)
        else:
            torch.manual_seed(seed)
        self.state = self._get()
        self._set(orig)

    def _get(self) -> torch.Tensor:
        return torch.cuda.get_rng_state(self.device) if self._cuda else torch.get_rng_state()

    def _set(self, state: torch.Tensor) -> None:
        if self._cuda:
            torch.cuda.set_rng_state(state, self.device)
        else:
            torch.set_rng_state(state)

    @contextlib.contextmanager
    def fork(self) -> Iterator[None]:
        orig = self._get()
        self._set(self.state)
        try:
            yield
        finally:
            self.state = self._get()
            self._set(orig)

    def state_dict(self) -> Dict[str, Any]:
        return {"seed": self.seed, "state": self.state}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.seed = sd["seed"]
        self.state = sd["state"]


# --------------------------------------------------------------------------------------------------------------
# topology
# --------------------------------------------------------------------------------------------------------------
_exit_hook_installed = False


def _destroy_gloo_group_at_exit() -> None:
    """A script that ends without ``destroy_process_group()`` leaves gloo's worker threads to the interpreter's
    teardown, which sporadically aborts ("terminate called without an active exception") and turns a finished rank
    into a failed one.  gloo's destroy is local (no peer has to answer), so it is safe to run it from an exit hook."""
    global _exit_hook_installed
    if _exit_hook_installed:
        return
    _exit_hook_installed = True
    import atexit

    def _shutdown() -> None:
        try:
            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001  (shutting down anyway)
            pass

    atexit.register(_shutdown)


class Topology:
    def __init__(self, config: TopologyConfig):
        self.config = config
        c = config
        self._grid = np.arange(c.world_size).reshape(c.pipe_parallel_size, c.data_parallel_size, c.model_parallel_size)
        self._pp_rank = self._dp_rank = self._mp_rank = None
        if c.global_rank is not None:
            pp, dp, mp = np.argwhere(self._grid == c.global_rank)[0]
            self._pp_rank, self._dp_rank, self._mp_rank = int(pp), int(dp), int(mp)
        self._device: Optional[torch.device] = None
        self._pp_group = self._dp_group = self._mp_group = None
        self._rng: Optional[CudaRNGStateTracker] = None
        self.is_distributed_initialized = False
        self.dropout_seed = 0
        self.dropout_offset = 0  # Philox counter of the next dropout call; advances identically on all TP ranks
        self.backend: Optional[str] = None

    # ---- device / distributed init ---------------------------------------------------------------
    def initialize_device(self) -> None:
        if torch.cuda.is_available():
            slot = self.config.local_slot
            assert slot is not None, "cannot initialize a CUDA device without local_slot"
            assert slot < torch.cuda.device_count(), f"cannot assign gpu {slot} of {torch.cuda.device_count()}"
            torch.cuda.set_device(slot)
            self._device = torch.device("cuda", slot)
        else:
            self._device = torch.device("cpu")

    def initialize_distributed(self, master_addr: str, master_port: str, torch_distributed_timeout_minutes: int = 20,
                               seed: int = 42, backend: Optional[str] = None) -> None:
        assert self.config.global_rank is not None, "cannot initialize distributed without global_rank"
        self.initialize_device()
        assert self._device is not None
        self.backend = backend or ("nccl" if self._device.type == "cuda" else "gloo")
        # one seed per (pp, dp) cell -> constant inside a TP group
        self._seed_rng(seed)
        if not dist.is_initialized():
            kwargs: Dict[str, Any] = {}
            if self.backend == "nccl":
                kwargs["device_id"] = self._device
            dist.init_process_group(
                backend=self.backend,
                init_method=f"tcp://{master_addr}:{master_port}",
                world_size=self.config.world_size,
                rank=self.config.global_rank,
                timeout=datetime.timedelta(minutes=torch_distributed_timeout_minutes),
                **kwargs,
            )
            if self.backend == "gloo":
                _destroy_gloo_group_at_exit()
        # every rank creates every group in the same order (collective-consistent)
        for ranks in self.all_pipe_parallel_groups():
            g = dist.new_group(ranks)
            if self.config.global_rank in ranks:
                self._pp_group = g
        for ranks in self.all_data_parallel_groups():
            g = dist.new_group(ranks)
            if self.config.global_rank in ranks:
                self._dp_group = g
        for ranks in self.all_model_parallel_groups():
            g = dist.new_group(ranks)
            if self.config.global_rank in ranks:
                self._mp_group = g
        # pipeline p2p gets communicators of its own — activations one way, gradients the other — so that a send waiting
        # for its receiver serialises neither with the collectives on the world group nor with traffic in the other
        # direction (the reference sends everything through the world group, communicator.py:381-500)
        self._p2p_groups = (None, None)
        if self.config.pipe_parallel_size > 1 and os.environ.get("SCALING_B200_P2P_GROUPS", "0") == "1":
            everyone = list(range(self.config.world_size))
            self._p2p_groups = (dist.new_group(everyone), dist.new_group(everyone))
        self.is_distributed_initialized = True

    @property
    def pipe_p2p_groups(self) -> tuple:
        """(forward-traffic group, backward-traffic group) spanning all ranks, or (None, None) without pipelining."""
        return getattr(self, "_p2p_groups", (None, None))

    def _seed_rng(self, seed: int) -> None:
        base = seed + int(self._grid[self._pp_rank, self._dp_rank, 0])
        self._rng = CudaRNGStateTracker(seed=base, device=self.device)
        self.dropout_seed = base * 2654435761 % (2**63)

    # ---- RNG -------------------------------------------------------------------------------------
    def has_model_parallel_constant_rng(self) -> bool:
        return self._rng is not None

    def model_parallel_constant_rng(self) -> Any:
        if self._rng is None:
            return contextlib.nullcontext()
        return self._rng.fork()

    def next_dropout_offset(self, numel: int = 0) -> int:
        """Stream id of the next Philox dropout call (the kernels hash (seed, element index, stream id), so every call
        gets an independent mask).  TP ranks of one (pp, dp) cell share the seed and make the same calls in the same
        order; activation recomputation replays the ids."""
        off = self.dropout_offset
        self.dropout_offset += 1
        return off

    def state_dict(self) -> Optional[Dict[str, Any]]:
        if self._rng is None:
            return None
        return {
            "dropout_offset": self.dropout_offset,
            "model_parallel_constant_rng": self._rng.state_dict(),
            "model_parallel_size": self.config.model_parallel_size,
            "pipe_parallel_size": self.config.pipe_parallel_size,
        }

    def load_state_dict(self, state_dict: Optional[Dict[str, Any]]) -> None:
        if state_dict is None or self._rng is None:
            return
        # the stream is only meaningful for the layout it was recorded under
        if (state_dict.get("model_parallel_size") == self.config.model_parallel_size
                and state_dict.get("pipe_parallel_size") == self.config.pipe_parallel_size):
            self._rng.load_state_dict(state_dict["model_parallel_constant_rng"])
            self.dropout_offset = int(state_dict.get("dropout_offset", 0))

    # ---- coordinates -----------------------------------------------------------------------------
    @property
    def device(self) -> torch.device:
        if self._device is None:
            self.initialize_device()
        assert self._device is not None
        return self._device

    @property
    def pipe_parallel_rank(self) -> int:
        return self._pp_rank  # type: ignore[return-value]

    @property
    def data_parallel_rank(self) -> int:
        return self._dp_rank  # type: ignore[return-value]

    @property
    def model_parallel_rank(self) -> int:
        return self._mp_rank  # type: ignore[return-value]

    @property
    def is_first_pipe_parallel_rank(self) -> bool:
        return self._pp_rank == 0

    @property
    def is_last_pipe_parallel_rank(self) -> bool:
        return self._pp_rank == self.config.pipe_parallel_size - 1

    @property
    def is_first_model_parallel_rank(self) -> bool:
        return self._mp_rank == 0

    @property
    def is_io_rank(self) -> bool:
        """First and last pipeline stage load data (inputs / targets); only TP rank 0 reads, the rest receive."""
        return (self.is_first_pipe_parallel_rank or self.is_last_pipe_parallel_rank) and self.is_first_model_parallel_rank

    @property
    def previous_pipe_parallel_rank(self) -> Optional[int]:
        if self.is_first_pipe_parallel_rank:
            return None
        return int(self._grid[self._pp_rank - 1, self._dp_rank, self._mp_rank])

    @property
    def next_pipe_parallel_rank(self) -> Optional[int]:
        if self.is_last_pipe_parallel_rank:
            return None
        return int(self._grid[self._pp_rank + 1, self._dp_rank, self._mp_rank])

    @property
    def pipe_parallel_ranks(self) -> List[int]:
        return [int(r) for r in self._grid[:, self._dp_rank, self._mp_rank]]

    @property
    def data_parallel_ranks(self) -> List[int]:
        return [int(r) for r in self._grid[self._pp_rank, :, self._mp_rank]]

    @property
    def model_parallel_ranks(self) -> List[int]:
        return [int(r) for r in self._grid[self._pp_rank, self._dp_rank, :]]

    @property
    def pipe_parallel_group(self) -> Any:
        return self._pp_group

    @property
    def data_parallel_group(self) -> Any:
        return self._dp_group

    @property
    def model_parallel_group(self) -> Any:
        return self._mp_group

    def all_pipe_parallel_groups(self) -> List[List[int]]:
        c = self.config
        return [[int(r) for r in self._grid[:, d, m]] for d in range(c.data_parallel_size) for m in range(c.model_parallel_size)]

    def all_data_parallel_groups(self) -> List[List[int]]:
        c = self.config
        return [[int(r) for r in self._grid[p, :, m]] for p in range(c.pipe_parallel_size) for m in range(c.model_parallel_size)]

    def all_model_parallel_groups(self) -> List[List[int]]:
        c = self.config
        return [[int(r) for r in self._grid[p, d, :]] for p in range(c.pipe_parallel_size) for d in range(c.data_parallel_size)]

    def get_global_rank(self, pipe_parallel_rank: Optional[int] = None, data_parallel_rank: Optional[int] = None,
                        model_parallel_rank: Optional[int] = None) -> int:
        p = self._pp_rank if pipe_parallel_rank is None else pipe_parallel_rank
        d = self._dp_rank if data_parallel_rank is None else data_parallel_rank
        m = self._mp_rank if model_parallel_rank is None else model_parallel_rank
        return int(self._grid[p, d, m])

    def get_global_rank_group(self, pipe_parallel_rank: Optional[int] = None, data_parallel_rank: Optional[int] = None,
                              model_parallel_rank: Optional[int] = None) -> List[int]:
        """All global ranks matching the given coordinates (None = any)."""
        sl = tuple(slice(None) if x is None else x for x in (pipe_parallel_rank, data_parallel_rank, model_parallel_rank))
        return [int(r) for r in np.atleast_1d(self._grid[sl]).flatten()]

    def coordinates_of(self, global_rank: int) -> tuple[int, int, int]:
        pp, dp, mp = np.argwhere(self._grid == global_rank)[0]
        return int(pp), int(dp), int(mp)


def local_slot_from_env() -> int:
    return int(os.environ.get("LOCAL_SLOT", os.environ.get("LOCAL_RANK", "0")))
