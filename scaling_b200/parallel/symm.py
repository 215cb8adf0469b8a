"""Symmetric (peer-mapped) workspaces for the fused GEMM + collective kernels.

A :class:`TPComm` belongs to one tensor-parallel process group.  It owns one CUDA-VMM allocation per rank that every
rank of the group maps into its own address space (``torch.distributed._symmetric_memory`` does the allocation and the
handle exchange; the kernels only ever see raw pointers), laid out as

    [ parity 0 | parity 1 ] x [ sender 0 | … | sender n-1 ] x partial tiles (bf16)     "data"
    [ parity 0 | parity 1 ] x [ sender 0 | … | sender n-1 ] x one u32 per (CTA tile, epilogue warp)   "flags"

The fused kernels (``csrc/gemm2_sm100.cu``, epilogue ``kAllReduce``) push partial tiles into slot ``rank`` of every
peer and publish ``flag = epoch``; ``epoch`` is the per-workspace call counter (identical on all ranks because TP ranks
execute the same program), ``parity = epoch & 1`` double-buffers the slots so a fast rank that already started call
c+1 cannot overwrite tiles a slow rank is still reducing for call c (it cannot reach c+2 before the slow rank has
pushed — and therefore finished launching — c+1).

Reference counterpart: none — the reference reduces with ``torch.distributed.all_reduce`` after the GEMM
(core/nn/linear/row_parallel_linear.py:150-160, core/nn/linear/utils.py:20-60).
"""
from __future__ import annotations

import ctypes
import os
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist

from ..ops import _lib

_COMMS: Dict[int, "TPComm"] = {}
_ENABLED = os.environ.get("SCALING_B200_FUSED_TP", "1") == "1"
_FLAG_ALIGN = 1024


def _enable_symm_mem(group_name: str) -> None:
    """Older torch builds need the group registered for symmetric memory before the first rendezvous; newer ones
    register every group themselves and only warn about the call."""
    import warnings

    import torch.distributed._symmetric_memory as symm_mem

    fn = getattr(symm_mem, "enable_symm_mem_for_group", None)
    if fn is None:
        return
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", FutureWarning)
        try:
            fn(group_name)
        except Exception:  # noqa: BLE001  (already enabled / not required by this torch)
            pass


def fused_tp_enabled() -> bool:
    return _ENABLED


def set_fused_tp(enabled: bool) -> None:
    global _ENABLED
    _ENABLED = bool(enabled)


class TPComm:
    """Peer-memory communicator of one tensor-parallel group (one per process)."""

    def __init__(self, group: Any, device: torch.device):
        self.group = group
        self.device = device
        self.size = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.slot_bytes = 0
        self.flag_slot = 0
        self.calls = 0
        self._need: Dict[tuple, int] = {}
        self._buf: Optional[torch.Tensor] = None
        self._hdl: Any = None
        self._data_ptrs = [None, None]
        self._flag_ptrs = [None, None]
        self._xbufs: Dict[str, Any] = {}
        self._barrier: Optional["GroupBarrier"] = None

    # ---- workspace -------------------------------------------------------------------------------------------
    def _allocate(self, slot_bytes: int) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        torch.cuda.synchronize(self.device)
        name = self.group.group_name
        _enable_symm_mem(name)
        n = self.size
        slot_bytes = (slot_bytes + 65535) // 65536 * 65536
        flag_slot = (slot_bytes // 65536 * 4 + _FLAG_ALIGN - 1) // _FLAG_ALIGN * _FLAG_ALIGN
        data_bytes = 2 * n * slot_bytes
        flag_bytes = 2 * n * flag_slot * 4
        buf = symm_mem.empty(data_bytes + flag_bytes, dtype=torch.uint8, device=self.device)
        hdl = symm_mem.rendezvous(buf, name)
        buf.zero_()
        torch.cuda.synchronize(self.device)
        hdl.barrier()
        torch.cuda.synchronize(self.device)
        self._buf, self._hdl = buf, hdl
        self.slot_bytes, self.flag_slot = slot_bytes, flag_slot
        self.calls = 0
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        arr = ctypes.c_ulonglong * n
        for parity in (0, 1):
            self._data_ptrs[parity] = arr(*[p + parity * n * slot_bytes for p in ptrs])
            self._flag_ptrs[parity] = arr(*[p + data_bytes + parity * n * flag_slot * 4 for p in ptrs])

    def reserve(self, m: int, n: int) -> None:
        """Make sure an [m, n] fused all-reduce fits.  Collective when it has to grow (all ranks ask for the same
        shapes in the same order, so they grow together)."""
        need = self._need.get((m, n))
        if need is None:
            need = self._need[(m, n)] = int(_lib.host_lib().sb_gemm2_allreduce_slot_bytes(m, n))
        if need > self.slot_bytes:
            self._allocate(need)

    def next_call(self):
        """(epoch, data_ptrs, flag_ptrs) for the next fused call on this workspace."""
        self.calls += 1
        parity = self.calls & 1
        return self.calls, self._data_ptrs[parity], self._flag_ptrs[parity]

    # ---- exchange buffers (embedding rows, sequence-parallel gathers) ---------------------------------------------
    _GATE_BYTES = 64 * 1024  # u32 arrival counters per buffer half (16 Ki row tiles)

    def exchange_buffer_gated(self, tag: str, nbytes: int):
        """As :meth:`exchange_buffer`, plus per-half arrival counters for the gated all-gather -> GEMM path: returns
        (local tensor, peer data pointers, local counter address, peer counter pointers, uses of this half so far
        including this one).  Every use of a tag must have the same shape (counters count uses)."""
        view, ptrs = self.exchange_buffer(tag, nbytes, _gates=True)
        ent = self._xbufs[tag]
        half = ent["calls"] & 1
        off = 2 * ent["bytes"] + half * self._GATE_BYTES
        arr = ctypes.c_ulonglong * self.size
        uses = (ent["calls"] + 1) // 2 if half else ent["calls"] // 2
        return view, ptrs, ent["ptrs"][self.rank] + off, arr(*[p + off for p in ent["ptrs"]]), uses

    def exchange_buffer(self, tag: str, nbytes: int, _gates: bool = False):
        """(local tensor, peer pointer array) of a double-buffered symmetric byte buffer named ``tag``: peers store into
        it with plain NVLink stores, a :meth:`barrier` publishes them.  Successive calls alternate between two halves,
        so a rank that is one call ahead never overwrites data a slower peer is still reading (it cannot get two calls
        ahead: every call contains a barrier).  Collective when it has to (re)allocate."""
        import torch.distributed._symmetric_memory as symm_mem

        ent = self._xbufs.get(tag)
        if ent is None or ent["bytes"] < nbytes:
            torch.cuda.synchronize(self.device)
            cap = (max(nbytes, 1 << 20) + 65535) // 65536 * 65536
            buf = symm_mem.empty(2 * cap + 2 * self._GATE_BYTES, dtype=torch.uint8, device=self.device)
            hdl = symm_mem.rendezvous(buf, self.group.group_name)
            buf.zero_()  # arrival counters start at zero
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
            ent = self._xbufs[tag] = {"bytes": cap, "buf": buf, "hdl": hdl, "ptrs": [int(p) for p in hdl.buffer_ptrs],
                                      "calls": 0}
        ent["calls"] += 1
        half = ent["calls"] & 1
        off = half * ent["bytes"]
        arr = ctypes.c_ulonglong * self.size
        return ent["buf"][off:off + nbytes], arr(*[p + off for p in ent["ptrs"]])

    def barrier(self) -> None:
        """Device-side barrier over the TP group on the current stream (own kernel)."""
        if self._barrier is None:
            self._barrier = GroupBarrier.get(self.group, self.device)
        self._barrier()

    # ---- torch.distributed expressions of the same collectives (guards / oracle) ---------------------------------
    def all_reduce_(self, x: torch.Tensor) -> torch.Tensor:
        dist.all_reduce(x, group=self.group)
        return x


_FUSED_ZERO = os.environ.get("SCALING_B200_FUSED_ZERO", "1") == "1"


def set_fused_zero(enabled: bool) -> None:
    global _FUSED_ZERO
    _FUSED_ZERO = bool(enabled)


def fused_zero_enabled(topology: Any) -> bool:
    """Data-parallel gradient reduction as peer loads / stores inside this library's kernels (ZeRO-1: reduce-scatter in
    ``sb_zero_reduce`` + all-gather in ``sb_adamw_zero``; plain DP: ``sb_dp_allreduce``): NCCL job on CUDA, 2..8
    data-parallel ranks (one NVLink domain), native library present."""
    if topology is None or not _FUSED_ZERO or not topology.is_distributed_initialized:
        return False
    dp = topology.config.data_parallel_size
    return (2 <= dp <= 8 and topology.backend == "nccl" and topology.device.type == "cuda" and _lib.available())


class GroupBarrier:
    """Device-side barrier of one process group: this library's own kernel (``sb_device_barrier``, csrc/runtime.cu) over
    a few flag words in symmetric memory — one launch on the current stream, no host synchronisation, no NCCL."""

    _by_group: Dict[str, "GroupBarrier"] = {}

    def __init__(self, group: Any, device: torch.device):
        import torch.distributed._symmetric_memory as symm_mem

        name = group.group_name
        _enable_symm_mem(name)
        self._flags = symm_mem.empty(64, dtype=torch.int32, device=device)
        self._hdl = symm_mem.rendezvous(self._flags, name)
        self._flags.zero_()
        torch.cuda.synchronize(device)
        dist.barrier(group=group)  # every rank's flags are zero before anybody may post epoch 1
        self.rank, self.world = self._hdl.rank, self._hdl.world_size
        arr = ctypes.c_ulonglong * self.world
        self._ptrs = arr(*[int(p) for p in self._hdl.buffer_ptrs])
        self.epoch = 0

    @classmethod
    def get(cls, group: Any, device: torch.device) -> "GroupBarrier":
        key = f"{group.group_name}@{device.index}"
        b = cls._by_group.get(key)
        if b is None:
            b = cls._by_group[key] = GroupBarrier(group, device)
        return b

    def __call__(self) -> None:
        self.epoch = (self.epoch + 1) & 0xFFFFFFFF
        _lib.call("sb_device_barrier", self._ptrs, self.rank, self.world, self.epoch)


class SymmetricArena:
    """A flat tensor whose storage every rank of ``group`` can address: ``peer_ptrs[r]`` is rank r's copy as mapped
    into this process.  Used for the ZeRO parameter and gradient arenas."""

    def __init__(self, numel: int, dtype: torch.dtype, group: Any, device: torch.device):
        import torch.distributed._symmetric_memory as symm_mem

        name = group.group_name
        _enable_symm_mem(name)
        self._buf = symm_mem.empty(numel, dtype=dtype, device=device)
        self._hdl = symm_mem.rendezvous(self._buf, name)
        self._buf.zero_()
        self.tensor = self._buf
        self.peer_ptrs = [int(p) for p in self._hdl.buffer_ptrs]
        self.rank = self._hdl.rank
        self.world = self._hdl.world_size
        self.itemsize = self._buf.element_size()
        self._barrier = GroupBarrier.get(group, device)

    def barrier(self) -> None:
        """Device-side barrier over the group on the current stream (own kernel, see :class:`GroupBarrier`)."""
        self._barrier()

    def ptr_array(self, element_offset: int):
        arr = ctypes.c_ulonglong * self.world
        return arr(*[p + element_offset * self.itemsize for p in self.peer_ptrs])


def comm_for_group(group: Any) -> Optional[TPComm]:
    """The already-created communicator of a process group (None when the fused paths are off or not set up)."""
    if not _ENABLED or group is None:
        return None
    for comm in _COMMS.values():
        if comm.group is group:
            return comm
    return None


def get_tp_comm(topology: Any, sequence_parallel: bool = False) -> Optional[TPComm]:
    """The fused-collective communicator of ``topology``'s TP group, or None when the fused path does not apply
    (single TP rank, CPU / gloo, disabled by SCALING_B200_FUSED_TP=0).  Plain tensor parallelism uses it for the fused
    GEMM + all-reduce; ``sequence_parallel=True`` returns it for sequence-parallel jobs (fused GEMM + reduce-scatter)."""
    if topology is None or not _ENABLED:
        return None
    c = topology.config
    if c.model_parallel_size == 1 or not topology.is_distributed_initialized:
        return None
    if bool(c.sequence_parallel) != bool(sequence_parallel):  # callers ask for the communicator of THEIR mode
        return None
    if topology.backend != "nccl" or topology.device.type != "cuda" or not _lib.available():
        return None
    if c.model_parallel_size > 8:
        return None
    key = id(topology)
    comm = _COMMS.get(key)
    if comm is None:
        comm = TPComm(topology.model_parallel_group, topology.device)
        _COMMS[key] = comm
    return comm
