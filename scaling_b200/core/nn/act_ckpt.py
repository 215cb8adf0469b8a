"""Activation checkpointing that is exact under dropout (reference:
core/nn/parallel_module/activation_checkpointing.py:1-190).

Non-reentrant: ``saved_tensors_hooks`` replace every tensor autograd wants to save by a slot; at the first unpack the
forward is replayed once with the device RNG *and* the tensor-parallel-constant RNG restored to their values at the
original forward, and the slots are filled in order.  The Philox dropout kernels carry no RNG state — their mask
depends only on (seed, offset); the replay restores the topology's offset counter so the same ranges are drawn.
"""
from __future__ import annotations

import weakref
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch


class _Holder:
    pass


def _rng_snapshot(topology: Any, device: Optional[torch.device]) -> Dict[str, Any]:
    snap: Dict[str, Any] = {"cpu": torch.get_rng_state()}
    if device is not None and device.type == "cuda":
        snap["cuda"] = torch.cuda.get_rng_state(device)
    if topology is not None and topology.has_model_parallel_constant_rng():
        snap["mp"] = topology._rng.state.clone()
    if topology is not None and hasattr(topology, "dropout_offset"):
        snap["dropout_offset"] = topology.dropout_offset  # Philox dropout: replay the same counter ranges
    return snap


def _rng_restore(snap: Dict[str, Any], topology: Any, device: Optional[torch.device]) -> None:
    torch.set_rng_state(snap["cpu"])
    if "cuda" in snap:
        torch.cuda.set_rng_state(snap["cuda"], device)
    if "mp" in snap:
        topology._rng.state = snap["mp"].clone()
    if "dropout_offset" in snap:
        topology.dropout_offset = snap["dropout_offset"]


def _checkpoint_without_reentrant(function: Callable[..., Any], topology: Any, *args: Any) -> Any:
    device = next((a.device for a in args if torch.is_tensor(a)), None)
    snap = _rng_snapshot(topology, device)
    autocast = torch.is_autocast_enabled()
    storage: "weakref.WeakKeyDictionary[_Holder, Any]" = weakref.WeakKeyDictionary()
    holders: List["weakref.ReferenceType[_Holder]"] = []

    def pack(_t: torch.Tensor) -> _Holder:
        h = _Holder()
        holders.append(weakref.ref(h))
        return h

    def unpack(h: _Holder) -> torch.Tensor:
        if len(storage) == 0:
            counter = [0]

            def inner_pack(t: torch.Tensor) -> None:
                i = counter[0]
                counter[0] += 1
                ref = holders[i]() if i < len(holders) else None
                if ref is not None:
                    storage[ref] = t
                return None

            def inner_unpack(_x: Any) -> torch.Tensor:
                raise RuntimeError("backward through a recomputation is not supported")

            now = _rng_snapshot(topology, device)
            _rng_restore(snap, topology, device)
            try:
                with torch.enable_grad(), torch.autograd.graph.saved_tensors_hooks(inner_pack, inner_unpack), \
                        torch.autocast(device_type=device.type if device is not None else "cpu", enabled=autocast):
                    detached = tuple(a.detach().requires_grad_(a.requires_grad) if torch.is_tensor(a) else a for a in args)
                    function(*detached)
            finally:
                _rng_restore(now, topology, device)
        if h not in storage:
            raise RuntimeError("checkpointed tensor was not recomputed (function is not deterministic in its saves)")
        return storage[h]

    with torch.autograd.graph.saved_tensors_hooks(pack, unpack):
        return function(*args)


def checkpoint(function: Callable[..., Any], topology: Any, *args: Any) -> Any:
    return _checkpoint_without_reentrant(function, topology, *args)
