"""Stage-to-stage transfer of arbitrary pytrees (reference: core/nn/parallel_module/communicator.py:23-511).

Protocol: the first message between two ranks carries pickled metadata (tree structure, per-leaf shape/dtype/
requires_grad, and the values of non-tensor leaves); afterwards only tensor payloads travel — one ``isend``/``irecv``
per tensor leaf, in leaf order.  ``use_continuous_recommunication=True`` repeats the handshake on every message
(variable shapes, e.g. inference with growing sequences).  Sends are asynchronous: the work handles (and the tensors
they read) are parked until the next send on the same communicator or ``drain()``, so the sender's next compute
overlaps the transfer; receives block because their data is needed immediately.  Gradients flow the other way for
exactly the leaves that were marked ``requires_grad``.
"""
from __future__ import annotations

import pickle
from typing import Any, List, NamedTuple, Optional, Tuple

import torch
import torch.distributed as dist
from torch.utils._pytree import tree_flatten, tree_unflatten


class GradOutput(NamedTuple):
    tensors: List[torch.Tensor]
    grad_tensors: List[torch.Tensor]


class CommunicationMetaBase(NamedTuple):
    treespec: Any
    leaves: Tuple[Any, ...]  # ("t", shape, dtype, requires_grad) | ("o", value)


def _wire(t: torch.Tensor, byte_wire: bool) -> torch.Tensor:
    t = t.detach().contiguous()
    if byte_wire:
        if t.dtype == torch.bool:
            t = t.to(torch.uint8)
        return t.reshape(-1).view(torch.uint8)
    if t.dtype == torch.bool:
        return t.to(torch.uint8)
    return t


class PipeCommunicator:
    def __init__(self, local_device: torch.device, recv_grads: bool, recv_data: bool,
                 use_continuous_recommunication: bool = False, group: Any = None, grad_group: Any = None):
        self.local_device = local_device
        self.recv_grads_enabled = recv_grads
        self.recv_data_enabled = recv_data
        self.use_continuous_recommunication = use_continuous_recommunication
        self.group = group  # activations (and handshakes)
        self.grad_group = grad_group if grad_group is not None else group  # gradients travel the other way
        self.byte_wire = local_device.type != "cuda"  # gloo: ship raw bytes, every dtype works
        # cached handshakes, one per mode: training and evaluation messages differ in their requires_grad flags (and
        # possibly shapes), and every rank switches mode at the same step, so each mode negotiates once
        self._metas: dict = {}
        self.mode = "train"
        self._pending: List[Tuple[Any, torch.Tensor]] = []

    @property
    def _send_meta(self) -> Optional[CommunicationMetaBase]:
        return self._metas.get((self.mode, "send"))

    @_send_meta.setter
    def _send_meta(self, value: Optional[CommunicationMetaBase]) -> None:
        self._metas[(self.mode, "send")] = value

    @property
    def _recv_meta(self) -> Optional[CommunicationMetaBase]:
        return self._metas.get((self.mode, "recv"))

    @_recv_meta.setter
    def _recv_meta(self, value: Optional[CommunicationMetaBase]) -> None:
        self._metas[(self.mode, "recv")] = value

    def set_mode(self, mode: str) -> None:
        self.mode = mode

    def reset_communication_meta(self) -> None:
        self._metas.clear()

    # ---- low level ----------------------------------------------------------------------------------
    def drain(self) -> None:
        for work, _keep in self._pending:
            work.wait()
        self._pending.clear()

    def _isend(self, t: torch.Tensor, dst: int, grad: bool = False) -> None:
        w = _wire(t, self.byte_wire)
        self._pending.append((dist.isend(w, dst=dst, group=self.grad_group if grad else self.group), w))

    def _recv_into(self, shape: Tuple[int, ...], dtype: torch.dtype, src: int, grad: bool = False) -> torch.Tensor:
        group = self.grad_group if grad else self.group
        wire_dtype = torch.uint8 if dtype == torch.bool else dtype
        if self.byte_wire:
            itemsize = torch.empty((), dtype=wire_dtype).element_size()
            n = int(torch.Size(shape).numel())
            buf = torch.empty(n * itemsize, dtype=torch.uint8, device=self.local_device)
            dist.recv(buf, src=src, group=group)
            out = buf.view(wire_dtype).reshape(shape)
        else:
            out = torch.empty(shape, dtype=wire_dtype, device=self.local_device)
            dist.irecv(out, src=src, group=group).wait()
        return out.to(torch.bool) if dtype == torch.bool else out

    def _send_object(self, obj: Any, dst: int) -> None:
        payload = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8).to(self.local_device)
        self._isend(torch.tensor([payload.numel()], dtype=torch.long, device=self.local_device), dst)
        self._isend(payload, dst)
        self.drain()

    def _recv_object(self, src: int) -> Any:
        n = self._recv_into((1,), torch.long, src)
        payload = self._recv_into((int(n.item()),), torch.uint8, src)
        return pickle.loads(payload.cpu().numpy().tobytes())

    # ---- activations --------------------------------------------------------------------------------
    def send_data(self, data: Any, target_global_rank: int) -> None:
        self.drain()  # previous message must have left before its buffers may be recycled
        leaves, spec = tree_flatten(data)
        if self._send_meta is None or self.use_continuous_recommunication:
            described = tuple(("t", tuple(x.shape), x.dtype, bool(x.requires_grad)) if torch.is_tensor(x) else ("o", x)
                              for x in leaves)
            self._send_meta = CommunicationMetaBase(spec, described)
            self._send_object(self._send_meta, target_global_rank)
        for x in leaves:
            if torch.is_tensor(x):
                self._isend(x, target_global_rank)

    def recv_data(self, origin_global_rank: int) -> Any:
        if self._recv_meta is None or self.use_continuous_recommunication:
            self._recv_meta = self._recv_object(origin_global_rank)
        leaves = []
        for d in self._recv_meta.leaves:
            if d[0] == "t":
                _, shape, dtype, requires_grad = d
                t = self._recv_into(shape, dtype, origin_global_rank)
                if requires_grad and t.is_floating_point():
                    t.requires_grad_(True)
                leaves.append(t)
            else:
                leaves.append(d[1])
        return tree_unflatten(leaves, self._recv_meta.treespec)

    # ---- gradients ----------------------------------------------------------------------------------
    def send_gradients(self, data: Any, target_global_rank: int) -> None:
        """``data`` is the pytree that was received as stage input; ships ``.grad`` of its differentiable leaves."""
        self.drain()
        leaves, _ = tree_flatten(data)
        for x in leaves:
            if torch.is_tensor(x) and x.requires_grad:
                g = x.grad if x.grad is not None else torch.zeros_like(x)
                self._isend(g, target_global_rank, grad=True)

    def recv_gradients(self, data: Any, origin_global_rank: int) -> GradOutput:
        """``data`` is the pytree that was sent as stage output; returns the tensors to backpropagate with their
        received gradients."""
        leaves, _ = tree_flatten(data)
        tensors, grads = [], []
        for x in leaves:
            if torch.is_tensor(x) and x.requires_grad:
                tensors.append(x)
                grads.append(self._recv_into(tuple(x.shape), x.dtype, origin_global_rank, grad=True))
        return GradOutput(tensors, grads)


class ModelParallelCommunicator:
    """Broadcast an arbitrary pytree from TP rank 0 to its tensor-parallel group (reference communicator.py:513-759;
    unused by the training path, kept for inference-style input sync).

    Same protocol as :class:`PipeCommunicator` with broadcasts instead of send/recv: the first call (and every call
    under ``use_continuous_recommunication``, or after :meth:`reset_communication_meta`) ships the pickled structure
    (treespec + per-leaf shape/dtype/requires_grad, non-tensor leaves by value), then one broadcast per tensor leaf.
    """

    def __init__(self, topology: Any, use_continuous_recommunication: bool = False):
        self.topology = topology
        self.local_device = topology.device
        self._use_continuous_recommunication = use_continuous_recommunication
        self.byte_wire = self.local_device.type != "cuda"
        self._meta: Optional[CommunicationMetaBase] = None

    @property
    def _src(self) -> int:
        return self.topology.model_parallel_ranks[0]

    def _bcast(self, t: torch.Tensor) -> None:
        dist.broadcast(t, src=self._src, group=self.topology.model_parallel_group)

    def _bcast_object(self, obj: Any = None) -> Any:
        sender = self.topology.model_parallel_rank == 0
        if sender:
            payload = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8).to(self.local_device)
            n = torch.tensor([payload.numel()], dtype=torch.long, device=self.local_device)
        else:
            n = torch.zeros(1, dtype=torch.long, device=self.local_device)
        self._bcast(n)
        if not sender:
            payload = torch.empty(int(n.item()), dtype=torch.uint8, device=self.local_device)
        self._bcast(payload)
        return obj if sender else pickle.loads(payload.cpu().numpy().tobytes())

    def _needs_meta(self) -> bool:
        return self._meta is None or self._use_continuous_recommunication

    def send_data(self, data: Any) -> None:
        leaves, spec = tree_flatten(data)
        if self._needs_meta():
            described = tuple(("t", tuple(x.shape), x.dtype, bool(x.requires_grad)) if torch.is_tensor(x) else ("o", x)
                              for x in leaves)
            self._meta = CommunicationMetaBase(spec, described)
            self._bcast_object(self._meta)
        for x in leaves:
            if torch.is_tensor(x):
                self._bcast(_wire(x.to(self.local_device), self.byte_wire))

    def recv_data(self) -> Any:
        if self._needs_meta():
            self._meta = self._bcast_object()
        assert self._meta is not None
        leaves = []
        for d in self._meta.leaves:
            if d[0] != "t":
                leaves.append(d[1])
                continue
            _, shape, dtype, requires_grad = d
            wire_dtype = torch.uint8 if dtype == torch.bool else dtype
            if self.byte_wire:
                n = int(torch.Size(shape).numel()) * torch.empty((), dtype=wire_dtype).element_size()
                buf = torch.empty(n, dtype=torch.uint8, device=self.local_device)
                self._bcast(buf)
                t = buf.view(wire_dtype).reshape(shape)
            else:
                t = torch.empty(shape, dtype=wire_dtype, device=self.local_device)
                self._bcast(t)
            t = t.to(torch.bool) if dtype == torch.bool else t
            if requires_grad and t.is_floating_point():
                t.requires_grad_(True)
            leaves.append(t)
        return tree_unflatten(leaves, self._meta.treespec)

    def sync_data(self, data: Any) -> Any:
        """TP rank 0 passes ``data`` and gets it back; every other rank passes anything (ignored) and receives it."""
        if self.topology.config.model_parallel_size == 1:
            return data
        if self.topology.model_parallel_rank == 0:
            self.send_data(data)
            return data
        return self.recv_data()

    def reset_communication_meta(self, use_continuous_recommunication: Optional[bool] = None) -> None:
        self._meta = None
        if use_continuous_recommunication is not None:
            self._use_continuous_recommunication = use_continuous_recommunication
