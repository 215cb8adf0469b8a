"""Parameter group = one flat arena (reference: core/optimizer/parameter_group.py:1-667).

The reference flattens a group's bf16 parameters into one buffer, keeps an fp32 copy of the rank's ZeRO partition and
walks Python lists of tensors for everything else.  Here the arena idea is taken to its end:

  * ``param_arena``  — all parameters of the group in model dtype, each 128-element aligned; ``p.data`` are views, so
                       adjacent projections (q|k|v, gate|up) can be used as one fused GEMM operand with zero copies;
  * ``grad_arena``   — same layout in fp32 (or the model dtype): ``p.main_grad`` views that the weight-gradient GEMMs
                       accumulate into directly from TMEM, across micro batches;
  * the rank's ZeRO-1 shard ``[shard_start, shard_end)`` of it carries the fp32 master weights and Adam moments.

A step is then a handful of whole-arena operations: reduce-scatter (or all-reduce) of ``grad_arena`` over DP, one fused
Σg² + non-finite kernel, one fused unscale·clip·AdamW·cast kernel, one all-gather of ``param_arena`` — independent of
the number of parameters.  Parameters whose gradient must be counted once in the global norm (not TP/PP duplicates)
are laid out first so de-duplication is a single boundary, not a per-parameter test.
"""
from __future__ import annotations

from typing import Any, Dict, Generator, List, Optional, Tuple

import torch
import torch.distributed as dist

from ...ops import _lib as native
from ...ops import kernels as K
from ..nn.parameter_meta import CoreParameterMeta
from ..topology import Topology
from .configs import OptimizerParamGroupConfig
from .lr import LearningRateScheduler

ALIGN = 128  # elements
NCCL_START_ALIGNMENT_FACTOR = ALIGN

NamedParameterMeta = Tuple[str, torch.Tensor, CoreParameterMeta]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def flatten_dense_tensors_aligned(tensor_list: List[torch.Tensor], alignment: int) -> Tuple[torch.Tensor, int]:
    """Concatenate flattened tensors and zero-pad the total to a multiple of ``alignment``; returns (flat, padding)."""
    total = sum(t.numel() for t in tensor_list)
    padded = _round_up(max(total, 1), alignment)
    flat = torch.zeros(padded, dtype=tensor_list[0].dtype, device=tensor_list[0].device)
    off = 0
    for t in tensor_list:
        flat[off:off + t.numel()].copy_(t.reshape(-1))
        off += t.numel()
    return flat, padded - total


def pack_frozen_parameters(parameters: List[torch.Tensor]) -> List[torch.Tensor]:
    """Move parameters that no optimizer group owns (finetuning: the frozen base weights) into one flat arena per
    (device, dtype) with the trained arenas' layout — registration order, ``ALIGN``-element slots.  The layers' fused
    forms (q|k|v as one GEMM, gate|up dual GEMM) take adjacent weights as one zero-copy [ΣN, K] view; without this a
    frozen model would fall back to one GEMM per projection.  ``p.data`` become views, values are kept; parameters
    are moved one at a time so the peak extra memory is one arena, not a second copy of the model.  Returns the arenas."""
    by_kind: Dict[Tuple[torch.device, torch.dtype], List[torch.Tensor]] = {}
    seen: set = set()
    for p in parameters:
        if id(p) in seen or p.requires_grad or p.numel() == 0:
            continue
        seen.add(id(p))
        by_kind.setdefault((p.device, p.dtype), []).append(p)
    arenas = []
    for (device, dtype), plist in by_kind.items():
        if len(plist) < 2:
            continue
        total = sum(_round_up(p.numel(), ALIGN) for p in plist)
        arena = torch.zeros(total, dtype=dtype, device=device)
        off = 0
        for p in plist:
            n = p.numel()
            view = arena[off:off + n].view(p.shape)
            view.copy_(p.data)
            p.data = view
            off += _round_up(n, ALIGN)
        arenas.append(arena)
    return arenas


def get_data_parallel_partitions(tensor: torch.Tensor, topology: Topology) -> List[torch.Tensor]:
    dp = topology.config.data_parallel_size
    assert tensor.numel() % dp == 0
    n = tensor.numel() // dp
    return [tensor.narrow(0, i * n, n) for i in range(dp)]


class OptimizerParamGroup:
    def __init__(self, named_parameters_with_meta: List[NamedParameterMeta], config: OptimizerParamGroupConfig):
        self.config = config
        self.learning_rate_scheduler = LearningRateScheduler(config.learning_rate_scheduler)
        self._named = list(named_parameters_with_meta)
        self.is_dummy = len(self._named) == 0
        self.initialized = False
        self._lr = 0.0

    # ------------------------------------------------------------------------------------------------
    def initialize(self, topology: Topology, zero: bool, grad_dtype: Optional[torch.dtype] = torch.float32) -> None:
        self.topology = topology
        self.zero = zero and topology.config.data_parallel_size > 1
        if self.is_dummy:
            # keeps every rank issuing the same collectives when this group is empty here (reference :118-134)
            device = topology.device
            dummy = torch.nn.Parameter(torch.zeros(1, dtype=torch.bfloat16 if device.type == "cuda" else torch.float32,
                                                   device=device))
            meta = CoreParameterMeta.register_on_parameter(dummy, is_model_parallel=False, layer_index=-1,
                                                           parameter_name="dummy_parameter")
            self._named = [("dummy_parameter", dummy, meta)]
        dtypes = {p.dtype for _, p, _ in self._named}
        assert len(dtypes) == 1, f"all parameters of a group must share a dtype, got {dtypes}"
        # the arenas live where the parameters live (a model kept on the host is optimised on the host, as with the
        # reference's torch optimizer); the topology's device is only the default for the empty group
        devices = {p.device for _, p, _ in self._named}
        assert len(devices) == 1, f"all parameters of a group must share a device, got {devices}"
        device = devices.pop()
        self.model_dtype = dtypes.pop()
        self.grad_dtype = grad_dtype or self.model_dtype
        mp_rank = topology.model_parallel_rank
        pp_rank = topology.pipe_parallel_rank

        def counted(meta: CoreParameterMeta) -> bool:
            if meta.is_model_parallel_duplicate and mp_rank != 0:
                return False
            if getattr(meta, "kv_replicas", 1) > 1 and mp_rank % meta.kv_replicas != 0:
                return False
            if meta.is_tied and getattr(meta, "tied_main_stage", pp_rank) != pp_rank:
                return False
            return True

        ordered = sorted(range(len(self._named)), key=lambda i: 0 if counted(self._named[i][2]) else 1)
        self.parameter_names = [self._named[i][0] for i in ordered]
        self.parameters_original: List[torch.Tensor] = [self._named[i][1] for i in ordered]
        self.parameter_metas = [self._named[i][2] for i in ordered]
        self.offsets: List[int] = []
        off = 0
        counted_end = None
        for p, m in zip(self.parameters_original, self.parameter_metas):
            if counted_end is None and not counted(m):
                counted_end = off
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.used = off
        self.counted_end = off if counted_end is None else counted_end
        dp = topology.config.data_parallel_size if self.zero else 1
        dp_all = topology.config.data_parallel_size
        self.total = _round_up(max(off, ALIGN), ALIGN * dp_all)  # divisible into dp chunks with or without ZeRO
        self._sym_params = self._sym_grads = None
        if dp_all > 1:
            from ...parallel.symm import SymmetricArena, fused_zero_enabled

            if fused_zero_enabled(topology) and self.grad_dtype in native.DTYPE_CODE and self.model_dtype in native.DTYPE_CODE:
                # arenas every DP rank can address (collective: all DP ranks build their groups in the same order).
                # ZeRO: the reduce-scatter becomes peer loads inside the reduce kernel and the all-gather peer stores
                # inside the AdamW kernel.  Plain DP: one kernel reads this rank's chunk from every peer and stores the
                # sum back into every peer (sb_dp_allreduce) — only the gradient arena has to be shared.
                if self.zero:
                    self._sym_params = SymmetricArena(self.total, self.model_dtype, topology.data_parallel_group, device)
                self._sym_grads = SymmetricArena(self.total, self.grad_dtype, topology.data_parallel_group, device)
        self.param_arena = self._sym_params.tensor if self._sym_params is not None else \
            torch.zeros(self.total, dtype=self.model_dtype, device=device)
        self.grad_arena = self._sym_grads.tensor if self._sym_grads is not None else \
            torch.zeros(self.total, dtype=self.grad_dtype, device=device)
        for p, o in zip(self.parameters_original, self.offsets):
            n = p.numel()
            view = self.param_arena[o:o + n].view(p.shape)
            view.copy_(p.data)
            p.data = view
            if not getattr(p, "_sb_no_main_grad", False):
                p.main_grad = self.grad_arena[o:o + n].view(p.shape)  # type: ignore[attr-defined]
        n_shard = self.total // dp
        r = topology.data_parallel_rank if self.zero else 0
        self.shard_start, self.shard_end = r * n_shard, (r + 1) * n_shard
        self.master = self.param_arena[self.shard_start:self.shard_end].float().clone()
        self.exp_avg = torch.zeros_like(self.master)
        self.exp_avg_sq = torch.zeros_like(self.master)
        self._use_native = device.type == "cuda" and native.available() and self.grad_dtype in native.DTYPE_CODE \
            and self.model_dtype in native.DTYPE_CODE
        self.initialized = True

    # ---- views ----------------------------------------------------------------------------------------
    @property
    def grad_shard(self) -> torch.Tensor:
        return self.grad_arena[self.shard_start:self.shard_end]

    @property
    def param_shard(self) -> torch.Tensor:
        return self.param_arena[self.shard_start:self.shard_end]

    def trainable_parameters_original(self) -> Generator[torch.Tensor, None, None]:
        yield from self.parameters_original

    def get_learning_rate(self) -> float:
        return self._lr

    # ---- step pieces ------------------------------------------------------------------------------------
    def fold_autograd_grads(self) -> None:
        """Gradients that arrived through plain autograd (``p.grad``) are added to the arena and released."""
        for p in self.parameters_original:
            if p.grad is not None:
                mg = getattr(p, "main_grad", None)
                if mg is None:
                    o = self.offsets[self._index_of(p)]
                    mg = self.grad_arena[o:o + p.numel()].view(p.shape)
                mg.add_(p.grad.to(mg.dtype))
                p.grad = None

    def _index_of(self, p: torch.Tensor) -> int:
        for i, q in enumerate(self.parameters_original):
            if q is p:
                return i
        raise KeyError

    def grad_views(self, predicate) -> List[torch.Tensor]:
        return [self.grad_arena[o:o + p.numel()] for p, o, m in zip(self.parameters_original, self.offsets, self.parameter_metas)
                if predicate(p, m)]

    def reduce_data_parallel(self) -> None:
        """Sum gradients over the DP group: true reduce-scatter into the owned shard under ZeRO (half the bytes of the
        reference's all-reduce-then-slice), all-reduce otherwise.  The 1/dp average is folded into the AdamW scale."""
        topo = self.topology
        if topo.config.data_parallel_size == 1 or not topo.is_distributed_initialized:
            return
        if self._sym_grads is not None and self.zero:
            n = self.shard_end - self.shard_start
            self._sym_grads.barrier()  # every rank's backward (its writes to its own arena) is complete
            native.call("sb_zero_reduce", self._sym_grads.ptr_array(self.shard_start), self._sym_grads.world,
                        native.ptr(self.grad_shard), n, native.DTYPE_CODE[self.grad_dtype])
        elif self._sym_grads is not None:
            # plain DP: this rank sums chunk `dp_rank` of every arena and stores it into every arena
            dp = topo.config.data_parallel_size
            n = self.total // dp
            self._sym_grads.barrier()
            native.call("sb_dp_allreduce", self._sym_grads.ptr_array(topo.data_parallel_rank * n), dp, n,
                        native.DTYPE_CODE[self.grad_dtype])
            self._sym_grads.barrier()  # every chunk of this rank's arena has been written by its owner
        elif self.zero and topo.backend == "nccl":
            dist.reduce_scatter_tensor(self.grad_shard, self.grad_arena, group=topo.data_parallel_group)
        else:
            dist.all_reduce(self.grad_arena, group=topo.data_parallel_group)

    def local_sumsq(self, out: torch.Tensor, nonfinite: torch.Tensor) -> None:
        """out[0] += Σ g² over the part of this rank's shard that is counted once globally."""
        lo, hi = self.shard_start, min(self.shard_end, self.counted_end)
        lo2, hi2 = max(self.shard_start, self.counted_end), self.shard_end  # duplicates: overflow check only
        if self._use_native:
            if hi > lo:
                K.grad_sumsq(self.grad_arena[lo:hi], out, nonfinite, accumulate=True)
            if hi2 > lo2:
                scratch = torch.zeros(1, dtype=torch.float32, device=out.device)
                K.grad_sumsq(self.grad_arena[lo2:hi2], scratch, nonfinite, accumulate=False)
        else:
            if hi > lo:
                g = self.grad_arena[lo:hi].float()
                out += (g * g).sum()
            g_all = self.grad_shard
            nonfinite |= (~torch.isfinite(g_all.float()).all()).to(nonfinite.dtype)

    def apply_adamw(self, step_index: int, beta1: float, beta2: float, eps: float, inv_scale: float, clip: float,
                    norm_sq: Optional[torch.Tensor], skip: Optional[torch.Tensor]) -> None:
        self._lr = self.learning_rate_scheduler.get_lr(step_index)
        wd = self.config.weight_decay
        if self._sym_params is not None:
            import math

            n = self.shard_end - self.shard_start
            native.call("sb_adamw_zero", native.ptr(self.master), native.ptr(self.exp_avg), native.ptr(self.exp_avg_sq),
                        native.ptr(self.grad_shard), self._sym_params.ptr_array(self.shard_start), self._sym_params.world, n,
                        native.DTYPE_CODE[self.grad_dtype], native.DTYPE_CODE[self.model_dtype], float(self._lr), float(beta1),
                        float(beta2), float(eps), float(wd), int(step_index), float(inv_scale), float(clip),
                        native.ptr(norm_sq), native.ptr(skip), 1)
            return
        if self._use_native:
            K.adamw_step(self.master, self.exp_avg, self.exp_avg_sq, self.grad_shard, self.param_shard, lr=self._lr,
                         beta1=beta1, beta2=beta2, eps=eps, weight_decay=wd, step=step_index,
                         inv_loss_scale=inv_scale, clip=clip, norm_sq=norm_sq, skip=skip, zero_grad=True)
            return
        g = self.grad_shard.float() * inv_scale
        if clip > 0 and norm_sq is not None:
            coef = clip / (norm_sq.sqrt() * inv_scale + 1e-6)
            g = g * torch.clamp(coef, max=1.0)
        keep = None if skip is None else (skip != 0)
        m = self.exp_avg * beta1 + (1 - beta1) * g
        v = self.exp_avg_sq * beta2 + (1 - beta2) * g * g
        bc1 = 1 - beta1 ** step_index
        bc2 = 1 - beta2 ** step_index
        p = self.master * (1 - self._lr * wd) - (self._lr / bc1) * m / (v.sqrt() / (bc2 ** 0.5) + eps)
        if keep is not None:
            m = torch.where(keep, self.exp_avg, m)
            v = torch.where(keep, self.exp_avg_sq, v)
            p = torch.where(keep, self.master, p)
        self.exp_avg.copy_(m)
        self.exp_avg_sq.copy_(v)
        self.master.copy_(p)
        self.param_shard.copy_(p.to(self.model_dtype))
        self.grad_shard.zero_()

    def zero_grad(self, set_to_none: bool = True) -> None:
        self.grad_arena.zero_()
        for p in self.parameters_original:
            p.grad = None

    def zero_unowned_grads(self) -> None:
        if self.shard_start > 0:
            self.grad_arena[:self.shard_start].zero_()
        if self.shard_end < self.total:
            self.grad_arena[self.shard_end:].zero_()

    def gather_params(self) -> None:
        """ZeRO: all-gather the updated model-dtype shards in place (the cast happened inside the AdamW kernel)."""
        topo = self.topology
        if not self.zero or not topo.is_distributed_initialized:
            return
        if self._sym_params is not None:
            # the AdamW kernels already stored every shard into every arena; wait until the peers' stores have landed
            # here (and until nobody reads this rank's gradient arena any more)
            self._sym_params.barrier()
            return
        if topo.backend == "nccl":
            dist.all_gather_into_tensor(self.param_arena, self.param_shard, group=topo.data_parallel_group)
        else:
            parts = [torch.empty_like(self.param_shard) for _ in range(topo.config.data_parallel_size)]
            dist.all_gather(parts, self.param_shard.contiguous(), group=topo.data_parallel_group)
            self.param_arena.copy_(torch.cat(parts))

    def refresh_optimized_params(self, topology: Optional[Topology] = None) -> None:
        """Master weights <- current model weights (after a checkpoint load or a LoRA merge changed them)."""
        self.master.copy_(self.param_shard.float())

    # ---- full (DP-gathered) state for checkpoints ----------------------------------------------------------
    def gather_full_state(self, which: torch.Tensor) -> torch.Tensor:
        """fp32 arena-sized tensor assembled from every DP rank's shard of ``which`` (master / exp_avg / exp_avg_sq)."""
        topo = self.topology
        if not self.zero or not topo.is_distributed_initialized:
            return which
        full = torch.empty(self.total, dtype=which.dtype, device=which.device)
        if topo.backend == "nccl":
            dist.all_gather_into_tensor(full, which.contiguous(), group=topo.data_parallel_group)
        else:
            parts = [torch.empty_like(which) for _ in range(topo.config.data_parallel_size)]
            dist.all_gather(parts, which.contiguous(), group=topo.data_parallel_group)
            full.copy_(torch.cat(parts))
        return full

    def scatter_full_state(self, which: torch.Tensor, offset: int, value: torch.Tensor) -> None:
        """Write ``value`` (flat, full parameter) at arena ``offset`` into this rank's shard of ``which``."""
        lo, hi = max(offset, self.shard_start), min(offset + value.numel(), self.shard_end)
        if hi > lo:
            which[lo - self.shard_start:hi - self.shard_start].copy_(value[lo - offset:hi - offset].to(which.dtype))

    def state_dict(self) -> Dict[str, Any]:
        return {"parameter_names": self.parameter_names, "offsets": self.offsets, "master": self.master,
                "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "shard": (self.shard_start, self.shard_end)}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        assert tuple(state_dict["shard"]) == (self.shard_start, self.shard_end), "static optimizer state needs the same layout"
        for k in ("master", "exp_avg", "exp_avg_sq"):
            getattr(self, k).copy_(state_dict[k].to(self.master.device))

    def log_state(self) -> None:
        from ..logging import logger

        logger.debug(f"param group {self.config.name}: {len(self.parameters_original)} parameters, arena {self.total}, "
                     f"shard [{self.shard_start}, {self.shard_end})")
