"""Mixed-precision AdamW with ZeRO-1 over flat arenas, and its layout-independent checkpoints.

Reference: core/optimizer/optimizer.py:1-734 (step order :107-208, checkpoint :335-734), allreduce.py, base.py.
Step order is the reference's — overflow check → SP/TP-replica gradient reduce → DP reduce → global norm → clip →
AdamW on fp32 master → cast back / ZeRO all-gather → zero grads — executed as whole-arena operations (see
``param_group.py``) with at most one host synchronisation (reading the norm / overflow flag for the step output).

Checkpoint contract: ``optimizer_state_layer_{i}.pt`` with ``step_index``, ``loss_scaler``,
``optimizer_param_groups`` and ``parameters[meta.key] = {parameter (fp32, TP-merged), meta, optimizer_state{step,
exp_avg, exp_avg_sq}}`` — independent of (mp, pp, dp, zero), so a run can resume under a different layout.
"""
from __future__ import annotations

import math
from pathlib import Path
from typing import Any, Dict, List, NamedTuple, Optional, Union

import torch
import torch.distributed as dist

from ..config import BaseConfig
from ..logging import logger
from ..nn.parameter_meta import CoreParameterMeta
from ..topology import Topology
from ..utils.param_merge import merge_parameter, split_parameter
from .configs import OptimizerConfig
from .loss_scaler import LossScaler
from .param_group import OptimizerParamGroup


def _weights_changed() -> None:
    """The arena kernels rewrite parameters without autograd noticing: drop per-step caches derived from weights."""
    from ...ops.functional import bump_weight_epoch

    bump_weight_epoch()


class OptimizerStepOutput(NamedTuple):
    global_grad_norm: Optional[float]
    global_grad_norm_clipped: Optional[float]
    learning_rates: Optional[Dict[str, float]]
    overflow: Optional[bool]
    no_overflow_steps: Optional[int]
    current_loss_scale: Optional[float]
    debug_dict: Optional[Dict[str, float]]


class BaseOptimizer:
    def __init__(self, config: BaseConfig) -> None:
        self.config = config

    def __repr__(self) -> str:
        return f"{self.__class__.__name__} (config: {self.config})"

    def step(self) -> OptimizerStepOutput:
        raise NotImplementedError

    def backward(self, loss: torch.Tensor) -> None:
        raise NotImplementedError

    def log_state(self) -> None:
        raise NotImplementedError

    def state_dict(self) -> Dict[str, Any]:
        raise NotImplementedError

    def save_checkpoint(self, dir: Path) -> None:
        raise NotImplementedError

    def load_checkpoint(self, dir: Path) -> None:
        raise NotImplementedError

    def refresh_optimizer_after_model_change(self) -> None:
        raise NotImplementedError


def allreduce_tensor_in_float32(tensor: torch.Tensor, group: Any) -> None:
    """In-place sum over ``group`` with an fp32 wire (used for the small tied-weight / replica reductions)."""
    if tensor.dtype == torch.float32:
        dist.all_reduce(tensor, group=group)
    else:
        t32 = tensor.float()
        dist.all_reduce(t32, group=group)
        tensor.copy_(t32)


def allreduce_no_retain(tensors: List[torch.Tensor], group: Any, bucket_elements: int = 500000000) -> None:
    """Bucketed fp32 all-reduce of a list of tensors (reference allreduce.py:16-66)."""
    bucket: List[torch.Tensor] = []
    n = 0

    def flush() -> None:
        nonlocal bucket, n
        if not bucket:
            return
        flat = torch.cat([t.reshape(-1).float() for t in bucket])
        dist.all_reduce(flat, group=group)
        off = 0
        for t in bucket:
            t.copy_(flat[off:off + t.numel()].view_as(t))
            off += t.numel()
        bucket, n = [], 0

    for t in tensors:
        bucket.append(t)
        n += t.numel()
        if n >= bucket_elements:
            flush()
    flush()


class Optimizer(BaseOptimizer):
    def __init__(self, config: OptimizerConfig, parameter_groups: List[OptimizerParamGroup], topology: Topology,
                 main_grad_dtype: Optional[torch.dtype] = torch.float32):
        super().__init__(config)
        self.config: OptimizerConfig = config
        self.topology = topology
        self.parameter_groups = parameter_groups
        self._assert_no_parameter_duplicates()
        for g in parameter_groups:
            g.initialize(topology=topology, zero=config.zero, grad_dtype=main_grad_dtype)
        self.loss_scaler = LossScaler(config.loss_scaler, parameter_groups)
        self.step_index = 0
        dev = parameter_groups[0].param_arena.device if parameter_groups else topology.device
        self._norm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self._nonfinite = torch.zeros(1, dtype=torch.int32, device=dev)

    def _assert_no_parameter_duplicates(self) -> None:
        seen = set()
        for g in self.parameter_groups:
            for _, p, _ in g._named:
                assert id(p) not in seen, "parameter occurs in more than one optimizer group"
                seen.add(id(p))

    # ------------------------------------------------------------------------------------------------
    def zero_grad(self, set_to_none: bool = True) -> None:
        for g in self.parameter_groups:
            g.zero_grad(set_to_none)

    def backward(self, loss: torch.Tensor) -> None:
        """Backward pass of one micro batch: scaled by 1/grad-acc and by the loss scale."""
        loss = loss.float()
        gas = self.topology.config.gradient_accumulation_steps
        if gas > 1:
            loss = loss / gas
        self.loss_scaler.scale_loss(loss).backward()

    def _reduce_tensor_parallel_replicas(self) -> None:
        """Gradients of TP-replicated parameters that were computed from a *part* of the data on each TP rank must be
        summed over the TP group: everything replicated under sequence parallelism (norm weights, row-parallel biases
        see only s/tp tokens; reference optimizer.py:253-260 matches the substring "norm"), per-head QK-norm weights
        always, and replicated KV projections (true MQA) over their replica set."""
        topo = self.topology
        if topo.config.model_parallel_size == 1 or not topo.is_distributed_initialized:
            return
        sp = topo.config.sequence_parallel

        def needs(p: torch.Tensor, m: CoreParameterMeta) -> bool:
            mode = getattr(m, "tp_grad_reduce", None)
            if mode == "always":
                return True
            if mode == "never":
                return False
            return sp and m.is_model_parallel_duplicate

        views = [v for g in self.parameter_groups for v in g.grad_views(needs)]
        if views:
            allreduce_no_retain(views, topo.model_parallel_group, self.config.allreduce_bucket_size)
        # replicated KV projections (fewer KV heads than TP ranks): rank r holds head r // replicas, so the gradient is
        # summed over the `replicas` consecutive ranks that share a head — gather over the TP group, add the own block
        tp, rank = topo.config.model_parallel_size, topo.model_parallel_rank
        for g in self.parameter_groups:
            for p, o, m in zip(g.parameters_original, g.offsets, g.parameter_metas):
                replicas = getattr(m, "kv_replicas", 1)
                if replicas <= 1:
                    continue
                view = g.grad_arena[o:o + p.numel()]
                gathered = torch.empty(tp * view.numel(), dtype=view.dtype, device=view.device)
                dist.all_gather_into_tensor(gathered, view.contiguous(), group=topo.model_parallel_group)
                first = (rank // replicas) * replicas
                view.copy_(gathered.view(tp, -1)[first:first + replicas].sum(0))

    def step(self) -> OptimizerStepOutput:
        self.step_index += 1
        _weights_changed()
        topo = self.topology
        c = self.config
        for g in self.parameter_groups:
            g.fold_autograd_grads()
        self._reduce_tensor_parallel_replicas()
        for g in self.parameter_groups:
            g.reduce_data_parallel()
        # global Σg² and overflow flag: local fused kernel, then tiny reductions (DP only if sharded, TP, PP)
        self._norm_sq.zero_()
        self._nonfinite.zero_()
        for g in self.parameter_groups:
            g.local_sumsq(self._norm_sq, self._nonfinite)
        if topo.is_distributed_initialized and topo.config.world_size > 1:
            packed = torch.cat([self._norm_sq, self._nonfinite.float()])
            if any(g.zero for g in self.parameter_groups):
                dist.all_reduce(packed, group=topo.data_parallel_group)
            if topo.config.model_parallel_size > 1:
                dist.all_reduce(packed, group=topo.model_parallel_group)
            if topo.config.pipe_parallel_size > 1:
                dist.all_reduce(packed, group=topo.pipe_parallel_group)
            if c.loss_scaler.enable and not any(g.zero for g in self.parameter_groups) and topo.config.data_parallel_size > 1:
                pass  # non-ZeRO: every DP rank holds identical reduced gradients, flag already agrees
            self._norm_sq.copy_(packed[:1])
            self._nonfinite.copy_((packed[1:] > 0).to(torch.int32))
        dp = topo.config.data_parallel_size
        scale = self.loss_scaler.scale
        inv_scale = 1.0 / (scale * dp)
        skip = self._nonfinite if c.loss_scaler.enable else None
        for g in self.parameter_groups:
            g.apply_adamw(self.step_index, c.beta1, c.beta2, c.eps, inv_scale, c.gradient_clipping, self._norm_sq, skip)
        for g in self.parameter_groups:
            g.zero_unowned_grads()
            g.gather_params()
        # the only host read of the step
        norm_sq, bad = self._norm_sq.item(), int(self._nonfinite.item())
        overflow = bool(bad) if c.loss_scaler.enable else None
        if bad and not c.loss_scaler.enable:
            raise RuntimeError(f"non-finite gradients at step {self.step_index} and no loss scaler is active")
        ls = self.loss_scaler.step(overflow)
        if overflow:
            logger.warning("loss scaler encountered overflow, skipping step")
            for g in self.parameter_groups:
                g.grad_arena.zero_()
            return OptimizerStepOutput(None, None, None, True, ls.no_overflow_steps, ls.current_loss_scale, None)
        global_grad_norm = math.sqrt(max(norm_sq, 0.0)) * inv_scale
        learning_rates = {}
        for i, g in enumerate(self.parameter_groups):
            learning_rates[g.config.name or f"parameter_group_{i}"] = g.get_learning_rate()
        debug = None
        if c.debug_log:
            debug = {}
            for g in self.parameter_groups:
                for p, m in zip(g.parameters_original, g.parameter_metas):
                    name = f"{m.parameter_name}-layer-{m.layer_index}"
                    debug[f"debug/{name}-norm"] = float(p.float().norm().item())
                    debug[f"debug/{name}-pct-zero"] = 1.0 - float((p.count_nonzero() / max(p.numel(), 1)).item())
        # norm after clipping (the reference declares the field and always returns None)
        clipped = min(global_grad_norm, c.gradient_clipping) if c.gradient_clipping and c.gradient_clipping > 0 \
            else global_grad_norm
        return OptimizerStepOutput(global_grad_norm, clipped, learning_rates, overflow, ls.no_overflow_steps,
                                   ls.current_loss_scale, debug)

    def refresh_optimizer_after_model_change(self) -> None:
        _weights_changed()
        for g in self.parameter_groups:
            g.refresh_optimized_params(self.topology)

    def log_state(self) -> None:
        for g in self.parameter_groups:
            g.log_state()

    # ---- checkpoint -------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        """Same keys as the reference (optimizer.py:297-307).  ``optimizer`` has the shape of a torch optimizer's state dict
        over the flat fp32 partitions — one entry per parameter group, as in the reference where every group is one
        flattened fp32 tensor: the moments are this rank's arena shards themselves (views, not copies)."""
        groups = self._torch_style_groups()
        state = {}
        for i, (g, tg) in enumerate(zip(self.parameter_groups, groups)):
            tg["params"] = [i]
            state[i] = {"step": torch.tensor(float(self.step_index)), "exp_avg": g.exp_avg, "exp_avg_sq": g.exp_avg_sq}
        return {"step_index": self.step_index, "parameter_groups": [g.state_dict() for g in self.parameter_groups],
                "optimizer": {"state": state, "param_groups": groups}, "loss_scaler": self.loss_scaler.state_dict()}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.step_index = sd["step_index"]
        self.loss_scaler.load_state_dict(sd["loss_scaler"])
        for g, s in zip(self.parameter_groups, sd["parameter_groups"]):
            g.load_state_dict(s)

    def _torch_style_groups(self) -> List[Dict[str, Any]]:
        c = self.config
        return [{"lr": g.get_learning_rate(), "betas": (c.beta1, c.beta2), "eps": c.eps,
                 "weight_decay": g.config.weight_decay, "amsgrad": False, "maximize": False, "foreach": None,
                 "capturable": False, "differentiable": False, "fused": None, "params": []}
                for g in self.parameter_groups]

    def save_checkpoint(self, directory: Union[Path, str]) -> None:
        directory = Path(directory)
        topo = self.topology
        if self.config.zero_save_static:
            name = (f"optimizer_state_static_mp_{topo.model_parallel_rank}_pp_{topo.pipe_parallel_rank}_"
                    f"dp_{topo.data_parallel_rank}.pt")
            torch.save(self.state_dict(), str(directory / name))
            return
        by_layer: Dict[int, Dict[str, Any]] = {}
        for g in self.parameter_groups:
            full = {k: g.gather_full_state(getattr(g, k)) for k in ("master", "exp_avg", "exp_avg_sq")}
            for p, m, o in zip(g.parameters_original, g.parameter_metas, g.offsets):
                if m.layer_index is None or m.layer_index < 0:
                    continue
                n = p.numel()
                merged = {k: merge_parameter(v[o:o + n].view(p.shape), m, topo) for k, v in full.items()}
                entry = {"parameter": merged["master"], "meta": m.state_dict(),
                         "optimizer_state": {"step": torch.tensor(float(self.step_index)), "exp_avg": merged["exp_avg"],
                                             "exp_avg_sq": merged["exp_avg_sq"]}}
                layers = sorted(m.tied_layer_indices) if m.is_tied and m.tied_layer_indices else [m.layer_index]
                for li in layers:
                    by_layer.setdefault(li, {})[m.key_for_layer(li)] = entry
            del full
        if topo.model_parallel_rank == 0 and topo.data_parallel_rank == 0:
            for li, params in by_layer.items():
                torch.save({"step_index": self.step_index, "loss_scaler": self.loss_scaler.state_dict(),
                            "optimizer_param_groups": self._torch_style_groups(), "parameters": params},
                           str(directory / f"optimizer_state_layer_{li}.pt"))

    def load_checkpoint(self, directory: Union[Path, str]) -> None:
        directory = Path(directory)
        topo = self.topology
        if self.config.zero_save_static:
            name = (f"optimizer_state_static_mp_{topo.model_parallel_rank}_pp_{topo.pipe_parallel_rank}_"
                    f"dp_{topo.data_parallel_rank}.pt")
            self.load_state_dict(torch.load(str(directory / name), map_location="cpu", weights_only=False))
            return
        cache: Dict[int, Dict[str, Any]] = {}

        def layer_file(li: int) -> Optional[Dict[str, Any]]:
            if li not in cache:
                f = directory / f"optimizer_state_layer_{li}.pt"
                cache[li] = torch.load(str(f), map_location="cpu", weights_only=False) if f.is_file() else None
            return cache[li]

        loaded_any = False
        # Start from the weights the model checkpoint just put into the arenas: a parameter without an entry in the
        # optimizer files (frozen in the run that wrote them, files pruned, groups changed for finetuning) must keep
        # its LOADED weight, not the construction-time master, and begins with zero moments.
        for g in self.parameter_groups:
            g.refresh_optimized_params(topo)
            g.exp_avg.zero_()
            g.exp_avg_sq.zero_()
        for g in self.parameter_groups:
            for p, m, o in zip(g.parameters_original, g.parameter_metas, g.offsets):
                if m.layer_index is None or m.layer_index < 0:
                    continue
                entry = None
                layers = sorted(m.tied_layer_indices) if m.is_tied and m.tied_layer_indices else [m.layer_index]
                for li in layers:
                    sd = layer_file(li)
                    if sd is not None and m.key_for_layer(li) in sd["parameters"]:
                        entry = sd["parameters"][m.key_for_layer(li)]
                        self.step_index = sd["step_index"]
                        self.loss_scaler.load_state_dict(sd["loss_scaler"])
                        break
                if entry is None:
                    logger.warning(f"no optimizer state for {m.key}; keeping fresh state")
                    continue
                loaded_any = True
                dev = g.master.device
                g.scatter_full_state(g.master, o, split_parameter(entry["parameter"], m, topo).reshape(-1).to(dev))
                g.scatter_full_state(g.exp_avg, o, split_parameter(entry["optimizer_state"]["exp_avg"], m, topo).reshape(-1).to(dev))
                g.scatter_full_state(g.exp_avg_sq, o, split_parameter(entry["optimizer_state"]["exp_avg_sq"], m, topo).reshape(-1).to(dev))
        _weights_changed()
        if not loaded_any:
            logger.warning(f"no optimizer checkpoint files found in {directory}")
