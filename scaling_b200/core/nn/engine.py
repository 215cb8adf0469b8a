"""The 3-D parallel engine: interprets a pipeline schedule over this stage's layers.

Reference: core/nn/parallel_module/parallel_module.py:1-747 (train_step :331-414, evaluation_step :416-474,
run_instructions :476-532, handlers :534-732, loss return path :276-329), buffers.py.

Same instruction set and public surface; differences by design:
  * stage-to-stage sends are asynchronous (parked ``isend`` handles), receives block only where the data is needed;
  * no host synchronisation inside the micro-batch loop — loss and metrics stay on the device until the end of the
    step, where loss, grad-norm and overflow flag are read back together;
  * profiling records CUDA events instead of synchronising around every instruction;
  * runs on CUDA/NCCL or CPU/gloo unchanged.
"""
from __future__ import annotations

import collections
import time
from enum import Enum
from typing import TYPE_CHECKING, Any, Callable, Dict, Generic, List, NamedTuple, Optional, Tuple, Union

import torch
import torch.distributed as dist
from torch.utils._pytree import tree_map

from ..data.base import BaseDatasetBatch
from ..logging import logger
from ..profiler import Profiler, ProfilerConfig, SynchronizedTimer
from ..topology import ActivationCheckpointingType, Topology
from .act_ckpt import checkpoint
from .layers import BaseDatasetBatchGeneric, BaseLossInputGeneric, LayerSpec
from .p2p import PipeCommunicator
from .parameter_meta import CoreParameterMeta
from .partitioned_module import PipePartitionedModule
from .schedule import (
    InstructionBackwardPass,
    InstructionBase,
    InstructionForwardPass,
    InstructionLoadMicroBatch,
    InstructionLoss,
    InstructionOptimizerStep,
    InstructionRecvActivation,
    InstructionRecvGrad,
    InstructionReduceTiedGrads,
    InstructionSendActivation,
    InstructionSendGrad,
    InstructionStoreMicroBatch,
    PipelineScheduleInference,
    PipelineScheduleTrain,
)


if TYPE_CHECKING:  # only for annotations: core.optimizer imports core.nn
    from ..optimizer import BaseOptimizer


class BufferType(Enum):
    PIPELINE_STAGE_INPUT = "pipeline_stage_input"
    PIPELINE_STAGE_OUTPUT = "pipeline_stage_output"
    TARGET = "pipeline_stage_target"
    LOSS = "loss"
    METRICS = "metrics"
    GRAD = "grad"


class Buffers:
    """``{buffer type: {buffer id: object}}`` plus the running loss sum of the step."""

    def __init__(self) -> None:
        self.reset()

    def reset(self) -> None:
        self.data: Dict[BufferType, Dict[int, Any]] = collections.defaultdict(dict)
        self.accum_loss: Optional[torch.Tensor] = None

    def write(self, buffer_type: BufferType, buffer_id: int, data: Any) -> None:
        self.data[buffer_type][buffer_id] = data

    def get(self, buffer_type: BufferType, buffer_id: int) -> Any:
        return self.data[buffer_type][buffer_id]

    def take(self, buffer_type: BufferType, buffer_id: int) -> Any:
        d = self.data[buffer_type][buffer_id]
        self.data[buffer_type][buffer_id] = None
        return d

    def dump(self, buffer_type: BufferType) -> Dict[int, Any]:
        return self.data[buffer_type]

    def add_loss(self, loss: torch.Tensor) -> None:
        l = loss.detach().float()
        self.accum_loss = l.clone() if self.accum_loss is None else self.accum_loss + l


class TrainStepOutput(NamedTuple):
    loss: Optional[float]
    metrics: Optional[Dict[str, Union[int, float]]]
    global_grad_norm: Optional[float]
    global_grad_norm_clipped: Optional[float]
    learning_rates: Optional[Dict[str, float]]
    overflow: Optional[bool]
    no_overflow_steps: Optional[int]
    current_loss_scale: Optional[float]
    step_duration: float
    debug_dict: Optional[Dict[str, float]]


class EvaluationStepOutput(NamedTuple):
    loss: Optional[float]
    metrics: Optional[Dict[str, Union[int, float]]]
    step_duration: float


def _to_python(x: Any) -> Any:
    return x.tolist() if torch.is_tensor(x) else x


def _timer(instruction: InstructionBase) -> Tuple[str, int, int]:
    name = instruction.name.replace("Instruction", "")
    mb = instruction.micro_batch_id if instruction.micro_batch_id is not None else -1
    buf = instruction.buffer_id if instruction.buffer_id is not None else -1
    return name, mb, buf


class ParallelModule(PipePartitionedModule, Generic[BaseLossInputGeneric, BaseDatasetBatchGeneric]):
    def __init__(self, layer_specs: List[LayerSpec], topology: Topology, profiler_config: Optional[ProfilerConfig] = None,
                 use_continuous_recommunication: bool = False):
        super().__init__(layer_specs=layer_specs, topology=topology)
        assert self.topology is not None
        topo = self.topology
        self.train_instructions = PipelineScheduleTrain(topo).instructions()
        self.eval_instructions = PipelineScheduleInference(topo).instructions()
        self.pipe_buffer = Buffers()
        pp = topo.config.pipe_parallel_size
        dev = topo.device
        cont = use_continuous_recommunication
        # forward traffic comes from the previous stage, gradients from the next
        fwd_group, bwd_group = topo.pipe_p2p_groups
        groups = dict(group=fwd_group, grad_group=bwd_group)
        self.communicator_in = PipeCommunicator(dev, recv_grads=False, recv_data=True, use_continuous_recommunication=cont, **groups)
        self.communicator_out = PipeCommunicator(dev, recv_grads=True, recv_data=False, use_continuous_recommunication=cont, **groups)
        self.communicator_loss_in = self.communicator_loss_out = None
        if pp > 1:
            # loss + metrics keep one structure per mode (train / eval): the handshake is cached like the activations'
            # (the reference re-pickles it every step, communicator.py:256-281)
            if topo.is_first_pipe_parallel_rank:
                self.communicator_loss_in = PipeCommunicator(dev, False, True, use_continuous_recommunication=cont, group=bwd_group)
            if topo.is_last_pipe_parallel_rank:
                self.communicator_loss_out = PipeCommunicator(dev, False, False, use_continuous_recommunication=cont, group=bwd_group)
        self.profiler = Profiler(profiler_config or ProfilerConfig(), topo)
        self.step_timer = SynchronizedTimer()
        self._mark_tied_main_stage()
        self.broadcast_model()

    # ---- parameters -------------------------------------------------------------------------------------
    def _mark_tied_main_stage(self) -> None:
        tli = self.tied_layer_index
        if tli is None:
            return
        from .partitioned_module import _resolve

        for key, module in tli.local_main_modules.items():
            main_stage = tli.stages_by_key[key][0]
            for attr in sorted(tli.attrs_by_key[key]):
                mod, name = _resolve(module, attr)
                getattr(mod, name).core_parameter_meta.tied_main_stage = main_stage  # type: ignore[attr-defined]

    def named_parameters_with_meta(self) -> List[Tuple[str, torch.nn.Parameter, CoreParameterMeta]]:
        """Every distinct parameter of this stage with its metadata (shared/tied parameters once)."""
        out, seen = [], set()
        for _, layer in self.local_layers():
            for name, p in layer.named_parameters():
                if id(p) in seen:
                    continue
                seen.add(id(p))
                out.append((name, p, p.core_parameter_meta))
        return out

    def broadcast_model(self) -> None:
        """Make replicas identical at start-up: TP-replicated tensors from mp-rank 0, everything from dp-rank 0,
        tied copies from their first stage."""
        topo = self.topology
        if not topo.is_distributed_initialized:
            return
        tensors = []
        for _, layer in self.local_layers():
            tensors += [t for _, t in list(layer.named_parameters()) + list(layer.named_buffers())]
        if topo.config.model_parallel_size > 1:
            for t in tensors:
                if t.core_parameter_meta.is_model_parallel_duplicate:
                    dist.broadcast(t.data, src=topo.model_parallel_ranks[0], group=topo.model_parallel_group)
        if topo.config.data_parallel_size > 1:
            for t in tensors:
                dist.broadcast(t.data, src=topo.data_parallel_ranks[0], group=topo.data_parallel_group)
        if self.tied_layer_index is not None:
            for p, group, stages in self.tied_layer_index.local_parameters_and_process_groups():
                if group is not None and len(stages) > 1:
                    src = topo.get_global_rank(pipe_parallel_rank=stages[0])
                    dist.broadcast(p.data, src=src, group=group)

    def get_params_count(self) -> Tuple[int, int]:
        """(total incl. TP/tied duplicates, unique) parameter count over the whole model."""
        topo = self.topology
        total = unique = 0
        for layer_index, layer in self.local_layers():
            dup_names = self.tied_layer_index.layer_index_to_tied_local_duplicate_parameter_names(layer_index) \
                if self.tied_layer_index is not None else set()
            global_dup = self.tied_layer_index.layer_index_is_tied_global_duplicate(layer_index) \
                if self.tied_layer_index is not None else False
            tied = self.tied_layer_index.tied_attributes(layer_index) if self.tied_layer_index is not None else set()
            for name, p in layer.named_parameters(remove_duplicate=False):
                if name in dup_names:
                    continue
                n = p.numel()
                total += n
                m = p.core_parameter_meta
                if m.is_model_parallel_duplicate and topo.model_parallel_rank != 0:
                    continue
                if getattr(m, "kv_replicas", 1) > 1 and topo.model_parallel_rank % m.kv_replicas != 0:
                    continue
                if global_dup and name in tied:
                    continue
                unique += n
        if topo.data_parallel_rank != 0:
            total = unique = 0
        if topo.is_distributed_initialized:
            t = torch.tensor([total, unique], dtype=torch.long, device=topo.device)
            dist.all_reduce(t)
            total, unique = int(t[0].item()), int(t[1].item())
        return total, unique

    # ---- forward ------------------------------------------------------------------------------------------
    def forward(self, x: Any) -> Any:
        mode = self.topology.config.activation_checkpointing_type
        training = self.training and torch.is_grad_enabled()
        if not training or mode == ActivationCheckpointingType.DISABLED:
            for layer in self._layers:
                x = layer(x)
            return x
        if mode == ActivationCheckpointingType.EVERY_LAYER:
            for layer in self._layers:
                tup = layer.input_to_tuple(x)
                x = checkpoint(layer._forward_tuple_input, self.topology, *tup)
            return x
        # every pipe stage: one checkpoint around the whole stage
        first = self._layers[0]

        def run_stage(*tup: Any) -> Any:
            y = first.tuple_to_input(tuple(tup))
            for layer in self._layers:
                y = layer(y)
            return y

        return checkpoint(run_stage, self.topology, *first.input_to_tuple(x))

    # ---- steps --------------------------------------------------------------------------------------------
    def train_step(self, dataloader: Optional[Any], optimizer: "BaseOptimizer",
                   sync_batch_to_model_parallel: Callable[[Topology, Optional[Any]], Any],
                   loss_function: Callable[..., Any], metrics_aggregation_fn: Optional[Callable] = None) -> TrainStepOutput:
        self.step_timer.start()
        self._layers.train()
        self._set_comm_mode("train")
        self.pipe_buffer.reset()
        self.profiler.step()
        step_out = None
        for ins in self.train_instructions:
            with self.profiler.time(*_timer(ins)):
                step_out = self._execute(ins, dataloader, optimizer, sync_batch_to_model_parallel, loss_function) or step_out
        self._drain()
        loss, metrics = self.get_loss(metrics_aggregation_fn)
        self.profiler.flush()
        duration = self.step_timer.stop()
        assert step_out is not None
        return TrainStepOutput(loss, metrics, step_out.global_grad_norm, step_out.global_grad_norm_clipped,
                               step_out.learning_rates, step_out.overflow, step_out.no_overflow_steps,
                               step_out.current_loss_scale, duration, step_out.debug_dict)

    def evaluation_step(self, dataloader: Optional[Any], sync_batch_to_model_parallel: Callable, loss_function: Callable,
                        metrics_aggregation_fn: Optional[Callable] = None) -> EvaluationStepOutput:
        self.step_timer.start()
        self._layers.eval()
        self._set_comm_mode("eval")
        self.pipe_buffer.reset()
        with torch.no_grad():
            for ins in self.eval_instructions:
                self._execute(ins, dataloader, None, sync_batch_to_model_parallel, loss_function)
        self._drain()
        loss, metrics = self.get_loss(metrics_aggregation_fn)
        return EvaluationStepOutput(loss, metrics, self.step_timer.stop())

    def run_instructions(self, instructions: List[InstructionBase], dataloader: Optional[Any] = None,
                         optimizer: Optional["BaseOptimizer"] = None, sync_batch_to_model_parallel: Optional[Callable] = None,
                         loss_function: Optional[Callable] = None) -> Buffers:
        """Interpret an arbitrary instruction list (inference-style use); returns the buffers."""
        self.pipe_buffer.reset()
        for ins in instructions:
            self._execute(ins, dataloader, optimizer, sync_batch_to_model_parallel, loss_function)
        self._drain()
        return self.pipe_buffer

    def reset_activation_shape(self) -> None:
        """Forget cached p2p metadata (next message re-negotiates shapes)."""
        for c in (self.communicator_in, self.communicator_out):
            c.reset_communication_meta()

    def _set_comm_mode(self, mode: str) -> None:
        for c in (self.communicator_in, self.communicator_out, self.communicator_loss_in, self.communicator_loss_out):
            if c is not None:
                c.set_mode(mode)

    def _drain(self) -> None:
        for c in (self.communicator_in, self.communicator_out, self.communicator_loss_in, self.communicator_loss_out):
            if c is not None:
                c.drain()

    # ---- instruction handlers -------------------------------------------------------------------------------
    def _execute(self, ins: InstructionBase, dataloader, optimizer, sync_fn, loss_fn):
        topo = self.topology
        buf = self.pipe_buffer
        b = ins.buffer_id
        if isinstance(ins, InstructionLoadMicroBatch):
            batch = next(dataloader) if (dataloader is not None and topo.is_io_rank) else None
            batch = sync_fn(topo, batch)
            if topo.is_first_pipe_parallel_rank:
                buf.write(BufferType.PIPELINE_STAGE_INPUT, b, batch.only_inputs() if topo.config.pipe_parallel_size > 1 else batch)
            if topo.is_last_pipe_parallel_rank:
                buf.write(BufferType.TARGET, b, batch.only_targets() if topo.config.pipe_parallel_size > 1 else batch)
        elif isinstance(ins, InstructionRecvActivation):
            tup = self.communicator_in.recv_data(origin_global_rank=topo.previous_pipe_parallel_rank)
            buf.write(BufferType.PIPELINE_STAGE_INPUT, b, (tup, self._layers[0].tuple_to_input(tup)))
        elif isinstance(ins, InstructionForwardPass):
            x = buf.get(BufferType.PIPELINE_STAGE_INPUT, b)
            if isinstance(x, tuple) and len(x) == 2 and isinstance(x[0], tuple):
                x = x[1]
            buf.write(BufferType.PIPELINE_STAGE_OUTPUT, b, self(x))
        elif isinstance(ins, InstructionLoss):
            out = buf.get(BufferType.PIPELINE_STAGE_OUTPUT, b)
            target = buf.take(BufferType.TARGET, b)
            res = loss_fn(out, target)
            if isinstance(res, tuple):
                loss, metrics = res
                buf.write(BufferType.METRICS, ins.micro_batch_id if ins.micro_batch_id is not None else b, metrics)
            else:
                loss = res
            assert torch.is_tensor(loss) and loss.dim() == 0, f"The loss needs to be a scalar, got shape {tuple(loss.shape)}"
            buf.add_loss(loss)
            buf.write(BufferType.LOSS, b, loss)
        elif isinstance(ins, InstructionSendActivation):
            out = buf.get(BufferType.PIPELINE_STAGE_OUTPUT, b)
            tup = self._layers[-1].output_to_tuple(out)
            buf.write(BufferType.PIPELINE_STAGE_OUTPUT, b, tup)  # what the gradients will refer to
            self.communicator_out.send_data(tup, target_global_rank=topo.next_pipe_parallel_rank)
        elif isinstance(ins, InstructionRecvGrad):
            tup = buf.get(BufferType.PIPELINE_STAGE_OUTPUT, b)
            buf.write(BufferType.GRAD, b, self.communicator_out.recv_gradients(tup, origin_global_rank=topo.next_pipe_parallel_rank))
        elif isinstance(ins, InstructionBackwardPass):
            if topo.is_last_pipe_parallel_rank:
                optimizer.backward(buf.take(BufferType.LOSS, b))
            else:
                g = buf.take(BufferType.GRAD, b)
                torch.autograd.backward(tensors=g.tensors, grad_tensors=g.grad_tensors)
            buf.write(BufferType.PIPELINE_STAGE_OUTPUT, b, None)
        elif isinstance(ins, InstructionSendGrad):
            x = buf.take(BufferType.PIPELINE_STAGE_INPUT, b)
            self.communicator_in.send_gradients(x[0], target_global_rank=topo.previous_pipe_parallel_rank)
        elif isinstance(ins, InstructionReduceTiedGrads):
            self._reduce_tied_grads()
        elif isinstance(ins, InstructionOptimizerStep):
            return optimizer.step()
        elif isinstance(ins, InstructionStoreMicroBatch):
            pass
        else:
            raise NotImplementedError(f"unknown instruction {ins}")
        return None

    def _reduce_tied_grads(self) -> None:
        """Sum the gradients of tied parameters over the stages that hold a copy (e.g. tied embedding / LM head)."""
        if self.tied_layer_index is None or not self.topology.is_distributed_initialized:
            return
        for p, group, stages in self.tied_layer_index.local_parameters_and_process_groups():
            if group is None or len(stages) < 2:
                continue
            g = getattr(p, "main_grad", None)
            if g is None:
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
                g = p.grad
            elif p.grad is not None:
                g.add_(p.grad.to(g.dtype))
                p.grad = None
            if g.dtype == torch.float32:
                dist.all_reduce(g, group=group)
            else:
                g32 = g.float()
                dist.all_reduce(g32, group=group)
                g.copy_(g32)

    # ---- loss return path ---------------------------------------------------------------------------------------
    def get_loss(self, metrics_aggregation_fn: Optional[Callable]) -> Tuple[Optional[float], Optional[Dict[str, Any]]]:
        topo = self.topology
        data = None
        if topo.is_last_pipe_parallel_rank:
            assert self.pipe_buffer.accum_loss is not None
            n_micro = topo.config.gradient_accumulation_steps
            loss = self.pipe_buffer.accum_loss / n_micro
            metrics_list = [m for m in self.pipe_buffer.dump(BufferType.METRICS).values() if m is not None]
            metrics = None
            if metrics_list:
                assert metrics_aggregation_fn is not None, "loss function returned metrics but no aggregation fn is set"
                metrics = metrics_aggregation_fn(topo, metrics_list)
            if topo.config.data_parallel_size > 1 and topo.is_distributed_initialized:
                dist.all_reduce(loss, group=topo.data_parallel_group)
                loss = loss / topo.config.data_parallel_size
            data = (loss, metrics)
        if topo.config.pipe_parallel_size > 1:
            if topo.is_last_pipe_parallel_rank:
                self.communicator_loss_out.send_data(data, target_global_rank=topo.get_global_rank(pipe_parallel_rank=0))
                self.communicator_loss_out.drain()
                if not topo.is_first_pipe_parallel_rank:
                    return None, None
            if topo.is_first_pipe_parallel_rank:
                last = topo.config.pipe_parallel_size - 1
                data = self.communicator_loss_in.recv_data(origin_global_rank=topo.get_global_rank(pipe_parallel_rank=last))
            else:
                return None, None
        loss, metrics = data
        return float(loss.item()), tree_map(_to_python, metrics)
