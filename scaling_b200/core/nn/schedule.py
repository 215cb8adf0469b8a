"""Pipeline schedules as per-stage instruction lists, plus a dependency simulator and PNG visualiser.

Reference: core/nn/pipeline_schedule/instructions.py:1-61, train.py:1-174 (1F1B), inference.py:1-75, base.py:1-697.

The 1F1B timetable is written in closed form: with S stages and M micro batches, stage ``s`` runs the forward of micro
batch ``m`` at tick ``2m + s`` and its backward at tick ``2m + 2S − 1 − s``; a tick is a forward tick for a stage iff
tick and stage have the same parity.  Communication at a forward tick is (recv activation, send previous grad) with the
previous stage, at a backward tick (send previous activation, recv grad) with the next stage — neighbours are always on
opposite ticks, so every blocking pair is posted in matching order and the schedule cannot deadlock.
"""
from __future__ import annotations

import json
from collections import defaultdict
from pathlib import Path
from typing import Any, Dict, List, NamedTuple, Optional, Tuple, Union

from ..topology import Topology, TopologyConfig


class InstructionBase(NamedTuple):
    buffer_id: Optional[int] = None
    micro_batch_id: Optional[int] = None
    is_first_pass: Optional[bool] = None

    @property
    def name(self) -> str:
        return self.__class__.__name__


def _instr(name: str):
    return type(name, (InstructionBase,), {"__slots__": ()})


class InstructionOptimizerStep(InstructionBase):
    __slots__ = ()


class InstructionReduceTiedGrads(InstructionBase):
    __slots__ = ()


class InstructionStoreMicroBatch(InstructionBase):
    __slots__ = ()


class InstructionLoadMicroBatch(InstructionBase):
    __slots__ = ()


class InstructionForwardPass(InstructionBase):
    __slots__ = ()


class InstructionLoss(InstructionBase):
    __slots__ = ()


class InstructionBackwardPass(InstructionBase):
    __slots__ = ()


class InstructionSendActivation(InstructionBase):
    __slots__ = ()


class InstructionRecvActivation(InstructionBase):
    __slots__ = ()


class InstructionSendGrad(InstructionBase):
    __slots__ = ()


class InstructionRecvGrad(InstructionBase):
    __slots__ = ()


_COMM = (InstructionSendActivation, InstructionRecvActivation, InstructionSendGrad, InstructionRecvGrad)
_PARTNER = {
    InstructionSendActivation: (InstructionRecvActivation, +1),
    InstructionRecvActivation: (InstructionSendActivation, -1),
    InstructionSendGrad: (InstructionRecvGrad, -1),
    InstructionRecvGrad: (InstructionSendGrad, +1),
}
_COLORS = {
    "ForwardPass": (86, 156, 214), "BackwardPass": (206, 145, 120), "Loss": (220, 220, 170),
    "LoadMicroBatch": (181, 206, 168), "SendActivation": (197, 134, 192), "RecvActivation": (156, 110, 190),
    "SendGrad": (215, 186, 125), "RecvGrad": (190, 160, 100), "ReduceTiedGrads": (120, 200, 200),
    "OptimizerStep": (244, 71, 71), "StoreMicroBatch": (150, 150, 150),
}


def _timer_name(instruction: InstructionBase) -> str:
    return instruction.name.replace("Instruction", "")


def _stage_topology(pipe_parallel_size: int, gradient_accumulation_steps: int, rank: int) -> Topology:
    return Topology(TopologyConfig(
        global_rank=rank, world_size=pipe_parallel_size, local_slot=0, model_parallel_size=1,
        pipe_parallel_size=pipe_parallel_size, data_parallel_size=1, micro_batch_size=1,
        gradient_accumulation_steps=gradient_accumulation_steps))


class PipelineScheduleBase:
    def __init__(self, topology: Topology):
        self.topology = topology
        self.stage = topology.pipe_parallel_rank
        self.stages = topology.config.pipe_parallel_size
        self.micro_batches = topology.config.gradient_accumulation_steps

    def instructions(self) -> List[InstructionBase]:
        raise NotImplementedError

    def required_buffer_count(self) -> int:
        raise NotImplementedError

    def _valid_micro_batch(self, micro_batch_id: int) -> bool:
        return 0 <= micro_batch_id < self.micro_batches

    def _is_valid_pipe_parallel_rank(self, pipe_parallel_rank: Optional[int]) -> bool:
        return pipe_parallel_rank is not None

    # ---- analysis ---------------------------------------------------------------------------------
    @classmethod
    def _all_instructions(cls, gradient_accumulation_steps: int, pipe_parallel_size: int) -> List[List[InstructionBase]]:
        return [cls(_stage_topology(pipe_parallel_size, gradient_accumulation_steps, r)).instructions()
                for r in range(pipe_parallel_size)]

    @classmethod
    def illustrate(cls, gradient_accumulation_steps: int, pipe_parallel_size: int) -> Dict[str, Any]:
        """Lock-step replay with unit-time instructions: every round, each stage executes its next instruction if its
        communication partner is ready.  Returns the per-stage timeline (None = idle slot) and the idle fraction;
        raises if no stage can advance (deadlock)."""
        prog = cls._all_instructions(gradient_accumulation_steps, pipe_parallel_size)
        pc = [0] * pipe_parallel_size
        timeline: List[List[Optional[InstructionBase]]] = [[] for _ in range(pipe_parallel_size)]
        while any(pc[r] < len(prog[r]) for r in range(pipe_parallel_size)):
            ready = []
            for r in range(pipe_parallel_size):
                if pc[r] >= len(prog[r]):
                    ready.append(False)
                    continue
                ins = prog[r][pc[r]]
                if isinstance(ins, _COMM):
                    partner_cls, delta = _PARTNER[type(ins)]
                    p = r + delta
                    ok = (0 <= p < pipe_parallel_size and pc[p] < len(prog[p]) and isinstance(prog[p][pc[p]], partner_cls)
                          and prog[p][pc[p]].micro_batch_id == ins.micro_batch_id)
                    ready.append(ok)
                else:
                    ready.append(True)
            if not any(ready):
                raise RuntimeError(f"pipeline schedule deadlocks at program counters {pc}")
            for r in range(pipe_parallel_size):
                if ready[r]:
                    timeline[r].append(prog[r][pc[r]])
                    pc[r] += 1
                else:
                    timeline[r].append(None)
        total = sum(len(t) for t in timeline)
        idle = sum(1 for t in timeline for x in t if x is None)
        return {"timeline": timeline, "steps": len(timeline[0]) if timeline else 0,
                "idle_fraction": idle / max(total, 1), "idle_percent": 100.0 * idle / max(total, 1)}

    @classmethod
    def visualize(cls, gradient_accumulation_steps: int, pipe_parallel_size: int):
        """PNG grid: one row per stage, one column per lock-step slot, coloured by instruction type."""
        from PIL import Image, ImageDraw

        info = cls.illustrate(gradient_accumulation_steps, pipe_parallel_size)
        timeline = info["timeline"]
        cell_w, cell_h, pad = 28, 22, 4
        steps = info["steps"]
        img = Image.new("RGB", (pad * 2 + max(steps, 1) * cell_w, pad * 2 + pipe_parallel_size * cell_h + 14), (30, 30, 30))
        draw = ImageDraw.Draw(img)
        for r, row in enumerate(timeline):
            for t, ins in enumerate(row):
                x0, y0 = pad + t * cell_w, pad + r * cell_h
                color = (55, 55, 55) if ins is None else _COLORS.get(_timer_name(ins), (200, 200, 200))
                draw.rectangle([x0, y0, x0 + cell_w - 2, y0 + cell_h - 2], fill=color)
                if ins is not None and ins.micro_batch_id is not None:
                    draw.text((x0 + 3, y0 + 4), str(ins.micro_batch_id), fill=(0, 0, 0))
        draw.text((pad, pad + pipe_parallel_size * cell_h), f"idle {info['idle_percent']:.1f}%", fill=(220, 220, 220))
        return img

    @classmethod
    def visualize_profile(cls, profile_file: Union[str, Path], milliseconds_per_pixel: float = 10,
                          pipe_pixels: int = 200):
        engine = SimulationEngine.from_profile(cls, profile_file)
        engine.simulate()
        return engine.summarize(), engine.visualize(milliseconds_per_pixel=milliseconds_per_pixel, pipe_pixels=pipe_pixels)


class PipelineScheduleTrain(PipelineScheduleBase):
    """1F1B with gradient accumulation."""

    def required_buffer_count(self) -> int:
        return max(2, min(self.stages - self.stage + 1, self.micro_batches))

    def _buffer(self, micro_batch_id: int) -> int:
        return micro_batch_id % self.required_buffer_count()

    def _tick(self, t: int) -> Tuple[int, bool]:
        """(micro batch, is_forward) of this stage at tick t; the micro batch may be out of range (bubble)."""
        if (t - self.stage) % 2 == 0:
            return (t - self.stage) // 2, True
        return (t - (2 * self.stages - 1 - self.stage)) // 2, False

    def instructions(self) -> List[InstructionBase]:
        out: List[InstructionBase] = []
        has_prev = self.stage > 0
        has_next = self.stage < self.stages - 1
        is_io = self.stage == 0 or self.stage == self.stages - 1
        total = 2 * (self.micro_batches + self.stages - 1)
        prev_mb = -1
        for t in range(total):
            mb, fwd = self._tick(t)
            valid, prev_valid = self._valid_micro_batch(mb), self._valid_micro_batch(prev_mb)
            if fwd:
                if valid and has_prev:
                    out.append(InstructionRecvActivation(self._buffer(mb), mb))
                if prev_valid and has_prev:
                    out.append(InstructionSendGrad(self._buffer(prev_mb), prev_mb))
            else:
                if prev_valid and has_next:
                    out.append(InstructionSendActivation(self._buffer(prev_mb), prev_mb))
                if valid and has_next:
                    out.append(InstructionRecvGrad(self._buffer(mb), mb))
            if is_io and fwd and valid:
                out.append(InstructionLoadMicroBatch(self._buffer(mb), mb))
            if valid:
                if fwd:
                    out.append(InstructionForwardPass(self._buffer(mb), mb))
                    if not has_next:
                        out.append(InstructionLoss(self._buffer(mb), mb, True))
                else:
                    out.append(InstructionBackwardPass(self._buffer(mb), mb))
            prev_mb = mb
        out.append(InstructionReduceTiedGrads())
        out.append(InstructionOptimizerStep())
        return out


class PipelineScheduleInference(PipelineScheduleBase):
    """Forward-only wavefront over ``micro_batches`` micro batches with two alternating buffers."""

    def required_buffer_count(self) -> int:
        return 2

    def instructions(self) -> List[InstructionBase]:
        out: List[InstructionBase] = []
        has_prev = self.stage > 0
        has_next = self.stage < self.stages - 1
        is_io = self.stage == 0 or self.stage == self.stages - 1
        for mb in range(self.micro_batches):
            buf = mb % 2
            if is_io:
                out.append(InstructionLoadMicroBatch(buf, mb))
            if has_prev:
                out.append(InstructionRecvActivation(buf, mb))
            out.append(InstructionForwardPass(buf, mb))
            if not has_next:
                out.append(InstructionLoss(buf, mb, True))
            if has_next:
                out.append(InstructionSendActivation(buf, mb))
        return out


# --------------------------------------------------------------------------------------------------------------
# replay of measured instruction durations (profile.json written by the profiler)
# --------------------------------------------------------------------------------------------------------------
class SimulationEngine:
    """Replays a schedule with measured per-instruction durations.

    Send/recv durations measured on a real run contain the time spent waiting for the partner; the replay replaces
    both sides of a transfer by ``min(send, recv)`` and lets the dependency graph recreate the waiting, which separates
    transfer time from pipeline bubbles.  ``summarize`` reports per-stage busy/idle time and the idle percentage.
    """

    def __init__(self, schedule_cls: type, pipe_parallel_size: int, gradient_accumulation_steps: int,
                 durations: Dict[Tuple[int, str, int], float]):
        self.schedule_cls = schedule_cls
        self.pipe_parallel_size = pipe_parallel_size
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.durations = durations  # (stage, timer name, micro batch) -> seconds
        self.programs = schedule_cls._all_instructions(gradient_accumulation_steps, pipe_parallel_size)
        self.events: List[List[Tuple[InstructionBase, float, float]]] = []

    @classmethod
    def from_profile(cls, schedule_cls: type, profile_file: Union[str, Path]) -> "SimulationEngine":
        data = json.loads(Path(profile_file).read_text())
        acc: Dict[Tuple[int, str, int], List[float]] = defaultdict(list)
        for ob in data["observations"]:
            if ob.get("data_parallel_rank", 0) != 0 or ob.get("model_parallel_rank", 0) != 0:
                continue
            acc[(ob["pipe_parallel_rank"], ob["timer_name"], ob["micro_batch_id"])].append(ob["duration"])
        durations = {k: sum(v) / len(v) for k, v in acc.items()}
        return cls(schedule_cls, data["pipe_parallel_size"], data["gradient_accumulation_steps"], durations)

    def get_duration(self, stage: int, ins: InstructionBase) -> float:
        name = _timer_name(ins)
        mb = ins.micro_batch_id if ins.micro_batch_id is not None else -1
        for key in ((stage, name, mb), (stage, name, -1)):
            if key in self.durations:
                return self.durations[key]
        same = [v for (s, n, _), v in self.durations.items() if s == stage and n == name]
        return sum(same) / len(same) if same else 0.0

    def transfer_duration(self, stage: int, ins: InstructionBase) -> float:
        partner_cls, delta = _PARTNER[type(ins)]
        mine = self.get_duration(stage, ins)
        theirs = self.get_duration(stage + delta, partner_cls(ins.buffer_id, ins.micro_batch_id))
        return min(mine, theirs) if theirs > 0 else mine

    def simulate(self) -> None:
        n = self.pipe_parallel_size
        pc = [0] * n
        clock = [0.0] * n
        self.events = [[] for _ in range(n)]
        guard = 0
        while any(pc[r] < len(self.programs[r]) for r in range(n)):
            progressed = False
            for r in range(n):
                if pc[r] >= len(self.programs[r]):
                    continue
                ins = self.programs[r][pc[r]]
                if isinstance(ins, _COMM):
                    partner_cls, delta = _PARTNER[type(ins)]
                    p = r + delta
                    if not (0 <= p < n and pc[p] < len(self.programs[p])):
                        continue
                    other = self.programs[p][pc[p]]
                    if not (isinstance(other, partner_cls) and other.micro_batch_id == ins.micro_batch_id):
                        continue
                    if r > p:  # handle each pair once, from the lower stage
                        continue
                    start = max(clock[r], clock[p])
                    dur = self.transfer_duration(r, ins)
                    self.events[r].append((ins, start, start + dur))
                    self.events[p].append((other, start, start + dur))
                    clock[r] = clock[p] = start + dur
                    pc[r] += 1
                    pc[p] += 1
                    progressed = True
                else:
                    dur = self.get_duration(r, ins)
                    self.events[r].append((ins, clock[r], clock[r] + dur))
                    clock[r] += dur
                    pc[r] += 1
                    progressed = True
            guard += 1
            if not progressed:
                raise RuntimeError(f"simulation deadlocks at program counters {pc}")
        self.makespan = max(clock) if clock else 0.0

    def summarize(self) -> Dict[str, Any]:
        stages = []
        for r, ev in enumerate(self.events):
            busy = sum(e - s for _, s, e in ev)
            stages.append({"pipe_parallel_rank": r, "busy": busy, "idle": self.makespan - busy,
                           "idle_percent": 100.0 * (self.makespan - busy) / self.makespan if self.makespan else 0.0})
        total_idle = sum(s["idle"] for s in stages)
        return {"duration": self.makespan, "stages": stages,
                "idle_percent": 100.0 * total_idle / (self.makespan * len(stages)) if self.makespan and stages else 0.0}

    def visualize(self, milliseconds_per_pixel: float = 10, pipe_pixels: int = 200):
        from PIL import Image, ImageDraw

        px_per_s = 1000.0 / milliseconds_per_pixel
        width = max(int(self.makespan * px_per_s) + 20, 40)
        width = min(width, 20000)
        scale = (width - 20) / (self.makespan * px_per_s) if self.makespan else 1.0
        img = Image.new("RGB", (width, pipe_pixels * self.pipe_parallel_size + 10), (30, 30, 30))
        draw = ImageDraw.Draw(img)
        for r, ev in enumerate(self.events):
            y0 = 5 + r * pipe_pixels
            for ins, s, e in ev:
                x0, x1 = 10 + s * px_per_s * scale, 10 + e * px_per_s * scale
                draw.rectangle([x0, y0, max(x1, x0 + 1), y0 + pipe_pixels - 6],
                               fill=_COLORS.get(_timer_name(ins), (200, 200, 200)))
        return img
