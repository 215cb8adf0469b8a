"""FLOP accounting and MFU (reference: transformer/utils/get_tflops.py:1-401).

Same function names and call signatures; the per-operation counters are derived from the *configured* architecture
(GQA/MQA key-value width, SwiGLU's three matrices, ``mlp_factor``) instead of assuming MHA + 4h GELU
(SURVEY §9.2 #2), B200 is a known device and an unknown device yields ``mfu = nan`` rather than a division by zero
(§9.2 #1).
"""
from __future__ import annotations

import math
from enum import Enum
from typing import Any

import torch

from ...core.logging import logger
from ..config import MLPType, TransformerArchitectureConfig


def _per_token_forward_matmul_flops(a: TransformerArchitectureConfig) -> float:
    """Forward matmul FLOPs per token for the whole model, attention score/value products included (causal halves
    them), LM head included."""
    h, L, s, V = a.hidden_size, a.num_layers, a.sequence_length, a.vocab_size
    kv = h if not a.attention_num_kv_heads else h * a.attention_num_kv_heads // a.num_attention_heads
    f = int(h * a.mlp_factor)
    proj = 2 * h * (h + 2 * kv) + 2 * h * h
    mlp = 2 * h * f * (3 if a.mlp_type == MLPType.SWIGLU else 2)
    attn = 2 * 2 * s * h * (0.5 if a.causal else 1.0)
    return L * (proj + mlp + attn) + 2 * h * V


def _tflops(a: TransformerArchitectureConfig, topology: Any, iter_time_s: float, backward_factor: float) -> float:
    tokens = topology.config.global_batch_size * a.sequence_length
    return _per_token_forward_matmul_flops(a) * backward_factor * tokens / (iter_time_s * topology.config.world_size * 1e12)


def _recompute(topology: Any) -> bool:
    return getattr(topology.config.activation_checkpointing_type, "value", "disabled") != "disabled"


def get_tflops_aleph_alpha(iter_time_s: float, topology: Any, transformer_architecture: TransformerArchitectureConfig) -> float:
    """Model FLOPs (forward + 2× backward, no recomputation) per GPU in TFLOP/s."""
    return _tflops(transformer_architecture, topology, iter_time_s, 3.0)


def get_tflops_electra(iter_time_s: float, topology: Any, transformer_architecture: TransformerArchitectureConfig) -> float:
    """Same operation count as :func:`get_tflops_aleph_alpha` (the two differ only in bookkeeping in the reference)."""
    return _tflops(transformer_architecture, topology, iter_time_s, 3.0)


def get_tflops_bloom(iter_time_s: float, topology: Any, transformer_architecture: TransformerArchitectureConfig) -> float:
    """Hardware FLOPs: counts the recomputed forward when activation checkpointing is on (×4 instead of ×3)."""
    return _tflops(transformer_architecture, topology, iter_time_s, 4.0 if _recompute(topology) else 3.0)


def get_tflops_megatron(parameter_count: int, iter_time_s: float, topology: Any,
                        transformer_architecture: TransformerArchitectureConfig) -> float:
    """(6·B·s·N + 60·B·s²·h·L) / (t · world) — the reference's parameter-count based estimate (:319-334)."""
    a = transformer_architecture
    B = topology.config.global_batch_size
    ff = B * a.sequence_length * parameter_count * 6
    attn = B * a.sequence_length * a.sequence_length * a.hidden_size * a.num_layers * 60
    return (ff + attn) / (iter_time_s * topology.config.world_size * 1e12)


class HardwareType(Enum):
    A100 = "a100"
    H100 = "h100"
    H200 = "h200"
    B200 = "b200"
    RTX3090 = "rtx3090"
    RTX4090 = "rtx4090"
    DEFAULT = "default"

    @property
    def max_tflops(self) -> float:
        """Dense bf16/fp16 tensor-core peak in FLOP/s."""
        return {"a100": 312e12, "h100": 989.4e12, "h200": 989.4e12, "b200": 2250e12, "rtx3090": 35.58e12,
                "rtx4090": 82.58e12, "default": 0.0}[self.value]

    @classmethod
    def from_name(cls, device_name: str) -> "HardwareType":
        clean = device_name.replace(" ", "").lower().replace("nvidia", "")
        return next((x for x in cls if x != cls.DEFAULT and x.value in clean), cls.DEFAULT)

    @classmethod
    def get_via_torch(cls) -> "HardwareType":
        try:
            name = torch.cuda.get_device_name()
        except Exception:  # noqa: BLE001  (no CUDA device / driver)
            return cls.DEFAULT
        hw = cls.from_name(name)
        if hw == cls.DEFAULT:
            logger.warning(f"cuda device {name} does not match any known HardwareType")
        return hw


def get_model_flop_utilization_palm(iter_time_s: float, parameter_count: int, topology: Any,
                                    transformer_architecture: TransformerArchitectureConfig,
                                    peak_flops: float | None = None) -> float:
    """PaLM MFU (arXiv:2204.02311): tokens/s ÷ (peak·world ÷ (6N + 12·L·h·s))."""
    a = transformer_architecture
    peak = HardwareType.get_via_torch().max_tflops if peak_flops is None else peak_flops
    if peak <= 0:
        return math.nan
    tokens_per_second = topology.config.global_batch_size * a.sequence_length / iter_time_s
    model_flops = 6 * parameter_count + 12 * a.num_layers * a.hidden_size * a.sequence_length
    return tokens_per_second * model_flops / (peak * topology.config.world_size)
