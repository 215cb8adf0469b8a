"""Masked softmax (reference: core/nn/masked_softmax/masked_softmax.py:8-30, masked_softmax_config.py).

Only used by the plain-PyTorch attention path (oracle / CPU / attention-score manipulation); the fused attention
kernels compute their softmax online in fp32.
"""
from __future__ import annotations

from enum import Enum

import torch
from pydantic import Field

from ..config import BaseConfig


class MaskedSoftmaxKernel(Enum):
    TORCH = "torch"
    FLASH_ATTENTION = "flash_attention"


class MaskedSoftmaxConfig(BaseConfig):
    kernel: MaskedSoftmaxKernel = Field(MaskedSoftmaxKernel.TORCH, description="'torch' = dense reference path, 'flash_attention' = fused attention kernel")
    softmax_in_fp32: bool = Field(False, description="upcast scores to fp32 for the softmax (dense path only; fused kernels always do)")
    scale: float = Field(1.0, description="scores are multiplied by this before the softmax")
    deterministic_flash_attn_bwd: bool = Field(False, description="deterministic backward of the fused attention kernel")


class MaskedSoftmax(torch.nn.Module):
    def __init__(self, config: MaskedSoftmaxConfig):
        super().__init__()
        self.config = config

    def forward(self, x: torch.Tensor, mask: torch.Tensor) -> torch.Tensor:
        """mask: True = masked out."""
        in_dtype = x.dtype
        if self.config.softmax_in_fp32 and in_dtype != torch.float32:
            x = x.float()
        if self.config.scale != 1.0:
            x = x * self.config.scale
        x = x.masked_fill(mask, -10000.0)
        p = torch.nn.functional.softmax(x, dim=-1)
        return p.to(in_dtype) if p.dtype != in_dtype else p
