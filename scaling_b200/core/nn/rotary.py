"""Rotary position embeddings (reference: core/nn/rotary.py:1-255, rotary_config.py).

Two flavours with one table: ``RotaryEmbedding`` rotates the pairs (x[i], x[i + d/2]) ("NeoX" half rotation, optionally
on the first ``dimensions`` features only), ``RotaryEmbeddingComplex`` the interleaved pairs (x[2i], x[2i+1]) (Llama).
cos/sin are kept in fp32 for both (the reference keeps the half-rotation tables in the model dtype and multiplies in
bf16 — SURVEY §9.2 #13).  The modules keep the reference's ``[sq, b, heads, d]`` call signature; the attention layer
uses ``apply_packed`` on token-major ``[tokens, heads, d]`` tensors, which runs the fused in-place kernel on B200.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from pydantic import Field

from ...ops import functional as OF
from ..config import BaseConfig


class RotaryConfig(BaseConfig):
    dimensions: int = Field(0, description="number of leading head features that are rotated")
    base: int = Field(10000, description="frequency base")
    max_seq_length: int = Field(2048, description="length of the precomputed cos/sin tables")


def _tables(config: RotaryConfig, device: torch.device) -> Tuple[torch.Tensor, torch.Tensor]:
    dim = config.dimensions
    inv_freq = 1.0 / (float(config.base) ** (torch.arange(0, dim, 2, device=device).float()[: dim // 2] / dim))
    ang = torch.outer(torch.arange(config.max_seq_length, device=device).float(), inv_freq)
    return ang.cos().contiguous(), ang.sin().contiguous()


class _RotaryBase(torch.nn.Module):
    mode = 0  # 0 interleaved pairs, 1 half rotation

    def __init__(self, config: RotaryConfig, device: torch.device, dtype: torch.dtype = torch.float32):
        super().__init__()
        assert config.dimensions > 1, "RotaryEmbedding cannot use `dim` == 1"
        self.dimensions = config.dimensions
        self.max_seq_length = config.max_seq_length
        cos, sin = _tables(config, device)
        self.register_buffer("cos_table", cos, persistent=False)
        self.register_buffer("sin_table", sin, persistent=False)

    # ---- math on [..., heads, d] with per-position cos/sin broadcast ---------------------------------
    def _rotate(self, x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
        rot = self.dimensions
        xf = x[..., :rot].float()
        if self.mode == 0:
            a, b = xf[..., 0::2], xf[..., 1::2]
            out = torch.stack((a * cos - b * sin, a * sin + b * cos), dim=-1).flatten(-2)
        else:
            a, b = xf[..., : rot // 2], xf[..., rot // 2:]
            out = torch.cat((a * cos - b * sin, a * sin + b * cos), dim=-1)
        out = out.to(x.dtype)
        if rot != x.shape[-1]:
            out = torch.cat((out, x[..., rot:]), dim=-1)
        return out

    def apply_packed(self, x: torch.Tensor, position_ids: Optional[torch.Tensor], seq: int) -> torch.Tensor:
        """x: [tokens, heads, d] token-major (tokens = batch * seq); position_ids: [tokens] or None (= t mod seq)."""
        if OF.native_enabled(x) and x.dtype == torch.bfloat16:
            pos = None if position_ids is None else position_ids.to(torch.int32).contiguous()
            return OF.rope(x, self.cos_table, self.sin_table, pos, seq, self.mode, self.dimensions)
        if position_ids is None:
            position_ids = torch.arange(x.shape[0], device=x.device) % seq
        cos = self.cos_table[position_ids.long()].unsqueeze(1)
        sin = self.sin_table[position_ids.long()].unsqueeze(1)
        return self._rotate(x, cos, sin)

    def forward(self, query: torch.Tensor, key: torch.Tensor, query_position_ids: Optional[torch.Tensor] = None,
                key_position_ids: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """query/key: [sq, b, heads, d]; position ids: [sq, b] or None (positions 0..sq-1)."""

        def one(x: torch.Tensor, pos: Optional[torch.Tensor]) -> torch.Tensor:
            sq = x.shape[0]
            if pos is None:
                cos = self.cos_table[:sq].view(sq, 1, 1, -1)
                sin = self.sin_table[:sq].view(sq, 1, 1, -1)
            else:
                cos = self.cos_table[pos.long()].unsqueeze(2)
                sin = self.sin_table[pos.long()].unsqueeze(2)
            return self._rotate(x, cos, sin)

        return one(query, query_position_ids), one(key, key_position_ids)


class RotaryEmbedding(_RotaryBase):
    mode = 1


class RotaryEmbeddingComplex(_RotaryBase):
    mode = 0

    def __init__(self, config: RotaryConfig, device: torch.device, dtype: torch.dtype = torch.float32):
        super().__init__(config, device, dtype)
