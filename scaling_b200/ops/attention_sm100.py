"""Binding of the sm_100a attention kernels (csrc/attention_sm100.cu)."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import call, ptr

_BWD_IMPL = os.environ.get("SCALING_B200_ATTENTION_BWD", "auto")  # auto | sm100 | flash


def sm100_attention_supported(q, k, v, cu_q, cu_k, causal, window, dropout_p) -> bool:
    lib = _lib.load() if torch.cuda.is_available() else None
    if lib is None or not hasattr(lib, "sb_attention_fwd"):
        return False
    return (q.dtype == torch.bfloat16 and q.shape[-1] == 128 and not window and dropout_p == 0.0 and cu_q is cu_k
            and q.shape[1] % k.shape[1] == 0 and q.stride(2) == 1 and q.stride(1) == 128 and k.stride(2) == 1
            and k.stride(1) == 128 and v.stride(2) == 1 and v.stride(1) == 128 and q.stride(0) % 8 == 0
            and k.stride(0) % 8 == 0 and v.stride(0) % 8 == 0)


def segment_bounds(cu_seqlens: torch.Tensor, total: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """int32 [total] start / end (exclusive) of the sequence every packed token belongs to — device side, no sync."""
    cu = cu_seqlens.to(torch.int64)
    pos = torch.arange(total, device=cu.device)
    seg = torch.searchsorted(cu[1:].contiguous(), pos, right=True).clamp_(max=cu.numel() - 2)
    return cu[seg].to(torch.int32).contiguous(), cu[seg + 1].to(torch.int32).contiguous()


def attention_fwd(q, k, v, seg_start, seg_end, causal: bool, scale: float):
    T, heads, d = q.shape
    out = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    lse = torch.empty((heads, T), dtype=torch.float32, device=q.device)
    call("sb_attention_fwd", ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse), ptr(seg_start), ptr(seg_end), T, heads, k.shape[1],
         q.stride(0), k.stride(0), v.stride(0), float(scale), int(causal))
    return out, lse


def has_native_bwd() -> bool:
    lib = _lib.load()
    return lib is not None and hasattr(lib, "sb_attention_bwd")


def attention_bwd(dout, q, k, v, out, lse, seg_start, seg_end, causal: bool, scale: float):
    T, heads, d = q.shape
    kvh = k.shape[1]
    dq = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    dk = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    dv = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    lib = _lib.load(required=True)
    delta = torch.zeros(lib.sb_attention_bwd_scratch_floats(T, heads), dtype=torch.float32, device=q.device)
    call("sb_attention_bwd", ptr(dout), ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse), ptr(delta), ptr(dq), ptr(dk), ptr(dv),
         ptr(seg_start), ptr(seg_end), T, heads, kvh, q.stride(0), k.stride(0), v.stride(0), float(scale), int(causal),
         launches=3)
    return dq, dk, dv


class AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens, max_seqlen, causal, scale):
        seg_start, seg_end = segment_bounds(cu_seqlens, q.shape[0])
        out, lse = attention_fwd(q, k, v, seg_start, seg_end, causal, scale)
        ctx.save_for_backward(q, k, v, out, lse, seg_start, seg_end, cu_seqlens)
        ctx.args = (max_seqlen, causal, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse, seg_start, seg_end, cu = ctx.saved_tensors
        max_seqlen, causal, scale = ctx.args
        dout = dout.contiguous()
        impl = _BWD_IMPL
        if impl == "auto":
            impl = "sm100" if has_native_bwd() else "flash"
        if impl == "sm100":
            dq, dk, dv = attention_bwd(dout, q, k, v, out, lse, seg_start, seg_end, causal, scale)
        else:
            # library backward on this kernel's forward results (same LSE convention: natural log, [heads, tokens])
            from flash_attn.flash_attn_interface import _flash_attn_varlen_backward

            qc, kc, vc = q.contiguous(), k.contiguous(), v.contiguous()
            dq, dk, dv = torch.empty_like(qc), torch.empty_like(kc), torch.empty_like(vc)
            cu32 = cu.to(torch.int32)
            _flash_attn_varlen_backward(dout, qc, kc, vc, out, lse, dq, dk, dv, cu32, cu32, max_seqlen, max_seqlen, 0.0,
                                        scale, causal, -1, -1, 0.0, None, False, None)
        return dq, dk, dv, None, None, None, None


def sm100_attention(q, k, v, cu_seqlens, max_seqlen, causal, scale):
    return AttentionFn.apply(q, k, v, cu_seqlens, max_seqlen, causal, scale)
