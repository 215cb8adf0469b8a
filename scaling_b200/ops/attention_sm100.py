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


_bounds_cache: "OrderedDict[tuple, tuple]" = None  # type: ignore[assignment]
_BOUNDS_CACHE_ENTRIES = 16  # micro batches in flight under 1F1B keep their own entry


def _compute_segment_bounds(cu_seqlens: torch.Tensor, total: int) -> Tuple[torch.Tensor, torch.Tensor]:
    cu = cu_seqlens.to(torch.int64)
    pos = torch.arange(total, device=cu.device)
    seg = torch.searchsorted(cu[1:].contiguous(), pos, right=True).clamp_(max=cu.numel() - 2)
    return cu[seg].to(torch.int32).contiguous(), cu[seg + 1].to(torch.int32).contiguous()


def segment_bounds(cu_seqlens: torch.Tensor, total: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """int32 [total] start / end (exclusive) of the sequence every packed token belongs to — device side, no sync.

    Every layer of a micro batch passes the SAME ``cu_seqlens`` tensor (it travels in the layer IO), so the result is
    computed once per micro batch and looked up by tensor identity + version counter afterwards."""
    global _bounds_cache
    from collections import OrderedDict

    if _bounds_cache is None:
        _bounds_cache = OrderedDict()
    key = (id(cu_seqlens), cu_seqlens._version, cu_seqlens.data_ptr(), total)
    hit = _bounds_cache.get(key)
    if hit is not None and hit[0]() is cu_seqlens:
        _bounds_cache.move_to_end(key)
        return hit[1], hit[2]
    import weakref

    start, end = _compute_segment_bounds(cu_seqlens, total)
    _bounds_cache[key] = (weakref.ref(cu_seqlens), start, end)
    while len(_bounds_cache) > _BOUNDS_CACHE_ENTRIES:
        _bounds_cache.popitem(last=False)
    return start, end


def attention_fwd(q, k, v, seg_start, seg_end, causal: bool, scale: float):
    T, heads, d = q.shape
    out = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    lse = torch.empty((heads, T), dtype=torch.float32, device=q.device)
    call("sb_attention_fwd", ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse), ptr(seg_start), ptr(seg_end), T, heads, k.shape[1],
         q.stride(0), k.stride(0), v.stride(0), float(scale), int(causal))
    return out, lse


_bwd_scratch_pool: dict = {}


def _bwd_scratch(device: torch.device, T: int, heads: int) -> torch.Tensor:
    """Persistent lse·log2e / delta scratch per (device, T, heads): zero-filled once (its padding is read but masked and
    must stay finite), then reused by every backward call of that shape on the stream."""
    key = (device.index, T, heads)
    buf = _bwd_scratch_pool.get(key)
    if buf is None:
        lib = _lib.load(required=True)
        if len(_bwd_scratch_pool) >= 8:
            _bwd_scratch_pool.clear()
        buf = torch.zeros(lib.sb_attention_bwd_scratch_floats(T, heads), dtype=torch.float32, device=device)
        _bwd_scratch_pool[key] = buf
    return buf


def has_native_bwd() -> bool:
    lib = _lib.load()
    return lib is not None and hasattr(lib, "sb_attention_bwd")


def attention_bwd(dout, q, k, v, out, lse, seg_start, seg_end, causal: bool, scale: float):
    T, heads, d = q.shape
    kvh = k.shape[1]
    dq = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    dk = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    dv = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    delta = _bwd_scratch(q.device, T, heads)
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
