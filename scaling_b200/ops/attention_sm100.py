"""Binding of the sm_100a attention kernels (csrc/attention_sm100.cu)."""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import call, ptr

_BWD_IMPL = os.environ.get("SCALING_B200_ATTENTION_BWD", "auto")  # auto | sm100 | flash


def sm100_attention_supported(q, k, v, cu_q, cu_k, causal, window, dropout_p) -> bool:
    """bf16 or fp16, head dim <= 128 (smaller heads are zero-padded to the kernel's 128), GQA/MQA, packed sequences, causal or
    bidirectional, sliding windows, attention-probability dropout (stateless hash mask); keys may differ from queries
    (KV-cache decoding) when no gradient is needed."""
    lib = _lib.load() if torch.cuda.is_available() else None
    if lib is None or not hasattr(lib, "sb_attention_fwd_ex"):
        return False
    d = q.shape[-1]
    if not (q.dtype in (torch.bfloat16, torch.float16) and k.dtype == q.dtype and v.dtype == q.dtype and d <= 128 and d % 8 == 0 and 0.0 <= dropout_p < 1.0 and q.shape[1] % k.shape[1] == 0):
        return False
    if cu_q is not cu_k and (q.requires_grad or k.requires_grad or v.requires_grad) and torch.is_grad_enabled():
        return False  # decoding against a cache is forward only
    if cu_q is not cu_k and (cu_q.numel() != 2 or cu_k.numel() != 2):
        return False  # one sequence (the reference's KV cache is batch size 1 as well)
    return True


def _strided_ok(t: torch.Tensor) -> bool:
    return t.stride(2) == 1 and t.stride(1) == 128 and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0


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


def window_bounds(cu_seqlens: torch.Tensor, total: int, causal: bool, window: int):
    """Per-token key bounds under a sliding window of ``window`` tokens to the left (and, when not causal, to the right):
    (lo, hi, key_qend) int32 [total].  Row t sees keys [lo[t], hi[t]) — the kernels take these in place of the sequence
    bounds — and key t is seen by queries up to key_qend[t] (lets the dK/dV blocks stop early)."""
    start, end = segment_bounds(cu_seqlens, total)
    pos = torch.arange(total, device=start.device, dtype=torch.int32)
    lo = torch.maximum(start, pos - window)
    qend = torch.minimum(end, pos + window + 1)
    hi = end if causal else qend
    return lo.contiguous(), hi.contiguous(), qend.contiguous()


def attention_fwd(q, k, v, seg_start, seg_end, causal: bool, scale: float, dropout=(0.0, 0, 0)):
    """``dropout`` = (p, seed, global index of the first local head)."""
    T, heads, d = q.shape
    out = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    lse = torch.empty((heads, T), dtype=torch.float32, device=q.device)
    call("sb_attention_fwd_f16" if q.dtype == torch.float16 else "sb_attention_fwd_ex", ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse), ptr(seg_start), ptr(seg_end), T, k.shape[0],
         heads, k.shape[1], q.stride(0), k.stride(0), v.stride(0), float(scale), int(causal), float(dropout[0]),
         int(dropout[1]) & 0xFFFFFFFFFFFFFFFF, int(dropout[2]))
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


def attention_bwd(dout, q, k, v, out, lse, seg_start, seg_end, causal: bool, scale: float, key_qend=None,
                  dropout=(0.0, 0, 0)):
    T, heads, d = q.shape
    kvh = k.shape[1]
    dq = torch.empty((T, heads, d), dtype=q.dtype, device=q.device)
    dk = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    dv = torch.empty((T, kvh, d), dtype=q.dtype, device=q.device)
    delta = _bwd_scratch(q.device, T, heads)
    call("sb_attention_bwd_f16" if q.dtype == torch.float16 else "sb_attention_bwd_ex", ptr(dout), ptr(q), ptr(k), ptr(v), ptr(out), ptr(lse), ptr(delta), ptr(dq), ptr(dk), ptr(dv),
         ptr(seg_start), ptr(seg_end), ptr(key_qend), T, heads, kvh, q.stride(0), k.stride(0), v.stride(0), float(scale), int(causal),
         float(dropout[0]), int(dropout[1]) & 0xFFFFFFFFFFFFFFFF, int(dropout[2]), launches=3)
    return dq, dk, dv


class AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens, max_seqlen, causal, scale, window, dropout=(0.0, 0, 0)):
        if window:
            seg_start, seg_end, key_qend = window_bounds(cu_seqlens, q.shape[0], causal, int(window))
        else:
            seg_start, seg_end = segment_bounds(cu_seqlens, q.shape[0])
            key_qend = None
        out, lse = attention_fwd(q, k, v, seg_start, seg_end, causal, scale, dropout)
        ctx.save_for_backward(q, k, v, out, lse, seg_start, seg_end, cu_seqlens)
        ctx.key_qend = key_qend
        ctx.dropout = dropout
        ctx.args = (max_seqlen, causal, scale, window)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse, seg_start, seg_end, cu = ctx.saved_tensors
        max_seqlen, causal, scale, window = ctx.args
        dout = dout.contiguous()
        impl = _BWD_IMPL
        if impl == "auto":
            impl = "sm100" if has_native_bwd() else "flash"
        if impl == "sm100" or ctx.dropout[0] > 0.0:
            dq, dk, dv = attention_bwd(dout, q, k, v, out, lse, seg_start, seg_end, causal, scale, ctx.key_qend, ctx.dropout)
        else:
            # library backward on this kernel's forward results (same LSE convention: natural log, [heads, tokens])
            from flash_attn.flash_attn_interface import _flash_attn_varlen_backward

            qc, kc, vc = q.contiguous(), k.contiguous(), v.contiguous()
            dq, dk, dv = torch.empty_like(qc), torch.empty_like(kc), torch.empty_like(vc)
            cu32 = cu.to(torch.int32)
            w = (-1, -1) if not window else (int(window), 0 if causal else int(window))
            _flash_attn_varlen_backward(dout, qc, kc, vc, out, lse, dq, dk, dv, cu32, cu32, max_seqlen, max_seqlen, 0.0,
                                        scale, causal, w[0], w[1], 0.0, None, False, None)
        return (dq, dk, dv) + (None,) * (len(ctx.needs_input_grad) - 3)


def _decode_forward(q, k, v, causal, scale, window):
    """Queries of ONE sequence against a longer key/value cache (right-aligned causal mask), forward only."""
    tq, tk = q.shape[0], k.shape[0]
    pos = torch.arange(tq, device=q.device, dtype=torch.int32) + (tk - tq)  # key index of each query token
    hi = (pos + 1) if causal else torch.full_like(pos, tk)
    lo = torch.clamp(pos - int(window), min=0) if window else torch.zeros_like(pos)
    if window and not causal:
        hi = torch.clamp(pos + int(window) + 1, max=tk)
    out, _ = attention_fwd(q, k, v, lo.contiguous(), hi.contiguous(), False, scale)
    return out


def sm100_attention(q, k, v, cu_seqlens, max_seqlen, causal, scale, window=None, cu_seqlens_k=None, dropout=(0.0, 0, 0)):
    d = q.shape[-1]
    if d < 128:  # the kernel's tiles are 128 wide: zero columns change neither the scores nor the first d outputs
        pad = (0, 128 - d)
        q, k, v = (torch.nn.functional.pad(t, pad) for t in (q, k, v))
    q, k, v = (t if _strided_ok(t) else t.contiguous() for t in (q, k, v))
    if cu_seqlens_k is not None and cu_seqlens_k is not cu_seqlens:
        out = _decode_forward(q, k, v, causal, scale, window)
    else:
        out = AttentionFn.apply(q, k, v, cu_seqlens, max_seqlen, causal, scale, window or None, tuple(dropout))
    return out[..., :d] if d < 128 else out
