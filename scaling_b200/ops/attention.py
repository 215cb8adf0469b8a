"""Scaled-dot-product attention over packed (varlen) token-major tensors.

    q: [Tq, heads, d]   k, v: [Tk, kv_heads, d]   cu_seqlens: int32 [n_seq + 1] boundaries inside the packed dimension

Paths, selected by device and shape (not a user-visible backend switch):
  * ``sm100``  — the hand-written sm_100a kernel (csrc/attention_sm100.cu: tcgen05 QKᵀ / PV with TMEM accumulators,
                 online softmax in fp32; causal / bidirectional, packed sequences, GQA/MQA, sliding windows, KV-cache
                 decoding), used for bf16 with head_dim <= 128 on a B200;
  * ``dense``  — plain PyTorch expression with an explicit mask: the oracle, the CPU path, and the only path that
                 supports attention-score manipulation (reference core/nn/attention/attention.py:96-201);
  * ``flash``  — the flash-attn wheel, only when selected by name: the library baseline for A/B measurements
                 (reference :204-259).  ``auto`` never picks it.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch

from . import _lib

_ATTN_IMPL = os.environ.get("SCALING_B200_ATTENTION", "auto")  # auto | sm100 | flash | dense


def set_attention_impl(name: str) -> None:
    global _ATTN_IMPL
    assert name in ("auto", "sm100", "flash", "dense")
    _ATTN_IMPL = name


def repeat_kv(x: torch.Tensor, n_rep: int) -> torch.Tensor:
    """[T, kv_heads, d] -> [T, kv_heads * n_rep, d] (each kv head serves n_rep consecutive query heads)."""
    if n_rep == 1:
        return x
    t, h, d = x.shape
    return x[:, :, None, :].expand(t, h, n_rep, d).reshape(t, h * n_rep, d)


def segment_ids(cu_seqlens: torch.Tensor, total: int) -> torch.Tensor:
    """Sequence index of every packed token, from the boundaries (no host sync)."""
    pos = torch.arange(total, device=cu_seqlens.device)
    return torch.searchsorted(cu_seqlens[1:].to(torch.int64).contiguous(), pos, right=True)


def dense_mask(cu_q: torch.Tensor, tq: int, cu_k: torch.Tensor, tk: int, causal: bool,
               window: Optional[int]) -> torch.Tensor:
    """[tq, tk] bool, True = masked out. Tokens attend inside their own sequence; causal/window are relative to the
    sequence starts, with keys right-aligned to queries (KV-cache decoding: tk >= tq)."""
    seg_q, seg_k = segment_ids(cu_q, tq), segment_ids(cu_k, tk)
    cu_q64, cu_k64 = cu_q.to(torch.int64), cu_k.to(torch.int64)
    pos_q = torch.arange(tq, device=cu_q.device) - cu_q64[seg_q.clamp(max=len(cu_q) - 2)]
    pos_k = torch.arange(tk, device=cu_k.device) - cu_k64[seg_k.clamp(max=len(cu_k) - 2)]
    len_q = (cu_q64[1:] - cu_q64[:-1])[seg_q.clamp(max=len(cu_q) - 2)]
    len_k = (cu_k64[1:] - cu_k64[:-1])[seg_q.clamp(max=len(cu_k) - 2)]  # key length of each QUERY's sequence
    pos_q = pos_q + (len_k - len_q)  # right-align
    mask = seg_q[:, None] != seg_k[None, :]
    if causal:
        mask = mask | (pos_k[None, :] > pos_q[:, None])
    if window is not None and window > 0:
        mask = mask | (pos_k[None, :] < pos_q[:, None] - window)
        if not causal:
            mask = mask | (pos_k[None, :] > pos_q[:, None] + window)
    return mask


def attention_dense(q, k, v, cu_q, cu_k, causal, scale, window=None, dropout_p=0.0, training=False,
                    softmax_fn=None, dropout_fn=None, scores_manipulation=None, manipulation_log_additive=True):
    tq, nh, d = q.shape
    tk, nkv, _ = k.shape
    k, v = repeat_kv(k, nh // nkv), repeat_kv(v, nh // nkv)
    mask = dense_mask(cu_q, tq, cu_k, tk, causal, window)
    scores = torch.einsum("qhd,khd->hqk", q, k) * scale
    if scores_manipulation is not None:
        if manipulation_log_additive:
            scores = scores + scores_manipulation
        else:
            scores = scores * scores_manipulation
    if softmax_fn is not None:
        probs = softmax_fn(scores, mask[None])
    else:
        probs = torch.softmax(scores.float().masked_fill(mask[None], float("-inf")), dim=-1).to(q.dtype)
        probs = torch.nan_to_num(probs)
    if dropout_fn is not None:
        probs = dropout_fn(probs)
    elif dropout_p > 0.0 and training:
        probs = torch.nn.functional.dropout(probs, dropout_p)
    return torch.einsum("hqk,khd->qhd", probs, v)


_flash_fn = None


def _flash():
    global _flash_fn
    if _flash_fn is None:
        try:
            from flash_attn import flash_attn_varlen_func

            _flash_fn = flash_attn_varlen_func
        except Exception:  # noqa: BLE001
            _flash_fn = False
    return _flash_fn


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: torch.Tensor, max_seqlen_q: int,
              causal: bool = True, scale: Optional[float] = None, cu_seqlens_k: Optional[torch.Tensor] = None,
              max_seqlen_k: Optional[int] = None, window: Optional[int] = None, dropout_p: float = 0.0,
              training: bool = False, deterministic: bool = False, softmax_fn=None, dropout_fn=None,
              scores_manipulation=None, manipulation_log_additive=True, dropout_seed: Optional[int] = None,
              head_offset: int = 0) -> torch.Tensor:
    """``dropout_seed`` / ``head_offset``: seed of the (stateless) dropout mask of the sm_100a kernel and the global index
    of the first local head — pass the topology's counter-based seed so that every replay (activation checkpointing)
    reproduces the mask; without a seed one is drawn from torch's CPU generator."""
    scale = 1.0 / math.sqrt(q.shape[-1]) if scale is None else scale
    cu_k = cu_seqlens_q if cu_seqlens_k is None else cu_seqlens_k
    max_k = max_seqlen_q if max_seqlen_k is None else max_seqlen_k
    p = dropout_p if training else 0.0
    impl = _ATTN_IMPL
    fusable = (q.is_cuda and q.dtype in (torch.bfloat16, torch.float16) and scores_manipulation is None
               and softmax_fn is None and dropout_fn is None)
    if impl == "auto":
        impl = "dense"
        if fusable:
            from .attention_sm100 import sm100_attention_supported

            # no library kernel behind "auto": what the sm_100a kernel does not cover (head dims above 128,
            # gradients through a KV cache) runs the plain PyTorch expression; the flash-attn
            # wheel is only used when asked for by name (SCALING_B200_ATTENTION=flash, A/B measurements)
            if sm100_attention_supported(q, k, v, cu_seqlens_q, cu_k, causal, window, p):
                impl = "sm100"
    if impl == "sm100" and fusable:
        from .attention_sm100 import sm100_attention

        drop = (0.0, 0, 0)
        if p > 0.0:
            seed = int(torch.randint(0, 2**62, (1,)).item()) if dropout_seed is None else int(dropout_seed)
            drop = (p, seed, head_offset)
        return sm100_attention(q, k, v, cu_seqlens_q, max_seqlen_q, causal, scale, window=window, cu_seqlens_k=cu_k,
                               dropout=drop)
    if impl == "flash" and fusable and _flash():
        w = (-1, -1) if not window else (window, 0 if causal else window)
        return _flash()(q.contiguous(), k.contiguous(), v.contiguous(), cu_seqlens_q=cu_seqlens_q.to(torch.int32),
                        cu_seqlens_k=cu_k.to(torch.int32), max_seqlen_q=max_seqlen_q, max_seqlen_k=max_k, dropout_p=p,
                        softmax_scale=scale, causal=causal, window_size=w, deterministic=deterministic)
    return attention_dense(q, k, v, cu_seqlens_q, cu_k, causal, scale, window, p, training, softmax_fn, dropout_fn,
                           scores_manipulation, manipulation_log_additive)
