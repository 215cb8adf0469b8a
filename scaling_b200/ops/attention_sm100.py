"""Binding of the sm_100a attention kernels (csrc/attention_sm100.cu)."""
from __future__ import annotations

import torch

from . import _lib


def sm100_attention_supported(q, k, v, cu_q, cu_k, causal, window, dropout_p) -> bool:
    lib = _lib.load() if torch.cuda.is_available() else None
    if lib is None or not hasattr(lib, "sb_attention_fwd"):
        return False
    return (q.dtype == torch.bfloat16 and q.shape[-1] == 128 and not window and dropout_p == 0.0 and cu_q is cu_k
            and q.shape[1] % k.shape[1] == 0)


def sm100_attention(q, k, v, cu_seqlens, max_seqlen, causal, scale):
    from .attention_sm100_fn import AttentionFn

    return AttentionFn.apply(q, k, v, cu_seqlens, max_seqlen, causal, scale)
