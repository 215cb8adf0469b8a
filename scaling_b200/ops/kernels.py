"""Thin tensor-level wrappers over the C ABI (no autograd here — see :mod:`scaling_b200.ops.functional`).

Every wrapper checks shapes/alignment, allocates outputs with torch (so memory stays in the caching allocator) and
launches on the current stream.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import DTYPE_CODE, call, ptr

_scratch: dict[tuple, torch.Tensor] = {}

# GEMM flavour: "2cta" = CTA pairs (cta_group::2, 256x256 tiles), "1cta" = single-CTA 128x256 tiles
_GEMM_IMPL = os.environ.get("SCALING_B200_GEMM", "2cta")


def set_gemm_impl(name: str) -> None:
    global _GEMM_IMPL
    assert name in ("1cta", "2cta")
    _GEMM_IMPL = name


_split_ws: dict[int, torch.Tensor] = {}


def _ensure_split_workspace(device: torch.device) -> None:
    """The 2-CTA GEMM's split-K tail parks fp32 partials in a workspace the CALLER owns (the library allocates nothing):
    one zero-initialised buffer per device from torch's allocator, registered once and kept alive here."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx in _split_ws:
        return
    lib = _lib.load(required=True)
    with torch.cuda.device(idx):
        n = int(lib.sb_gemm2_split_workspace_bytes())
        buf = torch.zeros(n, dtype=torch.uint8, device=torch.device("cuda", idx))
        torch.cuda.current_stream().synchronize()  # zeros must have landed before any stream's first GEMM uses it
        rc = lib.sb_gemm2_set_split_workspace(buf.data_ptr(), n)
    if rc != 0:
        raise RuntimeError(f"sb_gemm2_set_split_workspace failed with status {rc}")
    _split_ws[idx] = buf


def _gemm_entry(M: int, N: int, device: Optional[torch.device] = None) -> str:
    # tiny problems do not fill CTA pairs; keep them on the single-CTA kernel
    if _GEMM_IMPL == "2cta" and M >= 256 and N >= 128:
        if device is not None:
            _ensure_split_workspace(device)
        return "sb_gemm2_bf16"
    return "sb_gemm_bf16"


def _scratch_buf(device: torch.device, n_floats: int, tag: str) -> torch.Tensor:
    """Reusable fp32 scratch per (device, tag); grows geometrically, never shrinks."""
    key = (device.index, tag)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < n_floats:
        buf = torch.empty(max(n_floats, 1024), dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


# ----------------------------------------------------------------------------------------------------------- GEMM
def gemm_supported(M: int, N: int, K: int, lda: int, ldb: int, ldd: int) -> bool:
    return M > 0 and N > 0 and K > 0 and lda % 8 == 0 and ldb % 8 == 0 and ldd % 8 == 0


def gemm_f16_supported(M: int, N: int, a_mn: bool, b_mn: bool, out_dtype: torch.dtype = torch.float16,
                       accumulate: bool = False) -> bool:
    """fp16 operands: fprop (row-major x row-major), dgrad (b_mn) and wgrad (a_mn and b_mn) of a linear layer on the
    2-CTA kernel (M >= 256, N >= 128); fp32 output / accumulation for the wgrad only (main-grad arenas)."""
    if M < 256 or N < 128 or (a_mn and not b_mn):
        return False
    if out_dtype == torch.float32:
        return a_mn and b_mn
    return not accumulate


def gemm(
    a: torch.Tensor,
    b: torch.Tensor,
    *,
    a_mn: bool = False,
    b_mn: bool = False,
    out: Optional[torch.Tensor] = None,
    out_dtype: Optional[torch.dtype] = None,
    accumulate: bool = False,
    bias: Optional[torch.Tensor] = None,
    max_ctas: int = 0,
) -> torch.Tensor:
    """D[M,N] (+)= A · Bᵀ on the tcgen05 kernel.

    a: [M,K] row-major, or [K,M] when ``a_mn``;  b: [N,K] row-major, or [K,N] when ``b_mn``.  Rows may be strided
    (``stride(0)`` is the leading dimension) but the inner dimension must be contiguous.  bf16 operands take every
    layout; fp16 operands the three layouts of a linear layer (see :func:`gemm_f16_supported`).  The result is in the
    operands' dtype unless ``out`` / ``out_dtype`` say fp32.
    """
    assert a.dtype == b.dtype and a.dtype in (torch.bfloat16, torch.float16) and a.dim() == 2 and b.dim() == 2
    half = a.dtype == torch.float16
    if out_dtype is None:
        out_dtype = a.dtype
    assert a.stride(1) == 1 and b.stride(1) == 1
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"inner dimensions differ: {K} vs {Kb}"
    if out is None:
        assert not accumulate
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    assert out.shape == (M, N) and out.stride(1) == 1 and out.dtype in (a.dtype, torch.float32)
    assert a.data_ptr() % 16 == 0 and b.data_ptr() % 16 == 0 and out.data_ptr() % 16 == 0
    lda, ldb, ldd = a.stride(0), b.stride(0), out.stride(0)
    assert gemm_supported(M, N, K, lda, ldb, ldd), "leading dimensions must be multiples of 8 elements"
    if bias is not None:
        assert bias.dtype == a.dtype and bias.numel() == N and bias.is_contiguous()
    try:
        if half:
            assert gemm_f16_supported(M, N, a_mn, b_mn, out.dtype, accumulate), \
                "fp16 runs the linear-layer layouts of the 2-CTA kernel only"
            _ensure_split_workspace(a.device)
            entry = "sb_gemm2_f16"
        else:
            entry = _gemm_entry(M, N, a.device)
        if not a_mn:
            _apply_gate(a, entry)
        call(
            entry, ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, lda, ldb, ldd, int(a_mn), int(b_mn),
            int(out.dtype == torch.float32), int(accumulate), max_ctas,
        )
    except RuntimeError as e:
        raise RuntimeError(
            f"{e}: M={M} N={N} K={K} a={tuple(a.shape)}/{a.stride()}@{a.data_ptr():#x} "
            f"b={tuple(b.shape)}/{b.stride()}@{b.data_ptr():#x} out={tuple(out.shape)}/{out.stride()}@{out.data_ptr():#x} "
            f"a_mn={a_mn} b_mn={b_mn} acc={accumulate}") from e
    return out


def gemm_lowrank(a: torch.Tensor, b: torch.Tensor, a2: torch.Tensor, b2: torch.Tensor, *, b_mn: bool = False,
                 comm=None, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """D[M,N] = A·Bᵀ + A2·B2ᵀ in one launch: the low-rank pair (A2 [M,r]; B2 [N,r], or [r,N] when ``b_mn``) rides as
    extra k-blocks behind the main K range of the 2-CTA tcgen05 GEMM.  With ``comm`` the result is additionally summed
    over the tensor-parallel group inside the same launch (fused all-reduce)."""
    assert all(t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1 for t in (a, b, a2, b2))
    M, K = a.shape
    N = b.shape[1] if b_mn else b.shape[0]
    assert (b.shape[0] if b_mn else b.shape[1]) == K
    r = a2.shape[1]
    assert a2.shape[0] == M and (b2.shape == (r, N) if b_mn else b2.shape == (N, r)), (a2.shape, b2.shape, N, r)
    assert M >= 256 and N >= 128 and r % 8 == 0, "the low-rank slab exists on the 2-CTA kernel only"
    out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    lds = (a.stride(0), b.stride(0), a2.stride(0), b2.stride(0), out.stride(0))
    assert all(x % 8 == 0 for x in lds) and all(t.data_ptr() % 16 == 0 for t in (a, b, a2, b2, out))
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
    _ensure_split_workspace(a.device)
    if comm is None:
        call("sb_gemm2_lowrank_bf16", ptr(a), ptr(b), ptr(a2), ptr(b2), ptr(out), ptr(bias), M, N, K, r, lds[0], lds[1],
             lds[2], lds[3], lds[4], int(b_mn), 0, 0, 0)
    else:
        comm.reserve(M, N)
        epoch, data_ptrs, flag_ptrs = comm.next_call()
        call("sb_gemm2_allreduce_lowrank_bf16", ptr(a), ptr(b), ptr(a2), ptr(b2), ptr(out), ptr(bias), M, N, K, r, lds[0],
             lds[1], lds[2], lds[3], lds[4], int(b_mn), data_ptrs, flag_ptrs, comm.rank, comm.size, epoch, comm.slot_bytes,
             comm.flag_slot, 0)
    return out


def gemm_allreduce(a: torch.Tensor, b: torch.Tensor, comm, *, b_mn: bool = False, bias: Optional[torch.Tensor] = None,
                   out: Optional[torch.Tensor] = None, max_ctas: int = 0) -> torch.Tensor:
    """D[M,N] = Σ_ranks A_r · B_rᵀ (+ bias) in ONE launch: the tcgen05 GEMM pushes its partial tiles over NVLink into
    every TP rank's symmetric workspace while later tiles are still in the tensor cores, and reducer warps of the
    same kernel sum them (``comm`` is a :class:`scaling_b200.parallel.TPComm`).  Bit-identical on all ranks."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"inner dimensions differ: {K} vs {Kb}"
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    assert out.shape == (M, N) and out.stride(1) == 1 and out.dtype == torch.bfloat16
    lda, ldb, ldd = a.stride(0), b.stride(0), out.stride(0)
    assert gemm_supported(M, N, K, lda, ldb, ldd), "leading dimensions must be multiples of 8 elements"
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
    comm.reserve(M, N)
    epoch, data_ptrs, flag_ptrs = comm.next_call()
    call(
        "sb_gemm2_allreduce_bf16", ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, lda, ldb, ldd, 0, int(b_mn),
        data_ptrs, flag_ptrs, comm.rank, comm.size, epoch, comm.slot_bytes, comm.flag_slot, max_ctas,
    )
    return out


def reduce_scatter_supported(M: int, seq: int, tp: int) -> bool:
    """Sequence slices must be whole 256-row tiles."""
    return seq > 0 and M % seq == 0 and seq % tp == 0 and (seq // tp) % 256 == 0


def gemm_reduce_scatter(a: torch.Tensor, b: torch.Tensor, comm, seq: int, *, b_mn: bool = False,
                        bias: Optional[torch.Tensor] = None, max_ctas: int = 0) -> torch.Tensor:
    """D[M/tp, N] = this rank's sequence slice of Σ_ranks A_r · B_rᵀ (+ bias) in ONE launch (rows of A are [batch, seq];
    rank r keeps positions [r·seq/tp, (r+1)·seq/tp) of every batch item).  Partial tiles travel over NVLink to their
    owner only — half the bytes of the fused all-reduce at tp = 2 — while later tiles are in the tensor cores."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    M, K = a.shape
    N = b.shape[1] if b_mn else b.shape[0]
    assert (b.shape[0] if b_mn else b.shape[1]) == K
    assert reduce_scatter_supported(M, seq, comm.size), (M, seq, comm.size)
    out = torch.empty((M // comm.size, N), dtype=torch.bfloat16, device=a.device)
    lda, ldb, ldd = a.stride(0), b.stride(0), out.stride(0)
    assert gemm_supported(M, N, K, lda, ldb, ldd)
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
    comm.reserve(M, N)
    epoch, data_ptrs, flag_ptrs = comm.next_call()
    call(
        "sb_gemm2_reduce_scatter_bf16", ptr(a), ptr(b), ptr(out), ptr(bias), M, N, K, lda, ldb, ldd, 0, int(b_mn),
        data_ptrs, flag_ptrs, comm.rank, comm.size, epoch, comm.slot_bytes, comm.flag_slot, seq, max_ctas,
    )
    return out


def gemm_swiglu(x: torch.Tensor, w_gate_up: torch.Tensor, keep_pre: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """h = silu(x·W1ᵀ) ⊙ (x·W2ᵀ) with ``w_gate_up`` = [W1; W2] ([2F, K]); returns (h [M,F], pre [M,2F] or None)."""
    assert x.dtype == torch.bfloat16 and w_gate_up.dtype == torch.bfloat16
    M, K = x.shape
    F2, K2 = w_gate_up.shape
    assert K == K2 and F2 % 2 == 0 and x.stride(1) == 1 and w_gate_up.stride(1) == 1
    F = F2 // 2
    assert F % 8 == 0 and x.stride(0) % 8 == 0 and w_gate_up.stride(0) % 8 == 0
    h = torch.empty((M, F), dtype=torch.bfloat16, device=x.device)
    pre = torch.empty((M, 2 * F), dtype=torch.bfloat16, device=x.device) if keep_pre else None
    if _GEMM_IMPL == "2cta" and M >= 256:
        _ensure_split_workspace(x.device)
        _apply_gate(x, "sb_gemm2_swiglu_bf16")
    call(
        "sb_gemm2_swiglu_bf16" if (_GEMM_IMPL == "2cta" and M >= 256) else "sb_gemm_swiglu_bf16", ptr(x), ptr(w_gate_up), ptr(h), ptr(pre), M, F, K, x.stride(0), w_gate_up.stride(0), F,
        2 * F, 0,
    )
    return h, pre


def gemm_rope_supported(M: int, N: int, d: int, rot_dim: int, mode: int) -> bool:
    return _GEMM_IMPL == "2cta" and mode == 0 and M >= 256 and N >= 128 and d % 32 == 0 and rot_dim % 8 == 0 and 0 < rot_dim <= d


def gemm_rope(x: torch.Tensor, w: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: Optional[torch.Tensor], seq: int,
              rope_cols: int, d: int, rot_dim: int) -> torch.Tensor:
    """Y = x·Wᵀ with the rotary embedding (interleaved pairs) applied to the first ``rope_cols`` columns (q and k heads of
    ``d`` columns each) in the GEMM epilogue."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 2 and w.dim() == 2
    M, Kd = x.shape
    N = w.shape[0]
    assert w.shape[1] == Kd and x.stride(1) == 1 and w.stride(1) == 1 and gemm_supported(M, N, Kd, x.stride(0), w.stride(0), N)
    assert cos.dtype == torch.float32 and sin.dtype == torch.float32 and cos.is_contiguous() and sin.is_contiguous()
    assert pos is None or (pos.dtype == torch.int32 and pos.is_contiguous() and pos.numel() == M)
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    _ensure_split_workspace(x.device)
    _apply_gate(x, "sb_gemm2_rope_bf16")
    call("sb_gemm2_rope_bf16", ptr(x), ptr(w), ptr(y), M, N, Kd, x.stride(0), w.stride(0), N, ptr(cos), ptr(sin), ptr(pos),
         seq, rope_cols, d, rot_dim, 0)
    return y


def gemm_act(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], silu: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """(act(x·Wᵀ + bias), pre-activation) in ONE launch of the 2-CTA GEMM (activation in the TMEM epilogue)."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 2 and w.dim() == 2
    M, Kd = x.shape
    N = w.shape[0]
    assert w.shape[1] == Kd and M >= 256 and N >= 128 and gemm_supported(M, N, Kd, x.stride(0), w.stride(0), N)
    assert x.stride(1) == 1 and w.stride(1) == 1
    if bias is not None:
        assert bias.dtype == torch.bfloat16 and bias.numel() == N and bias.is_contiguous()
    h = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    pre = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    _ensure_split_workspace(x.device)
    call("sb_gemm2_act_bf16", ptr(x), ptr(w), ptr(bias), ptr(h), ptr(pre), M, N, Kd, x.stride(0), w.stride(0), N, N,
         int(silu), 0)
    return h, pre


# ----------------------------------------------------------------------------------------------------------- norms
def norm_fwd(
    x: torch.Tensor,
    weight: torch.Tensor,
    bias: Optional[torch.Tensor],
    eps: float,
    rms: bool,
    residual: Optional[torch.Tensor] = None,
) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor], torch.Tensor]:
    """Returns (y, x_plus_residual or None, mean or None, rstd). x: [..., cols] contiguous."""
    cols = x.shape[-1]
    rows = x.numel() // cols
    assert x.is_contiguous() and cols % 8 == 0 and weight.dtype == x.dtype
    y = torch.empty_like(x)
    summed = torch.empty_like(x) if residual is not None else None
    mean = None if rms else torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    if residual is not None:
        assert residual.is_contiguous() and residual.shape == x.shape
    call(
        "sb_norm_fwd", ptr(x), ptr(residual), ptr(weight), ptr(bias), ptr(y), ptr(summed), ptr(mean), ptr(rstd), rows,
        cols, float(eps), int(rms), DTYPE_CODE[x.dtype],
    )
    return y, summed, mean, rstd


def norm_bwd(
    dy: torch.Tensor,
    x: torch.Tensor,
    weight: torch.Tensor,
    mean: Optional[torch.Tensor],
    rstd: torch.Tensor,
    rms: bool,
    dres: Optional[torch.Tensor] = None,
    dw_out: Optional[torch.Tensor] = None,
    db_out: Optional[torch.Tensor] = None,
    accumulate: bool = False,
    has_bias: bool = False,
) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    """Returns (dx [+ dres], dw, db).  dw_out/db_out (+accumulate) write straight into a main-grad buffer."""
    cols = x.shape[-1]
    rows = x.numel() // cols
    assert dy.is_contiguous() and x.is_contiguous()
    dx = torch.empty_like(x)
    dw = dw_out if dw_out is not None else torch.empty(cols, dtype=x.dtype, device=x.device)
    db = db_out if db_out is not None else (torch.empty(cols, dtype=x.dtype, device=x.device) if has_bias else None)
    grad_fp32 = int(dw.dtype == torch.float32)
    lib = _lib.load(required=True)
    n = lib.sb_norm_bwd_scratch_floats(rows, cols, int(rms))
    scratch = _scratch_buf(x.device, n, "norm_bwd")
    call(
        "sb_norm_bwd", ptr(dy), ptr(x), ptr(weight), ptr(mean), ptr(rstd), ptr(dres), ptr(dx), ptr(dw), ptr(db),
        grad_fp32, int(accumulate), ptr(scratch), rows, cols, int(rms), DTYPE_CODE[x.dtype],
        launches=2 if (rms or db is None) else 3,
    )
    return dx, dw, db


# ----------------------------------------------------------------------------------------------------------- element-wise
def swiglu_fwd(pre: torch.Tensor) -> torch.Tensor:
    rows, F2 = pre.shape
    h = torch.empty((rows, F2 // 2), dtype=pre.dtype, device=pre.device)
    call("sb_swiglu_fwd", ptr(pre), ptr(h), rows, F2 // 2)
    return h


def swiglu_bwd(pre: torch.Tensor, dh: torch.Tensor, recompute_h: bool = False) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    rows, F2 = pre.shape
    assert pre.is_contiguous() and dh.is_contiguous()
    dpre = torch.empty_like(pre)
    h = torch.empty_like(dh) if recompute_h else None
    call("sb_swiglu_bwd", ptr(pre), ptr(dh), ptr(dpre), ptr(h), rows, F2 // 2)
    return dpre, h


def act_fwd(x: torch.Tensor, silu: bool) -> torch.Tensor:
    y = torch.empty_like(x)
    call("sb_act_fwd", ptr(x), ptr(y), x.numel(), int(silu))
    return y


def act_bwd(x: torch.Tensor, dy: torch.Tensor, silu: bool) -> torch.Tensor:
    dx = torch.empty_like(x)
    call("sb_act_bwd", ptr(x), ptr(dy), ptr(dx), x.numel(), int(silu))
    return dx


def rope_(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: Optional[torch.Tensor], seq: int, mode: int,
          rot_dim: int, backward: bool = False) -> torch.Tensor:
    """In-place rotary on x [tokens, heads, head_dim] (token stride may exceed heads*head_dim). mode 0 = interleaved
    pairs (rotary_complex), 1 = half rotation (rotary)."""
    assert x.dtype == torch.bfloat16 and x.dim() == 3 and x.stride(2) == 1 and x.stride(1) == x.shape[2]
    assert cos.dtype == torch.float32 and sin.dtype == torch.float32 and cos.is_contiguous() and sin.is_contiguous()
    if pos is not None:
        assert pos.dtype == torch.int32 and pos.is_contiguous()
    tokens, heads, hd = x.shape
    call("sb_rope", ptr(x), ptr(cos), ptr(sin), ptr(pos), tokens, heads, hd, rot_dim, x.stride(0), seq, mode,
         int(backward))
    return x


def qkv_rope(ins, outs, heads, rotate, cos: torch.Tensor, sin: torch.Tensor, pos: Optional[torch.Tensor], seq: int,
             mode: int, rot_dim: int, backward: bool = False) -> None:
    """One pass over up to three [tokens, heads_i, d] segments (q, k, v): copy ``ins[i]`` to ``outs[i]`` (either may be
    a strided slice of a fused buffer: unit stride inside a token, any token pitch), rotating the segments flagged in
    ``rotate`` (inverse rotation when ``backward``)."""
    import ctypes

    n = len(ins)
    tokens, _, d = ins[0].shape
    for t in list(ins) + list(outs):
        assert t.dtype == torch.bfloat16 and t.dim() == 3 and t.shape[0] == tokens and t.shape[2] == d
        assert t.stride(2) == 1 and t.stride(1) == d
    if pos is not None:
        assert pos.dtype == torch.int32 and pos.is_contiguous()
    U, L, I = ctypes.c_ulonglong * n, ctypes.c_longlong * n, ctypes.c_int * n
    call("sb_qkv_rope", U(*[t.data_ptr() for t in ins]), U(*[t.data_ptr() for t in outs]), L(*[t.stride(0) for t in ins]),
         L(*[t.stride(0) for t in outs]), I(*[int(h) for h in heads]), I(*[int(bool(r)) for r in rotate]), n, ptr(cos),
         ptr(sin), ptr(pos), tokens, d, rot_dim, seq, mode, int(backward))


def bias_dropout_add(x: torch.Tensor, bias: Optional[torch.Tensor], residual: Optional[torch.Tensor], p: float,
                     seed: int, offset: int) -> torch.Tensor:
    assert x.is_contiguous() and x.dtype == torch.bfloat16
    out = torch.empty_like(x)
    call("sb_bias_dropout_add", ptr(x), ptr(bias), ptr(residual), ptr(out), x.numel(), x.shape[-1], float(p), seed,
         offset)
    return out


def dropout_bwd(dout: torch.Tensor, p: float, seed: int, offset: int) -> torch.Tensor:
    dx = torch.empty_like(dout)
    call("sb_dropout_bwd", ptr(dout), ptr(dx), dout.numel(), float(p), seed, offset)
    return dx


def embedding_fwd(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, vocab_end: int) -> torch.Tensor:
    assert ids.dtype == torch.int64 and ids.is_contiguous() and weight.dtype == torch.bfloat16
    out = torch.empty((*ids.shape, weight.shape[1]), dtype=weight.dtype, device=weight.device)
    call("sb_embedding_fwd", ptr(ids), ptr(weight), ptr(out), ids.numel(), weight.shape[1], vocab_start, vocab_end)
    return out


_EMBEDDING_BWD = os.environ.get("SCALING_B200_EMBEDDING_BWD", "sorted")


def embedding_bwd(ids: torch.Tensor, dout: torch.Tensor, dweight: torch.Tensor, vocab_start: int, vocab_end: int) -> None:
    """dweight (fp32 or bf16, [V_local, hidden]) += scatter(dout) — accumulates in place.  Deterministic: the ids are
    sorted (stable) and every run of equal ids is summed in token order by one block; ``SCALING_B200_EMBEDDING_BWD=atomic``
    selects the atomicAdd scatter (faster for tiny batches, summation order varies from run to run)."""
    assert dout.is_contiguous() and dweight.is_contiguous()
    if _EMBEDDING_BWD == "atomic":
        call("sb_embedding_bwd", ptr(ids), ptr(dout), ptr(dweight), int(dweight.dtype == torch.float32), ids.numel(),
             dweight.shape[1], vocab_start, vocab_end)
        return
    sorted_ids, perm = torch.sort(ids.reshape(-1).long(), stable=True)
    call("sb_embedding_bwd_sorted", ptr(sorted_ids), ptr(perm), ptr(dout), ptr(dweight), int(dweight.dtype == torch.float32),
         ids.numel(), dweight.shape[1], vocab_start, vocab_end)


# ----------------------------------------------------------------------------------------------------------- sequence parallel
def seq_peer_ok(x: torch.Tensor, comm) -> bool:
    return (comm is not None and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 3 and x.shape[-1] % 8 == 0
            and x.numel() > 0)


def seq_all_gather(x: torch.Tensor, comm, clone: bool = True) -> torch.Tensor:
    """[b, s/tp, h] -> [b, s, h]: every rank stores its shard into every rank's exchange buffer over NVLink
    (``sb_seq_gather_push``), one device barrier, and the result is copied out of the (double-buffered) exchange buffer.
    ``clone=False`` hands out the exchange buffer itself: valid until the second next gather, for immediate consumers."""
    b, sl, h = x.shape
    x = x.contiguous()
    buf, ptrs = comm.exchange_buffer("seq_gather", b * sl * comm.size * h * 2)
    call("sb_seq_gather_push", ptr(x), ptrs, comm.size, comm.rank, b, sl, h)
    comm.barrier()
    out = buf.view(torch.bfloat16).view(b, sl * comm.size, h)
    return out.clone() if clone else out


_gates: dict[int, tuple] = {}  # data_ptr of a gated gather buffer -> (counter address, target)


def seq_gather_gated_supported(x: torch.Tensor, comm) -> bool:
    return (_GEMM_IMPL == "2cta" and seq_peer_ok(x, comm) and x.shape[1] % 256 == 0
            and reduce_scatter_supported(x.shape[0] * x.shape[1] * comm.size, x.shape[1] * comm.size, comm.size))


def seq_gather_gated(x: torch.Tensor, comm) -> torch.Tensor:
    """all-gather -> GEMM, the gather half: every rank stores its [b, s/tp, h] shard into every rank's
    gather buffer chunk by chunk and bumps a per-row-tile arrival counter (``sb_seq_gather_push_counted``).  NO barrier:
    the returned [b, s, h] view may ONLY be consumed by :func:`gemm` / :func:`gemm_swiglu` as the A operand — their TMA
    producer waits per 256-row tile for the counter, so tiles of the local shard start immediately and remote tiles as
    they land.  The buffer is reused by the second next gated gather of the same shape."""
    b, sl, h = x.shape
    tp = comm.size
    x = x.contiguous()
    buf, ptrs, gate_local, gate_ptrs, uses = comm.exchange_buffer_gated(f"sp_gate_{b}_{sl}_{h}", b * sl * tp * h * 2)
    call("sb_seq_gather_push_counted", ptr(x), ptrs, gate_ptrs, tp, comm.rank, b, sl, h)
    lib = _lib.load(required=True)
    out = buf.view(torch.bfloat16).view(b, sl * tp, h)
    _gates[out.data_ptr()] = (gate_local, (uses * int(lib.sb_seq_gather_chunks_per_tile())) & 0xFFFFFFFF)
    return out


def _apply_gate(a: torch.Tensor, entry: str) -> None:
    g = _gates.get(a.data_ptr())
    if g is not None:
        assert entry.startswith("sb_gemm2"), "a gated gather buffer can only feed the 2-CTA GEMM kernels"
        rc = _lib.load(required=True).sb_gemm2_set_gate(g[0], g[1])
        assert rc == 0


def seq_reduce_scatter(x: torch.Tensor, comm) -> torch.Tensor:
    """[b, s, h] partial sums -> [b, s/tp, h] summed over the group: slice j goes to rank j's slot buffer
    (``sb_seq_scatter_push``), one device barrier, then the tp slots are added in rank order (``sb_sum_slots``)."""
    b, s, h = x.shape
    tp = comm.size
    assert s % tp == 0
    sl = s // tp
    x = x.contiguous()
    buf, ptrs = comm.exchange_buffer("seq_scatter", tp * b * sl * h * 2)
    call("sb_seq_scatter_push", ptr(x), ptrs, tp, comm.rank, b, sl, h)
    comm.barrier()
    out = torch.empty((b, sl, h), dtype=torch.bfloat16, device=x.device)
    call("sb_sum_slots", ptr(buf), ptr(out), tp, b * sl * h)
    return out


# ----------------------------------------------------------------------------------------------------------- cross-entropy
def gemm_xent_stats(x: torch.Tensor, w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """LM-head GEMM with the cross-entropy statistics in its epilogue: (logits [M, N] bf16, partial [M, ceil(N/256), 4]
    fp32 = per 256-column tile (max, Σexp(x − max), unused, local arg-max column as int bits))."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dim() == 2 and w.dim() == 2
    assert x.stride(1) == 1 and w.stride(1) == 1 and x.shape[1] == w.shape[1]
    M, Kd = x.shape
    N = w.shape[0]
    assert M >= 256 and N >= 128 and gemm_supported(M, N, Kd, x.stride(0), w.stride(0), N)
    lib = _lib.load(required=True)
    logits = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    partial = torch.empty((M, int(lib.sb_gemm2_xent_tiles(N)), 4), dtype=torch.float32, device=x.device)
    _ensure_split_workspace(x.device)
    call("sb_gemm2_xent_stats_bf16", ptr(x), ptr(w), ptr(logits), ptr(partial), M, N, Kd, x.stride(0), w.stride(0), N, 0)
    return logits, partial


def xent_from_partials(partial: torch.Tensor, logits: torch.Tensor, targets: torch.Tensor, vocab_start: int,
                       comm=None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """(loss, lse, argmax) per token from the epilogue statistics: tiles are merged and the target logit picked by one
    small kernel that stores the per-token result into every TP rank's exchange buffer (16 bytes per token over
    NVLink), a device barrier, and a finalize kernel — no NCCL, no torch glue."""
    tokens, tiles, _ = partial.shape
    tp = 1 if comm is None else comm.size
    rank = 0 if comm is None else comm.rank
    dev = partial.device
    if comm is None:
        import ctypes

        allstats = torch.empty((1, tokens, 4), dtype=torch.float32, device=dev)
        ptrs = (ctypes.c_ulonglong * 1)(allstats.data_ptr())
    else:
        buf, ptrs = comm.exchange_buffer("xent", tp * tokens * 16)
        allstats = buf.view(torch.float32).view(tp, tokens, 4)
    call("sb_xent_combine", ptr(partial), tiles, ptr(logits), logits.stride(0), ptr(targets), tokens, logits.shape[1],
         vocab_start, ptrs, tp, rank)
    if comm is not None:
        comm.barrier()
    loss = torch.empty(tokens, dtype=torch.float32, device=dev)
    lse = torch.empty(tokens, dtype=torch.float32, device=dev)
    argmax = torch.empty(tokens, dtype=torch.int64, device=dev)
    call("sb_xent_finalize", ptr(allstats), tp, tokens, ptr(loss), ptr(lse), ptr(argmax))
    return loss, lse, argmax



def xent_stats(logits: torch.Tensor, targets: torch.Tensor, vocab_start: int) -> torch.Tensor:
    """[tokens, 4] fp32: local max, local Σexp(x-max), target logit (0 if remote), local argmax (int bits)."""
    tokens, v = logits.shape
    assert logits.stride(1) == 1 and targets.dtype == torch.int64 and targets.is_contiguous()
    stats = torch.empty((tokens, 4), dtype=torch.float32, device=logits.device)
    call("sb_xent_stats", ptr(logits), ptr(targets), ptr(stats), tokens, v, logits.stride(0), vocab_start,
         DTYPE_CODE[logits.dtype])
    return stats


def xent_grad_(logits: torch.Tensor, targets: torch.Tensor, lse: torch.Tensor, coeff: torch.Tensor,
               gscale: Optional[torch.Tensor], vocab_start: int) -> torch.Tensor:
    """Overwrites logits with (softmax - onehot) * coeff[token] * gscale."""
    tokens, v = logits.shape
    assert lse.dtype == torch.float32 and coeff.dtype == torch.float32
    call("sb_xent_grad", ptr(logits), ptr(targets), ptr(lse), ptr(coeff), ptr(gscale), tokens, v, logits.stride(0),
         vocab_start, DTYPE_CODE[logits.dtype])
    return logits


# ----------------------------------------------------------------------------------------------------------- optimizer
def grad_sumsq(g: torch.Tensor, out: torch.Tensor, nonfinite: torch.Tensor, accumulate: bool = False) -> None:
    lib = _lib.load(required=True)
    scratch = _scratch_buf(g.device, lib.sb_grad_sumsq_scratch_floats(), "sumsq")
    call("sb_grad_sumsq", ptr(g), g.numel(), DTYPE_CODE[g.dtype], ptr(scratch), ptr(out), int(accumulate),
         ptr(nonfinite), launches=2)


def adamw_step(p32, m, v, g, out, *, lr, beta1, beta2, eps, weight_decay, step, inv_loss_scale=1.0, clip=0.0,
               norm_sq=None, skip=None, zero_grad=False) -> None:
    n = p32.numel()
    assert m.numel() == n and v.numel() == n and g.numel() == n and (out is None or out.numel() == n)
    call("sb_adamw", ptr(p32), ptr(m), ptr(v), ptr(g), ptr(out), n, DTYPE_CODE[g.dtype],
         DTYPE_CODE[out.dtype] if out is not None else 0, float(lr), float(beta1), float(beta2), float(eps),
         float(weight_decay), int(step), float(inv_loss_scale), float(clip), ptr(norm_sq), ptr(skip), int(zero_grad))
