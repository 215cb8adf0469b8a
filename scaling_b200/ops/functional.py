"""Autograd-aware operators.  Each op has exactly one native path (the sm_100a kernels) plus a shape/device guard that
routes tiny, unaligned or CPU tensors to the plain PyTorch expression of the same math (linears of
1/2/8/17 features and CPU/gloo runs are first class in the tests) — the guard is part of the op, not a second backend.

Main-grad protocol: when a weight carries ``.main_grad`` (an fp32 or bf16 view into the optimizer's flat gradient
arena) the weight-gradient GEMM accumulates straight into it from TMEM (``D += dYᵀ·X``) and autograd receives no
weight gradient — no per-micro-batch ``.grad`` tensors, no separate accumulation kernels, no fp32 copy at step time.
"""
from __future__ import annotations

import os
import weakref
from typing import Any, Optional, Tuple

import torch
import torch.nn.functional as F

from . import _lib
from . import kernels as K

_FORCE_TORCH = os.environ.get("SCALING_B200_OPS", "sm100") == "torch"
# plain (epilogue-free) GEMMs may be routed to cuBLAS: "native" | "cublas"
_PLAIN_GEMM = os.environ.get("SCALING_B200_PLAIN_GEMM", "native")


def set_ops(mode: str) -> None:
    """``"sm100"`` (default on GPU) or ``"torch"`` (oracle / baseline-inside-the-product)."""
    global _FORCE_TORCH
    assert mode in ("sm100", "torch")
    _FORCE_TORCH = mode == "torch"


def set_plain_gemm(mode: str) -> None:
    global _PLAIN_GEMM
    assert mode in ("native", "cublas")
    _PLAIN_GEMM = mode


def native_enabled(x: torch.Tensor) -> bool:
    return (not _FORCE_TORCH) and x.is_cuda and _lib.available()


def _aligned(*dims: int) -> bool:
    return all(d % 8 == 0 and d > 0 for d in dims)


def _own_gemm(t: torch.Tensor, m: int, n: int) -> bool:
    """bf16 runs every shape on the own kernels; fp16 exists on the 2-CTA kernel (tiles of 256 x 256) only."""
    return _PLAIN_GEMM != "cublas" and (t.dtype == torch.bfloat16 or (m >= 256 and n >= 128))


def _gemm_nt(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not _own_gemm(x, x.shape[0], w.shape[0]):
        return F.linear(x, w, bias)
    return K.gemm(x, w, bias=bias)


def _gemm_nn(dy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    if not _own_gemm(dy, dy.shape[0], w.shape[1]):
        return dy @ w
    return K.gemm(dy, w, b_mn=True)


# ---- tensor-parallel reductions fused into the GEMM (scaling_b200.parallel.TPComm) --------------------------------
def _fusable(m: int, n: int) -> bool:
    return m >= 256 and n >= 128


def _gemm_nt_reduce(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], comm) -> torch.Tensor:
    """Σ_tp x_r·w_rᵀ (+ bias once): one fused launch, or GEMM + NCCL all-reduce for shapes below the 2-CTA tile."""
    if _fusable(x.shape[0], w.shape[0]) and x.dtype == torch.bfloat16:
        return K.gemm_allreduce(x, w, comm, bias=bias)
    y = _gemm_nt(x, w)
    comm.all_reduce_(y)
    return y if bias is None else y.add_(bias)


def _gemm_nn_reduce(dy: torch.Tensor, w: torch.Tensor, comm) -> torch.Tensor:
    if _fusable(dy.shape[0], w.shape[1]) and dy.dtype == torch.bfloat16:
        return K.gemm_allreduce(dy, w, comm, b_mn=True)
    return comm.all_reduce_(_gemm_nn(dy, w))


class _CopyToRegion(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, comm):
        ctx.comm = comm
        return x

    @staticmethod
    def backward(ctx, g):
        return ctx.comm.all_reduce_(g.contiguous()), None


class _ReduceFromRegion(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, comm):
        return comm.all_reduce_(x.contiguous().clone())

    @staticmethod
    def backward(ctx, g):
        return g, None


def _accumulate_wgrad(dy: torch.Tensor, x: torch.Tensor, weight: torch.Tensor) -> Optional[torch.Tensor]:
    """dW = dyᵀ·x. Into ``weight.main_grad`` (returns None) or as a fresh tensor."""
    main = getattr(weight, "main_grad", None)
    own = _own_gemm(dy, dy.shape[1], x.shape[1])
    if main is not None:
        if own or dy.dtype == torch.bfloat16:
            K.gemm(dy, x, a_mn=True, b_mn=True, out=main, accumulate=True)
        else:
            main.add_((dy.t() @ x).to(main.dtype))
        return None
    if not own:
        return dy.t() @ x
    return K.gemm(dy, x, a_mn=True, b_mn=True, out_dtype=weight.dtype)


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, comm=None, reduce_output=False, reduce_input_grad=False):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if comm is not None and reduce_output:
            y = _gemm_nt_reduce(x2, weight, bias, comm)
        else:
            y = _gemm_nt(x2, weight, bias)
        ctx.comm = comm if reduce_input_grad else None
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.x_shape = shape
        return y.view(*shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = (_gemm_nn_reduce(dy2, weight, ctx.comm) if ctx.comm is not None else _gemm_nn(dy2, weight))
            dx = dx.view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = _accumulate_wgrad(dy2, x2, weight)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = dy2.sum(0)
        return dx, dw, db, None, None, None


class _LinearReduceScatterFn(torch.autograd.Function):
    """Row-parallel projection under sequence parallelism: y[b, s/tp, N] = reduce-scatter_seq(x·Wᵀ) as ONE launch
    (reference: F.linear, then reduce_scatter along the sequence, attention.py:704-705 / mlp.py:164-165).  Backward
    all-gathers dY along the sequence (NCCL) and runs the ordinary dgrad / wgrad GEMMs."""

    @staticmethod
    def forward(ctx, x, weight, comm):
        b, s, k = x.shape
        x2 = x.reshape(b * s, k)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = K.gemm_reduce_scatter(x2, weight, comm, s)
        ctx.save_for_backward(x2, weight)
        ctx.comm, ctx.x_shape = comm, x.shape
        return y.view(b, s // comm.size, weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        from ..core.nn.tp_ops import _all_gather_seq

        x2, weight = ctx.saved_tensors
        dy_full = _all_gather_seq(dy.contiguous(), ctx.comm.size, ctx.comm.group)
        dy2 = dy_full.reshape(-1, dy_full.shape[-1])
        dx = _gemm_nn(dy2, weight).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _accumulate_wgrad(dy2, x2, weight) if ctx.needs_input_grad[1] else None
        return dx, dw, None


def linear_reduce_scatter(x: torch.Tensor, weight: torch.Tensor, comm) -> Optional[torch.Tensor]:
    """Fused row-parallel GEMM + sequence reduce-scatter, or None when the shapes do not fit the fused kernel (the
    caller then runs GEMM + NCCL reduce-scatter)."""
    if (comm is None or x.dim() != 3 or not native_enabled(x) or x.dtype != torch.bfloat16 or weight.dtype != torch.bfloat16
            or not _aligned(*weight.shape) or not K.reduce_scatter_supported(x.shape[0] * x.shape[1], x.shape[1], comm.size)
            or weight.shape[0] < 128):
        return None
    return _LinearReduceScatterFn.apply(x, weight, comm)


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, *, comm=None,
           reduce_output: bool = False, reduce_input_grad: bool = False) -> torch.Tensor:
    """y = x·Wᵀ (+ b) — replaces F.linear in Column/RowParallelLinear (reference column_parallel_linear.py:151).

    With a :class:`~scaling_b200.parallel.TPComm`: ``reduce_output`` sums y over the TP group (row-parallel forward;
    the bias is added once, after the sum) and ``reduce_input_grad`` sums dX over it (column-parallel backward) —
    each inside the GEMM launch that produces the operand, instead of an NCCL all-reduce behind it."""
    if (native_enabled(x) and x.dtype in (torch.bfloat16, torch.float16) and weight.dtype == x.dtype
            and _aligned(weight.shape[0], weight.shape[1]) and x.numel() > 0
            and (bias is None or bias.dtype == x.dtype)):
        return _LinearFn.apply(x, weight, bias, comm, reduce_output, reduce_input_grad)
    if comm is None:
        return F.linear(x, weight, bias)
    if reduce_input_grad:
        x = _CopyToRegion.apply(x, comm)
    if not reduce_output:
        return F.linear(x, weight, bias)
    y = _ReduceFromRegion.apply(F.linear(x, weight, None), comm)
    return y if bias is None else y + bias


# ------------------------------------------------------------------------------------------------ LoRA-fused linear
def _gather_dim(t: torch.Tensor, dim: int, tp) -> torch.Tensor:
    size, _, group = tp
    if size == 1:
        return t
    flat = torch.empty(size * t.numel(), dtype=t.dtype, device=t.device)  # flat: the layout every backend accepts
    torch.distributed.all_gather_into_tensor(flat, t.contiguous().view(-1), group=group)
    return torch.cat(list(flat.view(size, *t.shape).unbind(0)), dim=dim).contiguous()


_WEIGHT_EPOCH = 0
_ROW_LORA_CACHE: dict = {}


def bump_weight_epoch() -> None:
    """Called by whoever rewrites parameters behind autograd's back (the optimizer's arena kernels): invalidates
    per-step caches derived from weights."""
    global _WEIGHT_EPOCH
    _WEIGHT_EPOCH += 1


def _row_lora_operands(A: torch.Tensor, B: torch.Tensor, tp, kl: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Row-parallel LoRA needs this rank's COLUMNS of the full A and the full B; both are gathered over the TP group.
    The pair only changes at an optimizer step, so the gathers run once per step, not once per micro-batch."""
    size, rank, _ = tp
    if size == 1:
        return A[:, :kl], B
    key = (A.data_ptr(), B.data_ptr(), kl)
    stamp = (_WEIGHT_EPOCH, A._version, B._version)
    hit = _ROW_LORA_CACHE.get(key)
    # the entry belongs to these very tensor objects: another model allocated at the same addresses must not hit it
    if hit is not None and hit[0] == stamp and hit[3]() is A and hit[4]() is B:
        return hit[1], hit[2]
    with torch.no_grad():
        a_cols = _gather_dim(A, 0, tp)[:, rank * kl:(rank + 1) * kl].contiguous()
        b_full = _gather_dim(B, 1, tp)
    _ROW_LORA_CACHE[key] = (stamp, a_cols, b_full, weakref.ref(A), weakref.ref(B))
    return a_cols, b_full


class _LoraLinearFn(torch.autograd.Function):
    """y = x·Wᵀ + s·(x·Aᵀ)·Bᵀ with the low-rank pair riding as an extra K-slab of the base GEMM, forward AND backward
    (reference core/nn/lora.py:103-112 runs two extra linears and an add, and for the attention output
    projection all-gathers the whole [b, s, h] activation first, attention.py:690-701).

    ``kind == "column"`` (q / k / v): W [N/tp, K], A [r/tp, K] (rank dimension TP-split, as in the checkpoint), B [N/tp, r].
    Only the tiny [M, r] rank activations cross the TP group.  ``kind == "row"`` (attention output): W [N, K/tp],
    A [r/tp, K_full], B [N, r/tp]; each rank multiplies ITS input columns with the matching columns of the gathered A, so
    the low-rank term is a per-rank partial sum exactly like the base product and rides the same (fused) all-reduce —
    the activation is never gathered.  dX = dY·W + dT·A is again one launch with [dY | dT] x [W ; A].
    """

    @staticmethod
    def forward(ctx, x, W, A, B, scale, kind, tp, comm):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        size, rank, group = tp
        if kind == "column":
            t = _gather_dim(K.gemm(x2, A), 1, tp)  # [M, r]
            ts = t * scale
            y = K.gemm_lowrank(x2, W, ts, B)
            ctx.save_for_backward(x2, W, A, B, ts)
        else:
            a_cols, b_full = _row_lora_operands(A, B, tp, x2.shape[1])  # [r, K/tp], [N, r]
            ts = K.gemm(x2, a_cols) * scale  # this rank's partial of the rank activations
            if size > 1 and comm is not None:
                y = K.gemm_lowrank(x2, W, ts, b_full, comm=comm)
            else:
                y = K.gemm_lowrank(x2, W, ts, b_full)
                if size > 1:
                    torch.distributed.all_reduce(y, group=group)
            ctx.save_for_backward(x2, W, a_cols, b_full, ts)
        ctx.meta = (scale, kind, tp, comm, shape)
        return y.view(*shape[:-1], W.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, W, A_, B_, ts = ctx.saved_tensors
        scale, kind, tp, comm, shape = ctx.meta
        size, rank, group = tp
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dts = K.gemm(dy2, B_, b_mn=True)  # [M, r]: dY·B
        if kind == "column":
            if size > 1:  # every rank's output used the full rank activations: sum, keep the own rank slice
                torch.distributed.all_reduce(dts, group=group)
                rl = dts.shape[1] // size
                dts = dts[:, rank * rl:(rank + 1) * rl].contiguous()
            dtl = dts * scale
            dx = K.gemm_lowrank(dy2, W, dtl, A_, b_mn=True, comm=comm if size > 1 else None)
            if size > 1 and comm is None:
                torch.distributed.all_reduce(dx, group=group)
            dB = K.gemm(dy2, ts, a_mn=True, b_mn=True)  # [N/tp, r]
            dA = K.gemm(dtl, x2, a_mn=True, b_mn=True)  # [r/tp, K]
        else:
            dtl = dts * scale
            dx = K.gemm_lowrank(dy2, W, dtl, A_, b_mn=True)  # [M, K/tp], no reduction: columns are rank-local
            t_full = ts.clone()
            if size > 1:
                torch.distributed.all_reduce(t_full, group=group)
            dB_full = K.gemm(dy2, t_full, a_mn=True, b_mn=True)  # [N, r]
            rl = dB_full.shape[1] // size
            dB = dB_full[:, rank * rl:(rank + 1) * rl].contiguous()
            dA_cols = K.gemm(dtl, x2, a_mn=True, b_mn=True)  # [r, K/tp]
            dA = _gather_dim(dA_cols, 1, tp)[rank * rl:(rank + 1) * rl].contiguous()  # own rank rows, all columns
        dW = _accumulate_wgrad(dy2, x2, W) if ctx.needs_input_grad[1] else None
        return dx.view(shape), dW, dA, dB, None, None, None, None


_LORA_FUSED = os.environ.get("SCALING_B200_LORA_FUSED", "1") == "1"


def lora_fusion_enabled() -> bool:
    return _LORA_FUSED


def set_lora_fusion(enabled: bool) -> None:
    """A/B switch for tests and benchmarks: off = base GEMM plus separate LoRA linears (the reference's structure)."""
    global _LORA_FUSED
    _LORA_FUSED = bool(enabled)


def lora_linear(x: torch.Tensor, W: torch.Tensor, A: torch.Tensor, B: torch.Tensor, scale: float, kind: str,
                tp=(1, 0, None), comm=None) -> Optional[torch.Tensor]:
    """LoRA-fused projection (see :class:`_LoraLinearFn`), or None when the shapes do not fit the 2-CTA kernel's
    low-rank slab (the caller then adds the separate LoRA linears)."""
    m = x.numel() // x.shape[-1] if x.numel() else 0
    size = tp[0]
    r_full = A.shape[0] * size
    if not (native_enabled(x) and x.dtype == W.dtype == A.dtype == B.dtype == torch.bfloat16 and m >= 256
            and W.shape[0] >= 128 and W.shape[1] >= 128 and _aligned(*W.shape) and A.shape[0] % 8 == 0
            and r_full % 8 == 0 and _PLAIN_GEMM == "native"):
        return None
    return _LoraLinearFn.apply(x, W, A, B, float(scale), kind, tp, comm)


def qkv_lora_linear(x: torch.Tensor, weights: Tuple[torch.nn.Parameter, ...], loras: Tuple[Tuple[torch.Tensor, torch.Tensor], ...],
                    scale: float, tp=(1, 0, None), comm=None) -> Optional[torch.Tensor]:
    """q|k|v with a LoRA pair on every projection as ONE low-rank GEMM each way: [x | t]·[W_q;W_k;W_v | blockdiag(B)]ᵀ
    where t = x·[A_q;A_k;A_v]ᵀ.  Needs frozen base weights that are adjacent in memory (pack_frozen_parameters);
    then forward is one rank GEMM + one fused GEMM and backward ONE dX GEMM with ONE (fused) TP reduction instead of
    three of each, at the price of the block-diagonal's zero blocks (3r of K + 3r columns).  Returns [..., ΣN] or None.

    ``loras[i] = (A_i [r/tp, K], B_i [N_i(/tp), r])``; the gathered rank activations are rank-major
    ``[rank0: q k v | rank1: q k v | …]``, so block i of B is spread over the matching columns of every rank's group."""
    size = tp[0]
    if any(w.requires_grad for w in weights) or len({a.shape for a, _ in loras}) != 1:
        return None
    fw = fused_view(tuple(w.data for w in weights))
    m = x.numel() // x.shape[-1] if x.numel() else 0
    rl = loras[0][0].shape[0]  # r / tp
    n = len(loras)
    if fw is None or not (native_enabled(x) and x.dtype == fw.dtype == torch.bfloat16 and m >= 256 and fw.shape[0] >= 128
                          and _aligned(*fw.shape) and rl % 8 == 0 and _PLAIN_GEMM == "native"
                          and all(a.dtype == b.dtype == torch.bfloat16 and b.shape[1] == rl * size for a, b in loras)):
        return None
    a_cat = torch.cat([a for a, _ in loras], dim=0)  # [n·r/tp, K]
    blocks = []
    for i, (_, b) in enumerate(loras):
        own = b.reshape(b.shape[0], size, 1, rl)
        zero = own.new_zeros(own.shape)
        blocks.append(torch.cat([own if j == i else zero for j in range(n)], dim=2))  # [N_i, tp, n, r/tp]
    b_bd = torch.cat(blocks, dim=0).reshape(fw.shape[0], size * n * rl)
    return _LoraLinearFn.apply(x, fw, a_cat, b_bd, float(scale), "column", tp, comm)


# ------------------------------------------------------------------------------------------------ fused q|k|v etc.
class _MultiLinearFn(torch.autograd.Function):
    """Several column-parallel projections of the same input as ONE GEMM each way.

    The weights stay separate parameters (checkpoint names are frozen) but live back to back in the parameter arena,
    so ``fused_w`` is a zero-copy [ΣN_i, K] view; forward is one GEMM, backward one dgrad GEMM (instead of one per
    projection — the reference issues 3 + 2 dX all-reduces per layer for the same reason) and one
    wgrad GEMM accumulating into the equally contiguous main-grad arena.
    """

    @staticmethod
    def forward(ctx, x, fused_w, fused_main_grad, comm, sp, rope, *weights):
        shape = x.shape
        ctx.sp = sp

        def project(a2):  # plain GEMM, or (q|k|v) GEMM with the rotary embedding in its epilogue
            if rope is None:
                return K.gemm(a2, fused_w) if sp is not None else _gemm_nt(a2, fused_w)
            cos, sin, pos, seq, rope_cols, d, rot_dim = rope
            return K.gemm_rope(a2, fused_w, cos, sin, pos, seq, rope_cols, d, rot_dim)

        if sp is not None:
            # sequence parallel: x is this rank's [b, s/tp, h] shard.  all-gather -> GEMM as one unit: the
            # shard is pushed into every rank's gather buffer and the GEMM's TMA producer starts each 256-row tile when
            # its rows have arrived; only the SHARD is kept for backward.
            xl = x.contiguous()
            xg = K.seq_gather_gated(xl, sp)
            y = project(xg.view(-1, shape[-1]))
            ctx.save_for_backward(xl, fused_w)
            shape = xg.shape
        else:
            x2 = x.reshape(-1, shape[-1])
            if not x2.is_contiguous():
                x2 = x2.contiguous()
            y = project(x2)
            ctx.save_for_backward(x2, fused_w)
        ctx.comm = comm
        ctx.fused_main_grad = fused_main_grad
        ctx.sizes = [w.shape[0] for w in weights]
        ctx.x_shape = x.shape
        return y.view(*shape[:-1], fused_w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, fused_w = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = None
        if ctx.sp is not None:
            # dX: dgrad GEMM + reduce-scatter along the sequence in one launch; the wgrad re-gathers the saved shard
            seq = dy.shape[1]
            if ctx.needs_input_grad[0]:
                dx = K.gemm_reduce_scatter(dy2, fused_w, ctx.sp, seq, b_mn=True).view(ctx.x_shape)
            if any(ctx.needs_input_grad[6:]):
                x2 = K.seq_all_gather(x2, ctx.sp, clone=False).view(-1, x2.shape[-1])
        elif ctx.needs_input_grad[0]:
            dx = _gemm_nn_reduce(dy2, fused_w, ctx.comm) if ctx.comm is not None else _gemm_nn(dy2, fused_w)
            dx = dx.view(ctx.x_shape)
        grads: list = [None] * len(ctx.sizes)
        if not any(ctx.needs_input_grad[6:]):
            pass  # frozen weights: no wgrad GEMM at all
        elif ctx.fused_main_grad is not None:
            K.gemm(dy2, x2, a_mn=True, b_mn=True, out=ctx.fused_main_grad, accumulate=True)
        else:
            dw = K.gemm(dy2, x2, a_mn=True, b_mn=True, out_dtype=fused_w.dtype)
            grads = list(dw.split(ctx.sizes, dim=0))
        return (dx, None, None, None, None, None, *grads)


def fused_view(tensors: Tuple[torch.Tensor, ...]) -> Optional[torch.Tensor]:
    """[ΣN_i, K] view over tensors that are contiguous, same width and adjacent in memory; else None."""
    first = tensors[0]
    if first.dim() != 2 or not all(t.is_contiguous() and t.dim() == 2 and t.shape[1] == first.shape[1]
                                   and t.dtype == first.dtype for t in tensors):
        return None
    ptr = first.data_ptr()
    rows = 0
    for t in tensors:
        if t.data_ptr() != ptr:
            return None
        ptr += t.numel() * t.element_size()
        rows += t.shape[0]
    base = first.detach()
    try:
        return torch.as_strided(base, (rows, first.shape[1]), (first.shape[1], 1))
    except RuntimeError:  # storage too small: the tensors are adjacent but live in different storages
        return None


def sp_gather_fusable(x_local: torch.Tensor, sp_comm, n_out: int) -> bool:
    """The all-gather -> GEMM unit applies: peer communicator, [b, s/tp, h] bf16 shard whose sequence slice is whole
    256-row tiles, an output wide enough for the 2-CTA kernel."""
    return (sp_comm is not None and x_local.dim() == 3 and native_enabled(x_local) and _PLAIN_GEMM == "native"
            and K.seq_gather_gated_supported(x_local, sp_comm) and n_out >= 128)


def multi_linear_fused(x: torch.Tensor, weights: Tuple[torch.nn.Parameter, ...], comm=None, sp=None, rope=None) -> Optional[torch.Tensor]:
    """x·[W_0; W_1; …]ᵀ as one [..., ΣN_i] tensor when the weights are adjacent in the arena (see multi_linear), else None.
    ``sp``: x is a sequence-parallel shard [b, s/tp, h]; it is all-gathered INTO the GEMM (result is [b, s, ΣN_i]).
    ``rope`` = (cos, sin, pos, seq, rope_cols, head_dim, rot_dim): rotary embedding on the first rope_cols output
    columns in the GEMM epilogue (see :func:`rope_in_gemm_supported`)."""
    if native_enabled(x) and x.dtype == torch.bfloat16 and all(_aligned(*w.shape) for w in weights) and x.numel() > 0:
        fw = fused_view(tuple(w.data for w in weights))
        if fw is not None:
            mains = [getattr(w, "main_grad", None) for w in weights]
            fmg = fused_view(tuple(mains)) if all(m is not None for m in mains) else None
            if fmg is not None or all(m is None for m in mains):
                return _MultiLinearFn.apply(x, fw, fmg, comm, sp, rope, *weights)
    return None


def multi_linear(x: torch.Tensor, weights: Tuple[torch.nn.Parameter, ...], comm=None) -> Tuple[torch.Tensor, ...]:
    """(x·W_0ᵀ, x·W_1ᵀ, …) — one GEMM when the weights are adjacent in the arena, separate linears otherwise.
    ``comm``: sum dX over the TP group inside the dgrad GEMM (column-parallel projections of a replicated input)."""
    if native_enabled(x) and x.dtype == torch.bfloat16 and all(_aligned(*w.shape) for w in weights) and x.numel() > 0:
        fw = fused_view(tuple(w.data for w in weights))
        if fw is not None:
            mains = [getattr(w, "main_grad", None) for w in weights]
            fmg = fused_view(tuple(mains)) if all(m is not None for m in mains) else None
            if fmg is not None or all(m is None for m in mains):
                y = _MultiLinearFn.apply(x, fw, fmg, comm, None, None, *weights)
                return tuple(y.split([w.shape[0] for w in weights], dim=-1))
    if comm is not None:
        x = _CopyToRegion.apply(x, comm)
    return tuple(linear(x, w) for w in weights)


# ------------------------------------------------------------------------------------------------ SwiGLU MLP up
class _SwiGLUFn(torch.autograd.Function):
    """h = silu(x·W1ᵀ) ⊙ (x·W2ᵀ) — dual GEMM with the activation in the TMEM epilogue (reference mlp.py:157-160:
    two cuBLAS GEMMs + silu + mul).  Keeps only the bf16 pre-activations; backward rebuilds dgate/dup in one
    element-wise pass and runs one dgrad + one wgrad GEMM over the concatenated [2F, K] weight."""

    @staticmethod
    def forward(ctx, x, fused_w, fused_main_grad, comm, sp, w_gate, w_up):
        ctx.comm = comm
        ctx.sp = sp
        shape = x.shape
        if sp is not None:  # sequence-parallel shard: all-gather -> dual GEMM as one unit (see _MultiLinearFn)
            xl = x.contiguous()
            xg = K.seq_gather_gated(xl, sp)
            h, pre = K.gemm_swiglu(xg.view(-1, shape[-1]), fused_w, keep_pre=True)
            ctx.save_for_backward(xl, fused_w, pre)
            shape = xg.shape
        else:
            x2 = x.reshape(-1, shape[-1])
            if not x2.is_contiguous():
                x2 = x2.contiguous()
            h, pre = K.gemm_swiglu(x2, fused_w, keep_pre=True)
            ctx.save_for_backward(x2, fused_w, pre)
        ctx.fused_main_grad = fused_main_grad
        ctx.x_shape = x.shape
        return h.view(*shape[:-1], h.shape[-1])

    @staticmethod
    def backward(ctx, dh):
        x2, fused_w, pre = ctx.saved_tensors
        dh2 = dh.reshape(-1, dh.shape[-1])
        if not dh2.is_contiguous():
            dh2 = dh2.contiguous()
        dpre, _ = K.swiglu_bwd(pre, dh2)
        dx = None
        if ctx.sp is not None:
            if ctx.needs_input_grad[0]:
                dx = K.gemm_reduce_scatter(dpre, fused_w, ctx.sp, dh.shape[1], b_mn=True).view(ctx.x_shape)
            if any(ctx.needs_input_grad[5:]):
                x2 = K.seq_all_gather(x2, ctx.sp, clone=False).view(-1, x2.shape[-1])
        elif ctx.needs_input_grad[0]:
            dx = _gemm_nn_reduce(dpre, fused_w, ctx.comm) if ctx.comm is not None else _gemm_nn(dpre, fused_w)
            dx = dx.view(ctx.x_shape)
        dg = du = None
        if not any(ctx.needs_input_grad[5:]):
            pass
        elif ctx.fused_main_grad is not None:
            K.gemm(dpre, x2, a_mn=True, b_mn=True, out=ctx.fused_main_grad, accumulate=True)
        else:
            dw = K.gemm(dpre, x2, a_mn=True, b_mn=True, out_dtype=fused_w.dtype)
            dg, du = dw.split(dw.shape[0] // 2, dim=0)
        return dx, None, None, None, None, dg, du


def swiglu_linear(x: torch.Tensor, w_gate: torch.nn.Parameter, w_up: torch.nn.Parameter, comm=None, sp=None) -> torch.Tensor:
    """``sp``: x is a sequence-parallel shard [b, s/tp, h] that is all-gathered into the dual GEMM (callers check
    :func:`sp_gather_fusable` first; the result then covers the full sequence)."""
    if (native_enabled(x) and x.dtype == torch.bfloat16 and _aligned(*w_gate.shape) and w_gate.shape == w_up.shape
            and x.numel() > 0):
        fw = fused_view((w_gate.data, w_up.data))
        if fw is not None:
            mains = [getattr(w_gate, "main_grad", None), getattr(w_up, "main_grad", None)]
            fmg = fused_view(tuple(mains)) if all(m is not None for m in mains) else None
            if fmg is not None or all(m is None for m in mains):
                return _SwiGLUFn.apply(x, fw, fmg, comm, sp, w_gate, w_up)
    assert sp is None, "sequence-parallel shard given to a path that cannot gather it"
    if comm is not None:
        x = _CopyToRegion.apply(x, comm)
    return F.silu(linear(x, w_gate)) * linear(x, w_up)


# ------------------------------------------------------------------------------------------------ norms
class _NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual, eps, rms):
        xc = x.contiguous()
        rc = residual.contiguous() if residual is not None else None
        y, summed, mean, rstd = K.norm_fwd(xc, weight, bias, eps, rms, residual=rc)
        inp = summed if rc is not None else xc
        ctx.save_for_backward(inp, weight, mean if mean is not None else rstd, rstd)
        ctx.rms = rms
        ctx.has_bias = bias is not None
        ctx.has_res = rc is not None
        ctx.weight_ref = weight
        ctx.bias_ref = bias
        return y, summed

    @staticmethod
    def backward(ctx, dy, dsum):
        inp, weight, mean, rstd = ctx.saved_tensors
        w_main = getattr(ctx.weight_ref, "main_grad", None)
        b_main = getattr(ctx.bias_ref, "main_grad", None) if ctx.has_bias else None
        use_main = w_main is not None and (not ctx.has_bias or b_main is not None)
        dres = dsum.contiguous() if (ctx.has_res and dsum is not None) else None
        dx, dw, db = K.norm_bwd(
            dy.contiguous(), inp, weight, None if ctx.rms else mean, rstd, ctx.rms, dres=dres,
            dw_out=w_main if use_main else None, db_out=b_main if use_main else None, accumulate=use_main,
            has_bias=ctx.has_bias,
        )
        if use_main:
            dw = db = None
        return dx, dw, db, (dx if ctx.has_res else None), None, None


def _norm_native_ok(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (native_enabled(x) and x.dtype in (torch.bfloat16, torch.float32, torch.float16) and weight.dtype == x.dtype
            and x.shape[-1] % 8 == 0 and x.numel() > 0)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, residual: Optional[torch.Tensor] = None):
    """RMSNorm(x [+ residual]) · w with fp32 statistics.  Returns y, or (y, x + residual) when a residual is given
    (reference rms_norm.py:45-53: fp32 normalise → cast → · weight)."""
    if _norm_native_ok(x, weight):
        y, s = _NormFn.apply(x, weight, None, residual, eps, True)
        return y if residual is None else (y, s)
    h = x if residual is None else x + residual
    hf = h.float()
    y = (hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps)).to(h.dtype) * weight
    return y if residual is None else (y, h)


def layer_norm(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], eps: float,
               residual: Optional[torch.Tensor] = None):
    if _norm_native_ok(x, weight) and (bias is None or bias.dtype == x.dtype):
        y, s = _NormFn.apply(x, weight, bias, residual, eps, False)
        return y if residual is None else (y, s)
    h = x if residual is None else x + residual
    y = F.layer_norm(h, (h.shape[-1],), weight, bias, eps)
    return y if residual is None else (y, h)


# ------------------------------------------------------------------------------------------------ activations
class _ActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, silu):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        ctx.silu = silu
        return K.act_fwd(xc, silu)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return K.act_bwd(x, dy.contiguous(), ctx.silu), None


class _LinearActFn(torch.autograd.Function):
    """act(x·Wᵀ + b) with the activation in the GEMM epilogue (adapter bottlenecks, GELU / SiLU MLP up-projections,
    SURVEY §8 L2; reference: F.linear then gelu, core/nn/mlp.py:77-89).  Keeps the bf16 pre-activation for backward."""

    @staticmethod
    def forward(ctx, x, weight, bias, silu, comm):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        h, pre = K.gemm_act(x2, weight, bias, silu)
        ctx.save_for_backward(x2, weight, pre)
        ctx.meta = (silu, comm, shape, bias is not None)
        return h.view(*shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dh):
        x2, weight, pre = ctx.saved_tensors
        silu, comm, shape, has_bias = ctx.meta
        dh2 = dh.reshape(-1, dh.shape[-1])
        if not dh2.is_contiguous():
            dh2 = dh2.contiguous()
        dpre = K.act_bwd(pre, dh2, silu)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = (_gemm_nn_reduce(dpre, weight, comm) if comm is not None else _gemm_nn(dpre, weight)).view(shape)
        if ctx.needs_input_grad[1]:
            dw = _accumulate_wgrad(dpre, x2, weight)
        if has_bias and ctx.needs_input_grad[2]:
            db = dpre.sum(0)
        return dx, dw, db, None, None


def linear_act(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], name: str, comm=None) -> Optional[torch.Tensor]:
    """Fused column-parallel linear + activation, or None when the epilogue kernel does not apply.  ``comm``: sum dX over
    the TP group inside the dgrad GEMM (replicated input)."""
    m = x.numel() // x.shape[-1] if x.numel() else 0
    if not (native_enabled(x) and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and m >= 256
            and weight.shape[0] >= 128 and _aligned(*weight.shape) and _PLAIN_GEMM == "native" and name in ("gelu", "silu")
            and (bias is None or bias.dtype == torch.bfloat16)):
        return None
    return _LinearActFn.apply(x, weight, bias, name == "silu", comm)


def activation(x: torch.Tensor, name: str) -> torch.Tensor:
    if native_enabled(x) and x.dtype == torch.bfloat16 and x.numel() % 2 == 0 and x.numel() > 0:
        return _ActFn.apply(x, name == "silu")
    return F.silu(x) if name == "silu" else F.gelu(x)


# ------------------------------------------------------------------------------------------------ rotary
class _RopeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cos, sin, pos, seq, mode, rot_dim):
        ctx.save_for_backward(cos, sin, pos)
        ctx.args = (seq, mode, rot_dim)
        out = x.clone(memory_format=torch.contiguous_format)
        return K.rope_(out, cos, sin, pos, seq, mode, rot_dim)

    @staticmethod
    def backward(ctx, dy):
        cos, sin, pos = ctx.saved_tensors
        seq, mode, rot_dim = ctx.args
        dx = K.rope_(dy.clone(memory_format=torch.contiguous_format), cos, sin, pos, seq, mode, rot_dim, backward=True)
        return dx, None, None, None, None, None, None


def rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, pos: Optional[torch.Tensor], seq: int, mode: int,
         rot_dim: int) -> torch.Tensor:
    """x: [tokens, heads, head_dim] bf16; cos/sin fp32 [max_pos, rot_dim/2]; pos int32 [tokens] or None."""
    return _RopeFn.apply(x, cos, sin, pos, seq, mode, rot_dim)


class _QKVRopeFn(torch.autograd.Function):
    """(q, k, v) = split(y) with rotary applied to q and k, as ONE kernel each way.

    Forward reads the fused q|k|v GEMM output once and writes one contiguous [T, nh + 2·nkv, d] buffer (q, k, v are
    views of it).  Backward reads dq, dk, dv once and writes the fused dY [T, (nh + 2·nkv)·d] that the dgrad / wgrad
    GEMMs of the projections consume — replacing two strided clones + two rotary launches forward and two clones, two
    rotary launches and a cat backward (reference: rotary after separate projections, attention.py:560-600)."""

    @staticmethod
    def forward(ctx, y, cos, sin, pos, seq, mode, rot_dim, nh, nkv, d, pre_rotated=False):
        T = y.shape[0]
        ctx.save_for_backward(cos, sin, pos)
        ctx.args = (seq, mode, rot_dim, nh, nkv, d, y.shape)
        y2 = y.reshape(T, -1)
        if pre_rotated:  # the projection GEMM's epilogue already rotated q and k: the split is three strided views
            return _qkv_views(y2, nh, nkv, d)
        out = torch.empty((T, nh + 2 * nkv, d), dtype=y.dtype, device=y.device)
        ins = _qkv_views(y2, nh, nkv, d)
        outs = (out[:, :nh], out[:, nh:nh + nkv], out[:, nh + nkv:])
        K.qkv_rope(ins, outs, (nh, nkv, nkv), (True, True, False), cos, sin, pos, seq, mode, rot_dim)
        return outs

    @staticmethod
    def backward(ctx, dq, dk, dv):
        cos, sin, pos = ctx.saved_tensors
        seq, mode, rot_dim, nh, nkv, d, shape = ctx.args
        T = dq.shape[0]
        dy = torch.empty((T, (nh + 2 * nkv) * d), dtype=dq.dtype, device=dq.device)
        ins = tuple(g if (g.stride(2) == 1 and g.stride(1) == d) else g.contiguous() for g in (dq, dk, dv))
        K.qkv_rope(ins, _qkv_views(dy, nh, nkv, d), (nh, nkv, nkv), (True, True, False), cos, sin, pos, seq, mode,
                   rot_dim, backward=True)
        return (dy.view(shape),) + (None,) * (len(ctx.needs_input_grad) - 1)


def _qkv_views(y2: torch.Tensor, nh: int, nkv: int, d: int):
    T = y2.shape[0]
    q = y2[:, :nh * d].unflatten(1, (nh, d))
    k = y2[:, nh * d:(nh + nkv) * d].unflatten(1, (nkv, d))
    v = y2[:, (nh + nkv) * d:].unflatten(1, (nkv, d))
    return q, k, v


def qkv_rope_split(y: torch.Tensor, nh: int, nkv: int, d: int, cos: torch.Tensor, sin: torch.Tensor,
                   pos: Optional[torch.Tensor], seq: int, mode: int, rot_dim: int, pre_rotated: bool = False):
    """y: [..., (nh + 2·nkv)·d] fused projection output → (q [T, nh, d], k [T, nkv, d], v [T, nkv, d]), q and k rotated.
    ``pre_rotated``: the GEMM that produced y already applied the rotation (``rope=`` of :func:`multi_linear_fused`); the
    forward is then free (strided views), the backward still un-rotates dq / dk while assembling the fused dY."""
    return _QKVRopeFn.apply(y.reshape(-1, y.shape[-1]), cos, sin, pos, seq, mode, rot_dim, nh, nkv, d, pre_rotated)


def rope_in_gemm_supported(x: torch.Tensor, n_out: int, d: int, rot_dim: int, mode: int) -> bool:
    m = x.numel() // x.shape[-1] if x.numel() else 0
    return _PLAIN_GEMM == "native" and _ROPE_IN_GEMM and K.gemm_rope_supported(m, n_out, d, rot_dim, mode)


_ROPE_IN_GEMM = os.environ.get("SCALING_B200_ROPE_IN_GEMM", "1") == "1"  # measured: N = 1 step 1582.6 -> 1575.7 ms


def qkv_rope_fusable(x: torch.Tensor, d: int, rot_dim: int) -> bool:
    return native_enabled(x) and x.dtype == torch.bfloat16 and d % 8 == 0 and rot_dim % 16 == 0 and 0 < rot_dim <= d


# ------------------------------------------------------------------------------------------------ dropout + residual
class _BiasDropoutAddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, bias, residual, p, seed, offset):
        ctx.args = (p, seed, offset)
        ctx.has_bias = bias is not None
        ctx.has_res = residual is not None
        return K.bias_dropout_add(x.contiguous(), bias, residual.contiguous() if residual is not None else None, p,
                                  seed, offset)

    @staticmethod
    def backward(ctx, dout):
        p, seed, offset = ctx.args
        dout = dout.contiguous()
        dx = K.dropout_bwd(dout, p, seed, offset) if p > 0 else dout
        db = dx.reshape(-1, dx.shape[-1]).sum(0) if ctx.has_bias else None
        return dx, db, (dout if ctx.has_res else None), None, None, None


def bias_dropout_add(x: torch.Tensor, bias: Optional[torch.Tensor], residual: Optional[torch.Tensor], p: float,
                     training: bool, seed: int, offset: int) -> torch.Tensor:
    """residual + dropout(x + bias).  The mask is a pure function of (seed, offset, element index): identical on
    every TP rank and under activation recomputation (reference layer.py:211-216 forks a stateful RNG instead)."""
    p = p if training else 0.0
    if native_enabled(x) and x.dtype == torch.bfloat16 and x.numel() % 4 == 0 and x.shape[-1] % 4 == 0 and x.numel() > 0:
        if p == 0.0 and bias is None:
            return x if residual is None else x + residual
        return _BiasDropoutAddFn.apply(x, bias, residual, p, seed, offset)
    y = x if bias is None else x + bias
    if p > 0.0:
        y = F.dropout(y, p=p, training=True)
    return y if residual is None else y + residual


# ------------------------------------------------------------------------------------------------ embedding
class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight, vocab_start, vocab_end):
        ids = ids.contiguous()
        ctx.save_for_backward(ids)
        ctx.weight_ref = weight
        ctx.range = (vocab_start, vocab_end)
        return K.embedding_fwd(ids, weight, vocab_start, vocab_end)

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        w = ctx.weight_ref
        main = getattr(w, "main_grad", None)
        if main is not None:
            K.embedding_bwd(ids, dout.contiguous(), main, *ctx.range)
            return None, None, None, None
        dw = torch.zeros(w.shape, dtype=torch.float32, device=w.device)
        K.embedding_bwd(ids, dout.contiguous(), dw, *ctx.range)
        return None, dw.to(w.dtype), None, None


class _EmbeddingPeersFn(torch.autograd.Function):
    """Vocab-parallel embedding INCLUDING its tensor-parallel reduction, done by placement instead of arithmetic: the
    rank that owns a token's row stores it into every rank's output (or, under sequence parallelism, into the output of
    the rank that owns the token's sequence position) with NVLink peer stores — ``sb_embedding_fwd_peers`` — followed
    by one device barrier.  Replaces masked lookup + dense all-reduce / reduce-scatter of [b, s, h]
    (reference core/nn/linear/vocab_parallel_embedding.py:122-140)."""

    @staticmethod
    def forward(ctx, ids, weight, vocab_start, vocab_end, comm, sp_group_size):
        ids = ids.contiguous()
        b, s = (ids.shape[0], ids.shape[1]) if ids.dim() == 2 else (1, ids.numel())
        hidden = weight.shape[1]
        rows = ids.numel() // sp_group_size if sp_group_size else ids.numel()
        staging, ptrs = comm.exchange_buffer("embedding", rows * hidden * 2)
        _lib.call("sb_embedding_fwd_peers", _lib.ptr(ids), _lib.ptr(weight), ptrs, comm.size, ids.numel(), hidden,
                     vocab_start, vocab_end, s if sp_group_size else 0)
        comm.barrier()  # the peers' rows have landed here
        out_shape = (*ids.shape[:-1], ids.shape[-1] // sp_group_size if sp_group_size else ids.shape[-1], hidden)
        out = staging.view(torch.bfloat16).view(out_shape).clone()  # leave the exchange buffer to the next call
        ctx.save_for_backward(ids)
        ctx.weight_ref = weight
        ctx.range = (vocab_start, vocab_end)
        ctx.sp = (comm.size, comm.group) if sp_group_size else None
        return out

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        w = ctx.weight_ref
        dout = dout.contiguous()
        if ctx.sp is not None:  # gradient of the scatter along the sequence: gather it back (dim 1)
            from ..core.nn.tp_ops import _all_gather_seq

            size, group = ctx.sp
            dout = _all_gather_seq(dout, size, group)
        main = getattr(w, "main_grad", None)
        if main is not None:
            K.embedding_bwd(ids, dout, main, *ctx.range)
            return None, None, None, None, None, None
        dw = torch.zeros(w.shape, dtype=torch.float32, device=w.device)
        K.embedding_bwd(ids, dout, dw, *ctx.range)
        return None, dw.to(w.dtype), None, None, None, None


def vocab_embedding_reduced(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, vocab_end: int, comm: Any,
                            sequence_parallel: bool) -> Optional[torch.Tensor]:
    """The embedding already summed over the TP group (replicated, or scattered along the sequence when
    ``sequence_parallel``) through peer stores, or None when that path does not apply (the caller then reduces)."""
    if comm is None or not native_enabled(weight) or weight.dtype != torch.bfloat16 or weight.shape[1] % 8 or ids.numel() == 0:
        return None
    if sequence_parallel and (ids.dim() != 2 or ids.shape[1] % comm.size):
        return None
    return _EmbeddingPeersFn.apply(ids.long(), weight, vocab_start, vocab_end, comm, comm.size if sequence_parallel else 0)


def vocab_embedding(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int, vocab_end: int) -> torch.Tensor:
    """Rows of this rank's vocab shard for ids in [vocab_start, vocab_end), zeros elsewhere (reference
    vocab_parallel_embedding.py:119-135)."""
    if native_enabled(weight) and weight.dtype == torch.bfloat16 and weight.shape[1] % 8 == 0 and ids.numel() > 0:
        return _EmbeddingFn.apply(ids.long(), weight, vocab_start, vocab_end)
    mask = (ids < vocab_start) | (ids >= vocab_end)
    local = (ids - vocab_start).masked_fill(mask, 0)
    out = F.embedding(local, weight)
    return out.masked_fill(mask.unsqueeze(-1), 0.0)


# ------------------------------------------------------------------------------------------------ cross-entropy
class _VocabParallelXentFn(torch.autograd.Function):
    """Per-token CE loss and arg-max over a vocabulary that is sharded across the tensor-parallel group.

    Forward: one stats pass over the local [tokens, V/tp] logits (online max/Σexp, target logit, arg-max), a 16-byte
    per-token all-gather across TP (instead of the reference's full-logit all-gather + fp32 copy,
    lm_head_tied.py:41-43 / model.py:56-67).  Backward: overwrites the saved logits in place with
    (softmax − onehot)·dloss — no second [tokens, V] buffer ever exists.
    """

    @staticmethod
    def forward(ctx, logits, targets, vocab_start, group):
        assert logits.dim() == 2
        targets = targets.contiguous()
        stats = K.xent_stats(logits, targets, vocab_start)
        if group is not None and torch.distributed.get_world_size(group) > 1:
            ws = torch.distributed.get_world_size(group)
            gathered = torch.empty((ws * stats.shape[0], stats.shape[1]), dtype=stats.dtype, device=stats.device)
            torch.distributed.all_gather_into_tensor(gathered, stats.contiguous(), group=group)
            gathered = gathered.view(ws, *stats.shape)
            m, idx = gathered[..., 0].max(dim=0)
            s = (gathered[..., 1] * torch.exp(gathered[..., 0] - m)).sum(0)
            tl = gathered[..., 2].sum(0)
            am = gathered[..., 3].gather(0, idx.unsqueeze(0)).squeeze(0)
        else:
            m, s, tl, am = stats[:, 0], stats[:, 1], stats[:, 2], stats[:, 3]
        lse = m + torch.log(s)
        ctx.save_for_backward(logits, targets, lse)
        ctx.vocab_start = vocab_start
        argmax = am.contiguous().view(torch.int32).long()
        ctx.mark_non_differentiable(argmax)
        return lse - tl, argmax

    @staticmethod
    def backward(ctx, dloss, _dargmax):
        logits, targets, lse = ctx.saved_tensors
        K.xent_grad_(logits, targets, lse.contiguous(), dloss.float().contiguous(), None, ctx.vocab_start)
        return logits, None, None, None


class _LinearXentStatsFn(torch.autograd.Function):
    """LM-head projection whose GEMM epilogue also emits the cross-entropy statistics of every 256-column tile
    (``sb_gemm2_xent_stats_bf16``): returns (logits, partial statistics).  Gradients as :class:`_LinearFn`."""

    @staticmethod
    def forward(ctx, x, weight, comm):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y, partial = K.gemm_xent_stats(x2, weight)
        ctx.comm = comm
        ctx.save_for_backward(x2, weight)
        ctx.x_shape = shape
        ctx.mark_non_differentiable(partial)
        return y.view(*shape[:-1], weight.shape[0]), partial

    @staticmethod
    def backward(ctx, dy, _dpartial):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = (_gemm_nn_reduce(dy2, weight, ctx.comm) if ctx.comm is not None else _gemm_nn(dy2, weight)).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = _accumulate_wgrad(dy2, x2, weight)
        return dx, dw, None


def linear_with_xent_stats(x: torch.Tensor, weight: torch.Tensor, comm=None):
    """(logits, partial CE statistics) in one GEMM launch, or None when the epilogue kernel does not apply.  ``comm``:
    sum dX over the TP group inside the dgrad GEMM, as in :func:`linear`."""
    m = x.numel() // x.shape[-1] if x.numel() else 0
    if not (native_enabled(x) and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and m >= 256
            and weight.shape[0] >= 128 and _aligned(*weight.shape) and _PLAIN_GEMM == "native"):
        return None
    return _LinearXentStatsFn.apply(x, weight, comm)


class _XentFromStatsFn(torch.autograd.Function):
    """Cross-entropy on top of the LM-head epilogue statistics: forward never reads the logits again (only the one
    target logit per token); backward overwrites them in place with (softmax − onehot)·dloss."""

    @staticmethod
    def forward(ctx, logits, partial, targets, vocab_start, comm):
        targets = targets.contiguous()
        loss, lse, argmax = K.xent_from_partials(partial, logits, targets, vocab_start, comm)
        ctx.save_for_backward(logits, targets, lse)
        ctx.vocab_start = vocab_start
        ctx.mark_non_differentiable(argmax)
        return loss, argmax

    @staticmethod
    def backward(ctx, dloss, _dargmax):
        logits, targets, lse = ctx.saved_tensors
        K.xent_grad_(logits, targets, lse, dloss.float().contiguous(), None, ctx.vocab_start)
        return logits, None, None, None, None


def vocab_parallel_cross_entropy(logits: torch.Tensor, targets: torch.Tensor, vocab_start: int = 0, group=None,
                                 partial: Optional[torch.Tensor] = None, comm=None):
    if partial is not None and logits.dim() == 2 and logits.is_contiguous():
        return _XentFromStatsFn.apply(logits, partial, targets, vocab_start, comm)
    return _vocab_parallel_cross_entropy(logits, targets, vocab_start, group)


def _vocab_parallel_cross_entropy(logits: torch.Tensor, targets: torch.Tensor, vocab_start: int = 0, group=None):
    """logits: [tokens, V_local] (this rank's shard), targets: [tokens] global ids → (loss [tokens] fp32, argmax [tokens]).
    NOTE: the backward pass reuses the logits buffer for the gradient."""
    if native_enabled(logits) and logits.dtype in (torch.bfloat16, torch.float32) and logits.stride(-1) == 1 \
            and logits.stride(0) % 8 == 0 and logits.numel() > 0:
        return _VocabParallelXentFn.apply(logits, targets, vocab_start, group)
    lf = logits.float()
    ws = torch.distributed.get_world_size(group) if group is not None else 1
    if ws > 1:
        from ..core.nn.tp_ops import all_concat_with_group

        lf = all_concat_with_group(lf, dim=-1, group=group)
    loss = F.cross_entropy(lf, targets, reduction="none")
    return loss, lf.argmax(-1)
