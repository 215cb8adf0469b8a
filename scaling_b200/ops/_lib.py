"""ctypes binding of ``scaling_b200/_kernels.so`` (C ABI declared in ``include/scaling_b200.h``).

The library is built in-tree by :mod:`scaling_b200.ops.build`.  On a machine with a GPU a missing or broken library
is a hard error — the sm_100a kernels are the product, there is no silent fallback.  On a CPU-only machine
``available()`` is False and the modules use their plain PyTorch expressions (the oracle / gloo path).
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_float, c_int, c_longlong, c_uint, c_ulonglong, c_void_p
from pathlib import Path
from typing import Optional

import torch

_LIB_PATH = Path(__file__).resolve().parents[1] / "_kernels.so"
_lib: Optional[ctypes.CDLL] = None
_tls = threading.local()  # per-thread: which device's context this thread has been bound to
_launches = 0  # number of kernel launches issued through this binding (bench.py reports it as gpu_launches)

P = c_void_p
_SIGS = {
    "sb_gemm_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm_swiglu_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_f16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_lowrank_bf16": [P, P, P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                              c_int, c_int, P],
    "sb_gemm2_allreduce_lowrank_bf16": [P, P, P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                        c_int, P, P, c_int, c_int, c_uint, c_ulonglong, c_uint, c_int, P],
    "sb_gemm2_rope_bf16": [P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, P, P, P, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_act_bf16": [P, P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_swiglu_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_gemm2_allreduce_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P, P, c_int, c_int,
                                c_uint, c_ulonglong, c_uint, c_int, P],
    "sb_gemm2_reduce_scatter_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P, P, c_int, c_int,
                                     c_uint, c_ulonglong, c_uint, c_int, c_int, P],
    "sb_norm_fwd": [P, P, P, P, P, P, P, P, c_int, c_int, c_float, c_int, c_int, P],
    "sb_norm_bwd": [P, P, P, P, P, P, P, P, P, c_int, c_int, P, c_int, c_int, c_int, c_int, P],
    "sb_swiglu_fwd": [P, P, c_int, c_int, P],
    "sb_swiglu_bwd": [P, P, P, P, c_int, c_int, P],
    "sb_act_fwd": [P, P, c_longlong, c_int, P],
    "sb_act_bwd": [P, P, P, c_longlong, c_int, P],
    "sb_rope": [P, P, P, P, c_int, c_int, c_int, c_int, c_longlong, c_int, c_int, c_int, P],
    "sb_qkv_rope": [P, P, P, P, P, P, c_int, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_bias_dropout_add": [P, P, P, P, c_longlong, c_int, c_float, c_ulonglong, c_ulonglong, P],
    "sb_dropout_bwd": [P, P, c_longlong, c_float, c_ulonglong, c_ulonglong, P],
    "sb_embedding_fwd": [P, P, P, c_int, c_int, c_longlong, c_longlong, P],
    "sb_embedding_bwd": [P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, P],
    "sb_embedding_bwd_sorted": [P, P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, P],
    "sb_embedding_fwd_peers": [P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, c_int, P],
    "sb_seq_gather_push": [P, P, c_int, c_int, c_int, c_int, c_int, P],
    "sb_seq_gather_push_counted": [P, P, P, c_int, c_int, c_int, c_int, c_int, P],
    "sb_seq_scatter_push": [P, P, c_int, c_int, c_int, c_int, c_int, P],
    "sb_sum_slots": [P, P, c_int, c_longlong, P],
    "sb_xent_stats": [P, P, P, c_int, c_int, c_longlong, c_longlong, c_int, P],
    "sb_gemm2_xent_stats_bf16": [P, P, P, P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, P],
    "sb_xent_combine": [P, c_int, P, c_longlong, P, c_int, c_int, c_longlong, P, c_int, c_int, P],
    "sb_xent_finalize": [P, c_int, c_int, P, P, P, P],
    "sb_xent_grad": [P, P, P, P, P, c_int, c_int, c_longlong, c_longlong, c_int, P],
    "sb_attention_fwd": [P, P, P, P, P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong, c_float, c_int, P],
    "sb_attention_fwd_ex": [P, P, P, P, P, P, P, c_int, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong, c_float,
                            c_int, c_float, c_ulonglong, c_int, P],
    "sb_attention_bwd_ex": [P, P, P, P, P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong,
                            c_float, c_int, c_float, c_ulonglong, c_int, P],
    "sb_attention_fwd_f16": [P, P, P, P, P, P, P, c_int, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong, c_float,
                             c_int, c_float, c_ulonglong, c_int, P],
    "sb_attention_bwd_f16": [P, P, P, P, P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong,
                             c_float, c_int, c_float, c_ulonglong, c_int, P],
    "sb_attention_bwd": [P, P, P, P, P, P, P, P, P, P, P, P, c_int, c_int, c_int, c_longlong, c_longlong, c_longlong,
                         c_float, c_int, P],
    "sb_grad_sumsq": [P, c_longlong, c_int, P, P, c_int, P, P],
    "sb_adamw": [P, P, P, P, P, c_longlong, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_int, c_float,
                 c_float, P, P, c_int, P],
    "sb_cast_from_fp32": [P, P, c_longlong, c_int, P],
    "sb_adamw_zero": [P, P, P, P, P, c_int, c_longlong, c_int, c_int, c_float, c_float, c_float, c_float, c_float, c_int,
                      c_float, c_float, P, P, c_int, P],
    "sb_zero_reduce": [P, c_int, P, c_longlong, c_int, P],
    "sb_dp_allreduce": [P, c_int, c_longlong, c_int, P],
    "sb_device_barrier": [P, c_int, c_int, c_uint, P],
}
_RET_LL = {
    "sb_norm_bwd_scratch_floats": [c_int, c_int, c_int],
    "sb_blend_index": [P, c_int, P, c_longlong],
    "sb_attention_bwd_scratch_floats": [c_int, c_int],
    "sb_gather_window_i32": [P, P, P, c_int, P, c_longlong],
    "sb_gemm2_allreduce_slot_bytes": [c_int, c_int],
    "sb_gemm2_split_workspace_bytes": [],
    "sb_gemm2_xent_tiles": [c_int],
}
_RET_INT_NOSTREAM = {"sb_grad_sumsq_scratch_floats": [], "sb_bind_device": [c_int],
                     "sb_gemm2_set_split_workspace": [P, c_ulonglong], "sb_gemm2_set_profile_buffer": [P],
                     "sb_gemm2_set_gate": [P, c_uint], "sb_seq_gather_chunks_per_tile": []}


def lib_path() -> Path:
    return _LIB_PATH


def load(required: bool = False) -> Optional[ctypes.CDLL]:
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.is_file():
        if os.environ.get("SCALING_B200_AUTOBUILD", "1") == "1":
            try:
                from .build import build

                build()
            except Exception as e:  # noqa: BLE001
                if required:
                    raise RuntimeError(f"scaling_b200 native library could not be built: {e}") from e
                return None
        if not _LIB_PATH.is_file():
            if required:
                raise RuntimeError(f"{_LIB_PATH} is missing; run `python -m scaling_b200.ops.build`")
            return None
    lib = ctypes.CDLL(str(_LIB_PATH))
    for name, sig in _SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = sig
        fn.restype = c_int
    for name, sig in _RET_LL.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = sig
        fn.restype = c_longlong
    for name, sig in _RET_INT_NOSTREAM.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = sig
        fn.restype = c_int
    _lib = lib
    return lib


def host_lib() -> ctypes.CDLL:
    """Library handle for the host-only (C++) entry points; usable without a GPU."""
    lib = load(required=True)
    assert lib is not None
    return lib


def available() -> bool:
    """True when the sm_100a kernels can run: CUDA device present and the native library loads."""
    if not torch.cuda.is_available():
        return False
    return load(required=True) is not None


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_raw_device = getattr(torch._C, "_cuda_getDevice", None)


def stream() -> int:
    """Raw handle of the current stream of the current device (the C-level query skips building a Stream object)."""
    if _raw_stream is not None and _raw_device is not None:
        return _raw_stream(_raw_device())
    return torch.cuda.current_stream().cuda_stream


def call(name: str, *args, launches: int = 1) -> None:
    """Invoke a kernel entry point on the current stream and raise on a non-zero status."""
    global _launches
    lib = load(required=True)
    dev = _raw_device() if _raw_device is not None else torch.cuda.current_device()
    if getattr(_tls, "device", -1) != dev:  # e.g. an autograd worker thread making its first native call
        rc = lib.sb_bind_device(dev)
        if rc != 0:
            raise RuntimeError(f"sb_bind_device({dev}) failed with CUDA error {rc}")
        _tls.device = dev
    rc = getattr(lib, name)(*args, _raw_stream(dev) if _raw_stream is not None else stream())
    if rc != 0:
        raise RuntimeError(f"{name} failed with status {rc}")
    _launches += launches


def launch_count() -> int:
    return _launches


def reset_launch_count() -> None:
    global _launches
    _launches = 0


DTYPE_CODE = {torch.bfloat16: 0, torch.float32: 1, torch.float16: 2}
