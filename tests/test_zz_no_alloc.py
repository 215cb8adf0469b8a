"""The allocation-free norm backward (``sb_norm_bwd_noalloc``) against the regular entry point (``sb_norm_bwd``).

Some users of this library may not allocate GPU memory at the point where they need a kernel: they are recording a
CUDA graph, they have already given every free byte to something else (a KV cache, another framework's pool), or they
call the C ABI from a host that has no allocator to lend.  For them the library carries its own fixed scratch array;
this test drives that variant exactly as such a caller would — straight through the C ABI with ctypes, passing only
input and output pointers — and checks that it returns what the scratch-taking entry point returns.  It runs the pair
twice in a row on different data because the built-in array is shared between calls and still holds the previous
call's partial sums when the second one starts.

The file name makes pytest collect it last, so a run with ``-x`` has already exercised everything else if this fails.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

P = ctypes.c_void_p
I = ctypes.c_int


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


@pytest.fixture(scope="module")
def lib():
    from scaling_b200.ops import _lib

    handle = ctypes.CDLL(str(_lib.lib_path()))
    handle.sb_norm_bwd.argtypes = [P, P, P, P, P, P, P, P, P, I, I, P, I, I, I, I, P]
    handle.sb_norm_bwd.restype = I
    handle.sb_norm_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, P, P, I, I, I, I, I, I, P]
    handle.sb_norm_bwd_noalloc.restype = I
    handle.sb_norm_bwd_scratch_floats.argtypes = [I, I, I]
    handle.sb_norm_bwd_scratch_floats.restype = ctypes.c_longlong
    handle.sb_norm_bwd_noalloc_max_cols.argtypes = [I]
    handle.sb_norm_bwd_noalloc_max_cols.restype = I
    handle.sb_norm_bwd_noalloc_strerror.argtypes = [I]
    handle.sb_norm_bwd_noalloc_strerror.restype = ctypes.c_char_p
    return handle


def _rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.mark.parametrize("rms,dtype,code,with_dres,accumulate", [
    (1, torch.bfloat16, 0, False, 0), (0, torch.float32, 1, False, 0),
    (1, torch.float16, 2, True, 1), (0, torch.bfloat16, 0, True, 1)])
def test_noalloc_matches_regular_entry_point_twice(lib, rms, dtype, code, with_dres, accumulate):
    torch.manual_seed(0)
    dev = "cuda"
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    for call, (rows, cols) in enumerate([(1000, 4096), (333, 2048)]):
        x = torch.randn(rows, cols, device=dev, dtype=dtype)
        dy = torch.randn(rows, cols, device=dev, dtype=dtype)
        w = (1 + 0.1 * torch.randn(cols, device=dev)).to(dtype)
        xf = x.float()
        if rms:
            mean = None
            rstd = torch.rsqrt(xf.pow(2).mean(-1) + 1e-5)
        else:
            mean = xf.mean(-1)
            rstd = torch.rsqrt(xf.var(-1, unbiased=False) + 1e-5)
        # everything the two calls write is allocated up front by the caller; the calls themselves allocate nothing
        dx_ref, dx_new = torch.empty_like(x), torch.empty_like(x)
        # with accumulate=1 the gradients are added to what the buffers already hold (main-grad semantics)
        dw0, db0 = torch.randn(cols, device=dev), torch.randn(cols, device=dev)
        dw_ref, dw_new = dw0.clone(), dw0.clone()
        db_ref, db_new = db0.clone(), db0.clone()
        dres = torch.randn(rows, cols, device=dev, dtype=dtype) if with_dres else None  # gradient of the residual branch
        scratch = torch.empty(lib.sb_norm_bwd_scratch_floats(rows, cols, rms), device=dev, dtype=torch.float32)
        rc = lib.sb_norm_bwd(_ptr(dy), _ptr(x), _ptr(w), _ptr(mean), _ptr(rstd), _ptr(dres), _ptr(dx_ref), _ptr(dw_ref),
                             None if rms else _ptr(db_ref), 1, accumulate, _ptr(scratch), rows, cols, rms, code, stream)
        assert rc == 0
        before = torch.cuda.memory_allocated()
        rc = lib.sb_norm_bwd_noalloc(_ptr(dy), _ptr(x), _ptr(w), _ptr(mean), _ptr(rstd), _ptr(dres), _ptr(dx_new), _ptr(dw_new),
                                     None if rms else _ptr(db_new), 1, accumulate, rows, cols, rms, code, stream)
        assert rc == 0, lib.sb_norm_bwd_noalloc_strerror(rc).decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert _rel(dx_new, dx_ref) < tol, (call, _rel(dx_new, dx_ref))
        assert _rel(dw_new, dw_ref) < tol, (call, _rel(dw_new, dw_ref))
        if not rms:
            assert _rel(db_new, db_ref) < tol, (call, _rel(db_new, db_ref))
        assert torch.isfinite(dw_new).all()


def test_noalloc_empty_problem_is_a_no_op_like_the_regular_entry_point(lib):
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.sb_norm_bwd_noalloc(None, None, None, None, None, None, None, None, None, 1, 0, 0, 4096, 1, 0, stream) == 0
    assert lib.sb_norm_bwd(None, None, None, None, None, None, None, None, None, 1, 0, None, 0, 4096, 1, 0, stream) == 0


def test_noalloc_refuses_oversized_problems(lib):
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.sb_norm_bwd_noalloc_max_cols(1) == 8192 and lib.sb_norm_bwd_noalloc_max_cols(0) == 4096
    t = torch.zeros(8, 16384, device="cuda", dtype=torch.bfloat16)
    r = torch.ones(8, device="cuda")
    g = torch.zeros(16384, device="cuda")
    for rms, cols in ((1, 16384), (0, 8192)):
        rc = lib.sb_norm_bwd_noalloc(_ptr(t), _ptr(t), _ptr(t), _ptr(r), _ptr(r), None, _ptr(t), _ptr(g), _ptr(g), 1, 0, 8,
                                     cols, rms, 0, stream)
        assert rc == -3
        assert b"too large" in lib.sb_norm_bwd_noalloc_strerror(rc)
    torch.cuda.synchronize()
