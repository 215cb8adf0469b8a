"""Launch ONE hot kernel of the flagship step a few times at its headline shape, as a target for
`ncu --set full -k regex:<name> -s 2 -c 1` (profiling recipe: capture the top kernel once per change).

    python tools/ncu_target.py gemm_wgrad | gemm_fprop | gemm_dgrad | gemm_swiglu | attn | gemm_allreduce | gemm_allreduce_dgrad
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from scaling_b200.ops import kernels as K  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "gemm_wgrad"
dev = "cuda"
T, H, F = 4096, 4096, 14336
g = torch.Generator(device=dev).manual_seed(0)


def rnd(*shape):
    return (torch.randn(*shape, device=dev, generator=g) * 0.05).bfloat16()


if which == "gemm_wgrad":  # dW[2F, H] += dYᵀ[2F, T] · X[T, H], fp32 accumulate into the main-grad arena
    dy, x = rnd(T, 2 * F), rnd(T, H)
    main = torch.zeros(2 * F, H, device=dev, dtype=torch.float32)
    fn = lambda: K.gemm(dy, x, a_mn=True, b_mn=True, out=main, accumulate=True)  # noqa: E731
elif which == "gemm_fprop":  # Y[T, H] = X[T, F] · Wᵀ (mlp down projection)
    x, w = rnd(T, F), rnd(H, F)
    fn = lambda: K.gemm(x, w)  # noqa: E731
elif which == "gemm_dgrad":  # dX[T, H] = dY[T, 2F] · W[2F, H]
    dy, w = rnd(T, 2 * F), rnd(2 * F, H)
    fn = lambda: K.gemm(dy, w, b_mn=True)  # noqa: E731
elif which == "gemm_swiglu":
    x, w = rnd(T, H), rnd(2 * F, H)
    fn = lambda: K.gemm_swiglu(x, w)  # noqa: E731
elif which == "attn":
    from scaling_b200.ops.attention_sm100 import AttentionFn

    q = rnd(T, 32, 128).requires_grad_(True)
    k = rnd(T, 8, 128).requires_grad_(True)
    v = rnd(T, 8, 128).requires_grad_(True)
    do = rnd(T, 32, 128)
    cu = torch.tensor([0, T], device=dev, dtype=torch.int32)

    def fn():
        o = AttentionFn.apply(q, k, v, cu, T, True, 128 ** -0.5, None)
        torch.autograd.grad(o, (q, k, v), do)
elif which in ("gemm_allreduce", "gemm_allreduce_dgrad"):
    # The fused GEMM + all-reduce kernel on ONE GPU (ncu cannot follow a multi-rank job): "rank 0 of 2" whose peer
    # workspace is a second local buffer and whose peer flags are pre-set to the call's epoch, so the reducer warps do not
    # wait.  Pushes and reductions run exactly as in a real pair, only the NVLink hop is a local store.
    import ctypes

    from scaling_b200.ops import _lib

    lib = _lib.load(required=True)
    b_mn = which.endswith("dgrad")
    M, N, Kd = (T, H, 2 * F // 2) if b_mn else (T, H, F // 2)  # TP=2 halves: mlp-up dgrad / mlp-down fprop
    a = rnd(M, Kd)
    b = rnd(Kd, N) if b_mn else rnd(N, Kd)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    slot = (int(lib.sb_gemm2_allreduce_slot_bytes(M, N)) + 65535) // 65536 * 65536
    flag_slot = (slot // 65536 * 4 + 1023) // 1024 * 1024
    data = [torch.zeros(2 * slot, dtype=torch.uint8, device=dev) for _ in range(2)]
    flags = [torch.zeros(2 * flag_slot, dtype=torch.int32, device=dev) for _ in range(2)]
    dp = (ctypes.c_ulonglong * 2)(*[t.data_ptr() for t in data])
    fp = (ctypes.c_ulonglong * 2)(*[t.data_ptr() for t in flags])
    K._ensure_split_workspace(a.device)
    epoch = [0]

    def fn():
        epoch[0] += 1
        flags[0][flag_slot:].fill_(epoch[0])  # "the peer has pushed everything"
        _lib.call("sb_gemm2_allreduce_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), None, M, N, Kd, a.stride(0),
                  b.stride(0), N, 0, int(b_mn), dp, fp, 0, 2, epoch[0], slot, flag_slot, 0)
else:
    raise SystemExit(f"unknown target {which}")

for _ in range(4):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    fn()
e1.record()
torch.cuda.synchronize()
print(f"{which}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per call")
