"""Launch ONE hot kernel of the flagship step a few times at its headline shape, as a target for
`ncu --set full -k regex:<name> -s 2 -c 1` (profiling recipe: capture the top kernel once per change).

    python tools/ncu_target.py gemm_wgrad | gemm_fprop | gemm_dgrad | gemm_swiglu | attn
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
