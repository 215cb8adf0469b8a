import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from scaling_b200.ops.attention_sm100 import AttentionFn  # noqa: E402

dev = "cuda"
for Tt in (4096, 8192):
    heads, kvh = 32, 8
    q = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    cu = torch.tensor([0, Tt], device=dev, dtype=torch.int32)
    do = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16)
    for _ in range(3):
        o = AttentionFn.apply(q, k, v, cu, Tt, True, 128 ** -0.5, None)
        torch.autograd.grad(o, (q, k, v), do)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(5):
            o = AttentionFn.apply(q, k, v, cu, Tt, True, 128 ** -0.5, None)
            torch.autograd.grad(o, (q, k, v), do)
        torch.cuda.synchronize()
    print("T", Tt)
    for e in prof.key_averages():
        if "attention" in e.key or "delta" in e.key:
            print(f"  {e.key[:60]:60s} {e.device_time_total / e.count:10.1f} us x{e.count}")
