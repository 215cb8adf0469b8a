import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from scaling_b200.ops.attention_sm100 import AttentionFn  # noqa: E402

Tt = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
heads, kvh = 32, 8
dev = "cuda"
q = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
cu = torch.tensor([0, Tt], device=dev, dtype=torch.int32)
do = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16)
for _ in range(2):
    o = AttentionFn.apply(q, k, v, cu, Tt, True, 128 ** -0.5, None)
    torch.autograd.grad(o, (q, k, v), do)
torch.cuda.synchronize()
print("ok")
