"""Attention-only micro-benchmark (forward/backward vs flash-attn)."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from flash_attn import flash_attn_varlen_func  # noqa: E402

from scaling_b200.ops.attention_sm100 import AttentionFn  # noqa: E402

dev = "cuda"
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


out = []
for (Tt, heads, kvh, causal) in [(4096, 32, 8, True), (4096, 16, 4, True), (8192, 32, 8, True), (16384, 16, 4, True), (4096, 32, 8, False)]:
    q = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    cu = torch.tensor([0, Tt], device=dev, dtype=torch.int32)
    sc = 128 ** -0.5
    do = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16)
    ours_f = timeit(lambda: AttentionFn.apply(q, k, v, cu, Tt, causal, sc, None))
    lib_f = timeit(lambda: flash_attn_varlen_func(q, k, v, cu, cu, Tt, Tt, 0.0, sc, causal=causal))
    o1 = AttentionFn.apply(q, k, v, cu, Tt, causal, sc, None)
    o2 = flash_attn_varlen_func(q, k, v, cu, cu, Tt, Tt, 0.0, sc, causal=causal)
    err = float((o1.float() - o2.float()).norm() / o2.float().norm())
    ours_b = timeit(lambda: torch.autograd.grad(o1, (q, k, v), do, retain_graph=True))
    lib_b = timeit(lambda: torch.autograd.grad(o2, (q, k, v), do, retain_graph=True))
    fl = 4.0 * Tt * Tt * heads * 128 / (2 if causal else 1)
    r = {"name": f"attn_T{Tt}_h{heads}_kv{kvh}_{'causal' if causal else 'full'}", "rel_err_vs_flash": err,
         "fwd_ours_ms": ours_f, "fwd_lib_ms": lib_f, "fwd_ours_tflops": fl / ours_f / 1e9, "fwd_lib_tflops": fl / lib_f / 1e9,
         "bwd_ours_ms": ours_b, "bwd_lib_ms": lib_b, "bwd_ours_tflops": 2.5 * fl / ours_b / 1e9, "bwd_lib_tflops": 2.5 * fl / lib_b / 1e9}
    print(r, flush=True)
    out.append(r)
Path("gpurun_out").mkdir(exist_ok=True)
json.dump(out, open("gpurun_out/abench.json", "w"), indent=1)
