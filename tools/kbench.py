"""Device-timed micro-benchmarks of the sm_100a kernels against their library counterparts (cuBLAS via torch.matmul,
ATen element-wise).  Writes JSON to gpurun_out/kbench.json.  CUDA events, 3 warm-ups, L2 flushed between iterations."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from scaling_b200.ops import kernels as K  # noqa: E402

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


def bench_gemm(name, M, N, Kd, a_mn=False, b_mn=False, out_dtype=torch.bfloat16, accumulate=False):
    a = torch.randn((Kd, M) if a_mn else (M, Kd), device=dev, dtype=torch.bfloat16)
    b = torch.randn((Kd, N) if b_mn else (N, Kd), device=dev, dtype=torch.bfloat16)
    out = torch.zeros(M, N, device=dev, dtype=out_dtype)
    ours = timeit(lambda: K.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, accumulate=accumulate))
    A = a.t() if a_mn else a
    B = b if b_mn else b.t()
    if accumulate and out_dtype == torch.float32:
        tmp = torch.empty(M, N, device=dev, dtype=torch.bfloat16)

        def lib():
            torch.matmul(A, B, out=tmp)
            out.add_(tmp)
    elif accumulate:
        lib = lambda: torch.addmm(out, A, B, out=out)  # noqa: E731
    else:
        o2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        lib = lambda: torch.matmul(A, B, out=o2)  # noqa: E731
    ref = timeit(lib)
    fl = 2.0 * M * N * Kd
    return {"name": name, "M": M, "N": N, "K": Kd, "ours_ms": ours, "lib_ms": ref, "ours_tflops": fl / ours / 1e9,
            "lib_tflops": fl / ref / 1e9, "speedup_vs_lib": ref / ours}


import os
K.set_gemm_impl(os.environ.get("SCALING_B200_GEMM", "2cta"))
res = {"gemm": [], "fused": [], "gemm_impl": os.environ.get("SCALING_B200_GEMM", "2cta")}
T = 4096
shapes = [
    ("qkv_fprop", T, 6144, 4096, False, False, torch.bfloat16, False),
    ("attn_out_fprop", T, 4096, 4096, False, False, torch.bfloat16, False),
    ("mlp_down_fprop", T, 4096, 14336, False, False, torch.bfloat16, False),
    ("lm_head_fprop", T, 32000, 4096, False, False, torch.bfloat16, False),
    ("qkv_dgrad", T, 4096, 6144, False, True, torch.bfloat16, False),
    ("mlp_up_dgrad", T, 4096, 28672, False, True, torch.bfloat16, False),
    ("mlp_down_dgrad", T, 14336, 4096, False, True, torch.bfloat16, False),
    ("qkv_wgrad_fp32acc", 6144, 4096, T, True, True, torch.float32, True),
    ("mlp_up_wgrad_fp32acc", 28672, 4096, T, True, True, torch.float32, True),
    ("mlp_down_wgrad_fp32acc", 4096, 14336, T, True, True, torch.float32, True),
    ("square_8192", 8192, 8192, 8192, False, False, torch.bfloat16, False),
    # per-rank shapes at tensor parallel 2 (N or K halved)
    ("tp2_qkv_fprop", T, 3072, 4096, False, False, torch.bfloat16, False),
    ("tp2_attn_out_fprop", T, 4096, 2048, False, False, torch.bfloat16, False),
    ("tp2_mlp_down_fprop", T, 4096, 7168, False, False, torch.bfloat16, False),
    ("tp2_lm_head_fprop", T, 16000, 4096, False, False, torch.bfloat16, False),
    ("tp2_qkv_dgrad", T, 4096, 3072, False, True, torch.bfloat16, False),
    ("tp2_mlp_up_dgrad", T, 4096, 14336, False, True, torch.bfloat16, False),
    ("tp2_mlp_down_dgrad", T, 7168, 4096, False, True, torch.bfloat16, False),
    ("tp2_qkv_wgrad_fp32acc", 3072, 4096, T, True, True, torch.float32, True),
    ("tp2_mlp_up_wgrad_fp32acc", 14336, 4096, T, True, True, torch.float32, True),
    ("tp2_mlp_down_wgrad_fp32acc", 4096, 7168, T, True, True, torch.float32, True),
]
for s in shapes:
    try:
        r = bench_gemm(*s)
    except Exception as e:  # noqa: BLE001
        r = {"name": s[0], "error": str(e)[:200]}
    print(r, flush=True)
    res["gemm"].append(r)

# dual GEMM + SwiGLU epilogue vs 2 cuBLAS GEMMs + silu*mul
x = torch.randn(T, 4096, device=dev, dtype=torch.bfloat16)
w = torch.randn(2 * 14336, 4096, device=dev, dtype=torch.bfloat16) / 64
ours = timeit(lambda: K.gemm_swiglu(x, w, keep_pre=True))
ours_nopre = timeit(lambda: K.gemm_swiglu(x, w, keep_pre=False))


def lib_swiglu():
    g = x @ w[:14336].t()
    u = x @ w[14336:].t()
    return torch.nn.functional.silu(g) * u


ref = timeit(lib_swiglu)
fl = 2.0 * T * 2 * 14336 * 4096
r = {"name": "mlp_up_swiglu", "ours_ms": ours, "ours_nopre_ms": ours_nopre, "lib_ms": ref, "ours_tflops": fl / ours / 1e9,
     "lib_tflops": fl / ref / 1e9, "speedup_vs_lib": ref / ours}
print(r, flush=True)
res["fused"].append(r)

# RMSNorm fwd
xx = torch.randn(T, 4096, device=dev, dtype=torch.bfloat16)
ww = torch.ones(4096, device=dev, dtype=torch.bfloat16)
ours = timeit(lambda: K.norm_fwd(xx, ww, None, 1e-5, True))


def lib_rms():
    xf = xx.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5)).to(torch.bfloat16) * ww


ref = timeit(lib_rms)
r = {"name": "rmsnorm_fwd_4096x4096", "ours_ms": ours, "lib_ms": ref, "gbs": 2 * xx.numel() * 2 / ours / 1e6, "speedup_vs_lib": ref / ours}
print(r, flush=True)
res["fused"].append(r)

# cross entropy
logits = torch.randn(T, 32000, device=dev, dtype=torch.bfloat16)
tg = torch.randint(0, 32000, (T,), device=dev)
lse = torch.zeros(T, device=dev)
cf = torch.ones(T, device=dev)
ours = timeit(lambda: (K.xent_stats(logits, tg, 0), K.xent_grad_(logits, tg, lse, cf, None, 0)))


def lib_ce():
    lf = logits.float().requires_grad_(True)
    torch.nn.functional.cross_entropy(lf, tg).backward()


ref = timeit(lib_ce)
r = {"name": "xent_fwd_bwd_4096x32000", "ours_ms": ours, "lib_ms": ref, "speedup_vs_lib": ref / ours}
print(r, flush=True)
res["fused"].append(r)

# AdamW on 256M params
n = 256 * 1024 * 1024
p = torch.randn(n, device=dev)
m = torch.zeros_like(p)
v = torch.zeros_like(p)
g = torch.randn(n, device=dev, dtype=torch.bfloat16)
o = torch.empty(n, device=dev, dtype=torch.bfloat16)
ours = timeit(lambda: K.adamw_step(p, m, v, g, o, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1, step=3), iters=5)
pp = torch.nn.Parameter(p.clone())
opt = torch.optim.AdamW([pp], lr=1e-3, fused=True)
pp.grad = torch.randn(n, device=dev)


def lib_adam():
    opt.step()
    o.copy_(pp.data)


ref = timeit(lib_adam, iters=5)
r = {"name": "adamw_256M", "ours_ms": ours, "lib_ms": ref, "gbs": n * (4 * 3 * 2 + 2 + 2) / ours / 1e6, "speedup_vs_lib": ref / ours}
print(r, flush=True)
res["fused"].append(r)

# LM head with the CE statistics in the GEMM epilogue (+ combine / finalize) vs plain GEMM + the separate stats kernel
xh = torch.randn(T, 4096, device=dev, dtype=torch.bfloat16)
wh = torch.randn(32000, 4096, device=dev, dtype=torch.bfloat16) / 64


def fused_head():
    lg, part = K.gemm_xent_stats(xh, wh)
    return K.xent_from_partials(part, lg, tg, 0)


def unfused_head():
    lg = K.gemm(xh, wh)
    return K.xent_stats(lg, tg, 0)


ours = timeit(fused_head)
ref = timeit(unfused_head)


def lib_head():
    return torch.nn.functional.cross_entropy((xh @ wh.t()).float(), tg, reduction="none")


lib = timeit(lib_head)
r = {"name": "lm_head_plus_ce_forward_4096x32000x4096", "epilogue_stats_ms": ours, "gemm_plus_stats_kernel_ms": ref,
     "cublas_plus_aten_ce_ms": lib, "speedup_vs_unfused": ref / ours, "speedup_vs_lib": lib / ours}
print(r, flush=True)
res["fused"].append(r)

# LoRA: low-rank pair as extra K-slab vs base GEMM + two small GEMMs + add (rank 64, q-projection shape)
xl = torch.randn(T, 4096, device=dev, dtype=torch.bfloat16)
wl = torch.randn(4096, 4096, device=dev, dtype=torch.bfloat16) / 64
al = torch.randn(64, 4096, device=dev, dtype=torch.bfloat16) / 64
bl = torch.randn(4096, 64, device=dev, dtype=torch.bfloat16) / 8


def lora_fused():
    t = K.gemm(xl, al)
    return K.gemm_lowrank(xl, wl, t, bl)


def lora_unfused():
    return K.gemm(xl, wl) + K.gemm(K.gemm(xl, al), bl)


def lora_lib():
    return xl @ wl.t() + (xl @ al.t()) @ bl.t()


ours, ref, lib = timeit(lora_fused), timeit(lora_unfused), timeit(lora_lib)
r = {"name": "lora_q_proj_rank64_4096x4096x4096", "fused_ms": ours, "own_gemms_plus_add_ms": ref, "cublas_plus_add_ms": lib,
     "speedup_vs_unfused": ref / ours, "speedup_vs_lib": lib / ours}
print(r, flush=True)
res["fused"].append(r)

Path("gpurun_out").mkdir(exist_ok=True)
json.dump(res, open("gpurun_out/kbench.json", "w"), indent=1)

# attention forward / backward vs flash-attn (library)
try:
    from flash_attn import flash_attn_varlen_func

    from scaling_b200.ops.attention_sm100 import AttentionFn

    res["attention"] = []
    for (Tt, heads, kvh) in [(4096, 32, 8), (8192, 32, 8)]:
        q = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
        k = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
        v = torch.randn(Tt, kvh, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
        cu = torch.tensor([0, Tt], device=dev, dtype=torch.int32)
        sc = 128 ** -0.5
        do = torch.randn(Tt, heads, 128, device=dev, dtype=torch.bfloat16)
        ours_f = timeit(lambda: AttentionFn.apply(q, k, v, cu, Tt, True, sc, None))
        lib_f = timeit(lambda: flash_attn_varlen_func(q, k, v, cu, cu, Tt, Tt, 0.0, sc, causal=True))
        o1 = AttentionFn.apply(q, k, v, cu, Tt, True, sc, None)
        o2 = flash_attn_varlen_func(q, k, v, cu, cu, Tt, Tt, 0.0, sc, causal=True)
        ours_b = timeit(lambda: torch.autograd.grad(o1, (q, k, v), do, retain_graph=True))
        lib_b = timeit(lambda: torch.autograd.grad(o2, (q, k, v), do, retain_graph=True))
        fl = 4.0 * Tt * Tt * heads * 128 / 2
        r = {"name": f"attn_causal_T{Tt}_h{heads}_kv{kvh}", "fwd_ours_ms": ours_f, "fwd_lib_ms": lib_f, "fwd_ours_tflops": fl / ours_f / 1e9,
             "fwd_lib_tflops": fl / lib_f / 1e9, "bwd_ours_ms": ours_b, "bwd_lib_ms": lib_b, "bwd_ours_tflops": 2.5 * fl / ours_b / 1e9,
             "bwd_lib_tflops": 2.5 * fl / lib_b / 1e9}
        print(r, flush=True)
        res["attention"].append(r)
    json.dump(res, open("gpurun_out/kbench.json", "w"), indent=1)
except Exception as e:  # noqa: BLE001
    print("attention bench failed:", repr(e))
