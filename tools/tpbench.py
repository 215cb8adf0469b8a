"""Fused GEMM+all-reduce vs GEMM followed by NCCL all-reduce, at the flagship model's TP shapes.
Run: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/tpbench.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaling_b200.ops import kernels as K  # noqa: E402
from scaling_b200.parallel.symm import TPComm  # noqa: E402


try:
    PEAK_TFLOPS = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                   "MEASURED_PEAKS.json")))["bf16_tflops"])
except Exception:  # noqa: BLE001
    PEAK_TFLOPS = 1650.0


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def phase_profile(comm, shapes, rank):
    """In-kernel %globaltimer stamps of the fused kernel's phases (sb_gemm2_set_profile_buffer), per shape: microseconds
    after the earliest CTA's kernel entry, median and max over CTAs of this rank."""
    from scaling_b200.ops import _lib

    lib = _lib.load(required=True)
    buf = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")
    names = ["entry", "last_acc_ready", "last_push_complete", "last_flag_published", "first_tile_flags_seen",
             "last_tile_flags_seen", "reducer_done"]
    out = {}
    extra = dict(shapes)
    extra["one_wave_attn_out"] = (1024, 4096, 2048, False)
    for name, (M, N, Kd, b_mn) in extra.items():
        a = (torch.randn(M, Kd, device="cuda") * 0.5).bfloat16()
        b = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda") * 0.05).bfloat16()
        for _ in range(5):
            K.gemm_allreduce(a, b, comm, b_mn=b_mn)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        assert lib.sb_gemm2_set_profile_buffer(buf.data_ptr()) == 0
        buf.zero_()
        for _ in range(3):  # the third, steady-state call's stamps are the ones left in the buffer
            K.gemm_allreduce(a, b, comm, b_mn=b_mn)
        torch.cuda.synchronize()
        lib.sb_gemm2_set_profile_buffer(None)
        t = buf.view(148, 8).cpu().double()
        live = t[:, 0] > 0
        t = t[live]
        t0 = t[:, 0].min()
        rel = (t[:, :7] - t0) / 1e3
        out[name] = {n: {"median": float(rel[:, i].median()), "max": float(rel[:, i].max())} for i, n in enumerate(names)}
        out[name]["ctas"] = int(live.sum())
        if rank == 0:
            print("phases", name, json.dumps(out[name]), flush=True)
    return out


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    comm = TPComm(dist.group.WORLD, torch.device("cuda", rank))
    tp = world
    T, H, F, V = 4096, 4096, 14336, 32000
    shapes = {
        "attn_out_fprop": (T, H, H // tp, False),
        "mlp_down_fprop": (T, H, F // tp, False),
        "qkv_dgrad": (T, H, (H + 2 * 1024) // tp, True),
        "mlp_up_dgrad": (T, H, 2 * F // tp, True),
        "lm_head_dgrad": (T, H, V // tp, True),
    }
    res = {}
    ok = True
    for name, (M, N, Kd, b_mn) in shapes.items():
        g = torch.Generator(device="cuda").manual_seed(1234 + rank)
        a = (torch.randn(M, Kd, device="cuda", generator=g) * 0.5).bfloat16()
        b = (torch.randn((Kd, N) if b_mn else (N, Kd), device="cuda", generator=g) * 0.05).bfloat16()
        # numerics: fp32 reference of the summed product
        ref = (a.float() @ (b.float() if b_mn else b.float().t()))
        dist.all_reduce(ref)
        out = K.gemm_allreduce(a, b, comm, b_mn=b_mn)
        torch.cuda.synchronize()
        err = (out.float() - ref).abs().max().item() / ref.abs().max().item()
        gathered = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(gathered, out)
        same = all(torch.equal(gathered[0], x) for x in gathered)
        ok = ok and err < 2e-2 and same

        def base():
            y = K.gemm(a, b, b_mn=b_mn)
            dist.all_reduce(y)
            return y

        def base_cublas():
            y = a @ (b if b_mn else b.t())
            dist.all_reduce(y)
            return y

        t_f = timed(lambda: K.gemm_allreduce(a, b, comm, b_mn=b_mn))
        t_b = timed(base)
        t_c = timed(base_cublas)
        t_g = timed(lambda: K.gemm(a, b, b_mn=b_mn))
        t_cg = timed(lambda: a @ (b if b_mn else b.t()))
        # roofline of the fused op: the slower of the GEMM at the measured cuBLAS burst rate and this rank's NVLink
        # egress at 900 GB/s per direction (one-shot push: the partial goes to every other rank)
        flops = 2.0 * M * N * Kd
        nvlink_bytes = (tp - 1) * M * N * 2
        t_roof = max(flops / (PEAK_TFLOPS * 1e12), nvlink_bytes / 900e9) * 1e3
        res[name] = {"M": M, "N": N, "K": Kd, "fused_ms": t_f, "gemm_plus_nccl_ms": t_b, "cublas_plus_nccl_ms": t_c,
                     "gemm_only_ms": t_g, "cublas_only_ms": t_cg, "speedup_vs_gemm_plus_nccl": t_b / t_f,
                     "speedup_vs_cublas_plus_nccl": t_c / t_f, "exposed_over_gemm_only": t_f / t_g - 1.0,
                     "nvlink_bytes_out_per_rank": nvlink_bytes, "roofline_ms": t_roof, "fraction_of_roofline": t_roof / t_f,
                     "rel_err": err, "bit_identical": same}
        if rank == 0:
            print(name, json.dumps(res[name]), flush=True)
    phases = phase_profile(comm, shapes, rank) if "--profile" in sys.argv else None
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        if phases is not None:
            json.dump({"tp": tp, "phases_us": phases}, open(f"gpurun_out/tpbench_phases_tp{tp}.json", "w"), indent=1)
        json.dump({"tp": tp, "shapes": res, "ok": ok}, open(f"gpurun_out/tpbench_tp{tp}.json", "w"), indent=1)
        print("OK" if ok else "NUMERICS MISMATCH", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
