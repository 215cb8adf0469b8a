"""Probe the multi-GPU plumbing on the box: NCCL all-reduce latency at the model's sizes, symmetric memory
(peer pointers, multicast) availability, raw cudaIpc peer mapping."""
import os
import sys
import time

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    out = {}
    for mb in (1, 8, 32, 64):
        x = torch.randn(mb * 1024 * 1024 // 2, device="cuda", dtype=torch.bfloat16)
        for _ in range(5):
            dist.all_reduce(x)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            dist.all_reduce(x)
        e1.record()
        torch.cuda.synchronize()
        out[f"nccl_allreduce_{mb}MB_us"] = e0.elapsed_time(e1) / 20 * 1e3
    try:
        import torch.distributed._symmetric_memory as symm_mem

        g = dist.group.WORLD
        try:
            symm_mem.enable_symm_mem_for_group(g.group_name)
        except Exception as e:  # noqa: BLE001
            out["enable_err"] = repr(e)[:200]
        t = symm_mem.empty(1024 * 1024, dtype=torch.bfloat16, device="cuda")
        h = symm_mem.rendezvous(t, g.group_name)
        out["symm_rank"] = h.rank
        out["symm_world"] = h.world_size
        out["buffer_ptrs"] = [hex(p) for p in h.buffer_ptrs]
        out["signal_pad_ptrs"] = [hex(p) for p in h.signal_pad_ptrs]
        out["signal_pad_size"] = h.signal_pad_size
        try:
            out["multicast_ptr"] = hex(h.multicast_ptr)
        except Exception as e:  # noqa: BLE001
            out["multicast_err"] = repr(e)[:200]
        # peer write through the mapped pointer: rank r fills the peer's buffer with r+1
        t.zero_()
        h.barrier()
        peer = (rank + 1) % world
        pt = h.get_buffer(peer, (1024 * 1024,), torch.bfloat16)
        pt.fill_(rank + 1)
        torch.cuda.synchronize()
        h.barrier()
        torch.cuda.synchronize()
        out["peer_write_ok"] = bool((t == ((rank - 1) % world) + 1).all().item())
        # p2p copy bandwidth
        big = symm_mem.empty(64 * 1024 * 1024, dtype=torch.bfloat16, device="cuda")
        hb = symm_mem.rendezvous(big, g.group_name)
        src = torch.randn(64 * 1024 * 1024, device="cuda", dtype=torch.bfloat16)
        dst = hb.get_buffer(peer, (64 * 1024 * 1024,), torch.bfloat16)
        for _ in range(3):
            dst.copy_(src)
        torch.cuda.synchronize()
        hb.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dst.copy_(src)
        e1.record()
        torch.cuda.synchronize()
        out["p2p_push_GBps"] = 128 * 1024 * 1024 * 10 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    except Exception as e:  # noqa: BLE001
        import traceback

        out["symm_err"] = traceback.format_exc()[-1500:]
    print(f"[rank {rank}] {out}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
