"""Multi-process test harness: spawn N ranks with the launcher's env contract (reference tests/core/utils.py:225-307)."""
from __future__ import annotations

import os
import socket
import traceback
from typing import Any, Callable, Dict

import torch
import torch.multiprocessing as mp


from scaling_b200.core.utils.port import find_free_port  # noqa: E402,F401  (ports below the ephemeral range)


def _worker(rank: int, world: int, port: int, fn: Callable, kwargs: Dict[str, Any], results: Any) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), WORLD_SIZE=str(world), RANK=str(rank),
                      LOCAL_SLOT=str(rank))
    try:
        torch.set_num_threads(1)
        results[rank] = ("ok", fn(rank=rank, world_size=world, master_port=port, **kwargs))
    except Exception:  # noqa: BLE001
        results[rank] = ("error", traceback.format_exc())
    finally:
        try:
            import torch.distributed as dist

            if dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


def dist_launcher(fn: Callable, world_size: int, timeout: int = 240, **kwargs: Any) -> Dict[int, Any]:
    """Run ``fn(rank=, world_size=, master_port=, **kwargs)`` on ``world_size`` spawned processes; returns
    {rank: return value}.  Any failing rank fails the test with its traceback."""
    if world_size == 1:
        port = find_free_port()
        results: Dict[int, Any] = {}
        _worker(0, 1, port, fn, kwargs, results)
    else:
        ctx = mp.get_context("spawn")
        manager = ctx.Manager()
        results = manager.dict()
        port = find_free_port()
        procs = [ctx.Process(target=_worker, args=(r, world_size, port, fn, kwargs, results)) for r in range(world_size)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout)
        hung = [p for p in procs if p.is_alive()]
        for p in hung:
            p.terminate()
        results = dict(results)
        assert not hung, f"{len(hung)} worker(s) did not finish within {timeout}s; results so far: {results}"
    failures = [f"rank {r} failed:\n{results[r][1]}" for r in range(world_size) if r in results and results[r][0] != "ok"]
    # the root cause is usually not the rank that reports "connection closed by peer"
    failures.sort(key=lambda f: ("Connection closed" in f) or ("Connection reset" in f))
    assert not failures, "\n\n".join(failures)
    for r in range(world_size):
        assert r in results, f"rank {r} returned nothing"
    return {r: v[1] for r, v in results.items()}
