"""Per-node launcher: one training process per local slot, with supervision (reference: core/runner/launch.py).

    python -m scaling_b200.core.runner.launch --resource_pool=<b64> [--node_rank=N] --master_addr=A --master_port=P \
        <script> --payload <b64 config>

Each child gets ``MASTER_ADDR, MASTER_PORT, WORLD_SIZE, RANK, LOCAL_SLOT``.  The first child that exits non-zero takes
its siblings down and its code becomes the launcher's exit code; SIGINT/SIGTERM are forwarded.
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time
from typing import Any, List

from .config import decode_base64


def parse_args() -> argparse.Namespace:
    p = argparse.ArgumentParser(description="scaling_b200 node launcher")
    p.add_argument("--node_rank", type=int, default=0, help="index of this node in the resource pool")
    p.add_argument("--master_addr", type=str, default="127.0.0.1")
    p.add_argument("--master_port", type=int, default=29500)
    p.add_argument("--resource_pool", type=str, required=True, help="base64 json {nodes:[{name,slots}], world_size}")
    p.add_argument("script", type=str, help="training script")
    p.add_argument("--payload", type=str, default=None, help="base64 json config handed to the script")
    return p.parse_args()


def main() -> None:
    args = parse_args()
    pool = decode_base64(args.resource_pool)
    nodes = pool["nodes"]
    first_rank = sum(len(n["slots"]) for n in nodes[: args.node_rank])
    slots = nodes[args.node_rank]["slots"]
    world = sum(len(n["slots"]) for n in nodes)
    assert world == pool["world_size"], "derived world size differs from communicated world size"
    env = os.environ.copy()
    env.update(MASTER_ADDR=args.master_addr, MASTER_PORT=str(args.master_port), WORLD_SIZE=str(world))
    procs: List[subprocess.Popen] = []
    for i, slot in enumerate(slots):
        child_env = dict(env, RANK=str(first_rank + i), LOCAL_SLOT=str(slot))
        cmd = [sys.executable, "-u", args.script]
        if args.payload is not None:
            cmd += ["--payload", args.payload]
        procs.append(subprocess.Popen(cmd, env=child_env))

    def shutdown(code: int) -> None:
        for p in procs:
            if p.poll() is None:
                try:
                    p.kill()
                except Exception:  # noqa: BLE001
                    pass
        sys.exit(code)

    def on_signal(signum: int, _frame: Any) -> None:
        shutdown(1)

    signal.signal(signal.SIGINT, on_signal)
    signal.signal(signal.SIGTERM, on_signal)
    alive = list(procs)
    while alive:
        for p in list(alive):
            rc = p.poll()
            if rc is None:
                continue
            if rc != 0:
                shutdown(rc)
            alive.remove(p)
        time.sleep(0.2 if len(alive) else 0)


if __name__ == "__main__":
    main()
