"""Job runner: resource pool → (pdsh | docker) command line → per-node launcher (reference: core/runner/runner.py).

``runner_main(config, payload)`` resolves the hosts (list, hostsfile, or the local devices), finds the master address,
forwards NCCL*/PYTHON*/UCX*/MV2* environment variables and starts ``python -m scaling_b200.core.runner.launch`` on
every node with the whole validated config as a base64 payload.  On a single local node no pdsh/ssh is involved.
"""
from __future__ import annotations

import collections
import os
import subprocess
import sys
from shlex import quote
from typing import Any, Dict, List, Tuple

import torch

from .config import RunnerConfig, RunnerType, encode_base64

PDSH_MAX_FAN_OUT = 1024
EXPORT_ENVS = ["NCCL", "PYTHON", "MV2", "UCX", "SCALING_B200", "CUDA_VISIBLE_DEVICES"]
ENVIRONMENT_PATH = os.path.join(os.path.expanduser("~"), ".deepspeed_env")
LOCAL = ("localhost", "127.0.0.1")


def parse_host(config: RunnerConfig, host_str: str) -> Tuple[str, List[int]]:
    """``name`` | ``name slots=N`` | ``name slots=0,2,5``."""
    parts = host_str.split(" ")
    if len(parts) == 1:
        return parts[0], list(range(config.default_gpu_count))
    if len(parts) != 2:
        raise ValueError(f"cannot parse host '{host_str}'")
    name, slots = parts
    assert slots.startswith("slots="), "second part of host should start with 'slots'; e.g. 'nodename slots=8'"
    value = slots.split("=", 1)[1]
    assert len(value) > 0, f"no slots defined in '{host_str}'"
    if "," in value:
        return name, [int(s) for s in value.split(",") if s]
    return name, list(range(int(value)))


def hosts_str_to_resource_pool(config: RunnerConfig, hosts: List[str]) -> "collections.OrderedDict[str, List[int]]":
    pool: "collections.OrderedDict[str, List[int]]" = collections.OrderedDict()
    for host in hosts:
        name, slots = parse_host(config, host)
        if name in pool:
            raise ValueError(f"host '{name}' defined more than once")
        pool[name] = slots
    if not pool:
        n = torch.cuda.device_count()
        assert n > 0, "would use local resources for training without hosts defined; no local gpus found"
        pool["localhost"] = list(range(n))
    return pool


def get_resource_pool(config: RunnerConfig) -> Dict[str, Any]:
    if config.hostsfile is not None and config.hosts is not None:
        raise ValueError("Both hostsfile and hosts cannot be defined at the same time.")
    if config.hostsfile is not None:
        if not config.hostsfile.is_file():
            raise ValueError(f"Hostsfile not found: {config.hostsfile}")
        hosts = [line.strip() for line in config.hostsfile.read_text().splitlines() if line.strip()]
    else:
        hosts = list(config.hosts or [])
    pool = hosts_str_to_resource_pool(config, hosts)
    return {"nodes": [{"name": n, "slots": s} for n, s in pool.items()], "world_size": sum(len(s) for s in pool.values())}


class PDSHRunner:
    def __init__(self, master_addr: str, config: RunnerConfig, resource_pool: Dict[str, Any], payload: Dict[str, Any],
                 exports: Dict[str, str], environment: Dict[str, str]):
        self.master_addr, self.config, self.resource_pool = master_addr, config, resource_pool
        self.payload, self.exports, self.environment = payload, exports, environment

    def get_cmd(self) -> List[str]:
        nodes = self.resource_pool["nodes"]
        use_pdsh = not (len(nodes) == 1 and nodes[0]["name"] in LOCAL)
        docker = self.config.runner_type == RunnerType.PDSH_DOCKER
        cmd: List[str] = []
        if use_pdsh:
            cmd = ["pdsh", "-f", str(PDSH_MAX_FAN_OUT), "-w", ",".join(n["name"] for n in nodes)]
        if docker:
            dc = self.config.docker_config
            cmd += (["sudo"] if dc.docker_sudo else []) + ["docker", "run", "--rm", "--privileged"]
            for k, v in self.exports.items():
                if not k.lower().startswith("python"):
                    cmd += ["--env", f"{k}={quote(v)}"]
            for host_dir, container_dir in dc.docker_mounts or []:
                cmd += ["-v", f"{host_dir}:{container_dir}"]
            cmd += ["--network=host", "--gpus", "all", "--ipc=host", "--name=scaling_b200.core", "--ulimit", "memlock=-1",
                    "--ulimit", "stack=67108864", str(dc.docker_container)]
        elif use_pdsh:
            cmd += ["".join(f"export {k}={quote(v)}; " for k, v in self.exports.items()), f"cd {os.path.abspath('')};"]
        cmd += ["python" if docker else sys.executable, "-u", "-m", "scaling_b200.core.runner.launch",
                f"--resource_pool={encode_base64(self.resource_pool)}"]
        if use_pdsh:
            cmd += ["--node_rank=%n"]
        cmd += [f"--master_addr={self.master_addr}", f"--master_port={self.config.master_port}",
                str(self.config.script.absolute()), "--payload", encode_base64(self.payload)]
        return cmd


def runner_main(config: RunnerConfig, payload: Dict[str, Any]) -> int:
    assert config.script is not None, "must provide a script to run"
    pool = get_resource_pool(config)
    assert len(pool["nodes"]) > 0, "could not identify resources"
    if config.master_addr is not None:
        master_addr = config.master_addr
    else:
        first = pool["nodes"][0]["name"]
        if first in LOCAL:
            master_addr = "127.0.0.1"
        else:
            master_addr = subprocess.check_output([f"ssh {first} hostname -I"], shell=True).decode("utf-8").split()[0]
    env = os.environ.copy()
    here = os.path.abspath("")
    env["PYTHONPATH"] = here + (":" + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
    exports = {k.strip(): v.strip() for k, v in env.items() if any(k.startswith(p) for p in EXPORT_ENVS)}
    if os.path.isfile(ENVIRONMENT_PATH):
        for line in open(ENVIRONMENT_PATH):
            if "=" in line:
                k, v = line.split("=", 1)
                exports[k] = v.strip()
    cmd = PDSHRunner(master_addr, config, pool, payload, exports, env).get_cmd()
    proc = subprocess.Popen(cmd, env=env)
    proc.wait()
    if proc.returncode > 0:
        sys.exit(proc.returncode)
    return proc.returncode
