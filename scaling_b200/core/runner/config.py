"""Runner / launcher configs (reference: core/runner/runner_config.py:1-73, launch_config.py:1-89)."""
from __future__ import annotations

import base64
import json
import os
from argparse import REMAINDER, ArgumentParser
from enum import Enum
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple

from pydantic import Field

from ..config import BaseConfig


class RunnerType(Enum):
    PDSH = "pdsh"
    PDSH_DOCKER = "pdsh_docker"


class RunnerDockerConfig(BaseConfig):
    docker_container: Optional[str] = Field(None, description="image to run")
    docker_sudo: bool = Field(False, description="prefix the docker command with sudo")
    docker_mounts: Optional[List[Tuple[str, str]]] = Field(None, description="(host dir, container dir) pairs to mount")


class RunnerConfig(BaseConfig, populate_by_name=True):
    runner_type: RunnerType = Field(RunnerType.PDSH, description="how remote nodes are reached")
    hostsfile: Optional[Path] = Field(None, description="MPI-style hostsfile (e.g. `worker-0 slots=4`)", alias="hostfile")
    hosts: Optional[List[str]] = Field(None, description="hosts as a list, alternative to hostsfile")
    master_port: int = Field(29500, description="port of the torch.distributed rendezvous")
    master_addr: Optional[str] = Field(None, description="address of node 0; inferred with `hostname -I` when omitted")
    script: Optional[Path] = Field(None, description="training script started once per device")
    default_gpu_count: int = Field(8, description="slots per host when a host line does not say")
    docker_config: RunnerDockerConfig = Field(RunnerDockerConfig(), description="docker settings for pdsh_docker")
    use_determined: bool = Field(False, description="metrics and checkpoints go through Determined")


def encode_base64(d: Dict[Any, Any]) -> str:
    return base64.urlsafe_b64encode(json.dumps(d).encode("utf-8")).decode("utf-8")


def decode_base64(s: str) -> Dict[Any, Any]:
    return json.loads(base64.urlsafe_b64decode(s))


class LaunchConfig(BaseConfig):
    """What one training process learns from its launcher: env ``MASTER_ADDR, MASTER_PORT, WORLD_SIZE, RANK,
    LOCAL_SLOT`` (``LOCAL_RANK`` accepted for torchrun) and the ``--payload`` base64 config."""

    master_port: int = Field(description="port of the torch.distributed rendezvous")
    master_addr: str = Field(description="address of the master node")
    world_size: int = Field(description="number of processes in the job")
    global_rank: int = Field(description="rank of this process")
    local_slot: int = Field(description="device index of this process on its node")
    payload: Optional[Dict[Any, Any]] = Field(None, description="decoded config payload")

    @classmethod
    def from_launcher_args(cls) -> "LaunchConfig":
        parser = ArgumentParser(description="process launch")
        parser.add_argument("--payload", type=str, default=None, help="base64 encoded payload")
        parser.add_argument("remaining_args", nargs=REMAINDER)
        args = parser.parse_args()
        return cls(
            master_addr=os.environ["MASTER_ADDR"], master_port=int(os.environ["MASTER_PORT"]),
            world_size=int(os.environ["WORLD_SIZE"]), global_rank=int(os.environ["RANK"]),
            local_slot=int(os.environ.get("LOCAL_SLOT", os.environ.get("LOCAL_RANK", "0"))),
            payload=None if args.payload is None else decode_base64(args.payload),
        )

    def overwrite_config_dict_with_launcher_args(self, config_dict: Dict[str, Any]) -> Dict[str, Any]:
        topo = config_dict.setdefault("topology", {})
        topo["world_size"] = self.world_size
        topo["global_rank"] = self.global_rank
        topo["local_slot"] = self.local_slot
        log_dir = config_dict.get("logger", {}).get("log_dir")
        if config_dict.get("profiler", {}).get("profiler_output") is None and log_dir is not None:
            config_dict.setdefault("profiler", {})["profiler_output"] = str(Path(log_dir) / "profile.json")
        return config_dict
