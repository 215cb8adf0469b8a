"""Logger singleton, sinks and LoggerConfig (reference: core/logging/logging.py:46-209, logger_config.py:21-126).

One module-level ``logger`` proxy is re-configured per process with a rank name.  Sinks: coloured stream handler,
optional per-rank log file under a timestamped ``log_dir``, optional tensorboard / wandb / determined metrics, each with
its own rank allow-list (default: rank 0 only).  All third-party sinks are optional imports.
"""
from __future__ import annotations

import json
import logging as _pylog
import os
import socket
import sys
from datetime import datetime
from enum import Enum
from pathlib import Path
from typing import Any, Dict, List, Optional

from pydantic import Field, model_validator

from .config import BaseConfig


class LogLevel(Enum):
    DEBUG = "debug"
    INFO = "info"
    WARNING = "warning"
    ERROR = "error"
    CRITICAL = "critical"


_DATE_FORMATS = ("%Y-%m-%d-%H-%M-%S", "%Y-%m-%d", "%Y%m%d", "%Y-%m-%d_%H-%M-%S")


def is_date(text: str) -> bool:
    for fmt in _DATE_FORMATS:
        try:
            datetime.strptime(text, fmt)
            return True
        except ValueError:
            continue
    try:  # fall back to dateutil when present (the reference uses it)
        from dateutil.parser import parse

        parse(text, fuzzy=False)
        return True
    except Exception:  # noqa: BLE001
        return False


def _check_if_in_rank(target_rank: Optional[int], ranks: Optional[List[int]]) -> bool:
    """Name of this predicate in the reference (logger_config.py:106-111)."""
    return _rank_allowed(target_rank, ranks)


def _rank_allowed(rank: Optional[int], ranks: Optional[List[int]]) -> bool:
    if rank is None:
        return False
    return rank in ranks if ranks is not None else rank == 0


class LoggerConfig(BaseConfig):
    log_level: LogLevel = Field(LogLevel.INFO, description="minimum level that is emitted")
    log_dir: Optional[Path] = Field(None, description="directory for per-rank log files; a timestamp folder is appended")
    metrics_ranks: Optional[List[int]] = Field(None, description="global ranks that print metrics (default: rank 0)")
    use_wandb: bool = Field(False, description="log metrics to Weights & Biases")
    wandb_ranks: Optional[List[int]] = Field(None, description="global ranks that write to wandb (default: rank 0)")
    wandb_host: str = Field("https://api.wandb.ai", description="url of the wandb host")
    wandb_team: str = Field("aleph-alpha", description="wandb team (entity)")
    wandb_project: str = Field("aleph-alpha-scaling", description="wandb project")
    wandb_group: str = Field("debug", description="wandb group; a timestamp is appended")
    wandb_api_key: Optional[str] = Field(None, description="wandb api key; falls back to $WANDB_API_KEY")
    use_tensorboard: bool = Field(False, description="log metrics to tensorboard")
    tensorboard_ranks: Optional[List[int]] = Field(None, description="global ranks that write tensorboard files")
    determined_metrics_ranks: Optional[List[int]] = Field(None, description="global ranks that report to determined")

    @model_validator(mode="before")
    @classmethod
    def _timestamp_paths(cls, values: Dict[Any, Any]) -> Dict[Any, Any]:
        values = dict(values)
        now = datetime.now().strftime("%Y-%m-%d-%H-%M-%S")
        if values.get("log_dir") is not None:
            d = Path(values["log_dir"])
            values["log_dir"] = d if is_date(d.name) else d / now
        group = values.get("wandb_group")
        if group is not None and not is_date("-".join(group.split("-")[-6:])) and not is_date(group.split("-")[-1]):
            values["wandb_group"] = f"{group}-{now}"
        if values.get("use_wandb") and not values.get("wandb_api_key"):
            values["wandb_api_key"] = os.getenv("WANDB_API_KEY")
        return values

    @model_validator(mode="after")
    def _need_wandb_key(self) -> "LoggerConfig":
        if self.use_wandb and not self.wandb_api_key:
            raise ValueError("If 'use_wandb' is set to True a wandb api key needs to be provided.")
        return self

    def is_rank_in_tensorboard_ranks(self, rank: Optional[int]) -> bool:
        return _rank_allowed(rank, self.tensorboard_ranks)

    def is_rank_in_wandb_ranks(self, rank: Optional[int]) -> bool:
        return _rank_allowed(rank, self.wandb_ranks)

    def is_rank_in_metrics_ranks(self, rank: Optional[int]) -> bool:
        return _rank_allowed(rank, self.metrics_ranks)

    def is_rank_in_determined_metrics_ranks(self, rank: Optional[int]) -> bool:
        return _rank_allowed(rank, self.determined_metrics_ranks)


class _Color(_pylog.Formatter):
    COLORS = {"DEBUG": "\x1b[38;20m", "INFO": "\x1b[34;20m", "WARNING": "\x1b[33;20m", "ERROR": "\x1b[31;20m",
              "CRITICAL": "\x1b[31;1m"}

    def format(self, record: _pylog.LogRecord) -> str:
        base = f"[%(asctime)s] [%(levelname)s] %(message)s"
        return _pylog.Formatter(self.COLORS.get(record.levelname, "") + base + "\x1b[0m").format(record)


class Logger:
    """Concrete logger for one process."""

    def __init__(self, config: LoggerConfig, name: Optional[str] = None, global_rank: Optional[int] = None,
                 use_determined: bool = False, determined_context: Any = None):
        self.config = config
        self.name = name
        self.global_rank = global_rank
        self._use_determined = use_determined
        self.determined_context = determined_context
        self._log = _pylog.getLogger(f"scaling_b200.{name or 'main'}.{id(self)}")
        self._log.propagate = False
        self._log.setLevel(getattr(_pylog, config.log_level.value.upper()))
        handler = _pylog.StreamHandler(sys.stdout)
        handler.setFormatter(_Color())
        self._log.addHandler(handler)
        if config.log_dir is not None:
            config.log_dir.mkdir(parents=True, exist_ok=True)
            fh = _pylog.FileHandler(config.log_dir / f"log_{(name or 'main').replace(' ', '_')}.log")
            fh.setFormatter(_pylog.Formatter("[%(asctime)s] [%(levelname)s] %(message)s"))
            self._log.addHandler(fh)
        self._tb = None
        if config.use_tensorboard and config.log_dir is not None and config.is_rank_in_tensorboard_ranks(global_rank):
            try:
                from torch.utils.tensorboard import SummaryWriter

                self._tb = SummaryWriter(log_dir=str(config.log_dir / "tensorboard"))
            except Exception as e:  # noqa: BLE001
                self._log.warning(f"tensorboard unavailable: {e}")
        self._wandb = None
        if config.use_wandb and config.is_rank_in_wandb_ranks(global_rank):
            try:
                import wandb

                os.environ.setdefault("WANDB_BASE_URL", config.wandb_host)
                wandb.login(key=config.wandb_api_key)
                wandb.init(project=config.wandb_project, entity=config.wandb_team, group=config.wandb_group,
                           name=f"{socket.gethostname()}-{global_rank}")
                self._wandb = wandb
            except Exception as e:  # noqa: BLE001
                self._log.warning(f"wandb unavailable: {e}")

    def _fmt(self, msg: object) -> str:
        return f"[{self.name}] {msg}" if self.name else str(msg)

    def debug(self, msg: object) -> None:
        self._log.debug(self._fmt(msg))

    def info(self, msg: object) -> None:
        self._log.info(self._fmt(msg))

    def warning(self, msg: object) -> None:
        self._log.warning(self._fmt(msg))

    def error(self, msg: object) -> None:
        self._log.error(self._fmt(msg))

    def critical(self, msg: object) -> None:
        self._log.critical(self._fmt(msg))

    def log_metrics(self, metrics: Dict[str, Any], step: int) -> None:
        if self.config.is_rank_in_metrics_ranks(self.global_rank):
            self.info(json.dumps({"step": step, **{k: v for k, v in metrics.items()}}, default=float))
        if self._tb is not None:
            for k, v in metrics.items():
                if v is not None:
                    self._tb.add_scalar(k, v, step)
        if self._wandb is not None:
            self._wandb.log(metrics, step=step)
        if self._use_determined and self.determined_context is not None and \
                self.config.is_rank_in_determined_metrics_ranks(self.global_rank):
            train = {k: v for k, v in metrics.items() if not k.startswith("evaluation/")}
            evalm = {k: v for k, v in metrics.items() if k.startswith("evaluation/")}
            self.determined_context.train.report_training_metrics(steps_completed=step, metrics=train)
            if evalm:
                self.determined_context.train.report_validation_metrics(steps_completed=step, metrics=evalm)

    def log_config(self, config: BaseConfig) -> None:
        self.log_config_dict(config.as_dict())

    def log_config_dict(self, config_dict: Dict[str, Any]) -> None:
        if self._wandb is not None:
            self._wandb.config.update(config_dict, allow_val_change=True)
        if self._tb is not None:
            self._tb.add_text("config", json.dumps(config_dict, default=str))
        self.debug(f"config: {json.dumps(config_dict, default=str)}")


class DeterminedLogger(Logger):
    pass


class _LoggerProxy:
    """Module-level singleton; ``configure`` swaps the implementation for the current process."""

    def __init__(self) -> None:
        self._impl: Logger = Logger(LoggerConfig())

    def configure(self, config: LoggerConfig, name: Optional[str] = None, global_rank: Optional[int] = None) -> None:
        self._impl = Logger(config, name=name, global_rank=global_rank)

    def configure_determined(self, config: LoggerConfig, name: Optional[str] = None, global_rank: Optional[int] = None,
                             determined_context: Any = None) -> None:
        self._impl = DeterminedLogger(config, name=name, global_rank=global_rank, use_determined=True,
                                      determined_context=determined_context)

    def __getattr__(self, item: str) -> Any:
        return getattr(self._impl, item)


logger = _LoggerProxy()
