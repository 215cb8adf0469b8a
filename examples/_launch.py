"""Shared front door of the examples: `python -m examples.<name>.run <config.yml>` parses the YAML into the example's
config class and hands the validated dict to the job runner, which starts one training process per device."""
from __future__ import annotations

import sys
from typing import Type

from scaling_b200.core import BaseConfig, runner_main
from scaling_b200.core.logging import logger


def launch(config_class: Type[BaseConfig], argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if len(argv) != 1 or argv[0] in ("-h", "--help"):
        print(f"usage: python -m {sys.modules['__main__'].__spec__.name if sys.modules['__main__'].__spec__ else 'run'} CONFIG.yml")
        return 2
    cfg = config_class.from_yaml(argv[0])
    logger.configure(config=cfg.logger, name="runner")
    return runner_main(cfg.runner, payload=cfg.as_dict())
