from .config import LaunchConfig, RunnerConfig, RunnerDockerConfig, RunnerType, decode_base64, encode_base64  # noqa: F401
from .runner import PDSHRunner, get_resource_pool, parse_host, runner_main  # noqa: F401
