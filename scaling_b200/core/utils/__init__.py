from .port import find_free_port  # noqa: F401
