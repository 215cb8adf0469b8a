"""Optional Determined-AI glue (reference: core/determined/core.py, core/utils/determined_utils.py).  ``determined`` is
not required: every entry point degrades to a no-op when the package is missing."""
from __future__ import annotations

import contextlib
import faulthandler
from typing import Any, Iterator, Optional


def determined_available() -> bool:
    try:
        import determined  # type: ignore # noqa: F401

        return True
    except ImportError:
        return False


@contextlib.contextmanager
def init(**kwargs: Any) -> Iterator[Optional[Any]]:
    """``with init() as core_context`` — a Determined core context, or ``None`` outside a Determined cluster."""
    if not determined_available():
        yield None
        return
    import determined as det  # type: ignore

    with det.core.init(**kwargs) as ctx:
        yield ctx


def maybe_periodic_stacktraces(enabled: bool, interval_seconds: int = 30) -> None:
    """Dump all thread stacks periodically (debugging hangs)."""
    if enabled:
        faulthandler.dump_traceback_later(interval_seconds, repeat=True)
