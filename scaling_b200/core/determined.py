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


@contextlib.contextmanager
def maybe_periodic_stacktraces(enabled: bool, interval_seconds: int = 30) -> Iterator[None]:
    """While the block runs, dump all thread stacks every ``interval_seconds`` (debugging hangs)."""
    if enabled:
        faulthandler.dump_traceback_later(interval_seconds, repeat=True)
    try:
        yield
    finally:
        if enabled:
            faulthandler.cancel_dump_traceback_later()


@contextlib.contextmanager
def determined_profiler_from_ctx(core_context: Any, experiment_config: Any, info: Any) -> Iterator[Optional[Any]]:
    """Determined's system-metrics profiler agent for this trial, or ``None`` when Determined (or its profiler API,
    which newer releases dropped) is not available."""
    try:
        from determined.profiler import ProfilerAgent  # type: ignore
    except Exception:  # noqa: BLE001
        yield None
        return
    agent = ProfilerAgent.from_env(info, core_context.distributed.rank if core_context is not None else 0,
                                   core_context.distributed.local_rank if core_context is not None else 0)
    with agent:
        yield agent
