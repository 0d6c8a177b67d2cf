"""Rank-aware logger (reference ``src/neuronx_distributed/utils/logger.py:16-112``).

Env: ``NXD_LOG_LEVEL`` in {off, fatal, error, warning, info, debug, trace}
(default info), ``NXD_LOG_HIDE_TIME`` drops the timestamp.
"""
from __future__ import annotations

import logging
import os
import sys
from functools import lru_cache, wraps
from typing import Callable, Optional

# "trace" opens the DEBUG level as in the reference; ``logger.trace(...)`` lines are emitted (at DEBUG) only then
_LEVELS = {
    "trace": logging.DEBUG,
    "debug": logging.DEBUG,
    "info": logging.INFO,
    "warning": logging.WARNING,
    "error": logging.ERROR,
    "fatal": logging.FATAL,
    "off": logging.FATAL + 1,
}


class PackagePathFilter(logging.Filter):
    """Adds ``record.relativepath`` = the source path relative to the first ``sys.path`` entry that contains it, for
    compact ``%(relativepath)s:%(lineno)d`` log formats (reference ``utils/logger.py:16-32``)."""

    def filter(self, record: logging.LogRecord) -> bool:
        path = record.pathname
        record.relativepath = path
        for root in sorted((p for p in sys.path if p), key=len, reverse=True):
            root = root if root.endswith(os.sep) else root + os.sep
            if path.startswith(root):
                record.relativepath = os.path.relpath(path, root)
                break
        return True
_LOGGERS = {}


@lru_cache
def get_log_level() -> int:
    """``NXD_LOG_LEVEL`` -> ``logging`` level; cached (``get_log_level.cache_clear()`` after changing the variable)."""
    name = os.environ.get("NXD_LOG_LEVEL", "info").lower()
    if name not in _LEVELS:
        raise ValueError(f"Allowed NXD_LOG_LEVELS are: info, trace, debug, warning, error, fatal, off. Got: {name}")
    return _LEVELS[name]


def _trace_enabled() -> bool:
    return os.environ.get("NXD_LOG_LEVEL", "info").lower() == "trace"


def _is_rank0() -> bool:
    """Global rank 0: the process group's rank once it exists, the launcher's ``RANK`` variable before that."""
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank() == 0
    except Exception:  # noqa: BLE001
        pass
    return int(os.environ.get("RANK", "0")) == 0


def _rank0_only(fn: Callable, default=None, **extra_kwargs) -> Callable:
    """``fn`` that runs on global rank 0 only and returns ``default`` elsewhere (reference ``logger.py:91-112``)."""
    @wraps(fn)
    def wrapped_fn(*args, **kwargs):
        if _is_rank0():
            return fn(*args, **{**kwargs, **extra_kwargs})
        return default

    return wrapped_fn


class _Rank0Logger:
    """Forwards to a ``logging.Logger`` only on global rank 0."""

    def __init__(self, inner: logging.Logger, rank0_only: bool):
        self._inner = inner
        self._rank0_only = rank0_only

    def _ok(self) -> bool:
        return (not self._rank0_only) or _is_rank0()

    def trace(self, msg, *a, **k):
        if self._ok() and _trace_enabled():
            self._inner.debug(msg, *a, **k)

    def __getattr__(self, name):
        attr = getattr(self._inner, name)
        if name in ("debug", "info", "warning", "warn", "error", "critical", "fatal", "exception", "log"):
            def gated(*a, **k):
                if self._ok():
                    return attr(*a, **k)
            return gated
        return attr


def get_logger(name: str = "nxd_b200", rank0_only: bool = True) -> _Rank0Logger:
    key = (name, rank0_only)
    if key in _LOGGERS:
        return _LOGGERS[key]
    inner = logging.getLogger(name)
    if not inner.handlers:
        handler = logging.StreamHandler(sys.stderr)
        fmt = "[%(levelname)s] %(name)s: %(message)s"
        if not os.environ.get("NXD_LOG_HIDE_TIME"):
            fmt = "%(asctime)s " + fmt
        handler.setFormatter(logging.Formatter(fmt))
        inner.addHandler(handler)
        inner.propagate = False
    level = get_log_level()
    if level > logging.FATAL:
        inner.disabled = True           # NXD_LOG_LEVEL=off
    else:
        inner.setLevel(level)
    out = _Rank0Logger(inner, rank0_only)
    _LOGGERS[key] = out
    return out
