"""Rank-aware logger (reference ``src/neuronx_distributed/utils/logger.py:16-112``).

Env: ``NXD_LOG_LEVEL`` in {off, fatal, error, warning, info, debug, trace}
(default info), ``NXD_LOG_HIDE_TIME`` drops the timestamp.
"""
from __future__ import annotations

import logging
import os
import sys

_LEVELS = {
    "trace": logging.DEBUG - 5,
    "debug": logging.DEBUG,
    "info": logging.INFO,
    "warning": logging.WARNING,
    "error": logging.ERROR,
    "fatal": logging.FATAL,
    "off": logging.FATAL + 10,
}
logging.addLevelName(_LEVELS["trace"], "TRACE")


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


def get_log_level() -> int:
    name = os.environ.get("NXD_LOG_LEVEL", "info").lower()
    if name not in _LEVELS:
        raise ValueError(f"NXD_LOG_LEVEL={name!r} is not one of {sorted(_LEVELS)}")
    return _LEVELS[name]


def _is_rank0() -> bool:
    try:
        import torch.distributed as dist

        return (not dist.is_available()) or (not dist.is_initialized()) or dist.get_rank() == 0
    except Exception:
        return True


class _Rank0Logger:
    """Forwards to a ``logging.Logger`` only on global rank 0."""

    def __init__(self, inner: logging.Logger, rank0_only: bool):
        self._inner = inner
        self._rank0_only = rank0_only

    def _ok(self) -> bool:
        return (not self._rank0_only) or _is_rank0()

    def trace(self, msg, *a, **k):
        if self._ok():
            self._inner.log(_LEVELS["trace"], msg, *a, **k)

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
    inner.setLevel(get_log_level())
    out = _Rank0Logger(inner, rank0_only)
    _LOGGERS[key] = out
    return out
