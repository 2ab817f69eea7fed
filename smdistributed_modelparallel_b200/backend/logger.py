"""``smp`` logger: stdout, ``[ts: LEVEL file:line] msg`` format, environment-driven filters.

Environment (same names as the reference, smdistributed/modelparallel/backend/logger.py:13-122):
``SMP_LOG_LEVEL`` (debug/info/warning/error/off), ``SMP_LOG_HIDE_TIME``, ``SMP_LOG_ALLOW_FILES`` and
``SMP_LOG_BLOCK_FILES`` (comma-separated path fragments, mutually exclusive).
"""

from __future__ import annotations

import logging
import os
import sys

_LEVELS = {"debug": logging.DEBUG, "info": logging.INFO, "warning": logging.WARNING, "warn": logging.WARNING,
           "error": logging.ERROR, "critical": logging.CRITICAL, "off": logging.CRITICAL + 10}
_logger = None
_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class _PathFilter(logging.Filter):
    def __init__(self, allow, block):
        super().__init__()
        self.allow, self.block = allow, block

    def filter(self, record):
        rel = os.path.relpath(record.pathname, _PKG_ROOT) if record.pathname.startswith(_PKG_ROOT) else record.pathname
        record.relpath = rel
        if self.allow:
            return any(a in rel for a in self.allow)
        return not any(b in rel for b in self.block)


def _csv(name):
    return [s.strip() for s in os.environ.get(name, "").split(",") if s.strip()]


def get_logger() -> logging.Logger:
    global _logger
    if _logger is not None:
        return _logger
    from .exceptions import LoggingConfigError

    level_name = os.environ.get("SMP_LOG_LEVEL", "info").lower()
    if level_name not in _LEVELS:
        raise LoggingConfigError(f"SMP_LOG_LEVEL must be one of {sorted(_LEVELS)}, got {level_name!r}")
    allow, block = _csv("SMP_LOG_ALLOW_FILES"), _csv("SMP_LOG_BLOCK_FILES")
    if allow and block:
        raise LoggingConfigError("SMP_LOG_ALLOW_FILES and SMP_LOG_BLOCK_FILES cannot both be set")
    lg = logging.getLogger("smp")
    lg.setLevel(_LEVELS[level_name])
    lg.propagate = False
    if not lg.handlers:
        h = logging.StreamHandler(sys.stdout)
        hide_time = os.environ.get("SMP_LOG_HIDE_TIME", "").lower() in ("1", "true")
        fmt = "[%(levelname).1s %(relpath)s:%(lineno)d] %(message)s" if hide_time else \
            "[%(asctime)s.%(msecs)03d: %(levelname).1s %(relpath)s:%(lineno)d] %(message)s"
        h.setFormatter(logging.Formatter(fmt, datefmt="%Y-%m-%d %H:%M:%S"))
        h.addFilter(_PathFilter(allow, block))
        lg.addHandler(h)
    _logger = lg
    return lg


def rmsg(msg: str) -> str:
    """Prefix a message with this process' rank coordinates (reference torch/utils.py:209-210)."""
    try:
        from .. import state

        if state.initialized:
            c = state.core
            return f"[{c.rank()}: pp {c.pp_rank()} tp {c.tp_rank()} rdp {c.rdp_rank()}] {msg}"
    except Exception:
        pass
    return msg
