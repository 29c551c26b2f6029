"""Rank-aware logging.

``get_logger(__name__)`` returns a ``logging.LoggerAdapter`` whose emit methods take two extra keywords:

* ``main_process_only`` (default ``True``) — only global rank 0 writes the record;
* ``in_order`` (default ``False``) — every rank writes, one after the other in rank order, separated by
  barriers (so the output of several ranks does not interleave).

The process state must exist (``PartialState()`` / ``Accelerator()``) before the first record, because the
adapter needs the rank.  ``ACCELERATE_LOG_LEVEL`` sets the level when ``log_level`` is not given.
Parity target: reference ``src/accelerate/logging.py:23-126``.
"""

from __future__ import annotations

import logging
import os

from .state import PartialState

_STATE_MISSING = (
    "You must initialize the accelerate state by calling either `PartialState()` or `Accelerator()` "
    "before using the logging utility."
)


class MultiProcessAdapter(logging.LoggerAdapter):
    """``LoggerAdapter`` that decides per record which ranks emit it."""

    def __init__(self, logger, extra=None):
        super().__init__(logger, extra or {})
        self._warned = set()

    @staticmethod
    def _should_log(main_process_only: bool) -> bool:
        if not main_process_only:
            return True
        return bool(PartialState().is_main_process)

    def _emit(self, level, msg, args, kwargs):
        msg, kwargs = self.process(msg, kwargs)
        self.logger.log(level, msg, *args, **kwargs)

    def log(self, level, msg, *args, **kwargs):
        if not PartialState._shared_state:
            raise RuntimeError(_STATE_MISSING)
        only_main = kwargs.pop("main_process_only", True)
        ordered = kwargs.pop("in_order", False)
        # report the caller of .info()/.warning(), not this adapter: _emit <- log <- (logging internals, not counted) <- caller
        kwargs.setdefault("stacklevel", 3)
        if not self.isEnabledFor(level):
            return
        if ordered:
            state = PartialState()
            for turn in range(state.num_processes):
                if turn == state.process_index:
                    self._emit(level, msg, args, dict(kwargs))
                state.wait_for_everyone()
            return
        if self._should_log(only_main):
            self._emit(level, msg, args, kwargs)

    def warning_once(self, msg, *args, **kwargs):
        """Like ``warning`` but a given (message, args) pair is written at most once per adapter."""
        key = (msg, args)
        if key in self._warned:
            return
        self._warned.add(key)
        kwargs.setdefault("stacklevel", 4)
        self.warning(msg, *args, **kwargs)


def get_logger(name: str, log_level: str | None = None) -> MultiProcessAdapter:
    level = log_level if log_level is not None else os.environ.get("ACCELERATE_LOG_LEVEL")
    base = logging.getLogger(name)
    if level is not None:
        level = level.upper()
        base.setLevel(level)
        base.root.setLevel(level)
    return MultiProcessAdapter(base, {})
