"""Late-bound access to ``PartialState`` for the ``utils`` modules.

``state.py`` imports ``utils.dataclasses`` (which runs ``utils/__init__``), so the utils modules cannot
import ``state`` at import time; they resolve it on first call instead.
"""


def PartialState(*args, **kwargs):
    from ..state import PartialState as _PartialState

    return _PartialState(*args, **kwargs)
