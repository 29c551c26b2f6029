"""Importing this module switches uncaught-exception output to ``rich`` tracebacks (locals hidden).

Opt-in side-effect module, mirroring the reference's ``accelerate.utils.rich`` import hook.
"""

from .imports import is_rich_available

if not is_rich_available():
    raise ModuleNotFoundError("To use the rich extension, install rich with `pip install rich`")

import rich.traceback as _rich_traceback

_rich_traceback.install(show_locals=False)
