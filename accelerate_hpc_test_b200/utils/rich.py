"""Optional rich tracebacks (reference src/accelerate/utils/rich.py)."""

from .imports import is_rich_available

if is_rich_available():
    from rich.traceback import install

    install(show_locals=False)
else:
    raise ModuleNotFoundError("To use the rich extension, install rich with `pip install rich`")
