"""Progress bars that stay quiet on every process except the first one of each node.

``tqdm(iterable, main_process_only=True, **tqdm_kwargs)``; parity target: reference ``src/accelerate/utils/tqdm.py:25-43``.
"""

from ._lazy_state import PartialState
from .imports import is_tqdm_available


def tqdm(*args, main_process_only: bool = True, **kwargs):
    if not is_tqdm_available():
        raise ImportError("Accelerate's `tqdm` module requires `tqdm` to be installed. Please run `pip install tqdm`.")
    if args and isinstance(args[0], bool):
        # the pre-0.23 call style tqdm(True, iterable) is gone
        raise ValueError("Passing `True` or `False` as the first argument is no longer supported; use `main_process_only=`.")
    import tqdm.auto as _auto

    silent = bool(kwargs.pop("disable", False))
    if main_process_only and not silent:
        silent = PartialState().local_process_index != 0
    kwargs["disable"] = silent
    return _auto.tqdm(*args, **kwargs)
