"""``tqdm`` that only draws on the local main process (reference src/accelerate/utils/tqdm.py:25-43)."""

import warnings

from ._lazy_state import PartialState
from .imports import is_tqdm_available


def tqdm(*args, main_process_only: bool = True, **kwargs):
    if not is_tqdm_available():
        raise ImportError("Accelerate's `tqdm` module requires `tqdm` to be installed. Please run `pip install tqdm`.")
    from tqdm.auto import tqdm as _tqdm

    if len(args) > 0 and isinstance(args[0], bool):
        raise ValueError("Passing `True` or `False` as the first argument is no longer supported; use `main_process_only=`.")
    disable = kwargs.pop("disable", False)
    if main_process_only and not disable:
        disable = PartialState().local_process_index != 0
    return _tqdm(*args, **kwargs, disable=disable)
