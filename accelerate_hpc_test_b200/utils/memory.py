"""Device-memory housekeeping and the out-of-memory retry decorator.

* ``clear_device_cache`` / ``release_memory`` — drop references, run the garbage collector, return cached blocks to the driver;
* ``should_reduce_batch_size`` — does this exception mean "out of memory" (CUDA, cuDNN workspace or host allocator)?
* ``find_executable_batch_size`` — decorator for ``train(batch_size, ...)``: call it with ``starting_batch_size`` and, each
  time it dies of OOM, again with 90 % of the previous size, until it fits or the size reaches zero.

Parity target: reference ``utils/memory.py:39-180`` (CUDA and CPU only in this build).
"""

from __future__ import annotations

import functools
import gc
import inspect

import torch

_OOM_TEXT = (
    "CUDA out of memory.",
    "CUDA error: out of memory",
    "out of memory",
    "cuDNN error: CUDNN_STATUS_NOT_SUPPORTED.",
    "DefaultCPUAllocator: can't allocate memory",
)


def clear_device_cache(garbage_collection: bool = False) -> None:
    if garbage_collection:
        gc.collect()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()


def release_memory(*objects):
    """``a, b = release_memory(a, b)``: every name is rebound to ``None`` and the freed blocks leave the caching allocator."""
    cleared = [None for _ in objects]
    del objects
    clear_device_cache(garbage_collection=True)
    return cleared


def should_reduce_batch_size(exception: Exception) -> bool:
    if isinstance(exception, torch.cuda.OutOfMemoryError):
        return True
    if isinstance(exception, RuntimeError) and len(exception.args) == 1:
        message = str(exception.args[0])
        return any(marker in message for marker in _OOM_TEXT)
    return False


class _BatchSizeSearch:
    """State of one decorated function: the current candidate and how to shrink it."""

    def __init__(self, function, start: int, shrink):
        self.function = function
        self.batch_size = start
        self.shrink = shrink or self._by_ten_percent
        self.parameters = list(inspect.signature(function).parameters)

    def _by_ten_percent(self) -> int:
        return int(self.batch_size * 0.9)

    def _reject_explicit_batch_size(self, args) -> None:
        # the decorator supplies the first argument; a caller that passes it too has one positional too many
        if len(self.parameters) < len(args) + 1:
            shown = ", ".join(f"{name}={value}" for name, value in zip(self.parameters[1:], args[1:]))
            raise TypeError(
                f"Batch size was passed into `{self.function.__name__}` as the first argument when called."
                f"Remove this as the decorator already does so: `{self.function.__name__}({shown})`"
            )

    def __call__(self, *args, **kwargs):
        clear_device_cache(garbage_collection=True)
        self._reject_explicit_batch_size(args)
        while self.batch_size > 0:
            try:
                return self.function(self.batch_size, *args, **kwargs)
            except Exception as error:  # noqa: BLE001 - anything that is not an OOM is re-raised below
                if not should_reduce_batch_size(error):
                    raise
            clear_device_cache(garbage_collection=True)
            self.batch_size = self.shrink()
        raise RuntimeError("No executable batch size found, reached zero.")


def find_executable_batch_size(function: callable = None, starting_batch_size: int = 128, reduce_batch_size_fn: callable = None):
    """Use bare (``@find_executable_batch_size``) or configured (``@find_executable_batch_size(starting_batch_size=64)``).
    ``reduce_batch_size_fn()`` may replace the 10 % rule; it returns the next size to try."""
    if function is None:
        return functools.partial(find_executable_batch_size, starting_batch_size=starting_batch_size, reduce_batch_size_fn=reduce_batch_size_fn)
    search = _BatchSizeSearch(function, starting_batch_size, reduce_batch_size_fn)

    @functools.wraps(function)
    def retrying(*args, **kwargs):
        return search(*args, **kwargs)

    return retrying


def get_xpu_available_memory(device_index: int):  # pragma: no cover - API parity only
    raise NotImplementedError("XPU devices are not supported by this build")
