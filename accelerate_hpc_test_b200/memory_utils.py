"""Deprecated location of ``find_executable_batch_size`` (reference src/accelerate/memory_utils.py): import it from
``accelerate_hpc_test_b200.utils.memory``."""

import warnings

warnings.warn(
    "memory_utils has been reorganized to utils.memory. Import `find_executable_batch_size` from the main `__init__`: "
    "`from accelerate_hpc_test_b200 import find_executable_batch_size` to avoid this warning.",
    FutureWarning,
)

from .utils.memory import find_executable_batch_size  # noqa: E402,F401
