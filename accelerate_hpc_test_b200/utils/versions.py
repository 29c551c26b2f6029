"""Version comparison helpers (reference src/accelerate/utils/versions.py:1-56)."""

import importlib.metadata
from typing import Union

from packaging.version import Version, parse

from .constants import STR_OPERATION_TO_FUNC

try:
    import torch

    torch_version = parse(importlib.metadata.version("torch"))
except Exception:  # pragma: no cover
    torch_version = parse("0")


def compare_versions(library_or_version: Union[str, Version], operation: str, requirement_version: str) -> bool:
    """``compare_versions("torch", ">=", "2.6")`` or with an already parsed Version on the left."""
    if operation not in STR_OPERATION_TO_FUNC:
        raise ValueError(f"`operation` must be one of {list(STR_OPERATION_TO_FUNC)}, received {operation}")
    fn = STR_OPERATION_TO_FUNC[operation]
    if isinstance(library_or_version, str):
        library_or_version = parse(importlib.metadata.version(library_or_version))
    return fn(library_or_version, parse(requirement_version))


def is_torch_version(operation: str, version: str) -> bool:
    return compare_versions(torch_version, operation, version)
