"""Version predicates: ``compare_versions("torch", ">=", "2.6")`` and ``is_torch_version(">=", "2.6")``."""

from __future__ import annotations

import importlib.metadata
from typing import Union

from packaging.version import Version, parse

from .constants import STR_OPERATION_TO_FUNC


def _installed(package: str) -> Version:
    return parse(importlib.metadata.version(package))


try:
    torch_version = _installed("torch")
except importlib.metadata.PackageNotFoundError:  # pragma: no cover - torch is a hard dependency
    torch_version = parse("0")


def compare_versions(library_or_version: Union[str, Version], operation: str, requirement_version: str) -> bool:
    """Left side: a package name (its installed version is looked up) or an already parsed ``Version``."""
    try:
        test = STR_OPERATION_TO_FUNC[operation]
    except KeyError:
        raise ValueError(f"`operation` must be one of {list(STR_OPERATION_TO_FUNC)}, received {operation}") from None
    left = _installed(library_or_version) if isinstance(library_or_version, str) else library_or_version
    return test(left, parse(requirement_version))


def is_torch_version(operation: str, version: str) -> bool:
    return compare_versions(torch_version, operation, version)
