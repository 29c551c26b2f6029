"""Helpers for testing code written against this package: skip decorators keyed on hardware and optional libraries, launch-command
builders, subprocess runners that surface the child's stderr, tiny regression workloads, and the launched scripts ``accelerate test``
runs (``scripts/``).  Everything ``testing.py`` defines is offered at package level — the names the reference's ``accelerate.test_utils``
offers (``require_cuda``, ``slow``, ``execute_subprocess_async``, ``torch_device``, ...) — plus the regression workload."""

import importlib as _importlib

from . import testing as _testing
from .training import RegressionDataset, RegressionModel

# what `testing` defines itself (functions, decorators, classes) ...
globals().update({name: value for name, value in vars(_testing).items()
                  if not name.startswith("_") and getattr(value, "__module__", None) == _testing.__name__})
# ... and the values it computes once at import: the device under test and how to launch onto it
for _name in ("torch_device", "device_count", "memory_allocated_func", "DEFAULT_LAUNCH_COMMAND"):
    globals()[_name] = getattr(_testing, _name)

_LAUNCHED_SCRIPTS = ("test_script", "test_sync", "test_ops")


def __getattr__(name):
    """``test_utils.test_script`` / ``test_sync`` / ``test_ops`` load on first use: importing them with the package would put them in
    ``sys.modules`` before ``python -m ...scripts.test_script`` runs them."""
    if name in _LAUNCHED_SCRIPTS:
        return _importlib.import_module(f"{__name__}.scripts.{name}")
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
