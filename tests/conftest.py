"""pytest configuration: registers the ``gpu`` marker and resets framework singletons between tests."""

import os
import sys

# debug_launcher forks the test process: keep torch single-threaded so no OpenMP pool exists to deadlock the children
os.environ.setdefault("OMP_NUM_THREADS", "1")
os.environ.setdefault("MKL_NUM_THREADS", "1")

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run with -m gpu on a B200 box)")
    config.addinivalue_line("markers", "slow: long-running test")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _reset_state():
    yield
    try:
        from accelerate_hpc_test_b200.state import AcceleratorState, GradientState, PartialState

        AcceleratorState._reset_state(True)
        GradientState._reset_state()
        PartialState._reset_state()
    except Exception:
        pass
