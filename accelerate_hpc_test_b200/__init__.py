"""accelerate_hpc_test_b200 — a Blackwell (B200, sm_100a) native training / inference launcher with the
public API of Hugging Face Accelerate (reference: hang-zou/accelerate-hpc-test, src/accelerate/__init__.py:16-47).

    import accelerate_hpc_test_b200 as accelerate
    accelerator = accelerate.Accelerator(mixed_precision="bf16")

The framework layer (``Accelerator``, ``PartialState``, dataloaders, checkpointing, CLI) keeps the
reference's surface; the hot path underneath is hand-written sm_100a CUDA (``ops/``, ``csrc/``) and
native parallel engines over peer-mapped NVLink memory (``parallel/``).
"""

__version__ = "0.2.0"

from .accelerator import Accelerator
from .data_loader import skip_first_batches
from .parallelism_config import ParallelismConfig
from .state import AcceleratorState, GradientState, PartialState
from .utils import (
    AutocastKwargs,
    DataLoaderConfiguration,
    DDPCommunicationHookType,
    DeepSpeedPlugin,
    DistributedDataParallelKwargs,
    DistributedType,
    FullyShardedDataParallelPlugin,
    GradScalerKwargs,
    InitProcessGroupKwargs,
    ProfileKwargs,
    find_executable_batch_size,
    is_rich_available,
    set_seed,
    synchronize_rng_states,
)

_LAZY = {
    "init_empty_weights": "big_modeling", "init_on_device": "big_modeling", "cpu_offload": "big_modeling",
    "cpu_offload_with_hook": "big_modeling", "disk_offload": "big_modeling", "dispatch_model": "big_modeling",
    "load_checkpoint_and_dispatch": "big_modeling", "infer_auto_device_map": "utils.modeling",
    "load_checkpoint_in_model": "utils.modeling", "prepare_pippy": "inference", "debug_launcher": "launchers",
    "notebook_launcher": "launchers", "LocalSGD": "local_sgd",
}


def __getattr__(name):
    if name in _LAZY:
        import importlib

        mod = importlib.import_module(f"{__name__}.{_LAZY[name]}")
        return getattr(mod, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


if is_rich_available():  # pragma: no cover
    from .utils import rich  # noqa: F401
