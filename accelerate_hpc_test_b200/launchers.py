"""In-process launchers: ``notebook_launcher`` and ``debug_launcher``.

Parity with reference src/accelerate/launchers.py:41-309.  ``notebook_launcher`` forks ``num_processes``
workers through torchelastic (so ``max_restarts`` / ``monitor_interval`` / rendezvous options work as
with ``torchrun``); ``debug_launcher`` forks a CPU-only gloo job that rendezvous through a temp file —
the "cluster on a laptop" every distributed unit test of this repo runs on.
"""

from __future__ import annotations

import os
import sys
import tempfile

import torch

from .state import AcceleratorState, PartialState
from .utils.environment import are_libraries_initialized, patch_environment
from .utils.launch import PrepareForLaunch
from .utils.other import get_free_port, is_port_in_use


def notebook_launcher(function, args=(), num_processes=None, mixed_precision="no", use_port="29500", master_addr="127.0.0.1",
                      node_rank=0, num_nodes=1, rdzv_backend="static", rdzv_endpoint="", rdzv_conf=None, rdzv_id="none",
                      max_restarts=0, monitor_interval=0.1, log_line_prefix_template=None):
    """Run ``function(*args)`` on ``num_processes`` GPUs (forked from the current process), or in-process
    when only one device is requested/available."""
    in_colab = "google.colab" in sys.modules
    in_kaggle = "KAGGLE_KERNEL_RUN_TYPE" in os.environ
    if num_processes is None:
        num_processes = torch.cuda.device_count() if torch.cuda.is_available() else 1
    if num_processes is not None and int(num_processes) > 1:
        from torch.distributed.launcher.api import LaunchConfig, elastic_launch
        from torch.multiprocessing.spawn import ProcessRaisedException

        if len(AcceleratorState._shared_state) > 0:
            raise ValueError(
                "To launch a multi-GPU training from your notebook, the `Accelerator` should only be initialized "
                "inside your training function. Restart your notebook and make sure no cells initializes an `Accelerator`."
            )
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            raise ValueError(
                "To launch a multi-GPU training from your notebook, you need to avoid running any instruction "
                "using `torch.cuda` in any cell. Restart your notebook and make sure no cells use any CUDA function."
            )
        problematic_imports = are_libraries_initialized("bitsandbytes")
        if len(problematic_imports) > 0:
            raise RuntimeError(
                "Could not start distributed process. Libraries known to initialize CUDA upon import have been imported "
                f"already: {problematic_imports}. Please keep these imports inside your training function."
            )
        if use_port == "0" or int(use_port) == 0:
            use_port = str(get_free_port())
        elif is_port_in_use(int(use_port)):
            raise ConnectionError(f"Port {use_port} is already in use; pass another `use_port` (or 0 for a free one).")
        patched_env = dict(nproc=num_processes, node_rank=node_rank, world_size=num_nodes * num_processes,
                           master_addr=master_addr, master_port=use_port, mixed_precision=mixed_precision)
        with patch_environment(**patched_env):
            os.environ["ACCELERATE_MIXED_PRECISION"] = str(mixed_precision)
            use_cpu = not torch.cuda.is_available()
            if use_cpu:
                os.environ["ACCELERATE_USE_CPU"] = "yes"
            launcher = PrepareForLaunch(function, distributed_type="MULTI_CPU" if use_cpu else "MULTI_GPU")
            print(f"Launching training on {num_processes} {'CPUs' if use_cpu else 'GPUs'}.")
            try:
                if rdzv_conf is None:
                    rdzv_conf = {}
                if rdzv_backend == "static":
                    rdzv_conf["rank"] = node_rank
                    if not rdzv_endpoint:
                        rdzv_endpoint = f"{master_addr}:{use_port}"
                launch_config_kwargs = dict(
                    min_nodes=num_nodes, max_nodes=num_nodes, nproc_per_node=num_processes, run_id=rdzv_id, rdzv_endpoint=rdzv_endpoint,
                    rdzv_backend=rdzv_backend, rdzv_configs=rdzv_conf, max_restarts=max_restarts, monitor_interval=monitor_interval,
                    start_method="fork",
                )
                if log_line_prefix_template is not None:
                    launch_config_kwargs["log_line_prefix_template"] = log_line_prefix_template
                elastic_launch(config=LaunchConfig(**launch_config_kwargs), entrypoint=function)(*args)
            except ProcessRaisedException as e:
                if "Cannot re-initialize CUDA in forked subprocess" in e.args[0]:
                    raise RuntimeError(
                        "CUDA has been initialized before the `notebook_launcher` could create a forked subprocess. "
                        "This likely stems from an outside import causing issues once the `notebook_launcher()` is called. "
                        "Please review your imports and test them when running the `notebook_launcher()` to identify "
                        "which one is problematic and causing CUDA to be initialized."
                    ) from e
                raise RuntimeError(f"An issue was found when launching the training: {e}") from e
            finally:
                os.environ.pop("ACCELERATE_USE_CPU", None) if use_cpu else None
    else:
        if torch.cuda.is_available():
            print("Launching training on one GPU.")
        else:
            print("Launching training on CPU.")
        function(*args)


def debug_launcher(function, args=(), num_processes=2):
    """Run ``function(*args)`` on ``num_processes`` CPU processes over gloo with a file rendezvous."""
    from torch.multiprocessing import start_processes

    with tempfile.NamedTemporaryFile() as tmp_file:
        with patch_environment(world_size=num_processes, master_addr="127.0.0.1", master_port="29500", accelerate_mixed_precision="no",
                               accelerate_debug_rdv_file=tmp_file.name, accelerate_use_cpu="yes"):
            launcher = PrepareForLaunch(function, debug=True)
            start_processes(launcher, args=args, nprocs=num_processes, start_method="fork")
