"""Launching a distributed job from inside a running Python process.

``notebook_launcher`` is for notebooks: it forks one worker per GPU out of the current interpreter through torchelastic, so restarts,
monitoring and rendezvous options behave as with ``torchrun``.  Forking only works while CUDA has not been touched, so the launcher
first checks everything that would make the children die on their first CUDA call (``_preflight``).  ``debug_launcher`` forks a CPU-only
gloo job that rendezvous through a temporary file — the "cluster on a laptop" all multi-process unit tests of this repository run on.
Behavioural parity: reference ``src/accelerate/launchers.py:41-309`` (TPU branches excluded).
"""

from __future__ import annotations

import tempfile

import torch

from .state import AcceleratorState, PartialState
from .utils.environment import are_libraries_initialized, patch_environment
from .utils.launch import PrepareForLaunch
from .utils.other import get_free_port, is_port_in_use

_CUDA_FORK_ERROR = "Cannot re-initialize CUDA in forked subprocess"


def test_launch():
    """Verify a ``PartialState`` can be initialised (the smallest thing a launcher can be asked to run)."""
    _ = PartialState()


def _preflight() -> None:
    """Refuse to fork when the parent already holds state the children cannot share."""
    if AcceleratorState._shared_state:
        raise ValueError("To launch a multi-GPU training from your notebook, the `Accelerator` should only be initialized inside your "
                         "training function. Restart your notebook and make sure no cells initializes an `Accelerator`.")
    if torch.cuda.is_available() and torch.cuda.is_initialized():
        raise ValueError("To launch a multi-GPU training from your notebook, you need to avoid running any instruction using `torch.cuda` "
                         "in any cell. Restart your notebook and make sure no cells use any CUDA function.")
    touched = are_libraries_initialized("bitsandbytes")
    if touched:
        raise RuntimeError("Could not start distributed process. Libraries known to initialize CUDA upon import have been imported "
                           f"already: {touched}. Please keep these imports inside your training function.")


def _resolve_port(use_port) -> str:
    if int(use_port) == 0:
        return str(get_free_port())
    if is_port_in_use(int(use_port)):
        raise ConnectionError(f"Port {use_port} is already in use; pass another `use_port` (or 0 for a free one).")
    return str(use_port)


def notebook_launcher(function, args=(), num_processes=None, mixed_precision="no", use_port="29500", master_addr="127.0.0.1",
                      node_rank=0, num_nodes=1, rdzv_backend="static", rdzv_endpoint="", rdzv_conf=None, rdzv_id="none",
                      max_restarts=0, monitor_interval=0.1, log_line_prefix_template=None):
    """Run ``function(*args)`` on ``num_processes`` devices of this machine (default: every visible GPU).  With one process the
    function simply runs here; with more, workers are forked and the call returns when all of them have finished."""
    gpus_here = torch.cuda.is_available()
    if num_processes is None:
        num_processes = torch.cuda.device_count() if gpus_here else 1
    num_processes = int(num_processes)
    if num_processes <= 1:
        print("Launching training on one GPU." if gpus_here else "Launching training on CPU.")
        function(*args)
        return

    from torch.distributed.launcher.api import LaunchConfig, elastic_launch
    from torch.multiprocessing.spawn import ProcessRaisedException

    _preflight()
    port = _resolve_port(use_port)
    rdzv_conf = dict(rdzv_conf or {})
    if rdzv_backend == "static":
        rdzv_conf["rank"] = node_rank
        rdzv_endpoint = rdzv_endpoint or f"{master_addr}:{port}"
    config = dict(min_nodes=num_nodes, max_nodes=num_nodes, nproc_per_node=num_processes, run_id=rdzv_id, rdzv_endpoint=rdzv_endpoint,
                  rdzv_backend=rdzv_backend, rdzv_configs=rdzv_conf, max_restarts=max_restarts, monitor_interval=monitor_interval,
                  start_method="fork")
    if log_line_prefix_template is not None:
        config["log_line_prefix_template"] = log_line_prefix_template

    worker_env = dict(nproc=num_processes, node_rank=node_rank, world_size=num_nodes * num_processes, master_addr=master_addr,
                      master_port=port, mixed_precision=mixed_precision, accelerate_mixed_precision=mixed_precision)
    if not gpus_here:
        worker_env["accelerate_use_cpu"] = "yes"
    with patch_environment(**worker_env):
        # validates the mixed-precision / device choice in the parent, before anything is forked
        PrepareForLaunch(function, distributed_type="MULTI_GPU" if gpus_here else "MULTI_CPU")
        print(f"Launching training on {num_processes} {'GPUs' if gpus_here else 'CPUs'}.")
        try:
            elastic_launch(config=LaunchConfig(**config), entrypoint=function)(*args)
        except ProcessRaisedException as failure:
            if _CUDA_FORK_ERROR in failure.args[0]:
                raise RuntimeError("CUDA has been initialized before the `notebook_launcher` could create a forked subprocess. This "
                                   "likely stems from an outside import causing issues once the `notebook_launcher()` is called. Please "
                                   "review your imports and test them when running the `notebook_launcher()` to identify which one is "
                                   "problematic and causing CUDA to be initialized.") from failure
            raise RuntimeError(f"An issue was found when launching the training: {failure}") from failure


def debug_launcher(function, args=(), num_processes=2):
    """Run ``function(*args)`` on ``num_processes`` forked CPU processes that form a gloo group (file rendezvous, nothing on the network)."""
    from torch.multiprocessing import start_processes

    with tempfile.NamedTemporaryFile() as rendezvous:
        with patch_environment(world_size=num_processes, master_addr="127.0.0.1", master_port="29500", accelerate_mixed_precision="no",
                               accelerate_debug_rdv_file=rendezvous.name, accelerate_use_cpu="yes"):
            start_processes(PrepareForLaunch(function, debug=True), args=args, nprocs=num_processes, start_method="fork")
