"""What ``notebook_launcher`` has to get right, run as a plain program (it forks its workers itself, so it must NOT be started under
``accelerate launch``):

    python -m accelerate_hpc_test_b200.test_utils.scripts.test_notebook

Workers start and form a group (static and c10d rendezvous); a worker that crashes is restarted up to ``max_restarts`` times; the
agent notices a dead worker within the monitor interval instead of waiting for the others; imports that initialise CUDA are refused.
The number of workers comes from ``ACCELERATE_NUM_PROCESSES`` (default 2); without GPUs they are CPU / gloo workers.
"""

import os
import sys
import tempfile
import time

import torch

from accelerate_hpc_test_b200 import PartialState, notebook_launcher
from accelerate_hpc_test_b200.utils.environment import patch_environment

NUM_PROCESSES = int(os.environ.get("ACCELERATE_NUM_PROCESSES", 2))


def say_hello():
    state = PartialState()
    total = torch.ones(1, device=state.device)
    torch.distributed.all_reduce(total)
    assert int(total.item()) == state.num_processes
    print(f"hello from rank {state.process_index} of {state.num_processes}", flush=True)


def crash_until_marked(folder):
    """Rank 0 crashes on its first two lives and leaves a mark each time; the third life finds two marks and succeeds."""
    state = PartialState()
    if state.is_main_process:
        lives = len(os.listdir(folder))
        if lives < 2:
            open(os.path.join(folder, f"life_{lives}"), "w").close()
            raise RuntimeError(f"deliberate failure number {lives + 1}")
    state.wait_for_everyone()


def one_rank_dies_the_others_sleep(seconds):
    state = PartialState()
    if state.is_main_process:
        raise RuntimeError("rank 0 gives up at once")
    time.sleep(seconds)


def check_workers_start(**rendezvous):
    notebook_launcher(say_hello, num_processes=NUM_PROCESSES, use_port="0", **rendezvous)


def check_restarts():
    with tempfile.TemporaryDirectory() as folder:
        notebook_launcher(crash_until_marked, (folder,), num_processes=NUM_PROCESSES, use_port="0", max_restarts=3)
        assert len(os.listdir(folder)) == 2  # two crashes, then a clean run
        for name in os.listdir(folder):
            os.remove(os.path.join(folder, name))
        try:
            notebook_launcher(crash_until_marked, (folder,), num_processes=NUM_PROCESSES, use_port="0", max_restarts=1)
        except Exception as failure:  # torchelastic's ChildFailedError, or the launcher's RuntimeError around it
            assert "deliberate failure" in str(failure) or "crash_until_marked" in str(failure), failure
        else:
            raise AssertionError("one restart cannot absorb two crashes")


def check_monitoring(sleep_seconds=60, monitor_interval=0.05):
    started = time.time()
    try:
        notebook_launcher(one_rank_dies_the_others_sleep, (sleep_seconds,), num_processes=NUM_PROCESSES, use_port="0",
                          monitor_interval=monitor_interval)
    except Exception:
        pass
    else:
        raise AssertionError("a failed worker must fail the launch")
    assert time.time() - started < sleep_seconds / 2, "the agent waited for the sleeping workers instead of noticing the dead one"


def check_cuda_initialising_imports_are_refused():
    sys.modules.setdefault("bitsandbytes", type(sys)("bitsandbytes"))
    try:
        notebook_launcher(say_hello, num_processes=NUM_PROCESSES, use_port="0")
    except RuntimeError as refusal:
        assert "Please keep these imports" in str(refusal)
    else:
        raise AssertionError("an imported bitsandbytes must stop the launcher")
    finally:
        sys.modules.pop("bitsandbytes", None)


def main():
    print("Test basic notebook can be ran")
    check_workers_start()
    print("Test static rendezvous backend")
    check_workers_start(rdzv_backend="static")
    print("Test c10d rendezvous backend")
    check_workers_start(rdzv_backend="c10d", rdzv_endpoint="127.0.0.1:0")
    print("Test fault tolerance")
    check_restarts()
    print("Test monitoring")
    check_monitoring()
    if os.environ.get("ACCELERATE_TEST_PROBLEMATIC_IMPORTS", "1") != "0":
        print("Test problematic imports")
        check_cuda_initialising_imports_are_refused()
    print("test_notebook: all passed")


if __name__ == "__main__":
    with patch_environment(omp_num_threads=1):
        main()
