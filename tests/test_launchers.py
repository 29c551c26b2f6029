"""In-process launchers (behaviour of reference tests/test_notebook.py / launchers.py): single-process passthrough, preflight refusals,
a real 2-process CPU launch through torchelastic."""

import os

import pytest
import torch

from accelerate_hpc_test_b200 import launchers
from accelerate_hpc_test_b200.launchers import debug_launcher, notebook_launcher
from accelerate_hpc_test_b200.state import AcceleratorState, PartialState


def test_single_process_runs_in_place(capsys):
    seen = []
    notebook_launcher(lambda a, b: seen.append((a, b, os.getpid())), args=(1, 2), num_processes=1)
    assert seen == [(1, 2, os.getpid())]
    assert "Launching training on" in capsys.readouterr().out


def test_refuses_to_fork_after_an_accelerator_was_created():
    from accelerate_hpc_test_b200 import Accelerator

    Accelerator(cpu=True)
    assert AcceleratorState._shared_state
    with pytest.raises(ValueError, match="should only be initialized inside your training function"):
        notebook_launcher(lambda: None, num_processes=2)


def test_refuses_a_busy_port_and_picks_a_free_one_for_zero(monkeypatch):
    monkeypatch.setattr(launchers, "is_port_in_use", lambda port: True)
    with pytest.raises(ConnectionError, match="already in use"):
        launchers._resolve_port("29500")
    assert int(launchers._resolve_port("0")) > 0
    monkeypatch.setattr(launchers, "is_port_in_use", lambda port: False)
    assert launchers._resolve_port(29512) == "29512"


def test_refuses_when_a_cuda_initialising_library_is_loaded(monkeypatch):
    monkeypatch.setattr(launchers, "are_libraries_initialized", lambda *names: list(names))
    with pytest.raises(RuntimeError, match="bitsandbytes"):
        launchers._preflight()


def _report(folder):
    state = PartialState()
    t = torch.tensor([state.process_index + 1.0])
    torch.distributed.all_reduce(t)
    with open(os.path.join(folder, f"rank{state.process_index}.txt"), "w") as f:
        f.write(f"{state.num_processes} {t.item()} {os.environ.get('ACCELERATE_MIXED_PRECISION')} {state.distributed_type}")


@pytest.mark.skipif(torch.cuda.is_available(), reason="the CPU fork path")
def test_two_cpu_processes_through_torchelastic(tmp_path):
    notebook_launcher(_report, args=(str(tmp_path),), num_processes=2, use_port="0")
    lines = sorted((tmp_path / name).read_text() for name in os.listdir(tmp_path))
    assert len(lines) == 2 and all(line.startswith("2 3.0 no ") and "MULTI_CPU" in line for line in lines)
    assert "ACCELERATE_USE_CPU" not in os.environ  # the parent's environment is restored


def test_debug_launcher_passes_arguments(tmp_path):
    debug_launcher(_report, args=(str(tmp_path),), num_processes=2)
    assert sorted(os.listdir(tmp_path)) == ["rank0.txt", "rank1.txt"]
