"""Import hygiene: importing the package (or starting the CLI) must not drag in optional heavyweight packages — trackers, model zoos,
data libraries, DeepSpeed / TransformerEngine / bitsandbytes — and the kernel library is opened on first use, not at import.
Counterpart of reference tests/test_imports.py (which reads ``python -X importtime`` the same way)."""

import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OPTIONAL = {"transformers", "datasets", "deepspeed", "transformer_engine", "bitsandbytes", "torchao", "tensorboard", "wandb", "comet_ml",
            "mlflow", "aim", "clearml", "dvclive", "swanlab", "trackio", "pandas", "sklearn", "scipy", "torchvision", "timm", "peft", "triton",
            "flash_attn", "flashinfer", "vllm", "cutlass", "rich", "matplotlib"}


def _imported_by(statement: str) -> set:
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    res = subprocess.run([sys.executable, "-X", "importtime", "-c", statement], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    names = set()
    for line in res.stderr.splitlines():
        if line.startswith("import time:") and "|" in line:
            names.add(line.rsplit("|", 1)[1].strip().split(".")[0])
    assert "torch" in names or "argparse" in names     # the listing was really parsed
    return names


@pytest.mark.parametrize("statement", [
    "import accelerate_hpc_test_b200",
    # the user-facing names, the CLI, the trackers and the utilities in one interpreter (each start costs seconds)
    "from accelerate_hpc_test_b200 import Accelerator, ParallelismConfig, notebook_launcher, init_empty_weights; "
    "import accelerate_hpc_test_b200.commands.accelerate_cli, accelerate_hpc_test_b200.tracking, accelerate_hpc_test_b200.utils",
])
def test_optional_packages_are_not_imported(statement):
    dragged_in = _imported_by(statement) & OPTIONAL
    assert not dragged_in, f"`{statement}` imports optional packages at import time: {sorted(dragged_in)}"


def test_kernel_library_is_opened_on_first_use_not_at_import():
    code = ("import accelerate_hpc_test_b200, accelerate_hpc_test_b200.ops.functional, accelerate_hpc_test_b200.parallel.fsdp;"
            "maps = open('/proc/self/maps').read(); print('LOADED' if 'libb200' in maps else 'NOT_LOADED')")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    res = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and res.stdout.strip().endswith("NOT_LOADED"), res.stdout + res.stderr[-2000:]
