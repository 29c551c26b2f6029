"""Same script, same seeds, same data: this framework and the unmodified reference (installed offline under baseline/_ref)
produce the same losses, learning rate and final weights on CPU, with 1 process and with 2 gloo processes
(the reference's own CPU "cluster on a laptop" is the oracle for the framework layer)."""

import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
SCRIPT = os.path.join(ROOT, "accelerate_hpc_test_b200", "test_utils", "scripts", "parity_train.py")

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "accelerate")), reason="reference not installed under baseline/_ref")


def _run(pkg, out, nproc):
    extra = REF if pkg == "accelerate" else ROOT
    env = dict(os.environ, PYTHONPATH=extra + os.pathsep + os.environ.get("PYTHONPATH", ""), OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, SCRIPT, pkg, str(out), str(nproc)], env=env, cwd=str(out.parent), capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-1000:] + r.stderr[-3000:]
    return json.load(open(out))


@pytest.mark.parametrize("nproc", [1, 2])
def test_losses_match_the_reference(tmp_path, nproc):
    ref = _run("accelerate", tmp_path / "ref.json", nproc)
    ours = _run("accelerate_hpc_test_b200", tmp_path / "ours.json", nproc)
    assert ref["num_processes"] == ours["num_processes"] == nproc
    assert len(ref["losses"]) == len(ours["losses"]) > 0
    for i, (a, b) in enumerate(zip(ref["losses"], ours["losses"])):
        assert abs(a - b) <= 1e-6 * max(1.0, abs(a)), (i, a, b)
    assert abs(ref["lr"] - ours["lr"]) < 1e-12
    assert abs(ref["checksum"] - ours["checksum"]) <= 1e-6 * abs(ref["checksum"])
