"""In-tree build of the sm_100a kernel library (``lib/libb200.so``).

Plain ``nvcc`` (no torch headers): every kernel is exposed through a C ABI (``include/b200_ops.h``)
and loaded with ``ctypes``, so a full rebuild takes seconds and the ``.so`` travels with the
source tree.  nvcc cross-compiles without a GPU, which is what the driver's ``build()`` check
relies on.
"""

from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB_DIR = PKG_DIR / "lib"
OBJ_DIR = LIB_DIR / "obj"
LIB_PATH = LIB_DIR / "libb200.so"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-lineinfo",
    "--use_fast_math",
    "-Xcompiler",
    "-fPIC",
    "-Xcompiler",
    "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
]


def find_nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found; set NVCC or install the CUDA toolkit")
    return cand


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _stale(target: Path, deps: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(d.stat().st_mtime > t for d in deps)


def _compile_one(nvcc: str, src: Path, obj: Path, verbose: bool, ptxas_v: bool) -> str:
    cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
    if ptxas_v:
        cmd += ["-Xptxas", "-v"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout, res.stderr, file=sys.stderr)
    return res.stderr


def build(force: bool = False, verbose: bool = False, ptxas_v: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``lib/libb200.so``. Returns its path."""
    nvcc = find_nvcc()
    LIB_DIR.mkdir(exist_ok=True)
    OBJ_DIR.mkdir(exist_ok=True)
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h"))
    srcs = sources()
    todo = []
    objs = []
    for s in srcs:
        o = OBJ_DIR / (s.stem + ".o")
        objs.append(o)
        if force or _stale(o, [s, *headers]):
            todo.append((s, o))
    logs = {}
    if todo:
        with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            futs = {ex.submit(_compile_one, nvcc, s, o, verbose, ptxas_v): s for s, o in todo}
            for f in concurrent.futures.as_completed(futs):
                logs[futs[f].name] = f.result()
    if todo or not LIB_PATH.exists():
        cmd = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(LIB_PATH), *map(str, objs), "-lcudart"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    if ptxas_v:
        for name, log in sorted(logs.items()):
            print(f"==== {name}\n{log}")
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose=True, ptxas_v="--ptxas-v" in sys.argv)
    print(p)
