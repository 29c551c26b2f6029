"""Install the UNMODIFIED reference package into ``baseline/_ref`` (git-ignored) for ``bench.py --impl reference``.

Order tried, first success wins:
  1. ``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref <tmp copy>``
     (the source tree is read-only, so pip builds from a copy under /tmp);
  2. the same with ``--no-deps`` (the wheelhouse has no numpy wheel, so dependency resolution fails);
  3. a plain copy of ``<reference>/src/accelerate`` (the package is pure Python, ``setup.py`` declares no extensions).

The call is idempotent: an existing ``baseline/_ref/accelerate/__init__.py`` is left alone.
"""

from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
TARGET = os.path.join(HERE, "_ref")
REFERENCE_ROOTS = [os.environ.get("B200_REFERENCE_ROOT", ""), "/root/reference"]


def reference_root() -> str | None:
    for r in REFERENCE_ROOTS:
        if r and os.path.isfile(os.path.join(r, "src", "accelerate", "__init__.py")):
            return r
    return None


LOCK = os.path.join(TARGET, ".installing")


def installed() -> bool:
    """True once a complete install is there (a half-written one carries the ``.installing`` marker)."""
    return os.path.isfile(os.path.join(TARGET, "accelerate", "__init__.py")) and not os.path.exists(LOCK)


def _pip(src: str, extra: list[str]) -> bool:
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", "/opt/wheelhouse",
           "--target", TARGET, *extra, src]
    res = subprocess.run(cmd, capture_output=True, text=True)
    return res.returncode == 0 and os.path.isfile(os.path.join(TARGET, "accelerate", "__init__.py"))


def install(verbose: bool = False) -> str:
    """Returns how the reference got there: 'present' | 'pip' | 'pip --no-deps' | 'copy' | 'unavailable: ...'."""
    if installed():
        return "present"
    root = reference_root()
    if root is None:
        return "unavailable: no reference source tree visible (looked at " + ", ".join(r for r in REFERENCE_ROOTS if r) + ")"
    os.makedirs(TARGET, exist_ok=True)
    open(LOCK, "w").close()
    how = None
    tmp = tempfile.mkdtemp(prefix="refsrc_")
    try:
        src = os.path.join(tmp, "reference")
        shutil.copytree(root, src, ignore=shutil.ignore_patterns(".git", "docs", "manim_animations", "__pycache__"))
        for extra, name in (([], "pip"), (["--no-deps"], "pip --no-deps")):
            try:
                if _pip(src, extra):
                    how = name
                    break
            except Exception:  # noqa: BLE001 - pip missing / broken: fall through to the plain copy
                pass
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    if how is None:
        shutil.copytree(os.path.join(root, "src", "accelerate"), os.path.join(TARGET, "accelerate"), dirs_exist_ok=True)
        how = "copy"
    if os.path.exists(LOCK):
        os.remove(LOCK)
    if verbose:
        print(f"reference installed into {TARGET} ({how})")
    return how if installed() else "unavailable: install failed"


if __name__ == "__main__":
    print(install(verbose=True))
