"""``accelerate env`` — the facts a bug report needs: package / Python / torch versions, the GPU(s), whether the sm_100a kernel
library is built, and the launcher config in effect (``--config_file`` or the default one)."""

from __future__ import annotations

import argparse
import os
import platform
import shutil

import torch

from .. import __version__ as version
from .config import default_config_file, load_config_from_file


def _facts() -> dict:
    import numpy
    import psutil

    from ..ops import _lib

    lib = _lib.lib_path()
    facts = {
        "`Accelerate (b200)` version": version,
        "Platform": platform.platform(),
        "`accelerate` bash location": shutil.which("accelerate") or "Not found",
        "Python version": platform.python_version(),
        "Numpy version": numpy.__version__,
        "PyTorch version": torch.__version__,
        "PyTorch accelerator": "CUDA" if torch.cuda.is_available() else "N/A",
        "System RAM": f"{psutil.virtual_memory().total / 2**30:.2f} GB",
        "sm_100a kernel library": f"{lib} ({'built' if os.path.exists(lib) else 'not built'})",
    }
    if torch.cuda.is_available():
        major, minor = torch.cuda.get_device_capability()
        facts["GPU type"] = torch.cuda.get_device_name()
        facts["GPU count"] = torch.cuda.device_count()
        facts["Compute capability"] = f"{major}.{minor}"
    return facts


def _config_in_effect(path):
    if path is None and not os.path.isfile(default_config_file):
        return "Not found"
    return load_config_from_file(path).to_dict()


def env_command(args) -> dict:
    facts = _facts()
    config = _config_in_effect(args.config_file)
    lines = ["", "Copy-and-paste the text below in your GitHub issue", ""]
    lines += [f"- {name}: {value}" for name, value in facts.items()]
    lines.append("- `Accelerate` default config:" if args.config_file is None else "- `Accelerate` config passed:")
    if isinstance(config, dict):
        lines += [f"\t- {key}: {value}" for key, value in config.items()]
    else:
        lines.append(f"\t{config}")
    print("\n".join(lines))
    facts["`Accelerate` configs"] = config
    return facts


def env_command_parser(subparsers=None):
    parser = argparse.ArgumentParser("Accelerate env command") if subparsers is None else subparsers.add_parser("env")
    parser.add_argument("--config_file", default=None, help="The config file to use for the default values in the launching script.")
    if subparsers is not None:
        parser.set_defaults(func=env_command)
    return parser


def main() -> int:
    env_command(env_command_parser().parse_args())
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
