"""``accelerate config default`` / ``write_basic_config`` — a config for "every GPU of this machine" (or the CPU)
without asking anything.  An existing file is never overwritten."""

from __future__ import annotations

from pathlib import Path

import torch

from .config_args import ClusterConfig, default_json_config_file
from .config_utils import SubcommandHelpFormatter

description = "Create a default config file for Accelerate with only a few flags set."
_PRECISIONS = ("no", "fp16", "bf16", "fp8")


def _detect_hardware() -> dict:
    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n_gpu == 0:
        return {"use_cpu": True, "num_processes": 1, "distributed_type": "NO"}
    return {"use_cpu": False, "num_processes": n_gpu, "distributed_type": "MULTI_GPU" if n_gpu > 1 else "NO"}


def write_basic_config(mixed_precision: str = "no", save_location: str = default_json_config_file):
    """Returns the path written, or ``False`` when a config already exists there."""
    target = Path(save_location)
    if target.exists():
        print(f"Configuration already exists at {save_location}, will not override. Run `accelerate config` manually or pass a "
              "different `save_location`.")
        return False
    precision = mixed_precision.lower()
    if precision not in _PRECISIONS:
        raise ValueError(f"`mixed_precision` should be one of 'no', 'fp16', 'bf16', or 'fp8'. Received {mixed_precision}")
    target.parent.mkdir(parents=True, exist_ok=True)
    config = ClusterConfig(compute_environment="LOCAL_MACHINE", mixed_precision=precision, debug=False, enable_cpu_affinity=False,
                           **_detect_hardware())
    config.save(target)
    return target


def default_config_command(args) -> None:
    written = write_basic_config(args.mixed_precision, args.save_location)
    if written:
        print(f"accelerate configuration saved at {written}")


def default_command_parser(parser, parents):
    sub = parser.add_parser("default", parents=parents, help=description, formatter_class=SubcommandHelpFormatter)
    sub.add_argument("--config_file", dest="save_location", default=default_json_config_file,
                     help="The path to use to store the config file.")
    sub.add_argument("--mixed_precision", type=str, default="no", choices=["no", "fp16", "bf16"],
                     help="Whether or not to use mixed precision training. Choose between FP16 and BF16 (bfloat16) training.")
    sub.set_defaults(func=default_config_command)
    return sub
