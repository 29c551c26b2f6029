"""``accelerate config default``: write a config without asking questions (reference commands/config/default.py:36-163)."""

from pathlib import Path

import torch

from .config_args import ClusterConfig, default_json_config_file
from .config_utils import SubcommandHelpFormatter

description = "Create a default config file for Accelerate with only a few flags set."


def write_basic_config(mixed_precision="no", save_location: str = default_json_config_file):
    """One config for "all the GPUs of this machine" (or CPU). Returns the path, or False if the file exists."""
    path = Path(save_location)
    path.parent.mkdir(parents=True, exist_ok=True)
    if path.exists():
        print(f"Configuration already exists at {save_location}, will not override. Run `accelerate config` manually or pass a different `save_location`.")
        return False
    mixed_precision = mixed_precision.lower()
    if mixed_precision not in ["no", "fp16", "bf16", "fp8"]:
        raise ValueError(f"`mixed_precision` should be one of 'no', 'fp16', 'bf16', or 'fp8'. Received {mixed_precision}")
    config = {"compute_environment": "LOCAL_MACHINE", "mixed_precision": mixed_precision}
    if torch.cuda.is_available():
        num_gpus = torch.cuda.device_count()
        config["num_processes"] = num_gpus
        config["use_cpu"] = False
        config["distributed_type"] = "MULTI_GPU" if num_gpus > 1 else "NO"
    else:
        config["use_cpu"] = True
        config["num_processes"] = 1
        config["distributed_type"] = "NO"
    config["debug"] = False
    config["enable_cpu_affinity"] = False
    config = ClusterConfig(**config)
    if str(path).endswith(".json"):
        config.to_json_file(path)
    else:
        config.to_yaml_file(path)
    return path


def default_command_parser(parser, parents):
    parser = parser.add_parser("default", parents=parents, help=description, formatter_class=SubcommandHelpFormatter)
    parser.add_argument("--config_file", default=default_json_config_file, help="The path to use to store the config file.", dest="save_location")
    parser.add_argument("--mixed_precision", choices=["no", "fp16", "bf16"], type=str, default="no",
                        help="Whether or not to use mixed precision training. Choose between FP16 and BF16 (bfloat16) training.")
    parser.set_defaults(func=default_config_command)
    return parser


def default_config_command(args):
    config_file = write_basic_config(args.mixed_precision, args.save_location)
    if config_file:
        print(f"accelerate configuration saved at {config_file}")
