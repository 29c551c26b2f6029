"""``accelerate`` console entry point (reference src/accelerate/commands/accelerate_cli.py:28-50)."""

from .config import get_config_parser
from .env import env_command_parser
from .estimate import estimate_command_parser
from .launch import launch_command_parser
from .merge import merge_command_parser
from .test import test_command_parser
from .to_fsdp2 import to_fsdp2_command_parser
from .tpu import tpu_command_parser
from .utils import CustomArgumentParser


def main():
    parser = CustomArgumentParser("Accelerate CLI tool", usage="accelerate <command> [<args>]", allow_abbrev=False)
    subparsers = parser.add_subparsers(help="accelerate command helpers")
    get_config_parser(subparsers=subparsers)
    estimate_command_parser(subparsers=subparsers)
    env_command_parser(subparsers=subparsers)
    launch_command_parser(subparsers=subparsers)
    merge_command_parser(subparsers=subparsers)
    test_command_parser(subparsers=subparsers)
    to_fsdp2_command_parser(subparsers=subparsers)
    tpu_command_parser(subparsers=subparsers)
    args = parser.parse_args()
    if not hasattr(args, "func"):
        parser.print_help()
        exit(1)
    args.func(args)


if __name__ == "__main__":
    main()
