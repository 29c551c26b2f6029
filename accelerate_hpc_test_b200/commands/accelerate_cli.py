"""The ``accelerate`` console entry point: one sub-command per module of this package."""

from __future__ import annotations

import sys

from . import config, env, estimate, launch, merge, test, to_fsdp2, tpu
from .utils import CustomArgumentParser

# (module, name of its parser factory); every factory registers itself on the sub-parser object and sets `func`
_SUBCOMMANDS = (
    (config, "get_config_parser"),
    (estimate, "estimate_command_parser"),
    (env, "env_command_parser"),
    (launch, "launch_command_parser"),
    (merge, "merge_command_parser"),
    (test, "test_command_parser"),
    (to_fsdp2, "to_fsdp2_command_parser"),
    (tpu, "tpu_command_parser"),
)


def build_parser() -> CustomArgumentParser:
    parser = CustomArgumentParser("Accelerate CLI tool", usage="accelerate <command> [<args>]", allow_abbrev=False)
    commands = parser.add_subparsers(help="accelerate command helpers")
    for module, factory in _SUBCOMMANDS:
        getattr(module, factory)(subparsers=commands)
    return parser


def main():
    parser = build_parser()
    args = parser.parse_args()
    handler = getattr(args, "func", None)
    if handler is None:
        parser.print_help()
        sys.exit(1)
    handler(args)


if __name__ == "__main__":
    main()
