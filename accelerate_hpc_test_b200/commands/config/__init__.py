"""``accelerate config [default|update]`` parser assembly."""

import argparse
import sys

from .config import config_command_parser
from .config_args import default_config_file, load_config_from_file  # noqa: F401
from .default import default_command_parser
from .update import update_command_parser

_SUBCOMMAND_BUILDERS = (default_command_parser, update_command_parser)


def get_config_parser(subparsers=None):
    """The ``config`` parser with its ``default`` and ``update`` sub-commands attached."""
    parser = config_command_parser(subparsers)
    nested = parser.add_subparsers(title="subcommands", dest="subcommand")
    shared = argparse.ArgumentParser(add_help=False, allow_abbrev=False)
    for build in _SUBCOMMAND_BUILDERS:
        build(nested, parents=[shared])
    return parser


def main():
    parser = get_config_parser()
    ns = parser.parse_args()
    handler = getattr(ns, "func", None)
    if handler is None:
        parser.print_help()
        sys.exit(1)
    handler(ns)


if __name__ == "__main__":
    main()
