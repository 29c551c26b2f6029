"""``accelerate config update`` — load a config written by any earlier version and write it back in today's schema
(missing keys get their defaults, legacy keys are translated; unknown keys are an error, see ``config_args``)."""

from __future__ import annotations

import os

from . import config_args
from .config_utils import SubcommandHelpFormatter

description = "Update an existing config file with the latest defaults while maintaining the old configuration."


def update_config(args) -> str:
    """Rewrite the file in place and return its path."""
    path = args.config_file
    if path is None:
        if not os.path.exists(config_args.default_config_file):
            raise ValueError(f"No config file was passed and there is none at the default location {config_args.default_config_file}.")
        path = config_args.default_config_file
    elif not os.path.exists(path):
        raise ValueError(f"The passed config file located at {path} doesn't exist.")
    config_args.load_config_from_file(path).save(path)
    return path


def update_config_command(args) -> None:
    print(f"Sucessfully updated the configuration file at {update_config(args)}.")


def update_command_parser(parser, parents):
    sub = parser.add_parser("update", parents=parents, help=description, formatter_class=SubcommandHelpFormatter)
    sub.add_argument("--config_file", default=None,
                     help="The path to the config file to update. Will default to a file named default_config.yaml in the cache location.")
    sub.set_defaults(func=update_config_command)
    return sub
