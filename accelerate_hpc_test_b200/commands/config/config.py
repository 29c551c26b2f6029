"""``accelerate config`` — run the questionnaire and save the answers.

The file goes to ``--config_file`` when given, else to the default location (created on demand); the
extension picks the format (``.json`` → JSON, anything else → YAML).
"""

from __future__ import annotations

import argparse
import os

from . import config_args
from .cluster import get_cluster_input
from .config_args import cache_dir, default_config_file, default_yaml_config_file, load_config_from_file  # noqa: F401  (re-exported)
from .config_utils import _ask_field, _ask_options, _convert_compute_environment  # noqa: F401  (re-exported)

description = (
    "Launches a series of prompts to create and save a `default_config.yaml` configuration file for your training system. "
    "Should always be ran first on your machine"
)
_CONFIG_FILE_HELP = (
    "The path to use to store the config file. Will default to a file named default_config.yaml in the cache location, which is "
    "the content of the environment `HF_HOME` suffixed with 'accelerate', or if you don't have such an environment variable, your "
    "cache directory ('~/.cache' or the content of `XDG_CACHE_HOME`) suffixed with 'huggingface'."
)


def get_user_input():
    """Only local machines / clusters are configurable in this build, so the environment question has one answer."""
    _ask_options("In which compute environment are you running?", ["This machine"], _convert_compute_environment)
    return get_cluster_input()


def _destination(explicit: str | None) -> str:
    if explicit:
        return explicit
    os.makedirs(config_args.cache_dir, exist_ok=True)
    return config_args.default_yaml_config_file


def config_command(args) -> None:
    answers = get_user_input()
    target = _destination(args.config_file)
    answers.save(target)
    print(f"accelerate configuration saved at {target}")


def config_command_parser(subparsers=None):
    if subparsers is None:
        parser = argparse.ArgumentParser("Accelerate config command", description=description)
    else:
        parser = subparsers.add_parser("config", description=description)
        parser.set_defaults(func=config_command)
    parser.add_argument("--config_file", default=None, help=_CONFIG_FILE_HELP)
    return parser


def main():
    config_command(config_command_parser().parse_args())


if __name__ == "__main__":
    main()
