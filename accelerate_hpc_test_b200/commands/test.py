"""``accelerate test`` — sanity-check an installation / config: launches the in-package end-to-end script
(``test_utils/scripts/test_script.py``) through ``accelerate launch`` with the given or default config."""

from __future__ import annotations

import argparse
import subprocess
import sys

from ..utils.other import get_free_port

_SCRIPT_MODULE = "accelerate_hpc_test_b200.test_utils.scripts.test_script"
_LAUNCH_MODULE = "accelerate_hpc_test_b200.commands.launch"


def _command_line(config_file) -> list:
    line = [sys.executable, "-m", _LAUNCH_MODULE]
    if config_file is not None:
        line += ["--config_file", config_file]
    return line + ["--main_process_port", str(get_free_port()), "-m", _SCRIPT_MODULE]


def test_command(args) -> int:
    outcome = subprocess.run(_command_line(args.config_file))
    if outcome.returncode == 0:
        print("Test is a success! You are ready for your distributed training!")
    return outcome.returncode


def test_command_parser(subparsers=None):
    parser = argparse.ArgumentParser("Accelerate test command") if subparsers is None else subparsers.add_parser("test")
    parser.add_argument("--config_file", default=None,
                        help="The path to use to store the config file. Will default to a file named default_config.yaml in the cache location.")
    if subparsers is not None:
        parser.set_defaults(func=test_command)
    return parser


def main():
    raise SystemExit(test_command(test_command_parser().parse_args()))


if __name__ == "__main__":
    main()
