"""``accelerate test``: run the in-package end-to-end script under ``accelerate launch`` (reference commands/test.py:44-55)."""

import argparse
import sys

from ..utils.other import get_free_port


def test_command_parser(subparsers=None):
    if subparsers is not None:
        parser = subparsers.add_parser("test")
    else:
        parser = argparse.ArgumentParser("Accelerate test command")
    parser.add_argument("--config_file", default=None, help="The path to use to store the config file. Will default to a file named default_config.yaml in the cache location.")
    if subparsers is not None:
        parser.set_defaults(func=test_command)
    return parser


def test_command(args):
    import subprocess

    script_name = "accelerate_hpc_test_b200.test_utils.scripts.test_script"
    cmd = [sys.executable, "-m", "accelerate_hpc_test_b200.commands.launch"]
    if args.config_file is not None:
        cmd += ["--config_file", args.config_file]
    cmd += ["--main_process_port", str(get_free_port()), "-m", script_name]
    result = subprocess.run(cmd)
    if result.returncode == 0:
        print("Test is a success! You are ready for your distributed training!")
    return result.returncode


def main():
    parser = test_command_parser()
    args = parser.parse_args()
    raise SystemExit(test_command(args))


if __name__ == "__main__":
    main()
