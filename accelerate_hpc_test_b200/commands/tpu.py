"""``accelerate tpu-config`` (reference commands/tpu.py) — not available: this build targets B200 GPUs only."""

import argparse


def tpu_command_parser(subparsers=None):
    description = "Run commands across TPU VMs for initial setup before running `accelerate launch` (not supported on this build)."
    if subparsers is not None:
        parser = subparsers.add_parser("tpu-config", description=description)
    else:
        parser = argparse.ArgumentParser("Accelerate tpu-config command", description=description)
    if subparsers is not None:
        parser.set_defaults(func=tpu_command_launcher)
    return parser


def tpu_command_launcher(args):
    raise SystemExit("`accelerate tpu-config` is not supported: the B200-native build has no TPU back-end.")


def main():
    tpu_command_launcher(tpu_command_parser().parse_args())
