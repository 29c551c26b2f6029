"""Terminal I/O of the ``accelerate config`` questionnaire: one free-text prompt, one menu prompt, and the
parsers that turn what the user typed into config values.

Parsers are plain callables ``str -> value`` that raise on bad input; the prompt loops until one succeeds.
Menu parsers receive the *index* the menu returned.  (Behaviour parity: reference
``commands/config/config_utils.py``; the ``_convert_*`` names are kept because callers import them.)
"""

from __future__ import annotations

import argparse
from typing import Callable, Sequence

from ...utils.dataclasses import ComputeEnvironment, DistributedType, DynamoBackend, FP8BackendType, PrecisionType
from ..menu import BulletMenu

DYNAMO_BACKENDS = [
    "EAGER", "AOT_EAGER", "INDUCTOR", "AOT_TS_NVFUSER", "NVPRIMS_NVFUSER", "CUDAGRAPHS", "OFI", "FX2TRT", "ONNXRT",
    "TENSORRT", "AOT_TORCHXLA_TRACE_ONCE", "TORCHXLA_TRACE_ONCE", "IPEX", "TVM",
]
YES_NO_HINT = "Please enter yes or no."


def _ask_field(input_text: str, convert_value: Callable | None = None, default=None, error_message: str | None = None):
    """Free-text prompt.  Empty input returns ``default`` (when one is given); a parser failure prints
    ``error_message`` and asks again."""
    while True:
        typed = input(input_text)
        if typed == "" and default is not None:
            return default
        if convert_value is None:
            return typed
        try:
            return convert_value(typed)
        except Exception:  # noqa: BLE001 - any parser failure means "ask again"
            if error_message:
                print(error_message)


def _ask_options(input_text: str, options: Sequence = (), convert_value: Callable | None = None, default: int = 0):
    """Menu prompt: arrow keys on a TTY, numbered list otherwise.  Returns ``convert_value(index)``."""
    picked = BulletMenu(input_text, list(options)).run(default_choice=default)
    return picked if convert_value is None else convert_value(picked)


def pick_from(values: Sequence, wrap: Callable | None = None) -> Callable:
    """Parser factory for menus: index -> ``values[index]`` (optionally wrapped, e.g. into an Enum)."""

    def parse(index):
        item = values[int(index)]
        return item if wrap is None else wrap(item)

    return parse


_convert_compute_environment = pick_from(["LOCAL_MACHINE", "AMAZON_SAGEMAKER"], ComputeEnvironment)
_convert_distributed_mode = pick_from(["NO", "MULTI_CPU", "MULTI_GPU"], DistributedType)
_convert_dynamo_backend = pick_from(DYNAMO_BACKENDS, lambda name: DynamoBackend(name).value)
_convert_mixed_precision = pick_from(["no", "fp16", "bf16", "fp8"], PrecisionType)
_convert_fp8_backend = pick_from(["NATIVE", "TE", "AO", "MSAMP"], FP8BackendType)


def _convert_yes_no_to_bool(value: str) -> bool:
    answer = value.strip().lower()
    if answer in ("yes", "y"):
        return True
    if answer in ("no", "n"):
        return False
    raise ValueError(value)


class SubcommandHelpFormatter(argparse.RawDescriptionHelpFormatter):
    """``accelerate config <sub> -h`` without the parent's ``<command> [<args>]`` placeholder in the usage line."""

    def _format_usage(self, usage, actions, groups, prefix):
        text = super()._format_usage(usage, actions, groups, prefix)
        return text.replace("<command> [<args>] ", "")
