"""``accelerate to-fsdp2 --config_file old.yaml (--output_file new.yaml | --overwrite)`` — translate the ``fsdp_config``
section of a launcher config from FSDP1 vocabulary to FSDP2.

Each FSDP1 key falls in one of four classes (table ``ARGUMENT_KEY_MAPPING``, same content as the reference's): kept as is,
renamed (``fsdp_sharding_strategy`` → ``fsdp_reshard_after_forward``, with its value translated through
``ARGUMENT_VALUE_MAPPING``), gone in FSDP2 (dropped with a warning) or not available yet (dropped with a warning).  Keys the
table does not know are carried over untouched.
"""

from __future__ import annotations

import enum
import logging
import os

import yaml

from .utils import CustomArgumentParser

logger = logging.getLogger(__name__)


class ConversionStatus(enum.Enum):
    NOT_YET_IMPLEMENTED = 0
    REMOVED = -1


_SAME = object()  # marker: key keeps its name

ARGUMENT_KEY_MAPPING = {
    "fsdp_version": "fsdp_version",
    "fsdp_auto_wrap_policy": "fsdp_auto_wrap_policy",
    "fsdp_backward_prefetch": ConversionStatus.REMOVED,
    "fsdp_forward_prefetch": ConversionStatus.NOT_YET_IMPLEMENTED,
    "fsdp_cpu_ram_efficient_loading": "fsdp_cpu_ram_efficient_loading",
    "fsdp_offload_params": "fsdp_offload_params",
    "fsdp_sharding_strategy": "fsdp_reshard_after_forward",
    "fsdp_state_dict_type": "fsdp_state_dict_type",
    "fsdp_sync_module_states": ConversionStatus.REMOVED,
    "fsdp_transformer_layer_cls_to_wrap": "fsdp_transformer_layer_cls_to_wrap",
    "fsdp_min_num_params": "fsdp_min_num_params",
    "fsdp_use_orig_params": ConversionStatus.REMOVED,
    "fsdp_activation_checkpointing": "fsdp_activation_checkpointing",
}

_RESHARD = {"FULL_SHARD": True, "SHARD_GRAD_OP": False, "HYBRID_SHARD": True, "HYBRID_SHARD_ZERO2": False, "NO_SHARD": False}
ARGUMENT_VALUE_MAPPING = {"fsdp_sharding_strategy": _RESHARD, "fsdp_reshard_after_forward": _RESHARD}

_DROP_NOTES = {
    ConversionStatus.REMOVED: "Argument {key} has been removed in FSDP2, skipping this key...",
    ConversionStatus.NOT_YET_IMPLEMENTED: "Argument {key} is not yet implemented in FSDP2, skipping this key...",
}


def _translate(key: str, value):
    """(new key or None when dropped, new value)."""
    target = ARGUMENT_KEY_MAPPING.get(key, _SAME)
    if isinstance(target, ConversionStatus):
        logger.warning(_DROP_NOTES[target].format(key=key))
        return None, None
    if target is _SAME:
        return key, value
    return target, ARGUMENT_VALUE_MAPPING.get(key, {}).get(value, value)


def convert_config_to_fsdp2(config: dict) -> dict:
    """Returns ``config`` with its ``fsdp_config`` rewritten (in place); configs without one, or already at version 2, pass through."""
    old = config.get("fsdp_config") or {}
    if not old:
        logger.info("No FSDP config found in the config file, skipping conversion...")
        return config
    if old.get("fsdp_version", 1) == 2:
        logger.warning("Config already specifies FSDP2, skipping conversion...")
        logger.warning("If the config doesn't use new argument names, change `fsdp_version` to `1` and rerun the command.")
        return config
    new = {}
    for key, value in old.items():
        new_key, new_value = _translate(key, value)
        if new_key is not None:
            new[new_key] = new_value
    new["fsdp_version"] = 2
    config["fsdp_config"] = new
    return config


def _validate_to_fsdp2_args(args) -> None:
    if not os.path.exists(args.config_file):
        raise FileNotFoundError(f"Config file {args.config_file} not found")
    if args.overwrite:
        return
    if args.output_file is None:
        raise ValueError("If --overwrite is not set, --output_file must be provided")
    if os.path.exists(args.output_file):
        raise FileExistsError(f"Output file {args.output_file} already exists and --overwrite is not set")


def to_fsdp2_command(args) -> None:
    _validate_to_fsdp2_args(args)
    with open(args.config_file) as fh:
        config = yaml.safe_load(fh)
    destination = args.output_file if args.output_file is not None else args.config_file
    with open(destination, "w") as fh:
        yaml.dump(convert_config_to_fsdp2(config), fh)


def to_fsdp2_command_parser(subparsers=None):
    description = "Convert an Accelerate config from FSDP1 to FSDP2"
    standalone = subparsers is None
    parser = CustomArgumentParser(description=description) if standalone else subparsers.add_parser("to-fsdp2", description=description)
    parser.add_argument("--config_file", type=str, required=True, help="The config file to convert to FSDP2")
    parser.add_argument("--overwrite", action="store_true", default=False, help="Overwrite the config file if it exists")
    parser.add_argument("--output_file", type=str, default=None,
                        help="The path to the output file to write the converted config to. If not provided, the input file will be "
                             "overwritten (if --overwrite is set)")
    if not standalone:
        parser.set_defaults(func=to_fsdp2_command)
    return parser


def main():
    to_fsdp2_command(to_fsdp2_command_parser().parse_args())


if __name__ == "__main__":
    main()
