"""``accelerate merge-weights <checkpoint_dir> <output_path>`` — fold a sharded (``SHARDED_STATE_DICT``) FSDP checkpoint
directory into one ``model.safetensors`` (or ``pytorch_model.bin`` with ``--unsafe_serialization``).

Runs on one CPU process, needs no process group, and needs RAM for the whole state dict.  The directory is a
``torch.distributed.checkpoint`` one, whether this build or the reference wrote it.
"""

from __future__ import annotations

from ..utils.fsdp_utils import merge_fsdp_weights
from .utils import CustomArgumentParser

description = (
    "Utility to merge the weights from multiple FSDP checkpoints into a single combined checkpoint. Should be used if\n"
    "`SHARDED_STATE_DICT` was used for the model. Weights will be saved to `{output_path}`.\n\n"
    "This is a CPU-bound process and requires enough RAM to load the entire model state dict."
)

_ARGUMENTS = (
    (("checkpoint_directory",), dict(type=str, help="A directory containing sharded weights saved by FSDP.")),
    (("output_path",), dict(type=str, help="The path to save the merged weights. Defaults to the current directory. ")),
    (("--unsafe_serialization",), dict(action="store_true", default=False,
                                       help="Whether to save the merged weights as `.bin` rather than `.safetensors` (not recommended).")),
    (("--remove_checkpoint_dir",), dict(action="store_true", default=False,
                                        help="Whether to remove the checkpoint directory after merging.")),
)


def merge_command(args) -> None:
    merge_fsdp_weights(args.checkpoint_directory, args.output_path, safe_serialization=not args.unsafe_serialization,
                       remove_checkpoint_dir=args.remove_checkpoint_dir)


def merge_command_parser(subparsers=None):
    standalone = subparsers is None
    parser = CustomArgumentParser(description=description) if standalone else subparsers.add_parser("merge-weights", description=description)
    for flags, options in _ARGUMENTS:
        parser.add_argument(*flags, **options)
    if not standalone:
        parser.set_defaults(func=merge_command)
    return parser


def main():
    merge_command(merge_command_parser().parse_args())


if __name__ == "__main__":
    main()
