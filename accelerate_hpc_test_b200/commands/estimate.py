"""``accelerate estimate-memory <model>`` — how much memory do the weights need per dtype, what is the largest layer that must
fit on one device, and what does Adam training cost?  The model is built on the ``meta`` device, so nothing is allocated.

There is no hub access in this environment: ``<model>`` is one of the built-in native configs (``llama3-8b``, ``llama3-1b``,
``bert-base-cased``) or a local directory / ``config.json`` of a transformers model.
"""

from __future__ import annotations

import os

import torch

from ..big_modeling import init_empty_weights
from ..utils.modeling import calculate_maximum_sizes
from ..utils.other import convert_bytes
from .utils import CustomArgumentParser

BUILTIN = ("llama3-8b", "llama3-1b", "bert-base-cased")
_SHRINK = {"float32": 1, "float16": 2, "bfloat16": 2, "int8": 4, "int4": 8}  # bytes relative to fp32
_HEADERS = ["dtype", "Largest Layer", "Total Size", "Training using Adam"]


# --------------------------------------------------------------------------- model construction
def _builtin(name: str):
    from ..models import LlamaConfig, LlamaForCausalLM
    from ..models.bert import BertConfig, BertForSequenceClassification

    if name == "bert-base-cased":
        return BertForSequenceClassification(BertConfig.base_cased())
    cfg = LlamaConfig.llama3_8b() if name == "llama3-8b" else LlamaConfig.llama3_1b()
    return LlamaForCausalLM(cfg, dtype=torch.float32, device="meta")


def _from_local_config(path: str, trust_remote_code: bool):
    import transformers

    cfg = transformers.AutoConfig.from_pretrained(path, trust_remote_code=trust_remote_code)
    names = getattr(cfg, "architectures", None) or []
    cls = getattr(transformers, names[0], None) if names else None
    if cls is not None:
        return cls(cfg)
    return transformers.AutoModel.from_config(cfg, trust_remote_code=trust_remote_code)


def create_empty_model(model_name: str, library_name: str = None, trust_remote_code: bool = False, access_token=None):
    """``model_name`` instantiated with all parameters on ``meta``."""
    with init_empty_weights():
        if model_name in BUILTIN:
            return _builtin(model_name)
        if os.path.exists(model_name):
            return _from_local_config(model_name, trust_remote_code)
    raise OSError(
        f"Model `{model_name}` is neither a built-in config ({', '.join(BUILTIN)}) nor a local path; the Hub is not reachable "
        "from this environment."
    )


# --------------------------------------------------------------------------- numbers
def estimate_training_usage(bytes: int, mixed_precision: str, msamp_config: str = None) -> dict:
    """Bytes held at each stage of an Adam step at batch size 1 (``-1``: no estimate for that precision).

    fp32: weights + gradients of the same size, two moments, and all four live during the step.  16-bit mixed precision: fp32
    master weights, a 16-bit copy, gradients that are up-cast for the update, two fp32 moments.
    """
    full, half = bytes, bytes // 2
    if mixed_precision == "float32":
        return {"model": full, "optimizer": 2 * full, "gradients": full, "step": 4 * full}
    if mixed_precision in ("float16", "bfloat16") or (mixed_precision == "fp8" and msamp_config is None):
        return {"model": full, "optimizer": 2 * full, "gradients": full + half, "step": 2 * full}
    return {"model": -1, "optimizer": -1, "gradients": -1, "step": -1}


def gather_data(args) -> list:
    """One row per requested dtype: ``[dtype, largest layer bytes, total bytes, training-usage dict]``."""
    model = create_empty_model(args.model_name, library_name=args.library_name, trust_remote_code=args.trust_remote_code)
    total, (largest, _layer_names) = calculate_maximum_sizes(model)
    rows = []
    for dtype in args.dtypes:
        factor = _SHRINK[dtype]
        rows.append([dtype, largest / factor if factor > 1 else largest, total / factor if factor > 1 else total,
                     estimate_training_usage(total, dtype)])
    return rows


# --------------------------------------------------------------------------- presentation
def create_ascii_table(headers: list, rows: list, title: str) -> str:
    """Box-drawing table with a centred title line spanning all columns."""
    widths = [max(len(str(r[c])) for r in rows + [headers]) for c in range(len(headers))]
    inner = sum(widths) + len(widths) - 1
    if len(title) > inner:
        widths[-1] += len(title) - inner
        inner = len(title)

    def rule(left, mid, right):
        return left + mid.join("─" * w for w in widths) + right

    def line(cells):
        return "│" + "│".join(str(c).center(w) for c, w in zip(cells, widths)) + "│"

    out = [rule("┌", "─", "┐"), "│" + title.center(inner) + "│", rule("├", "┬", "┤"), line(headers), rule("├", "┼", "┤")]
    out += [line(r) for r in rows]
    out.append(rule("└", "┴", "┘"))
    return "\n".join(out)


def _humanise(cell):
    if isinstance(cell, dict):
        peak = max(cell.values())
        return "N/A" if peak == -1 else convert_bytes(peak)
    if isinstance(cell, (int, float)):
        return convert_bytes(cell)
    return cell


def estimate_command(args) -> None:
    rows = [[_humanise(cell) for cell in row] for row in gather_data(args)]
    print(create_ascii_table(_HEADERS, rows, f"Memory Usage for loading `{args.model_name}`"))


def estimate_command_parser(subparsers=None):
    if subparsers is None:
        parser = CustomArgumentParser(description="Model size estimator for fitting a model onto CUDA memory.")
    else:
        parser = subparsers.add_parser("estimate-memory")
        parser.set_defaults(func=estimate_command)
    parser.add_argument("model_name", type=str, help="A built-in config name or a local path with a transformers config.")
    parser.add_argument("--library_name", type=str, choices=["timm", "transformers"],
                        help="The library the model has an integration with, such as `transformers`.")
    parser.add_argument("--dtypes", type=str, nargs="+", default=["float32", "float16", "int8", "int4"], choices=list(_SHRINK),
                        help="The dtypes to use for the model, must be one (or many) of `float32`, `float16`, `int8`, and `int4`")
    parser.add_argument("--trust_remote_code", action="store_true", default=False,
                        help="Whether or not to allow for custom models defined on the Hub in their own modeling files.")
    return parser


def main():
    estimate_command(estimate_command_parser().parse_args())


if __name__ == "__main__":
    main()
