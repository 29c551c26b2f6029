"""``accelerate estimate-memory``: parameter / training footprint of a model built on the meta device
(reference commands/estimate.py:66-308).  There is no hub access here, so the model comes from a local
``config.json`` (transformers) or one of the built-in native configs (``llama3-8b``, ``llama3-1b``,
``bert-base-cased``)."""

import os

import torch

from ..big_modeling import init_empty_weights
from ..utils.modeling import calculate_maximum_sizes, compute_module_sizes
from ..utils.other import convert_bytes
from .utils import CustomArgumentParser

BUILTIN = ("llama3-8b", "llama3-1b", "bert-base-cased")


def create_empty_model(model_name: str, library_name: str = None, trust_remote_code: bool = False, access_token=None):
    """Instantiate ``model_name`` on the meta device."""
    if model_name in BUILTIN:
        from ..models import LlamaConfig, LlamaForCausalLM
        from ..models.bert import BertConfig, BertForSequenceClassification

        with init_empty_weights():
            if model_name == "llama3-8b":
                return LlamaForCausalLM(LlamaConfig.llama3_8b(), dtype=torch.float32, device="meta")
            if model_name == "llama3-1b":
                return LlamaForCausalLM(LlamaConfig.llama3_1b(), dtype=torch.float32, device="meta")
            return BertForSequenceClassification(BertConfig.base_cased())
    if os.path.isdir(model_name) or os.path.isfile(model_name):
        from transformers import AutoConfig, AutoModel

        config = AutoConfig.from_pretrained(model_name, trust_remote_code=trust_remote_code)
        with init_empty_weights():
            constructor = AutoModel
            if isinstance(getattr(config, "architectures", None), list) and config.architectures:
                import transformers

                constructor = getattr(transformers, config.architectures[0], AutoModel)
            return constructor.from_config(config, trust_remote_code=trust_remote_code) if constructor is AutoModel else constructor(config)
    raise OSError(f"Model `{model_name}` is neither a built-in config ({', '.join(BUILTIN)}) nor a local path; the Hub is not reachable from this environment.")


def create_ascii_table(headers: list, rows: list, title: str):
    sep_char, in_between = "│", "─"
    column_widths = []
    for i in range(len(headers)):
        column_values = [row[i] for row in rows] + [headers[i]]
        column_widths.append(len(max(column_values, key=len)))
    formats = [f"%{column_widths[i]}s" for i in range(len(rows[0]))]
    pattern = f"{sep_char}{sep_char.join(formats)}{sep_char}"
    diff = 0

    def make_row(left_char, middle_char, right_char):
        return f"{left_char}{middle_char.join([in_between * n for n in column_widths])}{in_between * diff}{right_char}"

    separator = make_row("├", "┼", "┤")
    if len(title) > sum(column_widths):
        diff = abs(len(title) - len(separator))
        column_widths[-1] += diff
    separator = make_row("├", "┼", "┤")
    initial_rows = [make_row("┌", in_between, "┐"), f"{sep_char}{title.center(len(separator) - 2)}{sep_char}", make_row("├", "┬", "┤")]
    table = "\n".join(initial_rows) + "\n"
    column_widths[-1] += diff
    centered_line = [text.center(column_widths[i]) for i, text in enumerate(headers)]
    table += f"{pattern % tuple(centered_line)}\n{separator}\n"
    for i, line in enumerate(rows):
        centered_line = [t.center(column_widths[i]) for i, t in enumerate(line)]
        table += f"{pattern % tuple(centered_line)}\n"
    table += f"└{'┴'.join([in_between * n for n in column_widths])}┘"
    return table


def estimate_command_parser(subparsers=None):
    if subparsers is not None:
        parser = subparsers.add_parser("estimate-memory")
    else:
        parser = CustomArgumentParser(description="Model size estimator for fitting a model onto CUDA memory.")
    parser.add_argument("model_name", type=str, help="A built-in config name or a local path with a transformers config.")
    parser.add_argument("--library_name", type=str, help="The library the model has an integration with, such as `transformers`.", choices=["timm", "transformers"])
    parser.add_argument("--dtypes", type=str, nargs="+", default=["float32", "float16", "int8", "int4"], help="The dtypes to use for the model, must be one (or many) of `float32`, `float16`, `int8`, and `int4`",
                        choices=["float32", "float16", "bfloat16", "int8", "int4"])
    parser.add_argument("--trust_remote_code", action="store_true", default=False, help="Whether or not to allow for custom models defined on the Hub in their own modeling files.")
    if subparsers is not None:
        parser.set_defaults(func=estimate_command)
    return parser


def estimate_training_usage(bytes: int, mixed_precision: str, msamp_config: str = None) -> dict:
    """Rough per-stage footprint of Adam training with batch size 1: model, gradients, optimizer, one step."""
    memory_sizes = {"model": -1, "optimizer": -1, "gradients": -1, "step": -1}
    fp32_size = bytes
    fp16_size = bytes // 2
    if mixed_precision == "float32":
        memory_sizes["model"] = fp32_size
        memory_sizes["gradients"] = fp32_size
        memory_sizes["optimizer"] = fp32_size * 2
        memory_sizes["step"] = fp32_size * 4
    elif mixed_precision in ("float16", "bfloat16") or (mixed_precision == "fp8" and msamp_config is None):
        # master fp32 weights + low-precision copy, fp32 gradients after the reduction, fp32 moments
        memory_sizes["model"] = fp32_size
        memory_sizes["gradients"] = fp32_size + fp16_size
        memory_sizes["optimizer"] = fp32_size * 2
        memory_sizes["step"] = memory_sizes["optimizer"]
    return memory_sizes


def gather_data(args):
    """Rows of [dtype, largest layer, total size, Adam training size]."""
    try:
        model = create_empty_model(args.model_name, library_name=args.library_name, trust_remote_code=args.trust_remote_code)
    except (RuntimeError, OSError) as e:
        library = getattr(e, "library_name", None)
        if library is None:
            raise
        raise OSError(f"Tried to load `{args.model_name}` with `{library}` but a possible model to load was not found inside the repo.")
    total_size, largest_layer = calculate_maximum_sizes(model)
    data = []
    for dtype in args.dtypes:
        dtype_total_size = total_size
        dtype_largest_layer = largest_layer[0]
        dtype_training_size = estimate_training_usage(dtype_total_size, dtype)
        if dtype in ("float16", "bfloat16"):
            dtype_total_size /= 2
            dtype_largest_layer /= 2
        elif dtype == "int8":
            dtype_total_size /= 4
            dtype_largest_layer /= 4
        elif dtype == "int4":
            dtype_total_size /= 8
            dtype_largest_layer /= 8
        data.append([dtype, dtype_largest_layer, dtype_total_size, dtype_training_size])
    return data


def estimate_command(args):
    data = gather_data(args)
    for row in data:
        for i, item in enumerate(row):
            if isinstance(item, (int, float)):
                row[i] = convert_bytes(item)
            elif isinstance(item, dict):
                training_usage = max(item.values())
                row[i] = convert_bytes(training_usage) if training_usage != -1 else "N/A"
    headers = ["dtype", "Largest Layer", "Total Size", "Training using Adam"]
    title = f"Memory Usage for loading `{args.model_name}`"
    table = create_ascii_table(headers, data, title)
    print(table)


def main():
    parser = estimate_command_parser()
    args = parser.parse_args()
    estimate_command(args)


if __name__ == "__main__":
    main()
