#!/usr/bin/env python
"""Write docs/package_reference/api.md from the package itself: public names per module with the first line of their docstring,
and the public methods of the classes a user talks to directly.

    python tools/gen_api_reference.py
"""
import importlib
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "accelerate_hpc_test_b200"
MODULES = [
    ("accelerator", "Accelerator"), ("state", "Process state"), ("data_loader", "Data loading"), ("optimizer", "Optimizer wrapper"),
    ("scheduler", "Scheduler wrapper"), ("checkpointing", "Checkpoint files"), ("parallelism_config", "Parallelism configuration"),
    ("big_modeling", "Big-model API"), ("hooks", "Hooks"), ("inference", "Pipeline inference"), ("launchers", "In-process launchers"),
    ("tracking", "Trackers"), ("logging", "Logging"), ("local_sgd", "Local SGD"), ("optim.adamw", "Fused optimizer"),
    ("models.llama", "Native Llama"), ("models.bert", "Native BERT"),
    ("ops.functional", "Fused ops (autograd functions over the sm_100a kernels)"), ("parallel.fsdp", "Sharded engine"),
    ("parallel.ddp", "Data-parallel engine"), ("parallel.tensor_parallel", "Tensor parallelism"),
    ("parallel.ulysses", "Ulysses sequence parallelism"), ("parallel.context_parallel", "Context parallelism"),
    ("parallel.peer_collectives", "Peer-memory collectives"), ("parallel.symm", "Symmetric workspace"), ("parallel.precision", "fp8 recipes"),
    ("utils.dataclasses", "Plugins and kwargs handlers"), ("utils.operations", "Tensor-tree collectives"), ("utils.modeling", "Model surgery"),
    ("utils.offload", "Offload store"), ("utils.fsdp_utils", "FSDP utilities"), ("utils.memory", "Memory helpers"),
    ("utils.random", "Randomness"), ("utils.environment", "Environment"), ("utils.diagnostics", "Diagnostics"),
    ("utils.deepspeed", "DeepSpeed-style configuration"), ("utils.megatron_lm", "Megatron-style engine"),
]
WITH_METHODS = {"Accelerator", "PartialState", "GradientState", "B200ShardedModule", "B200DataParallel", "FusedAdamW", "GeneralTracker"}


def first_line(obj) -> str:
    # a class's own docstring only: `inspect.getdoc` would hand every undocumented nn.Module torch's "Base class for all ..."
    text = obj.__dict__.get("__doc__") if inspect.isclass(obj) else inspect.getdoc(obj)
    return " ".join((text or "").strip().split("\n\n")[0].split())[:200]


def signature(obj) -> str:
    try:
        text = str(inspect.signature(obj))
    except (TypeError, ValueError):
        return "(...)"
    return text if len(text) <= 100 else text[:97] + "...)"


# topic pages: the same sections grouped the way the reference's docs/source/package_reference/ is laid out
TOPICS = [
    ("accelerator", "Accelerator", "The object a training script talks to.", ["accelerator"]),
    ("state", "Process and training state", "Shared-state singletons behind the `Accelerator`.", ["state", "parallelism_config"]),
    ("torch_wrappers", "Wrappers around torch objects", "What `prepare` returns for dataloaders, optimizers and schedulers.",
     ["data_loader", "optimizer", "scheduler"]),
    ("checkpointing", "Checkpoint files", "`save_state` / `load_state` layout and helpers.", ["checkpointing"]),
    ("big_modeling", "Big-model inference", "Empty-weight construction, device maps, offload, hooks.",
     ["big_modeling", "hooks", "utils.modeling", "utils.offload", "utils.memory"]),
    ("inference", "Pipeline inference", "`prepare_pippy` and the native stage machinery.", ["inference"]),
    ("launchers", "In-process launchers", "`notebook_launcher`, `debug_launcher`.", ["launchers"]),
    ("tracking", "Experiment trackers", "The tracker contract and the built-in trackers.", ["tracking"]),
    ("logging", "Logging", "Rank-aware logging.", ["logging"]),
    ("kwargs", "Plugins and kwargs handlers", "Dataclasses that configure engines, precision and process groups.", ["utils.dataclasses"]),
    ("fsdp", "Sharded training", "The native FSDP2-style engine and its utilities.", ["parallel.fsdp", "utils.fsdp_utils"]),
    ("ddp", "Data-parallel training", "The native bucketed all-reduce engine.", ["parallel.ddp", "local_sgd"]),
    ("parallelism", "Tensor / sequence / context parallelism", "Model-parallel layers over peer memory.",
     ["parallel.tensor_parallel", "parallel.ulysses", "parallel.context_parallel", "parallel.peer_collectives", "parallel.symm"]),
    ("fp8", "Low precision", "MXFP8 / e4m3 recipes on the tensor cores.", ["parallel.precision"]),
    ("deepspeed", "DeepSpeed-style configuration", "Configuration dialect executed by the native engines.", ["utils.deepspeed"]),
    ("megatron_lm", "Megatron-style engine", "Engine-style API and argument translation.", ["utils.megatron_lm"]),
    ("models_and_ops", "Native models, optimizer and ops", "What runs on the sm_100a kernels.",
     ["models.llama", "models.bert", "optim.adamw", "ops.functional"]),
    ("utilities", "Utilities", "Tensor-tree collectives, randomness, environment, diagnostics.",
     ["utils.operations", "utils.random", "utils.environment", "utils.diagnostics"]),
]


def section(short, title):
    """Markdown lines documenting one module, or [] when it cannot be imported (optional dependency missing)."""
    name = f"{PKG}.{short}"
    try:
        module = importlib.import_module(name)
    except Exception:
        return []
    out = [f"## {title} — `{name}`", ""]
    if first_line(module):
        out += [first_line(module), ""]
    members = [(n, v) for n, v in vars(module).items()
               if not n.startswith("_") and (inspect.isclass(v) or inspect.isfunction(v)) and getattr(v, "__module__", None) == name]
    for n, v in sorted(members):
        if inspect.isclass(v):
            out.append(f"* class `{n}` — {first_line(v)}")
            if n in WITH_METHODS:
                for m, f in sorted(vars(v).items()):
                    if m.startswith("_") or not (inspect.isfunction(f) or isinstance(f, property)):
                        continue
                    if isinstance(f, property):
                        out.append(f"    * `{m}` (property) — {first_line(f.fget) if f.fget else ''}")
                    else:
                        out.append(f"    * `{m}{signature(f)}` — {first_line(f)}")
        else:
            out.append(f"* `{n}{signature(v)}` — {first_line(v)}")
    out.append("")
    return out


def main():
    folder = os.path.join(ROOT, "docs", "package_reference")
    titles = dict(MODULES)
    out = ["# API reference", "", "Generated by `python tools/gen_api_reference.py`: public names per module with the first line of "
           "their docstring; the classes a user talks to directly also list their public methods.  The same content by topic: "
           + ", ".join(f"[{slug}]({slug}.md)" for slug, *_ in TOPICS) + ".", ""]
    for short, title in MODULES:
        out += section(short, title)
    with open(os.path.join(folder, "api.md"), "w") as handle:
        handle.write("\n".join(out))
    print(f"{folder}/api.md: {len(out)} lines")
    for slug, heading, blurb, modules in TOPICS:
        page = [f"# {heading}", "", blurb + "  (Generated by `python tools/gen_api_reference.py`; the whole surface on one page: [api.md](api.md).)", ""]
        for short in modules:
            page += section(short, titles.get(short, short))
        with open(os.path.join(folder, f"{slug}.md"), "w") as handle:
            handle.write("\n".join(page))
    print(f"{len(TOPICS)} topic pages")


if __name__ == "__main__":
    main()
