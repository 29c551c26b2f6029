"""argparse helpers of the CLI (reference src/accelerate/commands/utils.py:18-123).

``CustomArgumentParser`` accepts ``--foo-bar`` and ``--foo_bar`` interchangeably and records which
options the user typed explicitly in ``namespace.nondefault`` — that set decides whether a CLI flag
overrides the YAML config file (reference commands/launch.py:1267).
"""

import argparse


def _note(namespace, dest):
    if not hasattr(namespace, "nondefault") or namespace.nondefault is None:
        namespace.nondefault = set()
    namespace.nondefault.add(dest)


class _StoreAction(argparse._StoreAction):
    def __call__(self, parser, namespace, values, option_string=None):
        super().__call__(parser, namespace, values, option_string)
        _note(namespace, self.dest)


class _StoreTrueAction(argparse._StoreTrueAction):
    def __call__(self, parser, namespace, values, option_string=None):
        super().__call__(parser, namespace, values, option_string)
        _note(namespace, self.dest)


class _StoreConstAction(argparse._StoreConstAction):
    def __call__(self, parser, namespace, values, option_string=None):
        super().__call__(parser, namespace, values, option_string)
        _note(namespace, self.dest)


def _with_dash_aliases(args):
    """``("--mixed_precision",)`` -> ``("--mixed_precision", "--mixed-precision")``."""
    out = list(args)
    for a in args:
        if isinstance(a, str) and a.startswith("--") and "_" in a[2:]:
            alias = "--" + a[2:].replace("_", "-")
            if alias not in out:
                out.append(alias)
    return tuple(out)


def _patch_kwargs(kwargs):
    action = kwargs.get("action", None)
    if action is None or action == "store":
        kwargs["action"] = _StoreAction
    elif action == "store_true":
        kwargs["action"] = _StoreTrueAction
    elif action == "store_const":
        kwargs["action"] = _StoreConstAction
    return kwargs


class CustomArgumentGroup(argparse._ArgumentGroup):
    def add_argument(self, *args, **kwargs):
        return super().add_argument(*_with_dash_aliases(args), **_patch_kwargs(kwargs))


class CustomArgumentParser(argparse.ArgumentParser):
    def add_argument(self, *args, **kwargs):
        return super().add_argument(*_with_dash_aliases(args), **_patch_kwargs(kwargs))

    def add_argument_group(self, *args, **kwargs):
        group = CustomArgumentGroup(self, *args, **kwargs)
        self._action_groups.append(group)
        return group

    def parse_args(self, args=None, namespace=None):
        ns = super().parse_args(args, namespace)
        if not hasattr(ns, "nondefault"):
            ns.nondefault = set()
        return ns


class CustomHelpFormatter(argparse.HelpFormatter):
    """Help output that hides the sections not relevant to the flags already on the command line."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.titles = ["Hardware Selection Arguments", "Resource Selection Arguments", "Training Paradigm Arguments",
                       "positional arguments", "optional arguments", "options"]

    def add_argument(self, action: argparse.Action):
        import sys

        if "accelerate" in sys.argv[0] and "launch" in sys.argv[1:]:
            args = sys.argv[2:]
        else:
            args = sys.argv[1:]
        if len(args) > 1:
            args = list(map(lambda a: a.replace("-", "_") if a.startswith("--") else a, args))
            used_platforms = [arg for arg in args if arg in ("__multi_gpu", "__cpu", "__use_deepspeed", "__use_fsdp", "__use_megatron_lm")]
            titles = list(self.titles)
            if "__cpu" not in args and len(used_platforms) > 0:
                titles.append("Distributed GPUs")
            for plat, title in (("__use_deepspeed", "DeepSpeed Arguments"), ("__use_fsdp", "FSDP Arguments"), ("__use_megatron_lm", "Megatron-LM Arguments")):
                if plat in used_platforms:
                    titles.append(title)
            if action.container.title not in titles:
                action.help = argparse.SUPPRESS
            elif action.container.title == "Hardware Selection Arguments" and set(action.option_strings).isdisjoint(set(args)):
                action.help = argparse.SUPPRESS
        super().add_argument(action)

    def end_section(self):
        if len(self._current_section.items) < 2:
            self._current_section.items = []
            self._current_section.heading = ""
        super().end_section()
