"""Small modules that nothing else exercises: the arrow-key menu (driven through a fake key source and through the non-TTY
fall-back), progress bars that stay quiet off the main process, the rich-traceback switch, and the stubs that say what this build
does not support (TPU, SageMaker, XLA)."""

import io
import sys

import pytest


def test_bullet_menu_follows_keys_and_wraps_around(monkeypatch, capsys):
    from accelerate_hpc_test_b200.commands.menu import selection_menu as menu

    class Terminal(io.StringIO):
        def isatty(self):
            return True

    def play(keys, default=0, choices=("no", "fp16", "bf16", "fp8")):
        keys = iter(keys)
        monkeypatch.setattr(menu, "_read_key", lambda: next(keys))
        monkeypatch.setattr(sys, "stdin", Terminal())
        monkeypatch.delenv("ACCELERATE_DISABLE_RICH_MENU", raising=False)
        return menu.BulletMenu("Mixed precision?", list(choices)).run(default_choice=default)

    assert play([menu.KEY_ENTER], default=2) == 2
    assert play([menu.KEY_DOWN, menu.KEY_DOWN, menu.KEY_ENTER]) == 2
    assert play([menu.KEY_UP, menu.KEY_ENTER]) == 3                      # up from the first entry wraps to the last
    assert play([menu.KEY_DOWN, menu.KEY_ENTER], default=3) == 0         # and down from the last to the first
    assert play(["1", menu.KEY_ENTER]) == 1                              # a digit jumps
    assert play(["9", "x", "", menu.KEY_ENTER], default=1) == 1          # out of range digits and other keys change nothing
    drawn = capsys.readouterr().out
    assert "Mixed precision?" in drawn and "arrow or number keys" in drawn and "\033[32m" in drawn and menu.CLEAR_LINE in drawn
    with pytest.raises(KeyboardInterrupt):
        play([menu.KEY_INTERRUPT])


def test_bullet_menu_without_a_terminal_asks_for_a_number(monkeypatch, capsys):
    from accelerate_hpc_test_b200.commands.menu.selection_menu import BulletMenu

    answers = iter(["seven", "7", "2"])
    monkeypatch.setattr("builtins.input", lambda prompt="": next(answers))
    monkeypatch.setattr(sys, "stdin", io.StringIO())                    # not a TTY
    assert BulletMenu("Which?", ["a", "b", "c"]).run(default_choice=0) == 2
    shown = capsys.readouterr().out
    assert "[0] a" in shown and shown.count("Please enter a valid number.") == 2
    monkeypatch.setattr("builtins.input", lambda prompt="": "")
    assert BulletMenu("Which?", ["a", "b", "c"]).run(default_choice=1) == 1
    # a terminal can be told to use the plain prompt too
    class Terminal(io.StringIO):
        def isatty(self):
            return True

    monkeypatch.setattr(sys, "stdin", Terminal())
    monkeypatch.setenv("ACCELERATE_DISABLE_RICH_MENU", "1")
    assert BulletMenu("Which?", ["a", "b"]).run(default_choice=0) == 0


def test_progress_bars_are_silent_off_the_local_main_process(monkeypatch):
    import importlib

    tqdm_module = importlib.import_module("accelerate_hpc_test_b200.utils.tqdm")     # (`utils.tqdm` the attribute is the function)
    tqdm = tqdm_module.tqdm

    class OnRank:
        def __init__(self, rank):
            self.local_process_index = rank

    monkeypatch.setattr(tqdm_module, "PartialState", lambda: OnRank(0))
    bar = tqdm(range(3), file=io.StringIO())
    assert bar.disable is False and list(bar) == [0, 1, 2]
    monkeypatch.setattr(tqdm_module, "PartialState", lambda: OnRank(1))
    assert tqdm(range(3), file=io.StringIO()).disable is True
    assert tqdm(range(3), main_process_only=False, file=io.StringIO()).disable is False
    assert tqdm(range(3), main_process_only=False, disable=True).disable is True
    with pytest.raises(ValueError, match="no longer supported"):
        tqdm(True, range(3))


def test_rich_tracebacks_are_opt_in(monkeypatch):
    import importlib

    monkeypatch.delenv("ACCELERATE_ENABLE_RICH", raising=False)
    sys.modules.pop("accelerate_hpc_test_b200.utils.rich", None)
    with pytest.raises(ModuleNotFoundError, match="install rich"):
        importlib.import_module("accelerate_hpc_test_b200.utils.rich")
    pytest.importorskip("rich")
    monkeypatch.setenv("ACCELERATE_ENABLE_RICH", "1")
    hook_before = sys.excepthook
    try:
        importlib.import_module("accelerate_hpc_test_b200.utils.rich")
        assert sys.excepthook is not hook_before
    finally:
        sys.excepthook = hook_before
        sys.modules.pop("accelerate_hpc_test_b200.utils.rich", None)


def test_unsupported_back_ends_say_so():
    from accelerate_hpc_test_b200.commands.accelerate_cli import build_parser
    from accelerate_hpc_test_b200.commands.config.sagemaker import get_sagemaker_input
    from accelerate_hpc_test_b200.utils.torch_xla import install_xla

    with pytest.raises(NotImplementedError, match="SageMaker is not a supported compute environment"):
        get_sagemaker_input()
    with pytest.raises(NotImplementedError, match="TPU back-ends are not supported"):
        install_xla()
    args = build_parser().parse_args(["tpu-config"])
    with pytest.raises(SystemExit, match="no TPU back-end"):
        args.func(args)


def test_deprecated_memory_utils_module_forwards_with_a_warning():
    import importlib

    sys.modules.pop("accelerate_hpc_test_b200.memory_utils", None)
    with pytest.warns(FutureWarning, match="memory_utils has been reorganized to utils.memory"):
        module = importlib.import_module("accelerate_hpc_test_b200.memory_utils")
    from accelerate_hpc_test_b200.utils.memory import find_executable_batch_size

    assert module.find_executable_batch_size is find_executable_batch_size
