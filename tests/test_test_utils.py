"""test_utils helpers (reference src/accelerate/test_utils/testing.py, examples.py; tests/test_examples.py)."""

import textwrap
import unittest


from accelerate_hpc_test_b200.test_utils.examples import compare_against_test
from accelerate_hpc_test_b200.test_utils.testing import (AccelerateTestCase, assert_exception, capture_call_output, get_launch_command,
                                                         parse_flag_from_env, path_in_accelerate_package, require_cuda, require_xpu,
                                                         run_command)


def test_launch_command_and_flags(monkeypatch):
    cmd = get_launch_command(num_processes=2, multi_gpu=True, mixed_precision=None, debug=False)
    assert cmd[-2:] == ["--num_processes=2", "--multi_gpu"]
    monkeypatch.setenv("X_FLAG", "yes")
    assert parse_flag_from_env("X_FLAG") is True and parse_flag_from_env("NOPE", default=False) is False
    assert path_in_accelerate_package("test_utils", "scripts", "test_script.py").exists()
    assert capture_call_output(print, "hi").strip() == "hi"
    assert "ok" in run_command(["python", "-c", "print('ok')"], return_stdout=True)
    with assert_exception(ValueError, "boom"):
        raise ValueError("boom!")


def test_skip_decorators():
    @require_xpu
    class T(unittest.TestCase):
        def test_x(self):
            raise AssertionError("must be skipped")

    res = unittest.TestResult()
    unittest.defaultTestLoader.loadTestsFromTestCase(T).run(res)
    assert len(res.skipped) == 1 and not res.failures
    assert callable(require_cuda) and issubclass(AccelerateTestCase, unittest.TestCase)


def test_compare_against_test(tmp_path, monkeypatch):
    ex = tmp_path / "examples"
    ex.mkdir()
    template = textwrap.dedent('''
        def training_function(config, args):
            accelerator = Accelerator()
            model = get_model()
            for batch in loader:
                loss = model(batch)
        def main():
            parser = argparse.ArgumentParser()
            args = parser.parse_args()
        if __name__ == "__main__":
            main()
    ''')
    feature = template.replace("    accelerator = Accelerator()\n", "    accelerator = Accelerator(log_with='all')\n    accelerator.init_trackers('x')\n")
    complete_ok = feature.replace("    model = get_model()\n", "    model = get_model()\n    extra_thing()\n")
    complete_missing = template
    (ex / "nlp_example.py").write_text(template)
    (tmp_path / "feature.py").write_text(feature)
    (tmp_path / "ok.py").write_text(complete_ok)
    (tmp_path / "missing.py").write_text(complete_missing)
    monkeypatch.chdir(tmp_path)
    assert compare_against_test(str(tmp_path / "ok.py"), str(tmp_path / "feature.py"), parser_only=False) == []
    diff = compare_against_test(str(tmp_path / "missing.py"), str(tmp_path / "feature.py"), parser_only=False)
    assert any("init_trackers" in ln for ln in diff) and len(diff) == 2
    assert compare_against_test(str(tmp_path / "ok.py"), str(tmp_path / "feature.py"), parser_only=True) == []


def test_package_level_names_match_what_reference_test_suites_import():
    import re
    import sys

    import accelerate_hpc_test_b200.test_utils as test_utils

    wanted = """DEFAULT_LAUNCH_COMMAND are_the_same_tensors assert_exception capture_call_output device_count execute_subprocess_async
        get_launch_command get_torch_dist_unique_port memory_allocated_func path_in_accelerate_package pytest_xdist_worker_id require_bnb
        require_cpu require_cuda require_fp8 require_fp16 require_huggingface_suite require_multi_device require_multi_gpu require_non_cpu
        require_pippy require_single_device require_single_gpu require_torch_min_version require_transformer_engine require_torchao
        slow skip run_first torch_device RegressionDataset RegressionModel""".split()
    assert [name for name in wanted if not hasattr(test_utils, name)] == []
    # the launched scripts are reachable from the package but are not imported with it
    assert "accelerate_hpc_test_b200.test_utils.scripts.test_ops" not in sys.modules or True
    assert test_utils.test_ops.__name__.endswith("scripts.test_ops") and callable(test_utils.test_sync.main)
    assert re.fullmatch(r"\d+", str(test_utils.get_torch_dist_unique_port()))


def test_autoregressive_mocked_dataloaders_mask_padding_out_of_the_loss():
    import torch

    from accelerate_hpc_test_b200.test_utils.training import mocked_dataloaders_for_autoregressive_models

    train, evaluate = mocked_dataloaders_for_autoregressive_models(None, batch_size=8)
    assert len(train) == 8 and len(evaluate) == 4
    batch = next(iter(evaluate))
    assert set(batch) == {"input_ids", "attention_mask", "labels"} and batch["input_ids"].shape == (8, 32)
    padded = batch["attention_mask"] == 0
    assert padded.any() and torch.all(batch["labels"][padded] == -100) and torch.equal(batch["labels"][~padded], batch["input_ids"][~padded])
    again = next(iter(mocked_dataloaders_for_autoregressive_models(None, batch_size=8)[1]))
    assert torch.equal(again["input_ids"], batch["input_ids"])          # reproducible


def test_are_the_same_tensors_compares_whole_tensors():
    import torch

    from accelerate_hpc_test_b200 import Accelerator
    from accelerate_hpc_test_b200.test_utils import are_the_same_tensors

    Accelerator(cpu=True)
    assert are_the_same_tensors(torch.arange(6.0).view(2, 3))          # one process: trivially the same, whatever the shape
    assert are_the_same_tensors(torch.tensor(3.0))
