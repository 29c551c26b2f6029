"""Repository tools that run without a GPU: the test-report summariser (``tools/log_reports.py``) and the packaging metadata."""

import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    module = importlib.util.module_from_spec(spec)
    sys.modules[name] = module
    spec.loader.exec_module(module)
    return module


def test_report_summary_counts_junit_and_report_log_and_fails_on_failures(tmp_path, capsys):
    reports = _load(os.path.join(ROOT, "tools", "log_reports.py"), "b200_log_reports")
    (tmp_path / "core.xml").write_text(
        '<testsuites><testsuite name="pytest">'
        '<testcase classname="tests.test_a" name="test_ok" time="1.5"/>'
        '<testcase classname="tests.test_a" name="test_skip" time="0.0"><skipped message="needs a GPU"/></testcase>'
        '<testcase classname="tests.test_a" name="test_bad" time="2.5"><failure message="AssertionError: 1 != 2&#10;more">trace</failure></testcase>'
        '<testcase classname="tests.test_b" name="test_err" time="0.5"><error>\n  RuntimeError: boom\n</error></testcase>'
        "</testsuite></testsuites>")
    lines = [
        {"$report_type": "SessionStart"},
        {"$report_type": "TestReport", "nodeid": "t.py::one", "when": "setup", "outcome": "passed", "duration": 0.1},
        {"$report_type": "TestReport", "nodeid": "t.py::one", "when": "call", "outcome": "passed", "duration": 1.0},
        {"$report_type": "TestReport", "nodeid": "t.py::two", "when": "setup", "outcome": "skipped", "duration": 0.0, "longrepr": ["f", 1, "skip"]},
        {"$report_type": "TestReport", "nodeid": "t.py::three", "when": "call", "outcome": "failed", "duration": 3.0,
         "longrepr": {"reprcrash": {"message": "ValueError: nope"}}},
        {"$report_type": "TestReport", "nodeid": "t.py::three", "when": "teardown", "outcome": "passed", "duration": 0.0},
    ]
    (tmp_path / "cli.log").write_text("\n".join(json.dumps(line) for line in lines) + "\nnot json\n")

    groups = {g.name: g for g in reports.collect(tmp_path)}
    core, cli = groups["core"], groups["cli"]
    assert (core.passed, core.failed, core.skipped, core.total) == (1, 2, 1, 4) and abs(core.seconds - 4.5) < 1e-9
    assert core.failures == [("tests.test_a::test_bad", "AssertionError: 1 != 2"), ("tests.test_b::test_err", "RuntimeError: boom")]
    assert (cli.passed, cli.failed, cli.skipped) == (1, 1, 1) and cli.failures[0][0] == "t.py::three" and "nope" in cli.failures[0][1]

    summary = tmp_path / "summary.json"
    assert reports.main([str(tmp_path), "--json", str(summary), "--title", "nightly"]) == 1
    text = capsys.readouterr().out
    assert "## nightly" in text and "**3 failed** of 7 tests in 2 groups" in text and "| core | 1 | 2 | 1 |" in text
    assert "`tests.test_b::test_err` — RuntimeError: boom" in text
    assert [g["name"] for g in json.loads(summary.read_text())] == ["core", "cli"]

    (tmp_path / "core.xml").unlink()
    (tmp_path / "cli.log").unlink()
    (tmp_path / "ok.xml").write_text('<testsuite><testcase classname="a" name="b" time="60"/></testsuite>')
    assert reports.main([str(tmp_path)]) == 0 and "all 1 tests passed" in capsys.readouterr().out
    empty = tmp_path / "empty"
    empty.mkdir()
    assert reports.main([str(empty)]) == 0 and "no reports found" in capsys.readouterr().out


def test_packaging_declares_the_reference_console_scripts_and_the_package_version():
    import accelerate_hpc_test_b200 as pkg

    env = dict(os.environ, B200_SKIP_KERNEL_BUILD="1")
    version = subprocess.run([sys.executable, "setup.py", "--version"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert version.returncode == 0 and version.stdout.strip().splitlines()[-1] == pkg.__version__
    source = open(os.path.join(ROOT, "setup.py")).read()
    for script, target in [("accelerate", "commands.accelerate_cli:main"), ("accelerate-config", "commands.config:main"),
                           ("accelerate-estimate-memory", "commands.estimate:main"), ("accelerate-launch", "commands.launch:main"),
                           ("accelerate-merge-weights", "commands.merge:main")]:
        assert f'"{script}={{PACKAGE}}.{target}"' in source
        module_name, func = target.split(":")
        module = importlib.import_module(f"accelerate_hpc_test_b200.{module_name}")
        assert callable(getattr(module, func))


def test_built_library_contains_the_blackwell_instructions_it_claims():
    """``cuobjdump -sass`` on the built kernel library (no GPU needed): the GEMM and attention kernels issue tcgen05 MMAs fed by TMA
    tensor loads with accumulators read back from tensor memory, the NVLS collectives reduce inside the switch (``multimem.ld_reduce``),
    the peer-memory copy kernels move data with bulk asynchronous copies.  A refactor that silently falls back to ``mma.sync`` or to
    plain loads fails here, before any GPU time is spent."""
    import shutil

    import pytest

    lib = os.path.join(ROOT, "accelerate_hpc_test_b200", "lib", "libb200.so")
    if not os.path.exists(lib) or shutil.which("cuobjdump") is None:
        pytest.skip("needs the built kernel library and cuobjdump")
    listing = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_evidence.py")], cwd=ROOT, capture_output=True, text=True,
                             timeout=600)
    assert listing.returncode == 0, listing.stderr[-2000:]
    per_kernel = {}
    for line in listing.stdout.splitlines():
        name, _, evidence = line.partition(": ")
        per_kernel[name.strip()] = evidence

    def kernels(fragment):
        found = {name: evidence for name, evidence in per_kernel.items() if fragment in name}
        assert found, f"no kernel named *{fragment}* in the library"
        return found

    for name, evidence in kernels("gemm2_kernel").items():                      # the 2-CTA tcgen05 GEMM family
        assert ("UTCHMMA" in evidence or "UTCQMMA" in evidence) and "UTMALDG" in evidence and "LDTM" in evidence, (name, evidence)  # bf16 / fp8
        assert ".2CTA" in evidence, (name, "cta_group::2 expected")
        assert " HMMA" not in " " + evidence, (name, "legacy mma.sync in a tcgen05 kernel")
    for fragment in ("flash_fwd_kernel", "flash_bwd_kernel"):
        for name, evidence in kernels(fragment).items():
            if "prep" in name:
                continue
            assert "UTCHMMA" in evidence and "UTMALDG" in evidence and "MUFU.EX2" in evidence, (name, evidence)
    for fragment in ("allreduce_nvls_kernel", "reduce_scatter_adamw_nvls_kernel"):
        for name, evidence in kernels(fragment).items():
            assert "LDGMC.E.HPADD.BF16x8" in evidence, (name, evidence)            # multimem.ld_reduce: the NVSwitch adds
    for fragment in ("allgather_pull_kernel", "seq_head_exchange_kernel"):
        for name, evidence in kernels(fragment).items():
            assert "UBLKCP" in evidence and "SYNCS" in evidence, (name, evidence)      # cp.async.bulk completed through mbarriers
