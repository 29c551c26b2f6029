#!/usr/bin/env python
"""Summarise test reports into one table: which groups ran, how many tests passed / failed / were skipped, how long each took, and the
failing tests with the first line of their message.

Reads every ``*.xml`` (``pytest --junitxml``, what ``make test REPORT=1`` writes) and every ``*.log`` (``pytest --report-log`` JSON
lines, the format the reference's CI produces) in the given directory.  Prints GitHub-flavoured markdown, optionally writes the same
data as JSON, and exits 1 when anything failed — so a CI job can end with it.  The counterpart of the reference's
``utils/log_reports.py`` (which posts the table to a chat webhook; there is no network here, the table goes to stdout / a file).

    python tools/log_reports.py reports [--json reports/summary.json] [--title "nightly"]
"""

from __future__ import annotations

import argparse
import json
import sys
import xml.etree.ElementTree as ET
from dataclasses import asdict, dataclass, field
from datetime import date
from pathlib import Path


@dataclass
class Group:
    name: str
    passed: int = 0
    failed: int = 0
    skipped: int = 0
    seconds: float = 0.0
    failures: list = field(default_factory=list)     # (test id, first line of the message)

    @property
    def total(self) -> int:
        return self.passed + self.failed + self.skipped


def _first_line(text: str, limit: int = 160) -> str:
    for line in (text or "").splitlines():
        line = line.strip()
        if line:
            return line[:limit]
    return ""


def read_junit(path: Path) -> Group:
    group = Group(path.stem)
    root = ET.parse(path).getroot()
    for case in root.iter("testcase"):
        group.seconds += float(case.get("time") or 0.0)
        test_id = "::".join(part for part in (case.get("classname"), case.get("name")) if part)
        problem = case.find("failure")
        if problem is None:
            problem = case.find("error")
        if problem is not None:
            group.failed += 1
            group.failures.append((test_id, _first_line(problem.get("message") or problem.text)))
        elif case.find("skipped") is not None:
            group.skipped += 1
        else:
            group.passed += 1
    return group


def read_report_log(path: Path) -> Group:
    """One JSON object per line; a test is the set of its setup / call / teardown reports, failed if any phase failed."""
    group = Group(path.stem)
    outcome, message = {}, {}
    for raw in path.read_text().splitlines():
        try:
            entry = json.loads(raw)
        except json.JSONDecodeError:
            continue
        if entry.get("$report_type") != "TestReport":
            continue
        node = entry["nodeid"]
        group.seconds += float(entry.get("duration") or 0.0)
        result = entry.get("outcome")
        if result == "failed":
            outcome[node] = "failed"
            longrepr = entry.get("longrepr")
            text = longrepr if isinstance(longrepr, str) else json.dumps(((longrepr or {}).get("reprcrash") or {}).get("message", ""))
            message.setdefault(node, _first_line(text))
        elif result == "skipped" and outcome.get(node) != "failed":
            outcome[node] = "skipped"
        elif entry.get("when") == "call":
            outcome.setdefault(node, "passed")
    for node, result in outcome.items():
        setattr(group, result, getattr(group, result) + 1)
        if result == "failed":
            group.failures.append((node, message.get(node, "")))
    return group


def collect(directory: Path) -> list:
    groups = [read_junit(p) for p in sorted(directory.glob("*.xml"))]
    groups += [read_report_log(p) for p in sorted(directory.glob("*.log"))]
    return groups


def render(groups: list, title: str) -> str:
    lines = [f"## {title} — {date.today().isoformat()}", ""]
    if not groups:
        return "\n".join(lines + ["no reports found"])
    failed = sum(g.failed for g in groups)
    lines.append(f"**{failed} failed** of {sum(g.total for g in groups)} tests in {len(groups)} groups" if failed
                 else f"all {sum(g.passed for g in groups)} tests passed ({sum(g.skipped for g in groups)} skipped) in {len(groups)} groups")
    lines += ["", "| group | passed | failed | skipped | minutes |", "|---|---:|---:|---:|---:|"]
    for g in groups:
        lines.append(f"| {g.name} | {g.passed} | {g.failed} | {g.skipped} | {g.seconds / 60:.1f} |")
    for g in groups:
        if g.failures:
            lines += ["", f"### {g.name}", ""]
            lines += [f"* `{test}` — {why}" if why else f"* `{test}`" for test, why in g.failures]
    return "\n".join(lines)


def main(argv=None) -> int:
    parser = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    parser.add_argument("directory", nargs="?", default="reports")
    parser.add_argument("--json", dest="json_path", default=None, help="also write the summary as JSON to this path")
    parser.add_argument("--title", default="test report")
    args = parser.parse_args(argv)
    groups = collect(Path(args.directory))
    print(render(groups, args.title))
    if args.json_path:
        Path(args.json_path).write_text(json.dumps([{**asdict(g), "total": g.total} for g in groups], indent=1))
    return 1 if any(g.failed for g in groups) else 0


if __name__ == "__main__":
    sys.exit(main())
