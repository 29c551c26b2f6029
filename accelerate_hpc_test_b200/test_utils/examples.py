"""Example-script coverage check: does a "complete" example contain what a by-feature example adds?

``compare_against_test(base, feature, parser_only, secondary=None)`` returns the lines that the feature script adds on
top of the template script (``examples/nlp_example.py``) and that are *missing* from the complete script — an empty
list means the complete example covers the feature (reference test_utils/examples.py:63-148, used by
tests/test_examples.py:70-155).  The implementation works on sets of normalised source lines of one function
(``main`` when ``parser_only`` else ``training_function``).
"""

from __future__ import annotations

import os
from typing import Iterable, Optional

_DATALOADER_LINE = "train_dataloader, eval_dataloader = get_dataloaders(accelerator, batch_size)"
_MOCK_MARKER = "TESTING_MOCKED_DATALOADERS"


def get_function_contents_by_name(lines: list, name: str) -> list:
    """Source lines of ``training_function`` (up to ``def main``) or ``main`` (up to ``if __name__``)."""
    stops = {"training_function": "def main", "main": "if __name__"}
    if name not in stops:
        raise ValueError(f"Incorrect function name passed: {name}, choose either 'main' or 'training_function'")
    out, inside = [], False
    for line in lines:
        if not inside:
            if f"def {name}" in line:
                inside = True
                out.append(line)
            continue
        if stops[name] in line:
            break
        out.append(line)
    return out


def clean_lines(lines: Iterable[str]) -> list:
    """Drop comment-only and empty lines."""
    return [ln for ln in lines if ln != "\n" and not ln.lstrip().startswith("#")]


def _function_of(path, name):
    with open(path) as f:
        return clean_lines(get_function_contents_by_name(f.readlines(), name))


def _additions(func_lines: list, template: set) -> list:
    """Lines of ``func_lines`` that the template does not have (mocked-dataloader scaffolding excluded)."""
    added, skip_next = [], False
    for line in func_lines:
        if skip_next:
            skip_next = False
            continue
        if line in template or line.strip() == _DATALOADER_LINE:
            continue
        if _MOCK_MARKER in line:
            skip_next = True  # the statement guarded by the marker belongs to the test harness
            continue
        added.append(line)
    return added


def compare_against_test(base_filename: str, feature_filename: str, parser_only: bool, secondary_filename: Optional[str] = None):
    name = "main" if parser_only else "training_function"
    template_path = os.path.abspath(os.path.join("examples", "nlp_example.py"))
    with open(template_path) as f:
        template_all = f.readlines()
    template = set(clean_lines(get_function_contents_by_name(template_all, name)))
    feature_added = _additions(_function_of(feature_filename, name), template)
    complete_added = set(_additions(_function_of(base_filename, name), template))
    missing = [ln for ln in feature_added if ln not in complete_added]
    if secondary_filename is not None:
        secondary = set(_function_of(secondary_filename, name))
        template_only = {ln for ln in template_all if ln not in secondary}
        missing = [ln for ln in missing if ln not in template_only]
    return missing
