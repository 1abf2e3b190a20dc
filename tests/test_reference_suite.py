"""Acceptance test of the drop-in boundary: the reference's OWN tests of
the hot path -- tests/test_newton_polynomial.py, test_transformation.py, test_interpolation.py,
test_multi_index.py, test_multi_index_utils.py of casus/minterpy, 1051 cases -- run unmodified
in a child pytest with `minterpy_b200.install()` active.

* CPU (here): the device plan is the oracle-backed stand-in, so this checks the wiring --
  rebound names, operator classes (`isinstance(op, OperatorABC)`,
  reference tests/test_transformation.py:60-81), shapes and errors.
* GPU (`-m gpu`): the same files against the CUDA library, with the pip-installed reference in
  baseline/_ref (made by tools/make_baseline_ref.py; it travels to the GPU box).
"""
import os
import re
import subprocess
import sys
import tempfile

import pytest

from ref_shim import reference_location, reference_tests_dir

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOT_PATH_FILES = ["test_newton_polynomial.py", "test_transformation.py", "test_interpolation.py",
                  "test_multi_index.py", "test_multi_index_utils.py"]
N_CASES = 1051

needs_reference = pytest.mark.skipif(
    reference_location() is None or reference_tests_dir() is None,
    reason="reference not available (run tools/make_baseline_ref.py where /root/reference exists)")


def run_suite(mode, timeout, files=None, expect=None):
    """Run reference test files in a child pytest; returns (passed, sorted failing test ids).
    With `expect` the run must pass completely with that many cases."""
    files = HOT_PATH_FILES if files is None else files
    tests_dir = reference_tests_dir()
    env = dict(os.environ)
    env["MPX_REF_SUITE_PLAN"] = mode
    env["PYTHONPATH"] = os.pathsep.join(
        [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")] +
        ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
    with tempfile.TemporaryDirectory() as tmp:
        # copies: the child must not pick up the reference's setup.cfg (coverage options) nor
        # write __pycache__ / .pytest_cache beside the reference
        for f in files + ["conftest.py"]:
            with open(os.path.join(tests_dir, f)) as src, open(os.path.join(tmp, f), "w") as dst:
                dst.write(src.read())
        ini = os.path.join(tmp, "pytest.ini")
        with open(ini, "w") as f:
            f.write("[pytest]\n")
        cmd = [sys.executable, "-m", "pytest", "-q", "-rf", "-p", "ref_suite_plugin", "-p",
               "no:cacheprovider", "-c", ini, "--rootdir", tmp] + \
              [os.path.join(tmp, f) for f in files]
        res = subprocess.run(cmd, env=env, cwd=tmp, capture_output=True, text=True, timeout=timeout)
    tail = "\n".join((res.stdout + res.stderr).splitlines()[-40:])
    summary = re.search(r"(\d+) passed", res.stdout)
    failing = sorted(set(re.findall(r"^(?:FAILED|ERROR) (\S+)", res.stdout, flags=re.M)))
    assert summary, tail
    if expect is not None:
        assert res.returncode == 0 and not failing, tail
        assert int(summary.group(1)) == expect, tail
    return int(summary.group(1)), [f.split("/")[-1] for f in failing]


def all_reference_test_files():
    return sorted(f for f in os.listdir(reference_tests_dir()) if f.startswith("test_") and f.endswith(".py"))


@needs_reference
def test_reference_hot_path_suite_with_install_on_cpu():
    run_suite("oracle", timeout=1500, expect=N_CASES)


@needs_reference
def test_whole_reference_suite_changes_nothing_else_on_cpu():
    """Every test file of the reference (2250 cases): with install() exactly the cases fail that
    fail in the unpatched reference in this environment (its GMRES solver cases under the
    installed SciPy), nothing else, and as many pass."""
    files = all_reference_test_files()
    passed_plain, failing_plain = run_suite("unpatched", timeout=1500, files=files)
    passed, failing = run_suite("oracle", timeout=1500, files=files)
    assert failing == failing_plain
    assert passed == passed_plain and passed >= 2000


@needs_reference
@pytest.mark.gpu
def test_reference_hot_path_suite_with_install_on_gpu():
    import torch
    assert torch.cuda.is_available()
    run_suite("cuda", timeout=2400, expect=N_CASES)


@needs_reference
@pytest.mark.gpu
def test_whole_reference_suite_changes_nothing_else_on_gpu():
    import torch
    assert torch.cuda.is_available()
    files = all_reference_test_files()
    passed_plain, failing_plain = run_suite("unpatched", timeout=1500, files=files)
    passed, failing = run_suite("cuda", timeout=2400, files=files)
    assert failing == failing_plain
    assert passed == passed_plain and passed >= 2000
