"""Install the UNMODIFIED reference (casus/minterpy) into git-ignored ``baseline/_ref``.

``baseline/_ref`` is not product source and stays out of history, but gpurun ships it to the
GPU box like the built ``.so`` files, so that there

* ``bench.py --impl reference`` can time the reference's own ``eval_newton_polynomials``, and
* ``tests/test_reference_suite.py`` can run the reference's own hot-path tests with
  ``minterpy_b200.install()`` routing them through the CUDA library.

Recipe (the one BASELINE/the bench contract names): a pip install of ``/root/reference`` from a
scratch copy (the checkout is read-only and the build writes an egg-info), no index, no
dependency resolution (numpy, numba, scipy, attrs are in the image).  The reference's test files
are copied beside it as ``baseline/_ref/reference_tests``.  Nothing is edited; numpy-2 aliases
and the ``minterpy.version`` module setuptools_scm would have generated are supplied at import
time by ``tests/golden/ref_shim.py``.
"""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = "/root/reference"
TARGET = os.path.join(ROOT, "baseline", "_ref")


def _make_writable(top):
    for root, dirs, files in os.walk(top):
        for name in dirs + files:
            path = os.path.join(root, name)
            try:
                os.chmod(path, os.stat(path).st_mode | 0o200)
            except OSError:
                pass
    if os.path.isdir(top):
        os.chmod(top, os.stat(top).st_mode | 0o200)


def make(force=False):
    """Returns TARGET, or None when /root/reference is absent (GPU box: prebuilt copy is used)."""
    if not os.path.isdir(os.path.join(REFERENCE, "src", "minterpy")):
        return TARGET if os.path.isdir(os.path.join(TARGET, "minterpy")) else None
    if os.path.isdir(os.path.join(TARGET, "minterpy")) and not force:
        return TARGET
    _make_writable(TARGET)
    shutil.rmtree(TARGET, ignore_errors=True)
    os.makedirs(TARGET, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, "reference")
        shutil.copytree(REFERENCE, src, ignore=shutil.ignore_patterns(".git", "docs"))
        subprocess.check_call([sys.executable, "-m", "pip", "install", "--quiet", "--no-index",
                               "--no-build-isolation", "--no-deps", "--find-links",
                               "/opt/wheelhouse", "--target", TARGET, src])
    shutil.copytree(os.path.join(REFERENCE, "tests"), os.path.join(TARGET, "reference_tests"),
                    ignore=shutil.ignore_patterns("__pycache__"))
    _make_writable(TARGET)          # the checkout is read-only; the copy need not be
    return TARGET


if __name__ == "__main__":
    print(make(force="--force" in sys.argv))
