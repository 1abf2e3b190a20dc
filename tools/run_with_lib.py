"""Run a tool against an alternative build of the library: python tools/run_with_lib.py <lib.so> <script.py> [args]
(lab bench only: the product always loads minterpy_b200/libminterpy_b200.so)."""
import os
import runpy
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from minterpy_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.abspath(sys.argv[1])
sys.argv = sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
