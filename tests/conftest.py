import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_small():
    import golden_io
    return golden_io.load("small")


@pytest.fixture(scope="session")
def golden_cfg():
    import golden_io
    cache = {}

    def get(tag):
        if tag not in cache:
            cache[tag] = golden_io.load(tag)
        return cache[tag]
    return get
