"""CPU tests of the boundary: the shared library loads, exports every symbol the header
declares, and the host-side argument checks answer without a GPU (no compute calls)."""
import ctypes
import os
import re

import numpy as np
import pytest

from minterpy_b200 import _lib, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "minterpy_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mpx_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    lib = _lib.lib()
    names = declared_symbols()
    assert "mpx_eval" in names and "mpx_dds" in names and len(names) >= 11
    for name in names:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)
    assert lib.mpx_version() >= 100


def test_argument_errors_without_gpu():
    lib = _lib.lib()
    h = ctypes.c_void_p()
    e = np.zeros((1, 2), dtype=np.int64)
    gp = np.zeros((1, 2))
    assert lib.mpx_plan_create(0, e.ctypes.data, 0, 2, gp.ctypes.data, 1, ctypes.byref(h)) == _lib.MPX_ERR_ARG
    assert "at least 1 monomial" in _lib.last_error()
    big = np.full((1, 2), 300, dtype=np.int64)
    assert lib.mpx_plan_create(0, big.ctypes.data, 1, 2, gp.ctypes.data, 1, ctypes.byref(h)) == _lib.MPX_ERR_UNSUPPORTED
    assert lib.mpx_eval(None, None, 1, 1, None, 1, 1, None, 1, None) == _lib.MPX_ERR_ARG
    assert lib.mpx_dds(None, None, 1, 1, None) == _lib.MPX_ERR_ARG
    with pytest.raises(ValueError):
        _lib.check(_lib.MPX_ERR_ARG)
    with pytest.raises(NotImplementedError):
        _lib.check(_lib.MPX_ERR_UNSUPPORTED)


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    e = np.zeros((1, 1), dtype=np.int64)
    with pytest.raises(_lib.MinterpyB200Error):
        api.eval_newton_polynomials(np.zeros(3), np.ones(1), e, np.zeros((1, 1)))


def test_input_rules_match_reference():
    """rectify_* / convert_eval_output against the reference's own functions
    (src/minterpy/core/verification.py:48-163), when the reference is importable."""
    from ref_shim import load_reference
    if load_reference() is None:
        pytest.skip("reference not available")
    from minterpy.core import verification as ref
    rng = np.random.default_rng(0)
    e = rng.integers(0, 3, (6, 3))
    cases = [(rng.uniform(size=(5, 3)), rng.uniform(size=6)),
             (rng.uniform(size=(1, 3)), rng.uniform(size=(6, 1))),
             (rng.uniform(size=3), rng.uniform(size=(6, 4))),
             (rng.uniform(size=(5, 2)), rng.uniform(size=6)),
             (rng.uniform(size=4), rng.uniform(size=6)),
             (rng.uniform(size=(5, 3)), rng.uniform(size=(5, 2))),
             (rng.uniform(size=(5, 3)).astype(np.float32), rng.uniform(size=6)),
             (rng.uniform(size=(5, 3)), rng.integers(0, 2, 6))]
    for x, c in cases:
        try:
            want = ref.rectify_eval_input(x, c, e, True)
            err = None
        except Exception as ex:      # noqa: BLE001
            err = type(ex)
        if err is None:
            got = api.rectify_eval_input(x, c, e, True)
            assert got[0] == want[0] and got[2:5] == want[2:5]
            assert np.array_equal(got[1], want[1]) and np.array_equal(got[5], want[5])
        else:
            with pytest.raises(err):
                api.rectify_eval_input(x, c, e, True)
    for shape in ((5, 1), (1, 4), (1, 1), (5, 4)):
        a = rng.uniform(size=shape)
        assert api.convert_eval_output(a).shape == ref.convert_eval_output(a).shape
    with pytest.raises(ValueError):
        api.rectify_eval_input(np.zeros((2, 3)), np.zeros(0), e[:0], True)
