"""CPU oracle for the Newton hot path -- TEST INFRASTRUCTURE ONLY.

A plain-C restatement (``newton_oracle.c``) of the reference's algorithm, loaded through
ctypes.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs
may import this package; ``minterpy_b200`` never does.  Parity of the oracle itself is
pinned against outputs of the unmodified reference: ``tests/test_oracle_golden.py``.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None

_i64 = ctypes.c_int64
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int64)


def build(force=False):
    """Compile ``newton_oracle.c`` with gcc (recipe: ``oracle/Makefile``)."""
    src = os.path.join(_HERE, "newton_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= os.path.getmtime(src)):
        return _LIB_PATH
    subprocess.check_call(["make", "-s", "-C", _HERE, "-B"])
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        lib.orc_eval_newton_monomials.restype = ctypes.c_int
        lib.orc_eval_newton_monomials.argtypes = [_dp, _i64, _ip, _i64, ctypes.c_int, _dp,
                                                  _i64, ctypes.c_int, _dp]
        lib.orc_eval_newton_polynomials.restype = ctypes.c_int
        lib.orc_eval_newton_polynomials.argtypes = [_dp, _i64, _dp, ctypes.c_int, _ip, _i64,
                                                    ctypes.c_int, _dp, _i64, _dp,
                                                    ctypes.c_int]
        lib.orc_dds.restype = ctypes.c_int
        lib.orc_dds.argtypes = [_dp, _i64, ctypes.c_int, _ip, ctypes.c_int, _dp, _i64]
        lib.orc_num_threads.restype = ctypes.c_int
        lib.orc_deriv_newt_eval.restype = ctypes.c_int
        lib.orc_deriv_newt_eval.argtypes = [_dp, _i64, _dp, ctypes.c_int, _ip, _i64,
                                            ctypes.c_int, _dp, _i64, _ip, _dp]
        _lib = lib
    return _lib


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _ptr(a, t):
    return a.ctypes.data_as(t)


def _check(rc, what):
    if rc != 0:
        msg = {1: "bad sizes", 2: "generating_points has too few rows", 3: "out of memory",
               4: "index set is not lexicographically sorted and complete"}.get(rc, "error")
        raise ValueError(f"oracle.{what}: {msg} (code {rc})")


def eval_newton_monomials(xx, exponents, generating_points, triangular=False):
    """(P, N) Newton-monomial matrix -- ``minterpy/utils.py:177-244``."""
    xx = _f64(xx)
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    gp = _f64(generating_points)
    N, m = e.shape
    xx = xx.reshape(-1, m)
    out = np.empty((xx.shape[0], N))
    rc = _load().orc_eval_newton_monomials(_ptr(xx, _dp), xx.shape[0], _ptr(e, _ip), N, m,
                                           _ptr(gp, _dp), gp.shape[0], int(triangular),
                                           _ptr(out, _dp))
    _check(rc, "eval_newton_monomials")
    return out


def eval_newton_polynomials(xx, coefficients, exponents, generating_points, nthreads=1):
    """(P, k) polynomial values -- ``minterpy/utils.py:247-343`` (always 2-D, no squeeze)."""
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    gp = _f64(generating_points)
    N, m = e.shape
    xx = _f64(xx).reshape(-1, m)
    c = _f64(coefficients).reshape(N, -1)
    out = np.empty((xx.shape[0], c.shape[1]))
    rc = _load().orc_eval_newton_polynomials(_ptr(xx, _dp), xx.shape[0], _ptr(c, _dp),
                                             c.shape[1], _ptr(e, _ip), N, m, _ptr(gp, _dp),
                                             gp.shape[0], _ptr(out, _dp), int(nthreads))
    _check(rc, "eval_newton_polynomials")
    return out


def dds(fct_values, exponents, generating_points):
    """(N, k) Newton coefficients -- ``minterpy/dds.py:794-826`` (works on a copy)."""
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    gp = _f64(generating_points)
    N, m = e.shape
    v = np.array(fct_values, dtype=np.float64, order="C", copy=True).reshape(N, -1)
    rc = _load().orc_dds(_ptr(v, _dp), N, v.shape[1], _ptr(e, _ip), m, _ptr(gp, _dp),
                         gp.shape[0])
    _check(rc, "dds")
    return v


def deriv_newt_eval(xx, coefficients, exponents, generating_points, derivative_order_along):
    """(P, k) values of the partial derivative -- ``minterpy/polynomials/utils.py:8-112``
    (the reference returns (P,) for its single coefficient vector)."""
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    gp = _f64(generating_points)
    N, m = e.shape
    xx = _f64(xx).reshape(-1, m)
    c = _f64(coefficients).reshape(N, -1)
    orders = np.ascontiguousarray(derivative_order_along, dtype=np.int64).reshape(m)
    out = np.empty((xx.shape[0], c.shape[1]))
    rc = _load().orc_deriv_newt_eval(_ptr(xx, _dp), xx.shape[0], _ptr(c, _dp), c.shape[1],
                                     _ptr(e, _ip), N, m, _ptr(gp, _dp), gp.shape[0],
                                     _ptr(orders, _ip), _ptr(out, _dp))
    _check(rc, "deriv_newt_eval")
    return out


def num_threads():
    return int(_load().orc_num_threads())
