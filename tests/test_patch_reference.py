"""Host-logic test of `minterpy_b200.install()` against the REAL reference classes (CPU).

The container that has /root/reference has no GPU, the GPU box has no reference, so the CUDA
library is replaced here by an oracle-backed stand-in plan: what is under test is the drop-in
wiring -- which names get rebound, shape / squeeze / error behaviour through the reference's
own call stack (`NewtonPolynomial.__call__`, `interpolate`, `LagrangeToNewton`,
`NewtonToLagrange`) -- compared with the unpatched reference on the same inputs.
"""
import numpy as np
import pytest

import oracle
from ref_shim import load_reference

mp = load_reference()
pytestmark = pytest.mark.skipif(mp is None, reason="reference not available")


class OraclePlan:
    """Same interface as minterpy_b200.plan.Plan, computed by the CPU oracle."""

    def __init__(self, exponents, generating_points):
        self.e = np.ascontiguousarray(exponents, dtype=np.int64)
        self.gp = np.ascontiguousarray(generating_points, dtype=np.float64)

    is_complete = True

    def eval(self, xx, coeffs):
        return oracle.eval_newton_polynomials(xx, coeffs, self.e, self.gp)

    def monomials(self, xx, triangular=False):
        return oracle.eval_newton_monomials(xx, self.e, self.gp, triangular)

    def dds(self, values):
        return oracle.dds(values, self.e, self.gp)


@pytest.fixture()
def patched(monkeypatch):
    import minterpy_b200
    from minterpy_b200 import api, polynomials
    fake = lambda e, gp, device=None: OraclePlan(e, gp)          # noqa: E731
    monkeypatch.setattr(api, "get_plan", fake)
    monkeypatch.setattr(polynomials, "get_plan", fake)
    minterpy_b200.install()
    yield minterpy_b200
    minterpy_b200.uninstall()


def _poly_inputs(m, n, lp, k, seed=0):
    mi = mp.MultiIndexSet.from_degree(m, n, lp)
    rng = np.random.default_rng(seed)
    coeffs = rng.uniform(-1, 1, (len(mi), k)) if k > 1 else rng.uniform(-1, 1, len(mi))
    pts = rng.uniform(-1, 1, (17, m))
    return mi, coeffs, pts


@pytest.mark.parametrize("m,n,lp,k", [(1, 4, 2.0, 1), (3, 4, 2.0, 1), (3, 4, 1.0, 10), (2, 3, np.inf, 2)])
def test_newton_polynomial_call(patched, m, n, lp, k):
    mi, coeffs, pts = _poly_inputs(m, n, lp, k)
    poly = mp.NewtonPolynomial(mi, coeffs)
    got = poly(pts)
    got_one = poly(pts[0]) if m > 1 else None
    patched.uninstall()
    want = poly(pts)
    assert got.shape == want.shape
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-13 * np.max(np.abs(want)))
    if m > 1:                    # single point given as a 1-D vector, and the never-0-D rule
        want_one = poly(pts[0])
        assert got_one.shape == want_one.shape and got_one.ndim == 1
        np.testing.assert_allclose(got_one, want_one, rtol=0, atol=1e-13 * np.max(np.abs(want)))


def test_names_are_rebound_and_restored(patched):
    import minterpy.dds
    import minterpy.interpolation
    import minterpy.polynomials.newton_polynomial as npoly
    import minterpy.utils
    from minterpy_b200 import api
    assert minterpy.utils.eval_newton_polynomials is api.eval_newton_polynomials
    assert npoly.eval_newton_polynomials is api.eval_newton_polynomials
    assert minterpy.interpolation.dds is api.dds and minterpy.dds.dds is api.dds
    patched.uninstall()
    assert minterpy.utils.eval_newton_polynomials is not api.eval_newton_polynomials
    assert minterpy.dds.dds is not api.dds


def test_interpolate_and_transformations(patched):
    fct = lambda x: np.cos(2 * x[:, 0]) * x[:, 1] + x[:, 2] ** 2      # noqa: E731
    interpolant = mp.interpolate(fct, spatial_dimension=3, poly_degree=4, lp_degree=2.0)
    pts = np.random.default_rng(1).uniform(-1, 1, (9, 3))
    got = interpolant(pts)
    poly = interpolant._Interpolant__interpolation_poly
    lag = mp.NewtonToLagrange(poly)()
    back = mp.LagrangeToNewton(lag)()
    patched.uninstall()
    want = mp.interpolate(fct, spatial_dimension=3, poly_degree=4, lp_degree=2.0)(pts)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-13)
    lag_ref = mp.NewtonToLagrange(poly)()
    np.testing.assert_allclose(lag.coeffs, lag_ref.coeffs, rtol=0, atol=1e-13)
    np.testing.assert_allclose(back.coeffs, poly.coeffs, rtol=0, atol=1e-11)
    assert type(lag).__name__ == "LagrangePolynomial" and type(back).__name__ == "NewtonPolynomial"


def test_errors_through_reference_stack(patched):
    mi, coeffs, pts = _poly_inputs(3, 2, 2.0, 1)
    poly = mp.NewtonPolynomial(mi, coeffs)
    with pytest.raises(ValueError):
        poly(pts[:, :2])
    with pytest.raises(TypeError):
        poly(pts.astype(np.float32))
