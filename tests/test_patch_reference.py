"""Host-logic test of `minterpy_b200.install()` against the REAL reference classes (CPU).

The container that has /root/reference has no GPU, the GPU box has no reference, so the CUDA
library is replaced here by an oracle-backed stand-in plan: what is under test is the drop-in
wiring -- which names get rebound, shape / squeeze / error behaviour through the reference's
own call stack (`NewtonPolynomial.__call__`, `interpolate`, `LagrangeToNewton`,
`NewtonToLagrange`) -- compared with the unpatched reference on the same inputs.
"""
import numpy as np
import pytest

from oracle_plan import fake_get_plan
from ref_shim import load_reference

mp = load_reference()
pytestmark = pytest.mark.skipif(mp is None, reason="reference not available")


@pytest.fixture()
def patched(monkeypatch):
    import minterpy_b200
    from minterpy_b200 import api, polynomials
    monkeypatch.setattr(api, "get_plan", fake_get_plan)
    monkeypatch.setattr(polynomials, "get_plan", fake_get_plan)
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


def test_lagrange_arithmetic_through_reference_stack(patched):
    """polynomials/lagrange_polynomial.py:29-187 composes L2N + evaluation: with the path
    rebound the reference's own `+ - *` give the unpatched results."""
    rng = np.random.default_rng(3)
    mi_a = mp.MultiIndexSet.from_degree(2, 3, 1.0)
    mi_b = mp.MultiIndexSet.from_degree(2, 2, 2.0)
    a = mp.LagrangePolynomial(mi_a, rng.uniform(-1, 1, len(mi_a)))
    b = mp.LagrangePolynomial(mi_b, rng.uniform(-1, 1, len(mi_b)))
    got = [(a + b).coeffs, (a - b).coeffs, (a * b).coeffs]
    patched.uninstall()
    want = [(a + b).coeffs, (a - b).coeffs, (a * b).coeffs]
    for g, w in zip(got, want):
        assert g.shape == w.shape
        np.testing.assert_allclose(g, w, rtol=0, atol=1e-12 * max(1.0, np.max(np.abs(w))))


def test_diff_and_regression_through_reference_stack(patched):
    from minterpy.extras.regression import OrdinaryRegression
    rng = np.random.default_rng(4)
    mi = mp.MultiIndexSet.from_degree(3, 3, 2.0)
    poly = mp.NewtonPolynomial(mi, rng.uniform(-1, 1, len(mi)))
    xx = rng.uniform(-1, 1, (80, 3))
    yy = np.cos(xx @ np.array([1.0, 2.0, 0.5]))
    got_d = poly.diff(np.array([1, 0, 2])).coeffs
    got_p = poly.partial_diff(1).coeffs
    reg = OrdinaryRegression(mi)
    reg.fit(xx, yy)
    got_r, got_c = reg._regression_matrix, reg._coeffs
    regn = OrdinaryRegression(mi, origin_poly=mp.NewtonPolynomial)
    regn.fit(xx, yy)
    got_rn = regn._regression_matrix
    patched.uninstall()
    np.testing.assert_allclose(got_d, poly.diff(np.array([1, 0, 2])).coeffs, rtol=0, atol=1e-11)
    np.testing.assert_allclose(got_p, poly.partial_diff(1).coeffs, rtol=0, atol=1e-11)
    ref = OrdinaryRegression(mi)
    ref.fit(xx, yy)
    np.testing.assert_allclose(got_r, ref._regression_matrix, rtol=0, atol=1e-12)
    np.testing.assert_allclose(got_c, ref._coeffs, rtol=0, atol=1e-9)
    refn = OrdinaryRegression(mi, origin_poly=mp.NewtonPolynomial)
    refn.fit(xx, yy)
    assert np.array_equal(got_rn, refn._regression_matrix)


def test_completeness_check_is_rebound(patched):
    import minterpy.core.multi_index as mim
    from minterpy_b200.polynomials import is_lexicographically_complete
    assert mim.is_lexicographically_complete is is_lexicographically_complete
    e = mp.MultiIndexSet.from_degree(3, 3, 2.0).exponents
    holed = mp.MultiIndexSet(np.delete(e, 4, axis=0), lp_degree=2.0)
    assert mp.MultiIndexSet(e, lp_degree=2.0).is_complete and not holed.is_complete
    patched.uninstall()
    assert mim.is_lexicographically_complete is not is_lexicographically_complete
    assert not mp.MultiIndexSet(np.delete(e, 4, axis=0), lp_degree=2.0).is_complete
