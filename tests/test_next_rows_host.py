"""CPU tests of the host logic of the section-8(f) rows in the package's own classes
(differentiation, regression, Lagrange arithmetic, interpolate) with the device plan replaced
by the oracle-backed stand-in, against vectors written by the unmodified reference
(tests/golden/next_rows.npz).  The same vectors are checked through the CUDA path in
tests/test_gpu_next_rows.py."""
import numpy as np
import pytest

import golden_io
from oracle_plan import fake_get_plan


@pytest.fixture()
def mpb(monkeypatch):
    import minterpy_b200
    from minterpy_b200 import api, polynomials
    monkeypatch.setattr(api, "get_plan", fake_get_plan)
    monkeypatch.setattr(polynomials, "get_plan", fake_get_plan)
    return minterpy_b200


@pytest.fixture(scope="module")
def nxt():
    return golden_io.load("next_rows")


def close(a, ref, rtol):
    scale = max(np.max(np.abs(ref)), 1e-300)
    return np.max(np.abs(np.asarray(a) - ref)) <= rtol * scale


def test_diff_and_partial_diff(mpb, nxt):
    for tag in nxt["der/tags"]:
        e, gp = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"]
        mi = mpb.MultiIndexSet(e)
        poly = mpb.NewtonPolynomial(mi, nxt[f"der/{tag}/coeffs"], grid=mpb.Grid(mi, gp))
        m = e.shape[1]
        assert close(poly.diff(np.ones(m, dtype=np.int64)).coeffs, nxt[f"der/{tag}/diff_ones_ref"], 1e-10)
        ref = nxt[f"der/{tag}/partial_last2_ref"]
        got = poly.partial_diff(m - 1, 2).coeffs
        assert np.max(np.abs(got - ref)) <= 1e-10 * max(np.max(np.abs(ref)), 1.0)
        two = mpb.NewtonPolynomial(mi, np.stack([poly.coeffs, 2 * poly.coeffs], axis=1), grid=poly.grid)
        d2 = two.diff(np.ones(m, dtype=np.int64)).coeffs
        assert d2.shape == (e.shape[0], 2) and close(d2[:, 1], 2 * d2[:, 0], 1e-12)


def test_deriv_argument_rules(mpb, nxt):
    from minterpy_b200 import api
    tag = "m2_n5_lp1.0"
    e, gp = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"]
    x, c = nxt[f"der/{tag}/x"], nxt[f"der/{tag}/coeffs"]
    assert api.deriv_newt_eval(x, c, e, gp, [1, 0]).shape == (x.shape[0],)
    assert api.deriv_newt_eval(x, c[:, None], e, gp, [1, 0]).shape == (x.shape[0], 1)
    for bad, exc in (([1], ValueError), ([1, -1], ValueError), ([1.0, 0.0], TypeError)):
        with pytest.raises(exc):
            api.deriv_newt_eval(x, c, e, gp, bad)
    with pytest.raises(ValueError):
        api.deriv_newt_eval(x[:, :1], c, e, gp, [1, 0])
    with pytest.raises(ValueError):
        api.deriv_newt_eval(x, c[:-1], e, gp, [1, 0])


@pytest.mark.parametrize("basis", ["lag", "newt"])
def test_regression(mpb, nxt, basis):
    from minterpy_b200.regression import OrdinaryRegression
    for tag in nxt["reg/tags"]:
        e, gp = nxt[f"reg/{tag}/exponents"], nxt[f"reg/{tag}/generating_points"]
        mi = mpb.MultiIndexSet(e, lp_degree=float(nxt[f"reg/{tag}/lp"]))
        cls = mpb.LagrangePolynomial if basis == "lag" else mpb.NewtonPolynomial
        reg = OrdinaryRegression(mi, mpb.Grid(mi, gp), origin_poly=cls)
        xx, yy = nxt[f"reg/{tag}/xx"], nxt[f"reg/{tag}/yy"]
        with pytest.raises(TypeError):
            reg.predict(xx)
        reg.fit(xx, yy)
        assert close(reg.regression_matrix, nxt[f"reg/{tag}/{basis}/matrix_ref"], 1e-12)
        assert close(reg.coeffs, nxt[f"reg/{tag}/{basis}/coeffs_ref"], 1e-8)
        errs = np.array([*reg.regfit_l1_error, *reg.regfit_l2_error, *reg.loocv_error])
        assert np.allclose(errs, nxt[f"reg/{tag}/{basis}/errors_ref"], rtol=1e-6, atol=0)
        assert close(reg(nxt[f"reg/{tag}/xt"]), nxt[f"reg/{tag}/{basis}/predict_ref"], 1e-9)
        for solver in ("pinv", "qr", "inv"):
            reg.fit(xx, yy, lstsq_solver=solver)
            assert close(reg.coeffs, nxt[f"reg/{tag}/{basis}/coeffs_ref"], 1e-6), solver
        reg.fit(xx, yy, weights=nxt[f"reg/{tag}/weights"], lstsq_solver="inv")
        assert close(reg.coeffs, nxt[f"reg/{tag}/{basis}/weighted_inv_coeffs_ref"], 1e-7)
        assert np.allclose(reg.loocv_error, nxt[f"reg/{tag}/{basis}/weighted_loocv_ref"], rtol=1e-5)
    with pytest.raises(ValueError):
        OrdinaryRegression()
    with pytest.raises(NotImplementedError):
        reg.fit(xx, yy, lstsq_solver="nope")


def test_regression_edge_shapes(mpb):
    from minterpy_b200.regression import compute_regression_matrix
    mi = mpb.MultiIndexSet.from_degree(2, 2, 2.0)
    lag = mpb.LagrangePolynomial(mi, None)
    assert compute_regression_matrix(lag, np.zeros((1, 2))).shape == (1, len(mi))
    one = mpb.MultiIndexSet.from_degree(2, 0, 2.0)
    assert compute_regression_matrix(mpb.LagrangePolynomial(one, None), np.zeros((5, 2))).shape == (5, 1)
    with pytest.raises(TypeError):
        compute_regression_matrix(mpb.polynomials._Polynomial(mi, None), np.zeros((1, 2)))


def test_lagrange_arithmetic(mpb, nxt):
    for i in range(int(nxt["ari/count"])):
        m, n1, lp1, n2, lp2 = nxt[f"ari/{i}/spec"]
        mi_a = mpb.MultiIndexSet.from_degree(int(m), int(n1), float(lp1))
        mi_b = mpb.MultiIndexSet.from_degree(int(m), int(n2), float(lp2))
        a = mpb.LagrangePolynomial(mi_a, nxt[f"ari/{i}/coeffs_a"])
        b = mpb.LagrangePolynomial(mi_b, nxt[f"ari/{i}/coeffs_b"])
        for name, res in (("add", a + b), ("sub", a - b), ("mul", a * b)):
            assert np.array_equal(res.multi_index.exponents, nxt[f"ari/{i}/{name}/exponents"])
            assert res.multi_index.lp_degree == float(nxt[f"ari/{i}/{name}/lp"])
            assert np.array_equal(res.grid.generating_points, nxt[f"ari/{i}/{name}/generating_points"])
            ref = nxt[f"ari/{i}/{name}/coeffs_ref"]
            assert np.max(np.abs(res.coeffs - ref)) <= 1e-10 * max(np.max(np.abs(ref)), 1.0)
    with pytest.raises(NotImplementedError):
        a + mpb.LagrangePolynomial(mpb.MultiIndexSet.from_degree(int(m) + 1, 1, 1.0), np.ones(int(m) + 2))
    with pytest.raises(TypeError):
        a * 2.0


def test_interpolate_both_signatures(mpb, golden_cfg):
    g = golden_cfg("cfg1")
    fct = lambda x: x * np.sin(10 * x)  # noqa: E731
    poly = mpb.interpolate(fct, 1, 64)
    assert np.array_equal(poly.coeffs, g["quickstart_coeffs"])
    grid = mpb.Grid(mpb.MultiIndexSet.from_degree(1, 64, 2.0))
    assert np.array_equal(mpb.interpolate(fct, grid).coeffs, poly.coeffs)
    with pytest.raises(TypeError):
        mpb.interpolate(fct, 1)
