"""GPU parity tests for the callers either side of the hot path:
derivative evaluation, regression matrices and fits, Lagrange arithmetic -- against vectors
written by the unmodified reference (tests/golden/next_rows.npz) and the CPU oracle.

Tolerances:
  * derivative values: <= 1e-13 * max|ref| (another summation order than the reference's
    itertools.combinations sum; exact zeros must be exact zeros);
  * coefficients that went through divided differences after a tolerance-level step
    (diff, partial_diff, regression in the Lagrange basis): <= 1e-10 * max|ref|
    (dds is bit-exact itself but amplifies input differences);
  * regression matrix: Newton basis bit-exact, Lagrange basis <= 1e-12 * max|ref|;
  * least-squares coefficients and error measures: 1e-8 relative (LAPACK on both sides).
"""
import numpy as np
import pytest

import golden_io
import oracle

pytestmark = pytest.mark.gpu

VAL_RTOL = 1e-13
COEF_RTOL = 1e-10


@pytest.fixture(scope="module")
def mpb():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import minterpy_b200
    return minterpy_b200


@pytest.fixture(scope="module")
def nxt():
    return golden_io.load("next_rows")


def rel_err(a, ref):
    scale = np.max(np.abs(ref))
    return np.max(np.abs(np.asarray(a) - ref)) / (scale if scale > 0 else 1.0)


# ---- derivatives ------------------------------------------------------------------------------
def test_derivative_values_golden(mpb, nxt):
    from minterpy_b200 import api
    for tag in nxt["der/tags"]:
        e, gp = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"]
        x, c = nxt[f"der/{tag}/x"], nxt[f"der/{tag}/coeffs"]
        for o, want in zip(nxt[f"der/{tag}/orders"], nxt[f"der/{tag}/values_ref"]):
            got = api.deriv_newt_eval(x, c, e, gp, o)
            assert got.shape == want.shape
            if not want.any():
                assert not got.any(), (tag, o)
            else:
                assert rel_err(got, want) <= VAL_RTOL, (tag, o)


def test_derivative_many_columns_and_torch(mpb, nxt):
    import torch
    from minterpy_b200 import api
    tag = "m3_n4_lp2.0"
    e, gp, x = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"], nxt[f"der/{tag}/x"]
    c = golden_io.draw_uniform(91, (e.shape[0], 19))
    o = np.array([1, 0, 2])
    want = oracle.deriv_newt_eval(x, c, e, gp, o)
    got = api.deriv_newt_eval(x, c, e, gp, o)
    assert got.shape == want.shape and rel_err(got, want) <= VAL_RTOL
    got_t = api.deriv_newt_eval(torch.as_tensor(x, device="cuda"), torch.as_tensor(c, device="cuda"),
                                e, gp, o)
    assert got_t.is_cuda and np.array_equal(got_t.cpu().numpy(), got)


@pytest.mark.parametrize("cfg,orders", [("cfg2", (1, 0, 2, 0)), ("cfg5", (0, 3, 1)),
                                        ("cfg3", (1, 1, 0, 0, 0, 2))])
def test_derivative_at_baseline_shapes_against_oracle(mpb, golden_cfg, cfg, orders):
    from minterpy_b200 import api
    g = golden_cfg(cfg)
    e, gp = g["exponents"], g["generating_points"]
    x = g["xx"][:48]
    c = golden_io.draw_uniform(17, (e.shape[0], 3))
    want = oracle.deriv_newt_eval(x, c, e, gp, np.array(orders))
    got = api.deriv_newt_eval(x, c, e, gp, np.array(orders))
    assert rel_err(got, want) <= VAL_RTOL


def test_derivative_consistency_properties(mpb, golden_cfg):
    """Size-independent checks at a BASELINE shape: order zero is the evaluation kernel's
    answer; a derivative of higher order than any exponent is identically zero; the
    derivative along x0 matches a central difference of the polynomial."""
    from minterpy_b200 import api
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    rng = np.random.default_rng(4)
    x = rng.uniform(-0.9, 0.9, (4096, 4))
    c = golden_io.draw_uniform(23, (e.shape[0],)) / (1.0 + e.sum(1)) ** 4
    plain = api.eval_newton_polynomials(x, c, e, gp)
    zero = api.deriv_newt_eval(x, c, e, gp, np.zeros(4, dtype=np.int64))
    assert np.array_equal(zero, plain)
    assert not api.deriv_newt_eval(x, c, e, gp, np.array([0, 17, 0, 0])).any()
    h = 1e-5
    xp, xm = x.copy(), x.copy()
    xp[:, 0] += h
    xm[:, 0] -= h
    fd = (api.eval_newton_polynomials(xp, c, e, gp) - api.eval_newton_polynomials(xm, c, e, gp)) / (2 * h)
    d1 = api.deriv_newt_eval(x, c, e, gp, np.array([1, 0, 0, 0]))
    assert rel_err(d1, fd) <= 1e-7


def test_newton_diff_and_partial_diff(mpb, nxt):
    for tag in nxt["der/tags"]:
        e, gp = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"]
        mi = mpb.MultiIndexSet(e)
        poly = mpb.NewtonPolynomial(mi, nxt[f"der/{tag}/coeffs"], grid=mpb.Grid(mi, gp))
        m = e.shape[1]
        got = poly.diff(np.ones(m, dtype=np.int64))
        assert isinstance(got, mpb.NewtonPolynomial) and got.coeffs.shape == (e.shape[0],)
        assert rel_err(got.coeffs, nxt[f"der/{tag}/diff_ones_ref"]) <= COEF_RTOL
        got = poly.partial_diff(m - 1, 2)
        ref = nxt[f"der/{tag}/partial_last2_ref"]
        assert np.max(np.abs(got.coeffs - ref)) <= COEF_RTOL * max(np.max(np.abs(ref)), 1.0)
        same = poly.diff(np.zeros(m, dtype=np.int64))
        assert rel_err(same.coeffs, poly.coeffs) <= COEF_RTOL
    with pytest.raises(ValueError):
        poly.diff(np.ones(m + 1, dtype=np.int64))
    with pytest.raises(ValueError):
        poly.partial_diff(m, 1)


def test_derivative_argument_errors(mpb, nxt):
    from minterpy_b200 import api
    tag = "m2_n5_lp1.0"
    e, gp = nxt[f"der/{tag}/exponents"], nxt[f"der/{tag}/generating_points"]
    x, c = nxt[f"der/{tag}/x"], nxt[f"der/{tag}/coeffs"]
    with pytest.raises(ValueError):
        api.deriv_newt_eval(x, c, e, gp, np.array([1]))
    with pytest.raises(ValueError):
        api.deriv_newt_eval(x, c, e, gp, np.array([1, -1]))
    with pytest.raises(TypeError):
        api.deriv_newt_eval(x, c, e, gp, np.array([1.0, 0.0]))
    with pytest.raises(ValueError):
        api.deriv_newt_eval(x, c[:-1], e, gp, np.array([1, 0]))
    # more than 8 dimensions (no shared-memory-table kernel): few polynomials run on the tree
    # kernel with differentiated tables -- x0 * x8 differentiated along both is 1, along x1 is 0
    e9 = np.zeros((4, 9), dtype=np.int64)
    e9[1, 0] = e9[2, 8] = e9[3, 0] = e9[3, 8] = 1
    gp9 = np.tile(np.array([[1.0], [-1.0]]), (1, 9))
    c9 = np.array([0.0, 0.0, 0.0, 1.0])                     # (x0 - 1)(x8 - 1)
    x9 = np.random.default_rng(3).uniform(-1, 1, (5, 9))
    o = np.zeros(9, dtype=np.int64)
    o[0] = o[8] = 1
    assert np.allclose(api.deriv_newt_eval(x9, c9, e9, gp9, o), 1.0, rtol=0, atol=1e-15)
    o[8] = 0
    assert np.allclose(api.deriv_newt_eval(x9, c9, e9, gp9, o), x9[:, 8] - 1.0, rtol=0, atol=1e-15)
    o[1] = 1
    assert np.array_equal(api.deriv_newt_eval(x9, c9, e9, gp9, o), np.zeros(5))


# ---- regression ---------------------------------------------------------------------------------
def regression_case(mpb, nxt, tag, basis):
    from minterpy_b200.regression import OrdinaryRegression
    e, gp = nxt[f"reg/{tag}/exponents"], nxt[f"reg/{tag}/generating_points"]
    mi = mpb.MultiIndexSet(e, lp_degree=float(nxt[f"reg/{tag}/lp"]))
    cls = mpb.LagrangePolynomial if basis == "lag" else mpb.NewtonPolynomial
    return OrdinaryRegression(mi, mpb.Grid(mi, gp), origin_poly=cls)


@pytest.mark.parametrize("basis", ["lag", "newt"])
def test_regression_matrix_and_fit_golden(mpb, nxt, basis):
    for tag in nxt["reg/tags"]:
        reg = regression_case(mpb, nxt, tag, basis)
        xx, yy, xt = nxt[f"reg/{tag}/xx"], nxt[f"reg/{tag}/yy"], nxt[f"reg/{tag}/xt"]
        ref_matrix = nxt[f"reg/{tag}/{basis}/matrix_ref"]
        rr = reg.get_regression_matrix(xx)
        assert rr.shape == ref_matrix.shape
        if basis == "newt":
            assert np.array_equal(rr, ref_matrix)
        else:
            assert rel_err(rr, ref_matrix) <= 1e-12
        reg.fit(xx, yy)
        assert rel_err(reg.coeffs, nxt[f"reg/{tag}/{basis}/coeffs_ref"]) <= 1e-8
        errs = np.array([*reg.regfit_l1_error, *reg.regfit_l2_error, *reg.loocv_error])
        assert np.allclose(errs, nxt[f"reg/{tag}/{basis}/errors_ref"], rtol=1e-6, atol=0)
        assert rel_err(reg.predict(xt), nxt[f"reg/{tag}/{basis}/predict_ref"]) <= 1e-9
        reg.fit(xx, yy, weights=nxt[f"reg/{tag}/weights"], lstsq_solver="inv")
        assert rel_err(reg.coeffs, nxt[f"reg/{tag}/{basis}/weighted_inv_coeffs_ref"]) <= 1e-7
        assert np.allclose(reg.loocv_error, nxt[f"reg/{tag}/{basis}/weighted_loocv_ref"], rtol=1e-5)


def test_regression_on_device_tensors(mpb, nxt):
    """torch in, torch out: matrix, solve and prediction without leaving the device."""
    import torch
    tag = "m2_n5_lpinf"
    reg = regression_case(mpb, nxt, tag, "lag")
    xx = torch.as_tensor(nxt[f"reg/{tag}/xx"], device="cuda")
    yy = torch.as_tensor(nxt[f"reg/{tag}/yy"], device="cuda")
    reg.fit(xx, yy)
    assert reg.regression_matrix.is_cuda and reg.coeffs.is_cuda
    assert rel_err(reg.coeffs.cpu().numpy(), nxt[f"reg/{tag}/lag/coeffs_ref"]) <= 1e-8
    pred = reg.predict(torch.as_tensor(nxt[f"reg/{tag}/xt"], device="cuda"))
    assert pred.is_cuda
    assert rel_err(pred.cpu().numpy(), nxt[f"reg/{tag}/lag/predict_ref"]) <= 1e-9
    assert np.isclose(reg.regfit_l2_error[0], nxt[f"reg/{tag}/lag/errors_ref"][2], rtol=1e-6)


def test_lagrange_regression_matrix_properties_at_baseline_shape(mpb, golden_cfg):
    """k = N coefficient columns through the tensor-core kernel (cfg3, N = 8008): the Lagrange
    basis is a partition of unity (rows sum to 1) and interpolatory (identity on the nodes)."""
    import torch
    from minterpy_b200.regression import compute_regression_matrix
    g = golden_cfg("cfg3")
    e, gp = g["exponents"], g["generating_points"]
    mi = mpb.MultiIndexSet(e, lp_degree=1.0)
    grid = mpb.Grid(mi, gp)
    basis = mpb.LagrangePolynomial(mi, None, grid=grid)
    xx = torch.as_tensor(np.random.default_rng(8).uniform(-1, 1, (1000, 6)), device="cuda")
    rr = compute_regression_matrix(basis, xx)
    assert rr.shape == (1000, 8008) and rr.is_cuda
    assert float((rr.sum(1) - 1.0).abs().max()) <= 1e-11
    nodes = torch.as_tensor(grid.unisolvent_nodes[::11], device="cuda")
    at_nodes = compute_regression_matrix(basis, nodes)
    eye = torch.zeros_like(at_nodes)
    eye[torch.arange(nodes.shape[0]), torch.arange(0, 8008, 11)] = 1.0
    assert float((at_nodes - eye).abs().max()) <= 1e-11


# ---- Lagrange arithmetic ------------------------------------------------------------------------
def test_lagrange_arithmetic_golden(mpb, nxt):
    for i in range(int(nxt["ari/count"])):
        m, n1, lp1, n2, lp2 = nxt[f"ari/{i}/spec"]
        mi_a = mpb.MultiIndexSet.from_degree(int(m), int(n1), float(lp1))
        mi_b = mpb.MultiIndexSet.from_degree(int(m), int(n2), float(lp2))
        a = mpb.LagrangePolynomial(mi_a, nxt[f"ari/{i}/coeffs_a"], grid=mpb.Grid(mi_a))
        b = mpb.LagrangePolynomial(mi_b, nxt[f"ari/{i}/coeffs_b"], grid=mpb.Grid(mi_b))
        for name, res in (("add", a + b), ("sub", a - b), ("mul", a * b)):
            assert isinstance(res, mpb.LagrangePolynomial)
            assert np.array_equal(res.multi_index.exponents, nxt[f"ari/{i}/{name}/exponents"])
            assert res.multi_index.lp_degree == float(nxt[f"ari/{i}/{name}/lp"])
            assert np.array_equal(res.grid.generating_points,
                                  nxt[f"ari/{i}/{name}/generating_points"])
            ref = nxt[f"ari/{i}/{name}/coeffs_ref"]
            assert np.max(np.abs(res.coeffs - ref)) <= COEF_RTOL * max(np.max(np.abs(ref)), 1.0), (i, name)


def test_lagrange_arithmetic_on_device_and_identities(mpb):
    import torch
    mi_a = mpb.MultiIndexSet.from_degree(3, 5, 2.0)
    mi_b = mpb.MultiIndexSet.from_degree(3, 4, 1.0)
    ga, gb = mpb.Grid(mi_a), mpb.Grid(mi_b)
    fa = np.cos(ga.unisolvent_nodes @ np.array([1.0, 0.5, 2.0]))
    fb = np.sin(gb.unisolvent_nodes @ np.array([0.3, 1.5, 1.0]))
    a = mpb.LagrangePolynomial(mi_a, torch.as_tensor(fa, device="cuda"), grid=ga)
    b = mpb.LagrangePolynomial(mi_b, torch.as_tensor(fb, device="cuda"), grid=gb)
    s, p = a + b, a * b
    assert s.coeffs.is_cuda and p.coeffs.is_cuda
    a_h = mpb.LagrangePolynomial(mi_a, fa, grid=ga)
    b_h = mpb.LagrangePolynomial(mi_b, fb, grid=gb)
    assert np.array_equal((a_h + b_h).coeffs, s.coeffs.cpu().numpy())
    assert np.array_equal((a_h * b_h).coeffs, p.coeffs.cpu().numpy())
    # (a + b) - b has a's values on the result grid; a * 1 = a
    back = (a_h + b_h) - b_h
    newt_a = mpb.LagrangeToNewton(a_h)()
    assert rel_err(back.coeffs, newt_a(back.grid.unisolvent_nodes)) <= 1e-12
    one = mpb.LagrangePolynomial(mi_b, np.ones(len(mi_b)), grid=gb)
    prod = a_h * one
    assert rel_err(prod.coeffs, newt_a(prod.grid.unisolvent_nodes)) <= 1e-12
    with pytest.raises(NotImplementedError):
        a_h + mpb.LagrangePolynomial(mpb.MultiIndexSet.from_degree(2, 2, 2.0), np.ones(6))
    with pytest.raises(TypeError):
        a_h + 1.0


def test_interpolate_with_the_reference_signature(mpb, golden_cfg):
    g = golden_cfg("cfg1")
    fct = lambda x: x * np.sin(10 * x)  # noqa: E731
    poly = mpb.interpolate(fct, 1, 64)
    assert np.array_equal(poly.coeffs, g["quickstart_coeffs"])
    yq = poly(np.linspace(-1, 1, 1_000_000))
    assert rel_err(yq[::1000], g["quickstart_every1000"]) <= 1e-13


# ---- incomplete index sets: dense operator ------------------------------------------------------
def test_lagrange_to_newton_on_sets_with_holes(mpb, nxt):
    """transformations/utils.py:155-179: without completeness there are no divided differences;
    the operator is the dense inverse of the Newton->Lagrange matrix, applied by mpx_apply_dense.
    Inversion is LAPACK on the reference's side and cuSOLVER here: 1e-10 * max|ref|."""
    from minterpy_b200.polynomials import MatrixOperator
    assert int(nxt["inc/count"]) >= 5
    for i in range(int(nxt["inc/count"])):
        e, gp = nxt[f"inc/{i}/exponents"], nxt[f"inc/{i}/generating_points"]
        mi = mpb.MultiIndexSet(e, lp_degree=float(nxt[f"inc/{i}/lp"]))
        assert not mi.is_complete
        grid = mpb.Grid(mi)
        assert np.array_equal(grid.generating_points, gp)
        with pytest.raises(ValueError):
            grid.tree                                    # dds needs a complete set
        for key, ref_key in (("coeffs1", "l2n1_ref"), ("coeffs3", "l2n3_ref")):
            lag = mpb.LagrangePolynomial(mi, nxt[f"inc/{i}/{key}"], grid=grid)
            trafo = mpb.LagrangeToNewton(lag)
            op = trafo.transformation_operator
            assert isinstance(op, MatrixOperator)
            assert rel_err(op.array_repr_full, nxt[f"inc/{i}/operator_ref"]) <= 1e-10
            newt = trafo()
            ref = nxt[f"inc/{i}/{ref_key}"]
            assert newt.coeffs.shape == ref.shape and rel_err(newt.coeffs, ref) <= 1e-10
            back = mpb.NewtonToLagrange(newt)()
            assert rel_err(back.coeffs, lag.coeffs) <= 1e-11
    # operator algebra: (L2N) @ (N2L) is the identity operator
    n2l = mpb.NewtonToLagrange(newt).transformation_operator
    prod = op @ n2l
    assert isinstance(prod, MatrixOperator)
    # (entries of the Lagrange->Newton matrix reach 1e3..1e4 here: 1e-10 is ~1e-14 relative)
    assert np.max(np.abs(prod.array_repr_full - np.eye(len(mi)))) <= 1e-10
    with pytest.raises(ValueError):
        op @ np.ones(len(mi) + 1)


def test_matrix_operator_on_device_tensors(mpb, nxt):
    import torch
    from minterpy_b200.polynomials import MatrixOperator
    rng = np.random.default_rng(12)
    T = rng.uniform(-1, 1, (301, 301))
    c = rng.uniform(-1, 1, (301, 5))
    op = MatrixOperator(torch.as_tensor(T, device="cuda"))
    out = op @ torch.as_tensor(c, device="cuda")
    assert out.is_cuda and rel_err(out.cpu().numpy(), T @ c) <= 1e-13
    assert rel_err(op @ c[:, 0], T @ c[:, 0]) <= 1e-13              # NumPy in, NumPy out, 1-D
    assert np.array_equal(op.array_repr_full, T)
