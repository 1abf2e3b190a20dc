"""Golden vectors for the rows of SURVEY.md section 8(f) ("next"), from the UNMODIFIED reference.

Run here (the container that has /root/reference):  python tests/golden/make_golden_next.py
Writes tests/golden/next_rows.npz:

* ``gen/*``     index sets, polynomial degrees and default generating points
                (``MultiIndexSet.from_degree``, ``Grid``) for lp-degrees the other files lack
* ``cmp/*``     ``is_lexicographically_complete`` on complete sets and sets with holes
* ``der/*``     ``deriv_newt_eval`` and ``NewtonPolynomial.diff`` / ``partial_diff``
* ``reg/*``     ``OrdinaryRegression`` in the Lagrange and Newton bases: regression matrix,
                coefficients, error measures, predictions
* ``ari/*``     ``LagrangePolynomial`` ``+ - *``
* ``inc/*``     ``LagrangeToNewton`` on index sets WITH holes: the reference falls back to a dense
                ``MatrixOperator`` holding ``inv(N2L)`` (``transformations/utils.py:155-179``)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import load_reference  # noqa: E402

mp = load_reference()
assert mp is not None, "needs /root/reference"
from minterpy.core.utils import is_lexicographically_complete  # noqa: E402
from minterpy.extras.regression import OrdinaryRegression  # noqa: E402
from minterpy.polynomials.utils import deriv_newt_eval  # noqa: E402

out = {}


def tag_of(m, n, lp):
    return f"m{m}_n{n}_lp{lp}"


# ---- generators ---------------------------------------------------------------------------
gen_tags = []
for (m, n, lp) in [(1, 7, 2.0), (2, 6, 0.5), (2, 9, 1.5), (3, 5, 3.0), (3, 6, 1.0), (4, 4, 2.0),
                   (4, 3, np.inf), (5, 3, 1.0), (2, 0, 2.0), (3, 1, np.inf), (6, 4, 2.0)]:
    mi = mp.MultiIndexSet.from_degree(m, n, lp)
    grid = mp.Grid(mi)
    t = tag_of(m, n, lp)
    gen_tags.append(t)
    out[f"gen/{t}/args"] = np.array([m, n, lp], dtype=np.float64)
    out[f"gen/{t}/exponents"] = mi.exponents.astype(np.uint8)
    out[f"gen/{t}/poly_degree"] = np.int64(mi.poly_degree)
    out[f"gen/{t}/generating_points"] = grid.generating_points
out["gen/tags"] = np.array(gen_tags)
for n in (0, 1, 2, 3, 10, 33, 64):
    out[f"gen/leja/{n}"] = mp.utils.gen_chebychev_2nd_order_leja_ordered(n)

# ---- completeness ---------------------------------------------------------------------------
rng = np.random.default_rng(811)
cmp_n = 0
for (m, n, lp) in [(1, 4, 2.0), (2, 3, 1.0), (3, 3, 2.0), (3, 2, np.inf), (4, 2, 2.0)]:
    e = mp.MultiIndexSet.from_degree(m, n, lp).exponents
    cases = [e]
    for _ in range(5):
        drop = rng.choice(len(e), size=int(rng.integers(1, 3)), replace=False)
        cases.append(np.delete(e, drop, axis=0))
    for c in cases:
        out[f"cmp/{cmp_n}/exponents"] = c.astype(np.uint8)
        out[f"cmp/{cmp_n}/complete"] = np.bool_(is_lexicographically_complete(c))
        cmp_n += 1
out["cmp/count"] = np.int64(cmp_n)

# ---- derivatives ------------------------------------------------------------------------------
rng = np.random.default_rng(812)
der_tags = []
for (m, n, lp) in [(1, 6, 2.0), (2, 5, 1.0), (3, 4, 2.0), (3, 3, np.inf), (4, 5, 2.0), (2, 9, 2.0)]:
    mi = mp.MultiIndexSet.from_degree(m, n, lp)
    grid = mp.Grid(mi)
    e, gp = mi.exponents, grid.generating_points
    t = tag_of(m, n, lp)
    der_tags.append(t)
    N = len(mi)
    x = np.concatenate([rng.uniform(-1, 1, (14, m)), grid.unisolvent_nodes[:: max(1, N // 6)]])
    c = rng.uniform(-1, 1, N)
    orders = [np.zeros(m, int), np.ones(m, int), np.arange(m) % 3, np.full(m, 2),
              np.eye(m, dtype=int)[m - 1] * 3, np.full(m, n), np.full(m, n + 1),
              np.eye(m, dtype=int)[0] * (n + 4)]
    out[f"der/{t}/exponents"] = e.astype(np.uint8)
    out[f"der/{t}/generating_points"] = gp
    out[f"der/{t}/x"] = x
    out[f"der/{t}/coeffs"] = c
    out[f"der/{t}/orders"] = np.array(orders, dtype=np.int64)
    out[f"der/{t}/values_ref"] = np.array([deriv_newt_eval(x, c, e, gp, o) for o in orders])
    poly = mp.NewtonPolynomial(mi, c, grid=grid)
    out[f"der/{t}/diff_ones_ref"] = poly.diff(np.ones(m, dtype=np.int64)).coeffs
    out[f"der/{t}/partial_last2_ref"] = poly.partial_diff(m - 1, 2).coeffs
out["der/tags"] = np.array(der_tags)

# ---- regression -------------------------------------------------------------------------------
rng = np.random.default_rng(813)
reg_tags = []
for (m, n, lp, P) in [(1, 6, 2.0, 40), (2, 4, 2.0, 70), (3, 3, 1.0, 90), (2, 5, np.inf, 120)]:
    mi = mp.MultiIndexSet.from_degree(m, n, lp)
    grid = mp.Grid(mi)
    t = tag_of(m, n, lp)
    reg_tags.append(t)
    xx = rng.uniform(-1, 1, (P, m))
    w = rng.uniform(0.5, 2.5, m)
    yy = np.cos(xx @ w) + 0.01 * rng.standard_normal(P)
    xt = rng.uniform(-1, 1, (25, m))
    wts = rng.uniform(0.5, 1.5, P)
    out[f"reg/{t}/exponents"] = mi.exponents.astype(np.uint8)
    out[f"reg/{t}/generating_points"] = grid.generating_points
    out[f"reg/{t}/lp"] = np.float64(lp)
    out[f"reg/{t}/xx"] = xx
    out[f"reg/{t}/yy"] = yy
    out[f"reg/{t}/xt"] = xt
    out[f"reg/{t}/weights"] = wts
    for basis, cls in (("lag", mp.LagrangePolynomial), ("newt", mp.NewtonPolynomial)):
        reg = OrdinaryRegression(mi, grid, origin_poly=cls)
        reg.fit(xx, yy)
        out[f"reg/{t}/{basis}/matrix_ref"] = reg._regression_matrix
        out[f"reg/{t}/{basis}/coeffs_ref"] = reg._coeffs
        out[f"reg/{t}/{basis}/errors_ref"] = np.array(
            [*reg.regfit_l1_error, *reg.regfit_l2_error, *reg.loocv_error])
        out[f"reg/{t}/{basis}/predict_ref"] = reg.predict(xt)
        reg.fit(xx, yy, weights=wts, lstsq_solver="inv")
        out[f"reg/{t}/{basis}/weighted_inv_coeffs_ref"] = reg._coeffs
        out[f"reg/{t}/{basis}/weighted_loocv_ref"] = np.array(reg.loocv_error)
out["reg/tags"] = np.array(reg_tags)

# ---- Lagrange arithmetic ----------------------------------------------------------------------
rng = np.random.default_rng(814)
ari_n = 0
for (m, (n1, lp1), (n2, lp2)) in [(1, (3, 2.0), (5, 2.0)), (2, (3, 1.0), (2, 2.0)),
                                  (2, (4, 2.0), (4, 2.0)), (3, (2, 2.0), (3, 1.0)),
                                  (3, (2, np.inf), (1, 1.0))]:
    polys = []
    for (n, lp) in ((n1, lp1), (n2, lp2)):
        mi = mp.MultiIndexSet.from_degree(m, n, lp)
        grid = mp.Grid(mi)
        polys.append(mp.LagrangePolynomial(mi, rng.uniform(-1, 1, len(mi)), grid=grid))
    a, b = polys
    out[f"ari/{ari_n}/spec"] = np.array([m, n1, lp1, n2, lp2], dtype=np.float64)
    out[f"ari/{ari_n}/coeffs_a"] = a.coeffs
    out[f"ari/{ari_n}/coeffs_b"] = b.coeffs
    for name, res in (("add", a + b), ("sub", a - b), ("mul", a * b)):
        out[f"ari/{ari_n}/{name}/exponents"] = res.multi_index.exponents.astype(np.uint8)
        out[f"ari/{ari_n}/{name}/lp"] = np.float64(res.multi_index.lp_degree)
        out[f"ari/{ari_n}/{name}/generating_points"] = res.grid.generating_points
        out[f"ari/{ari_n}/{name}/coeffs_ref"] = res.coeffs
    ari_n += 1
out["ari/count"] = np.int64(ari_n)

# ---- incomplete index sets: dense operator ------------------------------------------------------
rng = np.random.default_rng(815)
inc_n = 0
for (m, n, lp, drop) in [(2, 3, 2.0, [4]), (2, 4, 1.0, [2, 7]), (3, 3, 2.0, [5, 11]),
                         (3, 2, np.inf, [13]), (2, 12, 2.0, [30, 31, 77])]:
    full = mp.MultiIndexSet.from_degree(m, n, lp).exponents
    mi = mp.MultiIndexSet(np.delete(full, drop, axis=0), lp_degree=lp)
    assert not mi.is_complete
    grid = mp.Grid(mi)
    N = len(mi)
    c1 = rng.uniform(-1, 1, N)
    c3 = rng.uniform(-1, 1, (N, 3))
    trafo = mp.LagrangeToNewton(mp.LagrangePolynomial(mi, c1, grid=grid))
    assert type(trafo.transformation_operator).__name__ == "MatrixOperator"
    out[f"inc/{inc_n}/exponents"] = mi.exponents.astype(np.uint8)
    out[f"inc/{inc_n}/lp"] = np.float64(lp)
    out[f"inc/{inc_n}/generating_points"] = grid.generating_points
    out[f"inc/{inc_n}/coeffs1"] = c1
    out[f"inc/{inc_n}/coeffs3"] = c3
    out[f"inc/{inc_n}/operator_ref"] = trafo.transformation_operator.array_repr_full
    out[f"inc/{inc_n}/l2n1_ref"] = trafo().coeffs
    out[f"inc/{inc_n}/l2n3_ref"] = mp.LagrangeToNewton(
        mp.LagrangePolynomial(mi, c3, grid=grid))().coeffs
    inc_n += 1
out["inc/count"] = np.int64(inc_n)

path = os.path.join(HERE, "next_rows.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")
