"""Pin the CPU oracle (oracle/newton_oracle.c) against the reference's own outputs.

The golden vectors were written by tests/golden/make_golden.py, which imports the
unmodified reference.  Tolerances: monomials and dds bit-exact; polynomial values
<= 1e-13 * max|ref| (the reference ends in a BLAS product, summation order unspecified).
"""
import numpy as np
import pytest

import golden_io
import oracle

EVAL_RTOL = 1e-13
SMALL_TAGS = [f"m{m}_n{n}_lp{lp}" for m in (1, 2, 3) for n in (0, 1, 4)
              for lp in (0.5, 1.0, 2.0, np.inf)]
CFGS = ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"]


def rel_err(a, ref):
    scale = np.max(np.abs(ref))
    return np.max(np.abs(a - ref)) / (scale if scale > 0 else 1.0)


@pytest.mark.parametrize("tag", SMALL_TAGS)
def test_small_monomials_bit_exact(golden_small, tag):
    g = golden_small
    e, gp, xx = g[f"{tag}/exponents"], g[f"{tag}/generating_points"], g[f"{tag}/xx"]
    mono = oracle.eval_newton_monomials(xx, e, gp)
    assert np.array_equal(mono, g[f"{tag}/monomials_ref"])
    tri = oracle.eval_newton_monomials(g[f"{tag}/unisolvent_nodes"], e, gp, triangular=True)
    assert np.array_equal(tri, g[f"{tag}/monomials_tri_ref"])


@pytest.mark.parametrize("k", [1, 10])
@pytest.mark.parametrize("tag", SMALL_TAGS)
def test_small_eval_and_dds(golden_small, tag, k):
    g = golden_small
    e, gp, xx = g[f"{tag}/exponents"], g[f"{tag}/generating_points"], g[f"{tag}/xx"]
    ev = oracle.eval_newton_polynomials(xx, g[f"{tag}/k{k}/coeffs"], e, gp)
    assert rel_err(ev, g[f"{tag}/k{k}/eval_ref"]) <= EVAL_RTOL
    out = oracle.dds(g[f"{tag}/k{k}/dds_in"], e, gp)
    assert np.array_equal(out, g[f"{tag}/k{k}/dds_ref"])          # bit-exact
    # LagrangeToNewton (barycentric operator in the reference) is the same map as dds
    assert rel_err(out, g[f"{tag}/k{k}/l2n_ref"]) <= 1e-12
    # NewtonToLagrange == evaluation at the unisolvent nodes
    n2l = oracle.eval_newton_polynomials(g[f"{tag}/unisolvent_nodes"],
                                         g[f"{tag}/k{k}/coeffs"], e, gp)
    assert rel_err(n2l, g[f"{tag}/k{k}/n2l_ref"]) <= EVAL_RTOL


@pytest.mark.parametrize("tag", CFGS)
def test_config_eval(golden_cfg, tag):
    g = golden_cfg(tag)
    e, gp = g["exponents"], g["generating_points"]
    k = int(g["k_eval"])
    c = golden_io.seeded(g, "coeffs_seed", "coeffs_sha", (e.shape[0], k))
    ev = oracle.eval_newton_polynomials(g["xx"], c, e, gp, nthreads=4)
    assert rel_err(ev, g["eval_ref"]) <= EVAL_RTOL


@pytest.mark.parametrize("tag", CFGS)
def test_config_dds_bit_exact(golden_cfg, tag):
    g = golden_cfg(tag)
    e, gp = g["exponents"], g["generating_points"]
    ref = g["dds_ref"]
    vals = golden_io.seeded(g, "dds_seed", "dds_in_sha", ref.shape)
    assert np.array_equal(oracle.dds(vals, e, gp), ref)
    sm = oracle.dds(g["smooth_in"], e, gp)
    assert np.array_equal(sm, g["smooth_dds_ref"])
    assert golden_io.sha256_of(gp[e, np.arange(e.shape[1])]) == str(g["unisolvent_sha"])


def test_cfg4_wide_dds_hash(golden_cfg):
    g = golden_cfg("cfg4")
    e, gp = g["exponents"], g["generating_points"]
    wide = golden_io.seeded(g, "dds_wide_seed", "dds_wide_in_sha", (e.shape[0], 256))
    assert golden_io.sha256_of(oracle.dds(wide, e, gp)) == str(g["dds_wide_ref_sha"])


def test_quickstart(golden_cfg):
    """README quickstart (cfg1): interpolate x sin(10x), degree 64."""
    g = golden_cfg("cfg1")
    e, gp = g["exponents"], g["generating_points"]
    coeffs = oracle.dds(g["quickstart_fvals"], e, gp).reshape(-1)
    assert np.array_equal(coeffs, g["quickstart_coeffs"])
    xq = np.linspace(-1, 1, 1_000_000)[::1000]
    y = oracle.eval_newton_polynomials(xq, coeffs, e, gp).reshape(-1)
    assert rel_err(y, g["quickstart_every1000"]) <= EVAL_RTOL


def test_incomplete_set_rejected():
    e = np.array([[0, 0], [2, 0]], dtype=np.int64)     # (1,0) missing
    gp = np.zeros((3, 2)); gp[:, 0] = [0.0, 1.0, -1.0]; gp[:, 1] = [0.0, -1.0, 1.0]
    with pytest.raises(ValueError):
        oracle.dds(np.ones(2), e, gp)


def test_derivative_evaluation_against_reference():
    """orc_deriv_newt_eval against deriv_newt_eval of the reference
    (src/minterpy/polynomials/utils.py:8-112); the final dot product is BLAS in the reference,
    so parity is <= 1e-13 * max|ref| per derivative (exact zeros must be exact zeros)."""
    d = golden_io.load("next_rows")
    assert len(d["der/tags"]) >= 6
    for tag in d["der/tags"]:
        e, gp = d[f"der/{tag}/exponents"], d[f"der/{tag}/generating_points"]
        x, c = d[f"der/{tag}/x"], d[f"der/{tag}/coeffs"]
        for o, want in zip(d[f"der/{tag}/orders"], d[f"der/{tag}/values_ref"]):
            got = oracle.deriv_newt_eval(x, c, e, gp, o)[:, 0]
            scale = np.abs(want).max()
            if scale == 0.0:
                assert not got.any()
            else:
                assert np.abs(got - want).max() <= 1e-13 * scale, (tag, o)
        # order zero is the polynomial itself
        plain = oracle.eval_newton_polynomials(x, c, e, gp)[:, 0]
        assert np.array_equal(oracle.deriv_newt_eval(x, c, e, gp, np.zeros(e.shape[1]))[:, 0], plain)


@pytest.mark.parametrize("tag", CFGS)
def test_config_wide_golden(tag):
    """4096 random points + all unisolvent nodes + corners per BASELINE config,
    values from the unmodified reference (tests/golden/make_golden_wide.py).  The oracle checks a
    spread of them here (every random point for the small sets; the GPU test checks them all)."""
    xx, c, ref, n2l = golden_io.wide_case(tag)
    g = golden_io.load(tag)
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    assert ref.shape == (4096 + N + 2 ** m, c.shape[1])
    step = max(1, (xx.shape[0] * N * c.shape[1]) // 400_000_000)       # bounded CPU time
    sel = np.unique(np.concatenate([np.arange(0, xx.shape[0], step), np.arange(xx.shape[0] - 2 ** m, xx.shape[0])]))
    ev = oracle.eval_newton_polynomials(xx[sel], c, e, gp, nthreads=8)
    assert rel_err(ev, ref[sel]) <= EVAL_RTOL
    if n2l is not None:
        # the reference's NewtonToLagrange (barycentric operator) on the first stored column is
        # evaluation at the unisolvent nodes, which the same reference evaluated above
        assert rel_err(ref[4096:4096 + N, 0], n2l) <= 1e-12
