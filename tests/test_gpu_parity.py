"""GPU parity tests: the CUDA path (through the C ABI) against the golden vectors written
by the unmodified reference and against the CPU oracle.

Tolerances (stated once, used everywhere):
  * polynomial values: max|gpu - ref| <= EVAL_RTOL * max|ref|, EVAL_RTOL = 1e-13
    (the contract in BASELINE.json is 1e-12; the reference's own BLAS product is only
    defined up to summation order);
  * Newton monomials and divided differences: bit-exact.
"""
import ctypes

import numpy as np
import pytest

import golden_io
import oracle

pytestmark = pytest.mark.gpu

EVAL_RTOL = 1e-13
SMALL_TAGS = [f"m{m}_n{n}_lp{lp}" for m in (1, 2, 3) for n in (0, 1, 4)
              for lp in (0.5, 1.0, 2.0, np.inf)]
CFGS = ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"]


@pytest.fixture(scope="module")
def mpb():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import minterpy_b200
    return minterpy_b200


def rel_err(a, ref):
    scale = np.max(np.abs(ref))
    return np.max(np.abs(np.asarray(a) - ref)) / (scale if scale > 0 else 1.0)


@pytest.mark.parametrize("tag", SMALL_TAGS)
def test_small_golden(mpb, golden_small, tag):
    g = golden_small
    e, gp, xx = g[f"{tag}/exponents"], g[f"{tag}/generating_points"], g[f"{tag}/xx"]
    plan = mpb.get_plan(e, gp)
    for k in (1, 10):
        ev = plan.eval(xx, g[f"{tag}/k{k}/coeffs"])
        assert rel_err(ev, g[f"{tag}/k{k}/eval_ref"]) <= EVAL_RTOL
        out = plan.dds(g[f"{tag}/k{k}/dds_in"])
        assert np.array_equal(out, g[f"{tag}/k{k}/dds_ref"])
    mono = plan.monomials(xx)
    assert np.array_equal(mono, g[f"{tag}/monomials_ref"])
    tri = plan.monomials(g[f"{tag}/unisolvent_nodes"], triangular=True)
    assert np.array_equal(tri, g[f"{tag}/monomials_tri_ref"])


@pytest.mark.parametrize("tag", CFGS)
def test_config_eval_golden(mpb, golden_cfg, tag):
    g = golden_cfg(tag)
    e, gp = g["exponents"], g["generating_points"]
    k = int(g["k_eval"])
    c = golden_io.seeded(g, "coeffs_seed", "coeffs_sha", (e.shape[0], k))
    ev = mpb.get_plan(e, gp).eval(g["xx"], c)
    assert rel_err(ev, g["eval_ref"]) <= EVAL_RTOL


@pytest.mark.parametrize("tag", CFGS)
def test_config_dds_golden_bit_exact(mpb, golden_cfg, tag):
    g = golden_cfg(tag)
    e, gp = g["exponents"], g["generating_points"]
    plan = mpb.get_plan(e, gp)
    ref = g["dds_ref"]
    vals = golden_io.seeded(g, "dds_seed", "dds_in_sha", ref.shape)
    assert np.array_equal(plan.dds(vals), ref)
    assert np.array_equal(plan.dds(g["smooth_in"].reshape(-1, 1)), g["smooth_dds_ref"])


@pytest.mark.parametrize("tag", CFGS)
def test_config_wide_golden(mpb, tag):
    """SURVEY.md 8c at full width: 4096 random points + ALL unisolvent nodes + the 2^m corners
    of every BASELINE config against the unmodified reference's values, and NewtonToLagrange at
    all nodes (cfg2: 24 809, cfg4: 59 711) against the reference's own transformation."""
    xx, c, ref, n2l = golden_io.wide_case(tag)
    g = golden_io.load(tag)
    e, gp = g["exponents"], g["generating_points"]
    N = e.shape[0]
    plan = mpb.get_plan(e, gp)
    got = plan.eval(xx, c)
    assert got.shape == ref.shape
    assert rel_err(got, ref) <= EVAL_RTOL
    if n2l is not None:
        mi = mpb.MultiIndexSet(e, lp_degree=2.0)
        newt = mpb.NewtonPolynomial(mi, np.ascontiguousarray(c[:, 0]), grid=mpb.Grid(mi, gp))
        lag = mpb.NewtonToLagrange(newt)()
        assert lag.coeffs.shape == (N,)
        assert rel_err(lag.coeffs, n2l) <= EVAL_RTOL
        # and back: divided differences of those values return the coefficients (up to the
        # conditioning of the round trip with random coefficients, 2.5e-8 at cfg2)
        back = mpb.LagrangeToNewton(lag)()
        assert rel_err(back.coeffs, c[:, 0]) <= 1e-6


def test_cfg4_wide_dds_hash(mpb, golden_cfg):
    """BASELINE config 4 at full size: 59 711 x 256 values, bit-exact through a sha256."""
    g = golden_cfg("cfg4")
    e, gp = g["exponents"], g["generating_points"]
    wide = golden_io.seeded(g, "dds_wide_seed", "dds_wide_in_sha", (e.shape[0], 256))
    out = mpb.get_plan(e, gp).dds(wide)
    assert golden_io.sha256_of(out) == str(g["dds_wide_ref_sha"])


def test_quickstart(mpb, golden_cfg):
    """README quickstart: interpolate x sin(10x) with degree 64, evaluate at 1e6 points."""
    g = golden_cfg("cfg1")
    mi = mpb.MultiIndexSet(g["exponents"])
    grid = mpb.Grid(mi, g["generating_points"])
    poly = mpb.interpolate(lambda x: (x * np.sin(10 * x)).reshape(-1), grid)
    assert np.array_equal(poly.coeffs, g["quickstart_coeffs"])
    y = poly(np.linspace(-1, 1, 1_000_000))
    assert y.shape == (1_000_000,)
    assert rel_err(y[::1000], g["quickstart_every1000"]) <= EVAL_RTOL


def test_against_oracle_irregular_sets(mpb):
    """Unsorted index sets with holes: evaluation must not depend on completeness."""
    rng = np.random.default_rng(11)
    for trial in range(12):
        m = int(rng.integers(1, 6))
        N = int(rng.integers(1, 200))
        e = np.unique(rng.integers(0, 9, (N, m)), axis=0).astype(np.int64)
        rng.shuffle(e)
        gp = rng.uniform(-1, 1, (10, m))
        k = int(rng.integers(1, 4))
        c = rng.uniform(-1, 1, (e.shape[0], k))
        xx = rng.uniform(-1, 1, (37, m))
        plan = mpb.Plan(e, gp)
        assert rel_err(plan.eval(xx, c), oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL
        assert np.array_equal(plan.monomials(xx), oracle.eval_newton_monomials(xx, e, gp))
        if not plan.is_complete:
            with pytest.raises(ValueError):
                plan.dds(np.ones((e.shape[0], 1)))


@pytest.mark.parametrize("P", [0, 1, 255, 256, 511, 513, 1025])
def test_ragged_point_counts(mpb, golden_cfg, P):
    g = golden_cfg("cfg2")
    e, gp = g["exponents"][:500], g["generating_points"]      # a lexicographic prefix is complete
    rng = np.random.default_rng(P)
    c = rng.uniform(-1, 1, (e.shape[0], 2))
    xx = rng.uniform(-1, 1, (P, 4))
    out = mpb.get_plan(e, gp).eval(xx, c)
    assert out.shape == (P, 2)
    if P:
        assert rel_err(out, oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL


def test_device_path_matches_host_path_and_strides(mpb, golden_cfg):
    import torch
    g = golden_cfg("cfg3")
    e, gp = g["exponents"], g["generating_points"]
    rng = np.random.default_rng(3)
    N, m = e.shape
    c = rng.uniform(-1, 1, (N, 5))
    xx = rng.uniform(-1, 1, (300, m))
    plan = mpb.get_plan(e, gp)
    host = plan.eval(xx, c)
    # zero-copy torch path with non-trivial leading dimensions
    xt = torch.zeros((300, m + 3), dtype=torch.float64, device="cuda")
    xt[:, :m] = torch.from_numpy(xx).cuda()
    ct = torch.zeros((N, 8), dtype=torch.float64, device="cuda")
    ct[:, :5] = torch.from_numpy(c).cuda()
    dev = plan.eval(xt[:, :m], ct[:, :5])
    assert dev.is_cuda and dev.shape == (300, 5)
    assert np.array_equal(dev.cpu().numpy(), host)
    assert rel_err(host, oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL
    # dds on a CUDA tensor: result on the device, input untouched
    vt = torch.from_numpy(c).cuda()
    keep = vt.clone()
    dt = plan.dds(vt)
    assert torch.equal(vt, keep)
    assert np.array_equal(dt.cpu().numpy(), oracle.dds(c, e, gp))


def test_api_shapes_and_errors(mpb, golden_small):
    """Shape, squeeze and error rules of src/minterpy/core/verification.py:48-163."""
    g = golden_small
    tag = "m3_n4_lp2.0"
    e, gp = g[f"{tag}/exponents"], g[f"{tag}/generating_points"]
    N = e.shape[0]
    c1, c10 = g[f"{tag}/k1/coeffs"][:, 0], g[f"{tag}/k10/coeffs"]
    xx = g[f"{tag}/xx"]
    f = mpb.eval_newton_polynomials
    assert f(xx, c1, e, gp).shape == (xx.shape[0],)
    assert f(xx, c10, e, gp).shape == (xx.shape[0], 10)
    assert f(xx[0], c10, e, gp).shape == (10,)              # 1-D x with m > 1: one point
    assert f(xx[0], c1, e, gp).shape == (1,)                # never 0-D
    assert f(xx[:1], c1.reshape(N, 1), e, gp).shape == (1,)
    for bs in (1, 100, 1000):                               # batch_size must not change results
        assert np.array_equal(f(xx, c10, e, gp, batch_size=bs), f(xx, c10, e, gp))
    with pytest.raises(ValueError):
        f(xx[:, :2], c1, e, gp)                             # wrong point dimension
    with pytest.raises(ValueError):
        f(xx[0, :2], c1, e, gp)
    with pytest.raises(ValueError):
        f(xx, c10[:-1], e, gp)                              # coefficient count mismatch
    with pytest.raises(ValueError):
        f(xx, c1[:0], e[:0], gp)                            # no monomials
    with pytest.raises(TypeError):
        f(xx.astype(np.float32), c1, e, gp)
    with pytest.raises(TypeError):
        f(xx, c1, e.astype(np.int32), gp)
    # m == 1: a 1-D x is a list of points
    e1, gp1 = g["m1_n4_lp2.0/exponents"], g["m1_n4_lp2.0/generating_points"]
    assert f(np.linspace(-1, 1, 7), np.ones(5), e1, gp1).shape == (7,)
    # dds: 1-D input -> (N, 1); works on a copy
    tree = mpb.Grid(mpb.MultiIndexSet(e), gp).tree
    vals = g[f"{tag}/k1/dds_in"][:, 0].copy()
    keep = vals.copy()
    out = mpb.dds(vals, tree)
    assert out.shape == (N, 1) and np.array_equal(vals, keep)
    assert np.array_equal(out, g[f"{tag}/k1/dds_ref"])


def test_transformations_round_trip(mpb, golden_small):
    """LagrangeToNewton / NewtonToLagrange against the reference's operators, and
    back-and-forth identity (tests/test_transformation.py:119-173,197-208)."""
    g = golden_small
    for tag in ("m3_n4_lp2.0", "m2_n4_lpinf", "m1_n4_lp1.0", "m3_n1_lp0.5", "m3_n0_lp2.0"):
        e, gp = g[f"{tag}/exponents"], g[f"{tag}/generating_points"]
        mi = mpb.MultiIndexSet(e)
        grid = mpb.Grid(mi, gp)
        assert np.array_equal(grid.unisolvent_nodes, g[f"{tag}/unisolvent_nodes"])
        for k in (1, 10):
            lag_c = g[f"{tag}/k{k}/dds_in"] if k > 1 else g[f"{tag}/k{k}/dds_in"][:, 0]
            lag = mpb.LagrangePolynomial(mi, lag_c, grid=grid)
            newt = mpb.LagrangeToNewton(lag)()
            assert isinstance(newt, mpb.NewtonPolynomial) and newt.coeffs.shape == lag_c.shape
            assert rel_err(newt.coeffs.reshape(len(mi), k), g[f"{tag}/k{k}/l2n_ref"]) <= 1e-12
            back = mpb.NewtonToLagrange(newt)()
            assert rel_err(back.coeffs, lag_c) <= 1e-12
            nc = g[f"{tag}/k{k}/coeffs"] if k > 1 else g[f"{tag}/k{k}/coeffs"][:, 0]
            n2l = mpb.NewtonToLagrange(mpb.NewtonPolynomial(mi, nc, grid=grid))()
            assert rel_err(n2l.coeffs.reshape(len(mi), k), g[f"{tag}/k{k}/n2l_ref"]) <= EVAL_RTOL
        op = mpb.LagrangeToNewton(lag).transformation_operator
        n2l_op = mpb.NewtonToLagrange(newt).transformation_operator
        assert rel_err(op.array_repr_full @ n2l_op.array_repr_full, np.eye(len(mi))) <= 1e-10


def test_large_properties_cfg2(mpb, golden_cfg):
    """BASELINE config 2 at a size the oracle cannot reach in seconds: size-independent
    properties -- linearity in the coefficients, and interpolation reproducing a polynomial
    (tests/test_interpolation.py:71-105)."""
    import torch
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    gen = torch.Generator(device="cuda").manual_seed(20261106)
    P = 1 << 20
    xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    c = torch.rand((N, 2), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    y = plan.eval(xx, c)
    ysum = plan.eval(xx, (c[:, 0] + 0.5 * c[:, 1]).reshape(N, 1))
    scale = y.abs().max().item()
    assert (ysum[:, 0] - (y[:, 0] + 0.5 * y[:, 1])).abs().max().item() <= 1e-12 * scale
    # spot-check 64 of the 2^20 points against the oracle
    idx = torch.randint(0, P, (64,), generator=gen, device="cuda")
    ref = oracle.eval_newton_polynomials(xx[idx].cpu().numpy(), c.cpu().numpy(), e, gp, nthreads=4)
    assert rel_err(y[idx].cpu().numpy(), ref) <= EVAL_RTOL
    # round trip: values at the unisolvent nodes -> dds -> the same coefficients
    grid = mpb.Grid(mpb.MultiIndexSet(e), gp)
    nodes = torch.from_numpy(grid.unisolvent_nodes).cuda()
    lag = plan.eval(nodes, c)
    back = plan.dds(lag)
    # dds at N = 24 809 amplifies rounding errors of its input by ~1e8 (the oracle shows the
    # same loss on the same values), so the round trip is a coarse check ...
    assert (back - c).abs().max().item() <= 1e-5 * c.abs().max().item()
    # ... and the sharp one is bit-exactness against the oracle on those values
    assert np.array_equal(back.cpu().numpy(), oracle.dds(lag.cpu().numpy(), e, gp))


def test_dense_apply(mpb):
    import torch
    from minterpy_b200 import _lib
    rng = np.random.default_rng(2)
    for rows, cols, k in ((1, 1, 1), (70, 70, 1), (130, 97, 5), (513, 513, 9), (3, 1025, 2),
                          (515, 700, 3), (1027, 1301, 4), (258, 257, 1)):
        T = np.tril(rng.uniform(-1, 1, (rows, cols)))
        c = rng.uniform(-1, 1, (cols, k))
        Tt, ct = torch.from_numpy(T).cuda(), torch.from_numpy(c).cuda()
        out = torch.empty((rows, k), dtype=torch.float64, device="cuda")
        rc = _lib.lib().mpx_apply_dense(0, Tt.data_ptr(), rows, cols, cols, ct.data_ptr(), k, k,
                                        out.data_ptr(), k, None)
        _lib.check(rc)
        assert rel_err(out.cpu().numpy(), T @ c) <= 1e-13
    # a view into a wider matrix: leading dimension larger than the column count, and a base
    # address that is only 8-byte aligned (first column dropped)
    wide = rng.uniform(-1, 1, (301, 1200))
    Wt = torch.from_numpy(wide).cuda()
    for first, cols in ((0, 1199), (1, 1199), (1, 700), (2, 1001)):
        c = rng.uniform(-1, 1, (cols, 3))
        ct = torch.from_numpy(c).cuda()
        out = torch.empty((301, 3), dtype=torch.float64, device="cuda")
        view = Wt[:, first:first + cols]
        rc = _lib.lib().mpx_apply_dense(0, view.data_ptr(), 301, cols, 1200, ct.data_ptr(), 3, 3,
                                        out.data_ptr(), 3, None)
        _lib.check(rc)
        assert rel_err(out.cpu().numpy(), wide[:, first:first + cols] @ c) <= 1e-13


def test_dense_apply_many_columns_and_lower_triangular(mpb):
    """k > 8 runs as one pass over T per 32 columns on the FP64 tensor cores (DMMA, contraction
    index permuted inside 16-column groups); MPX_DENSE_LOWER skips what lies right of the
    diagonal.  Odd and even leading dimensions, unaligned views, ragged last spans."""
    import torch
    from minterpy_b200 import _lib
    rng = np.random.default_rng(8)
    lib = _lib.lib()
    for rows, cols, k in ((1, 1, 9), (70, 70, 16), (130, 97, 12), (513, 513, 17), (33, 1025, 32),
                          (515, 700, 40), (1027, 1302, 9), (258, 257, 70), (400, 128, 24)):
        full = rng.uniform(-1, 1, (rows, cols))
        c = rng.uniform(-1, 1, (cols, k))
        ct = torch.from_numpy(c).cuda()
        for flags, T in ((0, full), (_lib.MPX_DENSE_LOWER, np.tril(full))):
            Tt = torch.from_numpy(T).cuda()
            out = torch.full((rows, k), float("nan"), dtype=torch.float64, device="cuda")
            _lib.check(lib.mpx_apply_dense_ex(0, Tt.data_ptr(), rows, cols, cols, ct.data_ptr(), k, k,
                                              out.data_ptr(), k, flags, None))
            assert rel_err(out.cpu().numpy(), T @ c) <= 1e-13, (rows, cols, k, flags)
    # lower-triangular flag on the few-column (FMA) kernels, odd sizes
    for n, k in ((1027, 1), (1301, 3), (514, 8), (96, 2)):
        T = np.tril(rng.uniform(-1, 1, (n, n)))
        c = rng.uniform(-1, 1, (n, k))
        Tt, ct = torch.from_numpy(T).cuda(), torch.from_numpy(c).cuda()
        out = torch.empty((n, k), dtype=torch.float64, device="cuda")
        _lib.check(lib.mpx_apply_dense_ex(0, Tt.data_ptr(), n, n, n, ct.data_ptr(), k, k,
                                          out.data_ptr(), k, _lib.MPX_DENSE_LOWER, None))
        assert rel_err(out.cpu().numpy(), T @ c) <= 1e-13
    # what lies right of the diagonal is NOT read beyond the CTA's own column block: poison it
    n, k = 777, 11
    T = np.tril(rng.uniform(-1, 1, (n, n)))
    poisoned = T.copy()
    iu = np.triu_indices(n, 64)
    poisoned[iu] = np.nan
    c = rng.uniform(-1, 1, (n, k))
    Tt, ct = torch.from_numpy(poisoned).cuda(), torch.from_numpy(c).cuda()
    for kk in (11, 2):
        out = torch.empty((n, kk), dtype=torch.float64, device="cuda")
        cc = ct[:, :kk].contiguous()
        _lib.check(lib.mpx_apply_dense_ex(0, Tt.data_ptr(), n, n, n, cc.data_ptr(), kk, kk,
                                          out.data_ptr(), kk, _lib.MPX_DENSE_LOWER, None))
        assert rel_err(out.cpu().numpy(), T @ c[:, :kk]) <= 1e-13
    # unaligned view with many columns
    wide = rng.uniform(-1, 1, (301, 1200))
    Wt = torch.from_numpy(wide).cuda()
    for first, cols in ((1, 1199), (2, 1001)):
        c = rng.uniform(-1, 1, (cols, 20))
        ct = torch.from_numpy(c).cuda()
        out = torch.empty((301, 20), dtype=torch.float64, device="cuda")
        view = Wt[:, first:first + cols]
        _lib.check(lib.mpx_apply_dense(0, view.data_ptr(), 301, cols, 1200, ct.data_ptr(), 20, 20,
                                       out.data_ptr(), 20, None))
        assert rel_err(out.cpu().numpy(), wide[:, first:first + cols] @ c) <= 1e-13


@pytest.mark.parametrize("k", [8, 9, 16, 31, 64, 100])
def test_many_polynomials_tensor_core_path(mpb, golden_cfg, k):
    """Many polynomials: the DMMA kernel (csrc/mpx_eval_mma.cu) from k = 8 (short leaves, cfg3)
    or k = 13 (long leaves, cfg2), the tree kernel below that; one tolerance for both."""
    for tag, npts in (("cfg3", 333), ("cfg2", 97)):
        g = golden_cfg(tag)
        e, gp = g["exponents"], g["generating_points"]
        rng = np.random.default_rng(k)
        c = rng.uniform(-1, 1, (e.shape[0], k))
        xx = rng.uniform(-1, 1, (npts, e.shape[1]))
        got = mpb.get_plan(e, gp).eval(xx, c)
        ref = oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4)
        assert rel_err(got, ref) <= EVAL_RTOL


def test_tensor_core_path_irregular_and_small(mpb):
    rng = np.random.default_rng(21)
    for trial in range(6):
        m = int(rng.integers(1, 8))
        e = np.unique(rng.integers(0, 7, (int(rng.integers(1, 150)), m)), axis=0).astype(np.int64)
        rng.shuffle(e)
        gp = rng.uniform(-1, 1, (8, m))
        k = int(rng.integers(8, 40))
        c = rng.uniform(-1, 1, (e.shape[0], k))
        xx = rng.uniform(-1, 1, (int(rng.integers(1, 300)), m))
        got = mpb.Plan(e, gp).eval(xx, c)
        assert rel_err(got, oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL


def test_long_dimension0_and_many_dimensions(mpb):
    """Shapes outside the tree kernel's register table (dimension-0 exponent > 64) run on the
    shared-memory-table kernel; m up to 16 runs on the tree kernel."""
    rng = np.random.default_rng(77)
    # 1-D, degree 150 (Chebyshev-like points), and 2-D with a long first dimension
    n = 150
    gp = np.cos(np.pi * np.arange(n + 1) / n).reshape(-1, 1)
    e = np.arange(n + 1, dtype=np.int64).reshape(-1, 1)
    c = rng.uniform(-1, 1, (n + 1, 2)) * np.exp(-0.05 * np.arange(n + 1))[:, None]
    xx = rng.uniform(-1, 1, (500, 1))
    plan = mpb.Plan(e, gp)
    assert rel_err(plan.eval(xx, c), oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL
    assert np.array_equal(plan.dds(c), oracle.dds(c, e, gp))
    e2 = np.array([[a, b] for b in range(4) for a in range(90 - 20 * b)], dtype=np.int64)
    gp2 = np.stack([np.cos(np.pi * np.arange(91) / 90), -np.cos(np.pi * np.arange(91) / 90)], axis=1)
    c2 = rng.uniform(-1, 1, (e2.shape[0], 1))
    xx2 = rng.uniform(-1, 1, (123, 2))
    assert rel_err(mpb.Plan(e2, gp2).eval(xx2, c2), oracle.eval_newton_polynomials(xx2, c2, e2, gp2)) <= EVAL_RTOL
    # many dimensions: total degree 2 in m = 9, 12, 16
    for m in (9, 12, 16):
        rows = [[0] * m]
        for i in range(m):
            for v in (1, 2):
                r = [0] * m; r[i] = v; rows.append(r)
            for j in range(i):
                r = [0] * m; r[i] = 1; r[j] = 1; rows.append(r)
        em = np.array(rows, dtype=np.int64)
        em = em[np.lexsort(em.T)]
        gpm = np.tile(np.array([[1.0], [-1.0], [0.0]]), (1, m)) * np.where(np.arange(m) % 2 == 0, -1, 1)
        cm = rng.uniform(-1, 1, (em.shape[0], 3))
        xm = rng.uniform(-1, 1, (77, m))
        planm = mpb.Plan(em, gpm)
        assert rel_err(planm.eval(xm, cm), oracle.eval_newton_polynomials(xm, cm, em, gpm)) <= EVAL_RTOL
        assert planm.is_complete
        assert np.array_equal(planm.dds(cm), oracle.dds(cm, em, gpm))


def test_long_higher_dimensions(mpb):
    """Small dimension-0 table but very long tables in dimensions 1 and 2 (tree kernel with few
    points per CTA), k below and above the tensor-core threshold."""
    rng = np.random.default_rng(78)
    rows = [[a, b, c] for c in range(0, 3) for b in range(0, 200 - 60 * c, 7) for a in range(3)]
    rows += [[0, b, 0] for b in range(200) if b % 7]
    e = np.unique(np.array(rows, dtype=np.int64), axis=0)
    e = e[np.lexsort(e.T)]
    pts = np.cos(np.pi * np.arange(201) / 200)
    gp = np.stack([pts, -pts, pts], axis=1)
    xx = rng.uniform(-1, 1, (60, 3))
    for k in (1, 3, 12):
        c = rng.uniform(-1, 1, (e.shape[0], k)) * 1e-3
        got = mpb.Plan(e, gp).eval(xx, c)
        assert rel_err(got, oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL


@pytest.fixture()
def tmma_mode():
    from minterpy_b200 import _lib
    yield _lib.lib().mpx_debug_set_tmma
    _lib.lib().mpx_debug_set_tmma(0)


@pytest.mark.parametrize("force", [1, 2])
def test_both_tree_kernels(mpb, golden_small, golden_cfg, tmma_mode, force):
    """Complete index sets with few polynomials can run on the DFMA tree kernel (mpx_eval.cu) or
    on the tensor-core tree kernel (mpx_eval_tmma.cu); the plan picks by padding overhead.  Both
    are forced here (test hook, read when the plan is built) on every golden shape."""
    tmma_mode(force)
    for tag in SMALL_TAGS:
        e, gp, xx = golden_small[f"{tag}/exponents"], golden_small[f"{tag}/generating_points"], golden_small[f"{tag}/xx"]
        plan = mpb.Plan(e, gp)
        for k in (1, 10):
            c = golden_small[f"{tag}/k{k}/coeffs"][:, :5]          # k < 8: one of the tree kernels
            ref = golden_small[f"{tag}/k{k}/eval_ref"][:, :5]
            assert rel_err(plan.eval(xx, c), ref) <= EVAL_RTOL
    for tag in ("cfg1", "cfg2", "cfg4", "cfg5"):
        g = golden_cfg(tag)
        e, gp = g["exponents"], g["generating_points"]
        rng = np.random.default_rng(4)
        c = rng.uniform(-1, 1, (e.shape[0], 2))
        xx = np.concatenate([g["xx"][:300], -np.ones((1, e.shape[1])), np.ones((1, e.shape[1]))])
        got = mpb.Plan(e, gp).eval(xx, c)
        assert rel_err(got, oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4)) <= EVAL_RTOL


def test_intermediate_shapes_against_oracle(mpb):
    """Index sets between the BASELINE configs (from the reference, tests/golden/shapes_extra.npz):
    long dimension-0 tables with staircase leaves -- the shapes where the plan switches between
    the DFMA and the tensor-core tree kernel."""
    g = golden_io.load("shapes_extra")
    rng = np.random.default_rng(31)
    for tag in ("m3_n24_lp2.0", "m2_n40_lp1.0", "m4_n20_lp1.0"):
        e, gp = g[f"{tag}/exponents"], g[f"{tag}/generating_points"]
        plan = mpb.Plan(e, gp)
        for k in (1, 3):
            c = rng.uniform(-1, 1, (e.shape[0], k))
            xx = rng.uniform(-1, 1, (257, e.shape[1]))
            assert rel_err(plan.eval(xx, c), oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4)) <= EVAL_RTOL
        v = rng.uniform(-1, 1, (e.shape[0], 2))
        assert np.array_equal(plan.dds(v), oracle.dds(v, e, gp))


@pytest.mark.parametrize("tag,P,k", [("cfg5", 1 << 20, 1), ("cfg3", 1 << 17, 64), ("cfg4", 1 << 15, 256)])
def test_large_properties_other_configs(mpb, golden_cfg, tag, P, k):
    """BASELINE configs 3-5 at sizes the oracle cannot reach in seconds (tensor-core tree kernel,
    DMMA contraction kernel): linearity in the coefficients and an oracle spot check."""
    import torch
    g = golden_cfg(tag)
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    gen = torch.Generator(device="cuda").manual_seed(5 + k)
    xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    c1 = torch.rand((N, k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    c2 = torch.rand((N, k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    y1, y2 = plan.eval(xx, c1), plan.eval(xx, c2)
    y12 = plan.eval(xx, c1 - 0.25 * c2)
    scale = max(y1.abs().max().item(), y2.abs().max().item())
    assert (y12 - (y1 - 0.25 * y2)).abs().max().item() <= 1e-12 * scale
    idx = torch.randint(0, P, (24,), generator=gen, device="cuda")
    ref = oracle.eval_newton_polynomials(xx[idx].cpu().numpy(), c1.cpu().numpy(), e, gp, nthreads=4)
    assert rel_err(y1[idx].cpu().numpy(), ref) <= EVAL_RTOL


def test_concurrent_host_threads(mpb, golden_cfg):
    """ctypes releases the GIL during the foreign call, so Python threads really do enter the
    library at once (SURVEY.md 8b "Threading"): several threads on ONE plan (serialised by the
    plan's mutex: staging buffers and packed coefficients are per plan) and threads on
    different plans (independent) must all get the single-threaded answers, bit for bit."""
    import threading
    g2, g5 = golden_cfg("cfg2"), golden_cfg("cfg5")
    plans = [mpb.get_plan(g2["exponents"], g2["generating_points"]),
             mpb.get_plan(g5["exponents"], g5["generating_points"])]
    rng = np.random.default_rng(77)
    jobs = []
    for t in range(8):
        plan = plans[t % 2] if t < 4 else plans[0]
        xx = rng.uniform(-1, 1, (3000 + 511 * t, plan.m))
        c = rng.uniform(-1, 1, (plan.N, 1 + t % 3))
        v = rng.uniform(-1, 1, (plan.N, 2))
        jobs.append((plan, xx, c, v))
    want = [(p.eval(xx, c), p.dds(v), p.monomials(xx[:64])) for p, xx, c, v in jobs]
    got = [None] * len(jobs)
    errors = []

    def work(i):
        # the two plans launch the same kernels with different shared-memory sizes: a
        # per-launch function attribute would race here (it did, once in five runs)
        try:
            p, xx, c, v = jobs[i]
            for _ in range(6):
                got[i] = (p.eval(xx, c), p.dds(v), p.monomials(xx[:64]))
        except Exception as ex:       # noqa: BLE001
            errors.append(ex)

    threads = [threading.Thread(target=work, args=(i,)) for i in range(len(jobs))]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    for (we, wd, wm), (ge, gd, gm) in zip(want, got):
        assert np.array_equal(we, ge) and np.array_equal(wd, gd) and np.array_equal(wm, gm)


def test_dds_fused_slice_kernel_still_bit_exact(mpb, golden_small, golden_cfg):
    """The default dds path is one warp per fiber; the fused slice kernel stays as the fallback
    for fibers whose buffers do not fit shared memory (a test hook routes every call through
    it).  Same goldens, same bit-exact bar."""
    from minterpy_b200 import _lib
    _lib.lib().mpx_debug_set_dds_group(1)
    try:
        n = 0
        for tag in SMALL_TAGS:
            plan = mpb.get_plan(golden_small[f"{tag}/exponents"], golden_small[f"{tag}/generating_points"])
            for k in (1, 10):
                assert np.array_equal(plan.dds(golden_small[f"{tag}/k{k}/dds_in"]),
                                      golden_small[f"{tag}/k{k}/dds_ref"]), tag
                n += 1
        for tag in ("cfg2", "cfg3", "cfg4", "cfg5"):
            g = golden_cfg(tag)
            e, gp = g["exponents"], g["generating_points"]
            vals = golden_io.seeded(g, "dds_seed", "dds_in_sha", g["dds_ref"].shape)
            assert np.array_equal(mpb.get_plan(e, gp).dds(vals), g["dds_ref"]), tag
            n += 1
        assert n == 76
    finally:
        _lib.lib().mpx_debug_set_dds_group(0)


def test_one_thread_per_gpu(mpb, golden_cfg):
    """one process per GPU or one THREAD per GPU must both work.  Needs two
    devices (skipped on a single-GPU box): a plan per device, threads that never call
    cudaSetDevice themselves, inputs as NumPy (the library selects the plan's device) and as
    tensors that live on the plan's device."""
    import threading
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    rng = np.random.default_rng(5)
    xx = rng.uniform(-1, 1, (20000, 4))
    c = rng.uniform(-1, 1, (e.shape[0], 2))
    v = rng.uniform(-1, 1, (e.shape[0], 3))
    ref_plan = mpb.get_plan(e, gp, device=0)
    want = (ref_plan.eval(xx, c), ref_plan.dds(v))
    results, errors = {}, []

    def work(dev):
        try:
            plan = mpb.get_plan(e, gp, device=dev)
            assert plan.device == dev
            for _ in range(3):
                a = plan.eval(xx, c)
                b = plan.dds(v)
                xt = torch.as_tensor(xx, device=f"cuda:{dev}")
                ct = torch.as_tensor(c, device=f"cuda:{dev}")
                t = plan.eval(xt, ct)
                assert t.device.index == dev
                results[dev] = (a, b, t.cpu().numpy())
        except Exception as ex:       # noqa: BLE001
            errors.append((dev, ex))

    threads = [threading.Thread(target=work, args=(d,)) for d in range(torch.cuda.device_count())]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    for dev, (a, b, t) in results.items():
        assert np.array_equal(a, want[0]) and np.array_equal(b, want[1]) and np.array_equal(t, want[0])


def test_kernel_choice_by_leaf_length(mpb, golden_cfg):
    """MPX_INFO_MMA_MIN_K: the longer the leaves, the longer the tree kernel run once per
    polynomial beats the many-polynomial tensor kernel (cfg5: up to k = 32, cfg2: 26, cfg4: 12);
    short-leaf sets (cfg3) go to the tensor kernel from k = 4; values on both sides of each
    threshold agree with the oracle to the one tolerance."""
    from minterpy_b200 import _lib
    want = {"cfg2": 27, "cfg4": 13, "cfg5": 33, "cfg3": 4}
    rng = np.random.default_rng(31)
    for tag, min_k in want.items():
        g = golden_cfg(tag)
        e, gp = g["exponents"], g["generating_points"]
        plan = mpb.get_plan(e, gp)
        assert plan.info(_lib.INFO_MMA_MIN_K) == min_k
        xx = rng.uniform(-1, 1, (70, e.shape[1]))
        for k in (min_k - 1, min_k):
            c = rng.uniform(-1, 1, (e.shape[0], k))
            ref = oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4)
            assert rel_err(plan.eval(xx, c), ref) <= EVAL_RTOL, (tag, k)
