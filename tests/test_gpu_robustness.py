"""GPU tests of the host-side machinery around the kernels: coefficient images shared between
the tensor path and the NumPy path of one plan, streams, pageable host memory, the
bulk-copy ring of the tree kernel under stress, the plan cache.  All through the C ABI."""
import ctypes
import threading

import numpy as np
import pytest

import golden_io
import oracle

pytestmark = pytest.mark.gpu
EVAL_RTOL = 1e-13


def rel_err(got, ref):
    return float(np.max(np.abs(np.asarray(got) - np.asarray(ref))) / max(np.max(np.abs(ref)), 1e-300))


@pytest.fixture(scope="module")
def mpb():
    import torch
    assert torch.cuda.is_available(), "the gpu tests need a CUDA device"
    import minterpy_b200
    return minterpy_b200


def test_tensor_call_keeps_its_coefficients_while_a_numpy_call_repacks(mpb, golden_cfg):
    """A long tensor-path evaluation is still running when a NumPy-path evaluation with other
    coefficients enters the same plan (plans are shared by content, so callers cannot know).
    Each path packs into an image of its own; both results must be right."""
    import torch
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    rng = np.random.default_rng(5)
    P = 3_000_000                                    # ~8 ms of kernel: the call returns long before
    xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
    c_t = torch.as_tensor(rng.uniform(-1, 1, (N, 1)), device="cuda")
    c_n = rng.uniform(-1, 1, (N, 1)) * 100.0
    xn = rng.uniform(-1, 1, (20_000, m))
    idx = torch.randint(0, P, (32,), device="cuda")
    for _ in range(3):
        y_t = plan.eval(xx, c_t)                     # enqueued, still running
        y_n = plan.eval(xn, c_n)                     # NumPy path: uploads + packs + evaluates
        torch.cuda.synchronize()
        ref_t = oracle.eval_newton_polynomials(xx[idx].cpu().numpy(), c_t.cpu().numpy(), e, gp, nthreads=4)
        assert rel_err(y_t[idx].cpu().numpy(), ref_t) <= EVAL_RTOL
        assert rel_err(y_t[-4:].cpu().numpy(),
                       oracle.eval_newton_polynomials(xx[-4:].cpu().numpy(), c_t.cpu().numpy(), e, gp)) <= EVAL_RTOL
        assert rel_err(y_n[:64], oracle.eval_newton_polynomials(xn[:64], c_n, e, gp)) <= EVAL_RTOL


def test_two_streams_on_one_plan(mpb, golden_cfg):
    """Two torch streams evaluate different polynomials on one plan back to back: the second
    call's repack waits (on the device) for the first call's kernel."""
    import torch
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    gen = torch.Generator(device="cuda").manual_seed(3)
    xx = torch.rand((2_000_000, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    ca = torch.rand((N, 1), generator=gen, dtype=torch.float64, device="cuda")
    cb = -7.0 * torch.rand((N, 1), generator=gen, dtype=torch.float64, device="cuda")
    want_a, want_b = plan.eval(xx, ca), plan.eval(xx, cb)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(4):
        with torch.cuda.stream(s1):
            ya = plan.eval(xx, ca)
        with torch.cuda.stream(s2):
            yb = plan.eval(xx, cb)
        torch.cuda.synchronize()
        assert torch.equal(ya, want_a) and torch.equal(yb, want_b)


def test_threads_mixing_tensor_and_numpy_calls(mpb, golden_cfg):
    import torch
    g = golden_cfg("cfg5")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    rng = np.random.default_rng(9)
    xs = [rng.uniform(-1, 1, (40_000 + 1000 * i, m)) for i in range(4)]
    cs = [rng.uniform(-1, 1, (N, 1 + i % 2)) * (i + 1) for i in range(4)]
    want = [plan.eval(x, c) for x, c in zip(xs, cs)]
    got, errors = [None] * 4, []

    def work(i):
        try:
            for _ in range(5):
                if i % 2:
                    got[i] = plan.eval(xs[i], cs[i])
                else:
                    st = torch.cuda.Stream()
                    with torch.cuda.stream(st):
                        y = plan.eval(torch.as_tensor(xs[i], device="cuda"), torch.as_tensor(cs[i], device="cuda"))
                    st.synchronize()
                    got[i] = y.cpu().numpy()
        except Exception as ex:       # noqa: BLE001
            errors.append(ex)

    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors
    for w, g_ in zip(want, got):
        assert np.array_equal(w, g_)


def _host_call(plan, xx, c, out):
    from minterpy_b200 import _lib
    rc = _lib.lib().mpx_eval_host(plan._h, xx.ctypes.data, xx.shape[0], xx.strides[0] // 8,
                                  c.ctypes.data, c.shape[1], c.strides[0] // 8, out.ctypes.data,
                                  out.strides[0] // 8)
    _lib.check(rc)


def test_pageable_and_page_locked_host_buffers_agree(mpb, golden_cfg):
    """mpx_eval_host / mpx_dds_host stage pageable arrays through page-locked blocks with host
    threads; page-locked arrays go to the copy engines directly.  Same bits either way, also
    with padded rows (leading dimensions larger than the row) and several blocks."""
    import torch
    from minterpy_b200 import _lib
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    rng = np.random.default_rng(11)
    P = 4_000_001                                    # 3 blocks of ~1.67e6 points
    xx = rng.uniform(-1, 1, (P, m + 1))[:, :m]       # pageable, ld = m + 1
    c = rng.uniform(-1, 1, (N, 1))
    out_pageable = np.full((P, 2), np.nan)[:, :1]    # pageable, ld = 2
    _host_call(plan, xx, c, out_pageable)
    xp = torch.empty((P, m), dtype=torch.float64, pin_memory=True).numpy()
    xp[:] = xx
    out_pinned = torch.empty((P, 1), dtype=torch.float64, pin_memory=True).numpy()
    _host_call(plan, xp, c, out_pinned)
    assert np.array_equal(out_pageable, out_pinned)
    idx = rng.integers(0, P, 24)
    assert rel_err(out_pinned[idx], oracle.eval_newton_polynomials(np.ascontiguousarray(xx[idx]), c, e, gp)) <= EVAL_RTOL
    # dds: pageable in/out (column blocks repacked by host threads) vs page-locked vs the
    # reference's result (sha256 of the 59 711 x 256 golden case)
    g4 = golden_cfg("cfg4")
    plan4 = mpb.get_plan(g4["exponents"], g4["generating_points"])
    vals = golden_io.seeded(g4, "dds_wide_seed", "dds_wide_in_sha", (g4["exponents"].shape[0], 256))
    lib = _lib.lib()
    out_a = np.empty_like(vals)
    _lib.check(lib.mpx_dds_host(plan4._h, vals.ctypes.data, vals.shape[1], vals.shape[1],
                                out_a.ctypes.data, vals.shape[1]))
    vp = torch.empty(vals.shape, dtype=torch.float64, pin_memory=True).numpy()
    vp[:] = vals
    out_b = torch.empty(vals.shape, dtype=torch.float64, pin_memory=True).numpy()
    _lib.check(lib.mpx_dds_host(plan4._h, vp.ctypes.data, vals.shape[1], vals.shape[1],
                                out_b.ctypes.data, vals.shape[1]))
    assert golden_io.sha256_of(out_a) == str(g4["dds_wide_ref_sha"])
    assert np.array_equal(out_a, out_b)
    k = 100                                          # ragged column blocks, padded rows
    wide_in = np.full((vals.shape[0], k + 3), np.nan)
    wide_in[:, :k] = vals[:, :k]
    wide_out = np.full((vals.shape[0], k + 5), np.nan)
    _lib.check(lib.mpx_dds_host(plan4._h, wide_in.ctypes.data, k, k + 3, wide_out.ctypes.data, k + 5))
    assert np.array_equal(wide_out[:, :k], out_a[:, :k]) and np.isnan(wide_out[:, k:]).all()


@pytest.mark.parametrize("stages", [2, 3, 4])
def test_tree_kernel_ring_under_stress(mpb, golden_cfg, stages):
    """The tree kernel streams its program through a shared-memory ring refilled by thread 0
    (cp.async.bulk + mbarrier, csrc/mpx_eval.cu).  Many chunks through few stages, point counts
    that leave tail CTAs and odd CTA counts, several polynomials per launch: every value is
    checked against the default configuration and a sample against the oracle."""
    import torch
    from minterpy_b200 import _lib
    lib = _lib.lib()
    for tag in ("cfg4", "cfg2", "cfg3"):             # 60+, 26 and ~20 chunks
        g = golden_cfg(tag)
        e, gp = g["exponents"], g["generating_points"]
        N, m = e.shape
        plan = mpb.Plan(e, gp)
        assert plan.info(_lib.INFO_CHUNKS) > 2 * stages
        gen = torch.Generator(device="cuda").manual_seed(17 + stages)
        for P, k in ((1, 1), (1023, 2), (151_553, 1), (37 * 1024 + 5, 3)):
            xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
            c = torch.rand((N, k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
            want = plan.eval(xx, c)
            torch.cuda.synchronize()
            lib.mpx_debug_set_eval_stages(stages)
            try:
                for _ in range(3):
                    got = plan.eval(xx, c)
                    torch.cuda.synchronize()
                    # different ring depth = different CTA shape, same arithmetic per point
                    assert torch.equal(got, want), (tag, P, k)
            finally:
                lib.mpx_debug_set_eval_stages(0)
            sel = torch.randint(0, P, (min(P, 8),), generator=gen, device="cuda")
            ref = oracle.eval_newton_polynomials(xx[sel].cpu().numpy(), c.cpu().numpy(), e, gp, nthreads=4)
            assert rel_err(want[sel].cpu().numpy(), ref) <= EVAL_RTOL


def test_plan_cache_confirms_the_arrays(mpb, golden_small, monkeypatch):
    """A cache hit is confirmed by comparing the arrays: with every checksum forced to collide,
    two different index sets still get their own plans."""
    import zlib
    from minterpy_b200 import plan as plan_mod
    plan_mod.clear_plan_cache()
    monkeypatch.setattr(zlib, "crc32", lambda a: 1)
    monkeypatch.setattr(zlib, "adler32", lambda a: 2)
    ta, tb = "m2_n4_lp1.0", "m2_n4_lpinf"
    ea, gpa = golden_small[f"{ta}/exponents"], golden_small[f"{ta}/generating_points"]
    eb = ea.copy()
    eb[-1, 0] += 1                                   # same shape, different set
    pa = plan_mod.get_plan(ea, gpa)
    pb = plan_mod.get_plan(eb, gpa)
    assert pa is not pb and np.array_equal(pb.exponents, eb)
    assert plan_mod.get_plan(eb, gpa) is pb
    xx, c = golden_small[f"{ta}/xx"], golden_small[f"{ta}/k1/coeffs"]
    assert rel_err(pb.eval(xx, c), oracle.eval_newton_polynomials(xx, c, eb, gpa)) <= EVAL_RTOL
    plan_mod.clear_plan_cache()


def test_regression_on_a_set_with_holes_with_tensors(mpb, nxt_rows):
    """Lagrange-basis regression on an incomplete index set, tensor inputs: the Lagrange->Newton
    operator is the dense MatrixOperator there and has to hand its matrix over on the device."""
    import torch
    from minterpy_b200.regression import compute_regression_matrix
    nxt = nxt_rows
    e, gp = nxt["inc/0/exponents"], nxt["inc/0/generating_points"]
    mi = mpb.MultiIndexSet(e, lp_degree=float(nxt["inc/0/lp"]))
    grid = mpb.Grid(mi)
    basis = mpb.LagrangePolynomial(mi, None, grid=grid)
    rng = np.random.default_rng(2)
    xx = rng.uniform(-1, 1, (50, e.shape[1]))
    r_np = compute_regression_matrix(basis, xx)
    r_t = compute_regression_matrix(basis, torch.as_tensor(xx, device="cuda"))
    assert r_t.is_cuda and rel_err(r_t.cpu().numpy(), r_np) <= 1e-12
    # interpolatory on the nodes: the Lagrange basis is the identity there
    ident = compute_regression_matrix(basis, torch.as_tensor(grid.unisolvent_nodes, device="cuda"))
    assert np.max(np.abs(ident.cpu().numpy() - np.eye(len(mi)))) <= 1e-10


@pytest.fixture(scope="module")
def nxt_rows():
    return golden_io.load("next_rows")


def _same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


def _same_bits_up_to_nan_payload(a, b):
    """Same bit patterns, except that a nan only has to be a nan (the CPU keeps a payload and a
    sign that the GPU's canonical nan does not)."""
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and _same_bits(np.where(na, 0.0, a), np.where(nb, 0.0, b))


def test_dds_kernels_agree_bit_for_bit_on_ragged_widths_and_special_values(mpb, golden_cfg, golden_small):
    """The several-columns-per-lane dds kernel (default for even k >= 34) against the
    one-column-per-lane kernel (test hook), bit patterns compared: column counts that leave the
    last column block ragged or nearly empty, and values the fast division must not take --
    zeros of both signs, equal neighbours, subnormals, huge numbers, inf and nan -- which send
    single lanes through the redo path."""
    import torch
    from minterpy_b200 import _lib
    lib = _lib.lib()
    rng = np.random.default_rng(123)
    cases = [("cfg2", golden_cfg("cfg2")), ("cfg5", golden_cfg("cfg5")), ("cfg3", golden_cfg("cfg3"))]
    small = golden_small
    cases.append(("m3_n4_lpinf", {"exponents": small["m3_n4_lpinf/exponents"],
                                  "generating_points": small["m3_n4_lpinf/generating_points"]}))
    for tag, g in cases:
        e, gp = g["exponents"], g["generating_points"]
        N = e.shape[0]
        plan = mpb.get_plan(e, gp)
        for k in (34, 48, 62, 64, 66, 100, 128, 130, 190, 254, 256, 320):
            vals = rng.uniform(-1, 1, (N, k))
            special = k in (48, 66, 130, 256)
            if special:
                idx = rng.integers(0, N, 40)
                col = rng.integers(0, k, 40)
                vals[idx[:8], col[:8]] = 0.0
                vals[idx[8:12], col[8:12]] = -0.0
                vals[idx[12:16], col[12:16]] = 1e-310
                vals[idx[16:20], col[16:20]] = 1e300
                vals[idx[20:22], col[20:22]] = np.inf
                vals[idx[22:24], col[22:24]] = np.nan
                vals[:, 3] = 0.25                               # a constant column: all zeros after
                vals[:, 5] = 0.0
            v = torch.as_tensor(vals, device="cuda")
            got = plan.dds(v).cpu().numpy()
            lib.mpx_debug_set_dds_group(2)                      # one column per lane
            try:
                want = plan.dds(v).cpu().numpy()
            finally:
                lib.mpx_debug_set_dds_group(0)
            # (a column holding both inf and nan: which nan comes out depends on the order of
            # fast and true divisions, so nans only have to sit in the same places)
            assert _same_bits_up_to_nan_payload(got, want), (tag, k)
            if not special:
                assert np.array_equal(got[:, :2], oracle.dds(np.ascontiguousarray(vals[:, :2]), e, gp)), (tag, k)
            elif tag != "cfg5":
                # and the true divisions of the oracle, signs of zeros and nan positions included
                with np.errstate(all="ignore"):
                    assert _same_bits_up_to_nan_payload(got, oracle.dds(vals, e, gp)), (tag, k)
        # columns of the identity (what an operator's matrix is built from): zeros of both signs
        # at every step
        for k in (32, 64, 256):
            eye = np.eye(N, k, k=-int(rng.integers(0, max(1, N - k))))
            got = plan.dds(torch.as_tensor(eye, device="cuda")).cpu().numpy()
            ref = oracle.dds(eye, e, gp)
            assert _same_bits(got, ref), (tag, "identity", k)
            # the MPX_DDS_MANY_ZEROS instantiation (mpx_dds_ex): the same bits, on zeros and on
            # ordinary and special values alike
            assert _same_bits(plan.dds(eye, many_zeros=True), ref), (tag, "identity, hint", k)
            hinted = plan.dds(v, many_zeros=True, inplace=False).cpu().numpy()
            assert hinted.shape == want.shape and _same_bits_up_to_nan_payload(hinted, want), (tag, "hint", k)


def test_uniform_control_tree_kernel_is_bit_identical_to_the_ring_kernel(mpb, golden_cfg, golden_small):
    """Index sets without irregular records run the tree kernel that takes its record stream as a
    kernel parameter (uniform control flow, csrc/mpx_eval_kernel.cuh); a test hook sends the same
    plan through the ring-fed kernel.  Same arithmetic per point: identical bits, on every golden
    shape, for several polynomials, ragged point counts, and with the tensor-core tree kernel
    switched off so that the DFMA kernels are what runs."""
    import torch
    from minterpy_b200 import _lib
    lib = _lib.lib()
    shapes = [(t, golden_small[f"{t}/exponents"], golden_small[f"{t}/generating_points"])
              for t in sorted({k.split("/")[0] for k in golden_small})]
    for tag in ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5"):
        g = golden_cfg(tag)
        shapes.append((tag, g["exponents"], g["generating_points"]))
    extra = golden_io.load("shapes_extra")
    for t in sorted({k.split("/")[0] for k in extra}):
        shapes.append((t, extra[f"{t}/exponents"], extra[f"{t}/generating_points"]))
    gen = torch.Generator(device="cuda").manual_seed(5)
    used = 0
    lib.mpx_debug_set_tmma(2)
    try:
        for tag, e, gp in shapes:
            N, m = e.shape
            plan = mpb.Plan(e, gp)
            used += plan.info(_lib.INFO_UNIFORM_PROGRAM)
            for P, k in ((1, 1), (777, 3), (40_001, 2)):
                xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
                c = torch.rand((N, k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
                a = plan.eval(xx, c)
                lib.mpx_debug_set_eval_uniform(2)
                try:
                    b = plan.eval(xx, c)
                finally:
                    lib.mpx_debug_set_eval_uniform(0)
                assert torch.equal(a, b), (tag, P, k)
    finally:
        lib.mpx_debug_set_tmma(0)
    assert used >= len(shapes) - 2          # complete index sets are regular: nearly all qualify


def _random_complete_set(mpb, rng):
    """A downward-closed index set in the reference's ordering: an l_p set cut by per-dimension
    caps (rows are dropped, never reordered, so the lexicographic order of core/utils.py:44-92
    is kept)."""
    from minterpy_b200 import generators
    m = int(rng.integers(1, 7))
    n = int(rng.integers(1, {1: 60, 2: 40, 3: 20, 4: 12, 5: 8, 6: 6}[m]))
    lp = float(rng.choice([0.5, 1.0, 2.0, 3.0, np.inf]))
    e = generators.get_exponent_matrix(m, n, lp)
    caps = rng.integers(0, n + 1, m)
    caps[int(rng.integers(0, m))] = n
    e = np.ascontiguousarray(e[np.all(e <= caps, axis=1)])
    rows = int(e.max()) + 2
    gp = rng.uniform(-1, 1, (rows, m))
    # distinct values per column, as a grid needs them
    gp += np.arange(rows)[:, None] * 1e-3
    return e, gp


def test_seeded_sweep_of_complete_sets_against_the_oracle(mpb):
    """40 seeded downward-closed sets (m = 1..6, l_p cut by caps, random generating points):
    evaluation on every kernel the polynomial count selects, divided differences bit for bit at
    widths on both sides of the column-kernel threshold, and one derivative each."""
    rng = np.random.default_rng(20261107)
    for trial in range(40):
        e, gp = _random_complete_set(mpb, rng)
        N, m = e.shape
        plan = mpb.Plan(e, gp)
        assert plan.is_complete
        P = int(rng.integers(1, 700))
        xx = rng.uniform(-1, 1, (P, m))
        for k in (1, int(rng.integers(2, 9)), int(rng.integers(9, 80))):
            c = rng.uniform(-1, 1, (N, k))
            ref = oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4)
            assert rel_err(plan.eval(xx, c).reshape(P, k), ref.reshape(P, k)) <= EVAL_RTOL, (trial, e.shape, k)
        for k in (1, int(rng.integers(2, 64)), 64 + 2 * int(rng.integers(0, 40)), 129):
            v = rng.uniform(-1, 1, (N, k))
            assert np.array_equal(plan.dds(v), oracle.dds(v, e, gp)), (trial, e.shape, k)
        orders = rng.integers(0, 3, m)
        k = int(rng.choice([1, 2, 11]))
        c = rng.uniform(-1, 1, (N, k))
        ref = oracle.deriv_newt_eval(xx, c, e, gp, orders)
        got = plan.eval_deriv(xx, c, orders)
        scale = max(np.max(np.abs(ref)), 1e-300)
        assert np.max(np.abs(np.asarray(got).reshape(P, k) - ref.reshape(P, k))) <= 1e-12 * scale, (trial, e.shape, orders)


def test_one_dimensional_polynomials_by_horner_rule(mpb):
    """m = 1 runs on the Horner kernel (mpx_eval_1d.cu) for any number of polynomials: against the
    oracle and against the tree kernel (test hook) on dense, gappy and unsorted exponent lists,
    ragged point counts, strided inputs, degrees beyond the tree kernel's 64, and duplicates
    (which must stay on the tree kernel)."""
    import torch
    from minterpy_b200 import _lib
    rng = np.random.default_rng(64)
    cases = [np.arange(65), np.arange(1), np.arange(2), np.array([0, 2, 5, 9]), np.array([7, 0, 3, 64, 12]),
             np.arange(200), np.array([255, 0, 100])]
    for ex in cases:
        e = ex.reshape(-1, 1).astype(np.int64)
        gp = np.sort(rng.uniform(-1, 1, (int(e.max()) + 2, 1)), axis=0)[::-1].copy()
        plan = mpb.Plan(e, gp)
        assert plan.info(_lib.INFO_SMALLK_KERNEL) == 3
        for k in (1, 2, 3, 4, 5, 32, 33):
            for P in (1, 255, 2049, 70001):
                xx = rng.uniform(-1, 1, (P, 1))
                c = rng.uniform(-1, 1, (e.shape[0], k))
                ref = oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=4).reshape(P, k)
                assert rel_err(plan.eval(xx, c).reshape(P, k), ref) <= EVAL_RTOL, (ex[:4], k, P)
        # strided device tensors: the points are one column of a wider array, the coefficients
        # and the result slices of wider ones
        P, k = 5000, 3
        wide = torch.rand((P, 4), dtype=torch.float64, device="cuda") * 2 - 1
        cw = torch.rand((e.shape[0], 8), dtype=torch.float64, device="cuda")
        ow = torch.zeros((P, 6), dtype=torch.float64, device="cuda")
        got = plan._eval_torch(wide[:, 2:3], cw[:, 1:1 + k], out=ow[:, 2:2 + k])
        ref = oracle.eval_newton_polynomials(wide[:, 2:3].cpu().numpy().copy(), cw[:, 1:1 + k].cpu().numpy().copy(), e, gp)
        assert rel_err(got.cpu().numpy(), ref.reshape(P, k)) <= EVAL_RTOL
        assert float(ow[:, :2].abs().max()) == 0.0 and float(ow[:, 5:].abs().max()) == 0.0
    # the tree kernel on the same inputs (what m = 1 ran on before): within rounding of Horner
    e = np.arange(65).reshape(-1, 1).astype(np.int64)
    gp = rng.uniform(-1, 1, (65, 1))
    xx = rng.uniform(-1, 1, (4097, 1))
    c = rng.uniform(-1, 1, (65, 2))
    horner = mpb.Plan(e, gp).eval(xx, c)
    _lib.lib().mpx_debug_set_eval_1d(2)
    try:
        tree_plan = mpb.Plan(e, gp)
        assert tree_plan.info(_lib.INFO_SMALLK_KERNEL) == 0
        tree = tree_plan.eval(xx, c)
    finally:
        _lib.lib().mpx_debug_set_eval_1d(0)
    assert rel_err(horner, tree) <= EVAL_RTOL
    # duplicate exponents: the reference adds both terms; such a list stays on the tree kernel
    e = np.array([0, 1, 1, 2]).reshape(-1, 1).astype(np.int64)
    gp = rng.uniform(-1, 1, (4, 1))
    plan = mpb.Plan(e, gp)
    assert plan.info(_lib.INFO_SMALLK_KERNEL) != 3
    c = rng.uniform(-1, 1, (4, 1))
    xx = rng.uniform(-1, 1, (33, 1))
    assert rel_err(plan.eval(xx, c), oracle.eval_newton_polynomials(xx, c, e, gp)) <= EVAL_RTOL
