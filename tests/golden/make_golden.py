"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference.

Run here (the container that has /root/reference):  python tests/golden/make_golden.py
The reference cannot travel to the GPU box, the vectors can.  Every array that defines
the hot path's inputs (exponents, generating points) comes from the reference's own
``MultiIndexSet.from_degree`` / ``Grid``; every ``*_ref`` array is an output of the
reference's ``eval_newton_polynomials`` / ``eval_newton_monomials`` / ``dds`` /
``LagrangeToNewton`` / ``NewtonToLagrange``.

Large random inputs are not stored: they are re-drawn from a stored seed with
``numpy.random.default_rng`` (PCG64) and guarded by a stored sha256, see ``golden_io.py``.
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import load_reference  # noqa: E402
from golden_io import draw_uniform, sha256_of  # noqa: E402

mp = load_reference()
assert mp is not None, "needs /root/reference"
from minterpy.dds import dds as ref_dds  # noqa: E402
from minterpy.utils import eval_newton_monomials, eval_newton_polynomials  # noqa: E402


def shape_arrays(m, n, lp):
    mi = mp.MultiIndexSet.from_degree(m, n, lp)
    grid = mp.Grid(mi)
    return mi, grid


def corners(m):
    return np.array([[1.0 if (i >> d) & 1 else -1.0 for d in range(m)] for i in range(2 ** m)])


def small_cases():
    """The reference's own test grid (tests/conftest.py:178-235), plus m=2."""
    out = {}
    for m in (1, 2, 3):
        for n in (0, 1, 4):
            for lp in (0.5, 1.0, 2.0, np.inf):
                tag = f"m{m}_n{n}_lp{lp}"
                mi, grid = shape_arrays(m, n, lp)
                e = mi.exponents
                gp = grid.generating_points
                N = e.shape[0]
                rng = np.random.default_rng(1000 * m + 10 * n + int(min(lp, 9)))
                xx = np.concatenate([rng.uniform(-1, 1, (21, m)), grid.unisolvent_nodes,
                                     corners(m)])
                for k in (1, 10):
                    c = rng.uniform(-1, 1, (N, k))
                    ev = eval_newton_polynomials(xx, c if k > 1 else c[:, 0], e, gp)
                    out[f"{tag}/k{k}/coeffs"] = c
                    out[f"{tag}/k{k}/eval_ref"] = np.asarray(ev).reshape(xx.shape[0], k)
                    vals = rng.uniform(-1, 1, (N, k))
                    out[f"{tag}/k{k}/dds_in"] = vals
                    out[f"{tag}/k{k}/dds_ref"] = ref_dds(vals if k > 1 else vals[:, 0],
                                                         grid.tree)
                    # transformation classes (barycentric operators for these complete sets)
                    lag = mp.LagrangePolynomial(mi, vals if k > 1 else vals[:, 0], grid=grid)
                    out[f"{tag}/k{k}/l2n_ref"] = np.asarray(
                        mp.LagrangeToNewton(lag)().coeffs).reshape(N, k)
                    newt = mp.NewtonPolynomial(mi, c if k > 1 else c[:, 0], grid=grid)
                    out[f"{tag}/k{k}/n2l_ref"] = np.asarray(
                        mp.NewtonToLagrange(newt)().coeffs).reshape(N, k)
                out[f"{tag}/exponents"] = e.astype(np.uint8)
                out[f"{tag}/generating_points"] = gp
                out[f"{tag}/unisolvent_nodes"] = grid.unisolvent_nodes
                out[f"{tag}/xx"] = xx
                out[f"{tag}/monomials_ref"] = eval_newton_monomials(xx, e, gp)
                out[f"{tag}/monomials_tri_ref"] = eval_newton_monomials(
                    grid.unisolvent_nodes, e, gp, triangular=True)
    return out


# BASELINE.json configs: tag -> (m, n, lp, k_eval, k_dds_stored)
CONFIGS = {
    "cfg1": (1, 64, 2.0, 1, 1),
    "cfg2": (4, 16, 2.0, 1, 2),
    "cfg3": (6, 10, 1.0, 64, 2),
    "cfg4": (5, 12, 2.0, 256, 2),
    "cfg5": (3, 32, np.inf, 1, 1),
}


def config_case(tag):
    m, n, lp, k_eval, k_dds = CONFIGS[tag]
    cfg = int(tag[3:])
    t0 = time.time()
    mi, grid = shape_arrays(m, n, lp)
    e = mi.exponents
    gp = grid.generating_points
    N = e.shape[0]
    out = {"exponents": e.astype(np.uint8), "generating_points": gp,
           "poly_degree": np.int64(mi.poly_degree),
           "unisolvent_sha": np.array(sha256_of(grid.unisolvent_nodes))}
    # evaluation: random points + a spread of unisolvent nodes + the corners
    rng = np.random.default_rng(20261106 + cfg)
    n_rand = 192 if N * k_eval > 4_000_000 else 1024
    pick = np.unique(np.concatenate([[0, N - 1], rng.integers(0, N, 62)]))
    xx = np.concatenate([rng.uniform(-1, 1, (n_rand, m)), grid.unisolvent_nodes[pick],
                         corners(m)])
    out["xx"] = xx
    seed_c = 7 + cfg
    c = draw_uniform(seed_c, (N, k_eval))
    out["coeffs_seed"] = np.int64(seed_c)
    out["coeffs_sha"] = np.array(sha256_of(c))
    out["k_eval"] = np.int64(k_eval)
    ev = eval_newton_polynomials(xx, c if k_eval > 1 else c[:, 0], e, gp)
    out["eval_ref"] = np.asarray(ev).reshape(xx.shape[0], k_eval)
    print(tag, "eval done", time.time() - t0, flush=True)
    # dds: stored columns (in + out), plus a hash-pinned wide case for cfg4
    tree = grid.tree
    print(tag, "tree done", time.time() - t0, flush=True)
    seed_v = 70 + cfg
    vals = draw_uniform(seed_v, (N, k_dds))
    out["dds_seed"] = np.int64(seed_v)
    out["dds_in_sha"] = np.array(sha256_of(vals))
    out["dds_ref"] = ref_dds(vals if k_dds > 1 else vals[:, 0], tree)
    # smooth values f(x) = cos(w.x + phi): decaying coefficients, the realistic case
    w = rng.uniform(0, 3, m)
    phi = rng.uniform(0, 2 * np.pi)
    f = np.cos(grid.unisolvent_nodes @ w + phi)
    out["smooth_in"] = f
    out["smooth_dds_ref"] = ref_dds(f, tree)
    if tag == "cfg4":
        wide = draw_uniform(700 + cfg, (N, 256))
        out["dds_wide_seed"] = np.int64(700 + cfg)
        out["dds_wide_in_sha"] = np.array(sha256_of(wide))
        out["dds_wide_ref_sha"] = np.array(sha256_of(ref_dds(wide, tree)))
    if tag == "cfg1":
        # README quickstart: interpolate x*sin(10x), degree 64, 1e6 points (README.md:95-121)
        fct = lambda x: x * np.sin(10 * x)  # noqa: E731
        interpolant = mp.interpolate(fct, spatial_dimension=1, poly_degree=64)
        xq = np.linspace(-1, 1, 1_000_000)
        yq = interpolant(xq)
        poly = interpolant._Interpolant__interpolation_poly
        out["quickstart_fvals"] = fct(grid.unisolvent_nodes).reshape(-1)
        out["quickstart_coeffs"] = poly.coeffs
        out["quickstart_every1000"] = yq[::1000].copy()
        out["quickstart_sha"] = np.array(sha256_of(yq))
    print(tag, "done", time.time() - t0, flush=True)
    return out


def extra_shapes():
    """Index sets and grids only (no reference outputs): shapes between the BASELINE configs,
    used by the GPU tests against the pinned oracle."""
    out = {}
    for (m, n, lp) in [(3, 24, 2.0), (2, 40, 1.0), (4, 20, 1.0)]:
        mi, grid = shape_arrays(m, n, lp)
        tag = f"m{m}_n{n}_lp{lp}"
        out[f"{tag}/exponents"] = mi.exponents.astype(np.uint8)
        out[f"{tag}/generating_points"] = grid.generating_points
    return out


if __name__ == "__main__":
    which = sys.argv[1:] or ["small"] + list(CONFIGS)
    for tag in which:
        path = os.path.join(HERE, f"{tag}.npz")
        data = (small_cases() if tag == "small" else extra_shapes() if tag == "shapes_extra"
                else config_case(tag))
        np.savez_compressed(path, **data)
        print("wrote", path, os.path.getsize(path), flush=True)
