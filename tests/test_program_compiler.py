"""CPU tests of the plan compiler (host logic of the C-ABI library, no GPU needed): the
program it emits, interpreted in Python, must reproduce the oracle for regular
(downward-closed) and irregular index sets."""
import numpy as np
import pytest

import oracle
from program_interp import compile_program, run_program

SMALL_TAGS = [f"m{m}_n{n}_lp{lp}" for m in (1, 2, 3) for n in (0, 1, 4)
              for lp in (0.5, 1.0, 2.0, np.inf)]


def _check(e, gp, n_points=3, seed=0):
    rng = np.random.default_rng(seed)
    N, m = e.shape
    coeffs = rng.uniform(-1, 1, N)
    xx = rng.uniform(-1, 1, (n_points, m))
    desc, slots, fma = compile_program(e)
    ref = oracle.eval_newton_polynomials(xx, coeffs, e, gp).reshape(-1)
    for p in range(n_points):
        val, fma_run = run_program(desc, slots, xx[p], coeffs, e, gp)
        assert abs(val - ref[p]) <= 1e-12 * max(1.0, np.max(np.abs(ref)))
        assert fma_run == fma          # the plan's FMA count is what the walk executes
    return fma


@pytest.mark.parametrize("tag", SMALL_TAGS)
def test_small_sets(golden_small, tag):
    _check(golden_small[f"{tag}/exponents"], golden_small[f"{tag}/generating_points"])


@pytest.mark.parametrize("tag", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_config_sets(golden_cfg, tag):
    g = golden_cfg(tag)
    e = g["exponents"]
    fma = _check(e, g["generating_points"], n_points=1)
    # tree form: one FMA per monomial (+ even padding) + one per tree node above the leaves
    assert e.shape[0] <= fma <= 2.2 * e.shape[0]


def test_irregular_sets():
    rng = np.random.default_rng(5)
    gp = rng.uniform(-1, 1, (9, 4))
    for trial in range(20):
        m = int(rng.integers(1, 5))
        N = int(rng.integers(1, 60))
        e = rng.integers(0, 8, (N, m))
        e = np.unique(e, axis=0)
        rng.shuffle(e)                    # unsorted, with holes, non-zero first leaf
        _check(e.astype(np.int64), gp[:, :m], n_points=2, seed=trial)


def test_duplicate_rows_and_many_chunks():
    rng = np.random.default_rng(6)
    gp = rng.uniform(-1, 1, (6, 3))
    e = np.array([[0, 0, 0], [1, 0, 0], [1, 0, 0], [0, 2, 1], [0, 2, 1], [3, 2, 1]], dtype=np.int64)
    _check(e, gp)
    # enough leaves to need several chunks (descriptor capacity 508 words per chunk)
    e = np.array([[0, a, b] for b in range(40) for a in range(40)], dtype=np.int64)
    gp = rng.uniform(-1, 1, (41, 3))
    desc, _, _ = compile_program(e)
    assert desc.shape[0] >= 2
    _check(e, gp, n_points=1)


# ---- tile program (tensor-core tree kernel) -------------------------------------------------
from program_interp import compile_tile_program, run_tile_program  # noqa: E402


def _check_tiles(e, gp, n_points=2, seed=0):
    rng = np.random.default_rng(seed)
    N, m = e.shape
    coeffs = rng.uniform(-1, 1, N)
    xx = rng.uniform(-1, 1, (n_points, m))
    desc, slots, fma = compile_tile_program(e)
    ref = oracle.eval_newton_polynomials(xx, coeffs, e, gp).reshape(-1)
    for p in range(n_points):
        val, fma_run = run_tile_program(desc, slots, xx[p], coeffs, e, gp)
        assert abs(val - ref[p]) <= 1e-12 * max(1.0, np.max(np.abs(ref)))
        assert fma_run == fma
    return fma


@pytest.mark.parametrize("tag", [t for t in SMALL_TAGS if not t.startswith("m1_")])
def test_tile_program_small_sets(golden_small, tag):
    _check_tiles(golden_small[f"{tag}/exponents"], golden_small[f"{tag}/generating_points"])


@pytest.mark.parametrize("tag", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_tile_program_config_sets(golden_cfg, tag):
    g = golden_cfg(tag)
    _check_tiles(g["exponents"], g["generating_points"], n_points=1)
