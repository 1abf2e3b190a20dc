"""CPU tests of the host-side input producers and index-set predicates (SURVEY.md 8a row A7,
8f rank 2) against arrays the unmodified reference produced (tests/golden/*.npz).  Everything
here is integer or bit-exact float work: comparisons are equalities."""
import re

import numpy as np
import pytest

import golden_io
from minterpy_b200 import generators
from minterpy_b200.polynomials import Grid, MultiIndexSet, is_lexicographically_complete

CFG = {"cfg1": (1, 64, 2.0), "cfg2": (4, 16, 2.0), "cfg3": (6, 10, 1.0), "cfg4": (5, 12, 2.0),
       "cfg5": (3, 32, np.inf)}


def parse_tag(tag):
    m, n, lp = re.fullmatch(r"m(\d+)_n(\d+)_lp(.+)", tag).groups()
    return int(m), int(n), float(lp)


def check_shape(m, n, lp, exponents, gp, poly_degree=None):
    mi = MultiIndexSet.from_degree(m, n, lp)
    assert mi.exponents.dtype == np.int64
    assert np.array_equal(mi.exponents, exponents)
    if poly_degree is not None:
        assert mi.poly_degree == int(poly_degree)
    got = Grid(mi).generating_points
    assert got.shape == gp.shape and np.array_equal(got, gp)      # bit for bit


def test_small_grid_of_the_reference_tests(golden_small):
    tags = sorted({k.split("/")[0] for k in golden_small})
    assert len(tags) == 36
    for tag in tags:
        m, n, lp = parse_tag(tag)
        check_shape(m, n, lp, golden_small[f"{tag}/exponents"],
                    golden_small[f"{tag}/generating_points"])
        nodes = Grid(MultiIndexSet.from_degree(m, n, lp)).unisolvent_nodes
        assert np.array_equal(nodes, golden_small[f"{tag}/unisolvent_nodes"])


@pytest.mark.parametrize("tag", sorted(CFG))
def test_baseline_configs(golden_cfg, tag):
    d = golden_cfg(tag)
    m, n, lp = CFG[tag]
    check_shape(m, n, lp, d["exponents"], d["generating_points"], d["poly_degree"])
    nodes = Grid(MultiIndexSet.from_degree(m, n, lp)).unisolvent_nodes
    assert golden_io.sha256_of(nodes) == str(d["unisolvent_sha"])


def test_poly_degree_quirk_of_config_3():
    # ceil of a float norm: 10.000000000000002 -> 11, and the grid has 12 points per dimension
    mi = MultiIndexSet.from_degree(6, 10, 1.0)
    assert mi.poly_degree == 11
    assert Grid(mi).generating_points.shape == (12, 6)


def test_other_lp_degrees_and_extra_shapes():
    d = golden_io.load("next_rows")
    for tag in d["gen/tags"]:
        m, n, lp = d[f"gen/{tag}/args"]
        check_shape(int(m), int(n), float(lp), d[f"gen/{tag}/exponents"],
                    d[f"gen/{tag}/generating_points"], d[f"gen/{tag}/poly_degree"])
    for n in (0, 1, 2, 3, 10, 33, 64):
        assert np.array_equal(generators.leja_ordered_chebyshev(n), d[f"gen/leja/{n}"])
    x = golden_io.load("shapes_extra")
    for tag in sorted({k.split("/")[0] for k in x}):
        m, n, lp = parse_tag(tag)
        check_shape(m, n, lp, x[f"{tag}/exponents"], x[f"{tag}/generating_points"])


def test_from_degree_argument_rules():
    with pytest.raises(TypeError):
        MultiIndexSet.from_degree(2, 3.0, 2.0)
    with pytest.raises(TypeError):
        MultiIndexSet.from_degree(2.0, 3, 2.0)
    assert MultiIndexSet.from_degree(2, 3).lp_degree == 2.0           # default lp-degree
    assert np.array_equal(MultiIndexSet.from_degree(3, 0, 1.0).exponents, np.zeros((1, 3)))
    with pytest.raises(TypeError):
        Grid(np.zeros((1, 2), dtype=np.int64))


def test_pruned_enumeration_equals_full_enumeration():
    # the reference filters all n^m candidates; the pruned enumeration must keep the same rows
    for (m, n, lp) in [(3, 6, 0.7), (4, 5, 1.0), (3, 7, 2.0), (2, 12, 4.0), (5, 3, 1.3)]:
        axes = np.meshgrid(*[np.arange(n, dtype=np.int64)] * m, indexing="ij")
        full = np.stack([a.reshape(-1) for a in axes], axis=1)
        full = np.vstack((full, np.diag([np.int64(n)] * m)))
        keep = full[np.sum(np.power(full, lp), axis=-1) <= n ** lp]
        keep = keep[np.lexsort(keep.T)]
        assert np.array_equal(generators.get_exponent_matrix(m, n, lp), keep)


def test_completeness_against_reference_answers():
    d = golden_io.load("next_rows")
    seen = set()
    for i in range(int(d["cmp/count"])):
        e = d[f"cmp/{i}/exponents"]
        want = bool(d[f"cmp/{i}/complete"])
        assert is_lexicographically_complete(e) == want
        assert MultiIndexSet(e).is_complete == want
        seen.add(want)
    assert seen == {True, False}


def test_completeness_of_the_baseline_sets_and_holes(golden_cfg):
    for tag in ("cfg2", "cfg4", "cfg5"):
        e = golden_cfg(tag)["exponents"]
        assert is_lexicographically_complete(e)
        # without the zero index every index of degree one lacks its smaller neighbour
        assert not is_lexicographically_complete(np.delete(e, 0, axis=0))
        # row 1 is (1, 0, .., 0); (2, 0, .., 0) needs it
        assert not is_lexicographically_complete(np.delete(e, 1, axis=0))
    # dropping a maximal element keeps the set downward closed
    e = golden_cfg("cfg5")["exponents"]
    assert is_lexicographically_complete(e[:-1])


def test_completeness_rejects_unsorted_input():
    e = MultiIndexSet.from_degree(2, 3, 2.0).exponents
    with pytest.raises(ValueError):
        is_lexicographically_complete(e[::-1])
    with pytest.raises(ValueError):
        is_lexicographically_complete(np.vstack([e, e[-1:]]))     # duplicate row
    with pytest.raises(ValueError):
        is_lexicographically_complete(-np.ones((1, 2), dtype=np.int64))


def test_against_the_imported_reference_when_present():
    from ref_shim import load_reference
    mp = load_reference()
    if mp is None:
        pytest.skip("/root/reference is not on this machine")
    from minterpy.core.utils import is_lexicographically_complete as ref_complete
    rng = np.random.default_rng(5)
    for (m, n, lp) in [(2, 8, 1.0), (3, 5, 2.0), (4, 3, np.inf), (3, 6, 0.5)]:
        ref = mp.MultiIndexSet.from_degree(m, n, lp)
        mine = MultiIndexSet.from_degree(m, n, lp)
        assert np.array_equal(ref.exponents, mine.exponents)
        assert ref.poly_degree == mine.poly_degree
        assert np.array_equal(mp.Grid(ref).generating_points, Grid(mine).generating_points)
        for _ in range(4):
            e = np.delete(ref.exponents, rng.integers(0, len(ref)), axis=0)
            assert ref_complete(e) == is_lexicographically_complete(e)
