"""Host-side producers of the hot path's inputs: index sets and generating points.

These arrays define the inputs of every kernel bit for bit, so they are restated here with
the same floating-point expressions the reference evaluates and
pinned against arrays the reference produced (tests/golden/*.npz, tests/test_generators.py):

* ``get_exponent_matrix``  ~ ``src/minterpy/core/utils.py:44-92``: the index set
  ``{e : sum_i e_i^p <= n^p}`` drawn from ``[0, n)^m`` plus the ``n * unit`` vectors, sorted
  with the last column as the slowest key.  The reference enumerates all ``n^m`` candidates;
  here the enumeration is pruned level by level (with slack) and only the survivors go
  through the reference's exact float test, which keeps m = 8 feasible.
* ``poly_degree_of``       ~ ``core/utils.py:23-42`` with ``lp_norm`` ``utils.py:18-36``
  (the ceil of a float norm: (6, 10, 1.0) gives 11, and the grid follows it).
* ``leja_ordered_chebyshev`` ~ ``utils.py:132-174`` (greedy Leja ordering of the
  Chebyshev-Lobatto points, ties resolved towards the LAST maximiser, products accumulated in
  the order the points were chosen).
* ``generating_points_from_values`` ~ ``core/grid.py:27-34`` (tile, negate columns 0, 2, ..).
"""
from math import ceil

import numpy as np

DEFAULT_LP_DEG = 2.0      # src/minterpy/global_settings.py


def verify_lp_degree(lp_degree):
    """``core/verification.py`` (verify_lp_deg): None -> default, must be positive."""
    if lp_degree is None:
        return DEFAULT_LP_DEG
    lp = float(lp_degree)
    if not lp > 0.0:
        raise ValueError("the lp-degree must be positive")
    return lp


def lp_norm(arr, p, axis=None):
    arr = np.asarray(arr)
    a = np.abs(arr).max()
    if a == 0.0:
        return 0.0
    return a * np.linalg.norm(arr / a, p, axis)


def poly_degree_of(exponents, lp_degree):
    return int(ceil(np.max(lp_norm(exponents, lp_degree, axis=1))))


def _pruned_candidates(m, n, lp):
    """All e in [0, n)^m whose power sum can still pass the test, as an (C x m) int64 array.
    Built from the last dimension down; a partial sum above the bound (with slack for the
    float rounding of any summation order) cannot recover because every term is >= 0."""
    bound = float(n) ** lp * (1.0 + 1e-9) + 1e-300
    pw = np.power(np.arange(n, dtype=np.float64), lp)
    cand = np.zeros((1, 0), dtype=np.int64)
    part = np.zeros(1)
    for _ in range(m):
        s = part[:, None] + pw[None, :]
        keep = s <= bound
        rows, vals = np.nonzero(keep)
        cand = np.concatenate([vals[:, None].astype(np.int64), cand[rows]], axis=1)
        part = s[keep]
    return cand


def get_exponent_matrix(spatial_dimension, poly_degree, lp_degree):
    m, n = int(spatial_dimension), poly_degree
    if isinstance(n, float):
        if not n.is_integer():
            raise ValueError(f"poly_degree needs to be a whole number! <{n} given.")
        n = int(n)
    if n == 0:
        return np.zeros((1, m), dtype=np.int64)
    if lp_degree == np.inf:
        axes = np.meshgrid(*[np.arange(n + 1, dtype=np.int64)] * m, indexing="ij")
        chosen = np.stack([a.reshape(-1) for a in axes], axis=1)
    else:
        inner = _pruned_candidates(m, n, lp_degree)
        cand = np.vstack((inner, np.diag([np.int64(n)] * m)))
        # the reference's own test, same expression, on the surviving candidates
        cond = np.sum(np.power(cand, lp_degree), axis=-1) <= n ** lp_degree
        chosen = cand[cond]
    return np.ascontiguousarray(chosen[np.lexsort(chosen.T)])


def chebyshev_lobatto(count):
    """``utils.py:100-129``: `count` points cos(i*pi/(count-1)); the one-point request
    returns the two end points, as the reference does."""
    if count == 0:
        return np.zeros(1)
    if count == 1:
        return np.array([-1.0, 1.0])
    return np.cos(np.arange(count, dtype=np.float64) * np.pi / (count - 1))


def leja_ordered_chebyshev(n):
    """(n+1) x 1 array.  Start from point 0 of the reversed Chebyshev set; then repeatedly take
    the remaining point with the largest product of distances to the chosen ones."""
    n = int(n)
    pts = chebyshev_lobatto(n + 1)[::-1]
    remaining = np.arange(1, n + 1)
    chosen = [0]
    dist = np.ones(n)            # running product per remaining point, chosen-order factors
    for _ in range(n):
        dist = dist * (pts[chosen[-1]] - pts[remaining])
        mag = np.abs(dist)
        # the reference scans upwards with `>=`: the last maximiser wins
        jj = len(mag) - 1 - int(np.argmax(mag[::-1]))
        chosen.append(int(remaining[jj]))
        remaining = np.delete(remaining, jj)
        dist = np.delete(dist, jj)
    return pts[np.asarray(chosen)].reshape(n + 1, 1).copy()


def generating_points_from_values(spatial_dimension, generating_values):
    gp = np.tile(np.asarray(generating_values, dtype=np.float64).reshape(-1, 1),
                 (1, spatial_dimension))
    gp[:, ::2] *= -1
    return gp


def default_generating_points(spatial_dimension, poly_degree):
    return generating_points_from_values(spatial_dimension, leja_ordered_chebyshev(poly_degree))
