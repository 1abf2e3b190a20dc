"""Helpers shared by make_golden.py (writer) and the tests (readers)."""
import hashlib
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def sha256_of(a):
    a = np.ascontiguousarray(a)
    return hashlib.sha256(a.tobytes()).hexdigest()


def draw_uniform(seed, shape):
    """uniform(-1, 1) doubles from PCG64(seed): the same stream on every machine."""
    return np.random.default_rng(int(seed)).uniform(-1.0, 1.0, shape)


def load(tag):
    """Load tests/golden/<tag>.npz as a dict; exponents come back as int64."""
    with np.load(os.path.join(HERE, f"{tag}.npz")) as z:
        d = {k: z[k] for k in z.files}
    for k in list(d):
        if k.endswith("exponents"):
            d[k] = d[k].astype(np.int64)
    return d


def seeded(d, seed_key, sha_key, shape):
    """Re-draw a seeded input and check it is the array the golden output was made from."""
    a = draw_uniform(d[seed_key], shape)
    if sha256_of(a) != str(d[sha_key]):
        raise RuntimeError(f"seeded input {seed_key} does not reproduce (numpy RNG drift)")
    return a


def corners(m):
    return np.array([[1.0 if (i >> d) & 1 else -1.0 for d in range(m)] for i in range(2 ** m)])


def wide_case(tag):
    """The wide golden case of a BASELINE config (make_golden_wide.py): 4096 seeded random
    points + ALL unisolvent nodes + the 2^m corners, the stored coefficient columns, and the
    reference's values there.  Returns (points, coefficient columns, eval_ref, n2l_ref or None)."""
    g, w = load(tag), load("wide")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    nodes = np.take_along_axis(gp, e, axis=0)
    if sha256_of(nodes) != str(g["unisolvent_sha"]):
        raise RuntimeError("unisolvent nodes do not reproduce")
    xr = draw_uniform(w[f"{tag}/xx_seed"], (4096, m))
    if sha256_of(xr) != str(w[f"{tag}/xx_sha"]):
        raise RuntimeError("seeded points do not reproduce (numpy RNG drift)")
    xx = np.concatenate([xr, nodes, corners(m)])
    c = seeded(g, "coeffs_seed", "coeffs_sha", (N, int(g["k_eval"])))
    cols = w[f"{tag}/columns"]
    return xx, np.ascontiguousarray(c[:, cols]), w[f"{tag}/eval_ref"], w.get(f"{tag}/n2l_ref")
