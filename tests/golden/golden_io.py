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
