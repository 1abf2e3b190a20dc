"""Wider golden vectors for the BASELINE configs: per config >= 4096
random points + ALL unisolvent nodes + the 2^m corners, evaluated by the UNMODIFIED reference;
and the reference's own Newton->Lagrange transformation at all nodes.

Run where /root/reference (or baseline/_ref) exists:  python tests/golden/make_golden_wide.py
Writes tests/golden/wide.npz.  Inputs are re-drawn from stored seeds (sha256-guarded) and the
nodes are rebuilt from the stored index set, so only reference OUTPUTS are stored: for configs
with many polynomials the values of a few coefficient columns (first, second, middle, last).
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import load_reference  # noqa: E402
from golden_io import draw_uniform, load, sha256_of  # noqa: E402

mp = load_reference()
assert mp is not None, "needs the reference"
from minterpy.utils import eval_newton_polynomials  # noqa: E402

N_RANDOM = 4096


def corners(m):
    return np.array([[1.0 if (i >> d) & 1 else -1.0 for d in range(m)] for i in range(2 ** m)])


def stored_columns(k):
    return sorted({0, min(1, k - 1), k // 2, k - 1})


def main():
    out = {}
    for cfg in (1, 2, 3, 4, 5):
        tag = f"cfg{cfg}"
        g = load(tag)
        e, gp = g["exponents"], g["generating_points"]
        N, m = e.shape
        k = int(g["k_eval"])
        t0 = time.time()
        nodes = np.take_along_axis(gp, e, axis=0)
        assert sha256_of(nodes) == str(g["unisolvent_sha"])
        seed = 5000 + cfg
        xr = draw_uniform(seed, (N_RANDOM, m))
        out[f"{tag}/xx_seed"] = np.int64(seed)
        out[f"{tag}/xx_sha"] = np.array(sha256_of(xr))
        xx = np.concatenate([xr, nodes, corners(m)])
        c = draw_uniform(int(g["coeffs_seed"]), (N, k))
        assert sha256_of(c) == str(g["coeffs_sha"])
        cols = stored_columns(k)
        out[f"{tag}/columns"] = np.array(cols, dtype=np.int64)
        cs = np.ascontiguousarray(c[:, cols])
        batch = max(64, int((1 << 30) // (8 * N)))
        ev = eval_newton_polynomials(xx, cs if len(cols) > 1 else cs[:, 0], e, gp, batch_size=batch)
        out[f"{tag}/eval_ref"] = np.asarray(ev).reshape(xx.shape[0], len(cols))
        print(tag, "eval at", xx.shape[0], "points x", len(cols), "columns:", round(time.time() - t0, 1), "s", flush=True)
        # the reference's Newton -> Lagrange transformation itself (barycentric operator), all nodes
        if cfg in (2, 4):
            mi = mp.MultiIndexSet.from_degree(*{2: (4, 16, 2.0), 4: (5, 12, 2.0)}[cfg])
            grid = mp.Grid(mi)
            assert np.array_equal(mi.exponents, e) and np.array_equal(grid.generating_points, gp)
            newt = mp.NewtonPolynomial(mi, cs[:, 0], grid=grid)
            lag = mp.NewtonToLagrange(newt)()
            out[f"{tag}/n2l_ref"] = np.asarray(lag.coeffs).reshape(N)
            print(tag, "NewtonToLagrange:", round(time.time() - t0, 1), "s", flush=True)
    np.savez_compressed(os.path.join(HERE, "wide.npz"), **out)
    print("wrote wide.npz", os.path.getsize(os.path.join(HERE, "wide.npz")) / 1e6, "MB")


if __name__ == "__main__":
    main()
