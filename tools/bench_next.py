#!/usr/bin/env python
"""Timings of the callers either side of the hot path, one JSON line
each: index-set completeness check and generation (host), derivative evaluation,
NewtonPolynomial.diff, the Lagrange-basis regression matrix (k = N columns through the
tensor-core kernel) and Lagrange multiplication.  Device times are CUDA-event medians after
3 warm-ups; host times are wall-clock medians."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402


def dev_ms(fn, iters=5):
    import torch
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def host_ms(fn, iters=5):
    fn()
    ts = []
    for _ in range(iters):
        t = time.perf_counter()
        fn()
        ts.append((time.perf_counter() - t) * 1e3)
    return float(np.median(ts))


def main():
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    from minterpy_b200 import api, generators
    from minterpy_b200.polynomials import is_lexicographically_complete
    from minterpy_b200.regression import compute_regression_matrix

    dmma_peak = 37.1        # TFLOP/s, tools/fp64_peak.cu on this pool's B200 (profiles/r01_fp64_peak.jsonl)

    # ---- host: completeness and generation -------------------------------------------------
    for (m, n, lp) in [(4, 16, 2.0), (5, 12, 2.0), (8, 8, 2.0)]:
        ms_gen = host_ms(lambda: generators.get_exponent_matrix(m, n, lp), 3)
        e = generators.get_exponent_matrix(m, n, lp)
        ms_cmp = host_ms(lambda: is_lexicographically_complete(e), 3)
        print(json.dumps({"bench": "index_set_host", "m": m, "n": n, "lp": lp, "N": int(len(e)),
                          "generate_ms": ms_gen, "is_complete_ms": ms_cmp,
                          "note": "reference is_lexicographically_complete: 25 s at (4,16,2.0), "
                                  "169 s at (5,12,2.0) (SURVEY.md 8f)"}))

    # ---- derivative evaluation ---------------------------------------------------------------
    for tag, P, k, orders in (("cfg2", 1_000_000, 1, (1, 0, 0, 0)), ("cfg2", 1_000_000, 1, (0, 1, 2, 1)),
                              ("cfg2", 1_000_000, 8, (1, 0, 2, 0)), ("cfg5", 1_000_000, 1, (2, 0, 1)),
                              ("cfg3", 1_000_000, 64, (1, 0, 0, 0, 0, 1))):
        g = golden_io.load(tag)
        e, gp = g["exponents"], g["generating_points"]
        N, m = e.shape
        plan = mpb.get_plan(e, gp)
        xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
        c = torch.rand((N, k), dtype=torch.float64, device="cuda") * 2 - 1
        o = np.array(orders)
        ms = dev_ms(lambda: plan.eval_deriv(xx, c, o))
        ms0 = dev_ms(lambda: plan.eval(xx, c))
        print(json.dumps({"bench": "deriv_eval", "workload": tag, "P": P, "k": k, "orders": list(orders),
                          "ms": ms, "pm_per_s": P * N / (ms * 1e-3),
                          "tensor_tflops": 2.0 * P * N * max(k, 8) / (ms * 1e-3) / 1e12,
                          "plain_eval_ms": ms0,
                          "kernel": ("shared-memory-table tensor kernel (mpx_eval_mma.cu)"
                                     if k >= plan.info(_lib.INFO_MMA_MIN_K) > 0 else
                                     "tree kernel with differentiated tables (mpx_eval_deriv.cu)")}))

    # ---- NewtonPolynomial.diff (derivative at the nodes + divided differences) -------------
    for tag in ("cfg2", "cfg4"):
        g = golden_io.load(tag)
        e, gp = g["exponents"], g["generating_points"]
        mi = mpb.MultiIndexSet(e)
        grid = mpb.Grid(mi, gp)
        c = torch.rand((len(mi),), dtype=torch.float64, device="cuda")
        poly = mpb.NewtonPolynomial(mi, c, grid=grid)
        o = np.zeros(e.shape[1], dtype=np.int64)
        o[0] = 1
        poly.diff(o)
        ms = host_ms(lambda: (poly.diff(o), torch.cuda.synchronize()), 5)
        print(json.dumps({"bench": "newton_diff", "workload": tag, "N": int(len(mi)), "wall_ms": ms,
                          "includes": "host copy of the node coordinates, derivative kernel, dds kernel"}))

    # ---- regression matrix in the Lagrange basis: k = N columns -----------------------------
    g = golden_io.load("cfg3")
    e, gp = g["exponents"], g["generating_points"]
    mi = mpb.MultiIndexSet(e, lp_degree=1.0)
    grid = mpb.Grid(mi, gp)
    basis = mpb.LagrangePolynomial(mi, None, grid=grid)
    N = len(mi)
    op = mpb.LagrangeToNewton(basis).transformation_operator
    l2n_ms = dev_ms(lambda: op.full_on_device(), 3)      # identity + dds of N columns, on device
    l2n_t = op.full_on_device()
    plan = mpb.get_plan(e, gp)
    for P in (20_000, 100_000):
        xx = torch.rand((P, 6), dtype=torch.float64, device="cuda") * 2 - 1
        out = torch.empty((P, N), dtype=torch.float64, device="cuda")
        ms = dev_ms(lambda: plan._eval_torch(xx, l2n_t, out=out), 3)
        tf = 2.0 * P * N * N / (ms * 1e-3) / 1e12
        print(json.dumps({"bench": "regression_matrix_lagrange", "workload": "cfg3", "P": P, "N": N,
                          "k": N, "ms": ms, "l2n_matrix_ms": l2n_ms,
                          "roofline": {"bound": "tensor", "achieved": tf, "peak": dmma_peak,
                                       "unit": "TFLOP/s", "frac": tf / dmma_peak,
                                       "peak_source": "tools/fp64_peak.cu (DMMA m8n8k4)"},
                          "out_bytes": P * N * 8}))
    del l2n_t, out

    # ---- Lagrange multiplication ----------------------------------------------------------------
    for (m, n, lp) in [(3, 12, 2.0), (4, 8, 1.0)]:
        mi = mpb.MultiIndexSet.from_degree(m, n, lp)
        grid = mpb.Grid(mi)
        fa = torch.as_tensor(np.cos(grid.unisolvent_nodes.sum(1)), device="cuda")
        fb = torch.as_tensor(np.sin(grid.unisolvent_nodes @ np.arange(1, m + 1)), device="cuda")
        a = mpb.LagrangePolynomial(mi, fa, grid=grid)
        b = mpb.LagrangePolynomial(mi, fb, grid=grid)
        res = a * b
        ms = host_ms(lambda: (a * b, torch.cuda.synchronize()), 3)
        print(json.dumps({"bench": "lagrange_mul", "m": m, "n": n, "lp": lp, "N_in": int(len(mi)),
                          "N_out": int(len(res.multi_index)), "wall_ms": ms,
                          "includes": "2 x dds, result index set + grid on the host, plan build, "
                                      "2 x evaluation at the result nodes"}))


if __name__ == "__main__":
    main()
