#!/usr/bin/env python
"""Time the two write/read-bound side kernels against the measured HBM peak:
  * mpx_monomials: the (P x N) Newton-monomial matrix, 8 B written per point*monomial;
  * mpx_apply_dense: out = T @ c, 8 B read per matrix entry.
One JSON line each."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402


def timed(fn, iters=10):
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


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="", help="comma list of: fill, monomials, dense")
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    only = set(filter(None, args.only.split(",")))

    def want(name):
        return not only or name in only
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    lib = _lib.lib()
    try:
        hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
        src = "MEASURED_PEAKS.json"
    except (OSError, KeyError):
        hbm, src = 6650.0, "fallback"
    stream = torch.cuda.current_stream().cuda_stream
    # context: what a pure write stream reaches on this GPU (torch fill of 8 GB)
    if want("fill"):
        buf = torch.empty(1 << 30, dtype=torch.float64, device="cuda")
        ms = timed(lambda: buf.fill_(1.0))
        print(json.dumps({"bench": "write_only_fill", "bytes": 8 << 30, "ms": ms,
                          "GBps": (8 << 30) / (ms * 1e-3) / 1e9, "hbm_copy_peak": hbm}))
        del buf
    for tag, P, pad in (("cfg2", 40000, 0), ("cfg2", 40000, 16), ("cfg3", 100000, 0)) if want("monomials") else ():
        g = golden_io.load(tag)
        e, gp = g["exponents"], g["generating_points"]
        N, m = e.shape
        plan = mpb.get_plan(e, gp)
        xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
        ld = N if not pad else (N + pad - 1) // pad * pad      # padded rows: 128-byte aligned
        out = torch.empty((P, ld), dtype=torch.float64, device="cuda")     # > L2

        def run():
            _lib.check(lib.mpx_monomials(plan._h, xx.data_ptr(), P, m, out.data_ptr(), ld, 0, stream))
        ms = timed(run, args.iters)
        gbs = P * N * 8 / (ms * 1e-3) / 1e9
        print(json.dumps({"bench": "monomials", "workload": tag, "P": P, "N": int(N), "ld_out": int(ld), "ms": ms,
                          "pm_per_s": P * N / (ms * 1e-3),
                          "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                                       "frac": gbs / hbm, "peak_source": src},
                          "l2": "output (8 B x P x N) larger than L2"}))
        del out
    cases = ((24809, 1, 0), (24809, 4, 0), (24809, 8, 0), (24809, 16, 0), (24809, 32, 0), (24808, 1, 0),
             (24808, 16, 0), (24809, 1, 1), (24809, 16, 1))
    for rows, k, lower in cases if want("dense") else ():
        T = torch.rand((rows, rows), dtype=torch.float64, device="cuda")
        if lower:
            T = T.tril()
        c = torch.rand((rows, k), dtype=torch.float64, device="cuda")
        o = torch.empty((rows, k), dtype=torch.float64, device="cuda")

        def run():
            _lib.check(lib.mpx_apply_dense_ex(0, T.data_ptr(), rows, rows, rows, c.data_ptr(), k, k,
                                              o.data_ptr(), k, lower, stream))
        ms = timed(run, args.iters)
        needed = rows * (rows + 1) // 2 * 8 if lower else rows * rows * 8     # bytes that must be read
        gbs = needed / (ms * 1e-3) / 1e9
        ref = (T[-64:] @ c)
        err = ((o[-64:] - ref).abs().max() / ref.abs().max()).item()
        print(json.dumps({"bench": "apply_dense", "rows": rows, "k": k, "lower_triangular": bool(lower),
                          "ms": ms, "rel_err_vs_torch": err,
                          "tflops": 2.0 * needed / 8 * k / (ms * 1e-3) / 1e12,
                          "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s",
                                       "frac": gbs / hbm, "peak_source": src,
                                       "algorithmic_bytes": needed},
                          "l2": "matrix (4.9 GB) larger than L2"}))
        del T

if __name__ == "__main__":
    main()
