#!/usr/bin/env python
"""One process, all visible GPUs: eval_newton_polynomials(numpy, devices="all") at a BASELINE
config -- what a NumPy caller of the reference gets from an 8-GPU box without torchrun.
Prints one JSON line per device count (1, 2, 4, ... visible)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--points", type=int, default=100_000_000)
    ap.add_argument("--pageable", action="store_true")
    args = ap.parse_args()
    import torch
    import minterpy_b200 as mpb
    g = golden_io.load(args.workload)
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    P = args.points
    c = golden_io.draw_uniform(7, (N,))
    if args.pageable:
        xx = np.empty((P, m))
    else:
        xx = torch.empty((P, m), dtype=torch.float64, pin_memory=True).numpy()
    rng = np.random.default_rng(1)
    for a in range(0, P, 1 << 22):
        xx[a:a + (1 << 22)] = rng.uniform(-1, 1, (min(1 << 22, P - a), m))
    n = torch.cuda.device_count()
    counts = [d for d in (1, 2, 4, 8) if d <= n]
    ref = None
    for d in counts:
        devs = list(range(d))
        for _ in range(2):
            y = mpb.eval_newton_polynomials(xx, c, e, gp, devices=devs)
        ts = []
        for _ in range(3):
            t = time.perf_counter()
            y = mpb.eval_newton_polynomials(xx, c, e, gp, devices=devs)
            ts.append((time.perf_counter() - t) * 1e3)
        if ref is None:
            ref = y[:: max(1, P // 4096)].copy()
        same = bool(np.array_equal(ref, y[:: max(1, P // 4096)]))
        ms = float(np.median(ts))
        print(json.dumps({"bench": "eval_multi_gpu", "workload": args.workload, "points": P,
                          "devices": d, "host_memory": "pageable" if args.pageable else "page-locked",
                          "ms": ms, "pm_per_s": P * N / (ms * 1e-3),
                          "identical_to_one_gpu": same,
                          "api": "minterpy_b200.eval_newton_polynomials(numpy, devices=[...])"}),
              flush=True)


if __name__ == "__main__":
    main()
