#!/usr/bin/env python
"""Evaluation throughput (k = 1, device-resident) of one shape from tests/golden/shapes_extra.npz
or a BASELINE config, for comparing the two tree kernels:
   python tools/bench_shape.py m3_n24_lp2.0 [points] [tmma: 0 auto | 1 force | 2 never]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402


def main():
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    tag = sys.argv[1]
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
    if tag.startswith("cfg"):
        g = golden_io.load(tag)
        e, gp = g["exponents"], g["generating_points"]
    else:
        g = golden_io.load("shapes_extra")
        e, gp = g[f"{tag}/exponents"], g[f"{tag}/generating_points"]
    N, m = e.shape
    if len(sys.argv) > 3:
        _lib.lib().mpx_debug_set_tmma(int(sys.argv[3]))
    plan = mpb.Plan(e, gp)
    xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
    c = torch.rand((N, 1), dtype=torch.float64, device="cuda") * 2 - 1
    out = torch.empty((P, 1), dtype=torch.float64, device="cuda")
    for _ in range(3):
        plan._eval_torch(xx, c, out=out)
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        plan._eval_torch(xx, c, out=out)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    kern = {0: "dfma-tree", 1: "tensor-tree", 2: "smem-table", 3: "horner-1d"}[plan.info(_lib.INFO_SMALLK_KERNEL)]
    print(f"{tag} N={N} P={P} kernel={kern} {ms:.2f} ms  {P * N / ms / 1e9:.3f}e12 pm/s")


if __name__ == "__main__":
    main()
