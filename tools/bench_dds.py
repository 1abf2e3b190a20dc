#!/usr/bin/env python
"""Time the divided-difference kernel (K3) on a BASELINE shape with k value columns.
Prints one JSON line: achieved = algorithmic bytes (2*N*k*8: one read + one write of the value
array) / CUDA-event time, against the measured HBM peak of MEASURED_PEAKS.json."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg4")
    ap.add_argument("--k", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--identity", action="store_true",
                    help="value array = the first k columns of the identity (what array_repr_full transforms): mostly zeros")
    ap.add_argument("--hint", action="store_true", help="pass MPX_DDS_MANY_ZEROS (mpx_dds_ex)")
    ap.add_argument("--mode", type=int, default=0, help="test hook: 0 auto, 1 fused slices, 2 one column per lane")
    args = ap.parse_args()
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    g = golden_io.load(args.workload)
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.get_plan(e, gp)
    gen = torch.Generator(device="cuda").manual_seed(1)
    vals = torch.rand((N, args.k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
    if args.identity:
        vals = torch.eye(N, args.k, dtype=torch.float64, device="cuda")
    work = vals.clone()
    lib = _lib.lib()
    lib.mpx_debug_set_dds_group(args.mode)
    stream = torch.cuda.current_stream().cuda_stream
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # > L2 (126 MB)

    def once():
        work.copy_(vals)
        flush.zero_()                   # evict the value array from L2 between iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.mpx_dds_ex(plan._h, work.data_ptr(), args.k, work.stride(0),
                                  _lib.MPX_DDS_MANY_ZEROS if args.hint else 0, stream))
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    for _ in range(3):
        once()
    ts = [once() for _ in range(args.iters)]
    ms = float(np.median(ts))
    # the same transformation through the NumPy API: upload, kernels, download inside the timing
    import time
    host_in = vals.cpu().numpy()
    plan.dds(host_in)
    hs = []
    for _ in range(5):
        t0 = time.perf_counter()
        plan.dds(host_in)
        hs.append((time.perf_counter() - t0) * 1e3)
    e2e_ms = float(np.median(hs))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0)
    algo = 2.0 * N * args.k * 8
    updates = plan.info(_lib.INFO_DDS_UPDATES) * args.k
    print(json.dumps({"bench": "dds", "workload": args.workload, "N": int(N), "m": int(m),
                      "k": args.k, "values": "identity columns" if args.identity else "uniform(-1, 1)", "many_zeros_hint": bool(args.hint), "ms": ms, "ms_min": float(min(ts)),
                      "coefficients_per_s": N * args.k / (ms * 1e-3),
                      "updates_per_s": updates / (ms * 1e-3),
                      "roofline": {"bound": "hbm", "achieved": algo / (ms * 1e-3) / 1e9,
                                   "peak": hbm, "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / hbm,
                                   "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650"},
                      "e2e": {"ms": e2e_ms, "coefficients_per_s": N * args.k / (e2e_ms * 1e-3),
                              "h2d_bytes": int(N * args.k * 8), "d2h_bytes": int(N * args.k * 8),
                              "api": "minterpy_b200.Plan.dds(numpy) -> mpx_dds_host"},
                      "l2": "value array evicted from L2 before each timed call"}))


if __name__ == "__main__":
    main()
