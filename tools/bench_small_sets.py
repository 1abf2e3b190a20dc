#!/usr/bin/env python
"""Latency of the tree kernel on SMALL index sets (one or two chunks of program), where a launch
is bound by the per-CTA prologue rather than by arithmetic: BASELINE config 1 (m = 1, n = 64) and
a few low-degree sets, 1e6 points, k = 1, device-resident.
   python tools/bench_small_sets.py            (shipped library)
   MPX_EVAL_VARIANT=18 python tools/run_with_lib.py tools/_bin/libmpx_tuning.so tools/bench_small_sets.py
prints one JSON line per set."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import generators
    P = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1_000_000
    shapes = [(1, 64, 2.0), (1, 16, 2.0), (2, 8, 2.0), (2, 16, 2.0), (3, 6, 2.0), (3, 8, 1.0), (2, 30, 2.0), (2, 60, 1.0)]
    if "--reverse" in sys.argv:
        shapes = shapes[::-1]
    for (m, n, lp) in shapes:
        e = generators.get_exponent_matrix(m, n, lp)
        gp = generators.default_generating_points(m, int(e.max()))
        N = e.shape[0]
        plan = mpb.Plan(e, gp)
        xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
        c = torch.rand((N, 1), dtype=torch.float64, device="cuda") * 2 - 1
        out = torch.empty((P, 1), dtype=torch.float64, device="cuda")
        for _ in range(5):
            plan._eval_torch(xx, c, out=out)
        ts = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            plan._eval_torch(xx, c, out=out)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        us = float(np.median(ts)) * 1e3
        # the same call queued 50 times: device time per call when the host runs ahead, and the
        # host's own cost of issuing it
        import time
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(50):
            plan._eval_torch(xx, c, out=out)
        e1.record()
        host_us = (time.perf_counter() - t0) / 50 * 1e6
        torch.cuda.synchronize()
        queued_us = e0.elapsed_time(e1) / 50 * 1e3
        # the C entry point alone (arguments prepared once): the library's share of the issue cost
        from minterpy_b200 import _lib
        fn = _lib.lib().mpx_eval
        args = (plan._h, xx.data_ptr(), P, xx.stride(0), c.data_ptr(), 1, c.stride(0), out.data_ptr(),
                out.stride(0), torch.cuda.current_stream().cuda_stream)
        t0 = time.perf_counter()
        for _ in range(200):
            fn(*args)
        abi_us = (time.perf_counter() - t0) / 200 * 1e6
        torch.cuda.synchronize()
        print(json.dumps({"set": f"m{m}_n{n}_lp{lp}", "N": N, "points": P,
                          "variant": os.environ.get("MPX_EVAL_VARIANT", "shipped"),
                          "us": round(us, 1), "queued_us": round(queued_us, 1), "host_issue_us": round(host_us, 1), "c_abi_issue_us": round(abi_us, 1), "pm_per_s": P * N / (us * 1e-6),
                          "hbm_gbs": round(P * 8 * (m + 1) / (us * 1e-6) / 1e9, 1)}))


if __name__ == "__main__":
    main()
