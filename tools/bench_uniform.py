"""Tree kernels with the record stream as a kernel parameter (uniform control flow) against the
same kernels reading it from the shared-memory ring (test hook), device-resident, k = 1."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch, golden_io
import minterpy_b200 as mpb
from minterpy_b200 import _lib
lib = _lib.lib()
for tag, P, tm in (("cfg2", 20_000_000, 2), ("cfg5", 20_000_000, 2), ("cfg4", 4_000_000, 2), ("cfg3", 2_000_000, 2)):
    g = golden_io.load(tag)
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    lib.mpx_debug_set_tmma(tm)         # 2: DFMA tree kernel, 1: tensor-core tree kernel
    plan = mpb.Plan(e, gp)
    xx = torch.rand((P, m), dtype=torch.float64, device="cuda") * 2 - 1
    c = torch.rand((N, 1), dtype=torch.float64, device="cuda")
    res = {}
    for mode in (0, 2):
        lib.mpx_debug_set_eval_uniform(mode)
        for _ in range(3): y = plan.eval(xx, c)
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); y = plan.eval(xx, c); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res[mode] = (float(np.median(ts)), y.clone())
    lib.mpx_debug_set_eval_uniform(0)
    same = torch.equal(res[0][1], res[2][1])
    print(tag, "tensor-tree" if tm == 1 else "dfma-tree", "uniform %.3f ms  %.3e pm/s | ring %.3f ms  %.3e pm/s | identical %s" % (
        res[0][0], P * N / res[0][0] * 1e3, res[2][0], P * N / res[2][0] * 1e3, same), flush=True)
