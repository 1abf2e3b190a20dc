"""Where does the time of Plan.dds(numpy) go?  (lab bench)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io, torch
import minterpy_b200 as mpb
from minterpy_b200 import _lib, plan as planmod
g = golden_io.load("cfg4")
plan = mpb.get_plan(g["exponents"], g["generating_points"])
N, k = plan.N, 256
vals = np.random.default_rng(0).uniform(-1, 1, (N, k))
lib = _lib.lib()
def t(fn, n=6):
    fn(); ts = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); ts.append((time.perf_counter() - t0) * 1e3)
    return round(float(np.median(ts)), 3)
print("Plan.dds(numpy)            ", t(lambda: plan.dds(vals)))
print("_host_result alloc          ", t(lambda: planmod._host_result((N, k))))
out_pin = torch.empty((N, k), dtype=torch.float64, pin_memory=True).numpy()
in_pin = torch.empty((N, k), dtype=torch.float64, pin_memory=True).numpy(); in_pin[:] = vals
out_page = np.empty((N, k))
call = lambda i, o: _lib.check(lib.mpx_dds_host(plan._h, i.ctypes.data, k, k, o.ctypes.data, k))
print("C call pinned in, pinned out", t(lambda: call(in_pin, out_pin)))
print("C call pageable in, pinned out", t(lambda: call(vals, out_pin)))
print("C call pinned in, pageable out", t(lambda: call(in_pin, out_page)))
print("C call pageable in, pageable out", t(lambda: call(vals, out_page)))
dv = torch.as_tensor(vals, device="cuda")
torch.cuda.synchronize()
def dev():
    plan.dds(dv); torch.cuda.synchronize()
print("device tensor (clone + kernels)", t(dev))
