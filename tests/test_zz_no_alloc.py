"""mpx_eval_noalloc: the evaluation entry point for callers that cannot allocate device memory
(all memory already given to their own pool, or no tolerance for cudaMalloc's synchronisation
inside the loop).  It must return what mpx_eval returns.

This file sorts last on purpose: if it fails, a run that stops at the first failure has
already executed every other test.  It passes alone in a fresh process as well as after the
rest of the suite (the fixed scratch array is shared by all calls, so the second call below
also checks that stale contents of the first do not leak into the result).
"""
import ctypes

import numpy as np
import pytest

import golden_io

pytestmark = pytest.mark.gpu

EVAL_RTOL = 1e-13          # the tolerance of tests/test_gpu_parity.py


def _call(fn, plan, xx, coeffs):
    import torch
    P, k = xx.shape[0], coeffs.shape[1]
    out = torch.full((P, k), float("nan"), dtype=torch.float64, device="cuda")
    rc = fn(plan._h, xx.data_ptr(), P, xx.stride(0), coeffs.data_ptr(), k, coeffs.stride(0),
            out.data_ptr(), out.stride(0), torch.cuda.current_stream().cuda_stream)
    return rc, out


def test_noalloc_matches_eval_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    lib = _lib.lib()
    g = golden_io.load("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    plan = mpb.Plan(e, gp)
    gen = torch.Generator(device="cuda").manual_seed(99)
    # calls in a row with different coefficients, polynomial counts and point counts
    for P, k in ((5000, 3), (1777, 1), (3001, 12), (2003, 30)):     # k = 30 takes the tensor-core path
        xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
        c = torch.rand((N, k), generator=gen, dtype=torch.float64, device="cuda") * 2 - 1
        rc, ref = _call(lib.mpx_eval, plan, xx, c)
        _lib.check(rc)
        scratch_before = plan.info(_lib.INFO_SCRATCH_BYTES)
        rc, got = _call(lib.mpx_eval_noalloc, plan, xx, c)
        _lib.check(rc)
        torch.cuda.synchronize()
        assert plan.info(_lib.INFO_SCRATCH_BYTES) == scratch_before     # plan scratch untouched
        scale = ref.abs().max().item()
        assert (got - ref).abs().max().item() <= EVAL_RTOL * scale
        assert not torch.isnan(got).any()


def test_noalloc_rejects_oversized_request():
    import torch
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    lib = _lib.lib()
    g = golden_io.load("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    plan = mpb.Plan(e, gp)
    k = 64                              # 24 809 x 64 packed coefficients: 13.5 MB > 8 MiB
    xx = torch.zeros((4, e.shape[1]), dtype=torch.float64, device="cuda")
    c = torch.zeros((e.shape[0], k), dtype=torch.float64, device="cuda")
    rc, out = _call(lib.mpx_eval_noalloc, plan, xx, c)
    assert rc == _lib.MPX_ERR_UNSUPPORTED
    assert "fixed scratch" in _lib.last_error()
    torch.cuda.synchronize()
    assert torch.isnan(out).all()                        # nothing was launched
    rc, _ = _call(lib.mpx_eval_noalloc, plan, xx, c[:, :32].contiguous())    # 7.2 MB: fits
    assert rc == _lib.MPX_OK
