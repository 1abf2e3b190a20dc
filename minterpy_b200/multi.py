"""Evaluation over several GPUs from ONE process: one host thread per GPU.

BASELINE.json's north_star (3) / SURVEY.md section 8e: query points are independent, so a
NumPy caller's (P, m) array is cut into contiguous shards, one per GPU; every GPU gets the
(small) index set, generating points and coefficients, evaluates its shard, and writes its
values straight into its slice of ONE page-locked result array -- there is nothing to exchange
and nothing to gather afterwards.  ``ctypes`` releases the GIL during the library call, so the
threads really run side by side (`mpx_eval_host` pipelines upload, kernel and download per GPU).

CUDA tensors are sharded with peer copies over NVLink and gathered back to the tensor's device.
The multi-PROCESS route (torchrun, NCCL) is in ``distributed.py``; both must work.
"""
import threading

import numpy as np

from . import _lib
from .distributed import shard_bounds
from .plan import _host_result, _is_torch, get_plan


def visible_devices():
    import torch
    if not torch.cuda.is_available():
        raise _lib.MinterpyB200Error("no CUDA device available; minterpy_b200 has no CPU fallback")
    return list(range(torch.cuda.device_count()))


def _run_threads(jobs):
    errors = []

    def guarded(fn):
        try:
            fn()
        except BaseException as ex:       # noqa: BLE001 -- re-raised in the caller's thread
            errors.append(ex)

    threads = [threading.Thread(target=guarded, args=(job,)) for job in jobs]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]


def eval_multi_gpu(xx, coeffs, exponents, generating_points, devices=None):
    """xx (P, m), coeffs (N, k) -> (P, k), the points sharded over `devices` (default: every
    visible GPU; a device may be listed twice, its shards then take turns).  NumPy in -> NumPy
    out (page-locked result), CUDA tensors in -> CUDA tensor out on the device of `xx`."""
    devices = visible_devices() if devices is None else [int(d) for d in devices]
    if not devices:
        raise ValueError("no devices given")
    world = len(devices)
    if _is_torch(xx) or _is_torch(coeffs):
        return _eval_multi_torch(xx, coeffs, exponents, generating_points, devices)
    xx = np.ascontiguousarray(xx, dtype=np.float64)
    coeffs = np.ascontiguousarray(coeffs, dtype=np.float64)
    P, k = xx.shape[0], coeffs.shape[1]
    out = _host_result((P, k))
    plans = [get_plan(exponents, generating_points, device=d) for d in devices]   # main thread

    def job(r):
        a, b = shard_bounds(P, r, world)
        if b > a:
            plans[r].eval_into(xx[a:b], coeffs, out[a:b])

    _run_threads([lambda r=r: job(r) for r in range(world)])
    return out


def _eval_multi_torch(xx, coeffs, exponents, generating_points, devices):
    import torch
    home = xx.device if _is_torch(xx) and xx.is_cuda else torch.device("cuda", devices[0])
    xx = torch.as_tensor(xx, dtype=torch.float64, device=home)
    coeffs = torch.as_tensor(coeffs, dtype=torch.float64, device=home)
    world = len(devices)
    P, k = xx.shape[0], coeffs.shape[1]
    out = torch.empty((P, k), dtype=torch.float64, device=home)
    plans = [get_plan(exponents, generating_points, device=d) for d in devices]
    torch.cuda.synchronize(home)              # the shards are read on other devices' streams

    def job(r):
        a, b = shard_bounds(P, r, world)
        if b <= a:
            return
        dev = torch.device("cuda", devices[r])
        with torch.cuda.device(dev):
            x_r = xx[a:b].to(dev, non_blocking=True)          # peer copy over NVLink
            c_r = coeffs.to(dev, non_blocking=True)           # "broadcast" of the parameters
            y_r = plans[r].eval(x_r, c_r)
            out[a:b].copy_(y_r, non_blocking=True)            # gather: peer copy back
            torch.cuda.current_stream(dev).synchronize()

    _run_threads([lambda r=r: job(r) for r in range(world)])
    return out
