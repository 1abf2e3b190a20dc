"""Multi-GPU paths: one process with one thread per GPU
(`eval_multi_gpu`, `eval_newton_polynomials(..., devices=...)`) and one process per GPU over NCCL
(`distributed.eval_sharded`, `dds_sharded`).  On a single-GPU box the thread path is exercised
with the same device listed several times; the NCCL test needs two GPUs."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mpb():
    import torch
    assert torch.cuda.is_available(), "the gpu tests need a CUDA device"
    import minterpy_b200
    return minterpy_b200


def rel_err(got, ref):
    return float(np.max(np.abs(np.asarray(got) - np.asarray(ref))) / np.max(np.abs(ref)))


def _device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0], [0, 0, 0]]
    if n >= 2:
        lists += [list(range(n)), [1, 0]]
    return lists


def test_threads_over_gpus_numpy(mpb, golden_cfg):
    g = golden_cfg("cfg2")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    rng = np.random.default_rng(21)
    for P, k in ((100_003, 1), (7, 3), (2, 2)):
        xx = rng.uniform(-1, 1, (P, m))
        c = rng.uniform(-1, 1, (N, k))
        want = mpb.get_plan(e, gp, device=0).eval(xx, c)
        for devices in _device_lists():
            got = mpb.eval_multi_gpu(xx, c, e, gp, devices=devices)
            assert got.shape == (P, k) and np.array_equal(got, want), devices
    # through the reference-shaped entry point: squeeze rules apply after the gather
    xx = rng.uniform(-1, 1, (5000, m))
    c1 = rng.uniform(-1, 1, N)
    y = mpb.eval_newton_polynomials(xx, c1, e, gp, devices="all")
    assert y.shape == (5000,)
    assert rel_err(y[:50], oracle.eval_newton_polynomials(xx[:50], c1[:, None], e, gp)[:, 0]) <= 1e-13
    with pytest.raises(ValueError):
        mpb.eval_multi_gpu(xx, c1[:, None], e, gp, devices=[])


def test_threads_over_gpus_tensors(mpb, golden_cfg):
    import torch
    g = golden_cfg("cfg5")
    e, gp = g["exponents"], g["generating_points"]
    N, m = e.shape
    gen = torch.Generator(device="cuda:0").manual_seed(4)
    xx = torch.rand((300_001, m), generator=gen, dtype=torch.float64, device="cuda:0") * 2 - 1
    c = torch.rand((N, 2), generator=gen, dtype=torch.float64, device="cuda:0")
    want = mpb.get_plan(e, gp, device=0).eval(xx, c)
    for devices in _device_lists():
        got = mpb.eval_multi_gpu(xx, c, e, gp, devices=devices)
        assert got.device == xx.device and torch.equal(got, want), devices


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_one_process_per_gpu_over_nccl():
    """torchrun, 2 ranks (all GPUs up to 4), NCCL: evaluation sharded by points, dds by value
    columns, both gathered and compared bit for bit with the single-GPU result."""
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = min(n, 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "dist_worker_gpu.py")]
    out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, (out.stdout + out.stderr)[-3000:]
    assert f"DIST_OK world={world}" in out.stdout
