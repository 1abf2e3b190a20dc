"""world_size-2 `gloo` tests of the multi-GPU host logic on CPU:
contiguous sharding of points / value columns, parameter broadcast, result gather.  The
per-shard compute is done by the oracle here -- what is under test is the plumbing."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from minterpy_b200 import distributed as mpd


def test_shard_bounds():
    assert [mpd.shard_bounds(10, r, 3) for r in range(3)] == [(0, 3), (3, 6), (6, 10)]
    assert mpd.shard_sizes(7, 2) == [3, 4]
    assert mpd.shard_bounds(2, 1, 4) == (0, 0)
    assert mpd.shard_bounds(2, 3, 4) == (0, 2)
    assert sum(mpd.shard_sizes(10**9, 8)) == 10**9
    with pytest.raises(ValueError):
        mpd.shard_bounds(5, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, P, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "tests", "golden")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import golden_io
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = golden_io.load("small")
        tag = "m3_n4_lp2.0"
        e, gp = g[f"{tag}/exponents"], g[f"{tag}/generating_points"]
        N, m = e.shape
        rng = np.random.default_rng(5)                       # same points on every rank
        xx = torch.from_numpy(rng.uniform(-1, 1, (P, m)))
        coeffs = torch.from_numpy(np.random.default_rng(100 + rank).uniform(-1, 1, (N, 3)))
        mpd.broadcast_parameters([coeffs], src=0)            # rank 0's coefficients win

        def evaluate(x, c):
            return torch.from_numpy(oracle.eval_newton_polynomials(x.numpy(), c.numpy(), e, gp))

        full = mpd.eval_sharded(evaluate, xx, coeffs, gather=True)
        ref_c = np.random.default_rng(100).uniform(-1, 1, (N, 3))
        ref = oracle.eval_newton_polynomials(xx.numpy(), ref_c, e, gp)
        assert full.shape == (P, 3)
        assert np.array_equal(full.numpy(), ref)
        local = mpd.eval_sharded(evaluate, xx, coeffs, gather=False)
        a, b = mpd.shard_bounds(P, rank, world)
        assert np.array_equal(local.numpy(), ref[a:b])
        # dds: columns split over the ranks
        vals = torch.from_numpy(np.random.default_rng(7).uniform(-1, 1, (N, 5)))
        out = mpd.dds_sharded(lambda v: torch.from_numpy(oracle.dds(v.numpy(), e, gp)), vals)
        assert np.array_equal(out.numpy(), oracle.dds(vals.numpy(), e, gp))
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("P", [64, 33, 1])
def test_two_rank_gloo(tmp_path, P):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, P, str(tmp_path)), nprocs=2, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]
