"""One rank of the multi-GPU tests (launched by torch.distributed.run from tests/test_gpu_multi.py):
NCCL process group, this rank's GPU, the CUDA library -- points sharded for evaluation, value
columns sharded for dds, parameters broadcast, results gathered; everything compared with the
single-GPU result computed on the same rank."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist
    import golden_io
    import minterpy_b200 as mpb
    from minterpy_b200 import distributed as mpd
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    try:
        for tag, P, k_eval, k_dds in (("cfg2", 200_003, 1, 64), ("cfg4", 20_001, 5, 256)):
            g = golden_io.load(tag)
            e, gp = g["exponents"], g["generating_points"]
            N, m = e.shape
            plan = mpb.get_plan(e, gp, device=local)
            gen = torch.Generator(device=dev).manual_seed(77)            # same data on every rank
            xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device=dev) * 2 - 1
            coeffs = torch.rand((N, k_eval), generator=gen, dtype=torch.float64, device=dev)
            if rank != 0:
                coeffs = coeffs * 0 - 1.0                                # rank 0's must win
            mpd.broadcast_parameters([coeffs], src=0)
            full = mpd.eval_sharded(plan.eval, xx, coeffs, gather=True)
            want = plan.eval(xx, coeffs)
            assert full.shape == (P, k_eval) and torch.equal(full, want), tag
            a, b = mpd.shard_bounds(P, rank, world)
            assert torch.equal(mpd.eval_sharded(plan.eval, xx, coeffs, gather=False), want[a:b])
            vals = torch.rand((N, k_dds), generator=gen, dtype=torch.float64, device=dev) * 2 - 1
            got = mpd.dds_sharded(plan.dds, vals, gather=True)
            assert torch.equal(got, plan.dds(vals)), tag                 # bit-exact, any split
        dist.barrier()
        if rank == 0:
            print(f"DIST_OK world={world}")
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
