#!/usr/bin/env python
"""Benchmark of the Newton-evaluation hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--scaling weak|strong]
                  [--impl reference]

One "step" = one pass of the evaluation kernel over one batch of synthetic query points.
At N = 1 the workload is BASELINE config 2 (m=4, n=16, lp=2, 1e8 points).  With N > 1 (torchrun,
one rank per GPU):
  --scaling weak   (default) every rank evaluates its own shard of the workload's size (no
                   data-path collective; the coefficients are broadcast once over NCCL before
                   the timed region);
  --scaling strong ONE job -- the workload's points (cfg5: 1e9, as BASELINE config 5 is written)
                   split over the ranks.

value  = point x monomial evaluations per second of the whole job, inputs resident in HBM (CUDA
         events on the launching stream, max over ranks);
e2e    = the same metric through the public API with HOST buffers: page-locked NumPy arrays
         (H2D of the points and D2H of the values inside the timed region); `e2e_pageable` the
         same with an ordinary NumPy array.  Strong scaling with N > 1: the whole job per step --
         NCCL broadcast of the coefficients, upload of the shard, evaluation, all_gather of the
         result shards, download;
roofline = executed FP64 flops of the evaluation kernel / its CUDA-event duration, against
         the DFMA (or DMMA) rate measured on this box by tools/fp64_peak.cu (MEASURED_PEAKS.json
         has no FP64 figure) -- nominal 37.2 TFLOP/s is reported beside it;
cpu_baseline = the UNMODIFIED reference (baseline/_ref, tools/make_baseline_ref.py) on a bounded
         sample of the same workload: its eval_newton_polynomials as-is in one process and over
         point shards in one forked process per host core; the C port of the oracle beside it,
         and instead of it if the reference does not import (the line then says why).
`--impl reference` times only that CPU path (rank 0; the other ranks exit).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import golden_io  # noqa: E402

WORKLOADS = {
    # tag: (description, points per GPU, k)
    "cfg1": ("m=1 n=64 lp=2 (README quickstart), 1e6 points", 1_000_000, 1),
    "cfg2": ("m=4 n=16 lp=2, 1e8 points in [-1,1]^4", 100_000_000, 1),
    "cfg3": ("m=6 n=10 lp=1, k=64 polynomials, 1e7 points", 10_000_000, 64),
    "cfg4": ("m=5 n=12 lp=2, k=256 polynomials, 1e6 points", 1_000_000, 256),
    "cfg5": ("m=3 n=32 lp=inf, 1.25e8 points per GPU (1e9 over 8)", 125_000_000, 1),
}
NOMINAL_FP64_TFLOPS = 37.2   # 148 SM x 64 DFMA/clk x 2 x 1.965 GHz


def load_workload(tag):
    g = golden_io.load(tag)
    return g["exponents"], g["generating_points"]


def model_flops_per_point(e, k):
    """F/P = N (1 + 2k) + 2 sum_i max_exp_i."""
    return e.shape[0] * (1 + 2 * k) + 2 * int(e.max(axis=0).sum())


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.lines, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ts, line in self.lines:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_fp64_peak():
    """Best DFMA rate of tools/fp64_peak.cu on this GPU (TFLOP/s), or None."""
    exe = os.path.join(ROOT, "tools", "_bin", "fp64_peak")
    if not os.path.exists(exe):
        return None, None
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout
    except (OSError, subprocess.TimeoutExpired):
        return None, None
    dfma, dmma = [], []
    for line in out.splitlines():
        try:
            d = json.loads(line)
        except ValueError:
            continue
        if d.get("bench", "").startswith("dfma_ch"):
            dfma.append(d["tflops"])
        if d.get("bench", "").startswith("dmma"):
            dmma.append(d["tflops"])
    return (max(dfma) if dfma else None), (max(dmma) if dmma else None)


def cpu_baseline(e, gp, k, target_seconds=12.0, seed=1):
    """Oracle port on all host threads over a bounded sample of the workload."""
    import oracle
    cores = oracle.num_threads()
    N, m = e.shape
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, (N, k))
    probe = max(cores * 8, 64)
    xx = rng.uniform(-1, 1, (probe, m))
    t = time.perf_counter()
    oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=cores)
    rate = probe / max(time.perf_counter() - t, 1e-6)
    P = int(max(probe, min(rate * target_seconds, 5_000_000)))
    xx = rng.uniform(-1, 1, (P, m))
    t = time.perf_counter()
    oracle.eval_newton_polynomials(xx, c, e, gp, nthreads=cores)
    dt = time.perf_counter() - t
    # the reference's Numba stage is single-threaded: also time one thread on a 3 s sample
    P1 = int(max(64, min(rate / cores * 3.0, P)))
    t = time.perf_counter()
    oracle.eval_newton_polynomials(xx[:P1], c, e, gp, nthreads=1)
    dt1 = time.perf_counter() - t
    return {"value": P * N / dt, "unit": "point*monomials/s", "cores": cores, "kind": "port",
            "sample": f"{P} points of the workload ({dt:.1f} s), oracle/newton_oracle.c on "
                      f"{cores} threads", "single_core_value": P1 * N / dt1,
            "seconds": dt, "points": P}


# ---- the unmodified reference on the host cores -------------------------------------------
# baseline/_ref holds the pip-installed reference (tools/make_baseline_ref.py; git-ignored, it
# travels to the GPU box).  What is timed is its own public callable,
# minterpy.utils.eval_newton_polynomials(xx, c, e, gp, batch_size=B)  (utils.py:247-343), as-is
# in one process, and process-parallel over point shards on all host cores.
_REF = {}


def _ref_eval(lo, hi):
    g = _REF
    return g["fn"](g["xx"][lo:hi], g["c"], g["e"], g["gp"], batch_size=g["batch"])


def _ref_worker(span):
    from threadpoolctl import threadpool_limits
    with threadpool_limits(1):             # one BLAS thread per worker process
        return float(_ref_eval(*span).sum())


def reference_available():
    from ref_shim import reference_location
    return reference_location() is not None


def reference_cpu_baseline(e, gp, k, target_seconds=12.0, seed=1, cores=None):
    """The reference itself on a bounded sample: `value` = process-parallel over all host
    cores, `single_core_value` = the call as-is in one process.  JIT warm-up excluded."""
    import multiprocessing as mp_
    from ref_shim import load_reference
    ref = load_reference()
    fn = ref.utils.eval_newton_polynomials
    cores = cores or os.cpu_count() or 1
    N, m = e.shape
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, (N, k))
    c = c[:, 0] if k == 1 else c
    batch = max(64, int((128 << 20) // (8 * N)))      # batch * N * 8 B <= 128 MB per process
    _REF.update(fn=fn, c=c, e=e, gp=gp, batch=batch, xx=rng.uniform(-1, 1, (64, m)))
    _ref_eval(0, 64)                                                   # JIT compile
    probe = 2048
    _REF["xx"] = rng.uniform(-1, 1, (probe, m))
    t = time.perf_counter()
    _ref_eval(0, probe)
    rate1 = probe / max(time.perf_counter() - t, 1e-6)                 # points/s, one process
    P1 = int(max(probe, rate1 * 3.0))
    P = int(max(cores * 256, min(rate1 * cores * target_seconds * 0.8, 20_000_000)))
    _REF["xx"] = xx = rng.uniform(-1, 1, (max(P, P1), m))
    t = time.perf_counter()
    _ref_eval(0, P1)
    dt1 = time.perf_counter() - t
    # shards: 4 per core so that the tail is short; children are forked (they inherit the
    # arrays and the compiled kernels; nothing in this process has touched CUDA)
    nshard = cores * 4
    bounds = np.linspace(0, P, nshard + 1).astype(np.int64)
    spans = [(int(a), int(b)) for a, b in zip(bounds[:-1], bounds[1:]) if b > a]
    with mp_.get_context("fork").Pool(cores) as pool:
        pool.map(_ref_worker, [(0, 64)] * cores)                       # start-up, untimed
        t = time.perf_counter()
        pool.map(_ref_worker, spans, chunksize=1)
        dt = time.perf_counter() - t
    del xx
    return {"value": P * N / dt, "unit": "point*monomials/s", "cores": cores, "kind": "reference",
            "sample": f"{P} points of the workload ({dt:.1f} s): minterpy.utils."
                      f"eval_newton_polynomials (unmodified, baseline/_ref), batch_size={batch}, "
                      f"point shards over {cores} forked processes",
            "single_core_value": P1 * N / dt1,
            "single_core_sample": f"{P1} points, the call as-is in one process ({dt1:.1f} s)",
            "seconds": dt, "points": P}


def run_reference(args):
    """`--impl reference`: the CPU path only (rank 0; other ranks exit).  The unmodified
    reference where baseline/_ref imports, else the oracle port (the import error is kept in
    the line)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    e, gp = load_workload(args.workload)
    desc, _, k = WORKLOADS[args.workload]
    N = e.shape[0]
    per_step = args.cpu_seconds or max(4.0, min(30.0, 150.0 / max(args.steps + args.warmup, 1)))
    runner, why_port = cpu_baseline, None
    if args.cpu_kind != "port":
        try:
            if not reference_available():
                raise ImportError("baseline/_ref is absent (tools/make_baseline_ref.py)")
            reference_cpu_baseline(e, gp, k, target_seconds=0.5)
            runner = reference_cpu_baseline
        except Exception as exc:       # noqa: BLE001 -- any import/JIT failure: fall back to the port
            why_port = f"{type(exc).__name__}: {exc}"
    for _ in range(args.warmup):
        runner(e, gp, k, target_seconds=1.0)
    pts, secs, last = 0, 0.0, None
    for s in range(args.steps):
        last = runner(e, gp, k, target_seconds=per_step, seed=100 + s)
        pts += last["points"]
        secs += last["seconds"]
    value = pts * N / secs
    base = {"value": value, "unit": "point*monomials/s", "cores": last["cores"],
            "kind": last["kind"], "sample": f"{pts} points over {args.steps} steps; " + last["sample"],
            "single_core_value": last["single_core_value"]}
    if why_port:
        base["reference_unavailable"] = why_port
    if last["kind"] == "reference" and not args.no_port:
        port = cpu_baseline(e, gp, k, target_seconds=min(per_step, 6.0))
        base["port_value"] = port["value"]
        base["port_single_core_value"] = port["single_core_value"]
    line = {"impl": "reference", "metric": "newton_evals_per_sec", "value": value,
            "unit": "point*monomials/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, N, k, desc, e.shape[1], args.gpus),
            "cpu_baseline": base,
            "e2e": {"value": value, "unit": "point*monomials/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def job_points(args, world):
    """(points of this job per GPU -- the largest shard --, points of the whole job)."""
    _, P, _ = WORKLOADS[args.workload]
    if args.scaling == "strong":
        if args.workload == "cfg5":
            P = 1_000_000_000                # BASELINE config 5 as written: 1e9 points in all
        if args.points:
            P = args.points
        return P - (world - 1) * (P // world), P
    if args.points:
        P = args.points
    return P, world * P


def workload_config(args, N, k, desc, m, world):
    """The `config` object, identical in both arms (the driver compares them)."""
    per_gpu, total = job_points(args, world)
    return {"workload": f"{args.workload}: {desc}", "monomials": int(N), "k": k,
            "points_per_gpu": int(per_gpu), "points_total": int(total),
            "parallelism": f"points sharded x{world}",
            "l2": ("inputs (points) larger than L2, streamed once per step"
                   if per_gpu * 8 * (m + k) >= (256 << 20) else
                   "256 MB buffer written between timed steps (inputs fit L2); "
                   "timed region = sum of per-step CUDA-event times")}


def cpu_baseline_subprocess(workload, seconds=12.0):
    """The CPU figure for the main line: `bench.py --impl reference` in a child process (it
    forks worker processes, which this process -- CUDA initialised -- must not do)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", workload,
           "--steps", "1", "--warmup", "0", "--cpu-seconds", str(seconds)]
    env = {k: v for k, v in os.environ.items()
           if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE")}
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env).stdout
        for line in reversed(out.splitlines()):
            if line.startswith("{"):
                return json.loads(line)["cpu_baseline"]
    except (OSError, ValueError, KeyError, subprocess.TimeoutExpired):
        pass
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: the workload's points on EVERY GPU; strong: the workload's points "
                         "(cfg5: 1e9) split over the GPUs, broadcast and gather inside e2e")
    ap.add_argument("--points", type=int, default=None,
                    help="points per GPU (weak) / of the whole job (strong); default: workload's")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=None, help="CPU sample size per step")
    ap.add_argument("--cpu-kind", default="auto", choices=["auto", "port"])
    ap.add_argument("--no-port", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import minterpy_b200 as mpb
    from minterpy_b200 import _lib
    from minterpy_b200 import distributed as mpd

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    desc, _, k = WORKLOADS[args.workload]
    strong = args.scaling == "strong"
    P, P_job = job_points(args, world)
    if strong:
        a, b = mpd.shard_bounds(P_job, rank, world)
        P = b - a                            # this rank's shard
    cfg = int(args.workload[3:])
    e, gp = load_workload(args.workload)
    N, m = e.shape
    plan = mpb.get_plan(e, gp, device=local)

    # synthetic inputs: points per rank from Philox on the device; coefficients drawn on rank 0
    # and broadcast once over NCCL (weak: outside the timed region; strong: inside e2e)
    gen = torch.Generator(device=dev).manual_seed(20261106 + 1000 * cfg + rank)
    xx = torch.rand((P, m), generator=gen, dtype=torch.float64, device=dev) * 2 - 1
    coeffs = torch.from_numpy(golden_io.draw_uniform(7 + cfg, (N, k))).to(dev)
    if world > 1:
        if rank != 0:
            coeffs.zero_()
        mpd.broadcast_parameters([coeffs], src=0)
    out = torch.empty((P, k), dtype=torch.float64, device=dev)

    def step():
        plan._eval_torch(xx, coeffs, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    in_bytes = P * 8 * (m + k)
    flush = None
    if in_bytes < (256 << 20):
        # inputs + outputs would stay in the 126 MB L2: evict them between timed steps
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    stops = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    t0 = time.time()
    for i in range(args.steps):
        if flush is not None:
            flush.zero_()
        starts[i].record()
        step()
        stops[i].record()
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    step_ms = [starts[i].elapsed_time(stops[i]) for i in range(args.steps)]
    # without a flush the steps run back to back: first start to last stop is the timed region
    total_ms = starts[0].elapsed_time(stops[-1]) if flush is None else float(sum(step_ms))
    total_ms = max_over_ranks(total_ms)
    ms_per_step = total_ms / args.steps
    value = P_job * N / (ms_per_step * 1e-3)

    # ---- end to end through the public NumPy API, HOST buffers ---------------------------
    # pinned: the caller's points are page-locked (torch pin_memory); pageable: an ordinary
    # NumPy array (np.random), which the library stages through its own pinned ring.
    # strong scaling: the step also holds the NCCL broadcast of the coefficients before and the
    # all_gather of the result shards after the evaluation (one job: broadcast, shard, gather).
    e2e = None
    e2e_pageable = None
    if not args.no_e2e:
        ch = coeffs.cpu().numpy()
        ch1 = ch[:, 0] if k == 1 else ch
        e2e_steps = max(1, min(args.steps, 3))
        def e2e_numpy(x_np):
            times = []
            for it in range(2 + e2e_steps):            # 2 warm-ups (staging buffers, pinned result)
                barrier()
                t = time.perf_counter()
                y = mpb.eval_newton_polynomials(x_np, ch1, e, gp)
                dt = time.perf_counter() - t
                assert y.shape[0] == x_np.shape[0]
                del y
                if it >= 2:
                    times.append(dt * 1e3)
            dt = max_over_ranks(float(np.mean(times)) * 1e-3)
            return {"value": P_job * N / dt, "unit": "point*monomials/s",
                    "h2d_bytes_per_step": int(P * m * 8 + N * k * 8),
                    "d2h_bytes_per_step": int(P * k * 8),
                    "ms_per_step": dt * 1e3, "step_ms": times,
                    "api": "minterpy_b200.eval_newton_polynomials(numpy)"}

        def e2e_one_job(xh, yh):
            """Strong scaling, N > 1: ONE job -- coefficients broadcast from rank 0 over NCCL,
            every rank uploads and evaluates its shard of the points (in pieces, the upload of
            piece i+1 on a copy stream under the evaluation of piece i), the result shards are
            all-gathered over NVLink while every rank reads its own shard back to the host."""
            times = []
            c1 = coeffs[:, 0] if k == 1 else coeffs
            pieces = 8
            cuts = [P * i // pieces for i in range(pieces + 1)]
            copy_stream, d2h_stream = torch.cuda.Stream(), torch.cuda.Stream()
            main = torch.cuda.current_stream()
            y = torch.empty((P, k), dtype=torch.float64, device=dev)
            for it in range(2 + e2e_steps):
                barrier()
                t = time.perf_counter()
                dist.broadcast(coeffs, src=0)
                ready = []
                with torch.cuda.stream(copy_stream):
                    for i in range(pieces):
                        xx[cuts[i]:cuts[i + 1]].copy_(xh[cuts[i]:cuts[i + 1]], non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(copy_stream)
                        ready.append(ev)
                for i in range(pieces):
                    main.wait_event(ready[i])
                    y[cuts[i]:cuts[i + 1]] = mpb.eval_newton_polynomials(
                        xx[cuts[i]:cuts[i + 1]], c1, e, gp).reshape(-1, k)
                done = torch.cuda.Event()
                done.record(main)
                with torch.cuda.stream(d2h_stream):
                    d2h_stream.wait_event(done)
                    yh.copy_(y, non_blocking=True)
                full = mpd.gather_values(y, P_job)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t
                assert full.shape[0] == P_job
                del full
                if it >= 2:
                    times.append(dt * 1e3)
            dt = max_over_ranks(float(np.mean(times)) * 1e-3)
            return {"value": P_job * N / dt, "unit": "point*monomials/s",
                    "h2d_bytes_per_step": int(P * m * 8), "d2h_bytes_per_step": int(P * k * 8),
                    "ms_per_step": dt * 1e3, "step_ms": times,
                    "api": "minterpy_b200.eval_newton_polynomials(torch) + distributed.gather_values",
                    "includes": "NCCL broadcast of the coefficients, H2D of the point shard "
                                "(8 pieces, overlapped with the evaluation), evaluation, all_gather "
                                "of the result shards, D2H of the shard"}

        xh = torch.empty((P, m), dtype=torch.float64, pin_memory=True)
        xh.copy_(xx)
        if strong and world > 1:
            yh = torch.empty((P, k), dtype=torch.float64, pin_memory=True)
            e2e = e2e_one_job(xh, yh)
            del yh
        else:
            e2e = e2e_numpy(xh.numpy())
        e2e["host_memory"] = "page-locked input (torch pin_memory), page-locked result"
        del xh
        if not (strong and world > 1):
            xp = np.empty((P, m))                       # ordinary pageable memory, same points
            blk = 1 << 22
            for a in range(0, P, blk):
                xp[a:a + blk] = xx[a:a + blk].cpu().numpy()
            e2e_pageable = e2e_numpy(xp)
            e2e_pageable["host_memory"] = ("pageable input (plain numpy array), result array "
                                           "allocated by the API")
            del xp

    gather_ms = None
    if world > 1:
        full = mpd.gather_values(out, world * P)          # NCCL all_gather of the result shards
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        full = mpd.gather_values(out, world * P)
        g1.record()
        barrier()
        gather_ms = max_over_ranks(g0.elapsed_time(g1))
        del full

    if rank == 0:
        fma = plan.info(_lib.INFO_EVAL_FMA)
        other = plan.info(_lib.INFO_EVAL_OTHER_FP64)
        model = model_flops_per_point(e, k)
        peak_dfma, peak_dmma = measured_fp64_peak()
        mma_min_k = plan.info(_lib.INFO_MMA_MIN_K)     # 0: never; else the smallest k on DMMA
        many_poly = mma_min_k > 0 and k >= mma_min_k
        tensor_path = many_poly                 # mpx_eval_mma.cu (DMMA), else the tree kernel
        if tensor_path:
            # per point: every monomial (padded to 32) is formed with m-1 DMUL and contracted
            # against k (padded to the CTA tile) coefficients on the tensor cores
            n_pad = -(-N // 32) * 32
            tile = 8 if k <= 8 else 16 if k <= 16 else 32 if k <= 32 else 64
            k_pad = -(-k // tile) * tile
            executed = n_pad * (2 * k_pad + (m - 1)) + 2 * int(e.max(axis=0).sum())
            fma = n_pad * k_pad
            peak = peak_dmma or NOMINAL_FP64_TFLOPS
        elif plan.info(_lib.INFO_SMALLK_KERNEL) == 1:
            # tensor-core tree kernel (mpx_eval_tmma.cu): leaf sums as DMMA tiles
            tensor_path = True
            fma = plan.info(_lib.INFO_TMMA_FMA) * k
            executed = 2 * fma + other * k
            peak = peak_dmma or NOMINAL_FP64_TFLOPS
        else:
            executed = (2 * fma + other) * k          # FP64 flops per point actually issued
            fma = fma * k
            peak = peak_dfma or NOMINAL_FP64_TFLOPS
        algo = min(model, executed)
        kernel_ms = float(np.median(step_ms))
        # DRAM bytes per launch from the committed ncu --set full capture of this workload
        # (profiles/traffic.json), scaled to the points of this run; null if never captured
        traffic = None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
            if tj and not tensor_path:
                traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) * P / tj["points"]
        except (OSError, ValueError, KeyError):
            traffic = None
        achieved = algo * P / (kernel_ms * 1e-3) / 1e12
        horner = (not many_poly) and plan.info(_lib.INFO_SMALLK_KERNEL) == 3
        roofline = {"bound": "tensor" if tensor_path else "fp64", "achieved": achieved,
                    "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                    "algorithmic_bytes": P * 8 * (m + k),
                    "peak_source": (("measured DMMA (mma.sync f64) rate" if tensor_path else
                                     "measured DFMA rate") + ", tools/fp64_peak.cu on this GPU"
                                    if peak_dfma else "nominal (fp64_peak tool missing)"),
                    "nominal_peak": NOMINAL_FP64_TFLOPS, "frac_of_nominal": achieved / NOMINAL_FP64_TFLOPS,
                    "measured_dmma_peak": peak_dmma,
                    "flops_per_point_model": model, "flops_per_point_executed": executed,
                    "fma_per_point": fma,
                    "kernel": ("eval_mma_kernel" if many_poly else
                               "horner1d_kernel" if horner else
                               "eval_tmma_kernel" if tensor_path else
                               "eval_tree_kernel_u" if plan.info(_lib.INFO_UNIFORM_PROGRAM) else
                               "eval_tree_kernel"),
                    "kernel_ms": kernel_ms,
                    "hbm_gbs": P * 8 * (m + k) / (kernel_ms * 1e-3) / 1e9}
        # (rank 0 at N = 1 only: with N > 1 the driver's reference arm carries the CPU figure)
        cpu = None if (args.no_cpu or world > 1) else cpu_baseline_subprocess(args.workload, args.cpu_seconds or 12.0)
        cfgd = workload_config(args, N, k, desc, m, world)
        line = {"metric": "newton_evals_per_sec", "value": value, "unit": "point*monomials/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": cfgd,
                "clocks": clocks, "e2e": e2e, "e2e_pageable": e2e_pageable,
                "gpu_launches": 2 * args.steps,
                "roofline": roofline, "cpu_baseline": cpu, "gather_ms": gather_ms,
                "step_ms": step_ms}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
