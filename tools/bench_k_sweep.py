#!/usr/bin/env python
"""Evaluation time against the number of polynomials k (device-resident inputs, CUDA events,
median of 3 after 2 warm-ups): where the tree kernel run once per polynomial and the
many-polynomial tensor kernel cross.  usage: bench_k_sweep.py <cfgN> <points>;
MPX_EVAL_MMA_MIN_K=9999 forces the tree kernel, =1 the tensor kernel (the variable is only
read by a library built with `python -m minterpy_b200.build --tuning`)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import golden_io  # noqa: E402
import minterpy_b200 as mpb  # noqa: E402

tag=sys.argv[1]; P=int(sys.argv[2])
g=golden_io.load(tag); e,gp=g['exponents'],g['generating_points']
plan=mpb.get_plan(e,gp)
xx=torch.rand((P,e.shape[1]),dtype=torch.float64,device='cuda')*2-1
for k in (2,3,4,5,6,7,8,10,12,14,16,20,24):
    c=torch.rand((e.shape[0],k),dtype=torch.float64,device='cuda')
    out=torch.empty((P,k),dtype=torch.float64,device='cuda')
    for _ in range(2): plan._eval_torch(xx,c,out=out)
    ts=[]
    for _ in range(3):
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(); plan._eval_torch(xx,c,out=out); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    print(tag,'mink',os.environ.get('MPX_EVAL_MMA_MIN_K','8'),'k',k,'%.2f ms'%np.median(ts))
