"""Experiment (GPU): do two contexts (two streams) running alternate C3 batches overlap the FP64-bound path / arc
kernels of one batch with the HBM/latency-bound toppra / param kernels of the other?"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import seacurrent_b200 as scb

w = scb.workloads.config_c3()
G = w["G"]; P, W = w["wx"].shape
dev = torch.device("cuda", 0)
opt = scb.default_opt(move_waypoints=1)
inputs = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev, dtype=torch.float64) for k, v in w.items() if k != "G"}
pls = [scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=W, max_gridpoints=G) for _ in range(4)]
bat = [pl.make_device_batch(inputs, G=G, opt=opt) for pl in pls]

def run(n_ctx, steps):
    for i in range(4):
        pls[i % n_ctx].run_device(bat[i % n_ctx][0], bat[i % n_ctx][1], sync=False)
    for pl in pls: pl.synchronize()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        pls[i % n_ctx].run_device(bat[i % n_ctx][0], bat[i % n_ctx][1], sync=False)
    for pl in pls: pl.synchronize()
    return (time.perf_counter() - t0) / steps * 1e3

for rep in range(2):
    for n in (1, 2):
        print("%d ctx   %.3f ms/step" % (n, run(n, 48)))
