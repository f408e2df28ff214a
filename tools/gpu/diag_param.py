"""Dump what the parametrizer kernel produced for a few problems of a parity batch (x, L in; t, v, a out) so that the
stage can be re-run on the CPU oracle off the GPU box."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import seacurrent_b200 as scb
out = {}
for mode, W, G, P in (("ref_cubic", 20, 500, 256), ("lau_quintic", 10, 1000, 256)):
    rng = np.random.default_rng(100 + W + G)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.05)
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=P, max_waypoints=W, max_gridpoints=G)
    res = pl.plan(wx=wx, wy=wy, G=G, **lim)
    pl.close()
    for k in ("status", "arclen_total", "T_total", "x", "t", "v", "a", "K_lo", "K_hi"):
        out["%s_%d_%s" % (mode, W, k)] = res[k][..., :16] if res[k].ndim == 2 else res[k][:16]
np.savez_compressed("gpurun_out/diag_param.npz", **out)
print("ok")
