"""Diagnostic (GPU): the worst FP64 omega point of test_batch_matches_oracle_fp64[lau_quintic-*]."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import seacurrent_b200 as scb
from oracle import oracle

for (W, G, P) in ((10, 1000, 256), (20, 500, 256)):
    rng = np.random.default_rng(100 + W + G)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.05)
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=4096, max_waypoints=64, max_gridpoints=1001, max_samples=0)
    res = pl.plan(wx=wx, wy=wy, G=G, **lim)
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=1, **lim)
    ok = ref["status"] == 0
    w = ref["omega"]; g = res["omega"]
    rel = np.abs(g - w) / np.maximum(np.abs(w), 1e-3 * np.abs(w).max(axis=0, keepdims=True))
    rel[:, ~ok] = 0
    for _ in range(4):
        i, p = np.unravel_index(np.argmax(rel), rel.shape)
        v = ref["v"][i, p]
        print("W=%d G=%d: p=%d i=%d rel=%.3g got=%.12g want=%.12g v=%.6g kappa=%.6g L=%.4g  kappa*64eps*L=%.3g" % (
            W, G, p, i, rel[i, p], g[i, p], w[i, p], v, w[i, p] / v, ref["arclen_total"][p],
            abs(w[i, p] / v) * 64 * 2.2e-16 * ref["arclen_total"][p]))
        print("   omega around:", w[max(i - 2, 0):i + 3, p], " px err %.3g py err %.3g" % (
            abs(res["px"][i, p] - ref["px"][i, p]), abs(res["py"][i, p] - ref["py"][i, p])))
        rel[i, p] = 0
    pl.close()
