"""Diagnostic (GPU): where does FP32-mode omega leave the tolerance of test_fp32_mode_within_1e4?"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import seacurrent_b200 as scb
from oracle import oracle
import test_gpu_parity as T

rng = np.random.default_rng(21)
P, W, G = 256, 10, 500
wx, wy = scb.workloads.random_walk_paths(rng, P, W)
wx = wx.astype(np.float32).astype(np.float64); wy = wy.astype(np.float32).astype(np.float64)
lim = scb.workloads.random_limits(rng, P, slow_fraction=0.0)
for mode in ("ref_cubic", "lau_quintic"):
    pl = scb.BatchPlanner(0, "fp32", mode, max_problems=4096, max_waypoints=64, max_gridpoints=1001, max_samples=0)
    res = pl.plan(wx=wx, wy=wy, G=G, **lim)
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0 if mode == "ref_cubic" else 1, **lim)
    ok = ref["status"] == 0
    for probes in (1, 3):
        tol = T.fp32_omega_tol(oracle, ref, wx, wy, G, lim, mode, n_probe=probes)[:, ok]
        w = ref["omega"][:, ok]; g = res["omega"][:, ok].astype(np.float64)
        r = np.abs(g - w) / tol
        idx = np.argwhere(r > 1.0)
        print(mode, "probes", probes, "violations", len(idx), "of", r.size, "max", r.max(), "problems", len(set(idx[:, 1])))
    if "cost" in res and "cost" in ref:
        print("  cost rel err max", np.max(np.abs(res["cost"][ok] - ref["cost"][ok]) / np.abs(ref["cost"][ok])))
    worst = np.argsort(-r.max(axis=0))[:6]
    for p in worst:
        i = int(np.argmax(r[:, p]))
        print("  p=%d i=%d ratio=%.2f got=%.6g want=%.6g tol=%.3g  ctrl err %.3g" % (
            p, i, r[i, p], g[i, p], w[i, p], tol[i, p], np.max(np.abs(res["ctrl_x"][ok][p] - ref["ctrl_x"][ok][p]))))
    pl.close()
