"""Diagnostic (GPU): per-field worst ratio of the const-accel parity test."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import seacurrent_b200 as scb
from oracle import oracle
import test_const_accel as T

wx, wy, lim = T._problem(P=160, W=10, seed=13)
G, S_cap = 300, 2048
for mode, sm in (("ref_cubic", 0), ("lau_quintic", 1)):
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=256, max_waypoints=16, max_gridpoints=512, max_samples=S_cap)
    res = pl.plan(wx=wx, wy=wy, G=G, S_cap=S_cap, param_mode=1, outputs=scb.ALL_OUTPUTS + ("s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega"), **lim)
    pl.close()
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=sm, param_mode=1, S_cap=S_cap, **lim)
    ok = ref["status"] == 0
    print(mode, "status equal", np.array_equal(res["status"], ref["status"]), "n_samples equal", np.array_equal(res["n_samples"], ref["n_samples"]),
          np.flatnonzero(res["n_samples"] != ref["n_samples"]), "T rel", np.max(np.abs(res["T_total"][ok] / ref["T_total"][ok] - 1)))
    for n in ("x", "t", "v", "a", "omega", "s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega"):
        g, w = res[n][:, ok], ref[n][:, ok]
        scale = np.abs(w).max(axis=0, keepdims=True)
        r = np.abs(g - w) / np.maximum(np.abs(w), 1e-3 * scale)
        i, p = np.unravel_index(np.argmax(r), r.shape)
        print("  %-8s worst %.3g at (i=%d, p=%d) got %.12g want %.12g" % (n, r[i, p], i, p, g[i, p], w[i, p]))
