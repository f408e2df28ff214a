"""Which control points differ between the pair mapping and the lane mapping?"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import seacurrent_b200 as scb
for W, iters in ((3, 0), (3, 32), (10, 0), (10, 32)):
    rng = np.random.default_rng(50 + W)
    P, G = 203, 40
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.0)
    out = {}
    for pair in ("0", "1"):
        os.environ["SCB_PATH_PAIR"] = pair
        pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=W, max_gridpoints=G)
        o = scb.default_opt(move_waypoints=0); o.iters = iters
        out[pair] = pl.plan(wx=wx, wy=wy, G=G, opt=o, **lim)
        pl.close()
    a, b = out["0"], out["1"]
    for k in ("ctrl_x", "ctrl_y", "cost"):
        d = a[k] != b[k]
        print("W", W, "iters", iters, k, "differ", int(d.sum()), "of", d.size)
        if k != "cost" and d.any():
            idx = np.argwhere(d)
            segs = idx[:, 1] // 6; js = idx[:, 1] % 6
            print("   (segment, j):", sorted(set(zip(segs.tolist(), js.tolist()))), " max rel", float(np.max(np.abs(a[k] - b[k])[d] / np.abs(a[k][d]))))
