"""Small batches through every entry point, for compute-sanitizer --tool memcheck (one tool per gpurun call)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import seacurrent_b200 as scb

rng = np.random.default_rng(1)
for mode, W, G, P in (("lau_quintic", 10, 120, 70), ("ref_cubic", 7, 60, 33), ("lau_quintic", 40, 80, 5)):
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.1)
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=P, max_waypoints=W, max_gridpoints=G, max_samples=512)
    a = pl.plan(wx=wx, wy=wy, G=G, S_cap=512, **lim)
    b = pl.plan(wx=wx, wy=wy, G=G, S_cap=512, record_f32=True, **lim)
    pl.set_host_chunk(16)
    c = pl.plan(wx=wx, wy=wy, G=G, record_f32=True, **lim)
    edges = np.array([[wx[0, 1] + 0.5, wy[0, 1] - 1.0, wx[0, 1] + 0.5, wy[0, 1] + 1.0], [3.0, 3.0, 4.0, 3.5]])
    d = pl.plan(wx=wx, wy=wy, G=G, obstacle_edges=edges, collision_chords=8, a_brake=1.0, t_react=0.1, omega_max=1.0, **lim)
    if mode == "lau_quintic" and W <= 32:
        s = pl.plan_summary(wx=wx, wy=wy, G=G, **lim)
        f = pl.fetch(np.array([0, P - 1, 3], dtype=np.int32), outputs=scb.ALL_OUTPUTS)
    pl.close()
    print(mode, W, "ok", int((a["status"] == 0).sum()), int((d["status"] == 19).sum()))
mp = scb.MultiPlanner(2, "fp64", "lau_quintic", max_problems=40, max_waypoints=10, max_gridpoints=64, devices=[0, 0])
wx, wy = scb.workloads.random_walk_paths(rng, 37, 10); lim = scb.workloads.random_limits(rng, 37)
r = mp.plan(wx=wx, wy=wy, G=64, **lim); s = mp.plan_summary(wx=wx, wy=wy, G=64, **lim); f = mp.fetch(np.array([36, 0, 18, 19], dtype=np.int32))
mp.close()
pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=8, max_waypoints=4, max_gridpoints=64)
co = np.zeros((3, 2, 4, 2)); co[:, :, 2, :] = 1.0; co[:, :, 3, :] = np.arange(2)[None, :, None]
pp = pl.toppra_ppoly(co, [0.0, 1.0, 2.0], 41, [-1, -1], [1, 1], [-0.5, -0.5], [0.5, 0.5], 0)
pl.close()
print("all ok")
