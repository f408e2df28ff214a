"""Diagnosis of two findings of the extended sweep (tools/gpu/fuzz_parity.py, FUZZ_EXT=1, seed 11):
(1) K_hi of one millimetre-scale problem differs by 1.3e-8 between the Seidel kernel and the oracle;
(2) s_acc of the const-accel stream differs by a factor at isolated samples."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import seacurrent_b200 as scb
import fuzz_cases
from oracle import oracle as ora

ora.build()
pl = {m: scb.BatchPlanner(0, "fp64", m, max_problems=256, max_waypoints=80, max_gridpoints=512, max_samples=2048) for m in ("ref_cubic", "lau_quintic")}

# ---- (1)
c = {'mode': 'ref_cubic', 'W': 18, 'G': 500, 'P': 97, 'lp': 2, 'iters': 7, 'move': 1, 'angle': 'none', 'ends': 'both', 'grid_limits': True,
     'curv': False, 'param_mode': 1, 'shape': 'tiny', 'slow': 0.3, 'seed': 912838130, 'obst': 4, 'brake': False, 'S_cap': 128, 'dt': 0.1}
kw = fuzz_cases.build(c)
ref = ora.plan_batch(**dict(kw, spline_mode=0))[0]
ok = np.flatnonzero(ref["status"] == 0)
p = int(ok[85])
rs = pl["ref_cubic"].plan(lp_mode=scb.abi.LP_SEIDEL, **kw)
rc = pl["ref_cubic"].plan(lp_mode=scb.abi.LP_CLOSED_FORM, **kw)
# the oracle's timing stage alone on the KERNEL's path lengths
tk = {k: v for k, v in kw.items() if k not in ("wx", "wy", "start_angle", "obstacle_edges", "collision_chords", "opt")}
rt = ora.plan_batch(path_len=rs["arclen_total"], **tk)[0]
print("problem", p, "L gpu/oracle", rs["arclen_total"][p], ref["arclen_total"][p], "rel", rs["arclen_total"][p] / ref["arclen_total"][p] - 1)
print("limits", {k: float(kw[k][p]) for k in ("vel_lo", "vel_hi", "acc_lo", "acc_hi")}, "vel_start/end", kw.get("vel_start", [0] * 200)[p], kw.get("vel_end", [0] * 200)[p])
for name, r in (("seidel kernel", rs), ("closed-form kernel", rc), ("oracle on the kernel's L", rt)):
    d = np.abs(r["K_hi"][:, p] - ref["K_hi"][:, p]) / np.maximum(np.abs(ref["K_hi"][:, p]), 1e-300)
    dx = np.abs(r["x"][:, p] - ref["x"][:, p]) / np.maximum(np.abs(ref["x"][:, p]), 1e-300)
    print("%-26s K_hi max rel %.3e at %d   x max rel %.3e at %d   status %d" % (name, d.max(), d.argmax(), dx.max(), dx.argmax(), r["status"][p]))
i = 470
for name, r in (("oracle", ref), ("seidel", rs), ("closed", rc), ("oracle@L", rt)):
    print(name, [float(r["K_hi"][k, p]) for k in (i - 2, i, i + 2, 490, 498, 499)])

# ---- (2)
c = {'mode': 'lau_quintic', 'W': 16, 'G': 4, 'P': 31, 'lp': 2, 'iters': 1, 'move': 0, 'angle': 'mixed', 'ends': 'none', 'grid_limits': False,
     'curv': False, 'param_mode': 1, 'shape': 'walk', 'slow': 0.05, 'seed': 834291615, 'obst': 0, 'brake': False, 'S_cap': 2048, 'dt': 0.5}
kw = fuzz_cases.build(c)
ref = ora.plan_batch(**dict(kw, spline_mode=0 if c["mode"] == "ref_cubic" else 1))[0]
res = pl[c["mode"]].plan(lp_mode=c["lp"], **kw)
S = c["S_cap"]
n = np.minimum(ref["n_samples"], S)
w = (np.arange(S)[:, None] < np.where(ref["status"] == 0, n, 0)[None, :])
err = np.where(w, np.abs(res["s_acc"] - ref["s_acc"]), 0.0)
ii, pp = np.unravel_index(err.argmax(), err.shape)
print("case", c)
print("worst s_acc sample", ii, "problem", pp, "n_samples", ref["n_samples"][pp], "got/want", res["s_acc"][ii, pp], ref["s_acc"][ii, pp])
print("sample time gpu/oracle", repr(res["s_time"][ii, pp]), repr(ref["s_time"][ii, pp]))
print("knot times gpu   ", [repr(float(v)) for v in res["t"][:, pp]])
print("knot times oracle", [repr(float(v)) for v in ref["t"][:, pp]])
print("neighbours got ", res["s_acc"][max(ii - 2, 0):ii + 3, pp], "want", ref["s_acc"][max(ii - 2, 0):ii + 3, pp])
