"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes view of oracle/liboracle.so (ora_capi.cpp).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (sea-current_b200/) never does.
"""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "liboracle.so")


def _abi():
    if "seacurrent_b200" not in sys.modules:
        spec = importlib.util.spec_from_file_location("seacurrent_b200_shim", os.path.join(ROOT, "seacurrent_b200.py"))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
    return sys.modules["seacurrent_b200"].abi


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".hpp", ".cpp", ".h"))]
    srcs.append(os.path.join(ROOT, "include", "seacurrent_b200.h"))
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["make", "-C", HERE, "-s"] + (["-B"] if force else []))
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
        _lib.ora_e1_batch_time.restype = C.c_double
        _lib.ora_lau_eval.restype = C.c_double
        _lib.ora_lau_optimise.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def num_threads():
    return lib().ora_num_threads()


def lp1d(v, A):
    A = _d(A).reshape(-1, 2)
    opt = C.c_double()
    ok = lib().ora_lp1d(C.c_double(v[0]), C.c_double(v[1]), _p(A), len(A), C.byref(opt))
    return bool(ok), opt.value


def lp2d(v, A, low, high):
    A = _d(A).reshape(-1, 3); v = _d(v); low = _d(low); high = _d(high)
    opt = np.zeros(2); act = np.zeros(2, dtype=np.int32)
    ok = lib().ora_lp2d(_p(v), _p(A), len(A), _p(low), _p(high), _p(opt), _p(act))
    return bool(ok), opt, act


def toppra_ppoly(coeffs, breaks, G, vel_lo, vel_hi, acc_lo, acc_hi, interpolation, sd_start=0.0, sd_end=0.0,
                 feasible=False):
    """coeffs [nseg][4][dof] (row 0 multiplies ds^3)."""
    coeffs = _d(coeffs); nseg, _, dof = coeffs.shape
    breaks = _d(breaks)
    K = np.zeros((G, 2)); x = np.zeros(G); feas = np.zeros((G, 2)) if feasible else None
    frc = C.c_int(0)
    rc = lib().ora_toppra_ppoly(nseg, dof, _p(coeffs), _p(breaks), G, _p(_d(vel_lo)), _p(_d(vel_hi)), _p(_d(acc_lo)),
                                _p(_d(acc_hi)), int(interpolation), C.c_double(sd_start), C.c_double(sd_end), _p(K),
                                _p(x), _p(feas), C.byref(frc))
    return rc, K, x, feas, frc.value


def toppra_ppoly_grid(coeffs, breaks, G, vel_lo_grid, vel_hi_grid, acc_lo, acc_hi, interpolation, sd_start=0.0, sd_end=0.0):
    """toppra_ppoly with velocity limits tabulated at the gridpoints: vel_*_grid [G][dof].  Returns (rc, K, x, fail_stage)."""
    coeffs = _d(coeffs); nseg, _, dof = coeffs.shape
    K = np.zeros((G, 2)); x = np.zeros(G); fs = C.c_int(-1)
    rc = lib().ora_toppra_ppoly_grid(nseg, dof, _p(coeffs), _p(_d(breaks)), G, _p(_d(vel_lo_grid)), _p(_d(vel_hi_grid)),
                                     _p(_d(acc_lo)), _p(_d(acc_hi)), int(interpolation), C.c_double(sd_start),
                                     C.c_double(sd_end), _p(K), _p(x), C.byref(fs))
    return rc, K, x, fs.value


def hermite_coeffs(pos_start, pos_end, vel_start, vel_end):
    """CubicHermiteSpline over [0, 1] (tp:geometric_path/piecewise_poly_path.cpp:364-375, dt = 1): [1][4][dof]."""
    p0, p1, v0, v1 = (_d(a) for a in (pos_start, pos_end, vel_start, vel_end))
    dt = 1.0
    c3 = p0; c2 = v0
    c0 = (v1 * dt - 2 * p1 + c2 * dt + 2 * c3) / dt ** 3
    c1 = (v1 - c2 - 3 * c0 * dt * dt) / (2 * dt)
    return np.ascontiguousarray(np.stack([c0, c1, c2, c3])[None])


def gen_vel_prof_ndof(pos_start, pos_end, vel_start, vel_end, vel_lo, vel_hi, acc_lo, acc_hi, G=101, dt=0.02, interpolation=1,
                      vel_lo_grid=None, vel_hi_grid=None, solve_mode=1):
    """gen_vel_prof<N> of ONE problem (sc:1191-1265): Hermite path, toppra with the interpolation discretisation, Spline
    parametrizer, uniform-dt record.  Limits [dof] (or grids [G][dof]).  Returns dict(status, fail_stage, K, x, t, time,
    pos, vel, acc) with pos / vel / acc [n][dof]."""
    dof = len(np.atleast_1d(pos_start))
    co = hermite_coeffs(pos_start, pos_end, vel_start if vel_start is not None else np.zeros(dof),
                        vel_end if vel_end is not None else np.zeros(dof))
    if vel_lo_grid is not None:
        rc, K, x, fs = toppra_ppoly_grid(co, [0.0, 1.0], G, vel_lo_grid, vel_hi_grid, acc_lo, acc_hi, interpolation)
    else:
        g = np.ones((G, 1))
        rc, K, x, fs = toppra_ppoly_grid(co, [0.0, 1.0], G, g * _d(vel_lo)[None, :], g * _d(vel_hi)[None, :], acc_lo, acc_hi, interpolation)
    out = dict(status=rc, fail_stage=fs, K=K, x=x)
    if rc != 0:
        return out
    gp = np.linspace(0.0, 1.0, G)
    ts, _, _, _ = spline_param(co, [0.0, 1.0], gp, x, solve_mode=solve_mode)
    n = int(np.ceil((ts[-1] - ts[0]) / dt))
    time = ts[0] + np.arange(n) * ((ts[-1] - ts[0]) / (n - 1)) if n > 1 else np.array([ts[0]])
    if n > 1:
        time[-1] = ts[-1]
    _, q, qd, qdd = spline_param(co, [0.0, 1.0], gp, x, solve_mode=solve_mode, t_eval=time)
    out.update(t=ts, time=time, pos=q, vel=qd, acc=qdd)
    return out


def spline_param(coeffs, breaks, gridpoints, vsq, t_eval_fn=None, solve_mode=0, t_eval=None):
    coeffs = _d(coeffs); nseg, _, dof = coeffs.shape
    gp = _d(gridpoints); G = len(gp)
    ts = np.zeros(G)
    te = _d(t_eval if t_eval is not None else [0.0])
    q = np.zeros((len(te), dof)); qd = np.zeros((len(te), dof)); qdd = np.zeros((len(te), dof))
    n = lib().ora_spline_param(nseg, dof, _p(coeffs), _p(_d(breaks)), G, _p(gp), _p(_d(vsq)), solve_mode, _p(ts), len(te),
                               _p(te), _p(q), _p(qd), _p(qdd))
    return ts[:n], q, qd, qdd


def const_accel_param(L, vel_start, vel_end, gridpoints, vsq, t_eval):
    """parametrizer::ConstAccel over gen_vel_prof's Hermite path: (ts, q, qd, qdd at t_eval)."""
    gp = _d(gridpoints); G = len(gp); te = _d(t_eval)
    ts = np.zeros(G); q = np.zeros(len(te)); qd = np.zeros(len(te)); qdd = np.zeros(len(te))
    lib().ora_const_accel_param(C.c_double(L), C.c_double(vel_start), C.c_double(vel_end), G, _p(gp), _p(_d(vsq)), _p(ts),
                                len(te), _p(te), _p(q), _p(qd), _p(qdd))
    return ts, q, qd, qdd


def cubic_spline(t, y, bc0, bc1, t_eval, solve_mode=0):
    t = _d(t); y = _d(y); te = _d(t_eval)
    out = np.zeros(len(te))
    lib().ora_cubic_spline(len(t), _p(t), _p(y), C.c_double(bc0), C.c_double(bc1), solve_mode, len(te), _p(te), _p(out))
    return out


def hermite_eval(pos, vel, times, s_eval, order):
    pos = _d(pos); vel = _d(vel); n, dof = pos.shape
    se = _d(s_eval)
    out = np.zeros((len(se), dof))
    lib().ora_hermite_eval(n, dof, _p(pos), _p(vel), _p(_d(times)), len(se), _p(se), order, _p(out))
    return out


def e1_single(wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, N=100, dt=0.02, mode=6, start_angle=float("nan"), cap=8192):
    """Reference-faithful (float32 spline side) E1 body for one problem."""
    wx = np.ascontiguousarray(wx, dtype=np.float32); wy = np.ascontiguousarray(wy, dtype=np.float32)
    W = len(wx)
    ctrl = np.zeros(8 * (W - 1), dtype=np.float32); tables = np.zeros((W - 1) * 101, dtype=np.float32)
    total = C.c_float(); ns = C.c_int(); nr = C.c_int()
    time = np.zeros(cap); f = lambda: np.zeros(cap, dtype=np.float32)
    pos, vel, acc, px, py, av = f(), f(), f(), f(), f(), f()
    rc = lib().ora_e1_single(_p(wx), _p(wy), W, C.c_float(start_angle), C.c_double(vel_lo), C.c_double(vel_hi),
                             C.c_double(acc_lo), C.c_double(acc_hi), N, C.c_float(dt), mode, _p(ctrl), _p(tables),
                             C.byref(total), cap, C.byref(ns), _p(time), _p(pos), _p(vel), _p(acc), _p(px), _p(py), _p(av),
                             C.byref(nr))
    S, Rn = min(ns.value, cap), min(nr.value, cap)
    return dict(rc=rc, ctrl=ctrl.reshape(W - 1, 4, 2), tables=tables.reshape(W - 1, 101), total=total.value,
                n_samples=ns.value, time=time[:S], pos=pos[:S], vel=vel[:S], acc=acc[:S], px=px[:Rn], py=py[:Rn],
                ang_vel=av[:Rn], n_resampled=nr.value)


def e1_batch_time(wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, N=100, dt=0.02, mode=7, nthreads=0):
    """Timed CPU baseline: the reference's E1 body (float32 spline side, FP64 toppra) over a batch,
    OpenMP over problems.  Returns (seconds, checksum)."""
    wx = np.ascontiguousarray(wx, dtype=np.float32); wy = np.ascontiguousarray(wy, dtype=np.float32)
    P, W = wx.shape
    cs = C.c_double()
    bc = lambda a: _d(np.broadcast_to(a, (P,)))
    s = lib().ora_e1_batch_time(P, W, _p(wx), _p(wy), _p(bc(vel_lo)), _p(bc(vel_hi)), _p(bc(acc_lo)), _p(bc(acc_hi)), N,
                                C.c_float(dt), mode, nthreads, C.byref(cs))
    return s, cs.value


def lau_eval(path_xy, Wp_xy, e, lam, start_angle=float("nan")):
    path = _d(path_xy); Wp = _d(Wp_xy); e = _d(e); W = len(path)
    ctrl = np.zeros(((W - 1), 6, 2)); ge = np.zeros(W); gW = np.zeros((W, 2))
    J = lib().ora_lau_eval(_p(path), W, C.c_double(start_angle), _p(Wp), _p(e), C.c_double(lam), _p(ctrl), _p(ge), _p(gW))
    return J, ctrl, ge, gW


def lau_optimise(path_xy, opt, start_angle=float("nan")):
    path = _d(path_xy); W = len(path)
    ctrl = np.zeros(((W - 1), 6, 2)); e = np.zeros(W); hist = np.zeros(opt.iters + 1)
    J = lib().ora_lau_optimise(_p(path), W, C.c_double(start_angle), C.byref(opt), _p(ctrl), _p(e), _p(hist))
    return J, ctrl, e, hist


def plan_batch(wx=None, wy=None, vel_lo=None, vel_hi=None, acc_lo=None, acc_hi=None, G=101, start_angle=None,
               vel_lo_grid=None, vel_hi_grid=None, vel_start=None, vel_end=None, spline_mode=0, solve_mode=1,
               S_cap=0, x_cap=0.0, dt=0.0, opt=None, nthreads=0, outputs=None, omega_max=0.0, acc_lat_max=0.0, obstacle_edges=None, param_mode=0, a_brake=0.0, t_react=0.0, collision_chords=0,
               path_len=None):
    """The product's batch on the CPU (FP64), same layouts as scb_plan_batch.  Returns (dict, seconds).
    path_len (and no waypoints) = SCB_STAGE_TIMING alone: gen_vel_prof on paths of those lengths."""
    abi = _abi()
    keep = []

    def arr(a, shape):
        if a is None:
            return None
        a = np.ascontiguousarray(np.broadcast_to(_d(a), shape))
        keep.append(a)
        return a.ctypes.data

    bi = abi.ScbBatchIn()
    if wx is None:
        path_len = _d(np.atleast_1d(path_len)); P, W = len(path_len), 2
        bi.path_len, bi.stages = arr(path_len, (P,)), abi.STAGE_TIMING
    else:
        wx = _d(wx); P, W = wx.shape
        bi.wx, bi.wy = arr(wx, (P, W)), arr(wy, (P, W))
    bi.P, bi.W, bi.G = P, W, G
    bi.start_angle = arr(start_angle, (P,))
    bi.vel_lo, bi.vel_hi = arr(vel_lo, (P,)), arr(vel_hi, (P,))
    bi.vel_lo_grid, bi.vel_hi_grid = arr(vel_lo_grid, (G, P)), arr(vel_hi_grid, (G, P))
    bi.acc_lo, bi.acc_hi = arr(acc_lo, (P,)), arr(acc_hi, (P,))
    bi.vel_start, bi.vel_end = arr(vel_start, (P,)), arr(vel_end, (P,))
    bi.x_cap, bi.dt = x_cap, dt
    bi.omega_max, bi.acc_lat_max = omega_max, acc_lat_max
    bi.a_brake, bi.t_react = a_brake, t_react
    bi.collision_chords = collision_chords
    bi.param_mode = param_mode
    if obstacle_edges is not None and len(obstacle_edges):
        ed = np.ascontiguousarray(_d(obstacle_edges)).reshape(-1, 4)
        keep.append(ed)
        bi.obstacle_edges, bi.n_obstacle_edges = ed.ctypes.data, ed.shape[0]
    if opt is None:
        opt = abi.ScbOpt(32, 0, 0.05, 1.0, 0.25, 4.0, 0.1, 0.05, 0.25)
    bi.opt = opt
    d = 3 if spline_mode == 0 else 5
    shapes = {"ctrl_x": (P, (W - 1) * (d + 1)), "ctrl_y": (P, (W - 1) * (d + 1)), "arclen_table": (P, (W - 1) * 101),
              "arclen_total": (P,), "T_total": (P,), "cost": (P,)}
    for n in ("K_lo", "K_hi", "x", "t", "px", "py", "v", "a", "omega"):
        shapes[n] = (G, P)
    if S_cap > 0:
        for n in ("s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega"):
            shapes[n] = (S_cap, P)
    res = {"status": np.zeros(P, dtype=np.int32), "fail_stage": np.zeros(P, dtype=np.int32)}
    bo = abi.ScbBatchOut()
    bo.status, bo.fail_stage, bo.S_cap = res["status"].ctypes.data, res["fail_stage"].ctypes.data, S_cap
    if S_cap > 0:
        res["n_samples"] = np.zeros(P, dtype=np.int32)
        bo.n_samples = res["n_samples"].ctypes.data
    for n, shp in shapes.items():
        if outputs is not None and n not in outputs:
            continue
        res[n] = np.zeros(shp)
        setattr(bo, n, res[n].ctypes.data)
    sec = C.c_double()
    lib().ora_plan_batch(C.byref(bi), C.byref(bo), spline_mode, solve_mode, nthreads, C.byref(sec))
    return res, sec.value


# ---------------------------------------------------------------- global planner front end (ora_fmt.hpp)
def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def halton(b, n, state=None):
    """sea_current.hpp:100-132.  state = [f, i] (mutated), None = a fresh sequence."""
    st = np.asarray(state if state is not None else [0, 0], dtype=np.int32)
    out = np.empty(n, dtype=np.float32)
    lib().ora_halton(int(b), int(n), _p(st), _p(out))
    if state is not None:
        state[0], state[1] = int(st[0]), int(st[1])
    return out


def make_obstacle(vertices, edges=None, closed=True):
    """The two obstacle constructors (sc:251-283): vertex loop (an even count appends the first vertex to the local
    copy the lines are built from) or explicit (i, j) edges.  `closed` is never set by the reference's constructors."""
    v = [tuple(map(np.float32, p)) for p in vertices]
    xmax = xmin = v[0][0]; ymax = ymin = v[0][1]
    lines = []

    def enclose(p):
        nonlocal xmax, xmin, ymax, ymin
        xmax = max(xmax, p[0]); xmin = min(xmin, p[0]); ymax = max(ymax, p[1]); ymin = min(ymin, p[1])
    if edges is None:
        loc = list(v) + ([v[0]] if len(v) % 2 == 0 else [])
        for i in range(len(loc) - 1):
            lines.append(loc[i] + loc[i + 1]); enclose(loc[i])
    else:
        for i, j in edges:
            lines.append(v[i] + v[j]); enclose(v[i]); enclose(v[j])
    return {"vertices": _f(v).reshape(-1, 2), "lines": _f(lines).reshape(-1, 4), "rect": _f([xmax, xmin, ymax, ymin]),
            "closed": bool(closed)}


def _pack_obstacles(obstacles):
    nv = np.asarray([len(o["vertices"]) for o in obstacles], dtype=np.int32)
    nl = np.asarray([len(o["lines"]) for o in obstacles], dtype=np.int32)
    verts = _f(np.concatenate([o["vertices"] for o in obstacles]) if obstacles else np.zeros((0, 2)))
    lines = _f(np.concatenate([o["lines"] for o in obstacles]) if obstacles else np.zeros((0, 4)))
    rects = _f(np.concatenate([o["rect"] for o in obstacles]) if obstacles else np.zeros(0))
    closed = np.asarray([1 if o["closed"] else 0 for o in obstacles], dtype=np.int32)
    return nv, verts, nl, lines, rects, closed


def obstacle_contains(o, p):
    return bool(lib().ora_obstacle_contains(len(o["vertices"]), _p(o["vertices"]), len(o["lines"]), _p(o["lines"]), _p(o["rect"]),
                                            int(o["closed"]), C.c_float(p[0]), C.c_float(p[1])))


def sample_free(n, rect, obstacles=(), free_rects=((-1e30, 1e30, -1e30, 1e30),), state=None, max_rounds=64):
    """planning_space::sample_free (sc:1294-1313).  rect = (x_min, x_max, y_min, y_max); free_rects = the free-space
    allocations as rectangles; state = [x.f, x.i, y.f, y.i] (mutated).  Returns float32 [m][2], (0, 0) first."""
    st = np.asarray(state if state is not None else [0, 0, 0, 0], dtype=np.int32)
    nv, verts, nl, lines, rects, closed = _pack_obstacles(list(obstacles))
    fr = _f(free_rects).reshape(-1, 4)
    out = np.empty((max(n, 1) + 1, 2), dtype=np.float32)
    m = lib().ora_sample_free(int(n), _p(_f(rect)), len(nv), _p(nv), _p(verts), _p(nl), _p(lines), _p(rects), _p(closed), len(fr),
                              _p(fr), _p(st), _p(out), int(max_rounds))
    if state is not None:
        state[:] = [int(x) for x in st]
    return out[:m].copy()


def fmt_batch(nodes, edges, init, goal, rn, max_path=256):
    """fast_marching_trees (sc:1339-1407) for Q queries over one node set.  nodes [n][2], edges [E][4], init/goal [Q][2]."""
    nodes = _f(nodes).reshape(-1, 2); init = _f(init).reshape(-1, 2); goal = _f(goal).reshape(-1, 2)
    ed = _f(edges).reshape(-1, 4) if edges is not None else np.zeros((0, 4), np.float32)
    Q, n = len(init), len(nodes)
    nx, ny = _f(nodes[:, 0]), _f(nodes[:, 1])
    res = {"status": np.empty(Q, np.int32), "path_len": np.empty(Q, np.int32), "closed": np.empty(Q, np.int32),
           "path_idx": np.full((Q, max_path), -1, np.int32), "path_x": np.zeros((Q, max_path), np.float32),
           "path_y": np.zeros((Q, max_path), np.float32), "cost": np.empty(Q, np.float32)}
    lib().ora_fmt_batch(Q, n, _p(nx), _p(ny), len(ed), _p(ed), _p(_f(init[:, 0])), _p(_f(init[:, 1])), _p(_f(goal[:, 0])),
                        _p(_f(goal[:, 1])), C.c_float(rn), max_path, _p(res["status"]), _p(res["path_len"]), _p(res["closed"]),
                        _p(res["path_idx"]), _p(res["path_x"]), _p(res["path_y"]), _p(res["cost"]))
    return res
