"""Randomised parity cases shared by tests/test_random_sweep.py (a pinned seed) and tools/gpu/fuzz_parity.py (a
time-boxed sweep): shapes, options and degenerate paths the parametrised tests do not enumerate, each checked through
tolerances.check at 1e-9 with identical verdicts.

Two input classes are ill-conditioned for the REFERENCE's algorithm itself and are checked at 1e-6 on the geometry only:
  * collinear waypoints - calc_tangent (sc:347-369) takes acos of the legs' normalised dot product, which is -1 there:
    acos has an infinite derivative at -1, so one ulp in the quotient turns the tangent by 1.5e-8 rad;
  * a straight path (also: two waypoints with no start angle) has an identically zero cost gradient, so the optimiser's
    iRPROP- moves are decided by rounding noise (the 1e-9 threshold is relative to the LARGEST component, itself noise):
    the elongations differ between any two implementations, the curve is the same straight line.
"""
import numpy as np

import seacurrent_b200 as scb
import tolerances

GRID = tolerances.GRID


def make_case(rng):
    c = {}
    c["mode"] = str(rng.choice(["ref_cubic", "lau_quintic"]))
    c["W"] = int(rng.choice([2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 15, 16, 17, 18, 24, 31, 32, 33, 34, 48, 65]))
    c["G"] = int(rng.choice([2, 3, 4, 5, 9, 17, 64, 101, 257, 500]))
    c["P"] = int(rng.choice([1, 2, 3, 31, 32, 33, 64, 97, 130]))
    c["lp"] = int(rng.choice([scb.abi.LP_CLOSED_FORM, scb.abi.LP_SEIDEL]))
    c["iters"] = int(rng.choice([0, 1, 2, 7, 32]))
    c["move"] = int(rng.integers(0, 2))
    c["angle"] = str(rng.choice(["none", "random", "mixed"]))
    c["ends"] = str(rng.choice(["none", "none", "start", "both"]))
    c["grid_limits"] = bool(rng.random() < 0.3)
    c["curv"] = bool(rng.random() < 0.3)
    c["param_mode"] = int(rng.random() < 0.25)
    c["shape"] = str(rng.choice(["walk", "walk", "walk", "line", "uturn", "dup", "far", "tiny", "zigzag"]))
    c["slow"] = float(rng.choice([0.0, 0.05, 0.3]))
    c["seed"] = int(rng.integers(0, 1 << 30))
    if c["W"] == 2 and c["angle"] != "random" and c["shape"] != "line":
        c["shape"] = "line"   # two waypoints and no start angle: a straight path
    return c


def build(c):
    rng = np.random.default_rng(c["seed"])
    P, W, G = c["P"], c["W"], c["G"]
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    sh = c["shape"]
    if sh == "line":       # collinear waypoints: curvature identically 0
        t = np.cumsum(rng.uniform(1.0, 5.0, (P, W)), axis=1)
        ang = rng.uniform(0, 2 * np.pi, (P, 1))
        wx, wy = t * np.cos(ang), t * np.sin(ang)
    elif sh == "uturn" and W >= 3:   # a 180-degree reversal at an inner waypoint
        k = W // 2
        wx[:, k + 1] = wx[:, k - 1] + 1e-3; wy[:, k + 1] = wy[:, k - 1]
    elif sh == "dup" and W >= 3:     # a repeated waypoint in a few problems: zero-length leg
        sel = rng.random(P) < 0.3
        wx[sel, 1] = wx[sel, 0]; wy[sel, 1] = wy[sel, 0]
    elif sh == "far":      # large offset: cancellation in every difference
        wx = wx + 1e6; wy = wy - 3e5
    elif sh == "tiny":     # millimetre scale
        wx = wx * 1e-3; wy = wy * 1e-3
    elif sh == "zigzag":
        wx = np.cumsum(np.ones((P, W)), axis=1) * rng.uniform(0.5, 3, (P, 1)); wy = (np.arange(W) % 2)[None, :] * rng.uniform(0.2, 4, (P, 1))
    lim = scb.workloads.random_limits(rng, P, slow_fraction=c["slow"])
    if sh == "tiny":
        lim = {k: v * 1e-3 for k, v in lim.items()}
    kw = dict(wx=np.ascontiguousarray(wx), wy=np.ascontiguousarray(wy), G=G, **lim)
    if c["angle"] == "random":
        kw["start_angle"] = rng.uniform(-np.pi, np.pi, P)
    elif c["angle"] == "mixed":
        a = rng.uniform(-np.pi, np.pi, P); a[rng.random(P) < 0.5] = np.nan
        kw["start_angle"] = a
    if c["ends"] in ("start", "both"):
        kw["vel_start"] = rng.uniform(0.0, 0.3, P) * lim["vel_hi"]
    if c["ends"] == "both":
        kw["vel_end"] = rng.uniform(0.0, 0.3, P) * lim["vel_hi"]
        kw.setdefault("vel_start", np.zeros(P))
    if c["grid_limits"]:
        f = rng.uniform(0.3, 1.0, (G, P))
        if G > 8:
            f[G // 3: G // 2, rng.random(P) < 0.3] *= 0.1
        kw["vel_hi_grid"] = f * lim["vel_hi"][None, :]; kw["vel_lo_grid"] = -kw["vel_hi_grid"]
    if c["curv"]:
        kw["omega_max"] = float(rng.uniform(0.3, 3.0)); kw["acc_lat_max"] = float(rng.uniform(0.2, 2.0)) if rng.random() < 0.7 else 0.0
    kw["param_mode"] = c["param_mode"]
    opt = scb.default_opt(move_waypoints=c["move"])
    opt.iters = c["iters"]
    kw["opt"] = opt
    if c.get("obst"):      # wall pieces beside random legs of the first path: tangent clipping (cubic), collision term and
        r2 = np.random.default_rng(c["seed"] + 1)   # verdict (quintic), optionally the obstacle-distance speed limit
        k = r2.integers(0, W - 1, 8)
        mx, my = (wx[0, k] + wx[0, k + 1]) / 2, (wy[0, k] + wy[0, k + 1]) / 2
        off = r2.uniform(-2.0, 2.0, (8, 2)); d = r2.uniform(-1.0, 1.0, (8, 2))
        kw["obstacle_edges"] = np.ascontiguousarray(np.stack([mx + off[:, 0] - d[:, 0], my + off[:, 1] - d[:, 1],
                                                              mx + off[:, 0] + d[:, 0], my + off[:, 1] + d[:, 1]], 1))
        kw["collision_chords"] = int(c["obst"])
        if c.get("brake"):
            kw["a_brake"], kw["t_react"] = 1.0, 0.2
    if c.get("S_cap"):
        kw["S_cap"] = c["S_cap"]; kw["dt"] = c["dt"]
    return kw


def make_ext_case(rng):
    """make_case plus obstacles (tangent clipping / collision term / obstacle-distance limit) and the uniform-dt stream;
    general-position shapes only."""
    c = make_case(rng)
    if c["shape"] in ("line", "far", "uturn"):
        c["shape"] = "walk"
    c["W"] = max(c["W"], 3)
    c["obst"] = int(rng.choice([0, 4, 8, 16])) if c["W"] <= 48 else 0
    c["brake"] = bool(rng.random() < 0.5)
    c["S_cap"] = int(rng.choice([0, 128, 2048]))
    c["dt"] = float(rng.choice([0.02, 0.1, 0.5]))
    return c


def straight(c):
    return c["shape"] == "line"


def run_case(c, planner, ora):
    """planner: a BatchPlanner of c["mode"], or None = the oracle against itself (no GPU: checks the harness)."""
    kw = build(c)
    okw = dict(kw, spline_mode=0 if c["mode"] == "ref_cubic" else 1)
    res = ora.plan_batch(**okw)[0] if planner is None else planner.plan(lp_mode=c["lp"], **kw)
    names = ("ctrl_x", "ctrl_y", "arclen_total", "arclen_table", "T_total", "cost") + GRID
    # the spread of EVERY quantity under 8-ulp input perturbations: the sweep feeds degenerate paths on purpose
    ref, spread = tolerances.conditioning(ora, names=names + (tuple("s_" + k for k in ("time", "pos", "px", "py", "vel", "acc", "omega")) if c.get("S_cap") else ()), **okw)
    ok = ref["status"] == 0
    # Millimetre scale with the closed-form LP: the reference's Seidel solver accepts a vertex that violates a row by
    # less than its ABSOLUTE tolerance (1e-13, tp:solver/seidel-internal.hpp:54-57) and stops there; the closed form
    # returns the exact optimum.  With x = sdot^2 ~ 1e-7 that slack is 1e-8 relative (measured: 1.35e-8 on K_hi, the
    # closed form agreeing with an 80-bit evaluation of the stage LP, the Seidel kernel with the oracle to 4e-15):
    # such cases are held to 1e-6 with the closed form and to 1e-9 with the Seidel back end.
    rtol = 1e-6 if (c["shape"] == "tiny" and c["lp"] == scb.abi.LP_CLOSED_FORM) else 1e-9
    if straight(c):
        # the curve is a straight line whatever the elongations: geometry, length and timing at 1e-6, curvature ~ 0
        np.testing.assert_array_equal(res["status"], ref["status"])
        np.testing.assert_array_equal(res["fail_stage"], ref["fail_stage"])
        # ... unless the optimiser may move the waypoints: its first move off the line is decided by noise, and from
        # there the two implementations descend from different points (verdicts only)
        noise_moves = c["mode"] == "lau_quintic" and c["move"] and c["iters"] > 0
        if ok.any() and not noise_moves:
            L = ref["arclen_total"][ok]
            for n in ("arclen_total", "T_total"):
                assert np.all(np.abs(res[n][ok] - ref[n][ok]) <= 1e-6 * np.abs(ref[n][ok])), n
            for n in ("px", "py"):
                assert np.all(np.abs(res[n][:, ok] - ref[n][:, ok]) <= 1e-6 * L[None, :]), n
            for n in ("x", "K_hi", "t"):
                sc = np.abs(ref[n][:, ok]).max(axis=0)
                assert np.all(np.abs(res[n][:, ok] - ref[n][:, ok]) <= 1e-6 * np.maximum(np.abs(ref[n][:, ok]), 1e-3 * sc[None, :])), n
            sa = kw.get("start_angle")
            line = ok & (np.ones(len(ok), dtype=bool) if sa is None else np.isnan(sa))   # no start angle: straight from end to end
            for r in (res, ref):
                if line.any():
                    assert np.all(np.abs(tolerances._kappa(r)[:, line]) * ref["arclen_total"][line][None, :] <= 1e-3), "curvature of a straight path"
    elif c["param_mode"] == 1:   # const-accel parametrizer: a, v are difference quotients of x, no spline solve
        tolerances.check(res, ref, rtol, tuple(n for n in names if n not in ("a", "v")), spread, spread_all=True)
        for n in ("a", "v"):
            if not ok.any():
                break
            sc = tolerances.scales(ref, ok)[n]
            tol = rtol * np.maximum(np.abs(ref[n][:, ok]), 1e-3 * sc[None, :]) + tolerances.SPREAD_FACTOR * spread[n][:, ok]
            err = np.abs(res[n][:, ok] - ref[n][:, ok])
            assert np.all(err <= tol), (n, float((err / tol).max()), int((err > tol).sum()))
    else:
        tolerances.check(res, ref, rtol, names, spread, spread_all=True)
    if c.get("S_cap") and not straight(c):
        check_stream(res, ref, spread, c["S_cap"], const_accel=c["param_mode"] == 1, rtol=rtol)
    return int(ok.sum()), int(len(ok))


def check_stream(res, ref, spread, S, const_accel=False, rtol=1e-9):
    """The uniform-dt sample stream: counts identical; time / position / curve points at 1e-9 of the stream's scale,
    the derivatives with the oracle's measured spread (tolerances.py).  const_accel: the acceleration of
    parametrizer::ConstAccel is piecewise constant, and bang-bang profiles put samples EXACTLY on knots (a profile that
    accelerates from rest over the first of three intervals has t_1 = 0.4 T, sample 126 of 316): which side such a
    sample takes is decided by the last bit of the knot time, so s_acc is not compared within 1e-9 T of a knot."""
    ok = ref["status"] == 0
    np.testing.assert_array_equal(res["n_samples"][ok], ref["n_samples"][ok])
    n = np.minimum(ref["n_samples"], S)
    written = (np.arange(S)[:, None] < np.where(ok, n, 0)[None, :])
    for k in ("s_time", "s_pos", "s_px", "s_py", "s_vel", "s_acc", "s_omega"):
        w = np.where(written, ref[k], 0.0); g = np.where(written, res[k], 0.0)
        sc = np.maximum(np.abs(w).max(axis=0, keepdims=True), 1e-300)
        tol = rtol * np.maximum(np.abs(w), 1e-3 * sc)
        if k in spread:
            tol = tol + tolerances.SPREAD_FACTOR * np.where(written, spread[k], 0.0)
        bad = np.abs(g - w) > tol
        if const_accel and k == "s_acc":
            T = ref["t"][-1][None, :]
            near = np.zeros_like(bad)
            for row in ref["t"][1:-1]:
                near |= np.abs(ref["s_time"] - row[None, :]) <= 1e-9 * T
            bad &= ~near
        assert not bad.any(), (k, float((np.abs(g - w) / tol)[bad].max()), int(bad.sum()))
