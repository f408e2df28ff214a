"""parametrizer::ConstAccel as the timing stage's parametrizer (second half; scb_batch_in.param_mode =
SCB_PARAM_CONST_ACCEL).  Reference: third-party/toppra/cpp/src/toppra/parametrizer/const_accel.cpp:19-115 — constant
path acceleration on every grid interval, knot times by the same trapezoid rule as the Spline parametrizer, nothing
solved.  gen_vel_prof (sea_current.hpp:1231) would construct it in place of parametrizer::Spline."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def _problem(P=48, W=8, seed=5):
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.0)
    return wx, wy, lim


def _hermite(L, s):
    """q, q', q'' of the 1-D Hermite path of gen_vel_prof (sc:1213-1222) with zero end velocities."""
    return L * (3 * s * s - 2 * s ** 3), 6 * L * s * (1 - s), 6 * L * (1 - 2 * s)


def test_oracle_const_accel_follows_the_reference_formulas(oracle):
    wx, wy, lim = _problem()
    G, S_cap = 101, 16384
    ca, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, param_mode=1, S_cap=S_cap, **lim)
    sp, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, param_mode=0, S_cap=S_cap, **lim)
    np.testing.assert_array_equal(ca["status"], sp["status"])
    np.testing.assert_array_equal(ca["x"], sp["x"])  # the toppra passes do not depend on the parametrizer
    ok = ca["status"] == 0
    assert ok.sum() >= 40
    s = np.linspace(0.0, 1.0, G)[:, None]
    sd = np.sqrt(ca["x"][:, ok])
    L = ca["arclen_total"][None, ok]
    # process_internals (const_accel.cpp:23-29)
    dt = 2 * np.diff(s, axis=0) / (sd[1:] + sd[:-1])
    t = np.concatenate([np.zeros((1, ok.sum())), np.cumsum(dt, axis=0)])
    np.testing.assert_allclose(ca["t"][:, ok], t, rtol=1e-12)
    np.testing.assert_allclose(ca["T_total"][ok], t[-1], rtol=1e-12)
    np.testing.assert_allclose(ca["t"][:, ok], sp["t"][:, ok], rtol=1e-12)  # same trapezoid rule as spline.cpp:14-24
    us = 0.5 * (sd[1:] ** 2 - sd[:-1] ** 2) / np.diff(s, axis=0)
    # evalParams at the knots: knot i >= 1 sits at the end of interval i - 1, where s = s_i and sdot = sd_i up to rounding
    _, q1, q2 = _hermite(L, s)
    np.testing.assert_allclose(ca["v"][:, ok], q1 * sd, rtol=1e-9, atol=1e-12 * L.max())
    a_want = q2 * sd ** 2 + q1 * np.concatenate([us[:1], us])
    np.testing.assert_allclose(ca["a"][:, ok], a_want, rtol=1e-8, atol=1e-9 * np.abs(a_want).max())
    # the uniform-dt stream: positions are monotone, end at L, and v = dq/dt of the same piecewise-quadratic s(t)
    for j, p in enumerate(np.flatnonzero(ok)[:8]):
        n = min(int(ca["n_samples"][p]), S_cap)
        tt, pos, vel = ca["s_time"][:n, p], ca["s_pos"][:n, p], ca["s_vel"][:n, p]
        assert n == ca["n_samples"][p] and np.all(np.diff(pos) >= -1e-12 * L[0, j]) and pos[0] == 0.0
        assert abs(pos[-1] - L[0, j]) <= 1e-9 * L[0, j]
        k = np.clip(np.searchsorted(t[1:-1, j], tt, side="left"), 0, G - 2)
        d = tt - t[k, j]
        ss = np.clip(s[k, 0] + d * sd[k, j] + 0.5 * d * d * us[k, j], 0, 1)
        q0, q1s, _ = _hermite(L[0, j], ss)
        np.testing.assert_allclose(pos, q0, rtol=1e-12, atol=1e-12 * L[0, j])
        np.testing.assert_allclose(vel, q1s * (sd[k, j] + d * us[k, j]), rtol=1e-10, atol=1e-12)
    # the two parametrizers describe the same motion: durations equal, knot velocities close
    assert np.max(np.abs(ca["v"][:, ok] - sp["v"][:, ok])) <= 0.05 * np.max(np.abs(sp["v"][:, ok]))


@pytest.mark.gpu
@pytest.mark.parametrize("mode,spline_mode", [("ref_cubic", 0), ("lau_quintic", 1)])
def test_gpu_matches_oracle_const_accel(oracle, mode, spline_mode):
    wx, wy, lim = _problem(P=160, W=10, seed=13)
    G, S_cap = 300, 2048
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=256, max_waypoints=16, max_gridpoints=512, max_samples=S_cap)
    res = pl.plan(wx=wx, wy=wy, G=G, S_cap=S_cap, param_mode=scb.abi.PARAM_CONST_ACCEL,
                  outputs=scb.ALL_OUTPUTS + ("s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega"), **lim)
    both = pl.plan(wx=wx, wy=wy, G=G, param_mode=scb.abi.PARAM_SPLINE, **lim)
    pl.close()
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=spline_mode, param_mode=1, S_cap=S_cap, **lim)
    np.testing.assert_array_equal(res["status"], ref["status"])
    np.testing.assert_array_equal(res["n_samples"], ref["n_samples"])
    ok = ref["status"] == 0
    assert ok.sum() >= 100
    np.testing.assert_array_equal(res["x"], both["x"])
    for n in ("t", "v", "a", "omega", "s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega"):
        g, w = res[n][:, ok], ref[n][:, ok]
        scale = np.abs(w).max(axis=0, keepdims=True)
        # omega = v * kappa carries the curvature's conditioning at a spike (see test_gpu_parity.fp32_omega_tol)
        rel = 1e-7 if n in ("omega", "s_omega") else 1e-9
        tol = rel * np.maximum(np.abs(w), 1e-3 * scale)
        assert np.all(np.abs(g - w) <= tol), (n, float(np.max(np.abs(g - w) / tol)))
    assert np.max(np.abs(res["T_total"][ok] / ref["T_total"][ok] - 1)) <= 1e-12


@pytest.mark.gpu
def test_gpu_const_accel_fp32_storage_and_bad_mode(oracle):
    wx, wy, lim = _problem(P=128, W=10, seed=17)
    f32 = lambda a: np.asarray(a, dtype=np.float32).astype(np.float64)
    wx, wy = f32(wx), f32(wy)
    lim = {k: f32(v) for k, v in lim.items()}
    G = 200
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, param_mode=1, **lim)
    pl = scb.BatchPlanner(0, "fp32", "ref_cubic", max_problems=128, max_waypoints=10, max_gridpoints=G)
    res = pl.plan(wx=wx, wy=wy, G=G, param_mode=scb.abi.PARAM_CONST_ACCEL, **lim)
    ok = (ref["status"] == 0) & (res["status"] == 0)
    assert ok.mean() > 0.95
    for n in ("t", "v", "a"):
        g, w = res[n][:, ok].astype(np.float64), ref[n][:, ok]
        scale = np.abs(w).max(axis=0, keepdims=True)
        # a = q'' sd^2 + q' u with u a difference quotient of the STORED float32 sdot^2: its rounding (eps32 x) comes back
        # multiplied by (G - 1) / 2
        floor = 2 * np.finfo(np.float32).eps * G * scale if n == "a" else 1e-6 * scale
        assert np.all(np.abs(g - w) <= 1e-4 * np.abs(w) + floor), n
    with pytest.raises(scb.ScbError):
        pl.plan(wx=wx, wy=wy, G=G, param_mode=7, **lim)
    pl.close()


def test_oracle_matches_python_toppra_const_accel(oracle, golden_dir):
    """The oracle's ConstAccel against the reference's own Python ParametrizeConstAccel on the pinned
    parametrizations (tests/golden/gen_pytoppra_const_accel.py).  Knot accelerations are not compared: the two
    siblings pick different intervals exactly at a knot (see the generator's docstring)."""
    import json
    import os
    gold = json.load(open(os.path.join(golden_dir, "pytoppra_const_accel_golden.json")))
    G = gold["G"]
    assert len(gold["cases"]) == 4
    for c in gold["cases"]:
        ts_w = np.array(c["ts"])
        t_all = np.concatenate([ts_w, c["t_eval"]])
        ts, q, qd, qdd = oracle.const_accel_param(c["L"], c["vel_start"], c["vel_end"], np.linspace(0, 1, G), c["x"], t_all)
        np.testing.assert_allclose(ts, ts_w, rtol=1e-13, err_msg=c["name"])
        L = c["L"]
        np.testing.assert_allclose(q[:G], c["q_knots"], rtol=1e-12, atol=1e-13 * L, err_msg=c["name"])
        np.testing.assert_allclose(qd[:G], c["qd_knots"], rtol=1e-10, atol=1e-12 * L, err_msg=c["name"])
        np.testing.assert_allclose(q[G:], c["q"], rtol=1e-12, atol=1e-13 * L, err_msg=c["name"])
        np.testing.assert_allclose(qd[G:], c["qd"], rtol=1e-11, atol=1e-13 * L, err_msg=c["name"])
        np.testing.assert_allclose(qdd[G:], c["qdd"], rtol=1e-10, atol=1e-12 * np.abs(c["qdd"]).max(), err_msg=c["name"])
