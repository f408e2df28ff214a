"""Curvature-aware velocity limits (scb_batch_in.omega_max / acc_lat_max).

The reference's hook for position-dependent limits is vel_lim_func (sea_current.hpp:1175-1188, used in
examples/test.cpp:194-209); Lau09 III-B names the two caps built here: |omega| <= omega_max and
v^2 |kappa| <= acc_lat_max, both as a velocity box per toppra gridpoint."""
import numpy as np
import pytest

import seacurrent_b200 as scb

OM, AL = 0.6, 0.4


def _problem(P=48, W=8, seed=3):
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.0)
    return wx, wy, lim


def _gridpoint_speed_and_kappa(res, ok):
    """Path speed q'(s_i) sqrt(x_i) of the toppra solution and the curvature omega_i / v_i at the gridpoints."""
    G = res["x"].shape[0]
    s = np.linspace(0.0, 1.0, G)[:, None]
    L = res["arclen_total"][None, ok]
    speed = 6.0 * L * s * (1.0 - s) * np.sqrt(np.maximum(res["x"][:, ok], 0.0))  # vel_start = vel_end = 0
    v = res["v"][:, ok]
    kappa = np.where(np.abs(v) > 1e-6 * np.abs(v).max(axis=0, keepdims=True), res["omega"][:, ok] / np.where(v == 0, 1, v), 0.0)
    return speed, np.abs(kappa)


def test_oracle_caps_hold_and_slow_the_trajectory(oracle):
    wx, wy, lim = _problem()
    G = 201
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=1, **lim)
    cap, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=1, omega_max=OM, acc_lat_max=AL, **lim)
    ok = (free["status"] == 0) & (cap["status"] == 0)
    assert ok.sum() >= 40
    speed, kappa = _gridpoint_speed_and_kappa(cap, ok)
    assert np.all(speed * kappa <= OM * (1 + 1e-6) + 1e-9)
    assert np.all(speed * speed * kappa <= AL * (1 + 1e-6) + 1e-9)
    # the caps bind somewhere on a random walk with 100-degree turns, and never make a trajectory faster
    assert np.all(cap["T_total"][ok] >= free["T_total"][ok] * (1 - 1e-12))
    assert np.mean(cap["T_total"][ok] > free["T_total"][ok] * 1.01) > 0.5
    # the geometry is untouched
    np.testing.assert_array_equal(cap["ctrl_x"], free["ctrl_x"])
    np.testing.assert_array_equal(cap["px"][:, ok], free["px"][:, ok])


@pytest.mark.gpu
@pytest.mark.parametrize("lp_mode", [scb.abi.LP_CLOSED_FORM, scb.abi.LP_SEIDEL])
@pytest.mark.parametrize("caps", [(OM, 0.0), (0.0, AL), (OM, AL)])
def test_gpu_matches_oracle_with_caps(oracle, lp_mode, caps):
    wx, wy, lim = _problem(P=160, W=10, seed=11)
    G = 300
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=256, max_waypoints=16, max_gridpoints=512)
    res = pl.plan(wx=wx, wy=wy, G=G, lp_mode=lp_mode, omega_max=caps[0], acc_lat_max=caps[1], **lim)
    pl.close()
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=1, omega_max=caps[0], acc_lat_max=caps[1], **lim)
    np.testing.assert_array_equal(res["status"], ref["status"])
    np.testing.assert_array_equal(res["fail_stage"], ref["fail_stage"])
    ok = ref["status"] == 0
    assert ok.sum() >= 100
    for n in ("K_lo", "K_hi", "x", "t"):
        g, w = res[n][:, ok], ref[n][:, ok]
        # where a cap binds, x = (c / q')^2 carries the curvature's own conditioning (d kappa / d q ~ kappa^2 at a spike,
        # arc position known to ~64 eps L: see test_gpu_parity.fp32_omega_tol), hence the per-problem floor
        scale = np.abs(ref["K_hi" if n == "K_lo" else n][:, ok]).max(axis=0, keepdims=True)  # K_lo is mostly 0
        tol = 1e-9 * np.maximum(np.abs(w), 1e-3 * scale) * 4
        assert np.all(np.abs(g - w) <= tol), (n, float(np.max(np.abs(g - w) / tol)))
    assert np.max(np.abs(res["T_total"][ok] / ref["T_total"][ok] - 1)) <= 1e-9


@pytest.mark.gpu
def test_gpu_caps_hold_at_c2_size():
    w = scb.workloads.config_c2()
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=4096, max_waypoints=20, max_gridpoints=500)
    free = pl.plan(**w)
    cap = pl.plan(omega_max=OM, acc_lat_max=AL, **w)
    pl.close()
    ok = (free["status"] == 0) & (cap["status"] == 0)
    assert ok.mean() > 0.95
    speed, kappa = _gridpoint_speed_and_kappa(cap, ok)
    assert np.all(speed * kappa <= OM * (1 + 1e-6) + 1e-9)
    assert np.all(speed * speed * kappa <= AL * (1 + 1e-6) + 1e-9)
    assert np.all(cap["T_total"][ok] >= free["T_total"][ok] * (1 - 1e-12))
    # invalid combination: caps without the geometry stage
    with pytest.raises(scb.ScbError):
        pl2 = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=64, max_waypoints=20, max_gridpoints=500)
        try:
            pl2.plan(path_len=np.full(4, 10.0), G=50, stages=scb.abi.STAGE_TIMING, vel_lo=[-1] * 4, vel_hi=[1] * 4,
                     acc_lo=[-1] * 4, acc_hi=[1] * 4, omega_max=1.0, outputs=("x",))
        finally:
            pl2.close()
