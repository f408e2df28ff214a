"""Property tests of the stages that have NO reference implementation (parity
unpinned): Lau09 quintic builder, curvature/arc-length cost + analytic gradient, iRPROP- loop.
CPU only (oracle)."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def _path(seed, W):
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, 1, W)
    return np.stack([wx[0], wy[0]], axis=1)


def _bez(ctrl, u, order=0):
    from math import comb
    P = np.array(ctrl)
    for _ in range(order):
        P = (len(P) - 1) * (P[1:] - P[:-1])
    n = len(P) - 1
    return sum(comb(n, k) * u ** k * (1 - u) ** (n - k) * P[k] for k in range(n + 1))


@pytest.mark.parametrize("W", [3, 6, 12])
def test_quintic_interpolates_and_is_c2(oracle, W):
    path = _path(W, W)
    e = np.linspace(0.6, 1.4, W)
    J, ctrl, ge, gW = oracle.lau_eval(path, path, e, 0.05)
    assert np.isfinite(J)
    for s in range(W - 1):
        np.testing.assert_allclose(ctrl[s, 0], path[s], atol=1e-12)
        np.testing.assert_allclose(ctrl[s, 5], path[s + 1], atol=1e-12)
    for s in range(W - 2):  # C1 and C2 at the joins
        np.testing.assert_allclose(_bez(ctrl[s], 1.0, 1), _bez(ctrl[s + 1], 0.0, 1), atol=1e-10)
        np.testing.assert_allclose(_bez(ctrl[s], 1.0, 2), _bez(ctrl[s + 1], 0.0, 2), atol=1e-9)


def test_gradient_matches_central_differences(oracle):
    W = 7
    path = _path(3, W)
    rng = np.random.default_rng(1)
    e = rng.uniform(0.7, 1.3, W)
    Wp = path.copy(); Wp[1:-1] += rng.normal(0, 0.05, (W - 2, 2))
    J, _, ge, gW = oracle.lau_eval(path, Wp, e, 0.05)
    h = 1e-6
    for i in range(W):
        ep, em = e.copy(), e.copy(); ep[i] += h; em[i] -= h
        fd = (oracle.lau_eval(path, Wp, ep, 0.05)[0] - oracle.lau_eval(path, Wp, em, 0.05)[0]) / (2 * h)
        assert abs(fd - ge[i]) <= 1e-6 * max(1.0, abs(fd)), (i, fd, ge[i])
    for i in range(1, W - 1):
        for k in range(2):
            a, b = Wp.copy(), Wp.copy(); a[i, k] += h; b[i, k] -= h
            fd = (oracle.lau_eval(path, a, e, 0.05)[0] - oracle.lau_eval(path, b, e, 0.05)[0]) / (2 * h)
            assert abs(fd - gW[i, k]) <= 1e-6 * max(1.0, abs(fd)), (i, k, fd, gW[i, k])


def test_cost_invariant_under_rigid_motion(oracle):
    W = 6
    path = _path(5, W)
    e = np.ones(W)
    J0 = oracle.lau_eval(path, path, e, 0.05)[0]
    th = 0.7
    R = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    p2 = path @ R.T + np.array([3.0, -2.0])
    J1 = oracle.lau_eval(p2, p2, e, 0.05)[0]
    assert abs(J0 - J1) <= 1e-9 * abs(J0)


@pytest.mark.parametrize("move", [0, 1])
def test_optimiser_best_cost_monotone(oracle, move):
    W = 8
    path = _path(11, W)
    opt = scb.default_opt(iters=40, move_waypoints=move)
    J, ctrl, e, hist = oracle.lau_optimise(path, opt)
    assert J <= hist[0] + 1e-15
    assert abs(J - hist.min()) <= 1e-15 * abs(J)
    assert J < 0.999 * hist[0]  # it actually improves
    assert np.all(e >= opt.e_min - 1e-15) and np.all(e <= opt.e_max + 1e-15)
    if not move:  # the returned (e, ctrl) reproduce the returned cost
        Jc, ctrl2, _, _ = oracle.lau_eval(path, path, e, opt.lambda_)
        assert abs(Jc - J) <= 1e-12 * abs(J)
        np.testing.assert_allclose(ctrl2, ctrl, atol=1e-12)
