"""Obstacle-aware tangent clipping (reference part): shrink_tangent, sea_current.hpp:575-596, with
intersects, sea_current.hpp:142-178.  from_path cuts every tangent back where W -> W + T or W -> W - T crosses an
obstacle edge, so that the cubic's control polygon (and with it the curve) stays on the waypoint's side."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def _cross_t(l0, l1, k0, k1):
    """Parameter t along l0 -> l1 at which it crosses k0 -> k1, or None (same algebra as the reference)."""
    dl, dk, e = l1 - l0, k1 - k0, k0 - l0
    b = dl[0] * dk[1] - dl[1] * dk[0]
    if b == 0:
        return None
    u = (e[0] * dl[1] - e[1] * dl[0]) / b
    t = (e[0] * dk[1] - e[1] * dk[0]) / b
    return t if (0 <= u <= 1 and 0 <= t <= 1) else None


def _random_case(P, W, E, seed):
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    # short edges scattered around the waypoints of the first paths: many cross a tangent, none (by rejection) touches
    # a waypoint itself
    k = rng.integers(0, W, size=E); p = rng.integers(0, P, size=E)
    c = np.stack([wx[p, k], wy[p, k]], axis=1) + rng.normal(0, 1.5, size=(E, 2))
    d = rng.normal(0, 1.0, size=(E, 2))
    edges = np.concatenate([c - d, c + d], axis=1)
    return wx, wy, edges


def test_oracle_hand_case(oracle):
    wx = np.array([[0.0, 10.0, 10.0]]); wy = np.array([[0.0, 0.0, 10.0]])
    lim = dict(vel_lo=[-1.0], vel_hi=[1.0], acc_lo=[-0.5], acc_hi=[0.5])
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=21, spline_mode=0, **lim)
    T1 = np.array([free["ctrl_x"][0][4 + 1] - 10.0, free["ctrl_y"][0][4 + 1] - 0.0])  # P1 of segment 1 = W1 + T1
    assert np.hypot(*T1) == pytest.approx(5.0)  # half the shorter leg
    # an edge across W1 -> W1 + T1 at 40 % of the tangent
    mid = np.array([10.0, 0.0]) + 0.4 * T1
    n = np.array([-T1[1], T1[0]]) / np.hypot(*T1)
    edge = np.concatenate([mid - 2 * n, mid + 2 * n])[None, :]
    clip, _ = oracle.plan_batch(wx=wx, wy=wy, G=21, spline_mode=0, obstacle_edges=edge, **lim)
    assert clip["status"][0] == 0
    T1c = np.array([clip["ctrl_x"][0][5] - 10.0, clip["ctrl_y"][0][5]])
    np.testing.assert_allclose(T1c, 0.4 * T1, rtol=1e-12, atol=1e-12)
    # the same tangent ends segment 0: its P2 = W1 - T1
    np.testing.assert_allclose([clip["ctrl_x"][0][2], clip["ctrl_y"][0][2]], np.array([10.0, 0.0]) - 0.4 * T1, atol=1e-12)
    # the end tangents are untouched, the path got shorter
    np.testing.assert_array_equal(clip["ctrl_x"][0][[0, 1, 6, 7]], free["ctrl_x"][0][[0, 1, 6, 7]])
    assert clip["arclen_total"][0] < free["arclen_total"][0]


def test_oracle_no_tangent_crosses_an_edge(oracle):
    P, W, E = 64, 8, 48
    wx, wy, edges = _random_case(P, W, E, seed=2)
    lim = scb.workloads.random_limits(np.random.default_rng(0), P, slow_fraction=0.0)
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=11, spline_mode=0, **lim)
    clip, _ = oracle.plan_batch(wx=wx, wy=wy, G=11, spline_mode=0, obstacle_edges=edges, **lim)
    cx, cy = clip["ctrl_x"].reshape(P, W - 1, 4), clip["ctrl_y"].reshape(P, W - 1, 4)
    changed = 0
    for p in range(P):
        for i in range(W):
            Wp = np.array([wx[p, i], wy[p, i]])
            T = (np.array([cx[p, i, 1], cy[p, i, 1]]) - Wp) if i < W - 1 else (Wp - np.array([cx[p, i - 1, 2], cy[p, i - 1, 2]]))
            for e in edges:
                for sgn in (1.0, -1.0):
                    t = _cross_t(Wp + sgn * T, Wp, e[:2], e[2:])
                    assert t is None or t <= 1e-9, (p, i, t)  # only the clipped end may touch an edge
    changed = np.sum(np.abs(clip["ctrl_x"] - free["ctrl_x"]) > 1e-9)
    assert changed > 20  # the case does exercise the clipping


@pytest.mark.gpu
@pytest.mark.parametrize("W,P", [(10, 300), (40, 24)])
def test_gpu_matches_oracle_with_obstacles(oracle, W, P):
    wx, wy, edges = _random_case(P, W, 64, seed=7 + W)
    lim = scb.workloads.random_limits(np.random.default_rng(1), P, slow_fraction=0.0)
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=64, spline_mode=0, obstacle_edges=edges, **lim)
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=64, spline_mode=0, **lim)
    assert np.sum(np.abs(ref["ctrl_x"] - free["ctrl_x"]) > 1e-9) > 20
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=P, max_waypoints=W, max_gridpoints=64)
    res = pl.plan(wx=wx, wy=wy, G=64, obstacle_edges=edges, **lim)
    pl.close()
    np.testing.assert_array_equal(res["status"], ref["status"])
    ok = ref["status"] == 0
    scale = float(np.max(np.abs(ref["ctrl_x"])))
    for n in ("ctrl_x", "ctrl_y"):
        assert np.max(np.abs(res[n][ok] - ref[n][ok])) <= 1e-9 * scale, n
    for n in ("arclen_total", "T_total"):
        assert np.max(np.abs(res[n][ok] / ref[n][ok] - 1)) <= 1e-9, n


@pytest.mark.gpu
def test_gpu_obstacles_fp32_storage_and_rejections(oracle):
    P, W = 128, 10
    wx, wy, edges = _random_case(P, W, 64, seed=3)
    f32 = lambda a: a.astype(np.float32).astype(np.float64)  # identical inputs on both sides
    wx, wy, edges = f32(wx), f32(wy), f32(edges)
    lim = scb.workloads.random_limits(np.random.default_rng(1), P, slow_fraction=0.0)
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=32, spline_mode=0, obstacle_edges=edges, **lim)
    pl = scb.BatchPlanner(0, "fp32", "ref_cubic", max_problems=P, max_waypoints=W, max_gridpoints=32)
    res = pl.plan(wx=wx, wy=wy, G=32, obstacle_edges=edges, **lim)
    pl.close()
    np.testing.assert_array_equal(res["status"], ref["status"])
    ok = ref["status"] == 0
    scale = float(np.max(np.abs(ref["ctrl_x"])))
    assert np.max(np.abs(res["ctrl_x"][ok] - ref["ctrl_x"][ok])) <= 1e-6 * scale
    assert np.max(np.abs(res["ctrl_y"][ok] - ref["ctrl_y"][ok])) <= 1e-6 * scale
    # the clipping argument is the cubic's: the quintic mode refuses obstacles
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=W, max_gridpoints=32)
    try:
        with pytest.raises(scb.ScbError):
            pl.plan(wx=wx, wy=wy, G=32, obstacle_edges=edges, **lim)
    finally:
        pl.close()
