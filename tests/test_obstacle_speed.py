"""Obstacle-distance speed limit (third "isolated constraint" of Lau09 III-B, eq. 2):
    |v_i| <= -a_brake t_react + sqrt(a_brake^2 t_react^2 + 2 a_brake d_i),
d_i = distance from the curve point of toppra gridpoint i to the closest obstacle edge.  The reference's hook for a
position-dependent limit is vel_lim_func (sea_current.hpp:1175-1188); here scb_batch_in.a_brake / t_react over the
obstacle_edges list.  CPU: the oracle against the formula written out in numpy.  GPU: the CUDA path against the
oracle at 1e-9 (FP64) / 1e-4 (FP32 storage) with identical verdicts."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def _seg_dist(px, py, edges):
    d = np.full(px.shape, np.inf)
    for x0, y0, x1, y1 in edges:
        ex, ey = x1 - x0, y1 - y0
        ee = ex * ex + ey * ey
        t = np.clip(((px - x0) * ex + (py - y0) * ey) / ee, 0, 1) if ee > 0 else 0.0
        d = np.minimum(d, np.hypot(px - x0 - t * ex, py - y0 - t * ey))
    return d


def _cap(d, ab, tr):
    return -ab * tr + np.sqrt((ab * tr) ** 2 + 2 * ab * d)


def _case(P, W, E, seed, clearance=0.4):
    """Random-walk paths and short wall pieces scattered beside them, none closer than `clearance` to a waypoint
    (so no edge crosses a tangent: the spline is the unobstructed one and only the speed limit acts)."""
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    edges = []
    while len(edges) < E:
        k = rng.integers(0, W); p = rng.integers(0, P)
        c = np.array([wx[p, k], wy[p, k]]) + rng.normal(0, 2.5, 2)
        d = rng.normal(0, 0.8, 2)
        e = np.concatenate([c - d, c + d])
        if _seg_dist(wx.ravel(), wy.ravel(), [e]).min() > clearance:
            edges.append(e)
    lim = dict(vel_lo=np.full(P, -2.0), vel_hi=np.full(P, 2.0), acc_lo=np.full(P, -1.0), acc_hi=np.full(P, 1.0))
    return wx, wy, np.asarray(edges), lim


def test_oracle_obeys_the_formula(oracle):
    wx, wy, edges, lim = _case(6, 6, 10, seed=2)
    G, ab, tr = 201, 0.8, 0.3
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, obstacle_edges=edges, **lim)
    got, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, obstacle_edges=edges, a_brake=ab, t_react=tr, **lim)
    assert (got["status"] == 0).all() and (free["status"] == 0).all()
    np.testing.assert_array_equal(got["ctrl_x"], free["ctrl_x"])  # the limit does not touch the geometry
    np.testing.assert_array_equal(got["px"], free["px"])
    c = _cap(_seg_dist(got["px"], got["py"], edges), ab, tr)
    sd = np.sqrt(got["x"])  # path speed at the gridpoints (|q'(s)| sdot with the arc length as the path coordinate)
    L = got["arclen_total"][None, :]
    s = np.linspace(0, 1, G)[:, None]
    qd = (6 * s * (1 - s)) * L  # q'(s) of the rest-to-rest Hermite path (vel_start = vel_end = 0)
    assert np.all(qd * sd <= c * (1 + 1e-9) + 1e-12)
    assert (c < 2.0).mean() > 0.2  # the walls do bind somewhere ...
    assert np.all(got["T_total"] >= free["T_total"] * (1 - 1e-12)) and np.any(got["T_total"] > free["T_total"] * 1.01)
    # a_brake = 0 or no edges: off
    off, _ = oracle.plan_batch(wx=wx, wy=wy, G=G, spline_mode=0, obstacle_edges=edges, a_brake=0.0, t_react=tr, **lim)
    np.testing.assert_array_equal(off["x"], free["x"])


def test_oracle_wall_beside_a_straight_path(oracle):
    """A straight path 0.5 beside a long wall: the cap is the same at every gridpoint, the plateau speed is that cap."""
    wx = np.array([[0.0, 10.0]]); wy = np.array([[0.0, 0.0]])
    lim = dict(vel_lo=[-5.0], vel_hi=[5.0], acc_lo=[-2.0], acc_hi=[2.0])
    ab, tr = 1.0, 0.5
    got, _ = oracle.plan_batch(wx=wx, wy=wy, G=401, spline_mode=0, obstacle_edges=np.array([[-5.0, 0.5, 15.0, 0.5]]), a_brake=ab,
                               t_react=tr, **lim)
    c = _cap(0.5, ab, tr)
    s = np.linspace(0, 1, 401)
    speed = 6 * s * (1 - s) * got["arclen_total"][0] * np.sqrt(got["x"][:, 0])  # q'(s) sdot at the gridpoints
    assert got["status"][0] == 0 and speed.max() <= c * (1 + 1e-9)
    assert np.median(speed) == pytest.approx(c, rel=1e-9)
    # the parametrizer's cubic spline through the knots overshoots the plateau by a fraction of a percent at its corners
    assert np.median(np.abs(got["v"][:, 0])) == pytest.approx(c, rel=1e-4) and np.abs(got["v"][:, 0]).max() <= c * 1.01


@pytest.fixture(scope="module")
def planners():
    made = {}

    def get(precision):
        if precision not in made:
            made[precision] = scb.BatchPlanner(0, precision, "ref_cubic", max_problems=4096, max_waypoints=64, max_gridpoints=1001)
        return made[precision]
    yield get
    for p in made.values():
        p.close()


def _rel(got, want, scale):
    return float(np.max(np.abs(got - want) / np.maximum(np.abs(want), scale)))


@pytest.mark.gpu
@pytest.mark.parametrize("lp_mode", [scb.abi.LP_CLOSED_FORM, scb.abi.LP_SEIDEL])
@pytest.mark.parametrize("W,P", [(8, 96), (40, 24)])
def test_gpu_matches_oracle_fp64(planners, oracle, lp_mode, W, P):
    wx, wy, edges, lim = _case(P, W, 24, seed=W)
    G, ab, tr = 301, 0.8, 0.3
    kw = dict(wx=wx, wy=wy, G=G, obstacle_edges=edges, a_brake=ab, t_react=tr, omega_max=1.5, **lim)  # with a curvature cap on top
    ref, _ = oracle.plan_batch(spline_mode=0, **kw)
    got = planners("fp64").plan(lp_mode=lp_mode, **kw)
    np.testing.assert_array_equal(got["status"], ref["status"])
    np.testing.assert_array_equal(got["fail_stage"], ref["fail_stage"])
    ok = ref["status"] == 0
    assert ok.mean() > 0.9
    for k in ("x", "K_lo", "K_hi"):
        assert _rel(got[k][:, ok], ref[k][:, ok], ref["x"][:, ok].max(axis=0) * 1e-3) <= 1e-9, k
    assert _rel(got["T_total"][ok], ref["T_total"][ok], 1e-300) <= 1e-9
    assert _rel(got["t"][:, ok], ref["t"][:, ok], ref["T_total"][ok] * 1e-3) <= 1e-9
    # the limit holds on the GPU's own output
    c = _cap(_seg_dist(got["px"], got["py"], edges), ab, tr)
    s = np.linspace(0, 1, G)[:, None]
    speed = 6 * s * (1 - s) * got["arclen_total"][None, :] * np.sqrt(got["x"])
    assert np.all(speed[:, ok] <= c[:, ok] * (1 + 1e-9) + 1e-12)
    # without px / py among the outputs the curve points live in scratch: same timing
    lean = planners("fp64").plan(lp_mode=lp_mode, outputs=("x", "T_total"), **kw)
    np.testing.assert_array_equal(lean["x"], got["x"])
    free = planners("fp64").plan(lp_mode=lp_mode, **dict(kw, a_brake=0.0))
    assert np.any(got["T_total"][ok] > free["T_total"][ok] * 1.01)


@pytest.mark.gpu
def test_gpu_fp32_storage_and_chunks_and_validation(planners, oracle):
    wx, wy, edges, lim = _case(700, 8, 24, seed=9)
    G, ab, tr = 201, 0.8, 0.3
    kw = dict(wx=wx, wy=wy, G=G, obstacle_edges=edges, a_brake=ab, t_react=tr, **lim)
    ref, _ = oracle.plan_batch(spline_mode=0, **kw)
    got = planners("fp32").plan(**kw)
    np.testing.assert_array_equal(got["status"], ref["status"])
    ok = ref["status"] == 0
    assert _rel(got["T_total"][ok], ref["T_total"][ok], 1e-300) <= 1e-4
    assert _rel(got["x"][:, ok], ref["x"][:, ok], ref["x"][:, ok].max(axis=0) * 1e-3) <= 1e-4
    # the pipelined host call (chunks of 256 problems) gives the same bits as one chunk
    pl = planners("fp64")
    one = pl.plan(**kw)
    try:
        pl.set_host_chunk(256)
        chunked = pl.plan(**kw)
    finally:
        pl.set_host_chunk(0)
    for k in ("status", "x", "t", "v", "T_total"):
        np.testing.assert_array_equal(chunked[k], one[k], err_msg=k)
    assert pl.timing()["launches"] == 5 * 3  # path, arc, obstacle cap, toppra, parametrizer per chunk
    with pytest.raises(scb.ScbError):  # the limit needs the geometry stage
        pl.plan(stages=scb.abi.STAGE_SPLINE | scb.abi.STAGE_TIMING, **kw)
    with pytest.raises(scb.ScbError):
        pl.plan(**dict(kw, a_brake=-1.0))
