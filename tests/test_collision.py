"""Collision term (SURVEY 8f N3, Lau09 IV-B: "if the trajectory is not collision-free according to the obstacle map, its
cost is set to infinity"): scb_batch_in.collision_chords = M cuts every segment into M chords and tests each against
the obstacle edges with the reference's segment test (intersects, sea_current.hpp:142-178).
  * verdict on the finished spline (both constructions): a hit -> status SCB_ST_ERR_COLLISION (19), cost = +inf, the
    problem is not timed;
  * inside the quintic optimiser's loop: a colliding trial iterate has infinite cost and is REJECTED (back to the best
    collision-free iterate, steps halved), so the optimiser keeps its best collision-free iterate instead of walking
    through a wall.
CPU: the oracle against a numpy restatement, and what the in-loop term buys.  GPU: verdicts identical to the oracle's,
control points and timing of the free problems at 1e-9, both spline constructions."""
import numpy as np
import pytest

import seacurrent_b200 as scb

COLLISION = 19


def _hit(l0, l1, k0, k1):
    dl, dk, e = l1 - l0, k1 - k0, k0 - l0
    b = dl[0] * dk[1] - dl[1] * dk[0]
    if b == 0:
        return False
    u = (e[0] * dl[1] - e[1] * dl[0]) / b
    t = (e[0] * dk[1] - e[1] * dk[0]) / b
    return bool(0 <= u <= 1 and 0 <= t <= 1)


def _spline_hits(cx, cy, nc, edges, M):
    """cx, cy: (W-1)*nc control points of one problem."""
    for s in range(len(cx) // nc):
        pts = np.stack([cx[s * nc:(s + 1) * nc], cy[s * nc:(s + 1) * nc]], 1)
        a = pts[0]
        for k in range(1, M + 1):
            u = 1.0 if k == M else k / M
            t = pts.copy()
            for r in range(1, nc):
                t[:nc - r] = (1 - u) * t[:nc - r] + u * t[1:nc - r + 1]
            if any(_hit(a, t[0], e[:2], e[2:]) for e in edges):
                return True
            a = t[0].copy()
    return False


def _case(P, W, E, seed):
    """Random-walk paths with short wall pieces strewn around their legs (midpoints of random legs, offset ~N(0, 0.7)):
    roughly a third of the paths end up crossing one."""
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    p = rng.integers(0, P, E); k = rng.integers(0, W - 1, E)
    c = np.stack([(wx[p, k] + wx[p, k + 1]) / 2, (wy[p, k] + wy[p, k + 1]) / 2], 1) + rng.normal(0, 0.7, (E, 2))
    d = rng.normal(0, 0.5, (E, 2))
    lim = dict(vel_lo=np.full(P, -2.0), vel_hi=np.full(P, 2.0), acc_lo=np.full(P, -1.0), acc_hi=np.full(P, 1.0))
    return wx, wy, np.concatenate([c - d, c + d], 1), lim


def test_oracle_hand_case(oracle):
    wx = np.array([[0.0, 10.0, 10.0]]); wy = np.array([[0.0, 0.0, 10.0]])
    lim = dict(vel_lo=[-1.0], vel_hi=[1.0], acc_lo=[-0.5], acc_hi=[0.5])
    for mode in (1,):  # the quintic takes an edge list for the verdict only (no tangent clipping)
        free, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=mode, **lim)
        wall = np.array([[3.0, -1.0, 3.0, 1.0]])  # across the first leg
        hit, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=mode, obstacle_edges=wall, collision_chords=16, **lim)
        assert hit["status"][0] == COLLISION and np.isinf(hit["cost"][0]) and hit["T_total"][0] == 0
        # the wall cuts the leg in two: every iterate collides, the best-so-far stays the initial spline
        start, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=mode, opt=scb.default_opt(iters=0), **lim)
        np.testing.assert_array_equal(hit["ctrl_x"], start["ctrl_x"])
        away = np.array([[3.0, 5.0, 3.0, 7.0]])
        ok, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=mode, obstacle_edges=away, collision_chords=16, **lim)
        assert ok["status"][0] == 0 and ok["T_total"][0] == free["T_total"][0] and ok["cost"][0] == free["cost"][0]
    # the cubic: the same wall first clips the start tangent back to it (shrink_tangent), and the curve still crosses it
    hit, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=0, obstacle_edges=np.array([[3.0, -1.0, 3.0, 1.0]]), collision_chords=16, **lim)
    assert hit["status"][0] == COLLISION and hit["ctrl_x"][0][1] == pytest.approx(3.0)
    off, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=0, obstacle_edges=np.array([[3.0, -1.0, 3.0, 1.0]]), **lim)
    assert off["status"][0] == 0  # collision_chords = 0: no verdict


@pytest.mark.parametrize("mode,nc", [(0, 4), (1, 6)])
def test_oracle_vs_numpy_restatement(oracle, mode, nc):
    wx, wy, edges, lim = _case(40, 6, 30, seed=4 + mode)
    M = 12
    got, _ = oracle.plan_batch(wx=wx, wy=wy, G=51, spline_mode=mode, obstacle_edges=edges, collision_chords=M, **lim)
    # control points of the same construction: with the cubic the edges also clip tangents, so take them from this run
    want = np.array([_spline_hits(got["ctrl_x"][p], got["ctrl_y"][p], nc, edges, M) for p in range(40)])
    np.testing.assert_array_equal(got["status"] == COLLISION, want)
    assert 4 <= want.sum() <= 36
    assert np.all(np.isinf(got["cost"][want])) and np.all(got["T_total"][want] == 0) and np.all(got["T_total"][~want] > 0)


def test_optimiser_keeps_its_best_collision_free_iterate(oracle):
    """What the term inside the loop buys (Lau09 IV-B): walls beside the legs, close enough that the unconstrained
    optimiser swings some splines through them.  With the term, a problem whose INITIAL spline is collision-free can no
    longer end in a collision, so the colliding fraction drops to the problems that collide from the start."""
    P, W, M = 160, 8, 12
    rng = np.random.default_rng(77)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    # E wall pieces, each 0.2 .. 1.2 m beside the middle of a random leg of a random problem and parallel to it (the paths
    # all leave the origin, so every piece lies near many of them); one list shared by the batch
    E = 14
    k = rng.integers(1, W - 2, E); p = rng.integers(0, P, E)
    mid = np.stack([(wx[p, k] + wx[p, k + 1]) / 2, (wy[p, k] + wy[p, k + 1]) / 2], 1)
    leg = np.stack([wx[p, k + 1] - wx[p, k], wy[p, k + 1] - wy[p, k]], 1)
    leg /= np.linalg.norm(leg, axis=1, keepdims=True)
    nrm = np.stack([-leg[:, 1], leg[:, 0]], 1)
    c = mid + nrm * (rng.uniform(0.2, 1.2, (E, 1)) * rng.choice([-1.0, 1.0], (E, 1)))
    edges = np.concatenate([c - 0.6 * leg, c + 0.6 * leg], 1)
    lim = dict(vel_lo=np.full(P, -2.0), vel_hi=np.full(P, 2.0), acc_lo=np.full(P, -1.0), acc_hi=np.full(P, 1.0))
    opt = scb.default_opt(move_waypoints=1)
    kw = dict(wx=wx, wy=wy, G=51, spline_mode=1, **lim)
    start, _ = oracle.plan_batch(opt=scb.default_opt(iters=0), **kw)           # the initial splines
    blind, _ = oracle.plan_batch(opt=opt, **kw)                                 # optimiser that knows no obstacles
    aware, _ = oracle.plan_batch(opt=opt, obstacle_edges=edges, collision_chords=M, **kw)
    hits = lambda r: np.array([_spline_hits(r["ctrl_x"][q], r["ctrl_y"][q], 6, edges, M) for q in range(P)])
    h0, hb, ha = hits(start), hits(blind), hits(aware)
    np.testing.assert_array_equal(aware["status"] == COLLISION, ha)      # the verdict is the finished spline's
    assert not np.any(ha & ~h0)                                          # free at the start -> free at the end
    assert (hb & ~h0).sum() >= 3                                         # ... which the blind optimiser does not give
    assert ha.sum() < hb.sum()
    # where neither run ever met a wall the term changes nothing; elsewhere the kept iterate costs at least the blind optimum's
    same = ~hb & ~ha & (aware["cost"] == blind["cost"])
    print("colliding: initial %d, blind optimiser %d, with the term %d of %d; untouched by the term: %d" % (h0.sum(), hb.sum(), ha.sum(), P, same.sum()))
    assert same.sum() >= P // 4
    np.testing.assert_array_equal(aware["ctrl_x"][same], blind["ctrl_x"][same])
    free = ~ha
    assert np.all(aware["cost"][free] <= start["cost"][free] * (1 + 1e-12))    # best-so-far never worse than the start


@pytest.fixture(scope="module")
def planners():
    made = {}

    def get(precision, mode):
        if (precision, mode) not in made:
            made[(precision, mode)] = scb.BatchPlanner(0, precision, mode, max_problems=4096, max_waypoints=64, max_gridpoints=1001)
        return made[(precision, mode)]
    yield get
    for p in made.values():
        p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mode,W,P", [("ref_cubic", 8, 300), ("ref_cubic", 40, 40), ("lau_quintic", 10, 300), ("lau_quintic", 40, 24)])
def test_gpu_verdicts_match_oracle(planners, oracle, mode, W, P):
    wx, wy, edges, lim = _case(P, W, 16, seed=W + P)
    kw = dict(wx=wx, wy=wy, G=201, obstacle_edges=edges, collision_chords=16, **lim)
    ref, _ = oracle.plan_batch(spline_mode=0 if mode == "ref_cubic" else 1, **kw)
    got = planners("fp64", mode).plan(**kw)
    np.testing.assert_array_equal(got["status"], ref["status"])
    hit = ref["status"] == COLLISION
    ok = ref["status"] == 0
    # the cubic has no optimiser: what collides stays colliding; the quintic's optimiser steers most candidates clear
    assert hit.sum() >= (P // 8 if mode == "ref_cubic" else 2) and ok.sum() >= P // 2
    assert np.all(np.isinf(got["cost"][hit])) and np.all(got["T_total"][hit] == 0)
    np.testing.assert_allclose(got["T_total"][ok], ref["T_total"][ok], rtol=1e-9)
    np.testing.assert_allclose(got["arclen_total"][ok], ref["arclen_total"][ok], rtol=1e-9)
    np.testing.assert_allclose(got["cost"][ok], ref["cost"][ok], rtol=1e-9)
    for k in ("ctrl_x", "ctrl_y"):  # in quintic mode the rejected trials steer the optimiser: same iterates, same control points
        sc = np.abs(ref[k][ok]).max(axis=1, keepdims=True)
        assert np.max(np.abs(got[k][ok] - ref[k][ok]) / np.maximum(np.abs(ref[k][ok]), 1e-3 * sc)) <= 1e-9, k
    # a summary-ranking outer loop sees the colliding candidates as infinitely dear
    assert np.all(got["cost"][ok] < np.inf)
    off = planners("fp64", mode).plan(**dict(kw, collision_chords=0)) if mode == "ref_cubic" else None
    if off is not None:
        assert (off["status"] == COLLISION).sum() == 0


@pytest.mark.gpu
def test_gpu_fp32_storage_and_validation(planners, oracle):
    wx, wy, edges, lim = _case(400, 8, 16, seed=21)
    kw = dict(wx=wx, wy=wy, G=101, obstacle_edges=edges, collision_chords=16, **lim)
    ref, _ = oracle.plan_batch(spline_mode=0, **kw)
    got = planners("fp32", "ref_cubic").plan(**kw)
    # float control points: a grazing chord may fall on the other side of an edge
    assert (got["status"] == ref["status"]).mean() >= 0.99
    pl = planners("fp64", "lau_quintic")
    with pytest.raises(scb.ScbError):  # the quintic takes an edge list only for the verdict
        pl.plan(**dict(kw, collision_chords=0))
    with pytest.raises(scb.ScbError):
        pl.plan(**dict(kw, collision_chords=-1))
