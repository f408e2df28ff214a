"""The rows of SURVEY 8(f) chained the way examples/test.cpp:250-300 chains them, batched: FMT* over the sampled map
(N4) -> its waypoints, perturbed into MPC candidate rollouts -> quintic spline + optimiser -> collision verdict (N3) ->
toppra under the curvature and obstacle-distance limits (N2) -> the outer loop keeps the fastest collision-free
candidate.  Checked against the oracle on a subset and through properties of the winner."""
import numpy as np
import pytest

import seacurrent_b200 as scb
from test_collision import COLLISION, _spline_hits
from test_fmt import _map

pytestmark = pytest.mark.gpu


def test_fmt_to_ranked_rollouts(oracle):
    obs, nodes, edges = _map(oracle, 400)
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=4096, max_waypoints=64, max_gridpoints=401)
    try:
        route = pl.fmt_batch(nodes, edges, [(-0.5, 1.0)], [(1.0, -1.0)], 0.45)
        assert route["status"][0] == 0
        W = int(route["path_len"][0])
        assert 4 <= W <= 32
        way = np.stack([route["path_x"][0, :W], route["path_y"][0, :W]], 1).astype(np.float64)
        # MPC candidates: the route itself (rollout 0) and perturbations of its inner waypoints
        P = 2048
        rng = np.random.default_rng(11)
        wx = np.repeat(way[None, :, 0], P, 0); wy = np.repeat(way[None, :, 1], P, 0)
        wx[1:, 1:-1] += rng.normal(0, 0.03, (P - 1, W - 2)); wy[1:, 1:-1] += rng.normal(0, 0.03, (P - 1, W - 2))
        lim = dict(vel_lo=np.full(P, -0.6), vel_hi=np.full(P, 0.6), acc_lo=np.full(P, -0.5), acc_hi=np.full(P, 0.5))
        opt = scb.default_opt(move_waypoints=1, w_radius=0.02)
        kw = dict(G=201, obstacle_edges=edges.astype(np.float64), collision_chords=16, a_brake=0.5, t_react=0.1, omega_max=2.0, opt=opt, **lim)
        res = pl.plan(wx=wx, wy=wy, **kw)
        ok = res["status"] == 0
        hit = res["status"] == COLLISION
        assert ok.sum() >= P // 10 and (ok | hit).all()
        assert np.all(np.isinf(res["cost"][hit])) and np.all(np.isfinite(res["cost"][ok]))
        # the outer loop: fastest collision-free candidate
        T = np.where(ok, res["T_total"], np.inf)
        best = int(np.argmin(T))
        assert not _spline_hits(res["ctrl_x"][best], res["ctrl_y"][best], 6, edges, 64)  # also at 4x the verdict's resolution
        if ok[0]:
            assert T[best] <= T[0]
        assert T[best] < np.median(T[ok])
        # limits hold along the winner, for the path speed toppra fixes at the gridpoints (the parametrizer's cubic
        # spline through the knots overshoots it by a few percent where the limit changes quickly)
        d = np.full(201, np.inf)
        for x0, y0, x1, y1 in edges:
            ex, ey = x1 - x0, y1 - y0
            t = np.clip(((res["px"][:, best] - x0) * ex + (res["py"][:, best] - y0) * ey) / (ex * ex + ey * ey), 0, 1)
            d = np.minimum(d, np.hypot(res["px"][:, best] - x0 - t * ex, res["py"][:, best] - y0 - t * ey))
        cap = -0.5 * 0.1 + np.sqrt((0.5 * 0.1) ** 2 + 2 * 0.5 * d)
        s = np.linspace(0, 1, 201)
        speed = 6 * s * (1 - s) * res["arclen_total"][best] * np.sqrt(res["x"][:, best])
        assert np.all(speed <= cap * (1 + 1e-9) + 1e-12)
        v = res["v"][:, best]
        kappa = np.divide(res["omega"][:, best], v, out=np.zeros(201), where=np.abs(v) > 1e-9)
        assert np.all(speed * np.abs(kappa) <= 2.0 * (1 + 1e-9)) and np.abs(res["omega"][:, best]).max() <= 2.0 * 1.15
        # the oracle on the first candidates: same verdicts, same durations
        n = 24
        ref, _ = oracle.plan_batch(wx=wx[:n], wy=wy[:n], spline_mode=1, **dict(kw, **{k: v[:n] for k, v in lim.items()}))
        np.testing.assert_array_equal(res["status"][:n], ref["status"])
        o = ref["status"] == 0
        np.testing.assert_allclose(res["T_total"][:n][o], ref["T_total"][o], rtol=1e-9)
    finally:
        pl.close()
