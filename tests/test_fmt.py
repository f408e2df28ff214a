"""Global planner front end: Halton sampling of the free space (sea_current.hpp:100-132, 1294-1313),
obstacle containment (sc:200-250) and FMT* (sc:1339-1407 with cost sc:1315-1326, near sc:1328-1337).

CPU: the oracle against known answers (the van der Corput values, hand-built maps) and against an independent
set-based Python restatement of the reference's loop.  GPU: scb_fmt_batch against the oracle, bit for bit (node
sequence of every path, float cost, expansion count, verdict)."""
import numpy as np
import pytest

import seacurrent_b200 as scb

F32_MAX = np.float32(np.finfo(np.float32).max)


# ------------------------------------------------------------------ independent restatement (sets and dicts, like the reference)
def _pt_dist(u, v):
    dx = float(np.float32(v[0]) - np.float32(u[0])); dy = float(np.float32(v[1]) - np.float32(u[1]))
    return np.float32(np.sqrt(dx * dx + dy * dy))


def _intersects(l0, l1, k0, k1):
    f = np.float32
    ex, ey, dlx, dly, dkx, dky = f(k0[0]) - f(l0[0]), f(k0[1]) - f(l0[1]), f(l1[0]) - f(l0[0]), f(l1[1]) - f(l0[1]), f(k1[0]) - f(k0[0]), f(k1[1]) - f(k0[1])
    a = f(f(ex * dly) - f(ey * dlx)); b = f(f(dlx * dky) - f(dly * dkx))
    if b == 0:
        return False
    u = f(a / b); c = f(f(ex * dky) - f(ey * dkx)); t = f(c / b)
    return bool(0 <= u <= 1 and 0 <= t <= 1)


def _fmt_sets(samples, edges, init, goal, rn):
    """fast_marching_trees with Python sets keyed by the point (ties: lowest insertion index, as the oracle)."""
    init = tuple(np.float32(init)); goal = tuple(np.float32(goal))
    pts = [init] + [tuple(p) for p in np.float32(samples) if tuple(p) != init and tuple(p) != goal] + [goal]
    order = {p: i for i, p in enumerate(pts)}
    rn2 = float(np.float32(rn)) ** 2

    def cost(a, b):
        for e in edges:
            if _intersects(a, b, e[:2], e[2:]):
                return F32_MAX
        return _pt_dist(a, b)

    def near(b, nodes):
        return sorted((a for a in nodes if float(_pt_dist(a, b)) <= rn2 and a != b), key=order.get)
    unvisited = set(pts[1:]); open_ = {init}; closed = set()
    cmap = {init: np.float32(0)}; parent = {init: init}
    z = init
    with np.errstate(over="ignore"):
        while z != goal:
            new = set()
            for x in near(z, unvisited):
                ys = near(x, open_)
                if not ys:
                    continue
                y_min = ys[0]
                for y in ys:
                    if np.float32(cmap[y] + cost(x, y)) < np.float32(cmap[y_min] + cost(x, y_min)):
                        y_min = y
                if cost(x, y_min) != F32_MAX:
                    parent[x] = y_min; new.add(x); cmap[x] = np.float32(cmap[y_min] + cost(x, y_min))
            unvisited -= new
            open_ |= new; open_.discard(z); closed.add(z)
            if not open_:
                return None, None, len(closed)
            z = min(sorted(open_, key=order.get), key=lambda p: cmap[p])
    path = [z]
    while path[-1] != init:
        path.append(parent[path[-1]])
    return path[::-1], cmap[goal], len(closed)


def _map(oracle, n, seed=0):
    """The planning space of examples/test.cpp:250-259 ([-1, 1]^2, three obstacles) sampled with n Halton points."""
    obs = [oracle.make_obstacle([(-0.5, 0), (1, 0), (1, 1), (0, 1)]), oracle.make_obstacle([(0, -0.5), (1, 0), (1, 1), (0, 1)]),
           oracle.make_obstacle([(-0.6, 0.148), (-1, 0.148), (-1, 0), (-0.6, 0)])]
    nodes = oracle.sample_free(n, (-1, 1, -1, 1), obs)
    edges = np.concatenate([o["lines"] for o in obs])
    return obs, nodes, edges


def _queries(rng, Q, obs, oracle):
    out = []
    while len(out) < 2 * Q:
        p = rng.uniform(-1, 1, 2).astype(np.float32)
        if not any(oracle.obstacle_contains(o, p) for o in obs):
            out.append(p)
    a = np.asarray(out, dtype=np.float32)
    return a[:Q], a[Q:]


# ------------------------------------------------------------------ CPU
def test_halton_known_values(oracle):
    np.testing.assert_array_equal(oracle.halton(2, 7), np.float32([1 / 2, 1 / 4, 3 / 4, 1 / 8, 5 / 8, 3 / 8, 7 / 8]))
    want3 = [np.float32(f) / np.float32(i) for f, i in [(1, 3), (2, 3), (1, 9), (4, 9), (7, 9), (2, 9), (5, 9), (8, 9), (1, 27)]]
    np.testing.assert_array_equal(oracle.halton(3, 9), np.float32(want3))
    # the state carries on where the last call stopped (sc:106-109, 126-127)
    st = [0, 0]
    parts = np.concatenate([oracle.halton(2, 5, st), oracle.halton(2, 6, st)])
    np.testing.assert_array_equal(parts, oracle.halton(2, 11))
    # radical inverse of k in base b
    def vdc(k, b):
        r, f = 0.0, 1.0 / b
        while k:
            r += f * (k % b); k //= b; f /= b
        return r
    np.testing.assert_allclose(oracle.halton(3, 500), [vdc(k, 3) for k in range(1, 501)], rtol=1e-6)


def test_obstacle_contains_and_sampling(oracle):
    sq = oracle.make_obstacle([(0, 0), (1, 0), (1, 1), (0, 1)])
    assert sq["lines"].shape == (4, 4)  # even vertex count: the loop is closed (sc:255-257)
    tri = oracle.make_obstacle([(0, 0), (1, 0), (0, 1)])
    assert tri["lines"].shape == (2, 4)  # odd count: the last edge is missing (the reference's quirk)
    assert oracle.obstacle_contains(sq, (0.5, 0.5)) and not oracle.obstacle_contains(sq, (1.5, 0.5))
    assert not oracle.obstacle_contains(sq, (0.5, -0.1))
    wall = oracle.make_obstacle([(0, 0), (1, 0), (1, 1), (0, 1)], closed=False)  # open: within 0.01 of a line (sc:239-248)
    assert oracle.obstacle_contains(wall, (0.5, 0.005)) and not oracle.obstacle_contains(wall, (0.5, 0.5))
    # the bounding rectangle of an odd vertex loop misses its last vertex (sc:261-265): (0.5, 0.005) is "outside" it
    assert not oracle.obstacle_contains(oracle.make_obstacle([(0, 0), (1, 0), (1, 1)], closed=False), (0.5, 0.005))
    pts = oracle.sample_free(300, (-1, 1, -1, 1), [sq])
    assert len(pts) == 300 and tuple(pts[0]) == (0.0, 0.0)  # the set starts as {(0, 0)} (sc:1297)
    assert len({tuple(p) for p in pts}) == 300
    inside = (pts[1:, 0] > 0) & (pts[1:, 0] < 1) & (pts[1:, 1] > 0) & (pts[1:, 1] < 1)
    assert not inside.any()
    # free-space allocations: only allocated space is sampled (sc:1290)
    left = oracle.sample_free(100, (-1, 1, -1, 1), [], free_rects=[(-1, 0, -1, 1)])
    assert len(left) == 100 and np.all(left[:, 0] <= 0)
    # nothing allocated: the reference loops forever; the restatement gives up after max_rounds
    assert len(oracle.sample_free(10, (-1, 1, -1, 1), [], free_rects=[])) == 1


def test_fmt_hand_cases(oracle):
    # a straight corridor: nodes on a line, spacing 0.5, near radius rn^2 = 0.64 -> one hop at a time
    nodes = np.float32([[0.5 * k, 0.0] for k in range(1, 6)])
    r = oracle.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], 0.8)
    assert r["status"][0] == 0 and r["path_len"][0] == 7
    np.testing.assert_array_equal(r["path_idx"][0, :7], [0, 1, 2, 3, 4, 5, 6])
    assert r["cost"][0] == np.float32(3.0)
    # a wall across the corridor: no path (std::nullopt, sc:1385-1387)
    r = oracle.fmt_batch(nodes, [(1.25, -5, 1.25, 5)], [(0, 0)], [(3, 0)], 0.8)
    assert r["status"][0] == 1 and r["cost"][0] == F32_MAX
    # ... and a detour node above the wall's end reopens it
    nodes2 = np.concatenate([nodes, np.float32([[1.0, 0.6], [1.5, 0.6]])])
    r = oracle.fmt_batch(nodes2, [(1.25, -5, 1.25, 0.3)], [(0, 0)], [(3, 0)], 0.8)
    assert r["status"][0] == 0
    ids = list(r["path_idx"][0, :r["path_len"][0]])
    assert 6 in ids and 7 in ids and ids[0] == 0 and ids[-1] == 8
    # x_init == x_goal: the loop never runs (sc:1355)
    r = oracle.fmt_batch(nodes, None, [(1, 1)], [(1, 1)], 0.8)
    assert r["status"][0] == 0 and r["path_len"][0] == 1 and r["closed"][0] == 0
    # path buffer too short
    r = oracle.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], 0.8, max_path=3)
    assert r["status"][0] == 2 and r["path_len"][0] == 7


def test_fmt_oracle_vs_set_restatement(oracle):
    obs, nodes, edges = _map(oracle, 120)
    init, goal = _queries(np.random.default_rng(3), 6, obs, oracle)
    r = oracle.fmt_batch(nodes, edges, init, goal, 0.55)
    found = 0
    for q in range(len(init)):
        path, cost, closed = _fmt_sets(nodes, edges, init[q], goal[q], 0.55)
        if path is None:
            assert r["status"][q] == 1
        else:
            found += 1
            n = r["path_len"][q]
            assert r["status"][q] == 0 and n == len(path)
            np.testing.assert_array_equal(np.stack([r["path_x"][q, :n], r["path_y"][q, :n]], 1), np.float32(path))
            assert r["cost"][q] == cost
        assert r["closed"][q] == closed
    assert found >= 3


def _check_paths(r, nodes, edges, init, goal, rn):
    """Size-independent properties: hops are near (<= rn^2), collision-free, and add up to the cost."""
    for q in range(len(init)):
        if r["status"][q] != 0:
            continue
        n = r["path_len"][q]
        px, py = r["path_x"][q, :n], r["path_y"][q, :n]
        assert (px[0], py[0]) == tuple(init[q]) and (px[-1], py[-1]) == tuple(goal[q])
        total = np.float32(0)
        for k in range(n - 1):
            a, b = (px[k], py[k]), (px[k + 1], py[k + 1])
            d = _pt_dist(a, b)
            assert float(d) <= float(np.float32(rn)) ** 2
            assert not any(_intersects(b, a, e[:2], e[2:]) for e in edges)
            total = np.float32(total + d)
        assert total == r["cost"][q]


def test_fmt_oracle_path_properties(oracle):
    obs, nodes, edges = _map(oracle, 400)
    init, goal = _queries(np.random.default_rng(5), 24, obs, oracle)
    r = oracle.fmt_batch(nodes, edges, init, goal, 0.45)
    assert (r["status"] == 0).sum() >= 12
    _check_paths(r, nodes, edges, init, goal, 0.45)


# ------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def planner():
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=64, max_waypoints=16, max_gridpoints=101)
    yield pl
    pl.close()


def _same(got, want):
    np.testing.assert_array_equal(got["status"], want["status"])
    np.testing.assert_array_equal(got["path_len"], want["path_len"])
    np.testing.assert_array_equal(got["closed"], want["closed"])
    np.testing.assert_array_equal(got["cost"], want["cost"])
    for q in np.nonzero(want["status"] == 0)[0]:
        n = want["path_len"][q]
        np.testing.assert_array_equal(got["path_x"][q, :n], want["path_x"][q, :n])
        np.testing.assert_array_equal(got["path_y"][q, :n], want["path_y"][q, :n])


@pytest.mark.gpu
@pytest.mark.parametrize("n,rn,Q", [(200, 0.6, 64), (200, 1.0, 16), (1000, 0.4, 96), (2560, 0.3, 48)])
def test_gpu_fmt_matches_oracle(planner, oracle, n, rn, Q):
    """examples/test.cpp:250-284's map (n = 200, rn = 1 is its literal call) and denser ones: bit-identical paths,
    costs, expansion counts and verdicts."""
    obs, nodes, edges = _map(oracle, n)
    init, goal = _queries(np.random.default_rng(n + Q), Q, obs, oracle)
    if n == 200 and rn == 1.0:
        init[0] = (-0.5, 1); goal[0] = (1, -1)  # the reference's own query
    want = oracle.fmt_batch(nodes, edges, init, goal, rn, max_path=128)
    got = planner.fmt_batch(nodes, edges, init, goal, rn, max_path=128)
    _same(got, want)
    assert (want["status"] == 0).sum() >= Q // 3
    _check_paths(got, nodes, edges, init, goal, rn)
    assert planner.timing()["launches"] == 1


@pytest.mark.gpu
def test_gpu_fmt_edge_cases(planner, oracle):
    nodes = np.float32([[0.5 * k, 0.0] for k in range(1, 6)])
    cases = [(None, (0, 0), (3, 0)), ([(1.25, -5, 1.25, 5)], (0, 0), (3, 0)), (None, (1, 1), (1, 1)),
             (None, (0, 0), (1.0, 0.0)),      # the goal coincides with a sample: one element of the set
             (None, (0.5, 0.0), (3, 0))]      # the start coincides with a sample
    for edges, a, b in cases:
        want = oracle.fmt_batch(nodes, edges, [a], [b], 0.8)
        got = planner.fmt_batch(nodes, edges, [a], [b], 0.8)
        _same(got, want)
    # hops of exactly 0.5 against rn^2 a hair below / above 0.5: the kernel's float screen cannot decide these, the
    # exact double-precision test does (one radius connects the corridor, the other does not)
    r_lo = np.float32(np.sqrt(0.5)); r_hi = np.nextafter(r_lo, np.float32(1))
    verdicts = []
    for rn in (r_lo, r_hi):
        want = oracle.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], rn)
        got = planner.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], rn)
        _same(got, want)
        verdicts.append(int(got["status"][0]))
    assert verdicts == [1, 0], (verdicts, float(r_lo) ** 2, float(r_hi) ** 2)
    want = oracle.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], 0.8, max_path=3)
    got = planner.fmt_batch(nodes, None, [(0, 0)], [(3, 0)], 0.8, max_path=3)
    assert got["status"][0] == want["status"][0] == 2 and got["path_len"][0] == want["path_len"][0] == 7
    # no samples at all: init and goal only
    got = planner.fmt_batch(np.zeros((0, 2), np.float32), None, [(0, 0), (0, 0)], [(0.5, 0), (5, 0)], 0.8)
    want = oracle.fmt_batch(np.zeros((0, 2), np.float32), None, [(0, 0), (0, 0)], [(0.5, 0), (5, 0)], 0.8)
    _same(got, want)
    assert list(got["status"]) == [0, 1]
    # the largest node set the kernel takes (8,192 samples: 172 KB of shared memory per query)
    rng = np.random.default_rng(1)
    big = rng.uniform(-1, 1, (8192, 2)).astype(np.float32)
    want = oracle.fmt_batch(big, [(0.0, -0.5, 0.0, 2.0)], [(-0.9, 0.9)], [(0.9, 0.9)], 0.2)
    got = planner.fmt_batch(big, [(0.0, -0.5, 0.0, 2.0)], [(-0.9, 0.9)], [(0.9, 0.9)], 0.2)
    _same(got, want)
    assert got["status"][0] == 0
    with pytest.raises(scb.ScbError):
        planner.fmt_batch(np.zeros((8193, 2), np.float32), None, [(0, 0)], [(1, 1)], 0.5)
