"""gen_vel_prof<N> for N degrees of freedom (sc:1191-1265, `requires (N >= 1)`) and toppra on a caller's piecewise-cubic
path with 1..6 degrees of freedom (scb_gen_vel_prof, scb_toppra_ppoly; toppra_ppoly_kernel<M, INTERP> with the generic-row
Seidel LP).  The oracle composes the same steps from its general-dof pieces (toppra_ppoly with the interpolation
discretisation, Spline parametrizer, uniform-dt record)."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def _problems(rng, P, m):
    p0 = rng.uniform(-2.0, 2.0, (P, m)); p1 = p0 + rng.uniform(0.5, 8.0, (P, m)) * rng.choice([-1.0, 1.0], (P, m))
    vs = rng.uniform(0.0, 0.3, (P, m)) * (p1 - p0); ve = rng.uniform(0.0, 0.3, (P, m)) * (p1 - p0)
    vhi = rng.uniform(0.5, 3.0, (P, m)); ahi = rng.uniform(0.25, 4.0, (P, m))
    return dict(pos_start=p0, pos_end=p1, vel_start=vs, vel_end=ve, vel_lo=-vhi, vel_hi=vhi, acc_lo=-ahi, acc_hi=ahi)


def _oracle_one(oracle, w, p, G, dt, **kw):
    return oracle.gen_vel_prof_ndof(w["pos_start"][p], w["pos_end"][p], w["vel_start"][p], w["vel_end"][p], w["vel_lo"][p],
                                    w["vel_hi"][p], w["acc_lo"][p], w["acc_hi"][p], G=G, dt=dt, **kw)


# ------------------------------------------------------------------ CPU: the oracle's composition
def test_oracle_one_dof_is_the_timing_stage(oracle):
    """N = 1 through the general-dof pieces == plan_batch's timing-only stage (the path gen_vel_prof<1> takes), bit for bit."""
    rng = np.random.default_rng(1)
    w = _problems(rng, 6, 1)
    for p in range(6):
        L = w["pos_end"][p, 0] - w["pos_start"][p, 0]
        if L < 0:
            continue
        r = _oracle_one(oracle, dict(w, pos_start=np.zeros((6, 1)), pos_end=w["pos_end"] - w["pos_start"]), p, 101, 0.02)
        ref = oracle.plan_batch(path_len=[L], vel_lo=w["vel_lo"][p], vel_hi=w["vel_hi"][p], acc_lo=w["acc_lo"][p], acc_hi=w["acc_hi"][p],
                                vel_start=w["vel_start"][p], vel_end=w["vel_end"][p], G=101, S_cap=8192, dt=0.02)[0]
        assert r["status"] == ref["status"][0] == 0
        n = len(r["time"])
        assert n == ref["n_samples"][0]
        np.testing.assert_array_equal(r["x"], ref["x"][:, 0])
        np.testing.assert_array_equal(r["t"], ref["t"][:, 0])
        np.testing.assert_array_equal(r["pos"][:, 0], ref["s_pos"][:n, 0])
        np.testing.assert_array_equal(r["acc"][:, 0], ref["s_acc"][:n, 0])


@pytest.mark.parametrize("m", [2, 3, 6])
def test_oracle_ndof_respects_every_limit(oracle, m):
    rng = np.random.default_rng(10 + m)
    w = _problems(rng, 4, m)
    G = 101
    s = np.linspace(0.0, 1.0, G)
    for p in range(4):
        r = _oracle_one(oracle, w, p, G, 0.05)
        assert r["status"] == 0
        co = oracle.hermite_coeffs(w["pos_start"][p], w["pos_end"][p], w["vel_start"][p], w["vel_end"][p])[0]
        qd = 3 * co[0][None, :] * s[:, None] ** 2 + 2 * co[1][None, :] * s[:, None] + co[2][None, :]
        v = qd * np.sqrt(r["x"])[:, None]
        assert np.all(v <= w["vel_hi"][p][None, :] * (1 + 1e-9) + 1e-12) and np.all(v >= w["vel_lo"][p][None, :] * (1 + 1e-9) - 1e-12)
        # the record starts and ends where it should, every degree of freedom
        np.testing.assert_allclose(r["pos"][0], w["pos_start"][p], atol=1e-12)
        np.testing.assert_allclose(r["pos"][-1], w["pos_end"][p], rtol=1e-12, atol=1e-12)
        # time-optimal: at every gridpoint but the ends some limit is (nearly) active somewhere on the path
        assert r["t"][-1] > 0


def test_oracle_identical_dofs_collapse_to_one(oracle):
    """Three copies of one degree of freedom have the constraints of one: same sdot^2, same record."""
    rng = np.random.default_rng(5)
    w1 = _problems(rng, 1, 1)
    w3 = {k: np.repeat(v, 3, axis=1) for k, v in w1.items()}
    a, b = _oracle_one(oracle, w1, 0, 64, 0.05), _oracle_one(oracle, w3, 0, 64, 0.05)
    np.testing.assert_allclose(b["x"], a["x"], rtol=1e-12, atol=1e-15)
    np.testing.assert_allclose(b["pos"][:, 2], a["pos"][:, 0], rtol=1e-12, atol=1e-12)


# ------------------------------------------------------------------ the reference's own end-to-end example
def _approach_case():
    """third-party/toppra/cpp/tests/test_toppra_approach.cpp:30-152 (the quickstart example): 2 joints, Hermite pieces
    through (0,0) -> (1,2) -> (2,0) with slopes (0,0), (1,1), (0,0) over s = 0, 1, 2; velocity box [[-1, 1], [-0.5, 0.5]],
    acceleration box [[-0.05, 0.2], [-0.1, 0.3]] with the INTERPOLATION discretisation; N = 100."""
    pos = np.array([[0.0, 0.0], [1.0, 2.0], [2.0, 0.0]]); vel = np.array([[0.0, 0.0], [1.0, 1.0], [0.0, 0.0]])
    breaks = np.array([0.0, 1.0, 2.0])
    co = np.zeros((2, 4, 2))
    for k in range(2):
        dt = breaks[k + 1] - breaks[k]
        c3 = pos[k]; c2 = vel[k]
        c0 = (vel[k + 1] * dt - 2 * pos[k + 1] + c2 * dt + 2 * c3) / dt ** 3
        c1 = (vel[k + 1] - c2 - 3 * c0 * dt * dt) / (2 * dt)
        co[k] = [c0, c1, c2, c3]
    return co, breaks, np.array([-1.0, -0.5]), np.array([1.0, 0.5]), np.array([-0.05, -0.1]), np.array([0.2, 0.3])


def _approach_check(co, breaks, x, vlo, vhi, alo, ahi):
    """What the reference's test asserts (l.133-151): ConstAccel parametrizer (tp:parametrizer/const_accel.cpp) evaluated
    at 11 times LinSpaced over its interval; every joint velocity and acceleration inside its box x (1 + 0.01)."""
    G = len(x); s = np.linspace(breaks[0], breaks[-1], G); sd = np.sqrt(x)
    ts = np.concatenate([[0.0], np.cumsum(2 * np.diff(s) / (sd[1:] + sd[:-1]))])
    us = 0.5 * np.diff(x) / np.diff(s)
    for t in np.linspace(ts[0], ts[-1], 11):
        k = min(max(int(np.searchsorted(ts[1:-1], t, side="left")), 0), G - 2)
        d = t - ts[k]
        v = sd[k] + d * us[k]
        ss = min(max(s[k] + d * sd[k] + 0.5 * d * d * us[k], breaks[0]), breaks[-1])
        seg = min(int(np.searchsorted(breaks, ss, side="right")) - 1, len(breaks) - 2)
        e = ss - breaks[seg]
        q1 = 3 * co[seg, 0] * e * e + 2 * co[seg, 1] * e + co[seg, 2]
        q2 = 6 * co[seg, 0] * e + 2 * co[seg, 1]
        vel, acc = q1 * v, q2 * v * v + q1 * us[k]
        assert np.all(vel >= vlo * 1.01) and np.all(vel <= vhi * 1.01), (t, vel)
        assert np.all(acc >= alo * 1.01) and np.all(acc <= ahi * 1.01), (t, acc)


def test_reference_quickstart_example_on_the_oracle(oracle):
    co, breaks, vlo, vhi, alo, ahi = _approach_case()
    rc, K, x, _, _ = oracle.toppra_ppoly(co, breaks, 101, vlo, vhi, alo, ahi, 1)
    assert rc == 0   # ReturnCode::OK (l.79, l.103)
    _approach_check(co, breaks, x, vlo, vhi, alo, ahi)


# ------------------------------------------------------------------ GPU
@pytest.fixture(scope="module")
def planner():
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=512, max_waypoints=4, max_gridpoints=512, max_samples=4096)
    yield pl
    pl.close()


@pytest.mark.gpu
@pytest.mark.parametrize("m", [1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("interp", [1, 0])
def test_gen_vel_prof_ndof_matches_oracle(planner, oracle, m, interp):
    rng = np.random.default_rng(100 + 10 * m + interp)
    P, G, S, dt = 40, 101, 4096, 0.02
    w = _problems(rng, P, m)
    w["acc_lo"][3] = 0.2 * w["acc_hi"][3]     # cannot brake: fails
    res = planner.gen_vel_prof(G=G, dt=dt, S_cap=S, interpolation=interp, **w)
    nfail = 0
    for p in range(P):
        r = _oracle_one(oracle, w, p, G, dt, interpolation=interp)
        assert res["status"][p] == r["status"], p
        if r["status"] != 0:
            assert res["fail_stage"][p] == r["fail_stage"], p
            nfail += 1
            continue
        # the LP passes are the reference's sequence operation for operation: same bits
        np.testing.assert_array_equal(res["x"][:, p], r["x"], err_msg="x %d" % p)
        np.testing.assert_array_equal(res["K_hi"][:, p], r["K"][:, 1]); np.testing.assert_array_equal(res["K_lo"][:, p], r["K"][:, 0])
        T = r["t"][-1]
        np.testing.assert_allclose(res["t"][:, p], r["t"], rtol=1e-12, atol=1e-12 * T)
        assert abs(res["T_total"][p] - T) <= 1e-12 * T
        n = len(r["time"])
        assert res["n_samples"][p] == n
        n = min(n, S)
        np.testing.assert_allclose(res["s_time"][:n, p], r["time"][:n], rtol=1e-12, atol=1e-12 * T)
        for k in range(m):
            span = abs(w["pos_end"][p, k] - w["pos_start"][p, k])
            np.testing.assert_allclose(res["s_pos"][k, :n, p], r["pos"][:n, k], rtol=0, atol=1e-9 * max(span, np.abs(r["pos"][:, k]).max()))
            np.testing.assert_allclose(res["s_vel"][k, :n, p], r["vel"][:n, k], rtol=0, atol=1e-9 * np.abs(r["vel"][:, k]).max() + 1e-9 * span / T)
            # the record's second derivative: the spline's conditioning (tests/tolerances.py), bounded here by 1e-6 of its range
            np.testing.assert_allclose(res["s_acc"][k, :n, p], r["acc"][:n, k], rtol=0, atol=1e-6 * np.abs(r["acc"][:, k]).max())
    assert nfail >= 1


@pytest.mark.gpu
def test_gen_vel_prof_ndof_with_a_varying_velocity_limit(planner, oracle):
    """vel_lim_func tabulated at the gridpoints (LinearJointVelocityVarying, sc:1177-1188): a limit that drops mid-path."""
    rng = np.random.default_rng(7)
    P, m, G = 24, 3, 201
    w = _problems(rng, P, m)
    f = np.ones((G, 1, 1)); f[G // 3: G // 2] = 0.25
    hi = f * w["vel_hi"].T[None, :, :]        # [G][m][P]
    res = planner.gen_vel_prof(G=G, dt=0.05, S_cap=2048, vel_lo_grid=-hi, vel_hi_grid=hi,
                               **{k: v for k, v in w.items() if k not in ("vel_lo", "vel_hi")}, vel_lo=None, vel_hi=None)
    for p in range(P):
        r = oracle.gen_vel_prof_ndof(w["pos_start"][p], w["pos_end"][p], w["vel_start"][p], w["vel_end"][p], None, None,
                                     w["acc_lo"][p], w["acc_hi"][p], G=G, dt=0.05, vel_lo_grid=-hi[:, :, p], vel_hi_grid=hi[:, :, p])
        assert res["status"][p] == r["status"]
        if r["status"] == 0:
            np.testing.assert_array_equal(res["x"][:, p], r["x"])
            assert np.all(res["x"][G // 3: G // 2, p] <= res["x"][:, p].max())


@pytest.mark.gpu
@pytest.mark.parametrize("m,interp", [(3, 0), (3, 1), (6, 0), (6, 1), (4, 1)])
def test_toppra_ppoly_many_dof_matches_oracle(planner, oracle, m, interp):
    """Random multi-segment cubic paths with m degrees of freedom: controllable sets, parametrization and feasible sets of
    the generic-row LP kernel against the oracle, bit for bit."""
    rng = np.random.default_rng(40 + m + interp)
    P, ns, G = 16, 4, 61
    breaks = np.concatenate([[0.0], np.cumsum(rng.uniform(0.5, 1.5, ns))])
    # a C1 path: cubic Hermite pieces through random knots with random slopes
    pos = rng.uniform(-1.0, 1.0, (P, ns + 1, m)).cumsum(axis=1); vel = rng.uniform(-1.0, 1.0, (P, ns + 1, m))
    co = np.zeros((P, ns, 4, m))
    for s in range(ns):
        dt = breaks[s + 1] - breaks[s]
        c3 = pos[:, s]; c2 = vel[:, s]
        c0 = (vel[:, s + 1] * dt - 2 * pos[:, s + 1] + c2 * dt + 2 * c3) / dt ** 3
        c1 = (vel[:, s + 1] - c2 - 3 * c0 * dt * dt) / (2 * dt)
        co[:, s, 0], co[:, s, 1], co[:, s, 2], co[:, s, 3] = c0, c1, c2, c3
    vhi = rng.uniform(1.0, 3.0, (P, m)); ahi = rng.uniform(0.5, 4.0, (P, m))
    res = planner.toppra_ppoly(co, breaks, G, -vhi, vhi, -ahi, ahi, interpolation=interp, feasible=True)
    seen = set()
    for p in range(P):
        rc, K, x, feas, frc = oracle.toppra_ppoly(co[p], breaks, G, -vhi[p], vhi[p], -ahi[p], ahi[p], interp, feasible=True)
        assert res["status"][p] == rc, p
        seen.add(rc)
        assert res["feas_status"][p] == frc
        if frc == 0:
            np.testing.assert_array_equal(res["feas_lo"][:, p], feas[:, 0]); np.testing.assert_array_equal(res["feas_hi"][:, p], feas[:, 1])
        if rc == 0:
            np.testing.assert_array_equal(res["K_lo"][:, p], K[:, 0]); np.testing.assert_array_equal(res["K_hi"][:, p], K[:, 1])
            np.testing.assert_array_equal(res["x"][:, p], x)
    assert 0 in seen


@pytest.mark.gpu
def test_reference_quickstart_example_on_the_kernel(planner, oracle):
    """test_toppra_approach.cpp on toppra_ppoly_kernel<2, true> (ten-row stage LPs): the reference's assertions on the
    kernel's own parametrization, and the oracle's bits."""
    co, breaks, vlo, vhi, alo, ahi = _approach_case()
    res = planner.toppra_ppoly(co[None], breaks, 101, vlo, vhi, alo, ahi, interpolation=1, feasible=False)
    assert res["status"][0] == 0
    _approach_check(co, breaks, res["x"][:, 0], vlo, vhi, alo, ahi)
    rc, K, x, _, _ = oracle.toppra_ppoly(co, breaks, 101, vlo, vhi, alo, ahi, 1)
    np.testing.assert_array_equal(res["x"][:, 0], x)
    np.testing.assert_array_equal(res["K_hi"][:, 0], K[:, 1]); np.testing.assert_array_equal(res["K_lo"][:, 0], K[:, 0])


@pytest.mark.gpu
def test_ndof_rejects_what_it_cannot_run(planner):
    w = _problems(np.random.default_rng(0), 2, 7)
    with pytest.raises(scb.ScbError):
        planner.gen_vel_prof(G=11, **w)
