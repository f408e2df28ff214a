"""Parity at the sizes and in the corners the benchmark configs name ("next" item 1):

  * C5 at its real shape (W = 200, G = 10,000) in FP64 (1e-9) and FP32 (1e-4) against the oracle's tridiagonal mode
    ("restated-fast": the reference's dense (G x G) column-pivoted QR, tp:geometric_path/piecewise_poly_path.cpp:112-163,
    is 1.3 TFLOP per problem at this G; the oracle's Thomas mode solves the same strictly diagonally dominant system);
  * verdicts at scale: 65,536 problems of which > 1 % FAIL (scb.workloads.hard_cases), closed form == Seidel ==
    oracle on every status / fail_stage;
  * the degeneracies of the reference's own LP tests (third-party/toppra/cpp/tests/test_solver.cpp:277-441: x-box
    inverted by 2e-14, equal bounds, rows with an infinite constant) rebuilt as 1-DOF stage problems and pushed through
    BOTH stage-LP back ends on the GPU;
  * the CTA-team path kernel (path_kernel.cu) where it is the only kernel: more than 512 waypoints.
"""
import numpy as np
import pytest

import seacurrent_b200 as scb
from test_gpu_parity import GRID, RTOL32, RTOL64, _check, rel_err
from tolerances import conditioning

pytestmark = pytest.mark.gpu

INF = np.inf


# ------------------------------------------------------------------ C5 at its real shape
@pytest.fixture(scope="module")
def c5_case(oracle):
    P, W, G = 64, 200, 10000
    rng = np.random.default_rng(2005)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    # the waypoints FP32 storage can represent, so that both precisions see identical inputs
    wx = wx.astype(np.float32).astype(np.float64); wy = wy.astype(np.float32).astype(np.float64)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.05)
    refs = {}

    def ref(mode):  # (oracle result, its measured own spread under 8-ulp input perturbations: tests/tolerances.py)
        if mode not in refs:
            refs[mode] = conditioning(oracle, wx=wx, wy=wy, G=G, spline_mode=0 if mode == "ref_cubic" else 1, solve_mode=1, **lim)
        return refs[mode]
    return dict(P=P, W=W, G=G, wx=wx, wy=wy, lim=lim, ref=ref)


@pytest.mark.parametrize("mode", ["ref_cubic", "lau_quintic"])
def test_c5_shape_fp64(c5_case, mode):
    c = c5_case
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=c["P"], max_waypoints=c["W"], max_gridpoints=c["G"])
    res = pl.plan(wx=c["wx"], wy=c["wy"], G=c["G"], **c["lim"])
    pl.close()
    ref, spread = c["ref"](mode)
    assert (ref["status"] == 0).sum() >= c["P"] - 2
    _check(res, ref, RTOL64, ("ctrl_x", "ctrl_y", "arclen_total", "arclen_table", "T_total", "cost") + GRID, spread)


@pytest.mark.parametrize("mode", ["ref_cubic", "lau_quintic"])
def test_c5_shape_fp32(c5_case, mode):
    """FP32 mode at C5's shape: identical verdicts, 1e-4 of the path's own scale on every record (positions against
    the extent of the path, the rest against 1 % of its range, as test_fp32_mode_within_1e4)."""
    c = c5_case
    pl = scb.BatchPlanner(0, "fp32", mode, max_problems=c["P"], max_waypoints=c["W"], max_gridpoints=c["G"])
    res = pl.plan(wx=c["wx"], wy=c["wy"], G=c["G"], **c["lim"])
    pl.close()
    ref, _ = c["ref"](mode)
    np.testing.assert_array_equal(res["status"], ref["status"])
    np.testing.assert_array_equal(res["fail_stage"], ref["fail_stage"])
    ok = ref["status"] == 0
    for n in ("arclen_total", "T_total", "ctrl_x", "ctrl_y", "x", "K_hi", "K_lo", "t", "px", "py", "v"):
        g, w = (res[n][:, ok], ref[n][:, ok]) if n in GRID else (res[n][ok], ref[n][ok])
        frac = 1.0 if n in ("ctrl_x", "ctrl_y", "px", "py") else 1e-2
        e = rel_err(g, w, floor=frac * float(np.max(np.abs(ref["K_hi" if n == "K_lo" else n]))))
        assert e <= RTOL32, (mode, n, e)


# ------------------------------------------------------------------ verdicts at scale
def test_failing_problems_at_scale(oracle):
    """65,536 problems in C4's shape (W = 10, G = 500) with SURVEY 8-D2's hard cases switched on: more than 1 % of them
    fail, at stages spread over the whole path.  The closed-form and the Seidel kernels and the oracle agree on EVERY
    status and fail_stage; K, x and T of a subsample that contains all failures' neighbours agree to 1e-9."""
    P = 65536
    w = scb.workloads.config_c4(P=P)
    G = w["G"]
    outs = ("K_lo", "K_hi", "x", "T_total", "arclen_total")
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=P, max_waypoints=10, max_gridpoints=G)
    a = pl.plan(outputs=outs, lp_mode=scb.abi.LP_CLOSED_FORM, **w)
    b = pl.plan(outputs=outs, lp_mode=scb.abi.LP_SEIDEL, **w)
    pl.close()
    failed = a["status"] != 0
    assert failed.mean() > 0.01, failed.mean()
    assert len(np.unique(a["fail_stage"][failed])) > 20      # not one canned failure
    np.testing.assert_array_equal(a["status"], b["status"])
    np.testing.assert_array_equal(a["fail_stage"], b["fail_stage"])
    ref, _ = oracle.plan_batch(spline_mode=0, outputs=outs, **w)
    np.testing.assert_array_equal(a["status"], ref["status"])
    np.testing.assert_array_equal(a["fail_stage"], ref["fail_stage"])
    ok = ref["status"] == 0
    sc = ref["x"][:, ok].max(axis=0)[None, :]
    for got in (a, b):
        for n in ("K_lo", "K_hi", "x"):
            err = np.abs(got[n][:, ok] - ref[n][:, ok]) / np.maximum(np.abs(ref[n][:, ok]), 1e-3 * sc)
            assert err.max() <= RTOL64, (n, err.max())
        assert rel_err(got["T_total"][ok], ref["T_total"][ok]) <= RTOL64


# ------------------------------------------------------------------ degenerate stage LPs through both back ends
def _stage_cases():
    """(name, L, vel_lo, vel_hi, acc_lo, acc_hi) with scalar limits."""
    return [
        ("plain", 10.0, -1.0, 1.0, -1.0, 1.0),
        # equal bounds (test_solver.cpp:277-441 instance 5: low[1] == high[1]): the acceleration slab has zero width
        ("acc_lo == acc_hi == 0: only rest is feasible", 10.0, -1.0, 1.0, 0.0, 0.0),
        ("acc_lo == acc_hi != 0: infeasible", 10.0, -1.0, 1.0, 0.3, 0.3),
        # rows with an infinite constant (instance 5's last row; seidel.cpp:14-25 treats 0 * inf as 0)
        ("no acceleration limit", 10.0, -1.0, 1.0, -INF, INF),
        ("no velocity limit", 10.0, -INF, INF, -1.0, 1.0),
        ("no limits at all: the hidden cap x <= 100 binds everywhere", 10.0, -INF, INF, -INF, INF),
        ("one-sided infinite limits", 10.0, -INF, 1.0, -1.0, INF),
        # coefficients around Seidel's ABS_TOLERANCE (rows with |a| < 1e-13 count as constant, seidel-internal.hpp:82)
        ("path shorter than the row tolerance", 1e-14, -1.0, 1.0, -1.0, 1.0),
        ("path of 1e-12", 1e-12, -1.0, 1.0, -1.0, 1.0),
        ("path of 1e-6", 1e-6, -1.0, 1.0, -1.0, 1.0),
        ("long path, tiny speed", 1e6, -1e-3, 1e-3, -1.0, 1.0),
        ("zero velocity box", 10.0, 0.0, 0.0, -1.0, 1.0),
        ("hidden cap binds: short path, huge limits", 1e-3, -1e3, 1e3, -1e6, 1e6),
        # limits the reference rejects at construction / lets through as NaN
        ("acc box inverted by 2e-14", 10.0, -1.0, 1.0, -1.0 + 2e-14, -1.0),
        ("NaN velocity limit", 10.0, np.nan, 1.0, -1.0, 1.0),
        ("NaN acceleration limit", 10.0, -1.0, 1.0, -1.0, np.nan),
    ]


@pytest.mark.parametrize("lp_mode", [scb.abi.LP_CLOSED_FORM, scb.abi.LP_SEIDEL])
def test_degenerate_stage_lps_scalar_limits(oracle, lp_mode):
    cases = _stage_cases()
    L, vlo, vhi, alo, ahi = (np.array([c[k] for c in cases], dtype=np.float64) for k in range(1, 6))
    G = 101
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=64, max_waypoints=4, max_gridpoints=G)
    outs = ("K_lo", "K_hi", "x", "t", "T_total")
    res = pl.plan(path_len=L, G=G, stages=scb.abi.STAGE_TIMING, lp_mode=lp_mode, vel_lo=vlo, vel_hi=vhi, acc_lo=alo, acc_hi=ahi,
                  outputs=outs)
    pl.close()
    ref, _ = oracle.plan_batch(path_len=L, G=G, vel_lo=vlo, vel_hi=vhi, acc_lo=alo, acc_hi=ahi)
    names = [c[0] for c in cases]
    for p, nm in enumerate(names):
        assert res["status"][p] == ref["status"][p], (nm, res["status"][p], ref["status"][p])
        if ref["status"][p] < 16:
            assert res["fail_stage"][p] == ref["fail_stage"][p], (nm, res["fail_stage"][p], ref["fail_stage"][p])
    assert list(ref["status"][-3:]) == [17, 17, 17]
    ok = ref["status"] == 0
    assert ok.sum() >= len(cases) - 5
    for p in np.flatnonzero(ok):
        sc = max(ref["x"][:, p].max(), 1e-300)
        for n in ("K_lo", "K_hi", "x"):
            err = np.max(np.abs(res[n][:, p] - ref[n][:, p]) / np.maximum(np.abs(ref[n][:, p]), 1e-3 * sc))
            assert err <= RTOL64, (names[p], n, err)
        assert abs(res["T_total"][p] - ref["T_total"][p]) <= RTOL64 * ref["T_total"][p], (names[p], res["T_total"][p], ref["T_total"][p])


@pytest.mark.parametrize("lp_mode", [scb.abi.LP_CLOSED_FORM, scb.abi.LP_SEIDEL])
def test_degenerate_stage_lps_gridpoint_boxes(oracle, lp_mode):
    """Per-gridpoint velocity boxes (vel_lim_func, sc:1175-1188) that produce the x-box degeneracies of
    test_solver.cpp:277-441 inside the passes: low == high (instance 5) and low > high by ~2e-14 relative (instance 1:
    `high << 2.06944 - 2.22045e-14`), which Seidel collapses to a 1-D LP (seidel.cpp:73-106); plus a box inverted by
    more than REL_TOLERANCE (infeasible), an infinite box in the middle of a finite one, and a NaN at one gridpoint."""
    G, L = 101, 10.0
    s = np.linspace(0.0, 1.0, G)
    zone = (s >= 0.3) & (s <= 0.7)
    base_hi = np.full(G, 1.0); base_lo = -base_hi

    def col(lo_zone, hi_zone):
        return np.where(zone, lo_zone, base_lo), np.where(zone, hi_zone, base_hi)
    cols = [
        ("plain", base_lo, base_hi, 1.0),
        ("equal box, reachable", *col(0.2, 0.2), 1.0),
        ("equal box, not reachable", *col(0.2, 0.2), 1e-3),
        ("box inverted by 2e-14 relative", *col(0.2 * (1 + 1e-14), 0.2), 1.0),
        ("box inverted by 1e-9 relative", *col(0.2 * (1 + 1e-9), 0.2), 1.0),
        ("box inverted by 1e-11 relative (inside REL_TOLERANCE on x = sd^2: 2e-11)", *col(0.2 * (1 + 1e-11), 0.2), 1.0),
        ("infinite box inside a finite one", *col(-INF, INF), 1.0),
        ("one NaN gridpoint", np.where(np.arange(G) == 40, np.nan, base_lo), base_hi, 1.0),
        ("lower limit above upper by a lot", *col(0.5, 0.2), 1.0),
    ]
    P = len(cols)
    vlg = np.ascontiguousarray(np.stack([c[1] for c in cols], axis=1)); vhg = np.ascontiguousarray(np.stack([c[2] for c in cols], axis=1))
    acc = np.array([c[3] for c in cols])
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=64, max_waypoints=4, max_gridpoints=G)
    outs = ("K_lo", "K_hi", "x", "t", "T_total")
    res = pl.plan(path_len=np.full(P, L), G=G, stages=scb.abi.STAGE_TIMING, lp_mode=lp_mode, vel_lo_grid=vlg, vel_hi_grid=vhg,
                  acc_lo=-acc, acc_hi=acc, outputs=outs)
    pl.close()
    ref, _ = oracle.plan_batch(path_len=np.full(P, L), G=G, vel_lo_grid=vlg, vel_hi_grid=vhg, acc_lo=-acc, acc_hi=acc)
    for p, c in enumerate(cols):
        assert res["status"][p] == ref["status"][p], (c[0], res["status"][p], ref["status"][p])
        if ref["status"][p] < 16:
            assert res["fail_stage"][p] == ref["fail_stage"][p], (c[0], res["fail_stage"][p], ref["fail_stage"][p])
    assert ref["status"][0] == 0 and ref["status"][1] == 0 and ref["status"][3] == 0 and ref["status"][7] == 17
    assert ref["status"][4] == 2 and ref["status"][8] == 2
    for p in np.flatnonzero(ref["status"] == 0):
        sc = ref["x"][:, p].max()
        for n in ("K_lo", "K_hi", "x"):
            err = np.max(np.abs(res[n][:, p] - ref[n][:, p]) / np.maximum(np.abs(ref[n][:, p]), 1e-3 * sc))
            assert err <= RTOL64, (cols[p][0], n, err)
        assert abs(res["T_total"][p] - ref["T_total"][p]) <= RTOL64 * ref["T_total"][p], cols[p][0]


# ------------------------------------------------------------------ the CTA-team path kernel beyond the lane kernels
@pytest.mark.parametrize("mode,W", [("ref_cubic", 600), ("lau_quintic", 520)])
def test_path_kernel_beyond_512_waypoints(oracle, mode, W):
    """W > 512: path_kernel.cu (CTA team, Bernstein-form cost in shared memory) is the only path kernel.  Its optimiser
    state lives in one SM's shared memory, which bounds the quintic mode at 522 waypoints (seacurrent_b200.h)."""
    P, G = 4, 400
    rng = np.random.default_rng(600)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.0)
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=P, max_waypoints=W, max_gridpoints=G)
    res = pl.plan(wx=wx, wy=wy, G=G, **lim)
    pl.close()
    ref, spread = conditioning(oracle, wx=wx, wy=wy, G=G, spline_mode=0 if mode == "ref_cubic" else 1, **lim)
    assert np.all(ref["status"] == 0)
    _check(res, ref, RTOL64, ("ctrl_x", "ctrl_y", "arclen_total", "arclen_table", "T_total", "cost") + GRID, spread)
    if mode == "lau_quintic":  # one more waypoint than fits: refused, not truncated
        pl = scb.BatchPlanner(0, "fp64", mode, max_problems=P, max_waypoints=600, max_gridpoints=G)
        wx2, wy2 = scb.workloads.random_walk_paths(rng, P, 523)
        with pytest.raises(scb.ScbError):
            pl.plan(wx=wx2, wy=wy2, G=G, **lim)
        pl.close()


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_warp_specialised_toppra_is_bit_identical(monkeypatch, precision):
    """toppra_ws_kernel (chain warp + three producer warps per 32 problems; what batches below 37,888 problems run) against
    toppra_fast_kernel (one thread per problem): the same arithmetic operation for operation, so every bit of the verdicts,
    the controllable sets and the parametrization - with failing problems, per-gridpoint limits, curvature caps, end slopes,
    ragged CTAs (P = 1, 33, 4096 + 5) and gridpoint counts around the chunk size (N = 1, 2, 6, 7, 12, 13, 499)."""
    rng = np.random.default_rng(99)
    cases = []
    for P, G in ((1, 2), (33, 3), (70, 7), (64, 8), (5, 13), (97, 14), (4101, 500), (300, 101)):
        w = scb.workloads.config_c2(P=P, G=G, hard_fraction=0.1) if G >= 8 else None
        if w is None:
            wx, wy = scb.workloads.random_walk_paths(rng, P, 6)
            w = dict(wx=wx, wy=wy, G=G, **scb.workloads.random_limits(rng, P, slow_fraction=0.1))
        w["vel_start"] = rng.uniform(0.0, 0.2, P); w["vel_end"] = rng.uniform(0.0, 0.2, P)
        cases.append((w, {}))
        if G >= 13:
            cases.append((w, dict(omega_max=1.0, acc_lat_max=0.8)))
    cases.append((dict(path_len=rng.uniform(1.0, 30.0, 500), G=64, **scb.workloads.random_limits(rng, 500, slow_fraction=0.1)),
                  dict(stages=scb.abi.STAGE_TIMING, outputs=("K_lo", "K_hi", "x", "t", "v", "a", "T_total"))))
    w = scb.workloads.config_c2(P=4800, G=40, hard_fraction=0.1)   # more than 148 groups: two groups per CTA
    cases.append((w, {}))
    got = {}
    for ws in ("0", "1"):
        monkeypatch.setenv("SCB_TOPPRA_WS", ws)
        pl = scb.BatchPlanner(0, precision, "ref_cubic", max_problems=4800, max_waypoints=24, max_gridpoints=512)
        try:
            got[ws] = [pl.plan(**w, **kw) for w, kw in cases]
        finally:
            pl.close()
    nfail = 0
    for a, b in zip(got["0"], got["1"]):
        nfail += int((a["status"] != 0).sum())
        for k in ("status", "fail_stage", "K_lo", "K_hi", "x", "t", "T_total", "v", "a"):
            np.testing.assert_array_equal(a[k], b[k], err_msg=k)
    assert nfail > 0
