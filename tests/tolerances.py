"""Tolerances of the GPU-vs-oracle parity tests, backed by measurements of the oracle itself.

The parity bar is 1e-9 relative (FP64).  Three of the per-gridpoint quantities cannot be held to it END TO END by any
two implementations that round differently, because the oracle's own value moves by more than that when its inputs
move by a few ulp:
  a      knot accelerations = twice the second-derivative unknowns of a spline INTERPOLATING q(t_i): condition number
         ~ L / h_min^2 with respect to the knot data
  v      knot velocities, v_i = dq/h - h/3 (2 X_i + X_{i+1}), inherit it times h
  omega  = v * kappa(u(q)): at a curvature spike (a near-cusp of the spline) kappa changes by orders of magnitude over
         1e-3 of a segment, so the ~eps * L uncertainty of the arc position q is amplified
No modelled slack is used for them:
  * omega is checked factor by factor - v in the parametrizer stage (below), kappa = omega / v here: conditioning()
    RE-RUNS THE ORACLE on inputs perturbed by PROBE_ULPS ulp (random signs) and the tolerance is
    rtol * max(|want|, 1e-3 scale) + OMEGA_FACTOR * observed spread, pointwise (running maximum of the spread over
    +-8 gridpoints, so that a handful of random probes (four) leave no accidental zeros).  A backward-error statement: the GPU result
    must be as close to the oracle as the oracle is to itself under input perturbations of PROBE_ULPS * SPREAD_FACTOR ulp.
  * a, v (and t) are checked STAGE-WISE, where the bar can be held: the parametrizer's inputs are x (= sdot^2) and L,
    which the tests hold to 1e-9 - but not to an ulp, and a amplifies whatever x differs by (first GPU run of this
    scheme: the end-to-end difference in a was 40x what 32-ulp input perturbations explain, all of it inherited from x).
    param_stage_check() therefore feeds the KERNEL's own x and L to the oracle's parametrizer (tp:parametrizer/
    spline.cpp:10-45 restated) and compares t, v, a with rtol * max(|want|, 1e-3 scale) + SPREAD_FACTOR * the oracle's
    OWN rounding error in that stage, measured against an 80-bit extended-precision evaluation of the same formulas
    (param_stage_exact; per problem, the maximum over the knots).  The knot values q(s_i) ~ L enter the spline through
    differences q_i - q_{i-1} that cancel ~L / dq digits, most near the path's ends where q' -> 0: the rounding of q_i
    alone (half an ulp of L, independent per knot) puts ~10 eps L / h^2 of noise on a in ANY double-precision
    evaluation, the oracle's included (measured on the first GPU run: 1e-10 .. 1e-9 absolute near the end of the path).
Everything else (x, K, t, px, py, control points, tables,
L, T, J) gets NO such term: tests/test_tolerances.py shows on the CPU that they move by < 1e-10 relative.
"""
import numpy as np

PROBE_ULPS = 8
SPREAD_FACTOR = 4.0
# omega = v * kappa(u(q)): both implementations stop the Newton inversion of the arc-length table at |du| <= 1e-12 h,
# i.e. 45 ulp of the segment parameter, on top of their rounding: 8 probes' ulp x 8 = 64 ulp-equivalents
OMEGA_FACTOR = 8.0
ILL_CONDITIONED = ("a", "v", "omega", "s_acc", "s_vel", "s_omega")
GRID = ("K_lo", "K_hi", "x", "t", "px", "py", "v", "a", "omega")


def conditioning(oracle, n_probe=4, ulps=PROBE_ULPS, seed=4242, names=ILL_CONDITIONED, **kw):
    """Spread of the oracle's own outputs `names` over n_probe re-runs with every waypoint coordinate (or path length)
    multiplied by 1 +- ulps * eps.  kw = the arguments of oracle.plan_batch.  Returns (reference result, {name: spread})."""
    ref, _ = oracle.plan_batch(**kw)
    rng = np.random.default_rng(seed)
    d = ulps * np.finfo(np.float64).eps
    spread = {}
    for _ in range(n_probe):
        pk = dict(kw)
        for key in ("wx", "wy", "path_len"):
            if pk.get(key) is not None:
                arr = np.asarray(pk[key], dtype=np.float64)
                pk[key] = arr * (1.0 + d * rng.choice([-1.0, 1.0], size=arr.shape))
        pr, _ = oracle.plan_batch(**pk)
        same = pr["status"] == ref["status"]
        if "omega" in ref and "v" in ref:  # curvature at the gridpoints, the factor of omega the arc kernel owns
            ref["_kappa"], pr["_kappa"] = _kappa(ref), _kappa(pr)
        for n in tuple(names) + ("_kappa",):
            if n in ref:
                with np.errstate(invalid="ignore"):   # +inf costs of colliding splines: inf - inf
                    diff = np.nan_to_num(np.abs(pr[n] - ref[n]), nan=0.0)
                # a verdict that flips under the perturbation is a knife-edge problem, not a spread
                if diff.ndim == 2 and (n in GRID or n.startswith("s_") or n == "_kappa"):
                    diff[:, ~same] = 0.0     # [G][P] / [S][P]
                else:
                    diff[~same] = 0.0        # [P] / [P][n]
                spread[n] = np.maximum(spread.get(n, 0.0), diff)
    for n, s in spread.items():  # running maximum over +-8 gridpoints / samples
        if s.ndim == 2 and s.shape[0] > 16:
            pad = np.pad(s, ((8, 8), (0, 0)), mode="edge")
            m = pad[0:s.shape[0]].copy()
            for k in range(1, 17):
                np.maximum(m, pad[k:k + s.shape[0]], out=m)
            spread[n] = m
        elif s.ndim == 2 and s.shape[0] > 1 and (n in GRID or n.startswith("s_") or n == "_kappa"):
            spread[n] = np.broadcast_to(s.max(axis=0, keepdims=True), s.shape).copy()   # fewer than 17 rows: +-8 covers them all
    spread["_oracle"] = oracle
    spread["_kw"] = kw
    return ref, spread


def param_stage_exact(L, vs, ve, x):
    """The parametrizer stage (trapezoid knot times, clamped cubic spline through (t_i, q(s_i)), its first and second
    derivative at the knots: tp:parametrizer/spline.cpp:10-45, tp:geometric_path/piecewise_poly_path.cpp:102-187) in 80-bit
    extended precision (numpy longdouble, 64-bit mantissa): the yardstick both the oracle and the kernel are measured
    with.  Inputs are the doubles the stage receives."""
    LD = np.longdouble
    G = len(x); N = G - 1
    L, vs, ve = LD(L), LD(vs), LD(ve)
    c2 = vs; c0 = ve - 2 * L + vs; c1 = (ve - vs - 3 * c0) / 2
    s = np.arange(G, dtype=LD) / LD(N)
    q = c0 * s ** 3 + c1 * s ** 2 + c2 * s
    qd = lambda u: 3 * c0 * u * u + 2 * c1 * u + c2
    sd = np.sqrt(np.asarray(x, dtype=LD))
    avg = (sd[:-1] + sd[1:]) / 2
    h = np.where(avg > LD(1e-8), np.diff(s) / np.where(avg > 0, avg, 1), LD(5))
    t = np.concatenate([[LD(0)], np.cumsum(h)])
    slope = np.diff(q) / h
    bc0, bc1 = qd(LD(0)) * sd[0], qd(LD(1)) * sd[-1]
    lo = np.concatenate([[LD(0)], h]); up = np.concatenate([h, [LD(0)]])
    di = np.concatenate([[2 * h[0]], 2 * (h[:-1] + h[1:]), [2 * h[-1]]])
    B = np.concatenate([[3 * (slope[0] - bc0)], 3 * np.diff(slope), [3 * (bc1 - slope[-1])]])
    cp = np.zeros(G, dtype=LD); bp = np.zeros(G, dtype=LD)
    cp[0] = up[0] / di[0]; bp[0] = B[0] / di[0]
    for j in range(1, G):
        m = di[j] - lo[j] * cp[j - 1]
        cp[j] = up[j] / m; bp[j] = (B[j] - lo[j] * bp[j - 1]) / m
    X = np.zeros(G, dtype=LD); X[N] = bp[N]
    for j in range(N - 1, -1, -1):
        X[j] = bp[j] - cp[j] * X[j + 1]
    v = np.empty(G, dtype=LD)
    v[:-1] = slope - h / 3 * (2 * X[:-1] + X[1:])
    v[N] = slope[-1] + h[-1] / 3 * (X[N - 1] + 2 * X[N])
    return {"t": t, "v": v, "a": 2 * X}


def param_stage_check(oracle, res, ref, rtol, kw, names=("t", "v", "a"), max_problems=None):
    """The parametrizer stage on its real inputs: the oracle's Spline parametrizer applied to the kernel's own x and L
    (Hermite path of gen_vel_prof, sc:1213-1222) against the kernel's t, v, a.  Slack: the oracle's OWN rounding error,
    measured against the extended-precision evaluation of the same stage (per problem, its maximum over the knots)."""
    ok = np.flatnonzero(ref["status"] == 0)
    if not len(ok):
        return
    G, P = res["x"].shape
    if max_problems is None:
        max_problems = 32 if G <= 2500 else 4   # the yardstick is a Python loop over the knots
    pick = ok[np.unique(np.linspace(0, len(ok) - 1, min(len(ok), max_problems)).astype(int))]
    sc = scales(ref, ref["status"] == 0)
    pos = {p: i for i, p in enumerate(ok)}
    gp = np.linspace(0.0, 1.0, G)
    vs_all = np.broadcast_to(np.asarray(kw.get("vel_start") if kw.get("vel_start") is not None else 0.0, dtype=np.float64), (P,))
    ve_all = np.broadcast_to(np.asarray(kw.get("vel_end") if kw.get("vel_end") is not None else 0.0, dtype=np.float64), (P,))
    for p in pick:
        L, vs, ve = float(res["arclen_total"][p]), float(vs_all[p]), float(ve_all[p])
        c0 = ve - 2.0 * L + vs
        coeffs = np.array([c0, (ve - vs - 3.0 * c0) / 2.0, vs, 0.0]).reshape(1, 4, 1)
        x = np.ascontiguousarray(res["x"][:, p], dtype=np.float64)
        ts, _, _, _ = oracle.spline_param(coeffs, [0.0, 1.0], gp, x, solve_mode=1)
        if len(ts) != G:
            continue
        _, _, qd, qdd = oracle.spline_param(coeffs, [0.0, 1.0], gp, x, solve_mode=1, t_eval=ts)
        want = {"t": ts, "v": qd[:, 0], "a": qdd[:, 0]}
        exact = param_stage_exact(L, vs, ve, x)
        for n in names:
            if n not in res:
                continue
            own = float(np.max(np.abs(np.asarray(want[n], dtype=np.longdouble) - exact[n])))   # the oracle's own rounding error
            tol = rtol * np.maximum(np.abs(want[n]), 1e-3 * sc[n][pos[p]]) + SPREAD_FACTOR * own
            err = np.abs(np.asarray(res[n][:, p], dtype=np.float64) - want[n])
            assert np.all(err <= tol), ("parametrizer stage", n, int(p), float(np.max(err / tol)), own)


def _kappa(r):
    """omega / v where v is not (numerically) zero, else 0."""
    v = np.asarray(r["v"], dtype=np.float64)
    live = np.abs(v) > 1e-6 * np.max(np.abs(v), axis=0, keepdims=True)
    return np.where(live, np.asarray(r["omega"], dtype=np.float64) / np.where(live, v, 1.0), 0.0)


def scales(ref, ok):
    """Per-problem magnitude of every physical quantity (the floor of the relative error is 1e-3 of it)."""
    x = ref["x"][:, ok]; v = np.abs(ref["v"][:, ok]); a = np.abs(ref["a"][:, ok])
    L = ref["arclen_total"][ok]; T = np.maximum(ref["T_total"][ok], 1e-300)
    vs = np.maximum(v.max(axis=0), L / T)
    return {"K_lo": x.max(axis=0), "K_hi": x.max(axis=0), "x": x.max(axis=0), "t": T, "px": L, "py": L, "v": vs,
            "a": np.maximum(a.max(axis=0), vs / T), "omega": np.maximum(np.abs(ref["omega"][:, ok]).max(axis=0), vs / L)}


def check(res, ref, rtol, names, spread, spread_all=False):
    """|got - want| <= rtol * max(|want|, 1e-3 * per-problem scale) [+ SPREAD_FACTOR * measured spread for a, v, omega];
    statuses and fail stages identical.  spread_all (the randomised sweep, tools/gpu/fuzz_parity.py): the measured
    spread of EVERY quantity conditioning() was asked for is added - degenerate inputs (collinear waypoints, where the
    optimiser's gradient is rounding noise; reversals; offsets of 1e6) make quantities ill-conditioned that are not on
    ordinary paths, and the criterion stays what it is for a / v / omega: as close to the oracle as the oracle is to
    itself under perturbations of a few ulp."""
    np.testing.assert_array_equal(res["status"], ref["status"])
    np.testing.assert_array_equal(res["fail_stage"], ref["fail_stage"])
    ok = ref["status"] == 0
    if not ok.any():
        return
    sc = scales(ref, ok)
    if "a" in names or "v" in names:  # stage-wise, on the kernel's own x and L (see the header)
        param_stage_check(spread["_oracle"], res, ref, rtol, spread["_kw"])
    for n in names:
        if n in ("a", "v"):
            continue
        if n == "omega":
            # omega = v * kappa, checked factor by factor: v in the parametrizer stage above (it carries that stage's
            # rounding noise, which near the path's ends, where v -> 0, is all there is of omega), kappa = omega / v here
            live = np.abs(ref["v"][:, ok]) > 1e-6 * np.max(np.abs(ref["v"][:, ok]), axis=0, keepdims=True)
            g, w = np.where(live, _kappa(res)[:, ok], 0.0), np.where(live, ref["_kappa"][:, ok], 0.0)
            ksc = np.maximum(np.abs(w).max(axis=0), 1.0 / ref["arclen_total"][ok])   # a straight path has kappa == 0: its scale is 1 / L
            tol = rtol * np.maximum(np.abs(w), 1e-3 * ksc[None, :]) + OMEGA_FACTOR * spread["_kappa"][:, ok]
        elif n in GRID:
            g, w = res[n][:, ok], ref[n][:, ok]
            tol = rtol * np.maximum(np.abs(w), 1e-3 * sc[n][None, :])
            if n in ILL_CONDITIONED or (spread_all and n in spread):
                tol = tol + SPREAD_FACTOR * spread[n][:, ok]
        elif n in ("T_total", "cost", "arclen_total"):
            g, w = res[n][ok], ref[n][ok]
            tol = rtol * np.abs(w) + 1e-300
            if spread_all and n in spread:
                tol = tol + SPREAD_FACTOR * spread[n][ok]
        else:  # ctrl_x, ctrl_y, arclen_table: [P][n]
            g, w = res[n][ok], ref[n][ok]
            tol = rtol * np.maximum(np.abs(w), 1e-3 * np.abs(w).max(axis=1, keepdims=True)) + 1e-300
            if spread_all and n in spread:
                tol = tol + SPREAD_FACTOR * spread[n][ok]
        bad = np.abs(g - w) > tol
        if bad.any():  # say where: the worst offenders with what the tolerance was made of
            r = np.abs(g - w) / tol
            worst = [tuple(int(k) for k in np.unravel_index(j, r.shape)) for j in np.argsort(r, axis=None)[::-1][:4]]
            detail = [(ix, float(g[ix]), float(w[ix]), float(tol[ix]) if np.ndim(tol) else float(tol),
                       float(spread["_kappa" if n == "omega" else n][:, ok][ix]) if n in GRID and ("_kappa" if n == "omega" else n) in spread else None,
                       float(sc[n][ix[-1]]) if n in sc else None) for ix in worst]
            raise AssertionError((n, float(r.max()), int(bad.sum()), "worst (index, got, want, tol, spread, scale):", detail))
