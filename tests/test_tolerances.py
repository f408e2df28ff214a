"""The measurements behind tests/tolerances.py (CPU only).

Every slack term a GPU parity test uses is the oracle's OWN observed movement under input perturbations of a few ulp;
these tests pin what that measurement shows, so that nobody has to take a conditioning argument on trust:
  * the quantities held to a plain 1e-9 move by less than a tenth of that (< 1e-10 relative; x and K, which the passes
    propagate along the path with gains up to ~1e4, move most: 2e-11),
  * a, v, omega do move by more than 1e-9 at some gridpoints - the reason they carry the measured term,
  * the size of the movement of `a` is the eps * L / h_min^2 the conditioning argument predicts,
  * a second, equally legitimate rounding of the same mathematics (the reference's dense column-pivoted QR for the
    parametrizer spline instead of Thomas, tp:geometric_path/piecewise_poly_path.cpp:112-163) passes the very check the
    GPU has to pass - the tolerance is not tighter than two correct implementations can meet."""
import numpy as np
import pytest

import seacurrent_b200 as scb
import tolerances

EPS = np.finfo(np.float64).eps
WELL = ("x", "K_lo", "K_hi", "t", "px", "py", "ctrl_x", "ctrl_y", "arclen_table", "arclen_total", "T_total", "cost")


def _batch(P, W, seed):
    rng = np.random.default_rng(seed)
    wx, wy = scb.workloads.random_walk_paths(rng, P, W)
    return dict(wx=wx, wy=wy, **scb.workloads.random_limits(rng, P, slow_fraction=0.0))


@pytest.mark.parametrize("mode,P,W,G", [(0, 48, 20, 500), (1, 48, 10, 1000), (1, 6, 200, 10000)])
def test_what_moves_and_what_does_not(oracle, mode, P, W, G):
    kw = dict(G=G, spline_mode=mode, **_batch(P, W, 77 + W))
    ref, spread = tolerances.conditioning(oracle, names=WELL + ("a", "v", "omega"), **kw)
    ok = ref["status"] == 0
    assert ok.all()
    sc = tolerances.scales(ref, ok)
    # held to a plain 1e-9 in the parity tests: the oracle itself is 10x steadier than that
    for n in WELL:
        w = ref[n]
        if n in tolerances.GRID:
            rel = spread[n] / np.maximum(np.abs(w), 1e-3 * sc[n][None, :])
        elif w.ndim == 2:
            rel = spread[n] / np.maximum(np.abs(w), 1e-3 * np.abs(w).max(axis=1, keepdims=True))
        else:
            rel = spread[n] / np.maximum(np.abs(w), 1e-300)   # cost is identically 0 in cubic mode
        assert rel.max() < 1e-10, (n, rel.max())
    # carry the measured term: the oracle's own value moves by more than 1e-9 of the plain tolerance somewhere
    moved = {n: float((spread[n] / (1e-9 * np.maximum(np.abs(ref[n]), 1e-3 * sc[n][None, :]))).max()) for n in ("a", "v", "omega")}
    assert moved["a"] > 1.0, moved
    if W >= 200:
        assert moved["omega"] > 1.0, moved   # long quintic paths have near-cusps: curvature spikes
    # ... and for `a` by what the conditioning argument says: ~ eps * L / h_min^2 per ulp of perturbation
    h_min = np.diff(ref["t"], axis=0).min(axis=0)
    model = EPS * ref["arclen_total"] / h_min ** 2
    ratio = spread["a"].max(axis=0) / (tolerances.PROBE_ULPS * model)
    assert np.all(ratio < 300.0) and np.median(ratio) > 0.05, (ratio.min(), np.median(ratio), ratio.max())


@pytest.mark.parametrize("mode", [0, 1])
def test_dense_qr_and_thomas_pass_each_others_check(oracle, mode):
    """Two roundings of the same mathematics: the reference's dense QR solve of the parametrizer spline against the
    tridiagonal solve the kernels use.  They differ by far more than 1e-9 in `a`, and pass tolerances.check."""
    kw = dict(G=201, spline_mode=mode, **_batch(24, 8, 5))
    ref, spread = tolerances.conditioning(oracle, solve_mode=1, **kw)       # Thomas
    qr, _ = oracle.plan_batch(solve_mode=0, **kw)                            # dense column-pivoted QR, as the reference
    tolerances.check(qr, ref, 1e-9, ("T_total", "arclen_total") + tolerances.GRID, spread)
    sc = tolerances.scales(ref, ref["status"] == 0)
    plain = np.abs(qr["a"] - ref["a"]) / (1e-9 * np.maximum(np.abs(ref["a"]), 1e-3 * sc["a"][None, :]))
    assert plain.max() > 1e-3   # the two solves do not agree bit for bit: the check above is not vacuous


def test_recorded_kernel_outputs_against_the_extended_precision_yardstick(oracle, golden_dir):
    """tests/golden/gpu_param_stage_r2.npz holds what param_kernel produced on a B200 (round 2, tools/gpu/diag_param.py)
    for the first problems of two parity batches: x and L in, t / v / a out.  Against the 80-bit evaluation of the same
    stage the KERNEL is closer to the exact answer than the oracle is (Thomas with FMA contraction against the oracle's
    pow()-based knot values), and kernel and oracle differ by no more than twice the oracle's own error - which is why
    tolerances.param_stage_check allows SPREAD_FACTOR times that error and nothing modelled."""
    import os
    d = np.load(os.path.join(golden_dir, "gpu_param_stage_r2.npz"))
    closer = total = 0
    for tag in ("ref_cubic_20", "lau_quintic_10"):
        G = d[tag + "_x"].shape[0]
        gp = np.linspace(0.0, 1.0, G)
        for p in range(d[tag + "_x"].shape[1]):
            if d[tag + "_status"][p] != 0:
                continue
            L = float(d[tag + "_arclen_total"][p]); x = np.ascontiguousarray(d[tag + "_x"][:, p])
            coeffs = np.array([-2 * L, 3 * L, 0.0, 0.0]).reshape(1, 4, 1)
            ts, _, _, _ = oracle.spline_param(coeffs, [0.0, 1.0], gp, x, solve_mode=1)
            _, _, qd, qdd = oracle.spline_param(coeffs, [0.0, 1.0], gp, x, solve_mode=1, t_eval=ts)
            ex = tolerances.param_stage_exact(L, 0.0, 0.0, x)
            for n, ora in (("a", qdd[:, 0]), ("v", qd[:, 0])):
                own = float(np.max(np.abs(ora - ex[n])))
                gpu = float(np.max(np.abs(d[tag + "_" + n][:, p] - ex[n])))
                diff = float(np.max(np.abs(d[tag + "_" + n][:, p] - ora)))
                assert diff <= 2.0 * own + 1e-9 * 1e-3 * np.abs(ora).max(), (tag, p, n, diff, own)
                closer += gpu <= own; total += 1
            # the rounding noise on `a` is far above the plain bar where a crosses zero, and far below the bar elsewhere
            assert 1e-13 < np.max(np.abs(qdd[:, 0] - ex["a"])) < 1e-8
            assert np.max(np.abs(ts - ex["t"])) <= 1e-14 * ts[-1]
    assert closer >= 0.8 * total, (closer, total)
