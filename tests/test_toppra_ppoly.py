"""The reference's own known-answer vectors through a KERNEL (missing item 4).

third-party/toppra/cpp/tests/test_algorithm.cpp:64-165 pins toppra's controllable sets, parametrization and feasible sets
on a 2-DOF, three-segment cubic path (velocity +-1, acceleration +-0.2, collocation, N = 50, Seidel) to 1e-6.  With two
degrees of freedom and collocation the stage LP has 2 + 2*2 = 6 rows - exactly the LP the product's Seidel kernel
unrolls (solve_lp2d6, toppra_kernel.cu) - so scb_toppra_ppoly runs that same device function on the reference's
configuration: the 51-entry vectors reach the GPU code directly, not only the oracle."""
import json
import os

import numpy as np
import pytest

import seacurrent_b200 as scb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def planner():
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=64, max_waypoints=4, max_gridpoints=101)
    yield pl
    pl.close()


def test_reference_known_answer_vectors_on_the_gpu(planner, golden_dir, oracle):
    g = json.load(open(os.path.join(golden_dir, "toppra_cpp_golden.json")))["algorithm"]
    coeffs = np.array(g["coeffs"]).reshape(1, 3, 4, 2)
    G = g["N"] + 1
    assert g["dof"] == 2 and not g["interpolation"]
    res = planner.toppra_ppoly(coeffs, g["breaks"], G, [g["vel"][0]] * 2, [g["vel"][1]] * 2, [g["acc"][0]] * 2, [g["acc"][1]] * 2, 0)
    assert res["status"][0] == 0 and res["feas_status"][0] == 0
    # the reference's expectations, at the reference's tolerance (test_algorithm.cpp:106-113, 129-136, 158-165)
    np.testing.assert_allclose(res["K_hi"][:, 0], g["K_max"], atol=g["tol"])
    np.testing.assert_allclose(res["x"][:, 0], g["parametrization"], atol=g["tol"])
    np.testing.assert_allclose(res["feas_hi"][:, 0], g["feasible_max"], atol=g["tol"])
    # ... and the oracle's Seidel restatement bit for bit
    rc, K, x, feas, frc = oracle.toppra_ppoly(coeffs[0], g["breaks"], G, [g["vel"][0]] * 2, [g["vel"][1]] * 2, [g["acc"][0]] * 2,
                                              [g["acc"][1]] * 2, 0, feasible=True)
    assert rc == 0 and frc == 0
    np.testing.assert_array_equal(res["K_lo"][:, 0], K[:, 0]); np.testing.assert_array_equal(res["K_hi"][:, 0], K[:, 1])
    np.testing.assert_array_equal(res["x"][:, 0], x)
    np.testing.assert_array_equal(res["feas_lo"][:, 0], feas[:, 0]); np.testing.assert_array_equal(res["feas_hi"][:, 0], feas[:, 1])


@pytest.mark.parametrize("dof,interpolation", [(2, 0), (1, 0), (1, 1)])
def test_random_paths_bit_exact_vs_oracle(planner, oracle, dof, interpolation):
    """Random piecewise-cubic paths (C1 joins through random knots), random limits, non-zero end path velocities: every
    configuration the six-row LP covers, bit for bit against the oracle, verdicts included (tight limits make some fail)."""
    rng = np.random.default_rng(10 * dof + interpolation)
    P, ns, G = 40, 4, 61
    breaks = np.array([0.0, 0.7, 1.5, 2.1, 3.0])
    # cubic Hermite segments through random knots with random knot velocities (kept away from 0 in dof 0 so that
    # q' = 0 inside the path, where a velocity box that excludes 0 throws in the reference, does not occur)
    q = np.cumsum(rng.uniform(0.5, 2.0, (P, ns + 1, dof)), axis=1)
    v = rng.uniform(0.3, 2.0, (P, ns + 1, dof))
    coeffs = np.zeros((P, ns, 4, dof))
    for s in range(ns):
        dt = breaks[s + 1] - breaks[s]
        c3, c2 = q[:, s], v[:, s]
        c0 = (v[:, s + 1] * dt - 2 * q[:, s + 1] + c2 * dt + 2 * c3) / dt ** 3
        c1 = (v[:, s + 1] - c2 - 3 * c0 * dt * dt) / (2 * dt)
        coeffs[:, s, 0], coeffs[:, s, 1], coeffs[:, s, 2], coeffs[:, s, 3] = c0, c1, c2, c3
    vhi = rng.uniform(0.5, 3.0, (P, dof)); ahi = rng.uniform(0.05, 2.0, (P, dof))
    res = planner.toppra_ppoly(coeffs, breaks, G, -vhi, vhi, -ahi, ahi, interpolation, sd_start=0.1, sd_end=0.05)
    st = []
    for p in range(P):
        rc, K, x, feas, frc = oracle.toppra_ppoly(coeffs[p], breaks, G, -vhi[p], vhi[p], -ahi[p], ahi[p], interpolation, sd_start=0.1,
                                                  sd_end=0.05, feasible=True)
        st.append(rc)
        assert res["status"][p] == rc, p
        if rc == 0:
            np.testing.assert_array_equal(res["K_lo"][:, p], K[:, 0]); np.testing.assert_array_equal(res["K_hi"][:, p], K[:, 1])
            np.testing.assert_array_equal(res["x"][:, p], x)
        if frc == 0:
            assert res["feas_status"][p] == 0
            np.testing.assert_array_equal(res["feas_lo"][:, p], feas[:, 0]); np.testing.assert_array_equal(res["feas_hi"][:, p], feas[:, 1])
        else:
            assert res["feas_status"][p] == 5
    assert st.count(0) >= P // 4


def test_one_dof_interpolation_equals_the_product_kernel(planner):
    """gen_vel_prof's configuration (1-DOF Hermite path on [0, 1], interpolation) through BOTH entries: the generic-path
    kernel and the product's fused Seidel kernel give the same bits."""
    rng = np.random.default_rng(3)
    P, G = 32, 101
    L = rng.uniform(1.0, 50.0, P)
    coeffs = np.zeros((P, 1, 4, 1))
    coeffs[:, 0, 0, 0] = -2 * L; coeffs[:, 0, 1, 0] = 3 * L   # q = L (3 s^2 - 2 s^3), piecewise_poly_path.cpp:364-375 with zero end slopes
    vhi = rng.uniform(0.3, 2.0, P); ahi = rng.uniform(0.2, 2.0, P)
    a = planner.toppra_ppoly(coeffs, [0.0, 1.0], G, -vhi[:, None], vhi[:, None], -ahi[:, None], ahi[:, None], 1, feasible=False)
    b = planner.plan(path_len=L, G=G, stages=scb.abi.STAGE_TIMING, lp_mode=scb.abi.LP_SEIDEL, vel_lo=-vhi, vel_hi=vhi, acc_lo=-ahi, acc_hi=ahi,
                     outputs=("K_lo", "K_hi", "x"))
    assert np.all(a["status"] == 0) and np.all(b["status"] == 0)
    for k in ("K_lo", "K_hi", "x"):
        np.testing.assert_array_equal(a[k], b[k])
