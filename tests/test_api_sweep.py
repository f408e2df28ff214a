"""Randomised consistency sweep over the ENTRY POINTS of the C ABI (pinned seed): for shapes and options drawn by
tests/fuzz_cases.py plus obstacles, a uniform-dt sample stream and float records, every way of asking for the same batch
must return the same bits -
  one host chunk  ==  a random chunk size (pitched copies, ragged last chunk)
  full plan       ==  summary + fetch of random problem ids (repeats, both ends)
  host entry      ==  device entry (scb_plan_batch_device on the caller's arrays)
  one context     ==  scb_multi_* over two / three contexts on one device
  FP64 records    ->  float records are their one-time rounding
Problems are independent and every kernel's arithmetic is per problem, so equality is exact; what the sweep exercises is
indexing: strides, ragged tails, odd P against 32 / 64 / 128-thread CTAs, several paths per warp, chunk boundaries."""
import os

import numpy as np
import pytest

import fuzz_cases
import seacurrent_b200 as scb

# a longer one-off sweep: SCB_SWEEP_SEED=.. SCB_SWEEP_CASES=.. python -m pytest tests/test_api_sweep.py -m gpu
SEED, CASES = int(os.environ.get("SCB_SWEEP_SEED", 777)), int(os.environ.get("SCB_SWEEP_CASES", 60))


def make_api_case(rng):
    c = fuzz_cases.make_case(rng)
    c["P"] = int(rng.choice([1, 5, 31, 33, 63, 64, 65, 127, 129, 200, 257]))
    c["G"] = int(rng.choice([2, 3, 17, 64, 101, 130]))
    if c["shape"] in ("line", "far"):
        c["shape"] = "walk"
    c["api_S"] = int(rng.choice([0, 0, 64, 300]))
    c["chunk"] = int(rng.choice([1, 7, 32, 50, 64, 100]))
    c["api_obst"] = bool(rng.random() < 0.35) and c["W"] <= 48
    c["rec32"] = bool(rng.random() < 0.5)
    c["multi"] = int(rng.choice([0, 2, 3]))
    return c


def api_kwargs(c):
    kw = fuzz_cases.build(c)
    rng = np.random.default_rng(c["seed"] + 1)
    if c["api_obst"]:
        wx, wy = kw["wx"], kw["wy"]
        k = rng.integers(0, c["W"] - 1, 10)
        mx, my = (wx[0, k] + wx[0, k + 1]) / 2, (wy[0, k] + wy[0, k + 1]) / 2
        off = rng.uniform(-2.0, 2.0, (10, 2)); d = rng.uniform(-1.0, 1.0, (10, 2))
        kw["obstacle_edges"] = np.ascontiguousarray(np.stack([mx + off[:, 0] - d[:, 0], my + off[:, 1] - d[:, 1],
                                                              mx + off[:, 0] + d[:, 0], my + off[:, 1] + d[:, 1]], 1))
        kw["collision_chords"] = 8
        if rng.random() < 0.5:
            kw["a_brake"], kw["t_react"] = 1.0, 0.2
    if c["api_S"]:
        kw["S_cap"] = c["api_S"]
        kw["dt"] = float(rng.choice([0.02, 0.1, 0.5]))
    return kw


def _same(got, one, S, what):
    assert set(got) == set(one), what
    ns = one.get("n_samples")
    written = None if ns is None else np.arange(S)[:, None] < np.where(one["status"] == 0, np.minimum(ns, S), 0)[None, :]
    for k in one:
        if k.startswith("s_"):  # rows beyond n_samples[p] are not written
            np.testing.assert_array_equal(got[k][written], one[k][written], err_msg="%s %s" % (what, k))
        else:
            np.testing.assert_array_equal(got[k], one[k], err_msg="%s %s" % (what, k))


def run_api_case(c, planner):
    import torch
    kw = api_kwargs(c)
    P, W, G, S = c["P"], c["W"], c["G"], c["api_S"]
    lp = c["lp"]
    planner.set_host_chunk(-1)
    one = planner.plan(lp_mode=lp, **kw)
    # ---- chunks
    planner.set_host_chunk(c["chunk"])
    _same(planner.plan(lp_mode=lp, **kw), one, S, "chunk=%d" % c["chunk"])
    # ---- float records
    if c["rec32"]:
        rec = planner.plan(lp_mode=lp, record_f32=True, **kw)
        written = np.arange(max(S, 1))[:, None] < np.where(one["status"] == 0, np.minimum(one.get("n_samples", np.zeros(P, int)), S), 0)[None, :]
        for k in one:
            if k in scb.RECORD:
                m = written[:S] if k.startswith("s_") else np.ones_like(one[k], dtype=bool)
                np.testing.assert_array_equal(rec[k][m], one[k].astype(np.float32)[m], err_msg="rec32 " + k)
            elif not k.startswith("s_") and k != "n_samples":
                np.testing.assert_array_equal(rec[k], one[k], err_msg="rec32 " + k)
    planner.set_host_chunk(-1)
    # ---- summary + fetch
    skw = {k: v for k, v in kw.items() if k not in ("S_cap", "dt")}
    summ = planner.plan_summary(lp_mode=lp, record_f32=c["rec32"], **skw)
    np.testing.assert_array_equal(summ["status"], one["status"])
    np.testing.assert_array_equal(summ["fail_stage"], one["fail_stage"])
    np.testing.assert_array_equal(summ["T_total"], one["T_total"])
    rng = np.random.default_rng(c["seed"] + 2)
    ids = np.concatenate([rng.integers(0, P, 6), [0, P - 1, 0]]).astype(np.int32)
    names = tuple(k for k in scb.ALL_OUTPUTS if k in one)
    got = planner.fetch(ids, outputs=names)
    for k in names:
        want = one[k][:, ids] if k in scb.PER_GRID else one[k][ids]
        if c["rec32"] and k in scb.RECORD:
            want = want.astype(np.float32)
        np.testing.assert_array_equal(got[k], want, err_msg="fetch " + k)
    # ---- device entry
    dev = torch.device("cuda", 0)
    dkw = {k: v for k, v in skw.items() if not isinstance(v, np.ndarray)}
    R = planner.real   # the context's storage type: plan() converts host arrays the same way
    t = {k: torch.from_numpy(np.ascontiguousarray(v, dtype=R)).to(dev) for k, v in skw.items() if isinstance(v, np.ndarray) and k != "obstacle_edges"}
    if "obstacle_edges" in skw:
        dkw["obstacle_edges"] = torch.from_numpy(np.ascontiguousarray(skw["obstacle_edges"], dtype=R)).to(dev)
    bi, bo, res = planner.make_device_batch(t, lp_mode=lp, **dkw)
    planner.run_device(bi, bo)
    for k, v in res.items():
        if k in one and not k.startswith("_"):
            np.testing.assert_array_equal(v.cpu().numpy(), one[k], err_msg="device entry " + k)
    return one


def _cases():
    rng = np.random.default_rng(SEED)
    return [make_api_case(rng) for _ in range(CASES)]


def test_api_cases_are_reproducible():
    a, b = _cases(), _cases()
    assert a == b
    assert any(c["api_obst"] for c in a) and any(c["api_S"] for c in a) and any(c["rec32"] for c in a) and any(c["multi"] for c in a)
    for c in a[:5]:
        kw = api_kwargs(c)
        assert kw["wx"].shape == (c["P"], c["W"])


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_entry_points_agree_bit_for_bit(precision):
    planners = {m: scb.BatchPlanner(0, precision, m, max_problems=300, max_waypoints=80, max_gridpoints=160, max_samples=320)
                for m in ("ref_cubic", "lau_quintic")}
    failures = []
    feasible = total = 0
    try:
        for k, c in enumerate(_cases()):
            try:
                one = run_api_case(c, planners[c["mode"]])
                feasible += int((one["status"] == 0).sum()); total += c["P"]
                if c["multi"]:
                    kw = api_kwargs(c)
                    skw = {a: v for a, v in kw.items() if a not in ("S_cap", "dt")}
                    mp = scb.MultiPlanner(c["multi"], precision, c["mode"], max_problems=c["P"], max_waypoints=c["W"],
                                          max_gridpoints=c["G"], devices=[0] * c["multi"])
                    try:
                        names = tuple(n for n in scb.TRAJECTORY_OUTPUTS + ("x", "ctrl_x", "ctrl_y") if n in one)
                        got = mp.plan(lp_mode=c["lp"], outputs=names, **skw)
                        for n in got:
                            np.testing.assert_array_equal(got[n], one[n], err_msg="multi %d %s" % (c["multi"], n))
                    finally:
                        mp.close()
            except (AssertionError, scb.ScbError) as e:
                failures.append((k, c, str(e)[:500]))
                if isinstance(e, scb.ScbError):
                    break
    finally:
        for p in planners.values():
            p.close()
    assert not failures, failures
    assert feasible > 0.7 * total
