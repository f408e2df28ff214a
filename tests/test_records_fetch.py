"""Float32 records on the wire, rank-then-fetch, the caller-stream entry and the one-process multi-GPU planner
(include/seacurrent_b200.h: scb_batch_out.record_f32, scb_plan_batch_summary / scb_fetch, scb_plan_batch_async,
scb_multi_*).

The reference hands its caller float32 pos / vel / acc (velocity_profile holds VectorXf, sea_current.hpp:379-386,
1245-1262) and float points / angular velocities (resample sc:898, angular_velocity sc:1055): record_f32 ships exactly
that - the FP64 result rounded ONCE to float32 - so parity of the float records is equality with float32(FP64 record)."""
import numpy as np
import pytest

import seacurrent_b200 as scb


def test_shard_range_is_the_librarys():
    import ctypes as C
    lib = scb.abi.load_library()
    for P in (0, 1, 7, 1000, 65536, (1 << 20) + 3):
        for R in (1, 2, 3, 4, 8):
            for r in range(R):
                lo, hi = C.c_int32(), C.c_int32()
                lib.scb_shard_range(P, r, R, C.byref(lo), C.byref(hi))
                assert (lo.value, hi.value) == scb.shard_range(P, r, R)


def _batch(P=600, W=10, G=160, seed=3):
    w = scb.workloads.config_c3(P=P, W=W, G=G)
    rng = np.random.default_rng(seed)
    w["acc_hi"] = w["acc_hi"] * rng.uniform(0.2, 2.0, P); w["acc_lo"] = -w["acc_hi"]
    w["vel_hi"] = w["vel_hi"] * rng.uniform(0.2, 2.0, P); w["vel_lo"] = -w["vel_hi"]
    w["acc_lo"][5] = 0.2 * w["acc_hi"][5]     # one failing problem: cannot brake
    w["wx"][7, 3] = np.nan                     # one bad path
    return w


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["ref_cubic", "lau_quintic"])
def test_float_records_are_the_rounded_fp64_records(mode):
    w = _batch()
    P, G, S = len(w["wx"]), w["G"], 2048
    pl = scb.BatchPlanner(0, "fp64", mode, max_problems=P, max_waypoints=10, max_gridpoints=G, max_samples=S)
    try:
        pl.set_host_chunk(-1)
        full = pl.plan(S_cap=S, **w)
        rec = pl.plan(S_cap=S, record_f32=True, **w)
        pl.set_host_chunk(160)  # chunked: pitched copies of 4-byte records
        rec_chunked = pl.plan(S_cap=S, record_f32=True, **w)
    finally:
        pl.close()
    assert (full["status"] != 0).sum() == 2
    written = np.arange(S)[:, None] < np.where(full["status"] == 0, np.minimum(full["n_samples"], S), 0)[None, :]
    for k in full:
        if k in scb.RECORD:
            assert rec[k].dtype == np.float32
            m = written if k.startswith("s_") else np.ones_like(full[k], dtype=bool)
            np.testing.assert_array_equal(rec[k][m], full[k].astype(np.float32)[m], err_msg=k)
            np.testing.assert_array_equal(rec_chunked[k][m], rec[k][m], err_msg=k)
        else:  # status, K, x, control points, tables, scalars: untouched by the switch
            np.testing.assert_array_equal(rec[k], full[k], err_msg=k)
            np.testing.assert_array_equal(rec_chunked[k], full[k], err_msg=k)


@pytest.mark.gpu
def test_float_records_with_the_obstacle_speed_limit():
    """The obstacle-distance kernel reads the FP64 curve points: with float records they go to scratch AND to the
    caller's float arrays."""
    w = _batch(P=200)
    P, G = len(w["wx"]), w["G"]
    edges = np.array([[w["wx"][0, 2] + 1.0, w["wy"][0, 2] + 1.0, w["wx"][0, 3] + 1.5, w["wy"][0, 3] + 1.0]])
    kw = dict(obstacle_edges=edges, a_brake=1.0, t_react=0.2, collision_chords=8)
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=10, max_gridpoints=G)
    try:
        full = pl.plan(**kw, **w)
        rec = pl.plan(record_f32=True, **kw, **w)
    finally:
        pl.close()
    np.testing.assert_array_equal(rec["status"], full["status"])
    for k in ("t", "px", "py", "v", "a", "omega"):
        np.testing.assert_array_equal(rec[k], full[k].astype(np.float32), err_msg=k)
    np.testing.assert_array_equal(rec["x"], full["x"])


@pytest.mark.gpu
@pytest.mark.parametrize("record_f32", [True, False])
def test_rank_then_fetch(record_f32):
    w = _batch(P=1500)
    P, G = len(w["wx"]), w["G"]
    opt = scb.default_opt(move_waypoints=1)
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=10, max_gridpoints=G)
    try:
        full = pl.plan(opt=opt, **w)
        summ = pl.plan_summary(opt=opt, record_f32=record_f32, **w)
        np.testing.assert_array_equal(summ["status"], full["status"])
        np.testing.assert_array_equal(summ["fail_stage"], full["fail_stage"])
        for k in ("T_total", "arclen_total", "cost"):
            np.testing.assert_array_equal(summ[k], full[k])
        # what an MPC outer loop does: the k fastest feasible rollouts
        T = np.where(summ["status"] == 0, summ["T_total"], np.inf)
        ids = np.argsort(T, kind="stable")[:16].astype(np.int32)
        ids = np.concatenate([ids, [5, 7, P - 1, 0, 0]]).astype(np.int32)   # failed ones, the ends, a repeat
        got = pl.fetch(ids, outputs=scb.ALL_OUTPUTS)
        np.testing.assert_array_equal(got["status"], full["status"][ids])
        np.testing.assert_array_equal(got["fail_stage"], full["fail_stage"][ids])
        for k in scb.ALL_OUTPUTS:
            want = full[k][:, ids] if k in scb.PER_GRID else full[k][ids]
            if record_f32 and k in scb.RECORD:
                want = want.astype(np.float32)
            assert got[k].dtype == want.dtype, k
            np.testing.assert_array_equal(got[k], want, err_msg=k)
        # a second fetch from the same retained batch, other outputs
        again = pl.fetch(ids[:3], outputs=("T_total", "px"))
        np.testing.assert_array_equal(again["T_total"], full["T_total"][ids[:3]])
        with pytest.raises(scb.ScbError):
            pl.fetch(np.array([P], dtype=np.int32))
    finally:
        pl.close()


@pytest.mark.gpu
def test_caller_stream_entry():
    """scb_plan_batch_async on the caller's stream: same bits as the synchronous call; alternating streams on one ctx
    are ordered on the device (the ctx's scratch is shared)."""
    import torch
    w = _batch(P=900, G=120)
    P, G = len(w["wx"]), w["G"]
    dev = torch.device("cuda", 0)
    pl = scb.BatchPlanner(0, "fp64", "lau_quintic", max_problems=P, max_waypoints=10, max_gridpoints=G)
    try:
        t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in w.items() if k != "G"}
        bi, bo, ref = pl.make_device_batch(t, G=G)
        pl.run_device(bi, bo)
        want = {k: v.clone() for k, v in ref.items()}
        s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        outs = []
        for i, s in enumerate((s1, s2, s1, s2)):
            bi_i, bo_i, res_i = pl.make_device_batch(t, G=G)
            s.wait_stream(torch.cuda.current_stream(dev))
            pl.run_async(bi_i, bo_i, s)
            outs.append((bi_i, bo_i, res_i))
        s1.synchronize(); s2.synchronize()
        for _, _, res_i in outs:
            for k in ("status", "x", "t", "px", "omega", "T_total", "cost"):
                assert torch.equal(res_i[k], want[k]), k
    finally:
        pl.close()


def _multi_case(devices, P=2001):
    w = _batch(P=P, G=100)
    G = w["G"]
    opt = scb.default_opt(move_waypoints=1)
    one = scb.BatchPlanner(devices[0], "fp64", "lau_quintic", max_problems=P, max_waypoints=10, max_gridpoints=G)
    mp = scb.MultiPlanner(len(devices), "fp64", "lau_quintic", max_problems=P, max_waypoints=10, max_gridpoints=G, devices=devices)
    try:
        assert mp.lib.scb_multi_devices(mp.m) == len(devices)
        full = one.plan(opt=opt, outputs=scb.TRAJECTORY_OUTPUTS + ("x", "ctrl_x"), **w)
        got = mp.plan(opt=opt, outputs=scb.TRAJECTORY_OUTPUTS + ("x", "ctrl_x"), **w)
        for k in full:
            np.testing.assert_array_equal(got[k], full[k], err_msg=k)
        rec = mp.plan(opt=opt, record_f32=True, **w)
        np.testing.assert_array_equal(rec["px"], full["px"].astype(np.float32))
        summ = mp.plan_summary(opt=opt, **w)
        np.testing.assert_array_equal(summ["status"], full["status"])
        np.testing.assert_array_equal(summ["T_total"], full["T_total"])
        ids = np.array([P - 1, 0, 1000, 1001, 3, 1999, 1000], dtype=np.int32)   # both sides of every block boundary, unsorted
        f = mp.fetch(ids, outputs=("T_total", "t", "px", "omega", "x", "ctrl_x"))
        np.testing.assert_array_equal(f["status"], full["status"][ids])
        np.testing.assert_array_equal(f["T_total"], full["T_total"][ids])
        np.testing.assert_array_equal(f["x"], full["x"][:, ids])
        np.testing.assert_array_equal(f["ctrl_x"], full["ctrl_x"][ids])
        for k in ("t", "px", "omega"):
            np.testing.assert_array_equal(f[k], full[k][:, ids].astype(np.float32), err_msg=k)
    finally:
        one.close(); mp.close()


@pytest.mark.gpu
def test_multi_planner_blocks_on_one_gpu():
    """Two and three contexts on ONE device exercise everything of scb_multi_* but the second GPU: block boundaries,
    ragged last block, one host thread per context, routed fetch."""
    _multi_case([0, 0])
    _multi_case([0, 0, 0])


@pytest.mark.gpu
def test_multi_planner_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU on this box (covered on one device by test_multi_planner_blocks_on_one_gpu)")
    _multi_case([0, 1])
