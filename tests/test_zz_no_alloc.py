"""scb_param_static: the allocation-free entry point (include/seacurrent_b200.h) against the existing one.

No counterpart in the reference.  The spline parametrizer with its Thomas
scratch in a file-scope device array of the library (G * P <= SCB_STATIC_MAX_GP = 2^20), called through the C ABI on the
caller's device arrays, must return what the timing stage of scb_plan_batch_device returns for the same sdot^2 and path
lengths - on two calls in a row that share the scratch (a different, larger batch in between leaves stale contents
behind).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import seacurrent_b200 as scb

RTOL64 = 1e-9   # the FP64 tolerance of tests/test_gpu_parity.py


def _timing_batch(pl, P, G, seed):
    """The existing entry point: toppra + parametrizer on paths of given lengths, everything resident on the device."""
    import torch
    rng = np.random.default_rng(seed)
    dev = torch.device("cuda", 0)
    lim = scb.workloads.random_limits(rng, P, slow_fraction=0.05)
    host = dict(path_len=rng.uniform(0.5, 60.0, P), vel_start=rng.uniform(0.0, 0.2, P), vel_end=rng.uniform(0.0, 0.2, P), **lim)
    host["acc_lo"][1 % P] = 0.2 * host["acc_hi"][1 % P]        # cannot brake: toppra fails, the record is all zero
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in host.items()}
    bi, bo, res = pl.make_device_batch(t, G=G, stages=scb.abi.STAGE_TIMING, outputs=("K_lo", "K_hi", "x", "t", "v", "a", "T_total"))
    pl.run_device(bi, bo)
    return t, res


def _static_call(lib, t_in, res, P, G):
    import torch
    dev = res["x"].device
    status = res["status"].clone()
    out = {k: torch.full((G, P), float("nan"), dtype=torch.float64, device=dev) for k in ("t", "v", "a")}
    T = torch.full((P,), float("nan"), dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream(dev)
    rc = lib.scb_param_static(P, G, t_in["path_len"].data_ptr(), res["x"].data_ptr(), t_in["vel_start"].data_ptr(),
                              t_in["vel_end"].data_ptr(), status.data_ptr(), out["t"].data_ptr(), out["v"].data_ptr(),
                              out["a"].data_ptr(), T.data_ptr(), C.c_void_p(stream.cuda_stream))
    assert rc == 0, (rc, lib.scb_static_last_error())
    stream.synchronize()
    return status, out, T


def _compare(res, status, out, T):
    np.testing.assert_array_equal(status.cpu().numpy(), res["status"].cpu().numpy())
    ok = res["status"].cpu().numpy() == 0
    assert ok.any() and not ok.all()
    for k in ("t", "v", "a"):
        got, want = out[k].cpu().numpy(), res[k].cpu().numpy()
        assert np.all(np.isfinite(got)), k
        scale = np.maximum(np.abs(want).max(axis=0, keepdims=True), 1e-300)
        assert np.all(np.abs(got - want) <= RTOL64 * np.maximum(np.abs(want), 1e-3 * scale)), k
        assert np.all(got[:, ~ok] == 0), k      # failed problems: all-zero record, as from the existing entry
    got, want = T.cpu().numpy(), res["T_total"].cpu().numpy()
    assert np.all(np.abs(got - want) <= RTOL64 * np.abs(want))


@pytest.mark.gpu
def test_static_scratch_entry_matches_the_existing_entry_twice_in_a_row():
    lib = scb.abi.load_library()
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=4096, max_waypoints=4, max_gridpoints=1024)
    try:
        small = _timing_batch(pl, 257, 101, seed=1)
        large = _timing_batch(pl, 3000, 300, seed=2)      # 900,000 of the 1,048,576 scratch elements
        for (t_in, res), (P, G) in ((small, (257, 101)), (large, (3000, 300)), (small, (257, 101)), (small, (257, 101))):
            _compare(res, *_static_call(lib, t_in, res, P, G))
    finally:
        pl.close()


@pytest.mark.gpu
def test_static_scratch_entry_rejects_what_does_not_fit():
    import torch
    lib = scb.abi.load_library()
    d = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    s = torch.zeros(8, dtype=torch.int32, device="cuda:0")
    args = lambda P, G: (P, G, d.data_ptr(), d.data_ptr(), None, None, s.data_ptr(), d.data_ptr(), None, None, None, None)
    assert lib.scb_param_static(*args(1025, 1024)) == -5                       # SCB_E_TOO_LARGE: 1,049,600 > 2^20
    assert b"exceeds the static scratch" in lib.scb_static_last_error()
    assert b"static scratch" in lib.scb_strerror(-5)
    assert lib.scb_param_static(*args(4, 1)) == -1                             # G < 2
    assert lib.scb_param_static(4, 2, None, d.data_ptr(), None, None, s.data_ptr(), d.data_ptr(), None, None, None, None) == -1
    assert lib.scb_param_static(*args(0, 2)) == 0                              # empty batch: nothing to do
    assert lib.scb_param_static(*args(2, 4)) == 0                              # optional outputs may be NULL
    torch.cuda.synchronize()
