"""Batched planner over the C ABI (include/seacurrent_b200.h) — test / bench harness side.

The product's host layer for C++ callers is include/sea_current_b200.hpp (namespace turtle::sc,
the reference's signatures).  This module is the thin ctypes equivalent used by tests/ and
bench.py: it only marshals numpy / torch buffers into scb_batch_in / scb_batch_out and calls
scb_plan_batch (host buffers) or scb_plan_batch_device (HBM-resident torch tensors).  There is
no planning arithmetic here and no fallback: if the library cannot be loaded or no sm_100
device exists, construction raises.
"""
import ctypes as C
import os

import numpy as np

from . import abi

PER_PROBLEM = ("arclen_total", "T_total", "cost")
PER_GRID = ("K_lo", "K_hi", "x", "t", "px", "py", "v", "a", "omega")
PER_SAMPLE = ("s_time", "s_pos", "s_vel", "s_acc", "s_px", "s_py", "s_omega")
ALL_OUTPUTS = ("ctrl_x", "ctrl_y", "arclen_table") + PER_PROBLEM + PER_GRID
# what the reference's E1 call sequence hands back to its caller (gen_vel_prof + resample +
# angular_velocity, examples/json_test.cpp:28-58), one record per toppra gridpoint
TRAJECTORY_OUTPUTS = ("arclen_total", "T_total", "t", "px", "py", "v", "a", "omega")
# the arrays scb_batch_out.record_f32 turns into float32 whatever the ctx precision (the reference's velocity_profile,
# resample and angular_velocity hand their caller float32: sea_current.hpp:379-386, 1245-1262)
RECORD = ("t", "px", "py", "v", "a", "omega") + PER_SAMPLE


class ScbError(RuntimeError):
    pass


def default_opt(**kw):
    o = abi.ScbOpt()
    abi.load_library().scb_default_opt(C.byref(o))
    for k, v in kw.items():
        setattr(o, "lambda_" if k == "lambda" else k, v)
    return o


class BatchResult(dict):
    """dict of output arrays (+ 'status', 'fail_stage', optionally 'n_samples') with attribute access."""
    __getattr__ = dict.__getitem__


class BatchPlanner:
    def __init__(self, device=0, precision="fp64", spline_mode="ref_cubic", max_problems=65536, max_waypoints=32,
                 max_gridpoints=1001, max_samples=0):
        self.lib = abi.load_library()
        self.precision = {"fp64": abi.SCB_FP64, "fp32": abi.SCB_FP32}[precision]
        self.spline_mode = {"ref_cubic": abi.SCB_REF_CUBIC, "lau_quintic": abi.SCB_LAU_QUINTIC}[spline_mode]
        self.real = np.float64 if self.precision == abi.SCB_FP64 else np.float32
        self.degree = 3 if self.spline_mode == abi.SCB_REF_CUBIC else 5
        self.device = device
        cfg = abi.ScbConfig(device, self.precision, self.spline_mode, max_problems, max_waypoints, max_gridpoints,
                            max_samples)
        h = C.c_void_p()
        rc = self.lib.scb_create(C.byref(h), C.byref(cfg))
        if rc != 0:
            raise ScbError("scb_create: %s" % self.lib.scb_strerror(rc).decode())
        self.ctx = h

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.scb_destroy(self.ctx)
            self.ctx = None

    __del__ = close

    def _check(self, rc, what):
        if rc != 0:
            raise ScbError("%s: %s [%s]" % (what, self.lib.scb_strerror(rc).decode(),
                                            self.lib.scb_last_cuda_error(self.ctx).decode()))

    def timing(self):
        t = abi.ScbTiming()
        self.lib.scb_get_timing(self.ctx, C.byref(t))
        return {k: getattr(t, k) for k, _ in abi.ScbTiming._fields_}

    def stream(self):
        return self.lib.scb_stream(self.ctx)

    def synchronize(self):
        self._check(self.lib.scb_synchronize(self.ctx), "scb_synchronize")

    def set_host_chunk(self, problems):
        """Problems per pipelined chunk of plan() (scb_set_host_chunk): 0 = automatic, < 0 = one chunk."""
        self._check(self.lib.scb_set_host_chunk(self.ctx, int(problems)), "scb_set_host_chunk")

    def algorithmic_bytes(self, P, W, G):
        return self.lib.scb_algorithmic_bytes(self.precision, self.spline_mode, P, W, G)

    def out_shape(self, name, P, W, G, S_cap):
        nseg = W - 1
        if name in ("ctrl_x", "ctrl_y"):
            return (P, nseg * (self.degree + 1))
        if name == "arclen_table":
            return (P, nseg * abi.ARCLEN_TABLE)
        if name in PER_PROBLEM:
            return (P,)
        if name in PER_GRID:
            return (G, P)
        if name in PER_SAMPLE:
            return (S_cap, P)
        raise KeyError(name)

    # ------------------------------------------------------------------ global planner front end
    def fmt_batch(self, nodes, edges, init, goal, rn, max_path=256):
        """FMT* for Q (init, goal) queries over one shared node set and obstacle edge list (scb_fmt_batch; replaces
        planning_space::fast_marching_trees, sea_current.hpp:1339-1407).  nodes [n][2], edges [E][4] or None,
        init / goal [Q][2]; float32 whatever the planner's precision."""
        f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        nodes = f32(nodes).reshape(-1, 2); init = f32(init).reshape(-1, 2); goal = f32(goal).reshape(-1, 2)
        ed = f32(edges).reshape(-1, 4) if edges is not None and len(edges) else None
        Q = len(init)
        keep = [f32(nodes[:, 0]), f32(nodes[:, 1]), f32(init[:, 0]), f32(init[:, 1]), f32(goal[:, 0]), f32(goal[:, 1])]
        fi = abi.ScbFmtIn()
        fi.Q, fi.n_nodes, fi.n_edges, fi.max_path, fi.rn = Q, len(nodes), 0 if ed is None else len(ed), max_path, rn
        fi.node_x, fi.node_y, fi.init_x, fi.init_y, fi.goal_x, fi.goal_y = (a.ctypes.data for a in keep)
        fi.edges = None if ed is None else ed.ctypes.data
        res = BatchResult(status=np.empty(Q, np.int32), path_len=np.empty(Q, np.int32), closed=np.empty(Q, np.int32),
                          path_x=np.zeros((Q, max_path), np.float32), path_y=np.zeros((Q, max_path), np.float32),
                          cost=np.empty(Q, np.float32))
        fo = abi.ScbFmtOut(*(res[k].ctypes.data for k in ("status", "path_len", "closed", "path_x", "path_y", "cost")))
        self._check(self.lib.scb_fmt_batch(self.ctx, C.byref(fi), C.byref(fo)), "scb_fmt_batch")
        return res

    # ------------------------------------------------------------------ toppra on a caller-supplied path
    def toppra_ppoly(self, coeffs, breaks, G, vel_lo, vel_hi, acc_lo, acc_hi, interpolation=0, sd_start=0.0, sd_end=0.0, x_cap=0.0,
                     feasible=True, vel_lo_grid=None, vel_hi_grid=None):
        """scb_toppra_ppoly: coeffs [P][n_seg][4][dof] (row 0 multiplies ds^3), breaks [n_seg + 1], limits [P][dof];
        vel_*_grid [G][dof][P] replace the constant velocity limits when given."""
        d = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        coeffs = d(coeffs); P, ns, _, m = coeffs.shape
        keep = [coeffs, d(breaks)] + [None if a is None else d(np.broadcast_to(d(a), (P, m))) for a in (vel_lo, vel_hi, acc_lo, acc_hi)]
        pi = abi.ScbPpolyIn()
        pi.P, pi.n_seg, pi.dof, pi.G = P, ns, m, G
        pi.coeffs, pi.breaks, pi.vel_lo, pi.vel_hi, pi.acc_lo, pi.acc_hi = (None if a is None else a.ctypes.data for a in keep)
        if vel_lo_grid is not None:
            keep += [d(np.broadcast_to(d(vel_lo_grid), (G, m, P))), d(np.broadcast_to(d(vel_hi_grid), (G, m, P)))]
            pi.vel_lo_grid, pi.vel_hi_grid = keep[-2].ctypes.data, keep[-1].ctypes.data
        pi.interpolation, pi.sd_start, pi.sd_end, pi.x_cap = int(interpolation), sd_start, sd_end, x_cap
        res = BatchResult(status=np.empty(P, np.int32), fail_stage=np.empty(P, np.int32), feas_status=np.empty(P, np.int32),
                          K_lo=np.empty((G, P)), K_hi=np.empty((G, P)), x=np.empty((G, P)))
        if feasible:
            res["feas_lo"] = np.empty((G, P)); res["feas_hi"] = np.empty((G, P))
        po = abi.ScbPpolyOut()
        for k, v in res.items():
            setattr(po, k, v.ctypes.data)
        self._check(self.lib.scb_toppra_ppoly(self.ctx, C.byref(pi), C.byref(po)), "scb_toppra_ppoly")
        return res

    def gen_vel_prof(self, pos_start, pos_end, vel_lo, vel_hi, acc_lo, acc_hi, G=101, vel_start=None, vel_end=None,
                     vel_lo_grid=None, vel_hi_grid=None, interpolation=1, dt=0.0, x_cap=0.0, param_mode=0, S_cap=0):
        """scb_gen_vel_prof: gen_vel_prof<N> (sc:1191-1265) for a batch.  pos_* / vel_* / acc_* [P][dof]; returns status,
        fail_stage, K_lo, K_hi, x, t [G][P], T_total [P] and, with S_cap > 0, n_samples, s_time [S][P], s_pos / s_vel /
        s_acc [dof][S][P]."""
        d = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        ps = d(np.atleast_2d(pos_start)); P, m = ps.shape
        bc = lambda a: None if a is None else d(np.broadcast_to(d(a), (P, m)))
        keep = [ps, bc(pos_end), bc(vel_start), bc(vel_end), bc(vel_lo), bc(vel_hi), bc(acc_lo), bc(acc_hi)]
        ni = abi.ScbNdofIn()
        ni.P, ni.dof, ni.G, ni.interpolation = P, m, G, int(interpolation)
        (ni.pos_start, ni.pos_end, ni.vel_start, ni.vel_end, ni.vel_lo, ni.vel_hi, ni.acc_lo, ni.acc_hi) = (
            None if a is None else a.ctypes.data for a in keep)
        if vel_lo_grid is not None:
            keep += [d(np.broadcast_to(d(vel_lo_grid), (G, m, P))), d(np.broadcast_to(d(vel_hi_grid), (G, m, P)))]
            ni.vel_lo_grid, ni.vel_hi_grid = keep[-2].ctypes.data, keep[-1].ctypes.data
        ni.dt, ni.x_cap, ni.param_mode = dt, x_cap, param_mode
        res = BatchResult(status=np.empty(P, np.int32), fail_stage=np.empty(P, np.int32), K_lo=np.empty((G, P)),
                          K_hi=np.empty((G, P)), x=np.empty((G, P)), t=np.empty((G, P)), T_total=np.empty(P))
        no = abi.ScbNdofOut()
        no.S_cap = S_cap
        if S_cap > 0:
            res["n_samples"] = np.empty(P, np.int32); res["s_time"] = np.empty((S_cap, P))
            for k in ("s_pos", "s_vel", "s_acc"):
                res[k] = np.empty((m, S_cap, P))
        for k, v in res.items():
            setattr(no, k, v.ctypes.data)
        self._check(self.lib.scb_gen_vel_prof(self.ctx, C.byref(ni), C.byref(no)), "scb_gen_vel_prof")
        return res

    # ------------------------------------------------------------------ host buffers
    def plan(self, wx=None, wy=None, vel_lo=None, vel_hi=None, acc_lo=None, acc_hi=None, G=101, start_angle=None,
             vel_lo_grid=None, vel_hi_grid=None, vel_start=None, vel_end=None, path_len=None, stages=0,
             outputs=ALL_OUTPUTS, S_cap=0, x_cap=0.0, dt=0.0, opt=None, out_buffers=None, resample_at=None, lp_mode=0,
             nudge=False, omega_max=0.0, acc_lat_max=0.0, obstacle_edges=None, param_mode=0, a_brake=0.0, t_react=0.0, collision_chords=0,
             spline=None, record_f32=False):
        """Host arrays in, host arrays out (scb_plan_batch).  `out_buffers` lets a caller pass
        preallocated (e.g. pinned) numpy arrays keyed by output name.  `spline` = dict(ctrl_x, ctrl_y, arclen_table) of an
        earlier SCB_STAGE_SPLINE call: inputs of a SCB_STAGE_GEOMETRY-only call (no waypoints then)."""
        R = self.real
        keep = []

        def arr(a, shape=None):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=R)
            if shape is not None and a.shape != shape:
                a = np.ascontiguousarray(np.broadcast_to(a, shape))
            keep.append(a)
            return a.ctypes.data

        if wx is not None:
            wx = np.ascontiguousarray(wx, dtype=R)
            P, W = wx.shape
        elif spline is not None:
            P = spline["ctrl_x"].shape[0]
            W = spline["ctrl_x"].shape[1] // (self.degree + 1) + 1
        else:
            P, W = len(np.atleast_1d(path_len)), 2
        bi = abi.ScbBatchIn()
        bi.P, bi.W, bi.G = P, W, G
        bi.wx, bi.wy = arr(wx), arr(wy, (P, W) if wx is not None else None)
        bi.start_angle = arr(start_angle, (P,))
        bi.vel_lo, bi.vel_hi = arr(vel_lo, (P,)), arr(vel_hi, (P,))
        bi.vel_lo_grid, bi.vel_hi_grid = arr(vel_lo_grid, (G, P)), arr(vel_hi_grid, (G, P))
        bi.acc_lo, bi.acc_hi = arr(acc_lo, (P,)), arr(acc_hi, (P,))
        bi.vel_start, bi.vel_end = arr(vel_start, (P,)), arr(vel_end, (P,))
        bi.path_len = arr(path_len, (P,))
        bi.stages, bi.x_cap, bi.dt, bi.lp_mode = stages, x_cap, dt, lp_mode
        bi.omega_max, bi.acc_lat_max = omega_max, acc_lat_max
        bi.a_brake, bi.t_react = a_brake, t_react
        bi.collision_chords = collision_chords
        bi.nudge_positions = 1 if nudge else 0
        bi.param_mode = param_mode
        if obstacle_edges is not None and len(obstacle_edges):
            ed = np.ascontiguousarray(obstacle_edges, dtype=R).reshape(-1, 4)
            keep.append(ed)
            bi.obstacle_edges, bi.n_obstacle_edges = ed.ctypes.data, ed.shape[0]
        bi.opt = opt if opt is not None else default_opt()
        res = BatchResult()
        bo = abi.ScbBatchOut()
        ob = out_buffers or {}

        def outbuf(name, shape, dtype):
            a = ob.get(name)
            if a is None:
                a = np.empty(shape, dtype=dtype)
            assert a.shape == tuple(shape) and a.dtype == dtype and a.flags.c_contiguous, name
            res[name] = a
            return a.ctypes.data

        bo.status = outbuf("status", (P,), np.int32)
        bo.fail_stage = outbuf("fail_stage", (P,), np.int32)
        bo.S_cap = S_cap
        bo.record_f32 = 1 if record_f32 else 0
        rec = np.float32 if record_f32 else R
        names = list(outputs)
        if S_cap > 0:
            names += [n for n in PER_SAMPLE if n not in names]
            bo.n_samples = outbuf("n_samples", (P,), np.int32)
            if resample_at is not None:  # geometry only: positions (and velocities) are inputs
                pos, vel, ns = resample_at
                res["s_pos"] = np.ascontiguousarray(pos, dtype=rec)
                res["s_vel"] = np.ascontiguousarray(vel, dtype=rec) if vel is not None else None
                res["n_samples"][:] = ns
                ob = dict(ob, s_pos=res["s_pos"])
                if res["s_vel"] is not None:
                    ob["s_vel"] = res["s_vel"]
                else:
                    names.remove("s_vel")
                    del res["s_vel"]
        for n in names:
            setattr(bo, n, outbuf(n, self.out_shape(n, P, W, G, S_cap), rec if n in RECORD else R))
        if spline is not None:
            for n in ("ctrl_x", "ctrl_y", "arclen_table"):
                setattr(bo, n, arr(spline[n], self.out_shape(n, P, W, G, S_cap)))
        self._check(self.lib.scb_plan_batch(self.ctx, C.byref(bi), C.byref(bo)), "scb_plan_batch")
        return res

    # ------------------------------------------------------------------ rank, then fetch
    def _batch_in(self, keep, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G, opt=None, **kw):
        R = self.real

        def arr(a, shape):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=R)
            if a.shape != shape:
                a = np.ascontiguousarray(np.broadcast_to(a, shape))
            keep.append(a)
            return a.ctypes.data
        wx = np.ascontiguousarray(wx, dtype=R)
        P, W = wx.shape
        bi = abi.ScbBatchIn()
        bi.P, bi.W, bi.G = P, W, G
        bi.wx, bi.wy = arr(wx, (P, W)), arr(wy, (P, W))
        bi.vel_lo, bi.vel_hi, bi.acc_lo, bi.acc_hi = arr(vel_lo, (P,)), arr(vel_hi, (P,)), arr(acc_lo, (P,)), arr(acc_hi, (P,))
        for n in ("start_angle", "vel_start", "vel_end"):
            setattr(bi, n, arr(kw.pop(n, None), (P,)))
        for n in ("vel_lo_grid", "vel_hi_grid"):
            setattr(bi, n, arr(kw.pop(n, None), (G, P)))
        ed = kw.pop("obstacle_edges", None)
        if ed is not None and len(ed):
            ed = np.ascontiguousarray(ed, dtype=R).reshape(-1, 4)
            keep.append(ed)
            bi.obstacle_edges, bi.n_obstacle_edges = ed.ctypes.data, ed.shape[0]
        for n, v in kw.items():  # scalars: x_cap, dt, lp_mode, omega_max, acc_lat_max, a_brake, t_react, collision_chords, param_mode
            setattr(bi, n, v)
        bi.opt = opt if opt is not None else default_opt()
        return bi, P, W

    def plan_summary(self, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G=101, record_f32=True, opt=None, summaries=None, **kw):
        """scb_plan_batch_summary: host inputs in, the batch stays in HBM, only the 32-byte summaries come back
        (numpy structured array SUMMARY_DTYPE [P]; `summaries` = a preallocated, e.g. pinned, one)."""
        keep = []
        bi, P, W = self._batch_in(keep, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G, opt, **kw)
        out = summaries if summaries is not None else np.empty(P, dtype=SUMMARY_DTYPE)
        assert out.shape == (P,) and out.dtype == SUMMARY_DTYPE and out.flags.c_contiguous
        self._check(self.lib.scb_plan_batch_summary(self.ctx, C.byref(bi), 1 if record_f32 else 0, out.ctypes.data),
                    "scb_plan_batch_summary")
        self._kept = (P, W, G, bool(record_f32))
        return out

    def fetch(self, ids, outputs=TRAJECTORY_OUTPUTS, out_buffers=None):
        """scb_fetch: the records of problems `ids` of the last plan_summary, as arrays laid out like plan()'s with P := len(ids)."""
        P, W, G, rec32 = self._kept
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n = len(ids)
        res = BatchResult()
        bo = abi.ScbBatchOut()
        ob = out_buffers or {}
        for name in ("status", "fail_stage") + tuple(outputs):
            dt = np.int32 if name in ("status", "fail_stage") else (np.float32 if rec32 and name in RECORD else self.real)
            shape = (n,) if name in ("status", "fail_stage") else self.out_shape(name, n, W, G, 0)
            a = ob.get(name)
            if a is None:
                a = np.empty(shape, dtype=dt)
            assert a.shape == tuple(shape) and a.dtype == dt and a.flags.c_contiguous, name
            res[name] = a
            setattr(bo, name, a.ctypes.data)
        self._check(self.lib.scb_fetch(self.ctx, n, ids.ctypes.data, C.byref(bo)), "scb_fetch")
        return res

    # ------------------------------------------------------------------ HBM-resident buffers
    def make_device_batch(self, inputs, G, outputs=ALL_OUTPUTS, S_cap=0, stages=0, x_cap=0.0, dt=0.0, opt=None, lp_mode=0,
                          omega_max=0.0, acc_lat_max=0.0, param_mode=0, obstacle_edges=None, a_brake=0.0, t_react=0.0,
                          collision_chords=0, record_f32=False):
        """inputs: dict name -> torch CUDA tensor (dtype = ctx precision, contiguous).  Allocates
        the output tensors once and returns (batch_in, batch_out, result_tensors): the structs can be
        passed to run_device() any number of times without further allocation."""
        import torch
        tdt = torch.float64 if self.precision == abi.SCB_FP64 else torch.float32
        dev = torch.device("cuda", self.device)
        bi = abi.ScbBatchIn()
        wx = inputs.get("wx")
        if wx is not None:
            P, W = wx.shape
        else:
            P, W = inputs["path_len"].shape[0], 2
        bi.P, bi.W, bi.G = P, W, G
        for name in ("wx", "wy", "start_angle", "vel_lo", "vel_hi", "vel_lo_grid", "vel_hi_grid", "acc_lo", "acc_hi",
                     "vel_start", "vel_end", "path_len"):
            t = inputs.get(name)
            if t is not None:
                assert t.is_cuda and t.dtype == tdt and t.is_contiguous(), name
                setattr(bi, name, t.data_ptr())
        bi.stages, bi.x_cap, bi.dt, bi.lp_mode = stages, x_cap, dt, lp_mode
        bi.omega_max, bi.acc_lat_max = omega_max, acc_lat_max
        bi.param_mode = param_mode
        bi.a_brake, bi.t_react, bi.collision_chords = a_brake, t_react, collision_chords
        if obstacle_edges is not None:  # [E][4] CUDA tensor; the caller keeps it alive
            assert obstacle_edges.is_cuda and obstacle_edges.dtype == tdt and obstacle_edges.is_contiguous()
            bi.obstacle_edges, bi.n_obstacle_edges = obstacle_edges.data_ptr(), obstacle_edges.shape[0]
        bi.opt = opt if opt is not None else default_opt()
        res = {"status": torch.empty(P, dtype=torch.int32, device=dev),
               "fail_stage": torch.empty(P, dtype=torch.int32, device=dev)}
        bo = abi.ScbBatchOut()
        bo.status, bo.fail_stage, bo.S_cap = res["status"].data_ptr(), res["fail_stage"].data_ptr(), S_cap
        bo.record_f32 = 1 if record_f32 else 0
        names = list(outputs)
        if S_cap > 0:
            names += [n for n in PER_SAMPLE if n not in names]
            res["n_samples"] = torch.empty(P, dtype=torch.int32, device=dev)
            bo.n_samples = res["n_samples"].data_ptr()
        for n in names:
            res[n] = torch.empty(self.out_shape(n, P, W, G, S_cap), dtype=torch.float32 if record_f32 and n in RECORD else tdt, device=dev)
            setattr(bo, n, res[n].data_ptr())
        return bi, bo, res

    def run_device(self, bi, bo, sync=True):
        self._check(self.lib.scb_plan_batch_device(self.ctx, C.byref(bi), C.byref(bo), 0 if sync else 1),
                    "scb_plan_batch_device")

    def run_async(self, bi, bo, stream):
        """scb_plan_batch_async: enqueue on the caller's stream (a torch.cuda.Stream or a raw cudaStream_t), never synchronise."""
        raw = getattr(stream, "cuda_stream", stream)
        self._check(self.lib.scb_plan_batch_async(self.ctx, C.byref(bi), C.byref(bo), C.c_void_p(raw)), "scb_plan_batch_async")

    def pack_summary(self, P, res):
        """Device-side pack of the 32-byte per-problem summary records (what the multi-GPU path gathers).  The record
        buffer belongs to the result set (allocated once, inside torch's allocator but never returned to it while the
        batch lives), so the kernel that fills it on the ctx stream never races a reuse of the block by torch."""
        import torch
        dst = res.get("_summary")
        if dst is None or dst.numel() != P * C.sizeof(abi.ScbSummary):
            dst = res["_summary"] = torch.empty(P * C.sizeof(abi.ScbSummary), dtype=torch.uint8, device=res["status"].device)
        ptr = lambda n: res[n].data_ptr() if n in res else None
        self._check(self.lib.scb_pack_summary_device(self.ctx, P, ptr("status"), ptr("fail_stage"), ptr("T_total"),
                                                     ptr("arclen_total"), ptr("cost"), dst.data_ptr()),
                    "scb_pack_summary_device")
        return dst


SUMMARY_DTYPE = np.dtype([("status", np.int32), ("fail_stage", np.int32), ("T_total", np.float64),
                          ("arclen_total", np.float64), ("cost", np.float64)])


def shard_range(P, rank, world):
    """Contiguous block of problems owned by `rank`: [lo, hi).  Same rule as scb_shard_range."""
    per = (P + world - 1) // world
    lo = min(P, rank * per)
    return lo, min(P, lo + per)


class MultiPlanner:
    """scb_multi_*: one process, one scb_ctx + host thread per GPU, contiguous blocks of problems per device."""

    def __init__(self, n_devices, precision="fp64", spline_mode="lau_quintic", max_problems=65536, max_waypoints=32,
                 max_gridpoints=1001, devices=None):
        self.lib = abi.load_library()
        self.precision = {"fp64": abi.SCB_FP64, "fp32": abi.SCB_FP32}[precision]
        self.spline_mode = {"ref_cubic": abi.SCB_REF_CUBIC, "lau_quintic": abi.SCB_LAU_QUINTIC}[spline_mode]
        self.real = np.float64 if self.precision == abi.SCB_FP64 else np.float32
        self.degree = 3 if self.spline_mode == abi.SCB_REF_CUBIC else 5
        cfg = abi.ScbConfig(0, self.precision, self.spline_mode, max_problems, max_waypoints, max_gridpoints, 0)
        dv = None
        if devices is not None:
            dv = np.ascontiguousarray(devices, dtype=np.int32)
            n_devices = len(dv)
        h = C.c_void_p()
        rc = self.lib.scb_multi_create(C.byref(h), C.byref(cfg), n_devices, None if dv is None else dv.ctypes.data)
        if rc != 0:
            raise ScbError("scb_multi_create: %s" % self.lib.scb_strerror(rc).decode())
        self.m = h
        self.n_devices = n_devices

    def close(self):
        if getattr(self, "m", None):
            self.lib.scb_multi_destroy(self.m)
            self.m = None

    __del__ = close
    out_shape = BatchPlanner.out_shape
    _batch_in = BatchPlanner._batch_in

    def _check(self, rc, what):
        if rc != 0:
            raise ScbError("%s: %s" % (what, self.lib.scb_strerror(rc).decode()))

    def timing(self, r):
        t = abi.ScbTiming()
        self.lib.scb_multi_get_timing(self.m, r, C.byref(t))
        return {k: getattr(t, k) for k, _ in abi.ScbTiming._fields_}

    def plan(self, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G=101, outputs=TRAJECTORY_OUTPUTS, record_f32=False, opt=None,
             out_buffers=None, **kw):
        keep = []
        bi, P, W = self._batch_in(keep, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G, opt, **kw)
        res = BatchResult()
        bo = abi.ScbBatchOut()
        bo.record_f32 = 1 if record_f32 else 0
        ob = out_buffers or {}
        for name in ("status", "fail_stage") + tuple(outputs):
            dt = np.int32 if name in ("status", "fail_stage") else (np.float32 if record_f32 and name in RECORD else self.real)
            shape = (P,) if name in ("status", "fail_stage") else self.out_shape(name, P, W, G, 0)
            a = ob.get(name)
            if a is None:
                a = np.empty(shape, dtype=dt)
            assert a.shape == tuple(shape) and a.dtype == dt and a.flags.c_contiguous, name
            res[name] = a
            setattr(bo, name, a.ctypes.data)
        self._check(self.lib.scb_multi_plan_batch(self.m, C.byref(bi), C.byref(bo)), "scb_multi_plan_batch")
        return res

    def plan_summary(self, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G=101, record_f32=True, opt=None, summaries=None, **kw):
        keep = []
        bi, P, W = self._batch_in(keep, wx, wy, vel_lo, vel_hi, acc_lo, acc_hi, G, opt, **kw)
        out = summaries if summaries is not None else np.empty(P, dtype=SUMMARY_DTYPE)
        self._check(self.lib.scb_multi_plan_summary(self.m, C.byref(bi), 1 if record_f32 else 0, out.ctypes.data),
                    "scb_multi_plan_summary")
        self._kept = (P, W, G, bool(record_f32))
        return out

    def fetch(self, ids, outputs=TRAJECTORY_OUTPUTS):
        P, W, G, rec32 = self._kept
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        n = len(ids)
        res = BatchResult()
        bo = abi.ScbBatchOut()
        for name in ("status", "fail_stage") + tuple(outputs):
            dt = np.int32 if name in ("status", "fail_stage") else (np.float32 if rec32 and name in RECORD else self.real)
            shape = (n,) if name in ("status", "fail_stage") else self.out_shape(name, n, W, G, 0)
            res[name] = np.empty(shape, dtype=dt)
            setattr(bo, name, res[name].ctypes.data)
        self._check(self.lib.scb_multi_fetch(self.m, n, ids.ctypes.data, C.byref(bo)), "scb_multi_fetch")
        return res


def gather_summaries(local_bytes, world, group=None):
    """all-gather of per-rank summary-record buffers (uint8 torch tensors of equal length) — the one
    collective of the multi-GPU path.  NCCL when the tensors are CUDA, gloo on CPU."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local_bytes
    out = torch.empty(world * local_bytes.numel(), dtype=torch.uint8, device=local_bytes.device)
    dist.all_gather_into_tensor(out, local_bytes, group=group)
    return out


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_local_cpus(device=0):
    """CPUs of the NUMA node the GPU's PCIe root hangs off (sysfs; NVML when sysfs has no node), or None
    when the platform exposes no locality (single node, VM without topology)."""
    import torch
    pr = torch.cuda.get_device_properties(device)
    try:
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        if node >= 0:
            return _parse_cpulist(open("/sys/devices/system/node/node%d/cpulist" % node).read()) or None
    except (OSError, ValueError, AttributeError):
        pass
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(pr.uuid)).encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        return cpus if cpus and len(cpus) < os.cpu_count() else None
    except Exception:
        return None


class near_gpu:
    """Context manager: run the enclosed host code on the GPU's own NUMA node.  Host buffers pinned inside
    it are first-touched there, so the H2D / D2H DMA of scb_plan_batch never crosses the socket interconnect
    (which on a two-socket box costs about 3x in D2H bandwidth).  The previous affinity mask is restored on
    exit; pages stay where they were allocated.  A no-op when the platform shows no locality."""

    def __init__(self, device=0):
        self.device, self.prev, self.bound = device, None, None

    def __enter__(self):
        try:
            cpus = gpu_local_cpus(self.device)
            self.prev = os.sched_getaffinity(0)
            want = (cpus or set()) & self.prev
            if want and want != self.prev:
                os.sched_setaffinity(0, want)
                self.bound = want
        except (OSError, AttributeError):
            self.prev = None
        return self

    def __exit__(self, *exc):
        if self.bound and self.prev:
            os.sched_setaffinity(0, self.prev)
        return False
