#!/usr/bin/env python3
"""bench.py — smoothed+timed trajectories/s of the planning hot path on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3|C2|C4|C5]

A step = one pass of the whole path (Lau quintic spline + optimiser -> arc-length tables ->
toppra backward/forward -> parametrizer -> per-gridpoint trajectory record) over one batch of
synthetic planning problems.  N=1 workload = BASELINE.json configs[2] ("65,536 MPC candidate
rollouts, 1,000 gridpoints"), the batch the north_star's roofline target is quoted on; N>1 = the
same batch per rank (weak scaling), per-problem summary records all-gathered over NCCL;
--config C4 = 2^20 problems sharded over the ranks (strong scaling, the config that names the gather).

value      whole-job trajectories/s, inputs and outputs resident in HBM (CUDA events on the ctx stream,
           max over ranks); batches are independent, so two (four for small batches) are kept in flight on as many
           contexts (run.in_flight; step i runs on context i % in_flight) -- every step is still one full pass over one batch
e2e        same metric through scb_plan_batch with pinned HOST buffers (H2D + kernels + D2H timed), one host
           thread per context; the trajectory record comes back as float32 (scb_batch_out.record_f32: what the
           reference hands its caller, sea_current.hpp:379-386, 1245-1262).  `e2e_fp64_records` = the same with FP64
           records (round 1's definition), `e2e_topk` = rank-then-fetch (scb_plan_batch_summary + scb_fetch of the 16
           fastest feasible rollouts: what an MPC outer loop reads), reported beside e2e, never instead of it
roofline   dominant kernel (by duration): `frac` = ALGORITHMIC flops (SURVEY 8-D5's frozen model) / its mean launch
           duration / the measured FMA peak, as the contract defines it; `frac_executed` = the flops the kernel
           actually EXECUTED (DFMA x 2 + DMUL + DADD thread instructions from the committed ncu capture of this command,
           profiles/ncu_ops.json) / the same duration / the same peak -- the honest figure; `bound` comes from the ncu
           record (FP64-pipe utilisation, DRAM throughput), not from the model
cpu_baseline  the oracle timed on this box's host cores: `value` = the SAME work as the GPU arm (quintic + optimiser +
           toppra + Thomas parametrizer), plus the reference-faithful cubic port (preview sampling + dense QR) and
           its restated-fast form; `same_work_cubic` puts a GPU ref_cubic number next to the restated-fast cubic
--impl reference  times that same-work oracle port on all host cores, same config dict
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "smoothed+timed trajectories/sec"
UNIT = "trajectories/s"
TOPK = 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3", choices=["C2", "C3", "C4", "C5"])
    ap.add_argument("--precision", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--spline", default="lau_quintic", choices=["lau_quintic", "ref_cubic"],
                    help="lau_quintic = the north star's construction (the headline); ref_cubic = the reference's own")
    ap.add_argument("--problems", type=int, default=0, help="override the per-rank batch size")
    ap.add_argument("--parametrizer", default="spline", choices=["spline", "const_accel"],
                    help="spline = what the reference's gen_vel_prof builds (the headline); const_accel = toppra's ConstAccel")
    ap.add_argument("--in-flight", type=int, default=0,
                    help="independent contexts (streams) the batches alternate over; 0 = 2 (1 for C4, whose outputs are 50 GB; "
                         "4 for batches of at most 9,472 problems)")
    ap.add_argument("--obstacles", type=int, default=0,
                    help="E > 0: E wall pieces beside the nominal path, collision verdict (16 chords per segment) and the obstacle-"
                         "distance speed limit (a_brake 1.0, t_react 0.2) switched on (SURVEY 8f N2 / N3); not the headline")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--host-chunk", type=int, default=0,
                    help="e2e leg: problems per pipelined chunk of scb_plan_batch (0 = the library's default, -1 = one chunk)")
    ap.add_argument("--e2e-threads", type=int, default=0, help="e2e leg: host threads (= contexts) calling at once; 0 = in-flight")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip e2e_fp64_records and the GPU ref_cubic same-work figure")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="target CPU time of every cpu_baseline sample")
    return ap.parse_args()


def workload(name, rank, world, problems=0):
    import seacurrent_b200 as scb
    P0, W, G = scb.workloads.CONFIGS[name]
    if name == "C4":  # fixed global batch, sharded (strong)
        P = problems or P0
        lo, hi = scb.shard_range(P, rank, world)
        w = scb.workloads.config_c4(P=P, lo=lo, hi=hi)
        return w, dict(workload="C4: %d randomized problems sharded over the ranks, 1 %% of them hard cases (scb.workloads.hard_cases)" % P,
                       P_global=P, W=W, G=G), "strong"
    P = problems or P0
    w = scb.workloads.by_name(name, P=P)
    if name == "C3" and world > 1:  # each rank perturbs the nominal path with its own stream
        rng = np.random.Generator(np.random.PCG64([scb.workloads.SEED, 3, 1000 + rank]))
        w["wx"][:, 1:-1] += rng.normal(0, 0.05, size=(P, W - 2))
    desc = {"C2": "C2: batch of %d independent 20-waypoint paths, 500 gridpoints, 1 %% of them hard cases (scb.workloads.hard_cases)",
            "C3": "C3: %d MPC candidate rollouts (perturbed 10-waypoint path), 1,000 gridpoints, vel/accel limits",
            "C5": "C5: %d long-horizon 200-waypoint paths, 10,000 gridpoints"}[name] % P
    if world > 1:
        desc += " per rank"
    return w, dict(workload=desc, P_global=P * world, W=W, G=G), "weak"


def config_dict(a, cfg, P, opt_iters):
    """The same dict for both arms (the driver compares them)."""
    return dict(cfg, P_per_rank=P,
                spline="lau_quintic, optimiser K=%d move_waypoints=1" % opt_iters if a.spline == "lau_quintic" else "ref_cubic",
                parametrizer=a.parametrizer, precision=a.precision,
                l2="inputs + outputs streamed per step exceed the 126 MB L2; no flush")


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.th = index, [], False, None

    def _run(self):
        if self._run_nvml():
            return
        while not self.stop_flag:
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                               "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def _run_nvml(self):
        """The same readings through NVML (what nvidia-smi prints), every 2 ms: a 140 ms timed region gets ~60 samples
        instead of the two a subprocess per sample manages.  False = NVML unavailable, fall back to nvidia-smi."""
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM); reasons_fn(h)
        except Exception:
            return False
        bits = ((0x8, 2), (0x40, 3), (0x20, 4), (0x4, 5))   # hw_slowdown, hw_thermal_slowdown, sw_thermal_slowdown, sw_power_cap
        while not self.stop_flag:
            try:
                r = reasons_fn(h)
                row = [str(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)), str(mx), "", "", "", ""]
                for bit, col in bits:
                    row[col] = "Active" if r & bit else "Not Active"
                self.rows.append(row)
            except Exception:
                pass
            time.sleep(0.002)
        return True

    def start(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def stop(self):
        self.stop_flag = True
        if self.th:
            self.th.join(timeout=10)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def path_kernel_name(W, mode, obstacles):
    """The kernel launch_path_lane picks (path_lane_kernel.cu): the pair mapping where it packs more paths onto the lanes."""
    if W > 512:
        return "path_kernel"
    if W > 32:
        return "path_lane_cta_kernel"
    pair = mode == "lau_quintic" and not obstacles and W >= 3 and 64 // (W - 1) > 2 * (32 // W)
    env = os.environ.get("SCB_PATH_PAIR")
    if env is not None:
        pair = mode == "lau_quintic" and not obstacles and W >= 3 and int(env) != 0
    return "path_pair_kernel" if pair else "path_lane_kernel"


def host_cores():
    """Host threads this process may use (torchrun exports OMP_NUM_THREADS=1; the CPU arm must not inherit that)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _sized(run, n0, P, seconds):
    """Time run(n) on a sample sized for about `seconds` of wall time: probe with n0 problems, then scale."""
    t = max(run(n0), 1e-6)
    n = int(min(P, max(n0, n0 * seconds / t)))
    if n > n0:
        t = max(run(n), 1e-6)
    return n, t


def cpu_same_work(w, G, opt, spline_mode, seconds, cores):
    """The oracle's plan_batch: the GPU arm's own work (same spline construction, optimiser budget, toppra, Thomas
    parametrizer, per-gridpoint record) in FP64 on `cores` OpenMP threads."""
    from oracle import oracle as ora
    ora.build()
    P = w["wx"].shape[0]
    keys = [k for k in w if k != "G"]

    def run(n):
        sub = {k: (w[k][:, :n] if k.endswith("_grid") else w[k][:n]) for k in keys}
        return ora.plan_batch(G=G, spline_mode=spline_mode, opt=opt, nthreads=cores, outputs=("T_total", "x", "t", "px", "py", "v", "a", "omega"),
                              **sub)[1]
    n, t = _sized(run, min(P, 4 * cores), P, seconds)
    return n / t, n, t


def cpu_baseline_run(w, G, opt, spline_mode, seconds, nthreads):
    """`value`: the same work as the GPU arm on the oracle ('port').  Beside it the reference-faithful port of the E1
    body (float32 cubic spline incl. the 10,001-point preview sampling, Bernstein-Fourier evaluation, FP64 toppra with
    the dense column-pivoted QR parametrizer: what the reference itself would run) and its restated-fast form."""
    from oracle import oracle as ora
    ora.build()
    cores = nthreads or host_cores()
    P = w["wx"].shape[0]
    v, n, t = cpu_same_work(w, G, opt, spline_mode, seconds, cores)
    args = lambda m: (w["wx"][:m], w["wy"][:m], w["vel_lo"][:m], w["vel_hi"][:m], w["acc_lo"][:m], w["acc_hi"][:m])
    e1 = lambda mode: (lambda m: ora.e1_batch_time(*args(m), N=G - 1, mode=mode, nthreads=cores)[0])
    nf, tf = _sized(e1(7), min(P, cores), P, seconds) if G <= 2001 else (0, 0.0)   # O(G^3) dense QR: unusable beyond
    nr, tr = _sized(e1(0), min(P, 8 * cores), P, max(2.0, seconds / 4))
    out = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": "first %d problems of the workload, oracle plan_batch (the GPU arm's own work: %s, toppra, Thomas parametrizer, FP64), %.1f s"
                     % (n, "quintic + optimiser" if spline_mode else "reference cubic", t),
           "restated_fast_cubic": {"value": nr / tr, "sample": "first %d problems, reference cubic (float32), de Casteljau + Thomas parametrizer, no preview, %.1f s" % (nr, tr)}}
    if nf:
        out["reference_faithful_cubic"] = {"value": nf / tf, "sample": "first %d problems, oracle E1 port mode 7 (preview sampling + Bernstein-Fourier + dense "
                                                                         "col-piv QR parametrizer, as the reference runs it), %.1f s" % (nf, tf)}
    return out


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import seacurrent_b200 as scb
    w, cfg, scaling = workload(a.config, 0, 1, a.problems)
    G = cfg["G"]
    P = w["wx"].shape[0]
    from oracle import oracle as ora
    ora.build()
    cores = host_cores()
    opt = scb.default_opt(move_waypoints=1)
    sm = 1 if a.spline == "lau_quintic" else 0
    keys = [k for k in w if k != "G"]
    # each step = a bounded sample of the workload, sized once for about 0.4 s of wall time
    probe = min(P, 2 * cores)
    sub = lambda n: {k: (w[k][:, :n] if k.endswith("_grid") else w[k][:n]) for k in keys}
    run = lambda n: ora.plan_batch(G=G, spline_mode=sm, opt=opt, nthreads=cores, outputs=("T_total", "x", "t", "px", "py", "v", "a", "omega"), **sub(n))[1]
    t = max(run(probe), 1e-6)
    n = int(min(P, max(probe, probe * 0.4 / t)))
    for _ in range(a.warmup):
        run(n)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        run(n)
    dt = time.perf_counter() - t0
    v = n * a.steps / dt
    sample = ("each step = the first %d problems of the workload on %d host threads: oracle plan_batch, the GPU arm's own work (%s + toppra + Thomas "
              "parametrizer + per-gridpoint record, FP64)" % (n, cores, "quintic + optimiser K=%d" % opt.iters if sm else "reference cubic"))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64" if a.precision == "fp64" else "f32", "data": "synthetic",
            "config": config_dict(a, cfg, P, opt.iters),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def load_json(*parts):
    try:
        return json.load(open(os.path.join(ROOT, *parts)))
    except Exception:
        return {}


def main():
    a = parse()
    if a.impl == "reference":
        return reference_arm(a)
    import torch
    import torch.distributed as dist
    import seacurrent_b200 as scb

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries the one JSON line only
        dist.init_process_group("nccl", device_id=dev)
    w, cfg, scaling = workload(a.config, rank, world, a.problems)
    P, W = w["wx"].shape
    G = w["G"]
    mode = a.spline
    # Batches are independent, so a planner keeps `nfl` of them in flight, each on its own context (= stream + scratch,
    # the C ABI's unit of concurrency: one scb_ctx per host thread): the tail of one batch's FP64-bound path / arc
    # kernels overlaps the HBM- and latency-bound toppra / param kernels of the other.  Step i runs on context i % nfl.
    # Small batches (C2, C5: the toppra / param kernels leave most of the machine idle) keep four in flight
    # (measured: C5 18.47 / 18.02 / 17.64 ms per batch with 2 / 3 / 4, C2 0.980 / 0.896 / 0.893).
    nfl = a.in_flight if a.in_flight > 0 else (1 if a.config == "C4" else (4 if P <= 9472 else 2))
    pls = [scb.BatchPlanner(local, a.precision, mode, max_problems=P, max_waypoints=W, max_gridpoints=G) for _ in range(nfl)]
    pl = pls[0]
    tdt = torch.float64 if a.precision == "fp64" else torch.float32
    opt = scb.default_opt(move_waypoints=1)
    inputs = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev, dtype=tdt) for k, v in w.items() if k != "G"}
    pm = 1 if a.parametrizer == "const_accel" else 0
    extra, extra_host = {}, {}
    if a.obstacles > 0:  # wall pieces scattered 1-3 m beside the first path's legs
        rng = np.random.default_rng(7)
        k = rng.integers(0, W - 1, a.obstacles)
        mid = np.stack([(w["wx"][0, k] + w["wx"][0, k + 1]) / 2, (w["wy"][0, k] + w["wy"][0, k + 1]) / 2], 1)
        leg = np.stack([w["wx"][0, k + 1] - w["wx"][0, k], w["wy"][0, k + 1] - w["wy"][0, k]], 1)
        nrm = np.stack([-leg[:, 1], leg[:, 0]], 1) / np.linalg.norm(leg, axis=1, keepdims=True)
        c = mid + nrm * (rng.uniform(1.0, 3.0, (a.obstacles, 1)) * rng.choice([-1.0, 1.0], (a.obstacles, 1)))
        d = leg / np.linalg.norm(leg, axis=1, keepdims=True) * rng.uniform(0.3, 1.0, (a.obstacles, 1))
        edges = np.ascontiguousarray(np.concatenate([c - d, c + d], 1))
        extra_host = dict(obstacle_edges=edges, a_brake=1.0, t_react=0.2, collision_chords=16)
        extra = dict(extra_host, obstacle_edges=torch.from_numpy(edges).to(dev, dtype=tdt))
    batches = [q.make_device_batch(inputs, G=G, opt=opt, param_mode=pm, **extra) for q in pls]
    bi, bo, res = batches[0]
    streams = [torch.cuda.ExternalStream(q.stream(), device=dev) for q in pls]
    stream = streams[0]
    nstep = [0]

    def step():
        c = nstep[0] % nfl
        nstep[0] += 1
        pls[c].run_device(batches[c][0], batches[c][1], sync=False)
        rec = pls[c].pack_summary(P, batches[c][2])
        if world > 1:
            with torch.cuda.stream(streams[c]):
                rec = scb.gather_summaries(rec, world)
        return rec

    for _ in range(max(a.warmup, 3) + (max(a.warmup, 3) % nfl)):
        rec = step()
    for q in pls:
        q.synchronize()
    torch.cuda.synchronize()
    status = res["status"].cpu().numpy()
    ok_frac = float((status == 0).mean())
    # ---- device-resident timing: K steps, one event pair on the ctx stream
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)  # every context is idle here (synchronize above), so stream 0's clock starts the region
    for _ in range(a.steps):
        rec = step()
    for sx in streams[1:]:  # the region ends when the last context has finished: stream 0 waits for the others
        ev = torch.cuda.Event()
        ev.record(sx)
        stream.wait_event(ev)
    e1.record(stream)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    launches_per_step = pl.timing()["launches"] + 1
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / a.steps
    value = cfg["P_global"] / (ms_per_step * 1e-3)

    # ---- per-kernel durations (CUDA events recorded by the ctx around every kernel, one step at a time)
    keys = ("path_ms", "arc_ms", "toppra_ms", "param_ms", "sample_ms")
    stage_ms = {k: [] for k in keys}
    for _ in range(min(a.steps, 5)):
        pl.run_device(bi, bo, sync=True)
        tm = pl.timing()
        for k in keys:
            stage_ms[k].append(tm[k])
    stage_ms = {k: float(np.mean(v)) for k, v in stage_ms.items()}
    rs = 8 if a.precision == "fp64" else 4
    d, K_it, Q = (5, opt.iters, 32) if mode == "lau_quintic" else (3, 0, 32)
    grid_in = 2 * G if "vel_lo_grid" in w else 0
    # algorithmic bytes of each kernel = its mandatory inputs + outputs; algorithmic flops = SURVEY 8-D5's model,
    # F_par's 160 G split 100 G (arc-length inversion, curve, curvature: arc kernel) + 60 G (parametrizer)
    kern = {
        "path_ms": (path_kernel_name(W, mode, a.obstacles), P * rs * (2 * W + 1 + 2 * (d + 1) * (W - 1) + 1), P * K_it * (W - 1) * Q * 120.0),
        "arc_ms": ("arc_kernel", P * rs * (2 * (d + 1) * (W - 1) + 1 + 3 * G), P * ((W - 1) * 3200 * (4.0 * (d - 1) + 8.0) + 100.0 * G)),
        "toppra_ms": ("toppra_fast_kernel", P * (rs * (1 + 6 + 3 * G + grid_in) + 8), P * 640.0 * G),
        "param_ms": ("param_ca_kernel" if pm else "param_kernel", P * rs * (2 * G + 3 + 4 * G + 1), P * 60.0 * G),
    }
    peaks = load_json("MEASURED_PEAKS.json")
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # FP-pipe peaks: FMA microbenchmark on this pool's B200 (tools/fp_peak.cu -> profiles/r1_fp_peak.json)
    fp64_peak, fp32_peak, fp_src = 37.07, 70.48, "nominal-equivalent"
    fpj = load_json("profiles", "r1_fp_peak.json")
    if fpj:
        fp64_peak, fp32_peak, fp_src = float(fpj["fp64_peak_tflops"]), float(fpj["fp32_peak_tflops"]), "measured (tools/fp_peak.cu)"
    fp_peak = fp64_peak if a.precision == "fp64" else fp32_peak
    # executed work and DRAM traffic per launch from the committed ncu capture of this command (tools/ncu_ops.py)
    tag = "%s_%s_%s" % (a.config, a.precision, mode) if not (a.obstacles or pm or a.problems) else "none"
    ops_all = load_json("profiles", "ncu_ops.json")
    ops = ops_all.get(tag, {}).get("kernels", {})

    def executed(name):
        for k, v in ops.items():
            if k.startswith(name + "<") or k == name:
                return v
        return None

    kernels = []
    for k, (name, nbytes, nflops) in kern.items():
        ms_k = max(stage_ms[k], 1e-6)
        gbs, tfs = nbytes / (ms_k * 1e-3) / 1e9, nflops / (ms_k * 1e-3) / 1e12
        row = {"kernel": name, "ms": stage_ms[k], "algorithmic_bytes": int(nbytes), "algorithmic_flops": nflops,
               "hbm_gbs": gbs, "hbm_frac": gbs / hbm_peak, "fp_tflops": tfs, "fp_frac": tfs / fp_peak}
        ex = executed(name)
        if ex:
            # fraction of the launch's duration the executed flops would take at the measured FMA peaks (FP32 mode keeps the
            # decision-taking stages in FP64: both pipes count, each against its own peak)
            exf = ex["fp64_flops"] + ex["fp32_flops"]
            frac_ex = (ex["fp64_flops"] / fp64_peak + ex["fp32_flops"] / fp32_peak) / 1e12 / (ms_k * 1e-3)
            traffic_gbs = ex["dram_bytes"] / (ms_k * 1e-3) / 1e9
            pipe = ex.get("fp64_pipe_pct", 0.0)
            bound = "fp_pipe" if pipe >= 60.0 else ("hbm" if traffic_gbs >= 0.6 * hbm_peak else "issue/latency")
            row.update({"executed_flops_per_launch": exf, "executed_fp64_flops": ex["fp64_flops"], "executed_fp32_flops": ex["fp32_flops"],
                        "frac_executed": frac_ex, "traffic": int(ex["dram_bytes"]),
                        "traffic_frac_of_hbm_peak": traffic_gbs / hbm_peak, "threads_per_inst": ex.get("threads_per_inst"),
                        "bound": bound, "bound_evidence": "ncu: FP64 pipe %.0f %% busy, issue slots %.0f %%, DRAM traffic %.2f GB per launch = %.0f %% of the "
                                                          "measured HBM peak at this duration" % (pipe, ex.get("issue_pct", 0.0), ex["dram_bytes"] / 1e9,
                                                                                                     100 * traffic_gbs / hbm_peak)})
        else:
            row.update({"traffic": None, "bound": "fp_pipe" if tfs / fp_peak >= gbs / hbm_peak else "hbm",
                        "bound_evidence": "model only: no ncu capture committed for this command"})
        kernels.append(row)
    dom = max(kernels, key=lambda r: r["ms"])
    path_bytes = pl.algorithmic_bytes(P, W, G)
    path_flops = sum(r["algorithmic_flops"] for r in kernels)
    have_ex = all("executed_flops_per_launch" in r for r in kernels)
    path_exec = sum(r["executed_flops_per_launch"] for r in kernels) if have_ex else None
    path_exec_time = sum(r["executed_fp64_flops"] / fp64_peak + r["executed_fp32_flops"] / fp32_peak for r in kernels) / 1e12 if have_ex else None
    path_traffic = sum(r["traffic"] for r in kernels) if all(r["traffic"] for r in kernels) else None
    step_s = ms_per_step * 1e-3
    if dom["bound"] == "hbm":
        roof = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": dom["hbm_frac"], "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback"}
    else:  # no stage is a dense contraction: the binding pipe is the FP64 (FP32) vector pipe, not the tensor cores
        roof = {"bound": dom["bound"], "kernel": dom["kernel"], "achieved": dom["fp_tflops"], "peak": fp_peak, "unit": "TFLOP/s",
                "frac": dom["fp_frac"], "peak_source": fp_src}
    roof.update({"frac_definition": "algorithmic flops of SURVEY 8-D5's frozen model / measured duration / measured FMA peak (the contract's "
                                    "definition); frac_executed uses the flops the kernel executed (ncu DFMA x 2 + DMUL + DADD)",
                 "frac_executed": dom.get("frac_executed"), "executed_flops_per_launch": dom.get("executed_flops_per_launch"),
                 "bound_evidence": dom["bound_evidence"],
                 "traffic": dom["traffic"], "kernel_ms": dom["ms"], "algorithmic_bytes_per_launch": dom["algorithmic_bytes"],
                 "algorithmic_flops_per_launch": dom["algorithmic_flops"],
                 "hbm": {"achieved": dom["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": dom["hbm_frac"]},
                 "kernels": kernels,
                 "path": {"algorithmic_bytes_per_step": path_bytes, "hbm_gbs": path_bytes / step_s / 1e9,
                          "hbm_frac": path_bytes / step_s / 1e9 / hbm_peak, "algorithmic_flops_per_step": path_flops,
                          "fp_tflops": path_flops / step_s / 1e12, "fp_frac": path_flops / step_s / 1e12 / fp_peak,
                          "executed_flops_per_step": path_exec,
                          "fp_frac_executed": (path_exec_time / step_s) if have_ex else None,
                          "traffic_per_step": path_traffic, "traffic_over_algorithmic": (path_traffic / path_bytes) if path_traffic else None},
                 "ncu_capture": ops_all.get(tag, {}).get("source")})
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f64" if a.precision == "fp64" else "f32", "data": "synthetic",
            "config": config_dict(a, cfg, P, opt.iters),
            "run": dict(in_flight=nfl, ok_fraction=ok_frac, failed=int((status != 0).sum()),
                        statuses={str(k): int(v) for k, v in zip(*np.unique(status, return_counts=True))},
                        **({"obstacles": "%d wall pieces, collision_chords 16, a_brake 1.0, t_react 0.2 (arc_ms includes the collision "
                                         "kernel, toppra_ms the obstacle-distance kernel)" % a.obstacles} if a.obstacles > 0 else {})),
            "stage_ms": stage_ms, "roofline": roof, "clocks": clocks, "gpu_launches": launches_per_step * a.steps}

    # ---- e2e: host buffers in pinned memory through the C ABI
    if not a.no_e2e:
        names = scb.TRAJECTORY_OUTPUTS
        # host buffers are pinned (and the calls made) from the GPU's own NUMA node: with several ranks on a
        # two-socket box the scheduler otherwise puts some of them on the far socket and their D2H runs ~3x slower
        numa = scb.near_gpu(local) if os.environ.get("SCB_BENCH_NUMA", "1") != "0" else contextlib.nullcontext()
        with numa:
            pin = lambda arr: torch.from_numpy(np.ascontiguousarray(arr)).pin_memory().numpy()
            host_in = {k: pin(np.ascontiguousarray(v, dtype=pl.real)) for k, v in w.items() if k != "G"}
            nth = min(nfl, a.e2e_threads) if a.e2e_threads > 0 else nfl
            for q in pls:
                q.set_host_chunk(a.host_chunk)

            def timed(fn, n_each):
                """fn(c) called n_each times on each of nth host threads (one per context, as the C ABI prescribes;
                ctypes drops the GIL inside the call); wall time of the slowest rank."""
                def calls(c, n):
                    for _ in range(n):
                        fn(c)

                def run(n):
                    th = [threading.Thread(target=calls, args=(c, n)) for c in range(nth)]
                    for t_ in th:
                        t_.start()
                    for t_ in th:
                        t_.join()
                run(1)
                if world > 1:
                    dist.barrier()
                t0 = time.perf_counter()
                run(n_each)
                dt = time.perf_counter() - t0
                if world > 1:
                    t = torch.tensor([dt], device=dev, dtype=torch.float64)
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                    dt = float(t.item())
                return dt

            def record_leg(record_f32, n_each):
                rdt = np.float32 if record_f32 else pl.real
                hb = []
                for _ in range(nth):
                    bufs = {"status": pin(np.empty(P, np.int32)), "fail_stage": pin(np.empty(P, np.int32))}
                    for n in names:
                        bufs[n] = pin(np.empty(pl.out_shape(n, P, W, G, 0), dtype=rdt if n in scb.RECORD else pl.real))
                    hb.append(bufs)
                dt = timed(lambda c: pls[c].plan(G=G, outputs=names, out_buffers=hb[c], opt=opt, param_mode=pm, record_f32=record_f32,
                                                 **extra_host, **host_in), n_each)
                tm = pl.timing()
                return {"value": cfg["P_global"] * n_each * nth / dt, "unit": UNIT, "h2d_bytes_per_step": sum(v.nbytes for v in host_in.values()),
                        "d2h_bytes_per_step": sum(v.nbytes for v in hb[0].values()), "steps": n_each * nth,
                        "record": "float32 (scb_batch_out.record_f32)" if record_f32 else "float64",
                        "last_step_ms": {k: tm[k] for k in ("h2d_ms", "d2h_ms", "total_ms")}, "outputs": list(names), "host_threads": nth,
                        "host_chunk": a.host_chunk if a.host_chunk else "auto (8192 problems: kernels of chunk c+1 over the D2H of chunk c)"}

            n_each = max(2, min(a.steps, 10) // nth)
            line["e2e"] = record_leg(True, n_each)
            if not a.no_extras:
                line["e2e_fp64_records"] = record_leg(False, max(2, n_each // 2))
            # rank, then fetch: summaries of all P rollouts, then the records of the TOPK fastest feasible ones
            summ = [torch.empty(P * scb.SUMMARY_DTYPE.itemsize, dtype=torch.uint8).pin_memory().numpy().view(scb.SUMMARY_DTYPE) for _ in range(nth)]
            fetched = [None] * nth

            def topk(c):
                s = pls[c].plan_summary(G=G, opt=opt, param_mode=pm, record_f32=True, summaries=summ[c], **extra_host, **host_in)
                T = np.where(s["status"] == 0, s["T_total"], np.inf)
                ids = np.argpartition(T, TOPK)[:TOPK].astype(np.int32)
                fetched[c] = pls[c].fetch(ids, outputs=names)
            dt = timed(topk, n_each)
            line["e2e_topk"] = {"value": cfg["P_global"] * n_each * nth / dt, "unit": UNIT, "k": TOPK,
                                "h2d_bytes_per_step": sum(v.nbytes for v in host_in.values()) + 4 * TOPK,
                                "d2h_bytes_per_step": summ[0].nbytes + sum(v.nbytes for v in fetched[0].values()), "steps": n_each * nth,
                                "what": "scb_plan_batch_summary (32-byte summaries of all rollouts) + host argpartition + scb_fetch of the %d fastest "
                                        "feasible rollouts' float32 records" % TOPK, "host_threads": nth}
    # ---- one same-work ratio with the REFERENCE's construction: GPU ref_cubic beside the CPU restated-fast cubic
    if world == 1 and not a.no_extras and mode == "lau_quintic" and not a.obstacles:
        for q in pls:
            q.close()
        pls = []
        pc = scb.BatchPlanner(local, a.precision, "ref_cubic", max_problems=P, max_waypoints=W, max_gridpoints=G)
        bc = pc.make_device_batch(inputs, G=G, param_mode=pm, outputs=scb.TRAJECTORY_OUTPUTS)
        sc = torch.cuda.ExternalStream(pc.stream(), device=dev)
        for _ in range(3):
            pc.run_device(bc[0], bc[1], sync=False)
        pc.synchronize()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(sc)
        nrep = max(3, min(a.steps, 10))
        for _ in range(nrep):
            pc.run_device(bc[0], bc[1], sync=False)
        c1.record(sc)
        pc.synchronize()
        line["same_work_cubic"] = {"gpu_value": P * nrep / (c0.elapsed_time(c1) * 1e-3), "unit": UNIT,
                                   "what": "the reference's own construction (cubic segments from the tangent heuristics, no optimiser) + arc length + "
                                           "toppra + Spline parametrizer + record, device-resident, one context"}
        pc.close()
    if rank == 0 and world == 1 and not a.no_cpu:  # N=1 only (torchrun pins OMP_NUM_THREADS=1 per rank)
        line["cpu_baseline"] = cb = cpu_baseline_run(w, G, opt, 1 if mode == "lau_quintic" else 0, a.cpu_seconds, 0)
        if "same_work_cubic" in line:
            line["same_work_cubic"]["cpu_restated_fast_value"] = cb["restated_fast_cubic"]["value"]
            line["same_work_cubic"]["ratio"] = line["same_work_cubic"]["gpu_value"] / cb["restated_fast_cubic"]["value"]
    if rank == 0:
        print(json.dumps(line))
    for q in pls:
        q.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
