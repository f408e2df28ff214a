#!/usr/bin/env python3
"""bench.py — smoothed+timed trajectories/s of the planning hot path on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config C3|C2|C4|C5]

A step = one pass of the whole path (Lau quintic spline + optimiser -> arc-length tables ->
toppra backward/forward -> parametrizer -> per-gridpoint trajectory record) over one batch of
synthetic planning problems.  N=1 workload = BASELINE.json configs[2] ("65,536 MPC candidate
rollouts, 1,000 gridpoints"), the batch the north_star's roofline target is quoted on; N>1 = the
same batch per rank (weak scaling), per-problem summary records all-gathered over NCCL.

value      whole-job trajectories/s, inputs and outputs resident in HBM (CUDA events on the ctx stream,
           max over ranks); batches are independent, so two are kept in flight on two contexts
           (config.in_flight; step i runs on context i % 2) -- every step is still one full pass over one batch
e2e        same metric through scb_plan_batch with pinned HOST buffers (H2D + kernels + D2H timed), one host
           thread per context
roofline   dominant kernel (by duration): algorithmic flops (or bytes) / its mean launch duration against the pipe
           that binds it -- the FP64/FP32 vector pipe (measured FMA peak, profiles/r1_fp_peak.json) or HBM
           (MEASURED_PEAKS.json); `kernels` lists every kernel of the step against both, `path` the whole step
cpu_baseline  the oracle's reference-faithful port of the E1 body timed on this box's host cores
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C3", choices=["C2", "C3", "C4", "C5"])
    ap.add_argument("--precision", default="fp64", choices=["fp64", "fp32"])
    ap.add_argument("--problems", type=int, default=0, help="override the per-rank batch size")
    ap.add_argument("--parametrizer", default="spline", choices=["spline", "const_accel"],
                    help="spline = what the reference's gen_vel_prof builds (the headline); const_accel = toppra's ConstAccel")
    ap.add_argument("--in-flight", type=int, default=0,
                    help="independent contexts (streams) the batches alternate over; 0 = 2 (1 for C4, whose outputs are 50 GB)")
    ap.add_argument("--obstacles", type=int, default=0,
                    help="E > 0: E wall pieces beside the nominal path, collision verdict (16 chords per segment) and the obstacle-"
                         "distance speed limit (a_brake 1.0, t_react 0.2) switched on (SURVEY 8f N2 / N3); not the headline")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--host-chunk", type=int, default=0,
                    help="e2e leg: problems per pipelined chunk of scb_plan_batch (0 = the library's default, -1 = one chunk)")
    ap.add_argument("--e2e-threads", type=int, default=0, help="e2e leg: host threads (= contexts) calling at once; 0 = in-flight")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    return ap.parse_args()


def workload(name, rank, world, problems=0):
    import seacurrent_b200 as scb
    P0, W, G = scb.workloads.CONFIGS[name]
    if name == "C4":  # fixed global batch, sharded (strong)
        P = problems or P0
        lo, hi = scb.shard_range(P, rank, world)
        w = scb.workloads.config_c4(P=P, lo=lo, hi=hi)
        return w, dict(workload="C4: %d randomized problems sharded over %d ranks" % (P, world), P_global=P, W=W, G=G), "strong"
    P = problems or P0
    w = scb.workloads.by_name(name, P=P)
    if name == "C3" and world > 1:  # each rank perturbs the nominal path with its own stream
        rng = np.random.Generator(np.random.PCG64([scb.workloads.SEED, 3, 1000 + rank]))
        w["wx"][:, 1:-1] += rng.normal(0, 0.05, size=(P, W - 2))
    desc = {"C2": "C2: batch of %d independent 20-waypoint paths, 500 gridpoints",
            "C3": "C3: %d MPC candidate rollouts (perturbed 10-waypoint path), 1,000 gridpoints, vel/accel limits",
            "C5": "C5: %d long-horizon 200-waypoint paths, 10,000 gridpoints"}[name] % P
    if world > 1:
        desc += " per rank x %d ranks" % world
    return w, dict(workload=desc, P_global=P * world, W=W, G=G), "weak"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop_flag, self.th = index, [], False, None

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                               "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

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


def host_cores():
    """Host threads this process may use (torchrun exports OMP_NUM_THREADS=1; the CPU arm must not inherit that)."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_baseline_run(w, G, seconds, nthreads):
    """The oracle's reference-faithful port of the E1 body (float32 spline side with the 10,001-point
    preview sampling, Bernstein-Fourier evaluation, FP64 toppra with the dense column-pivoted QR
    parametrizer; 'port'), on a bounded sample of the same workload."""
    from oracle import oracle as ora
    ora.build()
    cores = nthreads or host_cores()
    P = w["wx"].shape[0]
    n = min(P, cores)
    sel = slice(0, n)
    args = lambda s: (w["wx"][s], w["wy"][s], w["vel_lo"][s], w["vel_hi"][s], w["acc_lo"][s], w["acc_hi"][s])
    t, _ = ora.e1_batch_time(*args(sel), N=G - 1, mode=7, nthreads=cores)
    per_wave = max(t, 1e-6)
    waves = int(max(1, min(seconds / per_wave, P // n)))
    m = n * waves
    if waves > 1:
        t, _ = ora.e1_batch_time(*args(slice(0, m)), N=G - 1, mode=7, nthreads=cores)
    faithful = m / t
    # 'restated-fast': same port with the tridiagonal parametrizer and no preview sampling
    m2 = min(P, max(cores * 8, 64))
    t2, _ = ora.e1_batch_time(*args(slice(0, m2)), N=G - 1, mode=0, nthreads=cores)
    return {"value": faithful, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "first %d problems of the workload, oracle E1 port mode 7 (preview sampling + Bernstein-Fourier + "
                      "dense col-piv QR parametrizer, as the reference), %.1f s" % (m, t),
            "restated_fast": {"value": m2 / t2, "sample": "first %d problems, de Casteljau + Thomas parametrizer, no preview, %.2f s" % (m2, t2)}}


def reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w, cfg, scaling = workload(a.config, 0, 1, a.problems)
    G = cfg["G"]
    from oracle import oracle as ora
    ora.build()
    cores = host_cores()
    n = min(w["wx"].shape[0], cores)
    args = (w["wx"][:n], w["wy"][:n], w["vel_lo"][:n], w["vel_hi"][:n], w["acc_lo"][:n], w["acc_hi"][:n])
    for _ in range(a.warmup):
        ora.e1_batch_time(*args, N=G - 1, mode=7, nthreads=cores)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        ora.e1_batch_time(*args, N=G - 1, mode=7, nthreads=cores)
    dt = time.perf_counter() - t0
    v = n * a.steps / dt
    sample = "each step = the first %d problems of the workload (one per host thread), oracle E1 port mode 7" % n
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f32 spline / f64 toppra (reference arithmetic)", "data": "synthetic",
            "config": dict(cfg, sample=sample),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


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
    mode = "lau_quintic"
    # Batches are independent, so a planner keeps `nfl` of them in flight, each on its own context (= stream + scratch,
    # the C ABI's unit of concurrency: one scb_ctx per host thread): the tail of one batch's FP64-bound path / arc
    # kernels overlaps the HBM- and latency-bound toppra / param kernels of the other.  Step i runs on context i % nfl.
    nfl = a.in_flight if a.in_flight > 0 else (1 if a.config == "C4" else 2)
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
    nrec = P * 32
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
    ok_frac = float((res["status"] == 0).float().mean().item())
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
    value = cfg["P_global"] / (ms_per_step * 1e-3) if scaling == "weak" else cfg["P_global"] / (ms_per_step * 1e-3)

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
    d, K_it, Q = 5, opt.iters, 32
    # algorithmic bytes of each kernel = its mandatory inputs + outputs; algorithmic flops = SURVEY 8-D5's model,
    # F_par's 160 G split 100 G (arc-length inversion, curve, curvature: arc kernel) + 60 G (parametrizer)
    kern = {
        "path_ms": ("path_lane_kernel" if W <= 32 else "path_kernel", P * rs * (2 * W + 1 + 2 * (d + 1) * (W - 1) + 1), P * K_it * (W - 1) * Q * 120.0),
        "arc_ms": ("arc_kernel", P * rs * (2 * (d + 1) * (W - 1) + 1 + 3 * G), P * ((W - 1) * 3200 * 24.0 + 100.0 * G)),
        "toppra_ms": ("toppra_fast_kernel", P * (rs * (1 + 6 + 3 * G) + 8), P * 640.0 * G),
        "param_ms": ("param_ca_kernel" if pm else "param_kernel", P * rs * (2 * G + 3 + 4 * G + 1), P * 60.0 * G),
    }
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # FP-pipe peaks: FMA microbenchmark on this pool's B200 (tools/fp_peak.cu -> profiles/r1_fp_peak.json)
    fp_peak, fp_src = (37.07, "nominal-equivalent") if a.precision == "fp64" else (70.48, "nominal-equivalent")
    try:
        fpj = json.load(open(os.path.join(ROOT, "profiles", "r1_fp_peak.json")))
        fp_peak = float(fpj["fp64_peak_tflops" if a.precision == "fp64" else "fp32_peak_tflops"]); fp_src = "measured (tools/fp_peak.cu)"
    except Exception:
        pass
    traffic = {}
    try:  # dram__bytes_read+write per launch from the committed ncu --set full capture of this command (C3, fp64)
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json"))).get("%s_%s" % (a.config, a.precision), {})
    except Exception:
        pass
    kernels = []
    for k, (name, nbytes, nflops) in kern.items():
        ms_k = max(stage_ms[k], 1e-6)
        gbs, tfs = nbytes / (ms_k * 1e-3) / 1e9, nflops / (ms_k * 1e-3) / 1e12
        kernels.append({"kernel": name, "ms": stage_ms[k], "algorithmic_bytes": int(nbytes), "algorithmic_flops": nflops,
                        "hbm_gbs": gbs, "hbm_frac": gbs / hbm_peak, "fp_tflops": tfs, "fp_frac": tfs / fp_peak,
                        "bound": "fp_pipe" if tfs / fp_peak >= gbs / hbm_peak else "hbm", "traffic": traffic.get(name)})
    dom = max(kernels, key=lambda r: r["ms"])
    path_bytes = pl.algorithmic_bytes(P, W, G)
    path_flops = sum(r["algorithmic_flops"] for r in kernels)
    step_s = ms_per_step * 1e-3
    if dom["bound"] == "hbm":
        roof = {"bound": "hbm", "kernel": dom["kernel"], "achieved": dom["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s",
                "frac": dom["hbm_frac"], "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback"}
    else:  # no stage is a dense contraction: the binding pipe is the FP64 (FP32) vector pipe, not the tensor cores
        roof = {"bound": "fp_pipe", "kernel": dom["kernel"], "achieved": dom["fp_tflops"], "peak": fp_peak, "unit": "TFLOP/s",
                "frac": dom["fp_frac"], "peak_source": fp_src}
    roof.update({"traffic": dom["traffic"], "kernel_ms": dom["ms"], "algorithmic_bytes_per_launch": dom["algorithmic_bytes"],
                 "algorithmic_flops_per_launch": dom["algorithmic_flops"],
                 "hbm": {"achieved": dom["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": dom["hbm_frac"]},
                 "kernels": kernels,
                 "path": {"algorithmic_bytes_per_step": path_bytes, "hbm_gbs": path_bytes / step_s / 1e9,
                          "hbm_frac": path_bytes / step_s / 1e9 / hbm_peak, "algorithmic_flops_per_step": path_flops,
                          "fp_tflops": path_flops / step_s / 1e12, "fp_frac": path_flops / step_s / 1e12 / fp_peak}})
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "f64" if a.precision == "fp64" else "f32", "data": "synthetic",
            "config": dict(cfg, P_per_rank=P, spline="lau_quintic, optimiser K=%d move_waypoints=1" % opt.iters,
                           parametrizer=a.parametrizer, in_flight=nfl,
                           l2="outputs streamed per step (%.2f GB) exceed the 126 MB L2; no flush" % (path_bytes / 1e9),
                           ok_fraction=ok_frac,
                           **({"obstacles": "%d wall pieces, collision_chords 16, a_brake 1.0, t_react 0.2 (arc_ms includes the collision "
                                            "kernel, toppra_ms the obstacle-distance kernel)" % a.obstacles} if a.obstacles > 0 else {})),
            "stage_ms": stage_ms, "roofline": roof, "clocks": clocks, "gpu_launches": launches_per_step * a.steps}

    # ---- e2e: host buffers in pinned memory through scb_plan_batch
    if not a.no_e2e:
        names = scb.TRAJECTORY_OUTPUTS
        # host buffers are pinned (and the calls made) from the GPU's own NUMA node: with several ranks on a
        # two-socket box the scheduler otherwise puts some of them on the far socket and their D2H runs ~3x slower
        numa = scb.near_gpu(local) if os.environ.get("SCB_BENCH_NUMA", "1") != "0" else contextlib.nullcontext()
        with numa:
            host_in = {k: torch.from_numpy(np.ascontiguousarray(v, dtype=pl.real)).pin_memory().numpy() for k, v in w.items() if k != "G"}
            hbufs = []
            for _ in range(nfl):
                bufs = {"status": torch.empty(P, dtype=torch.int32).pin_memory().numpy(),
                        "fail_stage": torch.empty(P, dtype=torch.int32).pin_memory().numpy()}
                for n in names:
                    bufs[n] = torch.empty(pl.out_shape(n, P, W, G, 0), dtype=tdt).pin_memory().numpy()
                hbufs.append(bufs)
            h2d = sum(v.nbytes for v in host_in.values()); d2h = sum(v.nbytes for v in hbufs[0].values())
            nth = min(nfl, a.e2e_threads) if a.e2e_threads > 0 else nfl
            k2 = max(2, min(a.steps, 10))
            k2 += (-k2) % nth  # the same number of calls on every context
            for q in pls:
                q.set_host_chunk(a.host_chunk)

            def calls(c, n):  # one host thread per context, as the C ABI prescribes; ctypes drops the GIL inside the call
                for _ in range(n):
                    pls[c].plan(G=G, outputs=names, out_buffers=hbufs[c], opt=opt, param_mode=pm, **extra_host, **host_in)

            def run(n_each):
                th = [threading.Thread(target=calls, args=(c, n_each)) for c in range(nth)]
                for t_ in th:
                    t_.start()
                for t_ in th:
                    t_.join()

            run(2)
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            run(k2 // nth)
            dt = time.perf_counter() - t0
            if world > 1:
                t = torch.tensor([dt], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t.item())
        tm = pl.timing()
        line["e2e"] = {"value": cfg["P_global"] * k2 / dt if scaling == "weak" else cfg["P_global"] * k2 / dt, "unit": UNIT,
                       "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": k2,
                       "last_step_ms": {k: tm[k] for k in ("h2d_ms", "d2h_ms", "total_ms")},
                       "outputs": list(names), "host_threads": nth,
                       "host_chunk": a.host_chunk if a.host_chunk else "auto (8192 problems: kernels of chunk c+1 over the D2H of chunk c)",
                       "numa_bound_cpus": len(getattr(numa, "bound", None) or ())}
    if rank == 0 and world == 1 and not a.no_cpu:  # N=1 only (torchrun pins OMP_NUM_THREADS=1 per rank)
        line["cpu_baseline"] = cpu_baseline_run(w, G, a.cpu_seconds, 0)
    if rank == 0:
        print(json.dumps(line))
    for q in pls:
        q.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
