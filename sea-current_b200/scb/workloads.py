"""Deterministic synthetic inputs for BASELINE.json's configs.

numpy only (PCG64 streams keyed by (seed, config id)); the same arrays feed the CUDA path, the
oracle and the CPU baseline.  Waypoint arrays are [P][W]; limits are [P].
"""
import numpy as np

SEED = 0x5EAC0DE

CONFIGS = {
    # name: (P, W, G)
    "C1": (2, 6, 101),
    "C2": (4096, 20, 500),
    "C3": (65536, 10, 1000),
    "C4": (1 << 20, 10, 500),
    "C5": (8192, 200, 10000),
}


def _rng(cfg_id, stream=0):
    return np.random.Generator(np.random.PCG64([SEED, cfg_id, stream]))


def random_walk_paths(rng, P, W, leg=(2.0, 10.0), turn_deg=100.0):
    """Random-walk waypoints: leg length U(leg), heading change U(-turn, turn) per waypoint."""
    legs = rng.uniform(leg[0], leg[1], size=(P, W - 1))
    dth = np.deg2rad(rng.uniform(-turn_deg, turn_deg, size=(P, W - 1)))
    dth[:, 0] = rng.uniform(-np.pi, np.pi, size=P)
    th = np.cumsum(dth, axis=1)
    wx = np.zeros((P, W)); wy = np.zeros((P, W))
    wx[:, 1:] = np.cumsum(legs * np.cos(th), axis=1)
    wy[:, 1:] = np.cumsum(legs * np.sin(th), axis=1)
    return wx, wy


def random_limits(rng, P, slow_fraction=0.01):
    vel_hi = rng.uniform(0.5, 3.0, size=P)
    acc_hi = rng.uniform(0.25, 4.0, size=P)
    slow = rng.random(P) < slow_fraction  # knife-edge / very slow problems
    acc_hi[slow] = rng.uniform(1e-4, 1e-3, size=int(slow.sum()))
    return dict(vel_lo=-vel_hi, vel_hi=vel_hi, acc_lo=-acc_hi, acc_hi=acc_hi)


def hard_cases(rng, G, lim, hard_fraction=0.04):
    """SURVEY 8-D2, second half: a fixed fraction of the problems gets position-dependent velocity limits (the
    reference's vel_lim_func of the path position, sc:1175-1188) or an acceleration box that cannot be held, so that
    a batch at scale contains knife-edge AND failing problems.  Returns per-gridpoint limits [G][P] (problem index
    fastest, the C ABI's layout) and `kind` [P]:
      0  ordinary problem: the per-problem box at every gridpoint
      1  the upper velocity limit drops 10x over the middle of the path (s in [0.4, 0.6]) and the lower one rises 10x:
         a corner in the controllable sets; feasible
      2  minimum-speed zone: lower limit = 0.5 * upper on s in [s0, 1 - s0], s0 ~ U(0.02, 0.2): ERR_FAIL_CONTROLLABLE
         where the acceleration limit cannot reach that speed from rest (or shed it before the final rest), at a stage
         that depends on the problem; feasible otherwise
      3  acceleration box that excludes 0 (acc_lo = +0.1 acc_hi: cannot brake to the final rest): ERR_FAIL_CONTROLLABLE
         at the last stage
      4  minimum-speed zone over the first half only (s in [s0, 0.5]): braking has room, accelerating may not:
         ERR_FAIL_FORWARD_PASS at stage 0 (the controllable set of gridpoint 0 excludes rest), feasible otherwise
    lim is modified in place for kind 3."""
    P = lim["vel_hi"].shape[0]
    kind = np.zeros(P, dtype=np.int32)
    hard = rng.random(P) < hard_fraction
    kind[hard] = rng.integers(1, 5, size=int(hard.sum()))
    s = np.linspace(0.0, 1.0, G)[:, None]
    hi = np.ascontiguousarray(np.broadcast_to(lim["vel_hi"][None, :], (G, P)))
    lo = np.ascontiguousarray(np.broadcast_to(lim["vel_lo"][None, :], (G, P)))
    k1 = np.flatnonzero(kind == 1)
    mid = (s >= 0.4) & (s <= 0.6)
    hi[:, k1] = np.where(mid, 0.1 * lim["vel_hi"][None, k1], lim["vel_hi"][None, k1])
    lo[:, k1] = np.where(mid, 0.1 * lim["vel_lo"][None, k1], lim["vel_lo"][None, k1])
    k2 = np.flatnonzero(kind == 2)
    s0 = rng.uniform(0.02, 0.2, size=k2.size)[None, :]
    zone = (s >= s0) & (s <= 1.0 - s0)
    lo[:, k2] = np.where(zone, 0.5 * lim["vel_hi"][None, k2], lim["vel_lo"][None, k2])
    k4 = np.flatnonzero(kind == 4)
    s0 = rng.uniform(0.02, 0.2, size=k4.size)[None, :]
    zone = (s >= s0) & (s <= 0.5)
    lo[:, k4] = np.where(zone, 0.5 * lim["vel_hi"][None, k4], lim["vel_lo"][None, k4])
    k3 = kind == 3
    lim["acc_lo"] = np.where(k3, 0.1 * lim["acc_hi"], lim["acc_lo"])
    return dict(vel_lo_grid=lo, vel_hi_grid=hi), kind


def config_c1():
    """The two literal cases in the reference tree: examples/json_test.cpp:19-26 and
    examples/zmq_test.py:7-10 (the latter padded to 6 waypoints is NOT done: two separate batches)."""
    a = dict(wx=np.array([[0., 10., 10., 10., 30., 0.]]), wy=np.array([[0., 0., 10., 20., 50., 0.]]),
             vel_lo=np.array([0.0]), vel_hi=np.array([1.0]), acc_lo=np.array([-0.25]), acc_hi=np.array([0.25]), G=101)
    b = dict(wx=np.array([[0., 10., 10.]]), wy=np.array([[0., 0., 10.]]),
             vel_lo=np.array([-0.25]), vel_hi=np.array([0.25]), acc_lo=np.array([-0.5]), acc_hi=np.array([0.5]), G=101)
    return a, b


def config_c2(P=4096, W=20, G=500, hard_fraction=0.04):
    """hard_fraction > 0 adds the per-gridpoint limits of hard_cases() (about half of them fail)."""
    rng = _rng(2)
    wx, wy = random_walk_paths(rng, P, W)
    lim = random_limits(rng, P)
    out = dict(wx=wx, wy=wy, G=G)
    if hard_fraction > 0:
        grids, _ = hard_cases(_rng(2, 1), G, lim, hard_fraction)
        out.update(grids)
    out.update(lim)
    return out


def config_c3(P=65536, W=10, G=1000, sigma=0.25):
    """MPC rollouts: one nominal W-waypoint path, inner waypoints perturbed by N(0, sigma^2)."""
    rng = _rng(3)
    nx, ny = random_walk_paths(rng, 1, W)
    wx = np.repeat(nx, P, axis=0); wy = np.repeat(ny, P, axis=0)
    wx[:, 1:-1] += rng.normal(0.0, sigma, size=(P, W - 2))
    wy[:, 1:-1] += rng.normal(0.0, sigma, size=(P, W - 2))
    one = np.ones(P)
    return dict(wx=wx, wy=wy, G=G, vel_lo=-2.0 * one, vel_hi=2.0 * one, acc_lo=-1.5 * one, acc_hi=1.5 * one)


def config_c4(P=1 << 20, W=10, G=500, lo=0, hi=None, hard_fraction=0.04):
    """C2's distributions; problems [lo, hi) of the global batch (shard of one rank).  Generated in
    blocks of 65,536 keyed by block index so any shard reproduces the same global problem set."""
    hi = P if hi is None else hi
    blk = 65536
    parts = []
    for b in range(lo // blk, (hi + blk - 1) // blk):
        rng = _rng(4, b)
        n = min(blk, P - b * blk)
        wx, wy = random_walk_paths(rng, n, W)
        lim = random_limits(rng, n)
        d = dict(wx=wx, wy=wy)
        if hard_fraction > 0:
            grids, _ = hard_cases(_rng(4, 1 << 20 | b), G, lim, hard_fraction)
            d.update(grids)
        d.update(lim)
        s, e = max(lo, b * blk) - b * blk, min(hi, (b + 1) * blk) - b * blk
        parts.append({k: (v[:, s:e] if k.endswith("_grid") else v[s:e]) for k, v in d.items()})
    out = {k: np.ascontiguousarray(np.concatenate([p[k] for p in parts], axis=1 if k.endswith("_grid") else 0)) for k in parts[0]}
    out["G"] = G
    return out


def config_c5(P=8192, W=200, G=10000):
    rng = _rng(5)
    wx, wy = random_walk_paths(rng, P, W)
    return dict(wx=wx, wy=wy, G=G, **random_limits(rng, P))


def by_name(name, **kw):
    return {"C2": config_c2, "C3": config_c3, "C4": config_c4, "C5": config_c5}[name](**kw)
