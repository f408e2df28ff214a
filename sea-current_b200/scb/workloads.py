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


def config_c1():
    """The two literal cases in the reference tree: examples/json_test.cpp:19-26 and
    examples/zmq_test.py:7-10 (the latter padded to 6 waypoints is NOT done: two separate batches)."""
    a = dict(wx=np.array([[0., 10., 10., 10., 30., 0.]]), wy=np.array([[0., 0., 10., 20., 50., 0.]]),
             vel_lo=np.array([0.0]), vel_hi=np.array([1.0]), acc_lo=np.array([-0.25]), acc_hi=np.array([0.25]), G=101)
    b = dict(wx=np.array([[0., 10., 10.]]), wy=np.array([[0., 0., 10.]]),
             vel_lo=np.array([-0.25]), vel_hi=np.array([0.25]), acc_lo=np.array([-0.5]), acc_hi=np.array([0.5]), G=101)
    return a, b


def config_c2(P=4096, W=20, G=500):
    rng = _rng(2)
    wx, wy = random_walk_paths(rng, P, W)
    return dict(wx=wx, wy=wy, G=G, **random_limits(rng, P))


def config_c3(P=65536, W=10, G=1000, sigma=0.25):
    """MPC rollouts: one nominal W-waypoint path, inner waypoints perturbed by N(0, sigma^2)."""
    rng = _rng(3)
    nx, ny = random_walk_paths(rng, 1, W)
    wx = np.repeat(nx, P, axis=0); wy = np.repeat(ny, P, axis=0)
    wx[:, 1:-1] += rng.normal(0.0, sigma, size=(P, W - 2))
    wy[:, 1:-1] += rng.normal(0.0, sigma, size=(P, W - 2))
    one = np.ones(P)
    return dict(wx=wx, wy=wy, G=G, vel_lo=-2.0 * one, vel_hi=2.0 * one, acc_lo=-1.5 * one, acc_hi=1.5 * one)


def config_c4(P=1 << 20, W=10, G=500, lo=0, hi=None):
    """C2's distributions; problems [lo, hi) of the global batch (shard of one rank).  Generated in
    blocks of 65,536 keyed by block index so any shard reproduces the same global problem set."""
    hi = P if hi is None else hi
    blk = 65536
    parts = []
    for b in range(lo // blk, (hi + blk - 1) // blk):
        rng = _rng(4, b)
        n = min(blk, P - b * blk)
        wx, wy = random_walk_paths(rng, n, W)
        d = dict(wx=wx, wy=wy, **random_limits(rng, n))
        s, e = max(lo, b * blk) - b * blk, min(hi, (b + 1) * blk) - b * blk
        parts.append({k: v[s:e] for k, v in d.items()})
    out = {k: np.concatenate([p[k] for p in parts], axis=0) for k in parts[0]}
    out["G"] = G
    return out


def config_c5(P=8192, W=200, G=10000):
    rng = _rng(5)
    wx, wy = random_walk_paths(rng, P, W)
    return dict(wx=wx, wy=wy, G=G, **random_limits(rng, P))


def by_name(name, **kw):
    return {"C2": config_c2, "C3": config_c3, "C4": config_c4, "C5": config_c5}[name](**kw)
