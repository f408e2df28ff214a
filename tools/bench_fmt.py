#!/usr/bin/env python3
"""Throughput of scb_fmt_batch (FMT* queries/s over one shared map) next to the CPU restatement on the host cores.
Map: examples/test.cpp:250-259 ([-1, 1]^2, three obstacles), n Halton samples (sample_free).  One JSON line.
  python tools/bench_fmt.py [--nodes 2560] [--rn 0.3] [--queries 1184] [--steps 5]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=2560)
    ap.add_argument("--rn", type=float, default=0.3)
    ap.add_argument("--queries", type=int, default=148 * 8)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--cpu-queries", type=int, default=64)
    a = ap.parse_args()
    import seacurrent_b200 as scb
    from oracle import oracle as ora  # the checker and the CPU baseline, never the thing measured
    from test_fmt import _map, _queries
    obs, nodes, edges = _map(ora, a.nodes)
    init, goal = _queries(np.random.default_rng(1), a.queries, obs, ora)
    pl = scb.BatchPlanner(0, "fp64", "ref_cubic", max_problems=64, max_waypoints=16, max_gridpoints=101)
    for _ in range(2):
        res = pl.fmt_batch(nodes, edges, init, goal, a.rn)
    kms, tms = [], []
    for _ in range(a.steps):
        res = pl.fmt_batch(nodes, edges, init, goal, a.rn)
        t = pl.timing(); kms.append(t["spline_ms"]); tms.append(t["total_ms"])
    nc = min(a.cpu_queries, a.queries)
    t0 = time.perf_counter()
    want = ora.fmt_batch(nodes, edges, init[:nc], goal[:nc], a.rn)
    cpu_s = time.perf_counter() - t0
    same = bool(np.array_equal(want["status"], res["status"][:nc]) and np.array_equal(want["cost"], res["cost"][:nc])
                and np.array_equal(want["closed"], res["closed"][:nc]))
    print(json.dumps({"metric": "FMT* queries/s", "nodes": int(len(nodes)), "edges": int(len(edges)), "rn": a.rn, "queries": a.queries,
                      "kernel_ms": float(np.mean(kms)), "call_ms": float(np.mean(tms)),
                      "queries_per_s_kernel": a.queries / (np.mean(kms) * 1e-3), "queries_per_s_call": a.queries / (np.mean(tms) * 1e-3),
                      "found_fraction": float((res["status"] == 0).mean()), "mean_expansions": float(res["closed"].mean()),
                      "cpu": {"queries_per_s": nc / cpu_s, "cores": ora.num_threads(), "sample": "first %d queries, oracle restatement (literal loop)" % nc},
                      "parity_on_cpu_sample": same}))
    pl.close()


if __name__ == "__main__":
    main()
