#!/usr/bin/env python3
"""Extract the reference's own golden vectors for the hot path into small JSON fixtures.

Run in the build container (needs /root/reference):  python tests/golden/gen_reference_fixtures.py
Writes
  tests/golden/toppra_cpp_golden.json   third-party/toppra/cpp/tests/test_algorithm.cpp:64-165
                                        (path coefficients + three 51-vectors) and
                                        test_solver.cpp:129-441 (solve_lp1d / solve_lp2d cases)
  tests/golden/output_json_fixture.json examples/output.json (arclength block in full,
                                        time/pos/vel/acc/pos_x/pos_y/ang_vel every 16th sample)
Nothing under tests/ reads /root/reference at run time.
"""
import json
import os
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
INF = float("inf")


def _nums(body):
    body = body.replace("-seidel::infinity", "-Infinity").replace("seidel::infinity", "Infinity")
    out = []
    for tok in body.strip().split(","):
        tok = tok.strip()
        if not tok:
            continue
        out.append(float(eval(tok, {"Infinity": INF})))
    return out


def _stream(src, name):
    m = re.search(name + r"\s*<<\s*(.*?);", src, re.S)
    return _nums(m.group(1))


def toppra_golden():
    base = os.path.join(REF, "third-party/toppra/cpp/tests")
    alg = open(os.path.join(base, "test_algorithm.cpp")).read()
    out = {"source": "third-party/toppra/cpp/tests/test_algorithm.cpp:64-165, test_solver.cpp:129-441"}
    out["algorithm"] = {
        "coeffs": [_stream(alg, "coeff%d" % i) for i in range(3)],  # each 4x2 row-major
        "breaks": [0, 1, 2, 3], "dof": 2, "N": 50,
        "vel": [-1.0, 1.0], "acc": [-0.2, 0.2], "interpolation": 0,
        "K_max": _stream(alg, "e_K_max"),
        "parametrization": _stream(alg, "expected_parametrization"),
        "feasible_max": _stream(alg, "expected_feasible_max"),
        "tol": 1e-6,
    }
    sol = open(os.path.join(base, "test_solver.cpp")).read()
    # ---- solve_lp1d known answers (test_solver.cpp:129-275), transcribed case by case
    ABS_TOL, TINY, BIG = 1e-13, 1e-10, 1e10
    eps = ABS_TOL / 2
    out["lp1d"] = [
        {"v": [1, 0], "A": [[0, -1], [1, -BIG * 10]], "feasible": True, "optvar": BIG * 10},
        {"v": [1, 0], "A": [[1, -2], [eps, -eps]], "feasible": True, "optvar": 2.0},
        {"v": [1, 0], "A": [[eps, -eps * eps]], "feasible": True, "optvar": "inf"},
        {"v": [1, 0], "A": [[1, -3], [-1, 4]], "feasible": False},
        {"v": [1, 0], "A": [[1, "inf"]], "feasible": False},
        {"v": [-1, 0], "A": [[1, -BIG * 10]], "feasible": True, "optvar": "-inf"},
        {"v": [-1, 0], "A": [[-1, 3]], "feasible": True, "optvar": 3.0},
        {"v": [0, 0], "A": [[-1, 3]], "feasible": True, "optvar": 3.0},
        {"v": [0, 0], "A": [[1, -1], [-1, -1]], "feasible": True, "optvar": 0.0},
        {"v": [TINY / 2, 0], "A": [[1, -1], [-1, -1]], "feasible": True, "optvar": 1.0},
    ]
    # ---- solve_lp2d instances (test_solver.cpp:277-441): parse each { ... check(true); } block
    seg = sol[sol.index("TEST(SeidelFunctions, seidel_2d)"):]
    cases = []
    for blk in re.finditer(r"A\.resize\((\d+), 3\);(.*?)check\((true|false)\);", seg, re.S):
        n = int(blk.group(1)); body = blk.group(2)
        mv = re.search(r"v\s*(?:<<|=)\s*\{?(.*?)\}?;", body, re.S)
        ma = re.search(r"\n\s*A\s*<<\s*(.*?);", body, re.S)
        ml = re.search(r"low\s*(?:<<|=)\s*\{?(.*?)\}?;", body, re.S)
        mh = re.search(r"high\s*(?:<<|=)\s*\{?(.*?)\}?;", body, re.S)
        A = _nums(ma.group(1))
        assert len(A) == 3 * n, (len(A), n)
        enc = lambda xs: ["inf" if x == INF else "-inf" if x == -INF else x for x in xs]
        cases.append({"v": _nums(mv.group(1)), "A": enc(A), "n": n, "low": enc(_nums(ml.group(1))),
                      "high": enc(_nums(mh.group(1))), "feasible": blk.group(3) == "true"})
    assert len(cases) == 5, len(cases)
    out["lp2d"] = cases
    json.dump(out, open(os.path.join(HERE, "toppra_cpp_golden.json"), "w"), indent=1)


def output_json():
    d = json.load(open(os.path.join(REF, "examples/output.json")))
    step = 16
    out = {
        "source": "examples/output.json (request of examples/zmq_test.py:7-10)",
        "request": {"acc": [-0.5, 0.5], "vel": [-0.25, 0.25], "waypoints": [[0, 0], [10, 0], [10, 10]]},
        "arclength": d["arclength"],
        "n_samples": len(d["time"]),
        "stride": step,
        "time_last": d["time"][-1],
    }
    for k in ("time", "pos_x", "pos_y", "ang_vel"):
        out[k] = d[k][::step]
    for k in ("pos", "vel", "acc"):
        out[k] = d[k][0][::step]
    out["n_pos_x"] = len(d["pos_x"])
    out["max_vel"] = max(d["vel"][0]); out["max_acc"] = max(d["acc"][0]); out["min_acc"] = min(d["acc"][0])
    json.dump(out, open(os.path.join(HERE, "output_json_fixture.json"), "w"))


if __name__ == "__main__":
    toppra_golden()
    output_json()
    print("fixtures written to", HERE)
