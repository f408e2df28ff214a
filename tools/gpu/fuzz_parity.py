"""Randomised GPU-vs-oracle parity sweep (test infrastructure, like tests/): shapes, options and degenerate paths the
parametrised tests do not enumerate.  Every case goes through tests/tolerances.check (1e-9, identical verdicts).
usage: python tools/gpu/fuzz_parity.py [seconds] [seed]   -> gpurun_out/fuzz_<seed>.json (failures with their recipe)"""
import json
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import seacurrent_b200 as scb
import tolerances
from oracle import oracle as ora

import fuzz_cases


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 300.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    ora.build()
    rng = np.random.default_rng(seed)
    dry = os.environ.get("FUZZ_DRY") == "1"
    ext = os.environ.get("FUZZ_EXT") == "1"   # obstacles and the sample stream as well
    planners = None if dry else {m: scb.BatchPlanner(0, "fp64", m, max_problems=256, max_waypoints=80, max_gridpoints=512, max_samples=2048) for m in ("ref_cubic", "lau_quintic")}
    t0 = time.time()
    n = 0; fails = []; okp = 0; totp = 0
    while time.time() - t0 < budget:
        c = (fuzz_cases.make_ext_case if ext else fuzz_cases.make_case)(rng)
        n += 1
        try:
            a, b = fuzz_cases.run_case(c, None if planners is None else planners[c["mode"]], ora)
            okp += a; totp += b
        except Exception as e:  # noqa: BLE001 - the sweep records everything
            msg = "".join(traceback.format_exception_only(type(e), e))[:1500]
            fails.append({"case": c, "error": msg})
            print("FAIL", json.dumps(c), msg[:600], flush=True)
            if isinstance(e, scb.ScbError) and "CUDA" in msg:   # a sticky CUDA error: the contexts are gone
                break
    out = {"seed": seed, "cases": n, "failures": fails, "problems": totp, "feasible": okp, "seconds": time.time() - t0}
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"fuzz_{'ext_' if ext else ''}{seed}.json"), "w"), indent=1)
    print(f"cases {n}  failures {len(fails)}  problems {totp} feasible {okp}")


if __name__ == "__main__":
    main()
