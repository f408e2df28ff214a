"""Host-side logic of the multi-GPU path on CPU: contiguous sharding, shard-reproducible workloads and the
one collective (all-gather of the 32-byte summary records) over gloo with world_size 2."""
import os
import socket
import subprocess
import sys

import numpy as np

import seacurrent_b200 as scb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_the_batch():
    for P in (0, 1, 7, 65536, (1 << 20) + 3):
        for world in (1, 2, 4, 8):
            spans = [scb.shard_range(P, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == P
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:])) and all(lo <= hi for lo, hi in spans)
            assert max(hi - lo for lo, hi in spans) <= (P + world - 1) // world


def test_c4_shards_reproduce_the_global_problem_set():
    P = 3 * 65536 + 100  # crosses generator blocks
    whole = scb.workloads.config_c4(P=P)
    for world in (2, 4):
        parts = [scb.workloads.config_c4(P=P, lo=lo, hi=hi) for lo, hi in (scb.shard_range(P, r, world) for r in range(world))]
        for k in ("wx", "wy", "vel_hi", "acc_hi", "acc_lo"):
            np.testing.assert_array_equal(np.concatenate([p[k] for p in parts], axis=0), whole[k])
        for k in ("vel_lo_grid", "vel_hi_grid"):  # [G][P]: shards are column blocks
            np.testing.assert_array_equal(np.concatenate([p[k] for p in parts], axis=1), whole[k])
            assert all(p[k].flags.c_contiguous for p in parts)


def test_summary_record_layout():
    assert scb.SUMMARY_DTYPE.itemsize == 32
    assert [scb.SUMMARY_DTYPE.fields[n][1] for n in ("status", "fail_stage", "T_total", "arclen_total", "cost")] == [0, 4, 8, 16, 24]


WORKER = r"""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %r)
import seacurrent_b200 as scb
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
P = 1001
lo, hi = scb.shard_range(P, rank, world)
per = (P + world - 1) // world
rec = np.zeros(per, dtype=scb.SUMMARY_DTYPE)            # equal-length buffers, the tail of the last shard is padding
idx = np.arange(lo, hi)
rec["status"][: hi - lo] = idx %% 4; rec["fail_stage"][: hi - lo] = -1
rec["T_total"][: hi - lo] = idx * 0.5; rec["arclen_total"][: hi - lo] = idx * 2.0; rec["cost"][: hi - lo] = idx + 0.25
rec["status"][hi - lo:] = -7
local = torch.from_numpy(rec.view(np.uint8).copy())
allrec = scb.gather_summaries(local, world)
got = np.frombuffer(allrec.numpy().tobytes(), dtype=scb.SUMMARY_DTYPE).reshape(world, per)
flat = np.concatenate([got[r][: scb.shard_range(P, r, world)[1] - scb.shard_range(P, r, world)[0]] for r in range(world)])
want = np.arange(P)
assert flat.shape[0] == P
assert np.array_equal(flat["status"], want %% 4) and np.array_equal(flat["T_total"], want * 0.5)
assert np.array_equal(flat["arclen_total"], want * 2.0) and np.array_equal(flat["cost"], want + 0.25)
best = int(np.argmin(np.where(flat["status"] == 0, flat["T_total"], np.inf)))   # what an MPC outer loop does
assert best == 0
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_gather_summaries_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=180)[0].decode() for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert all("ok" in o for o in outs)
