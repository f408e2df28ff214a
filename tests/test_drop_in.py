"""The C++ drop-in (include/sea_current_b200.hpp): the reference's call sequence compiled against
the shim header, linked to libseacurrent_b200.so.

CPU: the header compiles as C++17 and C++20, the JSON writer prints what nlohmann's dump(4) prints
(= Python's json.dumps(indent=4, sort_keys=True)), and the program fails loudly without a GPU.
GPU: what it produces equals the oracle / the reference's fixture."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "sea-current_b200", "csrc")
LIB = os.path.join(CSRC, "libseacurrent_b200.so")


def _cxx():
    return "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _build(tmp, src, std="c++20", link=True):
    exe = os.path.join(str(tmp), os.path.basename(src).replace(".cpp", "_" + std.replace("+", "p")))
    cmd = [_cxx(), "-std=" + std, "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", src),
           "-o", exe]
    if link:
        cmd += ["-L" + CSRC, "-lseacurrent_b200", "-Wl,-rpath," + CSRC]
    subprocess.check_call(cmd)
    return exe


@pytest.fixture(scope="module")
def built(tmp_path_factory):
    if not os.path.exists(LIB):
        pytest.fail("libseacurrent_b200.so is not built; run __graft_entry__.build()")
    return _build(tmp_path_factory.mktemp("dropin"), "drop_in_e1.cpp")


def test_header_compiles_as_cxx17_and_cxx20(tmp_path):
    _build(tmp_path, "drop_in_e1.cpp", "c++17")
    _build(tmp_path, "drop_in_e1.cpp", "c++20")


def test_json_writer_matches_nlohmann_layout(tmp_path):
    exe = _build(tmp_path, "json_writer.cpp")
    pretty, compact, empty = subprocess.check_output([exe]).decode().split("====\n")
    doc = json.loads(pretty)
    assert len(doc) == 3 and doc == json.loads(compact)
    # nlohmann::json::dump(4) == Python's indent=4 with sorted keys (objects are std::map) and shortest round-trip floats
    assert pretty.rstrip("\n") == json.dumps(doc, indent=4, sort_keys=True)
    assert compact.rstrip("\n") == json.dumps(doc, separators=(",", ":"), sort_keys=True)
    assert set(doc[1]) == {"time", "velocity", "acceleration", "angularVelocity", "pose", "holonomicRotation", "holonomicAngularVelocity"}
    assert doc[1]["pose"]["translation"] == {"x": 10.5, "y": -3.0} and doc[1]["time"] == float(np.float32(0.02))
    assert empty.strip() == "[]"


def test_fails_loudly_without_a_device(built, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([built, "json", str(tmp_path / "o.json")], capture_output=True)
    assert r.returncode == 3 and b"no CPU fallback" in r.stderr


CORNER_LINES = [[8, -2.5, 8, -0.5], [8, -0.5, 9, -0.5], [9, -0.5, 9, -2.5], [9, -2.5, 8, -2.5],  # 4 vertices: closed by the ctor
                [11, 2, 13, 2], [13, 2, 13, 3],  # 3 vertices: open polyline
                [20, 20, 21, 21], [21, 21, 21, 20]]  # explicit edge list


def test_obstacle_constructors_build_the_reference_edge_lists(built):
    """sea_current.hpp:251-281: an even vertex count closes the edge list (the first vertex is appended to the
    constructor's local copy, `vertices` keeps its size); an odd one stays an open polyline; runs without a GPU."""
    doc = json.loads(subprocess.check_output([built, "lines", "-"]).decode())
    assert doc["lines"] == CORNER_LINES and doc["n_vertices"] == [4, 3, 3]


@pytest.mark.gpu
def test_from_path_clips_tangents_at_obstacles(built, oracle):
    """from_path over a planning_space with obstacles (shrink_tangent, sc:575-596) vs the oracle on the same edges."""
    doc = json.loads(subprocess.check_output([built, "obst", "-"]).decode())
    assert doc["lines"] == CORNER_LINES
    wx = np.array([[0.0, 10.0, 10.0, 0.0]]); wy = np.array([[0.0, 0.0, 10.0, 10.0]])
    lim = dict(vel_lo=[-1.0], vel_hi=[1.0], acc_lo=[-0.5], acc_hi=[0.5])
    ref, _ = oracle.plan_batch(wx=wx, wy=wy, G=21, spline_mode=0, obstacle_edges=np.array(CORNER_LINES, dtype=np.float64), **lim)
    free, _ = oracle.plan_batch(wx=wx, wy=wy, G=21, spline_mode=0, **lim)
    assert np.abs(ref["ctrl_x"][0] - free["ctrl_x"][0]).max() > 0.5  # the obstacles do cut a tangent back
    ctrl = np.array(doc["ctrl"])
    np.testing.assert_allclose(ctrl[:, 0], ref["ctrl_x"][0], atol=2e-6)  # the drop-in holds control points as float32
    np.testing.assert_allclose(ctrl[:, 1], ref["ctrl_y"][0], atol=2e-6)
    assert abs(doc["arclength"] / ref["arclen_total"][0] - 1) <= 1e-6


@pytest.mark.gpu
def test_json_case_matches_oracle(built, tmp_path, oracle):
    """examples/json_test.cpp's body through the shim vs the oracle's FP64 plan of the same problem."""
    import seacurrent_b200 as scb
    out = tmp_path / "o.json"
    head = json.loads(subprocess.check_output([built, "json", str(out)]).decode())
    doc = json.load(open(out))
    a, _ = scb.workloads.config_c1()
    ref, _ = oracle.plan_batch(spline_mode=0, S_cap=8192, **a)
    n = int(ref["n_samples"][0])
    assert head["n_samples"] == n == len(doc) and head["n_segments"] == 5 and head["degree"] == 3
    assert abs(head["arclength"] - ref["arclen_total"][0]) <= 1e-6 * ref["arclen_total"][0]
    assert abs(head["seg0_1"] - ref["arclen_table"][0, 1]) <= 1e-6
    assert abs(head["ctrl_1_x"] - ref["ctrl_x"][0, 1]) <= 1e-5 and abs(head["ctrl_1_y"] - ref["ctrl_y"][0, 1]) <= 1e-5
    col = lambda *k: np.array([_get(d, k) for d in doc])
    f32 = lambda x: x.astype(np.float32).astype(np.float64)
    np.testing.assert_allclose(col("time"), f32(ref["s_time"][:n, 0]), rtol=1e-6)
    # the drop-in carries positions as float32 (VectorXf, as the reference): arc positions are good to eps32 * L
    L = ref["arclen_total"][0]
    np.testing.assert_allclose(col("velocity"), ref["s_vel"][:n, 0], atol=1e-6)
    np.testing.assert_allclose(col("acceleration"), ref["s_acc"][:n, 0], atol=2e-6)
    np.testing.assert_allclose(col("pose", "translation", "x"), ref["s_px"][:n, 0], atol=2e-6 * L)
    np.testing.assert_allclose(col("pose", "translation", "y"), ref["s_py"][:n, 0], atol=2e-6 * L)
    np.testing.assert_allclose(col("angularVelocity"), ref["s_omega"][:n, 0], atol=2e-3 * np.abs(ref["s_omega"][:n, 0]).max())
    assert all(d["holonomicRotation"] == 0.0 and d["holonomicAngularVelocity"] == 0.0 for d in doc)


def _get(d, keys):
    for k in keys:
        d = d[k]
    return d


@pytest.mark.gpu
def test_text_request_matches_reference_fixture(built, tmp_path, golden_dir):
    """The zmq request of examples/zmq_test.py:7-10 -> JSON, against examples/output.json (the
    reference's own float32 output, older dict-of-arrays schema)."""
    out = tmp_path / "o.json"
    subprocess.check_call([built, "text", str(out)])
    doc = json.load(open(out))
    fix = json.load(open(os.path.join(golden_dir, "output_json_fixture.json")))
    n, st = fix["n_samples"], fix["stride"]
    assert len(doc) == n
    t = np.array([d["time"] for d in doc]); v = np.array([d["velocity"] for d in doc]); a = np.array([d["acceleration"] for d in doc])
    x = np.array([d["pose"]["translation"]["x"] for d in doc]); y = np.array([d["pose"]["translation"]["y"] for d in doc])
    np.testing.assert_allclose(t[::st], fix["time"], rtol=1e-5, atol=1e-5)
    np.testing.assert_allclose(v[::st], fix["vel"], atol=2e-5)
    np.testing.assert_allclose(a[::st], fix["acc"], atol=2e-4)
    m = len(fix["pos_x"])
    np.testing.assert_allclose(x[::st][:m], fix["pos_x"], atol=5e-3)
    np.testing.assert_allclose(y[::st][:m], fix["pos_y"], atol=5e-3)


def _example_map(oracle):
    """examples/test.cpp:250-259"""
    return [oracle.make_obstacle([(-0.5, 0), (1, 0), (1, 1), (0, 1)]), oracle.make_obstacle([(0, -0.5), (1, 0), (1, 1), (0, 1)]),
            oracle.make_obstacle([(np.float32(-0.6), np.float32(0.148)), (-1, np.float32(0.148)), (-1, 0), (np.float32(-0.6), 0)])]


def test_sample_free_matches_oracle(built, oracle):
    """planning_space::sample_free / halton / obstacle::contains / cost / near of the drop-in header (host code,
    sea_current.hpp:100-132, 200-250, 1294-1337) against the oracle: the same points in the same order, the Halton
    states after two draws; runs without a GPU."""
    doc = json.loads(subprocess.check_output([built, "sample", "200"]).decode())
    obs = _example_map(oracle)
    st = [0, 0, 0, 0]
    want = oracle.sample_free(200, (-1, 1, -1, 1), obs, state=st)
    np.testing.assert_array_equal(np.float32(doc["pts"]), want)
    oracle.sample_free(200, (-1, 1, -1, 1), obs, state=st)
    assert doc["state"] == st and doc["second_set_size"] == 200
    assert doc["contains"] == [1, 0, 0, 1]
    assert doc["cost"][0] == pytest.approx(0.5, rel=1e-6) and doc["cost"][1] == 1
    pts = np.float32(doc["pts"])
    d = np.sqrt(pts[:, 0].astype(np.float64) ** 2 + pts[:, 1].astype(np.float64) ** 2).astype(np.float32)
    assert doc["near"] == int(((d <= 0.25) & (d > 0)).sum())  # near(): pt_dist <= dist^2, the point itself excluded


@pytest.mark.gpu
def test_gen_vel_prof_three_dof(built, oracle):
    """gen_vel_prof<3> through the reference's signature (sc:1191-1265): the float record of every degree of freedom
    against the oracle's composition of the same steps."""
    doc = json.loads(subprocess.check_output([built, "ndof", "-"]).decode())
    G = 101
    s = np.linspace(0.0, 1.0, G)
    f = np.where((s > 0.405) & (s < 0.595), 0.25, 1.0)
    hi = f[:, None] * (1.0 + 0.5 * np.arange(3))[None, :]
    r = oracle.gen_vel_prof_ndof([0.5, 1.0, 2.5], [4.0, -1.5, 2.0], [0.0, -0.1, 0.0], [0.2, 0.0, 0.0], None, None, [-0.5, -0.4, -0.3],
                                 [0.5, 0.6, 0.3], G=G, dt=float(np.float32(0.05)), vel_lo_grid=-hi, vel_hi_grid=hi)
    assert r["status"] == 0
    assert doc["dof"] == 3 and doc["n"] == len(r["time"])
    assert abs(doc["T"] - r["time"][-1]) <= 1e-8 * r["time"][-1]
    for name, key in (("pos", "pos"), ("vel", "vel"), ("acc", "acc")):
        got = np.array(doc[name]).T
        want = r[key].astype(np.float32)
        scale = np.abs(want).max(axis=0, keepdims=True)
        assert np.all(np.abs(got - want) <= (2e-6 if name != "acc" else 1e-5) * scale + 1e-9), name
    # the limit that drops to a quarter between s = 0.405 and 0.595 is obeyed by the record of every degree of freedom
    assert np.all(np.abs(np.array(doc["vel"])) <= (1.0 + 0.5 * np.arange(3))[:, None] * 1.02)


@pytest.mark.gpu
def test_fast_marching_trees_matches_oracle(built, oracle):
    """planning_space::fast_marching_trees as examples/test.cpp:284 calls it, and b200::fmt_batch, vs the oracle."""
    doc = json.loads(subprocess.check_output([built, "fmt", "200"]).decode())
    obs = _example_map(oracle)
    edges = np.concatenate([o["lines"] for o in obs])
    st = [0, 0, 0, 0]
    first = oracle.sample_free(200, (-1, 1, -1, 1), obs, state=st)
    want = oracle.fmt_batch(first, edges, [(-0.5, 1)], [(1, -1)], 1.0)
    assert doc["found"] == 1 and want["status"][0] == 0
    n = want["path_len"][0]
    np.testing.assert_array_equal(np.float32(doc["path"]), np.stack([want["path_x"][0, :n], want["path_y"][0, :n]], 1))
    second = oracle.sample_free(200, (-1, 1, -1, 1), obs, state=st)
    np.testing.assert_array_equal(np.float32(doc["samples"]), second)
    q = np.float32([[-0.9, -0.9, -0.9, 0.9], [-0.5, 1, 1, -1], [0.5, 0.5, -0.9, -0.9]])
    want = oracle.fmt_batch(second, edges, q[:, :2], q[:, 2:], np.float32(0.6))
    for k, got in enumerate(doc["batch"]):
        assert got["found"] == int(want["status"][k] == 0)
        if got["found"]:
            n = want["path_len"][k]
            np.testing.assert_array_equal(np.float32(got["path"]), np.stack([want["path_x"][k, :n], want["path_y"][k, :n]], 1))
            assert np.float32(got["cost"]) == want["cost"][k]
    assert sum(g["found"] for g in doc["batch"]) >= 2


@pytest.mark.gpu
def test_batched_text_request_equals_single_requests(built, tmp_path):
    """handle_text_request_batch (batched variant): five requests in one message, grouped by waypoint
    count into three scb_plan_batch calls per group; every element equals the response of the one-request handler
    (examples/zmq_test.cpp:26-101), value for value, and a request that handler throws for becomes an error object."""
    out = tmp_path / "batch.json"
    doc = json.loads(subprocess.check_output([built, "textbatch", str(out)]).decode())
    batch = json.loads(out.read_text())
    assert doc["n"] == len(batch) == 5
    for k in range(3):
        single = json.loads((tmp_path / ("batch.json.%d" % k)).read_text())
        assert len(single) > 50 and batch[k] == single, k
    for k, status in ((3, 16), (4, None)):
        single = json.loads((tmp_path / ("batch.json.%d" % k)).read_text())
        assert isinstance(single, str)  # the one-request handler threw
        assert set(batch[k]) == {"error", "status", "fail_stage"} and batch[k]["status"] != 0
        if status is not None:
            assert batch[k]["status"] == status
        else:
            assert str(batch[k]["status"]) in single and batch[k]["error"] == single


@pytest.mark.gpu
def test_cpp_multi_gpu_program(tmp_path):
    """tests/cpp/multi_gpu_c4.cpp: BASELINE config 4's shape planned from C++ through b200::multi_gpu_planner on every GPU
    of the box (summaries of all problems, then the records of the 16 fastest), no CUDA toolkit on the caller's side."""
    import torch
    exe = _build(tmp_path, "multi_gpu_c4.cpp")
    n = torch.cuda.device_count()
    doc = json.loads(subprocess.check_output([exe, str(n), "30000", "2"]).decode())
    assert doc["devices"] == n and doc["P"] == 30000 and doc["fetched_consistent"] is True
    assert 100 < doc["failed"] < 600          # ~1 % of the problems cannot brake
    assert doc["trajectories_per_s"] > 1e5
    if n == 1:  # the same batch cut into three blocks on one device: same verdict counts, same winner
        exe3 = exe  # blocks are a property of the call, not of the binary
        doc3 = json.loads(subprocess.check_output([exe3, "1", "30000", "1"]).decode())
        assert doc3["failed"] == doc["failed"] and doc3["best_T"] == doc["best_T"]
