"""The C-ABI library loads and exports every symbol include/seacurrent_b200.h declares; struct
layouts in the ctypes view match the header; argument validation fails loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import pytest

import seacurrent_b200 as scb
from seacurrent_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "seacurrent_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(scb_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    lib = abi.load_library()
    names = _declared()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(abi.SYMBOLS) == names
    assert lib.scb_version() == 202


def test_struct_layouts_match_header(tmp_path):
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "seacurrent_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                   "sizeof(scb_config),sizeof(scb_opt),sizeof(scb_batch_in),sizeof(scb_batch_out),sizeof(scb_timing),"
                   "sizeof(scb_summary),offsetof(scb_batch_in,opt),offsetof(scb_batch_out,S_cap),offsetof(scb_batch_in,x_cap));\n"
                   'printf("%zu %zu\\n",offsetof(scb_batch_out,record_f32),offsetof(scb_batch_out,n_samples));\n'
                   'printf("%zu %zu %zu %zu %zu %zu\\n",sizeof(scb_ppoly_in),offsetof(scb_ppoly_in,vel_lo_grid),sizeof(scb_ndof_in),'
                   "sizeof(scb_ndof_out),offsetof(scb_ndof_in,dt),offsetof(scb_ndof_out,n_samples));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    want = [C.sizeof(abi.ScbConfig), C.sizeof(abi.ScbOpt), C.sizeof(abi.ScbBatchIn), C.sizeof(abi.ScbBatchOut),
            C.sizeof(abi.ScbTiming), C.sizeof(abi.ScbSummary), abi.ScbBatchIn.opt.offset, abi.ScbBatchOut.S_cap.offset,
            abi.ScbBatchIn.x_cap.offset, abi.ScbBatchOut.record_f32.offset, abi.ScbBatchOut.n_samples.offset,
            C.sizeof(abi.ScbPpolyIn), abi.ScbPpolyIn.vel_lo_grid.offset, C.sizeof(abi.ScbNdofIn), C.sizeof(abi.ScbNdofOut),
            abi.ScbNdofIn.dt.offset, abi.ScbNdofOut.n_samples.offset]
    assert got == want


def test_header_is_plain_c():
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c",
                           os.path.join(ROOT, "include", "seacurrent_b200.h")])


def test_pure_functions():
    lib = abi.load_library()
    # C3 (W=10, G=1000, d=5, FP64) = 73,116 B per trajectory
    assert lib.scb_algorithmic_bytes(abi.SCB_FP64, abi.SCB_LAU_QUINTIC, 1, 10, 1000) == 73116
    assert lib.scb_algorithmic_bytes(abi.SCB_FP64, abi.SCB_LAU_QUINTIC, 4096, 20, 500) == 38236 * 4096
    assert lib.scb_algorithmic_bytes(abi.SCB_FP32, abi.SCB_LAU_QUINTIC, 1, 200, 10000) == 371200
    o = scb.default_opt()
    assert o.iters == 32 and o.lambda_ == 0.05
    assert b"fallback" in lib.scb_strerror(-2)


def test_no_device_fails_loudly():
    """No GPU in the build container: creation must fail with SCB_E_NO_DEVICE, never fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(scb.ScbError, match="no usable sm_100"):
        scb.BatchPlanner()


def test_create_rejects_bad_config():
    lib = abi.load_library()
    h = C.c_void_p()
    cfg = abi.ScbConfig(0, 7, 0, 1, 2, 2, 0)
    assert lib.scb_create(C.byref(h), C.byref(cfg)) == -1
    cfg = abi.ScbConfig(0, 0, 0, 1, 1, 2, 0)
    assert lib.scb_create(C.byref(h), C.byref(cfg)) == -1
    assert lib.scb_create(None, None) == -1


def test_product_never_imports_oracle():
    """Nothing shipped may import, link or execute the oracle (it is test infrastructure)."""
    pat = re.compile(r"(from\s+oracle|import\s+oracle|liboracle|ora_[a-z]|oracle/|oracle\.)")
    for sub in ("sea-current_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, sub)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", ".hpp")) or f == "Makefile":
                    m = pat.search(open(os.path.join(dp, f)).read())
                    assert m is None, (dp, f, m.group(0))


def test_graft_entry_build_agrees_with_the_header():
    """__graft_entry__.build() (the driver's build check) must not pin a stale ABI version."""
    import __graft_entry__ as g
    g.build()
