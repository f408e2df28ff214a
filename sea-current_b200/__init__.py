"""sea-current_b200 — B200-native implementation of sea-current's planning hot path.

The directory name carries a hyphen (it mirrors the reference's name), so import it through
the root-level shim:  `import seacurrent_b200 as scb`.

Only what the path needs lives here: csrc/ (CUDA kernels + the C ABI, libseacurrent_b200.so)
and scb/ (ctypes marshalling for tests and bench, synthetic workloads).  The C++ drop-in for
sea_current.hpp callers is include/sea_current_b200.hpp.
"""
from .scb import abi, workloads  # noqa: F401
from .scb.planner import (ALL_OUTPUTS, PER_GRID, PER_PROBLEM, PER_SAMPLE, RECORD, TRAJECTORY_OUTPUTS, SUMMARY_DTYPE, BatchPlanner, BatchResult, MultiPlanner,  # noqa: F401
                          ScbError, default_opt, gather_summaries, gpu_local_cpus, near_gpu, shard_range)
