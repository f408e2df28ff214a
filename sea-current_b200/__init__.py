"""sea-current_b200 — B200-native implementation of sea-current's planning hot path.

The directory name carries a hyphen (it mirrors the reference's name), so import it through
the root-level shim:  `import seacurrent_b200 as scb`.
"""
from .scb import abi  # noqa: F401
from .scb.planner import (BatchPlanner, BatchResult, bezier_spline, gen_vel_prof, planning_space,  # noqa: F401
                          serialize_path_to_json, velocity_profile, arclength_data, default_opt)
