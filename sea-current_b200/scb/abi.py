"""ctypes view of include/seacurrent_b200.h (struct layouts must match the header 1:1)."""
import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(PKG_DIR, "csrc", "libseacurrent_b200.so")

SCB_FP64, SCB_FP32 = 0, 1
SCB_REF_CUBIC, SCB_LAU_QUINTIC = 0, 1
ARCLEN_TABLE = 101
STAGE_SPLINE, STAGE_TIMING, STAGE_GEOMETRY = 1, 2, 4
LP_DEFAULT, LP_SEIDEL, LP_CLOSED_FORM = 0, 1, 2
PARAM_SPLINE, PARAM_CONST_ACCEL = 0, 1

ST_OK, ST_ERR_UNKNOWN, ST_ERR_FAIL_CONTROLLABLE, ST_ERR_FAIL_FORWARD_PASS = 0, 1, 2, 3
ST_ERR_UNINITIALIZED, ST_ERR_FAIL_FEASIBLE = 4, 5
ST_ERR_BAD_PATH, ST_ERR_BAD_LIMITS, ST_ERR_TIME_KNOTS, ST_ERR_COLLISION = 16, 17, 18, 19


class ScbConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("precision", C.c_int32), ("spline_mode", C.c_int32),
                ("max_problems", C.c_int32), ("max_waypoints", C.c_int32), ("max_gridpoints", C.c_int32),
                ("max_samples", C.c_int32)]


class ScbOpt(C.Structure):
    _fields_ = [("iters", C.c_int32), ("move_waypoints", C.c_int32), ("lambda_", C.c_double),
                ("e_init", C.c_double), ("e_min", C.c_double), ("e_max", C.c_double),
                ("step_e", C.c_double), ("step_w", C.c_double), ("w_radius", C.c_double)]


class ScbBatchIn(C.Structure):
    _fields_ = [("P", C.c_int32), ("W", C.c_int32), ("G", C.c_int32),
                ("wx", C.c_void_p), ("wy", C.c_void_p), ("start_angle", C.c_void_p),
                ("vel_lo", C.c_void_p), ("vel_hi", C.c_void_p),
                ("vel_lo_grid", C.c_void_p), ("vel_hi_grid", C.c_void_p),
                ("acc_lo", C.c_void_p), ("acc_hi", C.c_void_p),
                ("vel_start", C.c_void_p), ("vel_end", C.c_void_p),
                ("path_len", C.c_void_p), ("obstacle_edges", C.c_void_p), ("stages", C.c_int32), ("lp_mode", C.c_int32),
                ("nudge_positions", C.c_int32), ("n_obstacle_edges", C.c_int32),
                ("param_mode", C.c_int32), ("collision_chords", C.c_int32),
                ("x_cap", C.c_double), ("dt", C.c_double), ("omega_max", C.c_double), ("acc_lat_max", C.c_double),
                ("a_brake", C.c_double), ("t_react", C.c_double),
                ("opt", ScbOpt)]


class ScbBatchOut(C.Structure):
    _fields_ = [("status", C.c_void_p), ("fail_stage", C.c_void_p),
                ("ctrl_x", C.c_void_p), ("ctrl_y", C.c_void_p),
                ("arclen_total", C.c_void_p), ("arclen_table", C.c_void_p),
                ("T_total", C.c_void_p), ("cost", C.c_void_p),
                ("K_lo", C.c_void_p), ("K_hi", C.c_void_p), ("x", C.c_void_p), ("t", C.c_void_p),
                ("px", C.c_void_p), ("py", C.c_void_p), ("v", C.c_void_p), ("a", C.c_void_p),
                ("omega", C.c_void_p),
                ("S_cap", C.c_int32), ("record_f32", C.c_int32), ("n_samples", C.c_void_p),
                ("s_time", C.c_void_p), ("s_pos", C.c_void_p), ("s_vel", C.c_void_p), ("s_acc", C.c_void_p),
                ("s_px", C.c_void_p), ("s_py", C.c_void_p), ("s_omega", C.c_void_p)]


class ScbSummary(C.Structure):
    _fields_ = [("status", C.c_int32), ("fail_stage", C.c_int32), ("T_total", C.c_double),
                ("arclen_total", C.c_double), ("cost", C.c_double)]


class ScbTiming(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("spline_ms", C.c_float), ("toppra_ms", C.c_float),
                ("param_ms", C.c_float), ("sample_ms", C.c_float), ("d2h_ms", C.c_float),
                ("total_ms", C.c_float), ("launches", C.c_int32), ("path_ms", C.c_float), ("arc_ms", C.c_float)]


class ScbFmtIn(C.Structure):
    _fields_ = [("Q", C.c_int32), ("n_nodes", C.c_int32), ("n_edges", C.c_int32), ("max_path", C.c_int32),
                ("node_x", C.c_void_p), ("node_y", C.c_void_p), ("edges", C.c_void_p),
                ("init_x", C.c_void_p), ("init_y", C.c_void_p), ("goal_x", C.c_void_p), ("goal_y", C.c_void_p),
                ("rn", C.c_float), ("reserved0", C.c_int32)]


class ScbPpolyIn(C.Structure):
    _fields_ = [("P", C.c_int32), ("n_seg", C.c_int32), ("dof", C.c_int32), ("G", C.c_int32),
                ("coeffs", C.c_void_p), ("breaks", C.c_void_p), ("vel_lo", C.c_void_p), ("vel_hi", C.c_void_p),
                ("acc_lo", C.c_void_p), ("acc_hi", C.c_void_p), ("interpolation", C.c_int32), ("reserved0", C.c_int32),
                ("sd_start", C.c_double), ("sd_end", C.c_double), ("x_cap", C.c_double),
                ("vel_lo_grid", C.c_void_p), ("vel_hi_grid", C.c_void_p)]


MAX_DOF = 6


class ScbNdofIn(C.Structure):
    _fields_ = [("P", C.c_int32), ("dof", C.c_int32), ("G", C.c_int32), ("interpolation", C.c_int32),
                ("pos_start", C.c_void_p), ("pos_end", C.c_void_p), ("vel_start", C.c_void_p), ("vel_end", C.c_void_p),
                ("vel_lo", C.c_void_p), ("vel_hi", C.c_void_p), ("vel_lo_grid", C.c_void_p), ("vel_hi_grid", C.c_void_p),
                ("acc_lo", C.c_void_p), ("acc_hi", C.c_void_p), ("dt", C.c_double), ("x_cap", C.c_double),
                ("param_mode", C.c_int32), ("reserved0", C.c_int32)]


class ScbNdofOut(C.Structure):
    _fields_ = [("status", C.c_void_p), ("fail_stage", C.c_void_p), ("K_lo", C.c_void_p), ("K_hi", C.c_void_p),
                ("x", C.c_void_p), ("t", C.c_void_p), ("T_total", C.c_void_p), ("S_cap", C.c_int32), ("reserved0", C.c_int32),
                ("n_samples", C.c_void_p), ("s_time", C.c_void_p), ("s_pos", C.c_void_p), ("s_vel", C.c_void_p),
                ("s_acc", C.c_void_p)]


class ScbPpolyOut(C.Structure):
    _fields_ = [("status", C.c_void_p), ("fail_stage", C.c_void_p), ("feas_status", C.c_void_p),
                ("K_lo", C.c_void_p), ("K_hi", C.c_void_p), ("x", C.c_void_p), ("feas_lo", C.c_void_p), ("feas_hi", C.c_void_p)]


class ScbFmtOut(C.Structure):
    _fields_ = [("status", C.c_void_p), ("path_len", C.c_void_p), ("closed", C.c_void_p),
                ("path_x", C.c_void_p), ("path_y", C.c_void_p), ("cost", C.c_void_p)]


FMT_OK, FMT_NO_PATH, FMT_PATH_CAP = 0, 1, 2

# every symbol include/seacurrent_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "scb_version": (C.c_int, []),
    "scb_strerror": (C.c_char_p, [C.c_int]),
    "scb_default_opt": (None, [C.POINTER(ScbOpt)]),
    "scb_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(ScbConfig)]),
    "scb_destroy": (C.c_int, [C.c_void_p]),
    "scb_plan_batch": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.POINTER(ScbBatchOut)]),
    "scb_plan_batch_device": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.POINTER(ScbBatchOut), C.c_int]),
    "scb_plan_batch_async": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.POINTER(ScbBatchOut), C.c_void_p]),
    "scb_synchronize": (C.c_int, [C.c_void_p]),
    "scb_plan_batch_summary": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.c_int32, C.c_void_p]),
    "scb_fetch": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.POINTER(ScbBatchOut)]),
    "scb_multi_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(ScbConfig), C.c_int32, C.c_void_p]),
    "scb_multi_destroy": (C.c_int, [C.c_void_p]),
    "scb_multi_devices": (C.c_int32, [C.c_void_p]),
    "scb_multi_plan_batch": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.POINTER(ScbBatchOut)]),
    "scb_multi_plan_summary": (C.c_int, [C.c_void_p, C.POINTER(ScbBatchIn), C.c_int32, C.c_void_p]),
    "scb_multi_fetch": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.POINTER(ScbBatchOut)]),
    "scb_shard_range": (None, [C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "scb_multi_get_timing": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(ScbTiming)]),
    "scb_set_host_chunk": (C.c_int, [C.c_void_p, C.c_int32]),
    "scb_fmt_batch": (C.c_int, [C.c_void_p, C.POINTER(ScbFmtIn), C.POINTER(ScbFmtOut)]),
    "scb_toppra_ppoly": (C.c_int, [C.c_void_p, C.POINTER(ScbPpolyIn), C.POINTER(ScbPpolyOut)]),
    "scb_gen_vel_prof": (C.c_int, [C.c_void_p, C.POINTER(ScbNdofIn), C.POINTER(ScbNdofOut)]),
    "scb_param_static": (C.c_int, [C.c_int32, C.c_int32] + [C.c_void_p] * 10),
    "scb_static_last_error": (C.c_char_p, []),
    "scb_alloc_pinned": (C.c_void_p, [C.c_size_t]),
    "scb_free_pinned": (None, [C.c_void_p]),
    "scb_get_timing": (C.c_int, [C.c_void_p, C.POINTER(ScbTiming)]),
    "scb_last_cuda_error": (C.c_char_p, [C.c_void_p]),
    "scb_stream": (C.c_void_p, [C.c_void_p]),
    "scb_algorithmic_bytes": (C.c_int64, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "scb_pack_summary_device": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
}

_lib = None


def load_library(path=None):
    """dlopen libseacurrent_b200.so and bind every declared symbol.  Raises if the library has
    not been built: there is no CPU fallback behind this package."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError("libseacurrent_b200.so is not built (%s); run `python -c 'import __graft_entry__ as g; "
                           "g.build()'` or `make -C sea-current_b200/csrc`. There is no CPU fallback." % p)
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib
