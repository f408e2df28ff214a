/* seacurrent_b200.h — C ABI of the B200-native sea-current planning hot path.
 *
 * Drop-in boundary for the path  waypoints -> Bezier spline (reference cubic, or Lau09 quintic +
 * curvature/arc-length optimiser) -> toppra time parameterisation -> sampled trajectory,
 * batched over P independent planning problems.  Everything behind these entry points is
 * hand-written sm_100a CUDA; there is no CPU fallback: every entry point returns
 * SCB_E_NO_DEVICE / SCB_E_CUDA when no B200-class device can be used.
 *
 * Reference interfaces replaced (paths relative to the sea-current tree; "sc:" =
 * sea_current.hpp, "tp:" = third-party/toppra/cpp/src/toppra/):
 *   bezier_spline::from_path            sc:402, sc:599-683   -> ctrl_x / ctrl_y
 *   bezier_spline::arclength            sc:413, sc:767-896   -> arclen_total / arclen_table
 *   gen_vel_prof<1>                     sc:1191-1265         -> status, K_lo/K_hi, x, t, s_* streams
 *   gen_vel_prof<N>, N > 1              sc:1191-1265         -> scb_gen_vel_prof (below)
 *     TOPPRA::computePathParametrization tp:algorithm.cpp:14-26 (ReturnCode tp:algorithm.hpp:13-32)
 *     parametrizer::Spline              tp:parametrizer/spline.cpp:10-45
 *   bezier_spline::resample             sc:419, sc:898-1005  -> px / py, s_px / s_py
 *   bezier_spline::angular_velocity     sc:416, sc:1055-1067 -> omega, s_omega
 *   serialize_path_to_json              sc:1423-1457         -> host side, include/sea_current_b200.hpp
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types cross the ABI; nothing throws.
 *   - return value: 0 = OK, < 0 = API / CUDA error (scb_strerror).  A per-problem planning
 *     failure never fails the batch: it is reported in status[p] with toppra's ReturnCode
 *     numbering and fail_stage[p] = the gridpoint index at which the LP became infeasible.
 *   - "real" buffers hold double when the context precision is SCB_FP64 and float when it is
 *     SCB_FP32.
 *   - layouts: per-path arrays are problem-major  [P][n]  (one path's data contiguous: the
 *     spline kernels run one warp per path); per-gridpoint and per-sample arrays are
 *     gridpoint-major  [G][P]  (problem index fastest: the toppra kernels run one thread per
 *     problem, so adjacent threads touch adjacent addresses).
 *   - any output pointer may be NULL: that array is then not produced.
 *   - one scb_ctx per host thread and GPU; calls on a ctx are serialised on its stream.
 */
#ifndef SEACURRENT_B200_H
#define SEACURRENT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCB_VERSION 202 /* 0.2.2 */

/* API-level error codes (function return values) */
enum {
  SCB_OK = 0,
  SCB_E_INVALID = -1,   /* bad argument (NULL, P/W/G out of range, capacity exceeded) */
  SCB_E_NO_DEVICE = -2, /* no CUDA device / not an sm_100 part */
  SCB_E_CUDA = -3,      /* CUDA runtime error; see scb_last_cuda_error */
  SCB_E_NOMEM = -4,
  SCB_E_TOO_LARGE = -5  /* scb_param_static: G * P beyond the static scratch (SCB_STATIC_MAX_GP) */
};

/* per-problem status[p].  0-5 are toppra's ReturnCode (tp:algorithm.hpp:13-32). */
enum {
  SCB_ST_OK = 0,
  SCB_ST_ERR_UNKNOWN = 1,
  SCB_ST_ERR_FAIL_CONTROLLABLE = 2,
  SCB_ST_ERR_FAIL_FORWARD_PASS = 3,
  SCB_ST_ERR_UNINITIALIZED = 4,
  SCB_ST_ERR_FAIL_FEASIBLE = 5,
  /* conditions the reference asserts on or throws for (sc:34-52 SC_ASSERT; tp:constraint/
   * linear_joint_velocity.cpp:54; tp:solver.cpp:36-37) */
  SCB_ST_ERR_BAD_PATH = 16,     /* non-finite waypoint, zero-length leg, or arclength <= 0 */
  SCB_ST_ERR_BAD_LIMITS = 17,   /* velocity box excludes 0 where q'(s) = 0 (tp:constraint/linear_joint_velocity.cpp:54),
                                   acc_lo > acc_hi (LinearJointAcceleration::check, tp:constraint/
                                   linear_joint_acceleration.cpp:18-22), or any limit (scalar or per gridpoint) is NaN.
                                   Infinite limits are legal ("no limit").  A per-gridpoint velocity box with
                                   lo > hi is NOT rejected here: like the reference's LinearJointVelocityVarying
                                   (sc:1177-1188, no check) it makes the stage LP infeasible */
  SCB_ST_ERR_TIME_KNOTS = 18,   /* a knot interval < 1e-8 s (the reference drops such knots,
                                   tp:parametrizer/spline.cpp:25-33; unreachable for x_cap = 100) */
  SCB_ST_ERR_COLLISION = 19     /* the finished spline crosses an obstacle edge (scb_batch_in.collision_chords) */
};

/* SCB_FP32 is a STORAGE mode, not a faster compute mode: every `real` buffer (inputs, outputs, tables) is float and the
 * arc-length / geometry kernels compute in float; the stages that take decisions - toppra's passes (tolerances 1e-13 /
 * 1e-10 / 1e-9), the parametrizer recurrences and the sign-driven optimiser - compute in FP64 in both modes.  Parity:
 * 1e-4 of the FP64 results.  For float RECORDS with FP64 arithmetic use SCB_FP64 + scb_batch_out.record_f32. */
enum { SCB_FP64 = 0, SCB_FP32 = 1 };
enum { SCB_REF_CUBIC = 0,    /* the reference's construction, sc:599-683 */
       SCB_LAU_QUINTIC = 1 };/* Lau et al. IROS'09 sec. III-A quintic + optimiser */

#define SCB_ARCLEN_SUBDIV 100 /* sub-intervals per segment (precision 0.01, sc:767) */
#define SCB_ARCLEN_TABLE (SCB_ARCLEN_SUBDIV + 1)

typedef struct scb_ctx scb_ctx;

typedef struct scb_config {
  int32_t device;         /* CUDA ordinal */
  int32_t precision;      /* SCB_FP64 | SCB_FP32 */
  int32_t spline_mode;    /* SCB_REF_CUBIC | SCB_LAU_QUINTIC */
  int32_t max_problems;   /* capacities used to size device scratch and staging buffers */
  int32_t max_waypoints;
  int32_t max_gridpoints;
  int32_t max_samples;    /* S_cap of the uniform-dt stream; 0 = stream never requested */
} scb_config;

/* Optimiser options (LAU_QUINTIC only; DESIGN.md "A6/A7").  iters = 0 disables it. */
typedef struct scb_opt {
  int32_t iters;            /* K: fixed RPROP iteration budget */
  int32_t move_waypoints;   /* 1: inner waypoints are parameters too; 0: elongations only */
  double lambda;            /* weight of the arc-length term */
  double e_init, e_min, e_max; /* tangent elongation factor: start value and clamp */
  double step_e, step_w;    /* initial RPROP steps (elongation; waypoint, as a fraction of the
                               tangent magnitude) */
  double w_radius;          /* inner waypoints may move at most this far from their input */
} scb_opt;

/* stages of the path one call runs (scb_batch_in.stages; 0 = all three) */
enum {
  SCB_STAGE_SPLINE = 1,   /* from_path (+ optimiser) + arclength: wx/wy -> ctrl, arclen_*, cost */
  SCB_STAGE_TIMING = 2,   /* gen_vel_prof: L -> status, K, x, t, v, a, T_total, s_time/pos/vel/acc.
                             Without SCB_STAGE_SPLINE the path length is read from path_len. */
  SCB_STAGE_GEOMETRY = 4  /* resample + angular_velocity: arc length -> px, py, omega and
                             s_px, s_py, s_omega.
                             Without SCB_STAGE_TIMING, out->n_samples / s_pos (and s_vel, if
                             s_omega is wanted) are INPUTS: the positions to resample at.
                             Without SCB_STAGE_SPLINE (and then without SCB_STAGE_TIMING), out->ctrl_x /
                             ctrl_y / arclen_table are INPUTS too: the spline of an earlier
                             SCB_STAGE_SPLINE call (the reference's bezier_spline carries its control
                             points into resample / angular_velocity the same way, sc:898, sc:1055). */
};

/* Stage-LP back ends of the toppra passes (tp:solver/seidel.cpp:265-347).  Both return the same
 * optimum (it is unique) and the same feasibility verdicts; tests compare them LP by LP. */
enum {
  SCB_LP_DEFAULT = 0,      /* = SCB_LP_CLOSED_FORM */
  SCB_LP_SEIDEL = 1,       /* the reference's incremental sequence, operation for operation: given the
                              same path length, bit-identical to the reference's FP64 results; ~6x slower */
  SCB_LP_CLOSED_FORM = 2   /* eliminate u: three determinants and six quotients per stage.  Returns the exact
                              optimum; the reference's Seidel stops at a vertex within its ABSOLUTE row tolerance
                              (1e-13) of it, so the two agree to ~1e-13 absolute in x = sdot^2: below 1e-9
                              relative unless x itself is below ~1e-4 (millimetre-scale paths with acceleration
                              limits ~1e-7: 1.4e-8 measured).  Use SCB_LP_SEIDEL there for 1e-9 agreement. */
};

/* Parametrizer of the timing stage: how sdot^2 at the gridpoints becomes a function of time. */
enum {
  SCB_PARAM_SPLINE = 0,      /* toppra::parametrizer::Spline, what gen_vel_prof builds (sc:1231): trapezoid knot times,
                                clamped cubic spline q(t) through (t_i, q(s_i))  (tp:parametrizer/spline.cpp:10-45) */
  SCB_PARAM_CONST_ACCEL = 1  /* toppra::parametrizer::ConstAccel (tp:parametrizer/const_accel.cpp:19-115): constant path
                                acceleration on every grid interval, s(t) a quadratic per interval, nothing solved;
                                t, v, a, omega, T_total and the uniform-dt stream are then its values */
};

typedef struct scb_batch_in {
  int32_t P, W, G;          /* problems, waypoints per problem (>= 2), toppra gridpoints (>= 2;
                               the reference's default is 101, tp:algorithm.hpp:140).  W <= 522 in SCB_LAU_QUINTIC
                               (beyond 512 waypoints the optimiser state of one path lives in one SM's shared
                               memory), a few thousand in SCB_REF_CUBIC; a longer path is SCB_E_INVALID */
  const void* wx;           /* real [P][W]  waypoint x */
  const void* wy;           /* real [P][W]  waypoint y */
  const void* start_angle;  /* real [P] or NULL; NaN = atan2 of the first leg (sc:606-609) */
  const void* vel_lo;       /* real [P]  velocity box (vel_lim_func, sc:1175) ... */
  const void* vel_hi;       /* real [P] */
  const void* vel_lo_grid;  /* real [G][P] or NULL: per-gridpoint limits replace vel_lo/vel_hi */
  const void* vel_hi_grid;  /* real [G][P] or NULL */
  const void* acc_lo;       /* real [P]  acc_min */
  const void* acc_hi;       /* real [P]  acc_max */
  const void* vel_start;    /* real [P] or NULL (= 0): q'(0) of the Hermite path (sc:1213-1222) */
  const void* vel_end;      /* real [P] or NULL (= 0) */
  const void* path_len;     /* real [P]: pos_end - pos_start, read only when SCB_STAGE_SPLINE is off */
  const void* obstacle_edges; /* real [n_obstacle_edges][4] = (x0, y0, x1, y1) or NULL: the obstacle.lines of the
                               planning_space (sc:193-284, 298-319), obstacle after obstacle, shared by every problem
                               of the batch.  from_path cuts every tangent back where W -> W +- T crosses an edge
                               (shrink_tangent, sc:575-596).  The clipping is SCB_REF_CUBIC only (the convex-hull
                               argument behind it is the cubic's); SCB_LAU_QUINTIC takes an edge list only together
                               with collision_chords > 0. */
  int32_t stages;           /* SCB_STAGE_* bitmask; 0 = all */
  int32_t lp_mode;          /* SCB_LP_*: how the stage LPs of toppra are solved; 0 = default */
  int32_t nudge_positions;  /* geometry-only calls: 1 = repair the caller's s_pos in place before resampling
                               (ends pinned to 0 / L, non-monotone entries averaged, clamp; sc:902-912) */
  int32_t n_obstacle_edges; /* 0 = no obstacles */
  int32_t param_mode;       /* SCB_PARAM_*; 0 = the reference's Spline parametrizer */
  int32_t collision_chords; /* M > 0: collision verdict of the finished spline (Lau09 IV-B: a trajectory that is not
                               collision-free has infinite cost): every segment is cut into M chords and each chord
                               tested against obstacle_edges with the reference's segment test (sc:142-178); a hit
                               gives status SCB_ST_ERR_COLLISION and cost = +inf, and the problem is not timed.
                               With it, SCB_LAU_QUINTIC accepts an edge list too (no tangent clipping there).  0 = off */
  double x_cap;             /* upper bound of sdot^2 in the backward pass; 0 = reference's 100
                               (tp:algorithm.hpp:152 m_initXBound) */
  double dt;                /* uniform-dt stream period; 0 = reference's 0.02 (sc:1199) */
  /* Curvature-aware velocity limits (Lau09 III-B "isolated constraints"; the reference's hook for them is the
   * position-dependent vel_lim_func, sc:1175-1188, examples/test.cpp:194-209).  With kappa_i the curvature at
   * the arc length of gridpoint i, the velocity box of that gridpoint is intersected with [-c_i, c_i],
   *   c_i = min(omega_max / |kappa_i|, sqrt(acc_lat_max / |kappa_i|));  0 switches a term off.
   * Needs SCB_STAGE_SPLINE | SCB_STAGE_TIMING | SCB_STAGE_GEOMETRY in the same call. */
  double omega_max;         /* |angular velocity| <= omega_max  */
  double acc_lat_max;       /* centripetal acceleration v^2 |kappa| <= acc_lat_max */
  /* Obstacle-distance speed limit (Lau09 III-B eq. 2, the third "isolated constraint"): braking distance plus the
   * distance travelled during the reaction time must stay below the distance d_i from the curve point of gridpoint i
   * to the closest obstacle edge (point-to-segment distance over obstacle_edges):
   *   |v_i| <= -a_brake t_react + sqrt(a_brake^2 t_react^2 + 2 a_brake d_i).
   * a_brake = 0 switches it off.  Needs all three stages and n_obstacle_edges > 0 in the same call. */
  double a_brake;           /* maximum deceleration when braking, > 0 */
  double t_react;           /* reaction time of the system, >= 0 */
  scb_opt opt;
} scb_batch_in;

typedef struct scb_batch_out {
  int32_t* status;          /* [P] */
  int32_t* fail_stage;      /* [P]  -1 when status is OK */
  void* ctrl_x;             /* real [P][(W-1)*(d+1)]  d = 3 (REF_CUBIC) or 5 (LAU_QUINTIC) */
  void* ctrl_y;
  void* arclen_total;       /* real [P]  L */
  void* arclen_table;       /* real [P][(W-1)*SCB_ARCLEN_TABLE] cumulative length per segment */
  void* T_total;            /* real [P]  duration */
  void* cost;               /* real [P]  J after the optimiser (quintic) / J of the cubic */
  /* one record per toppra gridpoint, [G][P] each */
  void* K_lo; void* K_hi;   /* controllable sets */
  void* x;                  /* parametrization sdot^2 */
  void* t;                  /* knot times (trapezoid rule) */
  void* px; void* py;       /* curve point at arc length q(s_i) */
  void* v; void* a;         /* dq/dt, d2q/dt2 of the parametrizer spline at t_i */
  void* omega;              /* v * curvature */
  /* uniform-dt stream (what gen_vel_prof + resample + angular_velocity return), [S_cap][P] */
  int32_t S_cap;
  int32_t record_f32;       /* 1: the trajectory RECORD - t, px, py, v, a, omega and the s_* stream - is float whatever the
                               ctx precision: what the reference hands its caller (velocity_profile holds VectorXf
                               pos / vel / acc, sc:379-386, 1245-1262; resample and angular_velocity return float).
                               Arithmetic is unchanged (FP64 in SCB_FP64); values are rounded once, in the kernels'
                               stores, so a host call moves half the bytes.  K_lo / K_hi / x, control points, tables
                               and the [P] scalars stay `real`.  In a geometry-only call the s_pos / s_vel INPUTS are
                               float too.  0: everything is `real`. */
  int32_t* n_samples;       /* [P]  ceil(T/dt) (sc:1237); entries beyond S_cap are not written */
  void* s_time; void* s_pos; void* s_vel; void* s_acc; void* s_px; void* s_py; void* s_omega;
} scb_batch_out;

/* per-stage device time of the last scb_plan_batch* call on this ctx, milliseconds */
typedef struct scb_timing {
  float h2d_ms, spline_ms, toppra_ms, param_ms, sample_ms, d2h_ms, total_ms;
  int32_t launches;         /* kernels launched by that call */
  float path_ms, arc_ms;    /* spline_ms split into its two kernels (waypoints -> control points; tables + geometry) */
} scb_timing;

int scb_version(void);
const char* scb_strerror(int code);
void scb_default_opt(scb_opt* opt);

int scb_create(scb_ctx** out, const scb_config* cfg);
int scb_destroy(scb_ctx* ctx);

/* Host buffers in, host buffers out: H2D copy, kernels, D2H copy, synchronous.  A batch of more than one chunk of
 * problems is pipelined: the kernels of chunk c+1 overlap the D2H of chunk c (second stream, two sets of device
 * mirrors), so device memory is sized by the chunk, not by P.  scb_timing of such a call: total_ms = the call,
 * d2h_ms = first D2H byte to last, h2d_ms = 0 (folded into the chunks), stage times = the last chunk's. */
int scb_plan_batch(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out);
/* problems per chunk of scb_plan_batch: 0 = automatic (8,192 once P >= 16,384, else one chunk), < 0 = never chunk.
 * Results do not depend on it (problems are independent). */
int scb_set_host_chunk(scb_ctx* ctx, int32_t problems);

/* Device buffers in and out (already resident in HBM); enqueued on the ctx stream and
 * synchronised before returning unless `async` is non-zero. */
int scb_plan_batch_device(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, int async);
/* The same, enqueued on the CALLER's stream (a cudaStream_t, passed as void* so that this header needs no CUDA
 * header) and never synchronised: an MPC loop overlaps its own kernels and copies with the planning of the next
 * batch and orders them with its own events.  The ctx's scratch is shared by all calls on the ctx, so a call on a
 * different stream than the previous one first makes that stream wait (device side, cudaStreamWaitEvent) for the
 * previous call's kernels.  scb_get_timing after such a call needs the caller to have synchronised `stream`. */
int scb_plan_batch_async(scb_ctx* ctx, const scb_batch_in* in, const scb_batch_out* out, void* stream);
int scb_synchronize(scb_ctx* ctx);

/* Page-locked host memory for the host-buffer entry points (their copies run at PCIe speed only from / to pinned
 * memory), so that a caller needs no CUDA toolkit of its own.  NULL on failure. */
void* scb_alloc_pinned(size_t bytes);
void scb_free_pinned(void* p);

int scb_get_timing(scb_ctx* ctx, scb_timing* out);
const char* scb_last_cuda_error(scb_ctx* ctx);
/* the cudaStream_t the ctx launches on (so a caller can order its own work / CUDA events) */
void* scb_stream(scb_ctx* ctx);
/* algorithmic bytes one launch over this batch must move (SURVEY 8-D4 formula) */
int64_t scb_algorithmic_bytes(int32_t precision, int32_t spline_mode, int32_t P, int32_t W, int32_t G);

/* Fixed-size per-problem summary: what an MPC outer loop needs to rank rollouts, and the only
 * payload the multi-GPU path exchanges (NCCL all-gather of P_local records per rank; full
 * trajectories stay sharded on the GPU that planned them). */
typedef struct scb_summary {
  int32_t status, fail_stage;
  double T_total, arclen_total, cost;   /* always double, whatever the ctx precision */
} scb_summary;
/* Device pointers (outputs of a previous scb_plan_batch_device on this ctx; any of fail_stage /
 * T_total / arclen_total / cost may be NULL = 0) -> dst[P] on the device.  Enqueued on the ctx
 * stream; not synchronised. */
int scb_pack_summary_device(scb_ctx* ctx, int32_t P, const int32_t* status, const int32_t* fail_stage,
                            const void* T_total, const void* arclen_total, const void* cost, scb_summary* dst);

/* ---- Rank, then fetch: what an MPC outer loop does with 65,536 candidate rollouts is rank them by
 * the 32-byte summaries and read the trajectories of the few it keeps.  scb_plan_batch_summary plans a batch of HOST
 * inputs, keeps every output of scb_batch_out (except the uniform-dt stream) resident in HBM inside the ctx, and
 * copies back only summaries[P]; scb_fetch then gathers the records of the n problems the caller names (device-side
 * gather into compact [G][n] / [n][..] arrays, one small D2H per array) into host arrays laid out as scb_batch_out
 * with P := n (record arrays float if the plan was made with record_f32).  The retained results stay valid until
 * the next scb_plan_batch_summary on the ctx. */
int scb_plan_batch_summary(scb_ctx* ctx, const scb_batch_in* in, int32_t record_f32, scb_summary* summaries);
int scb_fetch(scb_ctx* ctx, int32_t n, const int32_t* problem_ids, const scb_batch_out* out);

/* ---- One process, several GPUs: problems share nothing, so a batch is cut into contiguous blocks of
 * ceil(P / R) problems, block r planned on device r by its own scb_ctx and host thread; there is no data-path
 * collective, and since the summaries land in the caller's host array no gather either (a multi-PROCESS job
 * all-gathers the same 32-byte records over NCCL: bench.py).  All pointers are HOST pointers. */
typedef struct scb_multi scb_multi;
/* cfg->device is ignored; devices = NULL means 0 .. n_devices-1; cfg->max_problems is the capacity PER DEVICE */
int scb_multi_create(scb_multi** out, const scb_config* cfg, int32_t n_devices, const int32_t* devices);
int scb_multi_destroy(scb_multi* m);
int32_t scb_multi_devices(const scb_multi* m);
/* full records: every device writes its block of columns of the caller's [G][P] arrays (scb_plan_batch per block) */
int scb_multi_plan_batch(scb_multi* m, const scb_batch_in* in, const scb_batch_out* out);
/* rank-then-fetch over all devices: summaries[P] in problem order; fetch routes every id to the device that owns it */
int scb_multi_plan_summary(scb_multi* m, const scb_batch_in* in, int32_t record_f32, scb_summary* summaries);
int scb_multi_fetch(scb_multi* m, int32_t n, const int32_t* problem_ids, const scb_batch_out* out);
/* [lo, hi) of the block device r plans out of P problems on R devices */
void scb_shard_range(int32_t P, int32_t r, int32_t R, int32_t* lo, int32_t* hi);
/* timing of the last call on device r's ctx */
int scb_multi_get_timing(scb_multi* m, int32_t r, scb_timing* out);

/* ---- toppra on a caller-supplied path (what the reference's toppra does beyond the one configuration gen_vel_prof<1>
 * uses): a piecewise-cubic path of one to SCB_MAX_DOF degrees of freedom
 * (PiecewisePolyPath, tp:geometric_path/piecewise_poly_path.cpp:215-252), box limits per degree of freedom, N + 1 =
 * G gridpoints LinSpaced over the path interval, collocation or interpolation discretisation
 * (tp:constraint.hpp:13-14), the reference's Seidel sequence operation for operation; controllable sets, forward
 * pass (tp:algorithm.cpp:14-68, tp:algorithm/toppra.cpp:15-46) and feasible sets (computeFeasibleSets,
 * tp:algorithm.cpp:70-103).  1 <= dof <= SCB_MAX_DOF, both discretisations: a stage LP has 2 + 2 dof rows (collocation)
 * or 2 + 4 dof (interpolation); six rows run the register-resident solver, more rows a looped one.
 * This is the entry the reference's own known-answer vectors (cpp/tests/test_algorithm.cpp:64-165: 2-DOF, 3 segments,
 * N = 50, collocation) reach a KERNEL through.  Always FP64; host pointers; outputs [G][P]. */
#define SCB_MAX_DOF 6
typedef struct scb_ppoly_in {
  int32_t P, n_seg, dof, G;
  const double* coeffs;     /* [P][n_seg][4][dof]: row 0 multiplies (s - s_seg)^3, row 3 is the value */
  const double* breaks;     /* [n_seg + 1], strictly increasing, shared by the batch */
  const double* vel_lo;     /* [P][dof]; may be NULL when the grids below are given */
  const double* vel_hi;
  const double* acc_lo;     /* [P][dof] */
  const double* acc_hi;
  int32_t interpolation;    /* 0: collocation, 1: interpolation */
  int32_t reserved0;
  double sd_start, sd_end;  /* computePathParametrization(vel_start, vel_end) */
  double x_cap;             /* 0 = the reference's 100 */
  const double* vel_lo_grid;/* [G][dof][P] or NULL: velocity limits that vary along the path (the reference's */
  const double* vel_hi_grid;/* LinearJointVelocityVarying, sc:1177-1188, tabulated at the gridpoints) */
} scb_ppoly_in;
typedef struct scb_ppoly_out {
  int32_t* status;          /* [P] toppra ReturnCode numbering (+ SCB_ST_ERR_BAD_LIMITS) */
  int32_t* fail_stage;      /* [P] or NULL */
  int32_t* feas_status;     /* [P] or NULL: SCB_ST_OK / SCB_ST_ERR_FAIL_FEASIBLE */
  double* K_lo; double* K_hi;       /* [G][P] controllable sets */
  double* x;                        /* [G][P] parametrization */
  double* feas_lo; double* feas_hi; /* [G][P] or NULL: feasible sets */
} scb_ppoly_out;
int scb_toppra_ppoly(scb_ctx* ctx, const scb_ppoly_in* in, const scb_ppoly_out* out);

/* ---- gen_vel_prof<N> for N degrees of freedom (sc:1191-1265; `requires (N >= 1)`): the straight Hermite path
 * pos_start -> pos_end with end slopes vel_start / vel_end over s in [0, 1] (CubicHermiteSpline, sc:1210-1219;
 * tp:geometric_path/piecewise_poly_path.cpp:352-379), velocity limits per degree of freedom (constant, or tabulated
 * at the gridpoints for a vel_lim_func), acceleration limits with the interpolation discretisation (sc:1207), toppra,
 * then the parametrizer and the uniform-dt record of every degree of freedom (sc:1231-1262).  The LP passes are
 * toppra_ppoly_kernel<N, interpolation>; the parametrizer and sampler are the timing stage's own kernels run once per
 * degree of freedom on the shared sdot^2 (the spline parametrizer is one cubic spline per degree of freedom over the
 * same knot times, tp:parametrizer/spline.cpp:10-45).  FP64; host pointers.
 * With dof = 1 this is scb_plan_batch's SCB_STAGE_TIMING with SCB_LP_SEIDEL. */
typedef struct scb_ndof_in {
  int32_t P, dof, G;        /* 1 <= dof <= SCB_MAX_DOF; G gridpoints LinSpaced over [0, 1] */
  int32_t interpolation;    /* 1 = what gen_vel_prof builds (sc:1207); 0 = collocation */
  const double* pos_start;  /* [P][dof] */
  const double* pos_end;
  const double* vel_start;  /* [P][dof] end slopes dq/ds of the Hermite path, or NULL = 0 */
  const double* vel_end;
  const double* vel_lo;     /* [P][dof], or NULL when the grids are given */
  const double* vel_hi;
  const double* vel_lo_grid;/* [G][dof][P] or NULL */
  const double* vel_hi_grid;
  const double* acc_lo;     /* [P][dof] */
  const double* acc_hi;
  double dt;                /* sample period; 0 = the reference's 0.02f */
  double x_cap;             /* 0 = the reference's 100 */
  int32_t param_mode;       /* SCB_PARAM_* */
  int32_t reserved0;
} scb_ndof_in;
typedef struct scb_ndof_out {
  int32_t* status;          /* [P] */
  int32_t* fail_stage;      /* [P] or NULL */
  double* K_lo; double* K_hi; /* [G][P] or NULL */
  double* x;                /* [G][P] or NULL: sdot^2 at the gridpoints */
  double* t;                /* [G][P] or NULL: knot times */
  double* T_total;          /* [P] or NULL */
  int32_t S_cap;            /* rows of the sample arrays (<= the ctx's max_samples); 0 = no record */
  int32_t reserved0;
  int32_t* n_samples;       /* [P] or NULL: ceil(T / dt), may exceed S_cap (then the record is truncated) */
  double* s_time;           /* [S_cap][P] or NULL */
  double* s_pos;            /* [dof][S_cap][P] or NULL: velocity_profile.pos[k] of problem p = s_pos[k][.][p] */
  double* s_vel;
  double* s_acc;
} scb_ndof_out;
int scb_gen_vel_prof(scb_ctx* ctx, const scb_ndof_in* in, const scb_ndof_out* out);

/* ---- Allocation-free parametrizer.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The spline parametrizer of the timing stage (what scb_plan_batch_device runs after the toppra passes:
 * tp:parametrizer/spline.cpp:10-45 as a Thomas solve) on the CALLER's device arrays, with its two [G][P] scratch arrays
 * in one file-scope device array of the library: G * P <= SCB_STATIC_MAX_GP (2^20; 16 MiB that every GPU which loads
 * the library holds until exit).  The call allocates nothing (no cudaMalloc / cudaFree, no context, no stream or event
 * creation), enqueues one kernel on `stream` (a cudaStream_t of the CURRENT device; NULL = the default stream) and
 * does not synchronise.  All pointers are device pointers of the current device, FP64, [G][P] arrays problem-fastest:
 *   path_len [P], x [G][P] (sdot^2 at the gridpoints), vel_start / vel_end [P] or NULL (= 0): inputs;
 *   status [P]: in / out - a problem whose status is not SCB_ST_OK on entry gets an all-zero record; the kernel
 *               sets SCB_ST_ERR_TIME_KNOTS where the knot times do not increase;
 *   t [G][P] (required: the knot times are the sweep's own state), v, a [G][P] or NULL, T_total [P] or NULL: outputs.
 * Results are bit-identical to the t / v / a / T_total of scb_plan_batch_device on the same x and path length.
 * Every call shares the scratch array, which keeps the last call's contents (nothing in it has to start at zero: a
 * call reads only what it wrote itself): calls that OVERLAP on the GPU race - order them on one stream or with events.
 * Returns SCB_E_TOO_LARGE beyond the maximum, SCB_E_INVALID for bad arguments, SCB_E_CUDA for a runtime error;
 * scb_static_last_error() has the text (per host thread). */
#define SCB_STATIC_MAX_GP (1 << 20)
int scb_param_static(int32_t P, int32_t G, const double* path_len, const double* x, const double* vel_start,
                     const double* vel_end, int32_t* status, double* t, double* v, double* a, double* T_total, void* stream);
const char* scb_static_last_error(void);

/* ---- Global planner front end: FMT* over a shared set of sampled nodes, one CTA per query.
 * Replaces planning_space::fast_marching_trees (sc:1339-1407) with cost (sc:1315-1326), near (sc:1328-1337),
 * intersects (sc:142-178), for Q (x_init, x_goal) pairs over ONE map: the free-space samples
 * (planning_space::sample_free, sc:1294-1313 - host side, include/sea_current_b200.hpp) and the obstacle edge list.
 * Always float, whatever the ctx precision (the reference's planner is float32).  Host pointers.
 * A node is "near" another when their distance is <= rn * rn (the reference compares pt_dist with pow(rn, 2),
 * sc:1331), and an edge's cost is the Euclidean distance, FLT_MAX when it crosses an obstacle edge. */
typedef struct scb_fmt_in {
  int32_t Q;                /* queries */
  int32_t n_nodes;          /* shared sample nodes, <= 8192; distinct (a point set) */
  int32_t n_edges;          /* obstacle edges */
  int32_t max_path;         /* capacity of one row of path_x / path_y */
  const float* node_x;      /* [n_nodes] */
  const float* node_y;
  const float* edges;       /* [n_edges][4] = (x0, y0, x1, y1), or NULL */
  const float* init_x;      /* [Q] */
  const float* init_y;
  const float* goal_x;      /* [Q] */
  const float* goal_y;
  float rn;                 /* connection radius parameter */
  int32_t reserved0;
} scb_fmt_in;

enum { SCB_FMT_OK = 0,        /* path found */
       SCB_FMT_NO_PATH = 1,   /* the open set ran empty: the reference returns std::nullopt */
       SCB_FMT_PATH_CAP = 2 };/* found, but longer than max_path: path_len and cost are set, the row is not written */

typedef struct scb_fmt_out {
  int32_t* status;          /* [Q] SCB_FMT_* */
  int32_t* path_len;        /* [Q] nodes on the path, x_init and x_goal included */
  int32_t* closed;          /* [Q] or NULL: nodes expanded (size of V_closed) */
  float* path_x;            /* [Q][max_path] or NULL: x_init ... x_goal */
  float* path_y;
  float* cost;              /* [Q] cost-to-come of x_goal (FLT_MAX when there is no path) */
} scb_fmt_out;

int scb_fmt_batch(scb_ctx* ctx, const scb_fmt_in* in, const scb_fmt_out* out);

#ifdef __cplusplus
}
#endif
#endif /* SEACURRENT_B200_H */
