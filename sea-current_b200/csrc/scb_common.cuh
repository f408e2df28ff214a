// Shared device helpers for the sea-current B200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/seacurrent_b200.h"

namespace scb {

constexpr int kWarp = 32;
constexpr int kTable = SCB_ARCLEN_TABLE;   // 101 cumulative entries per segment
constexpr int kSub = SCB_ARCLEN_SUBDIV;    // 100 sub-intervals per segment (sc:767, precision 0.01)
constexpr int kInvNewton = 3;              // Newton steps of the arc-length inversion, fixed-count form (sampler)
constexpr int kInvNewtonMax = 40;          // ... and the cap of the adaptive form (arc kernel; 1 step is the rule)

template <class R> struct Num;
template <> struct Num<double> {
  static __device__ __forceinline__ double inf() { return CUDART_INF; }
  static __device__ __forceinline__ double nan() { return CUDART_NAN; }
};
template <> struct Num<float> {
  static __device__ __forceinline__ float inf() { return CUDART_INF_F; }
  static __device__ __forceinline__ float nan() { return CUDART_NAN_F; }
};

template <class R> struct V2 { R x, y; };

template <class R> __device__ __forceinline__ R rsqrt_(R v);
template <> __device__ __forceinline__ double rsqrt_(double v) { return rsqrt(v); }
template <> __device__ __forceinline__ float rsqrt_(float v) { return rsqrtf(v); }

// 1/sqrt(x) for a positive normal x.  FP64: MUFU.RSQ64H seed (2^-22) and one third-order step
// y (1 + e/2 + 3 e^2/8), e = 1 - x y^2 -> full double precision in 5 FP64 instructions, no special-
// case branches (CUDA's rsqrt() costs 16 instructions, 7 of them FP64: profiles/r1b_*.lines.txt).
template <class R> __device__ __forceinline__ R fast_rsqrt(R x);
template <> __device__ __forceinline__ float fast_rsqrt<float>(float x) { return rsqrtf(x); }
template <> __device__ __forceinline__ double fast_rsqrt<double>(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  const double p = fma(0.375, e, 0.5);
  return fma(y * e, p, y);
}

// 1/sqrt(x) to ~1e-13 relative: the same seed and ONE second-order step y (1 + e/2), error 3/8 e^2 <= 8.5e-14
// (|e| <= 2^-21).  4 FP64 instructions and no constant to materialise (the 0.5 is an immediate operand); for
// the iterates of the arc-length inversion (a root's error is the quadrature's error over the speed: 1e-13 h), not
// for the table itself -- the error is one-sided (always low), so summed over a path it biases the length by ~3e-14,
// which the timing stage's difference quotients amplify past 1e-9 -- and not for anything a sign decision hangs on.
template <class R> __device__ __forceinline__ R rsqrt_n1(R x);
template <> __device__ __forceinline__ float rsqrt_n1<float>(float x) { return rsqrtf(x); }
template <> __device__ __forceinline__ double rsqrt_n1<double>(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x * y, y, 1.0);
  return fma(y * e, 0.5, y);
}
// Smallest addend that keeps |B'|^2 a positive normal number (a cusp gives exactly 0): added inside the FMA that
// forms dx^2 + dy^2, it replaces the compare-and-select guard around the reciprocal square root.
template <class R> __device__ __forceinline__ R speed_floor();
template <> __device__ __forceinline__ double speed_floor<double>() { return 1e-300; }
template <> __device__ __forceinline__ float speed_floor<float>() { return 1e-30f; }

// 1/x for a normal x.  FP64: MUFU.RCP64H seed and two Newton steps (4 FMAs) instead of the ~20-instruction
// IEEE division sequence; relative error <= ~1 ulp.  Used where the result feeds a 1e-9 tolerance, never in the
// bit-exact Seidel kernel.
template <class R> __device__ __forceinline__ R fast_rcp(R x);
template <> __device__ __forceinline__ float fast_rcp<float>(float x) { return __frcp_rn(x); }
template <> __device__ __forceinline__ double fast_rcp<double>(double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
}
// sqrt(x) for x >= 0 through fast_rsqrt
__device__ __forceinline__ double fast_sqrt(double x) { return x > 0 ? x * fast_rsqrt<double>(x) : 0.0; }

// Curvature-aware velocity cap of one gridpoint (scb_batch_in.omega_max / acc_lat_max): intersects [lo, hi] with
// [-c, c], c = min(omega_max / |kappa|, sqrt(acc_lat_max / |kappa|)).
__device__ __forceinline__ void curvature_cap(double kappa, double omega_max, double acc_lat_max, double& lo, double& hi) {
  const double k = fabs(kappa);
  if (k > 0) {
    const double rk = fast_rcp<double>(k);
    double c = CUDART_INF;
    if (omega_max > 0) c = omega_max * rk;
    if (acc_lat_max > 0) c = fmin(c, fast_sqrt(acc_lat_max * rk));
    hi = fmin(hi, c); lo = fmax(lo, -c);
  }
}

// Kernel-argument bundles (plain pointers; R-typed buffers passed as void* and cast inside).
struct PathArgs {
  int P, W, G, PB;                    // PB (paths per CTA) and table_in_smem are filled in by launch_path
  const void *wx, *wy, *start_angle;  // [P][W], [P][W], [P] or null
  const void *vel_start, *vel_end;    // [P] or null: end slopes of the Hermite path q(s)
  const void *edges; int n_edges;     // [n_edges][4] obstacle edges (x0, y0, x1, y1) or null: tangent clipping
  const void *coll_edges; int n_coll_edges, coll_chords;  // quintic optimiser: collision term inside the loop (Lau09 IV-B)
  void *ctrl_x, *ctrl_y;              // [P][(W-1)*(d+1)] or null
  void *table;                        // [P][(W-1)*101]; may be null only when the table fits in shared memory
  void *L, *cost;                     // [P]
  void *px, *py, *kappa;              // [G][P] or null: curve point / curvature at arc length q(s_i)
  float *px32, *py32;                 // [G][P] or null: the same curve points as float records (scb_batch_out.record_f32)
  int32_t* status;                    // [P]  written: SCB_ST_OK or SCB_ST_ERR_BAD_PATH
  int table_in_smem;
  scb_opt opt;
};

struct ToppraArgs {
  int P, G;
  const void *L;                       // [P] path length
  const void *vel_lo, *vel_hi;         // [P]
  const void *vel_lo_grid, *vel_hi_grid;  // [G][P] or null
  const void *acc_lo, *acc_hi;         // [P]
  const void *vel_start, *vel_end;     // [P] or null
  double x_cap;
  const void *kappa;                   // S [G][P] or null: curvature at the gridpoints, for the two caps below
  double omega_max, acc_lat_max;       // scb_batch_in: 0 = off
  void *K_lo, *K_hi;                   // double [G][P], always (internal state; the FP64 outputs themselves)
  void *K_lo_out, *K_hi_out;           // S [G][P] or null: FP32-mode copies for the caller
  void *x;                             // S [G][P]
  int32_t *status, *fail_stage;        // [P]; status is read (skip if already failed) and written
};

struct ParamArgs {
  int P, G;
  const void *L, *vel_start, *vel_end; // [P]
  const void *x;                       // [G][P]
  const void *kappa;                   // [G][P] or null
  void *t, *v, *a, *omega;             // [G][P] outputs (any may be null); float when rec_f32
  int rec_f32;                         // scb_batch_out.record_f32: the record arrays above are float whatever S
  void *t_buf, *X_buf, *cp_buf;        // [G][P] scratch: knot times, spline unknowns, Thomas c'
  void *T_total;                       // [P]
  int32_t* status;
  int keep_X;                          // 1: X_buf must hold the solved unknowns afterwards (the uniform-dt sampler reads them)
  int param_mode;                      // SCB_PARAM_*
};

struct SampleArgs {
  int P, G, W, deg, S_cap;
  double dt;
  const void *L, *vel_start, *vel_end;
  const void *t_buf, *X_buf;           // [G][P]
  const void *x;                       // [G][P] sdot^2 (const-accel sampling)
  int param_mode;                      // SCB_PARAM_*
  const void *ctrl_x, *ctrl_y, *table;
  int32_t* n_samples;                  // [P]
  void *s_time, *s_pos, *s_vel, *s_acc, *s_px, *s_py, *s_omega;  // [S_cap][P]; float when rec_f32
  int rec_f32;
  const int32_t* status;
  int timing, geometry;                // timing off: s_pos / s_vel / n_samples are inputs
  int nudge;                           // geometry-only: repair s_pos in place first (sc:902-912)
};

struct ObstArgs {                      // obst_kernel.cu: per-gridpoint velocity box under the obstacle-distance limit
  int P, G, E;
  const void *px, *py;                 // [G][P] curve points of the gridpoints
  const void* edges;                   // [E][4]
  const void *vel_lo, *vel_hi;         // [P] ...
  const void *vel_lo_grid, *vel_hi_grid;  // ... or [G][P] (then these are the box)
  double a_brake, t_react;
  void *out_lo, *out_hi;               // [G][P]
};
template <class R> void launch_obst_cap(const ObstArgs& a, cudaStream_t st);

struct CollideArgs {                   // obst_kernel.cu: collision verdict of the finished spline
  int P, W, E, M;                      // problems, waypoints, obstacle edges, chords per segment
  const void *ctrl_x, *ctrl_y;         // [P][(W-1)(d+1)]
  const void* edges;                   // [E][4]
  int32_t* status;                     // [P]: OK -> SCB_ST_ERR_COLLISION on a hit
  void* cost;                          // [P] or null: +inf on a hit
};
template <class R> void launch_collide(const CollideArgs& a, int degree, cudaStream_t st);

struct FmtArgs {                       // fmt_kernel.cu; float whatever the ctx precision (the reference's planner is float)
  int Q, n, E, max_path;               // queries, shared sample nodes, obstacle edges, capacity of a path row
  const float *nx, *ny;                // [n]
  const float* edges;                  // [E][4]
  const float *ix, *iy, *gx, *gy;      // [Q] x_init, x_goal
  float rn;
  int edges_in_smem;                   // set by launch_fmt
  int32_t *status, *path_len, *closed; // [Q]; closed may be null
  float *path_x, *path_y;              // [Q][max_path] or null
  float* cost;                         // [Q]
};
size_t fmt_smem_bytes(int n);
cudaError_t launch_fmt(const FmtArgs& a, cudaStream_t st);

struct PpolyArgs {                     // toppra_kernel.cu: toppra on a caller-supplied piecewise-cubic path, FP64
  int P, n_seg, dof, G, interpolation;
  const double *coeffs, *breaks;       // [P][n_seg][4][dof], [n_seg + 1]
  const double *vel_lo, *vel_hi, *acc_lo, *acc_hi;  // [P][dof]
  const double *vel_lo_grid, *vel_hi_grid;          // [G][dof][P] or null: velocity limits that vary along the path
  double sd_start, sd_end, x_cap;
  int32_t *status, *fail_stage, *feas_status;       // [P]
  double *K_lo, *K_hi, *x, *feas_lo, *feas_hi;      // [G][P]; K_lo / K_hi are required (state of the forward pass)
};
void launch_toppra_ppoly(const PpolyArgs& a, cudaStream_t st);

template <class R> void launch_path(const PathArgs& a, int spline_mode, cudaStream_t st, cudaEvent_t between = nullptr);
template <class R> void launch_toppra(const ToppraArgs& a, cudaStream_t st);
template <class R> void launch_toppra_fast(const ToppraArgs& a, cudaStream_t st);
template <class R> void launch_param(const ParamArgs& a, cudaStream_t st);
template <class R> void launch_sample(const SampleArgs& a, cudaStream_t st);
// lane-per-waypoint variant of the path kernel (path_lane_kernel.cu); false = more than 32 waypoints
bool path_lane_shape(int W, int& PB, size_t& smem, int spline_mode);
template <class S> void launch_path_lane(const PathArgs& a, int spline_mode, int PB, size_t smem, cudaStream_t st);
// arc stage (arc_kernel.cu): tables, length, geometry at the gridpoints; paths per CTA (0 = path too long)
int arc_paths_per_cta(int W, int spline_mode, int real_size, size_t& smem, bool* table_in_smem = nullptr);
template <class R> void launch_arc(const PathArgs& a, int spline_mode, cudaStream_t st);
bool path_launch_shape(int W, int spline_mode, int real_size, int& PB, int& threads, bool& table_in_smem, size_t& smem);

}  // namespace scb
