// toppra on the GPU, reference-exact variant: gridpoint constraint assembly + backward
// controllable sets + forward greedy pass, fused, one thread per planning problem (sm_100a).
// (Measured alternatives, both slower: two lanes per problem — the upper and lower LP of a stage
// diverge, so a warp pays for both anyway; LP rows in shared memory with rolled loops — the
// dependent chain gains load latency.  See DESIGN.md "toppra kernels".)
//
// Replaces (reference, "tp:" = third-party/toppra/cpp/src/toppra/):
//   constraint assembly   tp:constraint/linear_joint_velocity.cpp:24-61,
//                         tp:constraint/linear_joint_acceleration.cpp:22-49,
//                         tp:constraint.cpp:100-162 (collocation -> interpolation)
//   stage matrices        tp:solver.cpp:26-63, tp:solver/seidel.cpp:204-263
//   Seidel LP             tp:solver/seidel.cpp:50-198, 265-347; tp:solver/seidel-internal.hpp:54-133
//   backward / forward    tp:algorithm.cpp:28-68, tp:algorithm/toppra.cpp:15-46
// for the 1-DOF cubic Hermite path q(s), s in [0,1], that gen_vel_prof<1> builds
// (sea_current.hpp:1213-1222; coefficients tp:geometric_path/piecewise_poly_path.cpp:364-375).
//
// Arithmetic is FP64 in both precision modes (toppra's value_type is double and its tolerances,
// 1e-13 / 1e-10 / 1e-9, do not exist in FP32); `S` is only the storage type of the caller's
// arrays.  This file is compiled without FMA contraction: given the same path length the kernel
// reproduces the reference's operation sequence (as restated by the oracle) bit for bit.
// Nothing of the reference's 176 B/gridpoint stage matrices is materialised in HBM.  Gridpoint
// arrays are [G][P]: a warp's 32 problems touch 256 contiguous bytes per gridpoint.
#include "scb_common.cuh"

namespace scb {

namespace tol {  // seidel-internal.hpp:54-57; algorithm.cpp:33-34; linear_joint_velocity.hpp:33
constexpr double abs_tol = 1e-13, rel_tol = 1e-10, tie = 1e-9, maxsd = 1e8;
}
#define kInf CUDART_INF

// ---- solve_lp1d, streaming form (seidel-internal.hpp:73-133): rows are fed one at a time.
struct Lp1d {
  double cur_min, cur_max;
  bool infeasible;
  __device__ __forceinline__ void init() { cur_min = -kInf; cur_max = kInf; infeasible = false; }
  __device__ __forceinline__ void row(double a, double b) {
    if (fabs(a) < tol::abs_tol) {
      if (b > tol::abs_tol) infeasible = true;
      return;
    }
    if (a > 0) {
      if (a * cur_max + b > 0) cur_max = fmin(-b / a, cur_max);
    } else if (a < 0 && a * cur_min + b > 0) {
      cur_min = fmax(-b / a, cur_min);
    }
  }
  // returns feasibility; optimum of  max v0*x
  __device__ __forceinline__ bool finish(double v0, double& opt) const {
    if (infeasible) return false;
    if (cur_min - cur_max > fmax(fabs(cur_min), fabs(cur_max)) * tol::rel_tol || cur_min == kInf || cur_max == -kInf)
      return false;
    if (v0 == 0) {
      if (cur_min != -kInf) {
        opt = (cur_max != kInf) ? (cur_max / 2 + cur_min / 2) : cur_min;
        return true;
      }
      opt = cur_max;
      return true;
    }
    opt = (v0 > 0) ? cur_max : cur_min;
    return true;
  }
};

// seidel.cpp:14-25: row . [u x 1] with 0 * inf := 0
__device__ __forceinline__ double value3(double r0, double r1, double r2, double u, double x) {
  double res = r2;
  res += (r0 == 0 && isinf(u)) ? 0.0 : r0 * u;
  res += (r1 == 0 && isinf(x)) ? 0.0 : r1 * x;
  return res;
}

// One stage LP:  max v.(u,x)  s.t.  A[r].(u,x,1) <= 0 (6 rows),  u free,  lo1 <= x <= hi1.
// Faithful to seidel.cpp:50-198 for low[0] = -inf, high[0] = +inf (what toppra passes:
// solver.cpp:41-46 leaves the u column unbounded).  The two u-box rows the reference projects,
// {-1,0,-inf} and {1,0,-inf}, can never tighten a 1-D bound (their constant is -inf or NaN, and
// every comparison they enter is false), so they are not fed to the 1-D LP.  Rows stay in
// registers and the loops are fully unrolled.
__device__ __forceinline__ bool solve_lp2d6(double v0, double v1, const double (&A)[6][3], double lo1, double hi1,
                                            double& u_opt, double& x_opt) {
  // ---- incoherent / equal-within-tolerance x box: 1-D LP in u  (seidel.cpp:73-106, i = 1)
  if (lo1 > hi1) {
    if (lo1 - hi1 > fmax(fabs(lo1), fabs(hi1)) * tol::rel_tol || lo1 == kInf || hi1 == -kInf) return false;
    const double xf = (lo1 + hi1) / 2;
    Lp1d lp; lp.init();
#pragma unroll
    for (int r = 0; r < 6; ++r) lp.row(A[r][0], A[r][2] + xf * A[r][1]);
    double uo;
    if (!lp.finish(v0, uo)) return false;
    u_opt = uo; x_opt = xf;
    return true;
  }
  double cu = v0 > 0 ? kInf : -kInf;   // cur_optvar[0]
  double cx = v1 > 0 ? hi1 : lo1;      // cur_optvar[1]
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const double r0 = A[i][0], r1 = A[i][1], r2 = A[i][2];
    if (value3(r0, r1, r2, cu, cx) < tol::abs_tol) continue;
    // project onto the line of row i
    const double sq = r0 * r0 + r1 * r1;
    const double f = -r2 / sq;
    const double zx0 = r0 * f, zx1 = r1 * f;   // zero_prj
    const double d0 = -r1, d1 = r0;            // d_tan
    const double v1d = v0 * d0 + v1 * d1;
    Lp1d lp; lp.init();
#pragma unroll
    for (int j = 0; j < 6; ++j)
      if (j < i) lp.row(A[j][0] * d0 + A[j][1] * d1, value3(A[j][0], A[j][1], A[j][2], zx0, zx1));
    // the x-box rows {0,-1,low1} {0,1,-high1}
    lp.row(-d1, value3(0.0, -1.0, lo1, zx0, zx1));
    lp.row(d1, value3(0.0, 1.0, -hi1, zx0, zx1));
    double t;
    if (!lp.finish(v1d, t)) return false;
    cu = zx0 + t * d0;
    cx = zx1 + t * d1;
  }
  u_opt = cu; x_opt = cx;
  return true;
}

// The same LP with NR rows (more degrees of freedom than gen_vel_prof<1>: 2 + 2 dof rows, twice the dof part with the
// interpolation discretisation).  Same operation sequence; six rows go through solve_lp2d6 (registers, fully unrolled),
// more rows keep the matrix in local memory and loop.
template <int NR>
__device__ __noinline__ bool solve_lp2d_n(double v0, double v1, const double (&A)[NR][3], double lo1, double hi1,
                                          double& u_opt, double& x_opt) {
  if (lo1 > hi1) {
    if (lo1 - hi1 > fmax(fabs(lo1), fabs(hi1)) * tol::rel_tol || lo1 == kInf || hi1 == -kInf) return false;
    const double xf = (lo1 + hi1) / 2;
    Lp1d lp; lp.init();
#pragma unroll 1
    for (int r = 0; r < NR; ++r) lp.row(A[r][0], A[r][2] + xf * A[r][1]);
    double uo;
    if (!lp.finish(v0, uo)) return false;
    u_opt = uo; x_opt = xf;
    return true;
  }
  double cu = v0 > 0 ? kInf : -kInf;
  double cx = v1 > 0 ? hi1 : lo1;
#pragma unroll 1
  for (int i = 0; i < NR; ++i) {
    const double r0 = A[i][0], r1 = A[i][1], r2 = A[i][2];
    if (value3(r0, r1, r2, cu, cx) < tol::abs_tol) continue;
    const double sq = r0 * r0 + r1 * r1;
    const double f = -r2 / sq;
    const double zx0 = r0 * f, zx1 = r1 * f;
    const double d0 = -r1, d1 = r0;
    const double v1d = v0 * d0 + v1 * d1;
    Lp1d lp; lp.init();
#pragma unroll 1
    for (int j = 0; j < i; ++j) lp.row(A[j][0] * d0 + A[j][1] * d1, value3(A[j][0], A[j][1], A[j][2], zx0, zx1));
    lp.row(-d1, value3(0.0, -1.0, lo1, zx0, zx1));
    lp.row(d1, value3(0.0, 1.0, -hi1, zx0, zx1));
    double t;
    if (!lp.finish(v1d, t)) return false;
    cu = zx0 + t * d0;
    cx = zx1 + t * d1;
  }
  u_opt = cu; x_opt = cx;
  return true;
}
template <int NR>
__device__ __forceinline__ bool solve_lp2d(double v0, double v1, const double (&A)[NR][3], double lo1, double hi1,
                                           double& u_opt, double& x_opt) {
  if constexpr (NR == 6) return solve_lp2d6(v0, v1, A, lo1, hi1, u_opt, x_opt);
  else return solve_lp2d_n<NR>(v0, v1, A, lo1, hi1, u_opt, x_opt);
}

// Per-problem constants of the Hermite path and its limits.
struct PathK {
  double c0, c1, c2;       // q(s) = c0 s^3 + c1 s^2 + c2 s   (c3 = pos_start = 0)
  double acc_lo, acc_hi, vlo, vhi;
  double step;             // gridpoint spacing of LinSpaced(G, 0, 1)
  int N;
  __device__ __forceinline__ double s_at(int i) const { return i == N ? 1.0 : (double)i * step; }
  // q'(s), q''(s) in the reference's power-form summation order (piecewise_poly_path.cpp:215-226)
  __device__ __forceinline__ double qd(double s) const { return (3.0 * c0) * (s * s) + (2.0 * c1) * s + c2; }
  __device__ __forceinline__ double qdd(double s) const { return (6.0 * c0) * s + 2.0 * c1; }
};

// x-bounds from the velocity box (linear_joint_velocity.cpp:24-61).  Returns false for the
// reference's "BoxConstraint is infeasible" throw.
__device__ __forceinline__ bool vel_xbound(double v, double lo, double hi, double& xlo, double& xhi) {
  double sdmin = -tol::maxsd, sdmax = tol::maxsd;
  if (v > 0) {
    const double vinv = 1.0 / v;
    sdmax = fmin(hi * vinv, sdmax); sdmin = fmax(lo * vinv, sdmin);
  } else if (v < 0) {
    const double vinv = 1.0 / v;
    sdmax = fmin(lo * vinv, sdmax); sdmin = fmax(hi * vinv, sdmin);
  } else if (hi < 0 || lo > 0) {
    return false;
  }
  xlo = sdmin > 0 ? sdmin * sdmin : 0.0;
  xhi = sdmax * sdmax;
  return true;
}

template <class S>
__global__ void __launch_bounds__(64) toppra_kernel(ToppraArgs g) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  double* K_lo = (double*)g.K_lo; double* K_hi = (double*)g.K_hi;  // always double (internal + FP64 output)
  S* oK_lo = (S*)g.K_lo_out; S* oK_hi = (S*)g.K_hi_out;            // FP32 mode copies, or null
  S* X = (S*)g.x;
  auto put_K = [&](int i, double lo, double hi) {
    const size_t o = (size_t)i * P + p;
    K_lo[o] = lo; K_hi[o] = hi;
    if (oK_lo) oK_lo[o] = (S)lo;
    if (oK_hi) oK_hi[o] = (S)hi;
  };
  int status = g.status[p], fail_stage = -1;
  const bool pre_failed = status != SCB_ST_OK;  // the path stage already failed this problem
  PathK pk;
  {
    const double L = pre_failed ? 1.0 : (double)((const S*)g.L)[p];
    const double vs = g.vel_start ? (double)((const S*)g.vel_start)[p] : 0.0;
    const double ve = g.vel_end ? (double)((const S*)g.vel_end)[p] : 0.0;
    // piecewise_poly_path.cpp:364-375 with dt = 1, positions {0, L}
    pk.c2 = vs;
    pk.c0 = (ve - 2.0 * L + vs);
    pk.c1 = (ve - vs - 3.0 * pk.c0) / 2.0;
    pk.acc_lo = (double)((const S*)g.acc_lo)[p]; pk.acc_hi = (double)((const S*)g.acc_hi)[p];
    pk.vlo = g.vel_lo ? (double)((const S*)g.vel_lo)[p] : 0.0; pk.vhi = g.vel_hi ? (double)((const S*)g.vel_hi)[p] : 0.0;
    pk.N = N; pk.step = 1.0 / (double)N;
  }
  const S* vlg = (const S*)g.vel_lo_grid; const S* vhg = (const S*)g.vel_hi_grid;
  const bool grid_lim = vlg && vhg;
  const S* __restrict__ kap = (const S*)g.kappa;
  const bool curv = kap && (g.omega_max > 0 || g.acc_lat_max > 0);
  const double x_cap = g.x_cap;

  // stage rows 2..5 and x box of stage i (rows 0,1 are the x_next coupling, filled per LP)
  auto stage = [&](int i, double (&A)[6][3], double& xlo, double& xhi, double& delta) {
    const double s = pk.s_at(i), sn = pk.s_at(i + 1 > N ? N : i + 1);
    delta = sn - s;
    const double a = pk.qd(s), b = pk.qdd(s);
    double a2, b2;
    if (i < N) { const double an = pk.qd(sn), bn = pk.qdd(sn); a2 = an + 2.0 * delta * bn; b2 = bn; }
    else { a2 = a; b2 = b; }
    A[2][0] = a;   A[2][1] = b;   A[2][2] = -pk.acc_hi;
    A[3][0] = -a;  A[3][1] = -b;  A[3][2] = pk.acc_lo;
    A[4][0] = a2;  A[4][1] = b2;  A[4][2] = -pk.acc_hi;
    A[5][0] = -a2; A[5][1] = -b2; A[5][2] = pk.acc_lo;
    double lo = pk.vlo, hi = pk.vhi;
    if (grid_lim) { lo = (double)vlg[(size_t)i * P + p]; hi = (double)vhg[(size_t)i * P + p]; }
    if (curv) curvature_cap((double)kap[(size_t)i * P + p], g.omega_max, g.acc_lat_max, lo, hi);
    vel_xbound(a, lo, hi, xlo, xhi);
  };
  auto couple = [&](double (&A)[6][3], double delta, double nlo, double nhi) {
    if (nlo <= -kInf) { A[0][0] = 0; A[0][1] = 0; A[0][2] = -1; }
    else { A[0][0] = -2.0 * delta; A[0][1] = -1; A[0][2] = nlo; }
    if (nhi >= kInf) { A[1][0] = 0; A[1][1] = 0; A[1][2] = -1; }
    else { A[1][0] = 2.0 * delta; A[1][1] = 1; A[1][2] = -nhi; }
  };

  // the reference evaluates every gridpoint's velocity box before the passes
  // (solver.cpp:26-63) and throws on the first bad one: same verdict here.
  // A lower acceleration limit above the upper one is rejected when the reference constructs the constraint
  // (LinearJointAcceleration::check, tp:constraint/linear_joint_acceleration.cpp:18-22); a NaN limit, which every
  // comparison of the reference lets through, is reported the same way (SCB_ST_ERR_BAD_LIMITS).
  if (!pre_failed) {
    double xl, xh;
    if (!(pk.acc_lo <= pk.acc_hi)) status = SCB_ST_ERR_BAD_LIMITS;
    for (int i = 0; i < G && status == SCB_ST_OK; ++i) {
      double lo = pk.vlo, hi = pk.vhi;
      if (grid_lim) { lo = (double)vlg[(size_t)i * P + p]; hi = (double)vhg[(size_t)i * P + p]; }
      const bool nums = lo == lo && hi == hi;
      if (curv) curvature_cap((double)kap[(size_t)i * P + p], g.omega_max, g.acc_lat_max, lo, hi);
      if (!vel_xbound(pk.qd(pk.s_at(i)), lo, hi, xl, xh) || !nums) status = SCB_ST_ERR_BAD_LIMITS;
    }
  }

  // ---- backward pass: controllable sets (algorithm.cpp:28-68)
  double nlo = 0, nhi = 0;  // K[N] = (sd_end^2, sd_end^2) with sd_end = 0 (sc:1225)
  put_K(N, nlo, nhi);
  int i = N - 1;
  if (status == SCB_ST_OK) {
    for (; i >= 0; --i) {
      double A[6][3], xlo, xhi, delta;
      stage(i, A, xlo, xhi, delta);
      couple(A, delta, nlo, nhi);
      const double lo1 = fmax(xlo, 0.0), hi1 = fmin(xhi, x_cap);
      double u, xu, xl;
      // g_upper = (1e-9, -1): v = -g
      if (!solve_lp2d6(-tol::tie, 1.0, A, lo1, hi1, u, xu)) { status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i; break; }
      // K_hi is stored before the lower LP runs (algorithm.cpp:52)
      if (!solve_lp2d6(tol::tie, -1.0, A, lo1, hi1, u, xl)) {
        status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i;
        put_K(i, 0.0, xu); --i;
        break;
      }
      nhi = xu; nlo = fmax(0.0, xl);
      put_K(i, nlo, nhi);
    }
  }
  for (; i >= 0; --i) put_K(i, 0.0, 0.0);

  // ---- forward pass (toppra.cpp:15-46)
  double xi = 0;  // parametrization(0) = vel_start argument of computePathParametrization = 0
  X[p] = (S)xi;
  i = 0;
  if (status == SCB_ST_OK) {
    for (; i < N; ++i) {
      double A[6][3], xlo, xhi, delta;
      stage(i, A, xlo, xhi, delta);
      const double klo = K_lo[(size_t)(i + 1) * P + p], khi = K_hi[(size_t)(i + 1) * P + p];
      couple(A, delta, klo, khi);
      const double lo1 = fmax(xlo, xi), hi1 = fmin(xhi, xi);
      double u, xx;
      // g = (-2 delta, -1): v = -g
      if (!solve_lp2d6(2.0 * delta, 1.0, A, lo1, hi1, u, xx)) { status = SCB_ST_ERR_FAIL_FORWARD_PASS; fail_stage = i; break; }
      xi = fmin(khi, fmax(klo, xi + 2.0 * delta * u));
      X[(size_t)(i + 1) * P + p] = (S)xi;
    }
  }
  for (++i; i <= N; ++i) X[(size_t)i * P + p] = (S)0;
  g.status[p] = status;
  g.fail_stage[p] = fail_stage;
}

// ---- toppra on a caller-supplied piecewise-cubic path with one or two degrees of freedom ------------------------------
// (scb_toppra_ppoly).  The stage LP of toppra has 2 + 2 m rows with collocation and 2 + 4 m with interpolation
// (tp:solver/seidel.cpp:204-263): six rows - the LP solve_lp2d6 above unrolls - for m = 2 with collocation, which is
// the configuration of the reference's own known-answer test (cpp/tests/test_algorithm.cpp:64-165: 2-DOF, three cubic
// segments, N = 50, Seidel), and for m = 1 with interpolation, gen_vel_prof's.  m = 1 with collocation has four rows
// and is padded with the always-satisfied row (0, 0, -1) the reference itself uses for an absent coupling row
// (seidel.cpp:298-303): such a row is skipped by the incremental loop and projects to a constant <= 0 in every 1-D
// LP, so the optimum is that of the four-row problem bit for bit.  Besides the controllable sets and the forward pass
// the kernel computes the feasible sets (computeFeasibleSets, tp:algorithm.cpp:70-103).  One thread per problem.
struct PpolyPath {
  const double* c;       // this problem's [nseg][4][dof]; row 0 multiplies ds^3 (piecewise_poly_path.cpp:215-226)
  const double* breaks;  // [nseg + 1]
  int nseg, dof;
  // findSegmentIndex (piecewise_poly_path.cpp:240-252): upper_bound - 1, clamped
  __device__ __forceinline__ int segment(double pos) const {
    int idx = -1;
    for (int k = 0; k <= nseg; ++k) if (breaks[k] <= pos) idx = k;  // last break <= pos  ==  upper_bound - 1
    return idx < 0 ? 0 : (idx > nseg - 1 ? nseg - 1 : idx);
  }
  // first and second derivative of dof k at pos, in the reference's order of operations: coefficients differentiated
  // row by row (computeDerivativesCoefficients, piecewise_poly_path.cpp:190-198), power form summed from the highest
  // power down starting at 0 (eval_single; pow(ds, 2) == ds * ds, pow(ds, 1) == ds, pow(ds, 0) == 1 exactly)
  __device__ __forceinline__ void d12(double pos, int k, double& q1, double& q2) const {
    const int seg = segment(pos);
    const double ds = pos - breaks[seg];
    const double* cs = c + ((size_t)seg * 4) * dof + k;
    const double c0 = cs[0], c1 = cs[dof], c2 = cs[2 * dof];
    const double d1 = c0 * 3.0, d2 = c1 * 2.0, d3 = c2 * 1.0;
    q1 = ((0.0 + d1 * (ds * ds)) + d2 * ds) + d3;
    const double e2 = d1 * 2.0, e3 = d2 * 1.0;
    q2 = (0.0 + e2 * ds) + e3;
  }
};

// M degrees of freedom; INTERP: the interpolation discretisation (every acceleration row a second time, at gridpoint
// i + 1 seen from i).  Rows: 2 coupling rows + 2 M (+ 2 M).  One degree of freedom with collocation pads its four rows
// to six so that every six-row problem runs the register-resident solver.
template <int M, bool INTERP>
__global__ void __launch_bounds__(64) toppra_ppoly_kernel(PpolyArgs g) {
  constexpr int NRraw = 2 + 2 * M * (INTERP ? 2 : 1), NR = NRraw < 6 ? 6 : NRraw;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  constexpr int m = M;
  PpolyPath path{g.coeffs + (size_t)p * g.n_seg * 4 * m, g.breaks, g.n_seg, m};
  const double s0 = g.breaks[0], s1 = g.breaks[g.n_seg];
  // Eigen's LinSpaced(G, s0, s1): low + i * step, the last point exactly high
  const double step = (s1 - s0) / (double)N;
  auto s_at = [&](int i) { return i >= N ? s1 : s0 + (double)i * step; };
  double vlo[M], vhi[M], alo[M], ahi[M];
  const bool vgrid = g.vel_lo_grid && g.vel_hi_grid;  // [G][M][P]: a velocity limit that varies along the path
#pragma unroll
  for (int k = 0; k < m; ++k) {
    vlo[k] = vgrid ? 0.0 : g.vel_lo[(size_t)p * m + k]; vhi[k] = vgrid ? 0.0 : g.vel_hi[(size_t)p * m + k];
    alo[k] = g.acc_lo[(size_t)p * m + k]; ahi[k] = g.acc_hi[(size_t)p * m + k];
  }
  int status = SCB_ST_OK, fail_stage = -1;
  // x-box of gridpoint i from the velocity limits (linear_joint_velocity.cpp:24-61, all dofs)
  auto xbox = [&](int i, double& xlo, double& xhi) -> bool {
    const double s = s_at(i);
    double sdmin = -tol::maxsd, sdmax = tol::maxsd;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < m; ++k) {
      double v, vv;
      path.d12(s, k, v, vv);
      const double lo = vgrid ? g.vel_lo_grid[((size_t)i * m + k) * P + p] : vlo[k];
      const double hi = vgrid ? g.vel_hi_grid[((size_t)i * m + k) * P + p] : vhi[k];
      if (lo != lo || hi != hi) ok = false;
      const double vinv = 1.0 / v;
      if (v > 0) { sdmax = fmin(hi * vinv, sdmax); sdmin = fmax(lo * vinv, sdmin); }
      else if (v < 0) { sdmax = fmin(lo * vinv, sdmax); sdmin = fmax(hi * vinv, sdmin); }
      else if (hi < 0 || lo > 0) ok = false;
    }
    xlo = sdmin > 0 ? sdmin * sdmin : 0.0;
    xhi = sdmax * sdmax;
    return ok;
  };
  // rows 2.. of stage i (constraint.cpp:100-162, 185-204; linear_joint_acceleration.cpp:22-49; seidel.cpp:228-239):
  // F = [I; -I] per block, g = [acc_hi; -acc_lo]; interpolation appends the block of gridpoint i + 1
  auto rows = [&](int i, double (&A)[NR][3], double& delta) {
    const double s = s_at(i), sn = s_at(i + 1 > N ? N : i + 1);
    delta = sn - s;
    int r = 2;
    double a[M], b[M];
#pragma unroll
    for (int k = 0; k < m; ++k) path.d12(s, k, a[k], b[k]);
#pragma unroll
    for (int k = 0; k < m; ++k) { A[r][0] = a[k]; A[r][1] = b[k]; A[r][2] = -ahi[k]; ++r; }
#pragma unroll
    for (int k = 0; k < m; ++k) { A[r][0] = -a[k]; A[r][1] = -b[k]; A[r][2] = alo[k]; ++r; }
    if constexpr (INTERP) {
      double a2[M], b2[M];
#pragma unroll
      for (int k = 0; k < m; ++k) {
        if (i < N) { double an, bn; path.d12(sn, k, an, bn); a2[k] = an + 2 * delta * bn; b2[k] = bn; }
        else { a2[k] = a[k]; b2[k] = b[k]; }
      }
#pragma unroll
      for (int k = 0; k < m; ++k) { A[r][0] = a2[k]; A[r][1] = b2[k]; A[r][2] = -ahi[k]; ++r; }
#pragma unroll
      for (int k = 0; k < m; ++k) { A[r][0] = -a2[k]; A[r][1] = -b2[k]; A[r][2] = alo[k]; ++r; }
    }
    for (; r < NR; ++r) { A[r][0] = 0; A[r][1] = 0; A[r][2] = -1; }  // one dof with collocation: padding
  };
  auto couple = [&](double (&A)[NR][3], int i, double delta, double nlo, double nhi) {
    if (i >= N || nlo <= -kInf) { A[0][0] = 0; A[0][1] = 0; A[0][2] = -1; }
    else { A[0][0] = -2.0 * delta; A[0][1] = -1; A[0][2] = nlo; }
    if (i >= N || nhi >= kInf) { A[1][0] = 0; A[1][1] = 0; A[1][2] = -1; }
    else { A[1][0] = 2.0 * delta; A[1][1] = 1; A[1][2] = -nhi; }
  };
  auto put = [&](double* lo, double* hi, int i, double a, double b) {
    if (lo) lo[(size_t)i * P + p] = a;
    if (hi) hi[(size_t)i * P + p] = b;
  };
  for (int i = 0; i < G && status == SCB_ST_OK; ++i) {
    double xl, xh;
    if (!xbox(i, xl, xh)) status = SCB_ST_ERR_BAD_LIMITS;
  }
#pragma unroll
  for (int k = 0; k < m; ++k) if (!(alo[k] <= ahi[k]) || vlo[k] != vlo[k] || vhi[k] != vhi[k]) status = SCB_ST_ERR_BAD_LIMITS;
  // ---- controllable sets (algorithm.cpp:28-68)
  const double xe = g.sd_end * g.sd_end;
  double nlo = xe, nhi = xe;
  put(g.K_lo, g.K_hi, N, nlo, nhi);
  int i = N - 1;
  if (status == SCB_ST_OK) {
    for (; i >= 0; --i) {
      double A[NR][3], xlo, xhi, delta;
      rows(i, A, delta);
      xbox(i, xlo, xhi);
      couple(A, i, delta, nlo, nhi);
      const double lo1 = fmax(xlo, 0.0), hi1 = fmin(xhi, g.x_cap);
      double u, xu, xl;
      if (!solve_lp2d<NR>(-tol::tie, 1.0, A, lo1, hi1, u, xu)) { status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i; break; }
      if (!solve_lp2d<NR>(tol::tie, -1.0, A, lo1, hi1, u, xl)) { status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i; put(g.K_lo, g.K_hi, i, 0.0, xu); --i; break; }
      nhi = xu; nlo = fmax(0.0, xl);
      put(g.K_lo, g.K_hi, i, nlo, nhi);
    }
  }
  for (; i >= 0; --i) put(g.K_lo, g.K_hi, i, 0.0, 0.0);
  // ---- forward pass (toppra.cpp:15-46); parametrization(0) = sd_start, not squared (l.23)
  double xi = g.sd_start;
  if (g.x) g.x[p] = xi;
  i = 0;
  if (status == SCB_ST_OK) {
    for (; i < N; ++i) {
      double A[NR][3], xlo, xhi, delta;
      rows(i, A, delta);
      xbox(i, xlo, xhi);
      const double klo = g.K_lo[(size_t)(i + 1) * P + p], khi = g.K_hi[(size_t)(i + 1) * P + p];
      couple(A, i, delta, klo, khi);
      const double lo1 = fmax(xlo, xi), hi1 = fmin(xhi, xi);
      double u, xx;
      if (!solve_lp2d<NR>(2.0 * delta, 1.0, A, lo1, hi1, u, xx)) { status = SCB_ST_ERR_FAIL_FORWARD_PASS; fail_stage = i; break; }
      xi = fmin(khi, fmax(klo, xi + 2.0 * delta * u));
      if (g.x) g.x[(size_t)(i + 1) * P + p] = xi;
    }
  }
  if (g.x) for (++i; i <= N; ++i) g.x[(size_t)i * P + p] = 0;
  // ---- feasible sets (algorithm.cpp:70-103): every stage against the whole box [0, 100] on both sides
  int feas_status = SCB_ST_OK;
  if (g.feas_lo || g.feas_hi) {
    for (int j = 0; j <= N; ++j) {
      double A[NR][3], xlo, xhi, delta, u, xu = 0, xl = 0;
      bool ok = status != SCB_ST_ERR_BAD_LIMITS && feas_status == SCB_ST_OK;
      if (ok) {
        rows(j, A, delta);
        xbox(j, xlo, xhi);
        couple(A, j, delta, 0.0, 100.0);
        const double lo1 = fmax(xlo, 0.0), hi1 = fmin(xhi, 100.0);
        ok = solve_lp2d<NR>(-tol::tie, 1.0, A, lo1, hi1, u, xu) && solve_lp2d<NR>(tol::tie, -1.0, A, lo1, hi1, u, xl);
        if (!ok) feas_status = SCB_ST_ERR_FAIL_FEASIBLE;
      }
      put(g.feas_lo, g.feas_hi, j, ok ? xl : 0.0, ok ? xu : 0.0);
    }
  }
  g.status[p] = status;
  if (g.fail_stage) g.fail_stage[p] = fail_stage;
  if (g.feas_status) g.feas_status[p] = feas_status;
}

template <int M> static void launch_ppoly_m(const PpolyArgs& a, cudaStream_t st) {
  const int block = 64, grid = (a.P + block - 1) / block;
  if (a.interpolation) toppra_ppoly_kernel<M, true><<<grid, block, 0, st>>>(a);
  else toppra_ppoly_kernel<M, false><<<grid, block, 0, st>>>(a);
}

void launch_toppra_ppoly(const PpolyArgs& a, cudaStream_t st) {
  switch (a.dof) {
    case 1: launch_ppoly_m<1>(a, st); break;
    case 2: launch_ppoly_m<2>(a, st); break;
    case 3: launch_ppoly_m<3>(a, st); break;
    case 4: launch_ppoly_m<4>(a, st); break;
    case 5: launch_ppoly_m<5>(a, st); break;
    case 6: launch_ppoly_m<6>(a, st); break;
    default: break;  // scb_toppra_ppoly validates 1 <= dof <= SCB_MAX_DOF
  }
}

template <class S> void launch_toppra(const ToppraArgs& a, cudaStream_t st) {
  const int block = 64;
  toppra_kernel<S><<<(a.P + block - 1) / block, block, 0, st>>>(a);
}
template void launch_toppra<double>(const ToppraArgs&, cudaStream_t);
template void launch_toppra<float>(const ToppraArgs&, cudaStream_t);

}  // namespace scb
