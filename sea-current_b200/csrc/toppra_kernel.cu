// toppra on the GPU: gridpoint constraint assembly + backward controllable sets + forward
// greedy pass, fused, one thread per planning problem (sm_100a).
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
// Nothing of the reference's 176 B/gridpoint stage matrices is materialised: the six LP rows of
// a stage are rebuilt in registers from (L, v0, v1, limits).  Gridpoint arrays are [G][P] so a
// warp's 32 problems read and write 256 contiguous bytes per gridpoint.
#include "scb_common.cuh"

namespace scb {

template <class R> struct Tol;
// seidel-internal.hpp:54-57
template <> struct Tol<double> {
  static constexpr double abs_tol = 1e-13, rel_tol = 1e-10;
  static constexpr double tie = 1e-9, maxsd = 1e8;
};
// FP32 mode: the reference's tolerances sit below float epsilon; scaled to the same number of
// ulps (DESIGN.md "FP32 mode").
template <> struct Tol<float> {
  static constexpr float abs_tol = 1e-6f, rel_tol = 1e-5f;
  static constexpr float tie = 1e-6f, maxsd = 1e8f;
};

// ---- solve_lp1d, streaming form (seidel-internal.hpp:73-133): rows are fed one at a time.
template <class R> struct Lp1d {
  R cur_min, cur_max;
  bool infeasible;
  __device__ __forceinline__ void init() { cur_min = -Num<R>::inf(); cur_max = Num<R>::inf(); infeasible = false; }
  __device__ __forceinline__ void row(R a, R b) {
    if (fabs(a) < Tol<R>::abs_tol) {
      if (b > Tol<R>::abs_tol) infeasible = true;
      return;
    }
    if (a > 0) {
      if (a * cur_max + b > 0) cur_max = fmin(-b / a, cur_max);
    } else if (a < 0 && a * cur_min + b > 0) {
      cur_min = fmax(-b / a, cur_min);
    }
  }
  // returns feasibility; optimum of  max v0*x
  __device__ __forceinline__ bool finish(R v0, R& opt) const {
    if (infeasible) return false;
    if (cur_min - cur_max > fmax(fabs(cur_min), fabs(cur_max)) * Tol<R>::rel_tol || cur_min == Num<R>::inf() ||
        cur_max == -Num<R>::inf())
      return false;
    if (v0 == 0) {
      if (cur_min != -Num<R>::inf()) {
        opt = (cur_max != Num<R>::inf()) ? (cur_max / 2 + cur_min / 2) : cur_min;
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
template <class R> __device__ __forceinline__ R value3(R r0, R r1, R r2, R u, R x) {
  R res = r2;
  res += (r0 == 0 && isinf(u)) ? (R)0 : r0 * u;
  res += (r1 == 0 && isinf(x)) ? (R)0 : r1 * x;
  return res;
}

// One stage LP:  max v.(u,x)  s.t.  A[r].(u,x,1) <= 0 (6 rows),  u free,  lo1 <= x <= hi1.
// Faithful to seidel.cpp:50-198 for low[0] = -inf, high[0] = +inf (what toppra passes:
// solver.cpp:41-46 leaves the u column unbounded).
template <class R>
__device__ __forceinline__ bool solve_lp2d6(R v0, R v1, const R (&A)[6][3], R lo1, R hi1, R& u_opt, R& x_opt) {
  const R INF = Num<R>::inf();
  // ---- incoherent / equal-within-tolerance x box: 1-D LP in u  (seidel.cpp:73-106, i = 1)
  if (lo1 > hi1) {
    if (lo1 - hi1 > fmax(fabs(lo1), fabs(hi1)) * Tol<R>::rel_tol || lo1 == INF || hi1 == -INF) return false;
    const R xf = (lo1 + hi1) / 2;
    Lp1d<R> lp; lp.init();
#pragma unroll
    for (int r = 0; r < 6; ++r) lp.row(A[r][0], A[r][2] + xf * A[r][1]);
    lp.row((R)-1, -INF);  // {-1, low[0]}
    lp.row((R)1, -INF);   // { 1, -high[0]}
    R uo;
    if (!lp.finish(v0, uo)) return false;
    u_opt = uo; x_opt = xf;
    return true;
  }
  R cu = v0 > 0 ? INF : -INF;      // cur_optvar[0]
  R cx = v1 > 0 ? hi1 : lo1;       // cur_optvar[1]
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    if (value3(A[i][0], A[i][1], A[i][2], cu, cx) < Tol<R>::abs_tol) continue;
    // project onto the line of row i
    const R sq = A[i][0] * A[i][0] + A[i][1] * A[i][1];
    const R f = -A[i][2] / sq;
    const R zx0 = A[i][0] * f, zx1 = A[i][1] * f;   // zero_prj
    const R d0 = -A[i][1], d1 = A[i][0];            // d_tan
    const R v1d = v0 * d0 + v1 * d1;
    Lp1d<R> lp; lp.init();
#pragma unroll
    for (int j = 0; j < 6; ++j)
      if (j < i) lp.row(A[j][0] * d0 + A[j][1] * d1, value3(A[j][0], A[j][1], A[j][2], zx0, zx1));
    // the four box rows {-1,0,low0} {1,0,-high0} {0,-1,low1} {0,1,-high1}
    lp.row(-d0, value3((R)-1, (R)0, -INF, zx0, zx1));
    lp.row(d0, value3((R)1, (R)0, -INF, zx0, zx1));
    lp.row(-d1, value3((R)0, (R)-1, lo1, zx0, zx1));
    lp.row(d1, value3((R)0, (R)1, -hi1, zx0, zx1));
    R t;
    if (!lp.finish(v1d, t)) return false;
    cu = zx0 + t * d0;
    cx = zx1 + t * d1;
  }
  u_opt = cu; x_opt = cx;
  return true;
}

// Per-problem constants of the Hermite path and its limits.
template <class R> struct PathK {
  R c0, c1, c2;       // q(s) = c0 s^3 + c1 s^2 + c2 s   (c3 = pos_start = 0)
  R acc_lo, acc_hi, vlo, vhi;
  R step;             // gridpoint spacing of LinSpaced(G, 0, 1)
  int N;
  __device__ __forceinline__ R s_at(int i) const { return i == N ? (R)1 : (R)i * step; }
  // q'(s), q''(s) in the reference's power-form summation order (piecewise_poly_path.cpp:215-226)
  __device__ __forceinline__ R qd(R s) const { return ((R)3 * c0) * (s * s) + ((R)2 * c1) * s + c2; }
  __device__ __forceinline__ R qdd(R s) const { return ((R)6 * c0) * s + (R)2 * c1; }
};

// x-bounds from the velocity box (linear_joint_velocity.cpp:24-61).  Returns false for the
// reference's "BoxConstraint is infeasible" throw.
template <class R> __device__ __forceinline__ bool vel_xbound(R v, R lo, R hi, R& xlo, R& xhi) {
  R sdmin = -Tol<R>::maxsd, sdmax = Tol<R>::maxsd;
  if (v > 0) {
    const R vinv = (R)1 / v;
    sdmax = fmin(hi * vinv, sdmax); sdmin = fmax(lo * vinv, sdmin);
  } else if (v < 0) {
    const R vinv = (R)1 / v;
    sdmax = fmin(lo * vinv, sdmax); sdmin = fmax(hi * vinv, sdmin);
  } else if (hi < 0 || lo > 0) {
    return false;
  }
  xlo = sdmin > 0 ? sdmin * sdmin : (R)0;
  xhi = sdmax * sdmax;
  return true;
}

template <class R>
__global__ void __launch_bounds__(64) toppra_kernel(ToppraArgs g) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  R* K_lo = (R*)g.K_lo; R* K_hi = (R*)g.K_hi; R* X = (R*)g.x;
  const R INF = Num<R>::inf();
  if (g.status[p] != SCB_ST_OK) {  // spline stage already failed this problem
    for (int i = 0; i < G; ++i) { K_lo[(size_t)i * P + p] = 0; K_hi[(size_t)i * P + p] = 0; X[(size_t)i * P + p] = 0; }
    g.fail_stage[p] = -1;
    return;
  }
  PathK<R> pk;
  {
    const R L = ((const R*)g.L)[p];
    const R vs = g.vel_start ? ((const R*)g.vel_start)[p] : (R)0;
    const R ve = g.vel_end ? ((const R*)g.vel_end)[p] : (R)0;
    // piecewise_poly_path.cpp:364-375 with dt = 1, positions {0, L}
    pk.c2 = vs;
    pk.c0 = (ve - (R)2 * L + vs);
    pk.c1 = (ve - vs - (R)3 * pk.c0) / (R)2;
    pk.acc_lo = ((const R*)g.acc_lo)[p]; pk.acc_hi = ((const R*)g.acc_hi)[p];
    pk.vlo = g.vel_lo ? ((const R*)g.vel_lo)[p] : (R)0; pk.vhi = g.vel_hi ? ((const R*)g.vel_hi)[p] : (R)0;
    pk.N = N; pk.step = (R)1 / (R)N;
  }
  const R* vlg = (const R*)g.vel_lo_grid; const R* vhg = (const R*)g.vel_hi_grid;
  const bool grid_lim = vlg && vhg;
  const R x_cap = (R)g.x_cap;
  int status = SCB_ST_OK, fail_stage = -1;

  // stage rows 2..5 and x box of stage i (rows 0,1 are the x_next coupling, filled per LP)
  auto stage = [&](int i, R (&A)[6][3], R& xlo, R& xhi, R& delta) -> bool {
    const R s = pk.s_at(i), sn = pk.s_at(i + 1 > N ? N : i + 1);
    delta = sn - s;
    const R a = pk.qd(s), b = pk.qdd(s);
    R a2, b2;
    if (i < N) { const R an = pk.qd(sn), bn = pk.qdd(sn); a2 = an + (R)2 * delta * bn; b2 = bn; }
    else { a2 = a; b2 = b; }
    A[2][0] = a;   A[2][1] = b;   A[2][2] = -pk.acc_hi;
    A[3][0] = -a;  A[3][1] = -b;  A[3][2] = pk.acc_lo;
    A[4][0] = a2;  A[4][1] = b2;  A[4][2] = -pk.acc_hi;
    A[5][0] = -a2; A[5][1] = -b2; A[5][2] = pk.acc_lo;
    R lo = pk.vlo, hi = pk.vhi;
    if (grid_lim) { lo = vlg[(size_t)i * P + p]; hi = vhg[(size_t)i * P + p]; }
    return vel_xbound(a, lo, hi, xlo, xhi);
  };
  auto couple = [&](R (&A)[6][3], R delta, R nlo, R nhi) {
    if (nlo <= -INF) { A[0][0] = 0; A[0][1] = 0; A[0][2] = -1; }
    else { A[0][0] = (R)-2 * delta; A[0][1] = -1; A[0][2] = nlo; }
    if (nhi >= INF) { A[1][0] = 0; A[1][1] = 0; A[1][2] = -1; }
    else { A[1][0] = (R)2 * delta; A[1][1] = 1; A[1][2] = -nhi; }
  };

  // the reference evaluates every gridpoint's velocity box before the passes
  // (solver.cpp:26-63) and throws on the first bad one: same verdict here.
  {
    R xl, xh;
    for (int i = 0; i < G && status == SCB_ST_OK; ++i) {
      R lo = pk.vlo, hi = pk.vhi;
      if (grid_lim) { lo = vlg[(size_t)i * P + p]; hi = vhg[(size_t)i * P + p]; }
      if (!vel_xbound(pk.qd(pk.s_at(i)), lo, hi, xl, xh)) status = SCB_ST_ERR_BAD_LIMITS;
    }
  }

  // ---- backward pass: controllable sets (algorithm.cpp:28-68)
  R nlo = 0, nhi = 0;  // K[N] = (sd_end^2, sd_end^2) with sd_end = 0 (sc:1225)
  K_lo[(size_t)N * P + p] = nlo; K_hi[(size_t)N * P + p] = nhi;
  int i = N - 1;
  if (status == SCB_ST_OK) {
    for (; i >= 0; --i) {
      R A[6][3], xlo, xhi, delta;
      stage(i, A, xlo, xhi, delta);
      couple(A, delta, nlo, nhi);
      const R lo1 = fmax(xlo, (R)0), hi1 = fmin(xhi, x_cap);
      R u, xu, xl;
      // g_upper = (1e-9, -1): v = -g
      if (!solve_lp2d6<R>(-Tol<R>::tie, (R)1, A, lo1, hi1, u, xu)) { status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i; break; }
      K_hi[(size_t)i * P + p] = xu;  // written before the lower LP, as algorithm.cpp:52 does
      if (!solve_lp2d6<R>(Tol<R>::tie, (R)-1, A, lo1, hi1, u, xl)) {
        status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = i;
        K_lo[(size_t)i * P + p] = 0; --i;
        break;
      }
      nhi = xu; nlo = fmax((R)0, xl);
      K_lo[(size_t)i * P + p] = nlo;
    }
  }
  for (; i >= 0; --i) { K_lo[(size_t)i * P + p] = 0; K_hi[(size_t)i * P + p] = 0; }

  // ---- forward pass (toppra.cpp:15-46)
  R xi = 0;  // parametrization(0) = vel_start argument of computePathParametrization = 0
  X[p] = xi;
  i = 0;
  if (status == SCB_ST_OK) {
    for (; i < N; ++i) {
      R A[6][3], xlo, xhi, delta;
      stage(i, A, xlo, xhi, delta);
      const R klo = K_lo[(size_t)(i + 1) * P + p], khi = K_hi[(size_t)(i + 1) * P + p];
      couple(A, delta, klo, khi);
      const R lo1 = fmax(xlo, xi), hi1 = fmin(xhi, xi);
      R u, xx;
      // g = (-2 delta, -1): v = -g
      if (!solve_lp2d6<R>((R)2 * delta, (R)1, A, lo1, hi1, u, xx)) { status = SCB_ST_ERR_FAIL_FORWARD_PASS; fail_stage = i; break; }
      xi = fmin(khi, fmax(klo, xi + (R)2 * delta * u));
      X[(size_t)(i + 1) * P + p] = xi;
    }
  }
  for (++i; i <= N; ++i) X[(size_t)i * P + p] = 0;
  g.status[p] = status;
  g.fail_stage[p] = fail_stage;
}

template <class R> void launch_toppra(const ToppraArgs& a, cudaStream_t st) {
  const int block = 64;
  toppra_kernel<R><<<(a.P + block - 1) / block, block, 0, st>>>(a);
}
template void launch_toppra<double>(const ToppraArgs&, cudaStream_t);
template void launch_toppra<float>(const ToppraArgs&, cudaStream_t);

}  // namespace scb
