// toppra on the GPU, closed-form variant (default): the same constraint assembly, backward
// controllable sets and forward greedy pass as toppra_kernel.cu, one thread per planning problem,
// but each stage LP is solved by eliminating u instead of running Seidel's incremental sequence.
//
// For the 1-DOF path every stage has exactly three slabs in the (u, x) plane,
//     l_k <= alpha_k u + beta_k x <= h_k,   k = 0 (x_next coupling: 2 delta, 1, [K_lo, K_hi] of stage i+1)
//                                               1 (acceleration at s_i:      q', q'', [acc_min, acc_max])
//                                               2 (acceleration at s_{i+1}, interpolation scheme)
// plus the box lo1 <= x <= hi1  (rows of tp:solver/seidel.cpp:204-263, tp:constraint.cpp:100-162).
// With every slab oriented so that alpha_k >= 0, a u exists for a given x iff for every pair
// lower_j(x) <= upper_i(x), i.e.  x (beta_i alpha_j - beta_j alpha_i) <= h_i alpha_j - l_j alpha_i:
// three determinants and six quotients give  max x  and  min x  at once (what the reference's two
// Seidel LPs per stage return, tp:algorithm.cpp:28-68).  The forward LP has x fixed, so its optimum
// is the smallest of three upper bounds on u (tp:algorithm/toppra.cpp:15-46).  No quotient has a
// u-coefficient in the denominator, so stages where q' = 0 (both path ends) need no special case:
// slab 0 always has alpha = 2 delta > 0 and pairs with every other slab.
//
// Results equal the Seidel variant's to rounding (the LP optimum is unique); feasibility uses
// the reference's tolerances (relative 1e-10 on incoherent bounds, absolute 1e-13 on rows,
// tp:solver/seidel-internal.hpp:54-57).  tests/test_gpu_parity.py compares the two kernels LP by LP
// at full batch size.  FP64 arithmetic in both precision modes; `S` is the storage type.
#include "scb_common.cuh"

namespace scb {

namespace ftol {
constexpr double abs_tol = 1e-13, rel_tol = 1e-10, maxsd = 1e8;
}

struct Slab { double al, be, lo, hi; };  // lo <= al u + be x <= hi, al >= 0

__device__ __forceinline__ Slab make_slab(double al, double be, double lo, double hi) {
  Slab s;
  if (al < 0) { s.al = -al; s.be = -be; s.lo = -hi; s.hi = -lo; }
  else { s.al = al; s.be = be; s.lo = lo; s.hi = hi; }
  return s;
}

// x-range for which slabs a and b admit a common u; tightens [xl, xh]; false = parallel and disjoint.
// One reciprocal serves both quotients.
__device__ __forceinline__ bool pair_bounds(const Slab& a, const Slab& b, double& xl, double& xh) {
  const double D = a.be * b.al - b.be * a.al;
  const double E1 = a.hi * b.al - b.lo * a.al;   // D x <= E1   (lower_b <= upper_a)
  const double E2 = a.lo * b.al - b.hi * a.al;   // D x >= E2   (lower_a <= upper_b)
  if (D != 0) {
    const double r = fast_rcp<double>(D), q1 = E1 * r, q2 = E2 * r;
    const bool pos = D > 0;
    xh = fmin(xh, pos ? q1 : q2); xl = fmax(xl, pos ? q2 : q1);
    return true;
  }
  return !(E1 < -ftol::abs_tol || E2 > ftol::abs_tol);
}

__device__ __forceinline__ bool incoherent(double lo, double hi) {
  return lo - hi > fmax(ftol::abs_tol, fmax(fabs(lo), fabs(hi)) * ftol::rel_tol);
}

struct FPathK {
  double c0, c1, c2, acc_lo, acc_hi, vlo, vhi, step;
  int N;
  __device__ __forceinline__ double s_at(int i) const { return i == N ? 1.0 : (double)i * step; }
  __device__ __forceinline__ double qd(double s) const { return (3.0 * c0) * (s * s) + (2.0 * c1) * s + c2; }
  __device__ __forceinline__ double qdd(double s) const { return (6.0 * c0) * s + 2.0 * c1; }
};

// linear_joint_velocity.cpp:24-61
__device__ __forceinline__ bool fvel_xbound(double v, double lo, double hi, double& xlo, double& xhi) {
  double sdmin = -ftol::maxsd, sdmax = ftol::maxsd;
  if (v != 0) {
    const double vinv = fast_rcp<double>(v);
    const double p = hi * vinv, m = lo * vinv;
    sdmax = fmin(v > 0 ? p : m, sdmax); sdmin = fmax(v > 0 ? m : p, sdmin);
  } else if (hi < 0 || lo > 0) {
    return false;
  }
  xlo = sdmin > 0 ? sdmin * sdmin : 0.0;
  xhi = sdmax * sdmax;
  return true;
}

constexpr int kTU = 8;  // forward pass: controllable sets of kTU stages are fetched ahead of the dependent chain

template <class S>
__global__ void __launch_bounds__(64) toppra_fast_kernel(ToppraArgs g) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  double* __restrict__ K_lo = (double*)g.K_lo; double* __restrict__ K_hi = (double*)g.K_hi;
  S* __restrict__ oK_lo = (S*)g.K_lo_out; S* __restrict__ oK_hi = (S*)g.K_hi_out;
  S* __restrict__ X = (S*)g.x;
  auto put_K = [&](int i, double lo, double hi) {
    const size_t o = (size_t)i * P + p;
    K_lo[o] = lo; K_hi[o] = hi;
    if (oK_lo) oK_lo[o] = (S)lo;
    if (oK_hi) oK_hi[o] = (S)hi;
  };
  int status = g.status[p], fail_stage = -1;
  const bool pre_failed = status != SCB_ST_OK;
  FPathK pk;
  {
    const double L = pre_failed ? 1.0 : (double)((const S*)g.L)[p];
    const double vs = g.vel_start ? (double)((const S*)g.vel_start)[p] : 0.0;
    const double ve = g.vel_end ? (double)((const S*)g.vel_end)[p] : 0.0;
    pk.c2 = vs; pk.c0 = (ve - 2.0 * L + vs); pk.c1 = (ve - vs - 3.0 * pk.c0) / 2.0;
    pk.acc_lo = (double)((const S*)g.acc_lo)[p]; pk.acc_hi = (double)((const S*)g.acc_hi)[p];
    pk.vlo = g.vel_lo ? (double)((const S*)g.vel_lo)[p] : 0.0; pk.vhi = g.vel_hi ? (double)((const S*)g.vel_hi)[p] : 0.0;
    pk.N = N; pk.step = 1.0 / (double)N;
  }
  const S* __restrict__ vlg = (const S*)g.vel_lo_grid; const S* __restrict__ vhg = (const S*)g.vel_hi_grid;
  const bool grid_lim = vlg && vhg;
  const double x_cap = g.x_cap;

  // acceleration slabs and velocity x-box of stage i; (a, b) = (q', q'') at s_i and (an, bn) at s_{i+1}
  // are carried from stage to stage by the two passes
  auto stage = [&](int i, double a, double b, double an, double bn, double delta, Slab& s1, Slab& s2, double& xlo, double& xhi) {
    s1 = make_slab(a, b, pk.acc_lo, pk.acc_hi);
    s2 = make_slab(an + 2.0 * delta * bn, bn, pk.acc_lo, pk.acc_hi);  // interpolation rows, constraint.cpp:100-162
    double lo = pk.vlo, hi = pk.vhi;
    if (grid_lim) { lo = (double)vlg[(size_t)i * P + p]; hi = (double)vhg[(size_t)i * P + p]; }
    return fvel_xbound(a, lo, hi, xlo, xhi);
  };

  // ---- backward pass: controllable sets.  The velocity box of every gridpoint is validated on
  // the way (the reference throws before its passes, solver.cpp:26-63: same verdict; gridpoint N
  // is checked here, 0..N-1 inside the loop).
  double nlo = 0, nhi = 0;  // K[N] = (0, 0)  (sc:1225)
  put_K(N, nlo, nhi);
  double sn = pk.s_at(N), an = pk.qd(sn), bn = pk.qdd(sn);
  if (!pre_failed) {
    double lo = pk.vlo, hi = pk.vhi, xl, xh;
    if (grid_lim) { lo = (double)vlg[(size_t)N * P + p]; hi = (double)vhg[(size_t)N * P + p]; }
    if (!fvel_xbound(an, lo, hi, xl, xh)) status = SCB_ST_ERR_BAD_LIMITS;
  }
  bool alive = status == SCB_ST_OK;
  int first_fail = -1;
  for (int i = N - 1; i >= 0; --i) {
    const double s = pk.s_at(i), delta = sn - s, a = pk.qd(s), b = pk.qdd(s);
    Slab s1, s2; double xlo, xhi;
    const bool vel_ok = stage(i, a, b, an, bn, delta, s1, s2, xlo, xhi);
    an = a; bn = b; sn = s;
    if (!pre_failed && !vel_ok) { status = SCB_ST_ERR_BAD_LIMITS; alive = false; }
    if (!alive) { put_K(i, 0.0, 0.0); continue; }
    Slab s0; s0.al = 2.0 * delta; s0.be = 1.0; s0.lo = nlo; s0.hi = nhi;
    double xl = fmax(xlo, 0.0), xh = fmin(xhi, x_cap);
    bool ok = !incoherent(xl, xh);
    ok = pair_bounds(s0, s1, xl, xh) && ok;
    ok = pair_bounds(s0, s2, xl, xh) && ok;
    ok = pair_bounds(s1, s2, xl, xh) && ok;
    ok = ok && !incoherent(xl, xh) && s1.lo <= s1.hi + ftol::abs_tol;
    if (!ok) { first_fail = i; alive = false; put_K(i, 0.0, 0.0); continue; }
    nhi = xh; nlo = fmax(0.0, xl);
    put_K(i, nlo, nhi);
  }
  if (status == SCB_ST_ERR_BAD_LIMITS) {
    // a bad velocity box anywhere voids the whole problem, whatever the passes found before it
    for (int i = 0; i < N; ++i) put_K(i, 0.0, 0.0);
  } else if (first_fail >= 0) {
    status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = first_fail;
  }

  // ---- forward pass
  double xi = 0;
  X[p] = (S)xi;
  int done = 0;  // stages completed
  if (status == SCB_ST_OK) {
    double s = pk.s_at(0), a = pk.qd(s), b = pk.qdd(s);
    bool go = true;
    for (int i0 = 0; i0 < N && go; i0 += kTU) {
      double kl[kTU], kh[kTU];
#pragma unroll
      for (int u = 0; u < kTU; ++u) {
        const size_t o = (size_t)min(i0 + u + 1, N) * P + p;
        kl[u] = K_lo[o]; kh[u] = K_hi[o];
      }
#pragma unroll
      for (int u = 0; u < kTU; ++u) {
        const int i = i0 + u;
        if (go && i < N) {
          const double sn2 = pk.s_at(i + 1), delta = sn2 - s, an2 = pk.qd(sn2), bn2 = pk.qdd(sn2);
          Slab s1, s2; double xlo, xhi;
          stage(i, a, b, an2, bn2, delta, s1, s2, xlo, xhi);
          a = an2; b = bn2; s = sn2;
          const double klo = kl[u], khi = kh[u];
          // x is pinned to x_i; a box that misses it by more than the tolerance is infeasible (seidel.cpp:73-76)
          const double lo1 = fmax(xlo, xi), hi1 = fmin(xhi, xi);
          bool ok = !(lo1 > hi1 && (lo1 - hi1 > fmax(fabs(lo1), fabs(hi1)) * ftol::rel_tol));
          const double xf = lo1 > hi1 ? (lo1 + hi1) / 2 : xi;
          // u* = smallest upper bound; slabs without u (alpha = 0) only constrain x
          double uu = (khi - xf) * fast_rcp<double>(2.0 * delta);
          if (s1.al > 0) uu = fmin(uu, (s1.hi - s1.be * xf) * fast_rcp<double>(s1.al));
          if (s2.al > 0) uu = fmin(uu, (s2.hi - s2.be * xf) * fast_rcp<double>(s2.al));
          const double slack = fmax(ftol::abs_tol, fabs(uu) * ftol::rel_tol);
          ok = ok && (klo - xf - 2.0 * delta * uu) <= 2.0 * delta * slack + ftol::abs_tol;
          ok = ok && (s1.lo - s1.be * xf - s1.al * uu) <= s1.al * slack + ftol::abs_tol;
          ok = ok && (s2.lo - s2.be * xf - s2.al * uu) <= s2.al * slack + ftol::abs_tol;
          if (s1.al == 0) ok = ok && (s1.be * xf - s1.hi) <= ftol::abs_tol;
          if (s2.al == 0) ok = ok && (s2.be * xf - s2.hi) <= ftol::abs_tol;
          if (!ok) { status = SCB_ST_ERR_FAIL_FORWARD_PASS; fail_stage = i; go = false; }
          else {
            xi = fmin(khi, fmax(klo, xi + 2.0 * delta * uu));
            X[(size_t)(i + 1) * P + p] = (S)xi;
            done = i + 1;
          }
        }
      }
    }
  }
  for (int i = done + 1; i <= N; ++i) X[(size_t)i * P + p] = (S)0;
  g.status[p] = status;
  g.fail_stage[p] = fail_stage;
}

template <class S> void launch_toppra_fast(const ToppraArgs& a, cudaStream_t st) {
  const int block = 64;
  toppra_fast_kernel<S><<<(a.P + block - 1) / block, block, 0, st>>>(a);
}
template void launch_toppra_fast<double>(const ToppraArgs&, cudaStream_t);
template void launch_toppra_fast<float>(const ToppraArgs&, cudaStream_t);

}  // namespace scb
