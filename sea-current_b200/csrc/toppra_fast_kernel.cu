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
//
// Instruction budget (profiles/r1h_*): the first version of this kernel spent 424 instructions per
// backward stage and 280 per forward stage, ~65 of them FP64 -- the rest were 64-bit moves, selects and
// index arithmetic around a generic slab struct and data-dependent branches.  Here every stage is
// straight-line code: slabs are four scalars oriented with sign tricks, a pair of slabs is resolved
// with selects (a zero determinant selects infinite bounds and a feasibility flag), array positions are
// running pointers, and only what depends on the previous stage's result (two FMAs, two products, a few
// min/max per pair) sits on the serial chain; reciprocals of the determinants do not.
#include <cstdlib>

#include "scb_common.cuh"

namespace scb {

namespace ftol {
constexpr double abs_tol = 1e-13, rel_tol = 1e-10, maxsd = 1e8;
}

namespace {

// min / max of two non-NaN doubles: compare + two selects (fmin / fmax cost a fourth instruction for NaN quieting)
__device__ __forceinline__ double dmin(double a, double b) { return a < b ? a : b; }
__device__ __forceinline__ double dmax(double a, double b) { return a > b ? a : b; }

// lo <= al u + be x <= hi with al >= 0 (a slab with al < 0 is its own mirror image)
struct Slab { double al, be, lo, hi; };

__device__ __forceinline__ Slab orient(double al, double be, double lo, double hi, double nlo, double nhi) {
  // nlo = -lo, nhi = -hi (constants of the problem)
  // the sign of al is read and applied as a bit (integer compare, one XOR into be's high word): no FP64-pipe
  // compare and no negate-then-select; al = -0.0 mirrors a slab that has no u in it, which is the same slab
  const int sgn = __double2hiint(al) & 0x80000000;
  const bool neg = sgn != 0;
  Slab s;
  s.al = fabs(al);
  s.be = __hiloint2double(__double2hiint(be) ^ sgn, __double2loint(be));
  s.lo = neg ? nhi : lo; s.hi = neg ? nlo : hi;
  return s;
}

// x-range for which two slabs admit a common u, given D x <= E1 and D x >= E2 (see header): tightens
// [xl, xh]; returns false only for parallel slabs (D == 0) that are disjoint.  rD = 1/D is formed by the caller
// (off the serial chain).
__device__ __forceinline__ bool pair_apply(double D, double rD, double E1, double E2, double& xl, double& xh) {
  const double q1 = E1 * rD, q2 = E2 * rD;
  const bool pos = __double2hiint(D) >= 0, zero = D == 0;  // pos only matters when D != 0 (integer compare on the sign)
  const double hc = pos ? q1 : q2, lc = pos ? q2 : q1;
  xh = dmin(xh, zero ? CUDART_INF : hc);
  xl = dmax(xl, zero ? -CUDART_INF : lc);
  return !(zero && (E1 < -ftol::abs_tol || E2 > ftol::abs_tol));
}

__device__ __forceinline__ bool incoherent(double lo, double hi) {
  return lo - hi > dmax(ftol::abs_tol, dmax(fabs(lo), fabs(hi)) * ftol::rel_tol);
}

// linear_joint_velocity.cpp:24-61: x-box of a gridpoint from the velocity limits [lo, hi] and q' = v
__device__ __forceinline__ bool vel_xbound(double v, double lo, double hi, double& xlo, double& xhi) {
  const double vinv = fast_rcp<double>(v);  // unused when v == 0
  const double p = hi * vinv, m = lo * vinv;
  const bool pos = v > 0, zero = v == 0;
  const double sdmax = zero ? ftol::maxsd : dmin(pos ? p : m, ftol::maxsd);
  const double sdmin = zero ? -ftol::maxsd : dmax(pos ? m : p, -ftol::maxsd);
  xlo = sdmin > 0 ? sdmin * sdmin : 0.0;
  xhi = sdmax * sdmax;
  return !(zero && (hi < 0 || lo > 0));
}

// A NaN limit must not be dropped by the compare-and-select min / max above (it would read as "unlimited"):
// SCB_ST_ERR_BAD_LIMITS, like a lower acceleration limit above the upper one, which the reference rejects when it
// constructs the constraint (LinearJointAcceleration::check, tp:constraint/linear_joint_acceleration.cpp:18-22).
__device__ __forceinline__ bool both_numbers(double a, double b) { return a == a && b == b; }

constexpr int kTU = 8;  // forward pass: controllable sets of kTU stages are fetched ahead of the dependent chain

}  // namespace

// DB: the forward pass fetches the controllable sets of the NEXT block of stages while the current block's dependent
// chain runs (see the comment there); chosen by the launcher for batches that leave the SMs short of warps
template <class S, bool GRID, bool CURV, bool DB>
__global__ void __launch_bounds__(64) toppra_fast_kernel(ToppraArgs g) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const size_t P = (size_t)g.P;
  const int G = g.G, N = G - 1;
  constexpr bool kNarrow = sizeof(S) != sizeof(double);  // FP32 mode: K is also written in S for the caller
  double* __restrict__ K_lo = (double*)g.K_lo; double* __restrict__ K_hi = (double*)g.K_hi;
  S* __restrict__ oK_lo = kNarrow ? (S*)g.K_lo_out : nullptr; S* __restrict__ oK_hi = kNarrow ? (S*)g.K_hi_out : nullptr;
  S* __restrict__ X = (S*)g.x;
  auto put_K = [&](size_t o, double lo, double hi) {
    K_lo[o] = lo; K_hi[o] = hi;
    if (kNarrow) { if (oK_lo) oK_lo[o] = (S)lo; if (oK_hi) oK_hi[o] = (S)hi; }
  };
  int status = g.status[p], fail_stage = -1;
  const bool pre_failed = status != SCB_ST_OK;
  // q(s) = c0 s^3 + c1 s^2 + c2 s: the 1-DOF Hermite path of gen_vel_prof (piecewise_poly_path.cpp:364-375)
  double c0, c1, c2, Al, Ah, vlo, vhi;
  {
    const double L = pre_failed ? 1.0 : (double)((const S*)g.L)[p];
    const double vs = g.vel_start ? (double)((const S*)g.vel_start)[p] : 0.0;
    const double ve = g.vel_end ? (double)((const S*)g.vel_end)[p] : 0.0;
    c2 = vs; c0 = (ve - 2.0 * L + vs); c1 = (ve - vs - 3.0 * c0) / 2.0;
    Al = (double)((const S*)g.acc_lo)[p]; Ah = (double)((const S*)g.acc_hi)[p];
    vlo = g.vel_lo ? (double)((const S*)g.vel_lo)[p] : 0.0; vhi = g.vel_hi ? (double)((const S*)g.vel_hi)[p] : 0.0;
  }
  if (!pre_failed && (!(Al <= Ah) || (!GRID && !both_numbers(vlo, vhi)))) status = SCB_ST_ERR_BAD_LIMITS;
  // an infinite acceleration limit is "no limit" (the reference's rows with an infinite constant): the slab algebra
  // below multiplies limits by u-coefficients that are exactly 0 at the path ends (0 * inf), so it sees +-1e100
  Al = dmax(Al, -1e100); Ah = dmin(Ah, 1e100);
  const double nAl = -Al, nAh = -Ah, step = 1.0 / (double)N, x_cap = g.x_cap;
  const double c0_3 = 3.0 * c0, c1_2 = 2.0 * c1, c0_6 = 6.0 * c0;
  auto qd = [&](double s) { return c0_3 * (s * s) + c1_2 * s + c2; };
  auto qdd = [&](double s) { return c0_6 * s + c1_2; };
  const S* __restrict__ vlg = (const S*)g.vel_lo_grid; const S* __restrict__ vhg = (const S*)g.vel_hi_grid;
  const S* __restrict__ kap = (const S*)g.kappa;
  const double om_max = g.omega_max, al_max = g.acc_lat_max;

  // ---- backward pass: controllable sets.  The velocity box of every gridpoint is validated on
  // the way (the reference throws before its passes, solver.cpp:26-63: same verdict; gridpoint N
  // is checked here, 0..N-1 inside the loop).
  double nlo = 0, nhi = 0;  // K[N] = (0, 0)  (sc:1225)
  size_t o = (size_t)N * P + p;
  put_K(o, nlo, nhi);
  double sn = 1.0, an = qd(sn), bn = qdd(sn);
  if (!pre_failed) {
    double lo = vlo, hi = vhi, xl, xh;
    if (GRID) { lo = (double)vlg[o]; hi = (double)vhg[o]; }
    const bool nums = !GRID || both_numbers(lo, hi);
    if (CURV) curvature_cap((double)kap[o], om_max, al_max, lo, hi);
    if (!vel_xbound(an, lo, hi, xl, xh) || !nums) status = SCB_ST_ERR_BAD_LIMITS;
  }
  bool alive = status == SCB_ST_OK;
  int first_fail = -1;
  // Per-gridpoint limits / curvatures are fetched three stages ahead (rows i, i-1, i-2 are in registers or in flight
  // while stage i runs): the velocity box is off the stage's dependent chain, but an in-order warp still waited out
  // the DRAM round trip of every stage's own loads - on small batches that was the stage (C2, 4,096 problems with
  // per-gridpoint limits: 535 ns per stage-pass against 270 ns for C5's scalar limits).
  S glo0 = 0, ghi0 = 0, gk0 = 0, glo1 = 0, ghi1 = 0, gk1 = 0, glo2 = 0, ghi2 = 0, gk2 = 0;
  auto fetch = [&](int row, S& l, S& h, S& k) {
    const size_t r = (size_t)(row < 0 ? 0 : row) * P + p;
    if (GRID) { l = vlg[r]; h = vhg[r]; }
    if (CURV) k = kap[r];
  };
  if (GRID || CURV) { fetch(N - 1, glo0, ghi0, gk0); fetch(N - 2, glo1, ghi1, gk1); fetch(N - 3, glo2, ghi2, gk2); }
#pragma unroll 1
  for (int i = N - 1; i >= 0; --i) {
    o -= P;
    const double s = (double)i * step, delta = sn - s, twod = 2.0 * delta, a = qd(s), b = qdd(s);
    double lo = vlo, hi = vhi, xlo, xhi;
    if (GRID) { lo = (double)glo0; hi = (double)ghi0; }
    const double kap_i = CURV ? (double)gk0 : 0.0;
    if (GRID || CURV) {
      glo0 = glo1; ghi0 = ghi1; gk0 = gk1; glo1 = glo2; ghi1 = ghi2; gk1 = gk2;
      fetch(i - 3, glo2, ghi2, gk2);
    }
    const bool nums = !GRID || both_numbers(lo, hi);
    if (CURV) curvature_cap(kap_i, om_max, al_max, lo, hi);
    const bool vel_ok = vel_xbound(a, lo, hi, xlo, xhi) && nums;
    // rows of the stage (tp:constraint.cpp:100-162): acceleration at s_i and, interpolated, at s_{i+1}
    const Slab s1 = orient(a, b, Al, Ah, nAl, nAh);
    const Slab s2 = orient(fma(twod, bn, an), bn, Al, Ah, nAl, nAh);
    an = a; bn = b; sn = s;
    if (!pre_failed && !vel_ok) { status = SCB_ST_ERR_BAD_LIMITS; alive = false; }
    if (!alive) { put_K(o, 0.0, 0.0); continue; }
    // slab 0 is K[i+1] seen through x_next = x + 2 delta u: (al, be, lo, hi) = (2 delta, 1, nlo, nhi)
    const double D01 = fma(-twod, s1.be, s1.al), D02 = fma(-twod, s2.be, s2.al), D12 = fma(s1.be, s2.al, -(s2.be * s1.al));
    const double r01 = fast_rcp<double>(D01), r02 = fast_rcp<double>(D02), r12 = fast_rcp<double>(D12);
    double xl = dmax(xlo, 0.0), xh = dmin(xhi, x_cap);
    bool ok = !incoherent(xl, xh);
    ok = pair_apply(D12, r12, fma(s1.hi, s2.al, -(s2.lo * s1.al)), fma(s1.lo, s2.al, -(s2.hi * s1.al)), xl, xh) && ok;
    ok = pair_apply(D01, r01, fma(nhi, s1.al, -(s1.lo * twod)), fma(nlo, s1.al, -(s1.hi * twod)), xl, xh) && ok;
    ok = pair_apply(D02, r02, fma(nhi, s2.al, -(s2.lo * twod)), fma(nlo, s2.al, -(s2.hi * twod)), xl, xh) && ok;
    ok = ok && !incoherent(xl, xh);
    if (!ok) { first_fail = i; alive = false; put_K(o, 0.0, 0.0); continue; }
    nhi = xh; nlo = dmax(0.0, xl);
    put_K(o, nlo, nhi);
  }
  if (status == SCB_ST_ERR_BAD_LIMITS) {
    // a bad velocity box anywhere voids the whole problem, whatever the passes found before it
    for (int i = 0; i < N; ++i) put_K((size_t)i * P + p, 0.0, 0.0);
  } else if (first_fail >= 0) {
    status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = first_fail;
  }

  // ---- forward pass (tp:algorithm/toppra.cpp:15-46): x pinned to x_i, u* = the smallest upper bound
  double xi = 0;
  X[p] = (S)xi;
  int done = 0;  // stages completed
  if (status == SCB_ST_OK) {
    double s = 0.0, a = qd(s), b = qdd(s);
    bool go = true;
    size_t ox = p;  // position of gridpoint i in the [G][P] arrays
    // The velocity box of gridpoint i can only bite at i = 0: for i >= 1 the pinned x_i was clamped into K_i, and the
    // backward pass built K_i inside that very box (exact selects, no rounding), so max(xlo, x_i) = min(xhi, x_i) = x_i.
    // Stage 0 starts from x_0 = vel_start^2 = 0, which a positive lower velocity limit excludes (seidel.cpp:73-76).
    bool ok0;
    double xf0;
    {
      double lo = vlo, hi = vhi, xlo, xhi;
      if (GRID) { lo = (double)vlg[ox]; hi = (double)vhg[ox]; }
      if (CURV) curvature_cap((double)kap[ox], om_max, al_max, lo, hi);
      vel_xbound(a, lo, hi, xlo, xhi);  // validated by the backward pass
      const double lo1 = dmax(xlo, xi), hi1 = dmin(xhi, xi);
      ok0 = !(lo1 > hi1 && (lo1 - hi1 > dmax(fabs(lo1), fabs(hi1)) * ftol::rel_tol));
      xf0 = lo1 > hi1 ? (lo1 + hi1) / 2 : xi;
    }
    // The controllable sets of a block of kTU stages are fetched ahead of the block's dependent chain.  With the fetch
    // at the top of its own block the first stage waits out a DRAM round trip (9.5 % of the stall samples on C3, long
    // scoreboard: profiles/r2j per-line profile).  DB fetches the NEXT block while the current one runs: measured, that
    // wins where the SMs are short of warps (C5, 8,192 problems: 6.15 -> 5.40 ms; C2: 0.558 -> 0.537) and loses on C3
    // (14 warps per SM already hide most of it; 36 more registers and the copies cost more: 1.23 -> 1.39 ms).
    double kl[kTU], kh[kTU], nl[DB ? kTU : 1], nh[DB ? kTU : 1];
    if constexpr (DB) {
#pragma unroll
      for (int u = 0; u < kTU; ++u) {
        const size_t ok_ = (size_t)min(u + 1, N) * P + p;
        kl[u] = K_lo[ok_]; kh[u] = K_hi[ok_];
      }
    }
#pragma unroll 1
    for (int i0 = 0; i0 < N && go; i0 += kTU) {
#pragma unroll
      for (int u = 0; u < kTU; ++u) {
        if constexpr (DB) {
          const size_t ok_ = (size_t)min(i0 + kTU + u + 1, N) * P + p;
          nl[u] = K_lo[ok_]; nh[u] = K_hi[ok_];
        } else {
          const size_t ok_ = (size_t)min(i0 + u + 1, N) * P + p;
          kl[u] = K_lo[ok_]; kh[u] = K_hi[ok_];
        }
      }
#pragma unroll
      for (int u = 0; u < kTU; ++u) {
        const int i = i0 + u;
        if (go && i < N) {
          const double sn2 = i + 1 == N ? 1.0 : (double)(i + 1) * step, delta = sn2 - s, twod = 2.0 * delta;
          const double an2 = qd(sn2), bn2 = qdd(sn2);
          const Slab s1 = orient(a, b, Al, Ah, nAl, nAh);
          const Slab s2 = orient(fma(twod, bn2, an2), bn2, Al, Ah, nAl, nAh);
          a = an2; b = bn2; s = sn2;
          const double r0 = fast_rcp<double>(twod), r1 = fast_rcp<double>(s1.al), r2 = fast_rcp<double>(s2.al);
          const double klo = kl[u], khi = kh[u];
          bool ok = i == 0 ? ok0 : true;
          const double xf = i == 0 ? xf0 : xi;
          // slabs without u (al = 0) only constrain x
          const double u0 = (khi - xf) * r0;
          const double u1 = s1.al > 0 ? fma(-s1.be, xf, s1.hi) * r1 : CUDART_INF;
          const double u2 = s2.al > 0 ? fma(-s2.be, xf, s2.hi) * r2 : CUDART_INF;
          const double uu = dmin(u0, dmin(u1, u2));
          const double slack = dmax(ftol::abs_tol, fabs(uu) * ftol::rel_tol);
          ok = ok && (klo - xf - twod * uu) <= twod * slack + ftol::abs_tol;
          ok = ok && (s1.lo - s1.be * xf - s1.al * uu) <= s1.al * slack + ftol::abs_tol;
          ok = ok && (s2.lo - s2.be * xf - s2.al * uu) <= s2.al * slack + ftol::abs_tol;
          ok = ok && (s1.al > 0 || (s1.be * xf - s1.hi) <= ftol::abs_tol);
          ok = ok && (s2.al > 0 || (s2.be * xf - s2.hi) <= ftol::abs_tol);
          ox += P;
          if (!ok) { status = SCB_ST_ERR_FAIL_FORWARD_PASS; fail_stage = i; go = false; }
          else {
            xi = dmin(khi, dmax(klo, fma(twod, uu, xi)));
            X[ox] = (S)xi;
            done = i + 1;
          }
        }
      }
      if constexpr (DB) {
#pragma unroll
        for (int u = 0; u < kTU; ++u) { kl[u] = nl[u]; kh[u] = nh[u]; }
      }
    }
  }
  for (int i = done + 1; i <= N; ++i) X[(size_t)i * P + p] = (S)0;
  g.status[p] = status;
  g.fail_stage[p] = fail_stage;
}

// ---------------------------------------------------------------------------------------------------------------
// Warp-specialised variant for batches that leave the SMs short of warps (C2: 4,096 problems).
// There the kernel above is bound by ONE warp's instruction stream per scheduler: ~200 instructions per stage-pass, of
// which about half depend on the previous stage's result (two FMAs, two products and a few min / max per slab pair, the
// feasibility flags, the store).  Here four warps own 32 problems: warp 0 (the chain) runs exactly that half, warps 1-3
// (the producers) compute everything that does not depend on the chain - q', q'', the velocity box, the oriented slabs,
// the determinants and their reciprocals, the whole pair of the two acceleration slabs - for every third stage each, one
// chunk of stages ahead, and hand it over through a double-buffered shared-memory area ([stage][value][lane]:
// conflict-free), one __syncthreads per chunk.
// The arithmetic is the kernel's above operation for operation: results are bit-identical (tests/test_scale_parity.py).
// NG groups of 32 problems per CTA (four warps each), kWsChunk stages per hand-over.
constexpr int kWsVals = 13;

template <class S, bool GRID, bool CURV, int NG, int kWsChunk>
__global__ void __launch_bounds__(NG * 128) toppra_ws_kernel(ToppraArgs g) {
  __shared__ double bufs[NG][2][kWsChunk][kWsVals][32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int grp = wid % NG, warp = wid / NG;   // warp: 0 = the chain, 1..3 = the producers of group grp
  double (*buf)[kWsChunk][kWsVals][32] = bufs[grp];
  const int p_raw = (blockIdx.x * NG + grp) * 32 + lane;
  const bool valid = p_raw < g.P;
  const int p = valid ? p_raw : g.P - 1;   // lanes past the end compute problem P - 1 again and store nothing
  const size_t P = (size_t)g.P;
  const int G = g.G, N = G - 1;
  constexpr bool kNarrow = sizeof(S) != sizeof(double);
  double* __restrict__ K_lo = (double*)g.K_lo; double* __restrict__ K_hi = (double*)g.K_hi;
  S* __restrict__ oK_lo = kNarrow ? (S*)g.K_lo_out : nullptr; S* __restrict__ oK_hi = kNarrow ? (S*)g.K_hi_out : nullptr;
  S* __restrict__ X = (S*)g.x;
  auto put_K = [&](size_t o, double lo, double hi) {
    if (!valid) return;
    K_lo[o] = lo; K_hi[o] = hi;
    if (kNarrow) { if (oK_lo) oK_lo[o] = (S)lo; if (oK_hi) oK_hi[o] = (S)hi; }
  };
  int status = g.status[p], fail_stage = -1;
  const bool pre_failed = status != SCB_ST_OK;
  double c0, c1, c2, Al, Ah, vlo, vhi;
  {
    const double L = pre_failed ? 1.0 : (double)((const S*)g.L)[p];
    const double vs = g.vel_start ? (double)((const S*)g.vel_start)[p] : 0.0;
    const double ve = g.vel_end ? (double)((const S*)g.vel_end)[p] : 0.0;
    c2 = vs; c0 = (ve - 2.0 * L + vs); c1 = (ve - vs - 3.0 * c0) / 2.0;
    Al = (double)((const S*)g.acc_lo)[p]; Ah = (double)((const S*)g.acc_hi)[p];
    vlo = g.vel_lo ? (double)((const S*)g.vel_lo)[p] : 0.0; vhi = g.vel_hi ? (double)((const S*)g.vel_hi)[p] : 0.0;
  }
  if (!pre_failed && (!(Al <= Ah) || (!GRID && !both_numbers(vlo, vhi)))) status = SCB_ST_ERR_BAD_LIMITS;
  Al = dmax(Al, -1e100); Ah = dmin(Ah, 1e100);
  const double nAl = -Al, nAh = -Ah, step = 1.0 / (double)N, x_cap = g.x_cap;
  const double c0_3 = 3.0 * c0, c1_2 = 2.0 * c1, c0_6 = 6.0 * c0;
  auto qd = [&](double s) { return c0_3 * (s * s) + c1_2 * s + c2; };
  auto qdd = [&](double s) { return c0_6 * s + c1_2; };
  auto s_of = [&](int i) { return i >= N ? 1.0 : (double)i * step; };
  const S* __restrict__ vlg = (const S*)g.vel_lo_grid; const S* __restrict__ vhg = (const S*)g.vel_hi_grid;
  const S* __restrict__ kap = (const S*)g.kappa;
  const double om_max = g.omega_max, al_max = g.acc_lat_max;
  const int nchunk = (N + kWsChunk - 1) / kWsChunk;

  // velocity box of gridpoint i (validated): what both passes' producers and the chain's two end points need
  auto vel_box = [&](int i, double a, double& xlo, double& xhi) -> bool {
    double lo = vlo, hi = vhi;
    const size_t o = (size_t)i * P + p;
    if (GRID) { lo = (double)vlg[o]; hi = (double)vhg[o]; }
    const bool nums = !GRID || both_numbers(lo, hi);
    if (CURV) curvature_cap((double)kap[o], om_max, al_max, lo, hi);
    return vel_xbound(a, lo, hi, xlo, xhi) && nums;
  };

  // ---- backward pass
  // producer: stage i -> buf[b][j]: {al1, -(lo1 twod), -(hi1 twod), r01, D01, al2, -(lo2 twod), -(hi2 twod), r02, D02, xl, xh, flags}
  auto produce_back = [&](int b, int j, int i) {
    const double s = (double)i * step, sn = s_of(i + 1), delta = sn - s, twod = 2.0 * delta;
    const double a = qd(s), bq = qdd(s), an = qd(sn), bn = qdd(sn);
    double xlo, xhi;
    const bool vel_ok = vel_box(i, a, xlo, xhi);
    const Slab s1 = orient(a, bq, Al, Ah, nAl, nAh);
    const Slab s2 = orient(fma(twod, bn, an), bn, Al, Ah, nAl, nAh);
    const double D01 = fma(-twod, s1.be, s1.al), D02 = fma(-twod, s2.be, s2.al), D12 = fma(s1.be, s2.al, -(s2.be * s1.al));
    const double r01 = fast_rcp<double>(D01), r02 = fast_rcp<double>(D02), r12 = fast_rcp<double>(D12);
    double xl = dmax(xlo, 0.0), xh = dmin(xhi, x_cap);
    bool ok = !incoherent(xl, xh);
    ok = pair_apply(D12, r12, fma(s1.hi, s2.al, -(s2.lo * s1.al)), fma(s1.lo, s2.al, -(s2.hi * s1.al)), xl, xh) && ok;
    double (*v)[32] = buf[b][j];
    v[0][lane] = s1.al; v[1][lane] = -(s1.lo * twod); v[2][lane] = -(s1.hi * twod); v[3][lane] = r01; v[4][lane] = D01;
    v[5][lane] = s2.al; v[6][lane] = -(s2.lo * twod); v[7][lane] = -(s2.hi * twod); v[8][lane] = r02; v[9][lane] = D02;
    v[10][lane] = xl; v[11][lane] = xh;
    v[12][lane] = __longlong_as_double((long long)((vel_ok ? 1 : 0) | (ok ? 2 : 0)));
  };
  double nlo = 0, nhi = 0;  // K[N] = (0, 0)  (sc:1225)
  bool alive;
  int first_fail = -1;
  if (warp == 0) {
    put_K((size_t)N * P + p, nlo, nhi);
    if (!pre_failed) {
      double xl, xh;
      if (!vel_box(N, qd(1.0), xl, xh)) status = SCB_ST_ERR_BAD_LIMITS;
    }
  }
  alive = status == SCB_ST_OK;
  auto chunk_first = [&](int c) { return N - 1 - c * kWsChunk; };   // stages chunk_first(c) - j, j = 0 .. kWsChunk - 1, while >= 0
  if (warp > 0) {
    for (int j = warp - 1; j < kWsChunk; j += 3) { const int i = chunk_first(0) - j; if (i >= 0) produce_back(0, j, i); }
  }
  __syncthreads();
#pragma unroll 1
  for (int c = 0; c < nchunk; ++c) {
    const int b = c & 1;
    if (warp > 0) {
      if (c + 1 < nchunk)
        for (int j = warp - 1; j < kWsChunk; j += 3) { const int i = chunk_first(c + 1) - j; if (i >= 0) produce_back(b ^ 1, j, i); }
    } else {
      // The chain, branch-free: a failed problem keeps computing (on whatever its registers hold) and stores zeros, so
      // the only thing between one stage's sdot^2 bounds and the next is the arithmetic that produces them - two FMAs,
      // two products, the selects and two min / max per slab pair; the feasibility flags ride beside it, and the
      // shared-memory loads of the stages of a chunk can be issued ahead of the stages before them.
      auto stage = [&](int j, int i) {
        const size_t o = (size_t)i * P + p;
        const double (*v)[32] = buf[b][j];
        const int fl = (int)__double_as_longlong(v[12][lane]);
        const double al1 = v[0][lane], nlt1 = v[1][lane], nht1 = v[2][lane], r01 = v[3][lane], D01 = v[4][lane];
        const double al2 = v[5][lane], nlt2 = v[6][lane], nht2 = v[7][lane], r02 = v[8][lane], D02 = v[9][lane];
        double xl = v[10][lane], xh = v[11][lane];
        const bool bad = !pre_failed && !(fl & 1);
        status = bad ? SCB_ST_ERR_BAD_LIMITS : status;
        alive = alive && !bad;
        bool ok = (fl & 2) != 0;
        ok = pair_apply(D01, r01, fma(nhi, al1, nlt1), fma(nlo, al1, nht1), xl, xh) && ok;
        ok = pair_apply(D02, r02, fma(nhi, al2, nlt2), fma(nlo, al2, nht2), xl, xh) && ok;
        ok = ok && !incoherent(xl, xh);
        first_fail = (alive && !ok) ? i : first_fail;
        alive = alive && ok;
        nhi = xh; nlo = dmax(0.0, xl);
        put_K(o, alive ? nlo : 0.0, alive ? nhi : 0.0);
      };
      const int i0 = chunk_first(c);
      if (i0 >= kWsChunk - 1) {
#pragma unroll
        for (int j = 0; j < kWsChunk; ++j) stage(j, i0 - j);
      } else {
        for (int j = 0; j <= i0; ++j) stage(j, i0 - j);
      }
    }
    __syncthreads();
  }
  if (warp == 0) {
    if (status == SCB_ST_ERR_BAD_LIMITS) {
      for (int i = 0; i < N; ++i) put_K((size_t)i * P + p, 0.0, 0.0);
    } else if (first_fail >= 0) {
      status = SCB_ST_ERR_FAIL_CONTROLLABLE; fail_stage = first_fail;
    }
  }
  __syncthreads();   // the producers start on the forward pass's first chunk: the backward pass's last buffer is free

  // ---- forward pass
  // producer: stage i -> {twod, r0, s1.al, s1.be, s1.hi, s1.lo, r1, s2.al, s2.be, s2.hi, s2.lo, r2}
  auto produce_fwd = [&](int b, int j, int i) {
    const double s = (double)i * step, sn2 = s_of(i + 1), delta = sn2 - s, twod = 2.0 * delta;
    const double a = qd(s), bq = qdd(s), an2 = qd(sn2), bn2 = qdd(sn2);
    const Slab s1 = orient(a, bq, Al, Ah, nAl, nAh);
    const Slab s2 = orient(fma(twod, bn2, an2), bn2, Al, Ah, nAl, nAh);
    double (*v)[32] = buf[b][j];
    v[0][lane] = twod; v[1][lane] = fast_rcp<double>(twod);
    v[2][lane] = s1.al; v[3][lane] = s1.be; v[4][lane] = s1.hi; v[5][lane] = s1.lo; v[6][lane] = fast_rcp<double>(s1.al);
    v[7][lane] = s2.al; v[8][lane] = s2.be; v[9][lane] = s2.hi; v[10][lane] = s2.lo; v[11][lane] = fast_rcp<double>(s2.al);
  };
  double xi = 0;
  int done = 0;
  bool go = false, ok0 = true;
  double xf0 = 0;
  double kl[kWsChunk], kh[kWsChunk], nl[kWsChunk], nh[kWsChunk];
  if (warp == 0) {
    if (valid) X[p] = (S)xi;
    go = status == SCB_ST_OK;
    if (go) {
      double xlo, xhi;
      vel_box(0, qd(0.0), xlo, xhi);  // validated by the backward pass
      const double lo1 = dmax(xlo, xi), hi1 = dmin(xhi, xi);
      ok0 = !(lo1 > hi1 && (lo1 - hi1 > dmax(fabs(lo1), fabs(hi1)) * ftol::rel_tol));
      xf0 = lo1 > hi1 ? (lo1 + hi1) / 2 : xi;
    }
#pragma unroll
    for (int u = 0; u < kWsChunk; ++u) {
      const size_t ok_ = (size_t)min(u + 1, N) * P + p;
      kl[u] = K_lo[ok_]; kh[u] = K_hi[ok_];
    }
  } else {
    for (int j = warp - 1; j < kWsChunk; j += 3) { if (j < N) produce_fwd(0, j, j); }
  }
  __syncthreads();
#pragma unroll 1
  for (int c = 0; c < nchunk; ++c) {
    const int b = c & 1, i0 = c * kWsChunk;
    if (warp > 0) {
      if (c + 1 < nchunk)
        for (int j = warp - 1; j < kWsChunk; j += 3) { const int i = i0 + kWsChunk + j; if (i < N) produce_fwd(b ^ 1, j, i); }
    } else {
#pragma unroll
      for (int u = 0; u < kWsChunk; ++u) {   // the next chunk's controllable sets, in flight while this chunk's chain runs
        const size_t ok_ = (size_t)min(i0 + kWsChunk + u + 1, N) * P + p;
        nl[u] = K_lo[ok_]; nh[u] = K_hi[ok_];
      }
      auto stage = [&](int u, int i) {   // branch-free like the backward chain: a stopped problem computes on and stores nothing
        const double (*v)[32] = buf[b][u];
        const double twod = v[0][lane], r0 = v[1][lane];
        Slab s1, s2;
        s1.al = v[2][lane]; s1.be = v[3][lane]; s1.hi = v[4][lane]; s1.lo = v[5][lane];
        const double r1 = v[6][lane];
        s2.al = v[7][lane]; s2.be = v[8][lane]; s2.hi = v[9][lane]; s2.lo = v[10][lane];
        const double r2 = v[11][lane];
        const double klo = kl[u], khi = kh[u];
        bool ok = i == 0 ? ok0 : true;
        const double xf = i == 0 ? xf0 : xi;
        const double u0 = (khi - xf) * r0;
        const double u1 = s1.al > 0 ? fma(-s1.be, xf, s1.hi) * r1 : CUDART_INF;
        const double u2 = s2.al > 0 ? fma(-s2.be, xf, s2.hi) * r2 : CUDART_INF;
        const double uu = dmin(u0, dmin(u1, u2));
        const double slack = dmax(ftol::abs_tol, fabs(uu) * ftol::rel_tol);
        ok = ok && (klo - xf - twod * uu) <= twod * slack + ftol::abs_tol;
        ok = ok && (s1.lo - s1.be * xf - s1.al * uu) <= s1.al * slack + ftol::abs_tol;
        ok = ok && (s2.lo - s2.be * xf - s2.al * uu) <= s2.al * slack + ftol::abs_tol;
        ok = ok && (s1.al > 0 || (s1.be * xf - s1.hi) <= ftol::abs_tol);
        ok = ok && (s2.al > 0 || (s2.be * xf - s2.hi) <= ftol::abs_tol);
        const bool fail_now = go && !ok;
        status = fail_now ? SCB_ST_ERR_FAIL_FORWARD_PASS : status;
        fail_stage = fail_now ? i : fail_stage;
        go = go && ok;
        const double xn = dmin(khi, dmax(klo, fma(twod, uu, xi)));
        xi = go ? xn : xi;
        if (go && valid) X[(size_t)(i + 1) * P + p] = (S)xi;
        done = go ? i + 1 : done;
      };
      if (i0 + kWsChunk <= N) {
#pragma unroll
        for (int u = 0; u < kWsChunk; ++u) stage(u, i0 + u);
      } else {
#pragma unroll
        for (int u = 0; u < kWsChunk; ++u) if (i0 + u < N) stage(u, i0 + u);
      }
#pragma unroll
      for (int u = 0; u < kWsChunk; ++u) { kl[u] = nl[u]; kh[u] = nh[u]; }
    }
    __syncthreads();
  }
  if (warp == 0 && valid) {
    for (int i = done + 1; i <= N; ++i) X[(size_t)i * P + p] = (S)0;
    g.status[p] = status;
    g.fail_stage[p] = fail_stage;
  }
}

template <class S> void launch_toppra_fast(const ToppraArgs& a, cudaStream_t st) {
  const int block = 64, grid = (a.P + block - 1) / block;
  const bool gl = a.vel_lo_grid && a.vel_hi_grid, cv = a.kappa && (a.omega_max > 0 || a.acc_lat_max > 0);
  const bool db = a.P < 148 * 64 * 4;  // fewer than four CTAs per SM
  const char* ws_env = getenv("SCB_TOPPRA_WS");  // A/B switch of the tests: 1 = always, 0 = never
  const int groups = (a.P + 31) / 32;
  // at most one CTA per SM (two groups per CTA beyond 148); FP64 storage only: with float storage the chain warp also
  // converts and writes the caller's float copies of K, and that is slower than the kernel above (C5: 7.27 against 5.64 ms)
  const bool ws = ws_env ? atoi(ws_env) != 0 : (groups <= 2 * 148 && sizeof(S) == 8);
  if (ws) {
    auto go = [&](auto kern1, auto kern2) {
      if (groups > 148) kern2<<<(groups + 1) / 2, 256, 0, st>>>(a);
      else kern1<<<groups, 128, 0, st>>>(a);
    };
    if (gl && cv) go(toppra_ws_kernel<S, true, true, 1, 6>, toppra_ws_kernel<S, true, true, 2, 3>);
    else if (gl) go(toppra_ws_kernel<S, true, false, 1, 6>, toppra_ws_kernel<S, true, false, 2, 3>);
    else if (cv) go(toppra_ws_kernel<S, false, true, 1, 6>, toppra_ws_kernel<S, false, true, 2, 3>);
    else go(toppra_ws_kernel<S, false, false, 1, 6>, toppra_ws_kernel<S, false, false, 2, 3>);
    return;
  }
  if (db) {
    if (gl && cv) toppra_fast_kernel<S, true, true, true><<<grid, block, 0, st>>>(a);
    else if (gl) toppra_fast_kernel<S, true, false, true><<<grid, block, 0, st>>>(a);
    else if (cv) toppra_fast_kernel<S, false, true, true><<<grid, block, 0, st>>>(a);
    else toppra_fast_kernel<S, false, false, true><<<grid, block, 0, st>>>(a);
  } else {
    if (gl && cv) toppra_fast_kernel<S, true, true, false><<<grid, block, 0, st>>>(a);
    else if (gl) toppra_fast_kernel<S, true, false, false><<<grid, block, 0, st>>>(a);
    else if (cv) toppra_fast_kernel<S, false, true, false><<<grid, block, 0, st>>>(a);
    else toppra_fast_kernel<S, false, false, false><<<grid, block, 0, st>>>(a);
  }
}
template void launch_toppra_fast<double>(const ToppraArgs&, cudaStream_t);
template void launch_toppra_fast<float>(const ToppraArgs&, cudaStream_t);

}  // namespace scb
