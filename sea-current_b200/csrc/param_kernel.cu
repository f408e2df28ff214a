// Parametrizer + geometry on the GPU, one thread per planning problem (sm_100a).
//
// param_kernel   parametrizer::Spline  (tp:parametrizer/spline.cpp:10-45, tp:parametrizer.cpp:6-21):
//                knot times by the trapezoid rule, clamped cubic spline q(t) through (t_i, q(s_i))
//                (tp:geometric_path/piecewise_poly_path.cpp:102-187) — solved as the tridiagonal
//                system it is (Thomas) instead of the reference's dense column-pivoted QR —
//                then the per-gridpoint record: v, a (spline derivatives at the knots) and
//                omega = v * kappa (sea_current.hpp:1055-1067) with kappa at arc length q(s_i) as
//                written by path_kernel.
// param_ca_kernel  parametrizer::ConstAccel (tp:parametrizer/const_accel.cpp), scb_batch_in.param_mode = 1.
// sample_kernel  the uniform-dt stream of gen_vel_prof (sea_current.hpp:1235-1262), resample
//                (sc:898-1005) and angular_velocity (sc:1055-1067).
//
// Arc length -> curve parameter is an exact inversion of the arc-length table (search + Newton
// on a 4-point Gauss-Legendre integral of |B'|), not the reference's float Chebyshev fit
// (DESIGN.md "Deviations").  All gridpoint / sample arrays are [n][P], problem index fastest.
#include <cstdio>

#include "scb_common.cuh"

namespace scb {

__device__ __constant__ double kGL4X[4] = {-0.8611363115940525752, -0.3399810435848562648, 0.3399810435848562648,
                                           0.8611363115940525752};
__device__ __constant__ double kGL4W[4] = {0.3478548451374538574, 0.6521451548625461426, 0.6521451548625461426,
                                           0.3478548451374538574};

// Power-basis form of one Bezier segment: a_k = C(d,k) * (k-th forward difference of P at 0).
template <class R, int DEG> struct SegPoly {
  R ax[DEG + 1], ay[DEG + 1];
  __device__ __forceinline__ void load(const R* __restrict__ cx, const R* __restrict__ cy) {
    R dx[DEG + 1], dy[DEG + 1];
#pragma unroll
    for (int i = 0; i <= DEG; ++i) { dx[i] = cx[i]; dy[i] = cy[i]; }
    R binom = 1;
#pragma unroll
    for (int k = 0; k <= DEG; ++k) {
      ax[k] = binom * dx[0]; ay[k] = binom * dy[0];
#pragma unroll
      for (int j = 0; j + k < DEG; ++j) { dx[j] = dx[j + 1] - dx[j]; dy[j] = dy[j + 1] - dy[j]; }
      binom = binom * (R)(DEG - k) / (R)(k + 1);
    }
  }
  __device__ __forceinline__ void d1(R u, R& x, R& y) const {
    x = (R)DEG * ax[DEG]; y = (R)DEG * ay[DEG];
#pragma unroll
    for (int k = DEG - 2; k >= 0; --k) { x = x * u + (R)(k + 1) * ax[k + 1]; y = y * u + (R)(k + 1) * ay[k + 1]; }
  }
  __device__ __forceinline__ R speed(R u) const { R x, y; d1(u, x, y); return sqrt(x * x + y * y); }
  __device__ __forceinline__ void eval(R u, R& px, R& py, R& kappa) const {
    px = ax[DEG]; py = ay[DEG];
#pragma unroll
    for (int k = DEG - 1; k >= 0; --k) { px = px * u + ax[k]; py = py * u + ay[k]; }
    R x1, y1; d1(u, x1, y1);
    // B'' coefficients: (k+1) * [(k+2) a_{k+2}]
    R x2 = (R)(DEG - 1) * ((R)DEG * ax[DEG]), y2 = (R)(DEG - 1) * ((R)DEG * ay[DEG]);
#pragma unroll
    for (int k = DEG - 3; k >= 0; --k) {
      x2 = x2 * u + (R)(k + 1) * ((R)(k + 2) * ax[k + 2]);
      y2 = y2 * u + (R)(k + 1) * ((R)(k + 2) * ay[k + 2]);
    }
    const R n2 = x1 * x1 + y1 * y1;
    kappa = (x1 * y2 - y1 * x2) / (n2 * sqrt(n2));  // sc:1029
  }
};

// Walks the arc-length table of one problem: arc length q -> (segment, u) -> point, curvature.
template <class R, int DEG> struct ArcWalker {
  const R* cx; const R* cy; const R* table;  // this problem's rows
  int nseg, seg, k;
  R seg_start, seg_end, total;
  SegPoly<R, DEG> poly;
  __device__ __forceinline__ void set_segment(int s) {
    seg = s;
    R acc = 0;
    for (int j = 0; j < s; ++j) acc += table[j * kTable + kSub];
    seg_start = acc; seg_end = acc + table[s * kTable + kSub];
    poly.load(cx + s * (DEG + 1), cy + s * (DEG + 1));
    k = 0;
  }
  __device__ __forceinline__ void init(const R* cx_, const R* cy_, const R* table_, int nseg_) {
    cx = cx_; cy = cy_; table = table_; nseg = nseg_;
    R acc = 0;
    for (int j = 0; j < nseg; ++j) acc += table[j * kTable + kSub];
    total = acc;
    set_segment(0);
  }
  __device__ __forceinline__ void locate(R q, R& px, R& py, R& kappa) {
    q = fmin(fmax(q, (R)0), total);
    if (q > seg_end) {
      while (seg < nseg - 1 && q > seg_end) {  // forward: cumulative sums continue exactly
        ++seg; seg_start = seg_end; seg_end = seg_start + table[seg * kTable + kSub];
      }
      poly.load(cx + seg * (DEG + 1), cy + seg * (DEG + 1)); k = 0;
    } else if (seg > 0 && q <= seg_start) {
      int s = seg;
      R st = seg_start;
      // backward: recompute the prefix so it equals the forward cumulative sum bit for bit
      while (s > 0 && q <= st) { --s; st = 0; for (int j = 0; j < s; ++j) st += table[j * kTable + kSub]; }
      set_segment(s);
    }
    const R r = q - seg_start;
    const R* c = table + seg * kTable;
    while (k < kSub - 1 && c[k + 1] <= r) ++k;
    while (k > 0 && c[k] > r) --k;
    const R h = (R)0.01, a = fmin((R)k * h, (R)1);
    const R ck = c[k], target = r - ck;
    R du = h * target / (c[k + 1] - ck);
    du = fmin(fmax(du, (R)0), h);
    // Newton to convergence (linear, not quadratic, next to a cusp of the curve: a fixed count is not enough there)
    const R tol = (sizeof(R) == 8 ? (R)1e-12 : (R)1e-6) * h;
#pragma unroll 1
    for (int it = 0; it < kInvNewtonMax; ++it) {
      const R half = (R)0.5 * du, mid = a + half;
      R F = 0;
#pragma unroll
      for (int gq = 0; gq < 4; ++gq) F += (R)kGL4W[gq] * poly.speed(mid + half * (R)kGL4X[gq]);
      F *= half;
      const R delta = (F - target) / poly.speed(a + du);
      du = fmin(fmax(du - delta, (R)0), h);
      if (fabs(delta) <= tol) break;
    }
    poly.eval(fmin(a + du, (R)1), px, py, kappa);
  }
};

struct Hermite1 {  // FP64 in both precision modes, like the toppra kernel
  double c0, c1, c2, step; int N;
  __device__ __forceinline__ void init(double L, double vs, double ve, int N_) {
    c2 = vs; c0 = (ve - 2.0 * L + vs); c1 = (ve - vs - 3.0 * c0) / 2.0; N = N_; step = 1.0 / (double)N_;
  }
  __device__ __forceinline__ double s_at(int i) const { return i == N ? 1.0 : (double)i * step; }
  __device__ __forceinline__ double q(double s) const { return c0 * (s * s * s) + c1 * (s * s) + c2 * s; }
  __device__ __forceinline__ double qd(double s) const { return (3.0 * c0) * (s * s) + (2.0 * c1) * s + c2; }
  __device__ __forceinline__ double qdd(double s) const { return (6.0 * c0) * s + 2.0 * c1; }
};

// `S` is the storage type of the caller's arrays; the recurrences (knot times, Thomas) run in FP64
// and their scratch (t_buf, X_buf, cp_buf) is always double.  Both sweeps are sequential per problem,
// so the kernel lives on memory-level parallelism: every thread fetches kPU gridpoints of each input
// array before it touches the dependent arithmetic (a warp's 32 problems = 256 contiguous bytes per
// gridpoint), and quotients go through fast_rcp / fast_sqrt (no IEEE division sequences).
constexpr int kPU = 8;

// O: type of the record arrays t, v, a, omega (= S, or float when scb_batch_out.record_f32 asks for float records)
// DB: both sweeps fetch the NEXT block of gridpoints while the current block's dependent arithmetic runs (57 % of the
//     kernel's stall samples are long-scoreboard waits on the first use of a block it has just requested, profiles/r2q
//     per-line profile).  On C3 the kernel already moves its traffic at 84 % of the HBM peak, so the variant is for
//     batches that leave the SMs short of warps (the launcher picks it below four CTAs of toppra per SM, like toppra's).
template <class S, class O, bool DB>
__global__ void __launch_bounds__(32) param_kernel(ParamArgs g) {
  typedef double R;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  const S* __restrict__ X2 = (const S*)g.x;  // sdot^2
  R* __restrict__ tb = (R*)g.t_buf; R* __restrict__ Xb = (R*)g.X_buf; R* __restrict__ cpb = (R*)g.cp_buf;
  O* __restrict__ o_t = (O*)g.t; O* __restrict__ o_v = (O*)g.v; O* __restrict__ o_a = (O*)g.a; O* __restrict__ o_om = (O*)g.omega;
  const S* __restrict__ kap = (const S*)g.kappa;
  int status = g.status[p];
  // A problem that arrives failed (bad path, collision verdict, toppra's verdict) gets an all-zero record.  Its thread
  // walks both sweeps WITH its warp and stores zeros through the same store instructions as its healthy neighbours
  // (`live` selects the value): no divergent zero-fill loop, and no 32-byte sector written in two partial pieces at
  // different times (measured with 7 % of C3 colliding: a separate zero-fill loop doubled the kernel's time).
  const bool live = status == SCB_ST_OK;
  Hermite1 hp;
  hp.init(live ? (R)((const S*)g.L)[p] : 1.0, g.vel_start ? (R)((const S*)g.vel_start)[p] : 0.0,
          g.vel_end ? (R)((const S*)g.vel_end)[p] : 0.0, N);
  {
    // ---- pass 1: knot times (trapezoid rule, spline.cpp:14-24) + Thomas forward elimination of the
    //      clamped cubic spline through (t_i, q(s_i))  (piecewise_poly_path.cpp:102-176)
    const R sd0 = fast_sqrt((R)X2[p]);
    R sd_prev = sd0, t_prev = 0, h_prev = 0, slope_prev = 0;
    R q_prev = hp.q(hp.s_at(0)), s_prev = hp.s_at(0);
    R cp_prev = 0, b_prev = 0;
    const R bc0 = hp.qd(hp.s_at(0)) * sd0;
    tb[p] = 0;
    R xs[kPU], xn[DB ? kPU : 1];
    if constexpr (DB) {
#pragma unroll
      for (int u = 0; u < kPU; ++u) xs[u] = (R)X2[(size_t)min(1 + u, N) * P + p];
    }
    for (int i0 = 1; i0 <= N; i0 += kPU) {
#pragma unroll
      for (int u = 0; u < kPU; ++u) {
        if constexpr (DB) xn[u] = (R)X2[(size_t)min(i0 + kPU + u, N) * P + p];
        else xs[u] = (R)X2[(size_t)min(i0 + u, N) * P + p];
      }
#pragma unroll
      for (int u = 0; u < kPU; ++u) {
        const int i = i0 + u;
        if (i <= N) {
          const R s = hp.s_at(i), sd = fast_sqrt(xs[u]);
          const R avg = (sd_prev + sd) * 0.5;
          const R dt = avg > 1e-8 ? (s - s_prev) * fast_rcp<R>(avg) : 5.0;
          if (live && dt < 1e-8) status = SCB_ST_ERR_TIME_KNOTS;
          const R t = t_prev + dt;
          tb[(size_t)i * P + p] = live ? t : 0.0;
          const R qi = hp.q(s);
          const R h = t - t_prev;                          // h_{i-1}
          const R slope = (qi - q_prev) * fast_rcp<R>(h);  // (q_i - q_{i-1}) / h_{i-1}
          // row j = i - 1 of the tridiagonal system
          R lo, di, B;
          if (i == 1) { lo = 0; di = 2 * h; B = 3 * slope - 3 * bc0; }
          else { lo = h_prev; di = 2 * (h_prev + h); B = 3 * slope - 3 * slope_prev; }
          const R rm = fast_rcp<R>(di - lo * cp_prev);
          const R cpj = h * rm, bj = (B - lo * b_prev) * rm;
          cpb[(size_t)(i - 1) * P + p] = cpj; Xb[(size_t)(i - 1) * P + p] = bj;
          cp_prev = cpj; b_prev = bj;
          if (i == N) {  // last row j = N
            const R bc1 = hp.qd(s) * sd;
            const R B2 = 3 * bc1 - 3 * slope;
            Xb[(size_t)N * P + p] = (B2 - h * b_prev) * fast_rcp<R>(2 * h - h * cp_prev);  // = X_N
          }
          t_prev = t; h_prev = h; slope_prev = slope; q_prev = qi; s_prev = s; sd_prev = sd;
        }
      }
      if constexpr (DB) {
#pragma unroll
        for (int u = 0; u < kPU; ++u) xs[u] = xn[u];
      }
    }
  }
  if (live && status != SCB_ST_OK) {  // failed in pass 1 (knot times; rare): its knot times are already stored
    g.status[p] = status;
    for (int i = 0; i < G; ++i) tb[(size_t)i * P + p] = 0;
  }
  const bool good = status == SCB_ST_OK;
  // ---- pass 2 (backward): back substitution, knot derivatives, angular velocity
  R X_next = Xb[(size_t)N * P + p];
  R t_next = tb[(size_t)N * P + p];
  R q_next = hp.q(hp.s_at(N));
  if (g.T_total) ((S*)g.T_total)[p] = good ? (S)(t_next - tb[p]) : (S)0;
  const R kapN = kap ? (R)kap[(size_t)N * P + p] : 0.0;
  const R X_N = X_next, t_N = t_next;
  R xb[kPU], cb[kPU], tt[kPU], kk[kPU], xb2[DB ? kPU : 1], cb2[DB ? kPU : 1], tt2[DB ? kPU : 1], kk2[DB ? kPU : 1];
  if constexpr (DB) {
#pragma unroll
    for (int u = 0; u < kPU; ++u) {
      const size_t o = (size_t)max(N - 1 - u, 0) * P + p;
      xb[u] = Xb[o]; cb[u] = cpb[o]; tt[u] = tb[o]; kk[u] = kap ? (R)kap[o] : 0.0;
    }
  }
  for (int i0 = N - 1; i0 >= 0; i0 -= kPU) {
#pragma unroll
    for (int u = 0; u < kPU; ++u) {
      if constexpr (DB) {
        const size_t o = (size_t)max(i0 - kPU - u, 0) * P + p;
        xb2[u] = Xb[o]; cb2[u] = cpb[o]; tt2[u] = tb[o]; kk2[u] = kap ? (R)kap[o] : 0.0;
      } else {
        const size_t o = (size_t)max(i0 - u, 0) * P + p;
        xb[u] = Xb[o]; cb[u] = cpb[o]; tt[u] = tb[o]; kk[u] = kap ? (R)kap[o] : 0.0;
      }
    }
#pragma unroll
    for (int u = 0; u < kPU; ++u) {
      const int i = i0 - u;
      if (i >= 0) {
        const size_t o = (size_t)i * P + p;
        const R Xi = xb[u] - cb[u] * X_next;
        if (g.keep_X) Xb[o] = good ? Xi : 0.0;
        const R ti = tt[u], qv = hp.q(hp.s_at(i));
        const R h = t_next - ti;
        const R vi = (q_next - qv) * fast_rcp<R>(h) - h * (1.0 / 3.0) * (2 * Xi + X_next);
        if (i == N - 1) {
          // knot N is evaluated on the last segment at ds = h (piecewise_poly_path.cpp:240-252)
          const R k0 = (X_N - Xi) * fast_rcp<R>(3 * h), k1 = Xi;
          const R vN = (k0 * 3) * (h * h) + (k1 * 2) * h + vi, aN = ((k0 * 3) * 2) * h + k1 * 2;
          const size_t oN = (size_t)N * P + p;
          if (g.keep_X && !good) Xb[oN] = 0.0;
          if (o_t) o_t[oN] = good ? (O)t_N : (O)0;
          if (o_v) o_v[oN] = good ? (O)vN : (O)0;
          if (o_a) o_a[oN] = good ? (O)aN : (O)0;
          if (o_om) o_om[oN] = good ? (O)(vN * kapN) : (O)0;
        }
        if (o_t) o_t[o] = good ? (O)ti : (O)0;
        if (o_v) o_v[o] = good ? (O)vi : (O)0;
        if (o_a) o_a[o] = good ? (O)(Xi * 2) : (O)0;
        if (o_om) o_om[o] = good ? (O)(vi * kk[u]) : (O)0;
        X_next = Xi; t_next = ti; q_next = qv;
      }
    }
    if constexpr (DB) {
#pragma unroll
      for (int u = 0; u < kPU; ++u) { xb[u] = xb2[u]; cb[u] = cb2[u]; tt[u] = tt2[u]; kk[u] = kk2[u]; }
    }
  }
}

// parametrizer::ConstAccel (tp:parametrizer/const_accel.cpp:19-33 process_internals, 35-58 evalParams, 60-115
// eval_impl): constant path acceleration u_k on every grid interval, so one forward sweep gives the knot times and the
// record of every gridpoint, with no system to solve and no scratch besides the knot times the sampler reads.
// Knot i >= 1 is evaluated the way evalParams finds it: on interval k = i - 1 at dt = t_i - t_{i-1} (lower_bound over
// m_ts[1 .. G-2]); knot 0 on interval 0 at dt = 0.
template <class S, class O>
__global__ void __launch_bounds__(32) param_ca_kernel(ParamArgs g) {
  typedef double R;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, G = g.G, N = G - 1;
  const S* __restrict__ X2 = (const S*)g.x;
  R* __restrict__ tb = (R*)g.t_buf;
  O* __restrict__ o_t = (O*)g.t; O* __restrict__ o_v = (O*)g.v; O* __restrict__ o_a = (O*)g.a; O* __restrict__ o_om = (O*)g.omega;
  const S* __restrict__ kap = (const S*)g.kappa;
  int status = g.status[p];
  if (status == SCB_ST_OK) {
    Hermite1 hp;
    hp.init((R)((const S*)g.L)[p], g.vel_start ? (R)((const S*)g.vel_start)[p] : 0.0,
            g.vel_end ? (R)((const S*)g.vel_end)[p] : 0.0, N);
    R v_prev = sqrt((R)X2[p]), t_prev = 0, s_prev = hp.s_at(0);
    const R kap0 = kap ? (R)kap[p] : 0.0;
    tb[p] = 0;
    for (int i0 = 1; i0 <= N; i0 += kPU) {
      R xs[kPU], kk[kPU];
#pragma unroll
      for (int u = 0; u < kPU; ++u) {
        const size_t o = (size_t)min(i0 + u, N) * P + p;
        xs[u] = (R)X2[o]; kk[u] = kap ? (R)kap[o] : 0.0;
      }
#pragma unroll
      for (int u = 0; u < kPU; ++u) {
        const int i = i0 + u;
        if (i <= N) {
          const size_t o = (size_t)i * P + p;
          const R s = hp.s_at(i), v = sqrt(xs[u]), ds = s - s_prev;
          const R acc_k = 0.5 * (v * v - v_prev * v_prev) / ds;  // m_us[i-1]
          const R t = t_prev + 2 * ds / (v + v_prev);             // m_ts[i]
          tb[o] = t;
          if (i == 1) {  // knot 0: interval 0, dt = 0
            const R q1 = hp.qd(s_prev);
            const R vel = q1 * v_prev, acc = hp.qdd(s_prev) * v_prev * v_prev + q1 * acc_k;
            if (o_t) o_t[p] = (O)0;
            if (o_v) o_v[p] = (O)vel;
            if (o_a) o_a[p] = (O)acc;
            if (o_om) o_om[p] = (O)(vel * kap0);
          }
          const R dk = t - t_prev;
          const R vs = v_prev + dk * acc_k;
          const R ss = fmin(fmax(s_prev + dk * v_prev + 0.5 * dk * dk * acc_k, 0.0), 1.0);  // const_accel.cpp:80-81
          const R q1 = hp.qd(ss);
          const R vel = q1 * vs, acc = hp.qdd(ss) * vs * vs + q1 * acc_k;
          if (o_t) o_t[o] = (O)t;
          if (o_v) o_v[o] = (O)vel;
          if (o_a) o_a[o] = (O)acc;
          if (o_om) o_om[o] = (O)(vel * kk[u]);
          t_prev = t; s_prev = s; v_prev = v;
        }
      }
    }
    // two neighbouring gridpoints at rest: the reference divides by zero (const_accel.cpp:28); reported here
    if (!(t_prev < 1.0e300)) status = SCB_ST_ERR_TIME_KNOTS;
    else if (g.T_total) ((S*)g.T_total)[p] = (S)t_prev;
  }
  if (status != SCB_ST_OK) {
    if (g.status[p] == SCB_ST_OK) g.status[p] = status;
    for (int i = 0; i < G; ++i) {
      const size_t o = (size_t)i * P + p;
      tb[o] = 0;
      if (o_t) o_t[o] = 0; if (o_v) o_v[o] = 0; if (o_a) o_a[o] = 0; if (o_om) o_om[o] = 0;
    }
    if (g.T_total) ((S*)g.T_total)[p] = 0;
  }
}

// O: type of the stream arrays s_* (= R, or float when scb_batch_out.record_f32 asks for float records; in a
// geometry-only call the caller's s_pos / s_vel inputs have that type too)
template <class R, int DEG, class O>
__global__ void __launch_bounds__(64) sample_kernel(SampleArgs g) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= g.P) return;
  const int P = g.P, N = g.G - 1, S_cap = g.S_cap;
  O* o_time = (O*)g.s_time; O* o_pos = (O*)g.s_pos; O* o_vel = (O*)g.s_vel; O* o_acc = (O*)g.s_acc;
  O* o_px = (O*)g.s_px; O* o_py = (O*)g.s_py; O* o_om = (O*)g.s_omega;
  const bool ok = g.status[p] == SCB_ST_OK;
  int n = 0;
  if (ok) {
    ArcWalker<R, DEG> walk;
    const bool geo = g.geometry != 0;
    if (geo) {
      const int nseg = g.W - 1;
      walk.init((const R*)g.ctrl_x + (size_t)p * nseg * (DEG + 1), (const R*)g.ctrl_y + (size_t)p * nseg * (DEG + 1),
                (const R*)g.table + (size_t)p * nseg * kTable, nseg);
    }
    if (g.timing) {
      const double* tb = (const double*)g.t_buf; const double* Xb = (const double*)g.X_buf;
      Hermite1 hp;
      hp.init((double)((const R*)g.L)[p], g.vel_start ? (double)((const R*)g.vel_start)[p] : 0.0,
              g.vel_end ? (double)((const R*)g.vel_end)[p] : 0.0, N);
      const double t0 = tb[p], T = tb[(size_t)N * P + p];
      const int S = (int)ceil((T - t0) / g.dt);  // sc:1237
      g.n_samples[p] = S;
      n = S < S_cap ? S : S_cap;
      const double stp = S > 1 ? (T - t0) / (double)(S - 1) : 0.0;
      const bool ca = g.param_mode == SCB_PARAM_CONST_ACCEL;
      const R* X2 = (const R*)g.x;
      int j = -1;
      double tj = 0, k0 = 0, k1 = 0, k2 = 0, k3 = 0, t_up = t0;
      for (int i = 0; i < n; ++i) {
        const double time = (i == S - 1) ? T : t0 + (double)i * stp;  // Eigen LinSpaced
        double pos, vel, acc;
        if (ca) {
          // evalParams (const_accel.cpp:42-57): interval k with m_ts[k] < time <= m_ts[k+1]; k0 = u_k, k1 = v_k, k2 = s_k
          if (j < 0 || (j < N - 1 && t_up < time)) {
            if (j < 0) j = 0;
            while (j < N - 1 && tb[(size_t)(j + 1) * P + p] < time) ++j;
            tj = tb[(size_t)j * P + p]; t_up = tb[(size_t)(j + 1) * P + p];
            const double vj = sqrt((double)X2[(size_t)j * P + p]), vn = sqrt((double)X2[(size_t)(j + 1) * P + p]);
            k2 = hp.s_at(j); k1 = vj;
            k0 = 0.5 * (vn * vn - vj * vj) / (hp.s_at(j + 1) - k2);
          }
          const double d = time - tj, vs = k1 + d * k0;
          const double ss = fmin(fmax(k2 + d * k1 + 0.5 * d * d * k0, 0.0), 1.0);
          const double q1 = hp.qd(ss);
          pos = hp.q(ss); vel = q1 * vs; acc = hp.qdd(ss) * vs * vs + q1 * k0;
        } else {
          if (j < 0 || (j < N - 1 && t_up <= time)) {
            if (j < 0) j = 0;
            while (j < N - 1 && tb[(size_t)(j + 1) * P + p] <= time) ++j;
            tj = tb[(size_t)j * P + p]; t_up = tb[(size_t)(j + 1) * P + p];
            const double Xj = Xb[(size_t)j * P + p], Xn = Xb[(size_t)(j + 1) * P + p];
            const double qj = hp.q(hp.s_at(j)), qn = hp.q(hp.s_at(j + 1));
            const double h = t_up - tj;
            k0 = (Xn - Xj) / (3 * h); k1 = Xj; k2 = (qn - qj) / h - h / 3 * (2 * Xj + Xn); k3 = qj;
          }
          const double ds = time - tj;
          pos = k0 * (ds * ds * ds) + k1 * (ds * ds) + k2 * ds + k3;
          vel = (k0 * 3) * (ds * ds) + (k1 * 2) * ds + k2;
          acc = ((k0 * 3) * 2) * ds + k1 * 2;
        }
        const size_t o = (size_t)i * P + p;
        if (o_time) o_time[o] = (O)time;
        if (o_pos) o_pos[o] = (O)pos;
        if (o_vel) o_vel[o] = (O)vel;
        if (o_acc) o_acc[o] = (O)acc;
        if (geo) {
          R px, py, kap;
          walk.locate((R)pos, px, py, kap);
          if (o_px) o_px[o] = (O)px;
          if (o_py) o_py[o] = (O)py;
          if (o_om) o_om[o] = (O)(vel * (double)kap);
        }
      }
    } else if (geo) {  // resample / angular_velocity of caller-supplied positions
      const int S = g.n_samples[p];      // caller-supplied: clamp to [0, S_cap]
      n = S < S_cap ? S : S_cap;
      if (n < 0) n = 0;
      if (g.nudge && n > 0) {  // sc:902-912, in place like the reference (it takes profile_pos by reference)
        const O Ltot = (O)walk.total;
        o_pos[p] = 0;
        o_pos[(size_t)(n - 1) * P + p] = Ltot;
        O prev = 0;
        for (int i = 1; i + 1 < n; ++i) {
          const size_t o = (size_t)i * P + p;
          O cur = o_pos[o];
          const O nxt = o_pos[o + P];
          if (cur < prev || cur > nxt) cur = (prev + nxt) / 2;
          if (cur < 0) cur = 0;
          if (cur > Ltot) cur = Ltot;
          o_pos[o] = cur;
          prev = cur;
        }
      }
      for (int i = 0; i < n; ++i) {
        const size_t o = (size_t)i * P + p;
        R px, py, kap;
        walk.locate((R)o_pos[o], px, py, kap);
        if (o_px) o_px[o] = (O)px;
        if (o_py) o_py[o] = (O)py;
        if (o_om) o_om[o] = (O)((o_vel ? (R)o_vel[o] : (R)1) * kap);
      }
    }
  } else if (g.timing) {
    g.n_samples[p] = 0;
  }
  for (int i = n; i < S_cap; ++i) {
    const size_t o = (size_t)i * P + p;
    if (g.timing) { if (o_time) o_time[o] = 0; if (o_pos) o_pos[o] = 0; if (o_vel) o_vel[o] = 0; if (o_acc) o_acc[o] = 0; }
    if (o_px) o_px[o] = 0; if (o_py) o_py[o] = 0; if (o_om) o_om[o] = 0;
  }
}

template <class R> void launch_param(const ParamArgs& a, cudaStream_t st) {
  const int block = 32, grid = (a.P + block - 1) / block;
  const bool f32 = a.rec_f32 && sizeof(R) == 8;  // R = storage type
  if (a.param_mode == SCB_PARAM_CONST_ACCEL) {
    if (f32) param_ca_kernel<R, float><<<grid, block, 0, st>>>(a); else param_ca_kernel<R, R><<<grid, block, 0, st>>>(a);
  } else {
    const bool db = a.P < 148 * 64 * 4;  // measured on C3 with the rule overridden: 1.034 ms without, 1.039 ms with (HBM-bound either way)
    if (db) { if (f32) param_kernel<R, float, true><<<grid, block, 0, st>>>(a); else param_kernel<R, R, true><<<grid, block, 0, st>>>(a); }
    else { if (f32) param_kernel<R, float, false><<<grid, block, 0, st>>>(a); else param_kernel<R, R, false><<<grid, block, 0, st>>>(a); }
  }
}
template <class R> void launch_sample(const SampleArgs& a, cudaStream_t st) {
  const int block = 64, grid = (a.P + block - 1) / block;
  const bool f32 = a.rec_f32 && sizeof(R) == 8;
  if (a.deg == 3) { if (f32) sample_kernel<R, 3, float><<<grid, block, 0, st>>>(a); else sample_kernel<R, 3, R><<<grid, block, 0, st>>>(a); }
  else { if (f32) sample_kernel<R, 5, float><<<grid, block, 0, st>>>(a); else sample_kernel<R, 5, R><<<grid, block, 0, st>>>(a); }
}
template void launch_param<double>(const ParamArgs&, cudaStream_t);
template void launch_param<float>(const ParamArgs&, cudaStream_t);
template void launch_sample<double>(const SampleArgs&, cudaStream_t);
template void launch_sample<float>(const SampleArgs&, cudaStream_t);

}  // namespace scb

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free entry (include/seacurrent_b200.h: scb_param_static).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The Thomas scratch of the spline parametrizer (elimination coefficients c' and the
// right-hand sides / unknowns X, [G][P] doubles each) lives in ONE file-scope device array of fixed size; every GPU
// that loads the library holds its 16 MiB until exit.  The kernel is param_kernel<double, double, DB> itself: same
// arithmetic, same bits as the timing stage of scb_plan_batch_device.
//   * Nothing in the scratch has to start at zero: every element a problem's backward sweep reads (rows 0..N of X,
//     rows 0..N-1 of c') was written by the same thread's forward sweep of the SAME call, so the previous call's
//     contents are never observed.
//   * Every call shares the array: two calls that overlap on the GPU (different streams, or the same stream from two
//     host threads without ordering) race on it.  Order them on one stream or with events.
namespace scb {
namespace {
__device__ __align__(256) double g_param_static_scratch[2 * (size_t)SCB_STATIC_MAX_GP];

constexpr int kStaticMaxDevices = 64;
double* g_static_ptr[kStaticMaxDevices] = {};   // device address of the array on each device (set once per device)
thread_local char g_static_err[256] = {0};

int static_fail(cudaError_t e, const char* where) {
  snprintf(g_static_err, sizeof g_static_err, "%s: %s", where, cudaGetErrorString(e));
  cudaGetLastError();
  return SCB_E_CUDA;
}
}  // namespace
}  // namespace scb

extern "C" {

const char* scb_static_last_error(void) { return scb::g_static_err; }

int scb_param_static(int32_t P, int32_t G, const double* path_len, const double* x, const double* vel_start,
                     const double* vel_end, int32_t* status, double* t, double* v, double* a, double* T_total, void* stream) {
  using namespace scb;
  g_static_err[0] = 0;
  if (P < 0 || G < 2 || !path_len || !x || !status || !t) {
    snprintf(g_static_err, sizeof g_static_err, "scb_param_static: bad argument (P >= 0, G >= 2, path_len, x, status and t are required)");
    return SCB_E_INVALID;
  }
  if ((long long)P * (long long)G > (long long)SCB_STATIC_MAX_GP) {
    snprintf(g_static_err, sizeof g_static_err, "scb_param_static: G * P = %lld exceeds the static scratch (SCB_STATIC_MAX_GP = %d)",
             (long long)P * (long long)G, (int)SCB_STATIC_MAX_GP);
    return SCB_E_TOO_LARGE;
  }
  if (P == 0) return SCB_OK;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_fail(e, "cudaGetDevice");
  if (dev < 0 || dev >= kStaticMaxDevices) {
    snprintf(g_static_err, sizeof g_static_err, "scb_param_static: device ordinal %d beyond %d", dev, kStaticMaxDevices);
    return SCB_E_INVALID;
  }
  // The array's name is a host-side stand-in here, not a device address: ask the runtime for the address on THIS
  // device (the current one), once per device.  Racing first calls from two host threads store the same value.
  double* scratch = g_static_ptr[dev];
  if (!scratch) {
    e = cudaGetSymbolAddress((void**)&scratch, g_param_static_scratch);
    if (e != cudaSuccess || !scratch) return static_fail(e, "cudaGetSymbolAddress(g_param_static_scratch)");
    g_static_ptr[dev] = scratch;
  }
  const size_t gp = (size_t)G * (size_t)P;
  ParamArgs q{};
  q.P = P; q.G = G; q.L = path_len; q.vel_start = vel_start; q.vel_end = vel_end; q.x = x; q.kappa = nullptr;
  q.t = nullptr;           // the caller's FP64 t array is the knot-time state itself, as in scb_plan_batch_device
  q.v = v; q.a = a; q.omega = nullptr; q.rec_f32 = 0;
  q.t_buf = t; q.X_buf = scratch; q.cp_buf = scratch + gp;
  q.T_total = T_total; q.status = status; q.keep_X = 0; q.param_mode = SCB_PARAM_SPLINE;
  launch_param<double>(q, (cudaStream_t)stream);
  e = cudaGetLastError();
  if (e != cudaSuccess) return static_fail(e, "param_kernel launch");
  return SCB_OK;
}

}  // extern "C"
