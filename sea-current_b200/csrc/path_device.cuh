// Device helpers shared by the two path kernels (path_kernel.cu: CTA team, path_lane_kernel.cu:
// lane per waypoint).  Reference lines are sea_current.hpp ("sc:").
#pragma once
#include "gl32_device.h"
#include "scb_common.cuh"

namespace scb {

template <class R> __device__ __forceinline__ R dist2(R ax, R ay, R bx, R by) {  // sc:86-88 pt_dist
  const R dx = bx - ax, dy = by - ay;
  return sqrt(dx * dx + dy * dy);
}

// sc:347-369 (acos argument clamped: the one conscious fix, DESIGN.md "Deviations")
template <class R>
__device__ __forceinline__ void calc_tangent(R w0x, R w0y, R w1x, R w1y, R w2x, R w2y, R& tx, R& ty) {
  const R ux = w0x - w1x, uy = w0y - w1y, vx = w2x - w1x, vy = w2y - w1y;
  const R udv = ux * vx + uy * vy;
  const R du = dist2<R>(ux, uy, 0, 0), dv = dist2<R>(vx, vy, 0, 0);
  R arg = udv / (du * dv);
  arg = fmax((R)-1, fmin((R)1, arg));
  const R theta = acos(arg) / 2;
  const R offset = atan2(uy, ux), test_offset = atan2(vy, vx);
  const R mult = (test_offset - offset < 0) ? (R)-1 : (R)1;  // no angle wrapping, as the reference
  R sn, cs;
  sincos(offset + mult * theta, &sn, &cs);
  R lx = sn, ly = -cs;
  const R n2 = lx * lx + ly * ly;
  if (n2 > 0) { const R n = sqrt(n2); lx = lx / n; ly = ly / n; }
  const R m2 = dist2<R>(w1x + lx, w1y + ly, w2x, w2y) < dist2<R>(w1x - lx, w1y - ly, w2x, w2y) ? (R)1 : (R)-1;
  const R mag = (R)0.5 * fmin(du, dv);  // sc:343-345: |W0W1| = |u|, |W1W2| = |v|
  tx = mag * (m2 * lx); ty = mag * (m2 * ly);
}

// Tangent of waypoint i of one path (x, y: its W waypoints).  sc:371-377, 606-638.
template <class R>
__device__ __forceinline__ void waypoint_tangent(const R* x, const R* y, int W, int i, R start_angle, R& ox, R& oy) {
  if (i == 0) {  // sc:371-373, 606-609
    R ang = start_angle;
    if (isnan(ang)) ang = atan2(y[1] - y[0], x[1] - x[0]);
    const R mag = (R)0.5 * dist2<R>(x[0], y[0], x[1], y[1]);
    R sn, cs; sincos(ang, &sn, &cs);
    ox = mag * cs; oy = mag * sn;
  } else if (i == W - 1) {  // sc:375-377
    const R dx = x[i] - x[i - 1], dy = y[i] - y[i - 1];
    const R mag = (R)0.5 * dist2<R>(x[i - 1], y[i - 1], x[i], y[i]);
    const R n = sqrt(dx * dx + dy * dy);
    ox = mag * (dx / n); oy = mag * (dy / n);
  } else {
    calc_tangent<R>(x[i - 1], y[i - 1], x[i], y[i], x[i + 1], y[i + 1], ox, oy);
  }
}

// sc:137-178: does segment l0 -> l1 cross segment k0 -> k1, and where (collinear / parallel: no, as the reference)
template <class R>
__device__ __forceinline__ bool seg_intersects(R l0x, R l0y, R l1x, R l1y, R k0x, R k0y, R k1x, R k1y, R& px, R& py) {
  const R dlx = l1x - l0x, dly = l1y - l0y, dkx = k1x - k0x, dky = k1y - k0y, ex = k0x - l0x, ey = k0y - l0y;
  const R a = ex * dly - ey * dlx;    // cross(k0 - l0, l1 - l0)
  const R b = dlx * dky - dly * dkx;  // cross(l1 - l0, k1 - k0)
  if (b == 0) return false;
  const R u = a / b;
  const R c = ex * dky - ey * dkx;    // cross(k0 - l0, k1 - k0)
  const R t = c / b;
  if (0 <= u && u <= 1 && 0 <= t && t <= 1) { px = t * dlx + l0x; py = t * dly + l0y; return true; }
  return false;
}

// sc:575-596 (k = 1, sc:612): cut the tangent T of waypoint W back, edge after edge, where W -> W + T or W -> W - T
// crosses an obstacle edge.  edges: [n][4] = (x0, y0, x1, y1) in the storage type S.
template <class R, class S>
__device__ __forceinline__ void shrink_tangent(R wx, R wy, const S* edges, int n, R& tx, R& ty) {
  for (int k = 0; k < n; ++k) {
    const R ax = (R)edges[4 * k], ay = (R)edges[4 * k + 1], bx = (R)edges[4 * k + 2], by = (R)edges[4 * k + 3];
    R px, py;
    if (seg_intersects<R>(wx + tx, wy + ty, wx, wy, ax, ay, bx, by, px, py)) { tx = px - wx; ty = py - wy; }
    if (seg_intersects<R>(wx - tx, wy - ty, wx, wy, ax, ay, bx, by, px, py)) { tx = wx - px; ty = wy - py; }
  }
}

template <class R> struct Carve {
  R* base; size_t off;
  __device__ __forceinline__ Carve(R* b) : base(b), off(0) {}
  __device__ __forceinline__ R* take(size_t n) { R* p = base + off; off += n; return p; }
};

template <class R> struct HermiteQ {  // q(s) of gen_vel_prof's 1-DOF path (piecewise_poly_path.cpp:364-375)
  R c0, c1, c2;
  __device__ __forceinline__ void init(R L, R vs, R ve) { c2 = vs; c0 = (ve - (R)2 * L + vs); c1 = (ve - vs - (R)3 * c0) / (R)2; }
  __device__ __forceinline__ R q(R s) const { return c0 * (s * s * s) + c1 * (s * s) + c2 * s; }
};

// Gauss-Legendre rules on [-1, 1] (positive halves)
static __constant__ double kG4X[2] = {0.3399810435848562648, 0.8611363115940525752};
static __constant__ double kG4W[2] = {0.6521451548625461426, 0.3478548451374538574};
static __constant__ double kG6X[3] = {0.2386191860831969086, 0.6612093864662645137, 0.9324695142031520278};
static __constant__ double kG6W[3] = {0.4679139345726910473, 0.3607615730481386076, 0.1713244923791703450};
// Kronrod extension of the 3-point Gauss rule: nodes x1 < x2 (= sqrt(3/5), the Gauss node) < x3 are the roots of
// P3(x) (x^4 - 10/9 x^2 + 155/891); weights for {0, x1, x2, x3} from exactness on 1, x^2, x^4, x^6
static __constant__ double kK7X[3] = {0.43424374934680265, 0.7745966692414834, 0.96049126870802026};
static __constant__ double kK7W[4] = {0.4509165386584757, 0.40139741477596125, 0.26848808986833372, 0.10465622602646719};
// 4-point rule in the node order the arc-length inversion uses (DESIGN.md "Arc-length inversion")
static __constant__ double kPGL4X[4] = {-0.8611363115940525752, -0.3399810435848562648, 0.3399810435848562648,
                                            0.8611363115940525752};
static __constant__ double kPGL4W[4] = {0.3478548451374538574, 0.6521451548625461426, 0.6521451548625461426,
                                            0.3478548451374538574};

// iRPROP- update of one parameter (DESIGN.md "New stages")
template <class R> __device__ __forceinline__ R rprop_step(R gv, R& gprev, R& step, R smin, R smax) {
  // selects, not branches (the same values: a NaN product fails both comparisons and leaves everything as it is)
  const R s = gv * gprev;
  const R up = fmin(step * (R)1.2, smax), dn = fmax(step * (R)0.5, smin);
  step = s > 0 ? up : (s < 0 ? dn : step);
  gv = s < 0 ? (R)0 : gv;
  gprev = gv;
  return gv > 0 ? -step : (gv < 0 ? step : (R)0);
}

// Power basis of one segment: a_k = C(d,k) Delta^k P_0 (curve) and h_k likewise for the hodograph
// control points d (P[j+1] - P[j])  (sc:1041-1053).
template <class R, int DEG>
__device__ __forceinline__ void power_basis(const R* cx, const R* cy, R* ax, R* ay, R* hx, R* hy) {
  constexpr int NC = DEG + 1;
  R dx[NC], dy[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) { dx[j] = cx[j]; dy[j] = cy[j]; }
  R hdx[DEG], hdy[DEG];
#pragma unroll
  for (int j = 0; j < DEG; ++j) { hdx[j] = (R)DEG * (dx[j + 1] - dx[j]); hdy[j] = (R)DEG * (dy[j + 1] - dy[j]); }
  R binom = 1;
#pragma unroll
  for (int k = 0; k <= DEG; ++k) {
    ax[k] = binom * dx[0]; ay[k] = binom * dy[0];
#pragma unroll
    for (int j = 0; j + k < DEG; ++j) { dx[j] = dx[j + 1] - dx[j]; dy[j] = dy[j + 1] - dy[j]; }
    binom = binom * (R)(DEG - k) / (R)(k + 1);
  }
  binom = 1;
#pragma unroll
  for (int k = 0; k < DEG; ++k) {
    hx[k] = binom * hdx[0]; hy[k] = binom * hdy[0];
#pragma unroll
    for (int j = 0; j + k < DEG - 1; ++j) { hdx[j] = hdx[j + 1] - hdx[j]; hdy[j] = hdy[j + 1] - hdy[j]; }
    binom = binom * (R)(DEG - 1 - k) / (R)(k + 1);
  }
}

// Basis values of B' and B'' of a quintic at cost node `node` of the 32-point rule mapped to [0,1]:
// out[0..5] = dB'/dP_j, out[6..11] = dB''/dP_j, out[12] = quadrature weight (DESIGN.md "New stages").
template <class R> __device__ __forceinline__ void lau_node_basis(R* out, int node) {
  const double t = 0.5 * GL32_X[node] + 0.5, s = 1 - t;
  const double b4[5] = {s * s * s * s, 4 * t * s * s * s, 6 * t * t * s * s, 4 * t * t * t * s, t * t * t * t};
  const double b3[4] = {s * s * s, 3 * t * s * s, 3 * t * t * s, t * t * t};
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const double lo = j > 0 ? b4[j - 1] : 0, hi = j < 5 ? b4[j] : 0;
    out[j] = (R)(5 * (lo - hi));
    const double a = j >= 2 ? b3[j - 2] : 0, b = (j >= 1 && j <= 4) ? b3[j - 1] : 0, c = j <= 3 ? b3[j] : 0;
    out[6 + j] = (R)(20 * (a - 2 * b + c));
  }
  out[12] = (R)(0.5 * GL32_W[node]);
}

// Geometry at toppra gridpoint i of one path: arc length q(s_i) (FP64, Hermite path of gen_vel_prof)
// -> segment (first j with q <= segend[j]) -> table sub-interval -> Newton on a 4-point Gauss-Legendre
// integral of |B'| from a cubic-Hermite start (DESIGN.md "Arc-length inversion") -> curve point and curvature (sc:1029).
//   segend: cumulative segment ends of this path (FP64); tab: its [nseg][101] cumulative table;
//   ax, ay: its [nseg][DEG+1] power-basis coefficients.
template <class R, int DEG>
__device__ __forceinline__ void gridpoint_geometry(int i, int N, double step, const double* segend, int nseg, const double* hqc,
                                                   const R* inv_len, int& sg, const unsigned char* coarse, int coarse_stride,
                                                   const R* tab, const R* ax, const R* ay,
                                                   const R* hx, const R* hy, R& bx, R& by, R& kap) {
  // hqc: {c0, c1, c2} of q(s) = c0 s^3 + c1 s^2 + c2 s; inv_len[j] = 100 / length of segment j; sg: the caller's
  // guess of the segment (its previous gridpoint's), updated
  constexpr int NC = DEG + 1;
  const double Ltot = segend[nseg - 1];
  const double sq = i == N ? 1.0 : (double)i * step;
  double q = hqc[0] * (sq * sq * sq) + hqc[1] * (sq * sq) + hqc[2] * sq;
  q = fmin(fmax(q, 0.0), Ltot);
  // first j with q <= segend[j]: a thread's gridpoints advance along the path, so walk from the last answer
  while (sg > 0 && q <= segend[sg - 1]) --sg;
  while (sg < nseg - 1 && q > segend[sg]) ++sg;
  const R r = (R)(q - (sg > 0 ? segend[sg - 1] : 0.0));
  const R* c = tab + (size_t)sg * kTable;
  // largest kl in [0, 99] with c[kl] <= r: the table is close to linear, so start from the proportional
  // guess and walk (two independent loads, then rarely more) instead of seven dependent bisection loads
  int kl;
  if (coarse) {  // coarse inverse table (arc_kernel.cu): 32 cells per segment, linear interpolation of the piece index
    const R pc = r * inv_len[sg] * (R)0.32;  // position in cells: r / length * 32
    const int j = max(0, min(31, (int)pc));
    const unsigned char* cz = coarse + (size_t)sg * coarse_stride;
    const int k0 = cz[j], k1 = cz[j + 1];
    kl = k0 + (int)((pc - (R)j) * (R)(k1 - k0));
  } else {
    kl = (int)(r * inv_len[sg]);
  }
  kl = max(1, min(kSub - 2, kl));
  // four independent loads around the guess decide almost every gridpoint in one memory round trip
  // (the dependent walk that follows was 21 % of this kernel's stall samples, profiles/r1h_arc_kernel_f64.lines.txt)
  R ck, ck1;
  {
    const R cm = c[kl - 1], c0 = c[kl], c1 = c[kl + 1], c2 = c[kl + 2];
    if (r < c0) { --kl; ck = cm; ck1 = c0; }
    else if (r >= c1) { ++kl; ck = c1; ck1 = c2; }
    else { ck = c0; ck1 = c1; }
    if (ck > r || (ck1 <= r && kl < kSub - 1)) {  // the guess was off by more than one piece: walk
      while (kl > 0 && c[kl] > r) --kl;
      while (kl < kSub - 1 && c[kl + 1] <= r) ++kl;
      ck = c[kl]; ck1 = c[kl + 1];
    }
  }
  const R h = (R)0.01, a0 = fmin((R)kl * h, (R)1);
  const R target = r - ck, ds = ck1 - ck;
  const R* sax = ax + NC * sg; const R* say = ay + NC * sg;
  R kx[DEG], ky[DEG];  // B' in power basis (the hodograph coefficients the table was integrated with)
#pragma unroll
  for (int j = 0; j < DEG; ++j) { kx[j] = hx[DEG * sg + j]; ky[j] = hy[DEG * sg + j]; }
  auto speed2 = [&](R u) {
    R dx = kx[DEG - 1], dy = ky[DEG - 1];
#pragma unroll
    for (int j = DEG - 2; j >= 0; --j) { dx = dx * u + kx[j]; dy = dy * u + ky[j]; }
    return fma(dy, dy, fma(dx, dx, speed_floor<R>()));  // > 0 also at a cusp: no guard around the rsqrt
  };
  auto speed = [&](R u) { const R n2 = speed2(u); return n2 * rsqrt_n1<R>(n2); };
  auto inv_speed = [&](R u) { return rsqrt_n1<R>(speed2(u)); };
  // Start: the cubic Hermite inverse u(s) on [c_k, c_k+1] (end values 0 and h, end slopes 1/|B'|): its error
  // is ~ds^4 |d4u/ds4| / 384, i.e. ~1e-8 h for a smooth piece against ~1e-3 h for linear interpolation, so
  // that ONE Newton step on the quadrature normally reaches the root; near a cusp the clamps keep the
  // iterate inside the piece.
  R du;
  bool smooth;
  {
    const R tau = fmin(fmax(target * fast_rcp<R>(ds), (R)0), (R)1), om = (R)1 - tau;
    const R r0 = inv_speed(a0), r1 = inv_speed(fmin(a0 + h, (R)1));
    du = (tau * om) * ds * (om * r0 - tau * r1) + h * (tau * tau) * ((R)3 - (R)2 * tau);
    du = fmin(fmax(du, (R)0), h);
    const R m0 = r0 * ds, m1 = r1 * ds;  // mean speed / end speed
    smooth = m0 >= (R)0.8 * h && m0 <= (R)1.25 * h && m1 >= (R)0.8 * h && m1 <= (R)1.25 * h;
  }
  // One step is accepted only where the iteration is known to be quadratic and the step small:
  //   (a) the piece is smooth (end speeds within 25 % of the mean speed ds/h: |B'| has no dip inside, so the
  //       4-point quadrature F and its Newton derivative |B'| are consistent to ~1e-8),
  //   (b) the correction was small, and (c) the predicted next correction C delta^2 is below tolerance,
  //       C = |d|B'|/du| / (2 |B'|) = |B'.B''| / (2 |B'|^2) where the step lands.
  // Otherwise (near a cusp F is a poor quadrature, its derivative is not |B'| and Newton converges only
  // linearly) the loop runs to |delta| <= tolS like the restated definition does.
  const bool f64 = sizeof(R) == 8;
  const R tolS = (f64 ? (R)1e-12 : (R)2e-5) * h, tolB = (f64 ? (R)1e-5 : (R)1e-2) * h, tolC = (f64 ? (R)1e-12 : (R)1e-5) * h;
#pragma unroll 1
  for (int nit = 0; nit < kInvNewtonMax; ++nit) {
    const R half = (R)0.5 * du, mid = a0 + half;
    R F = 0;
#pragma unroll
    for (int gq = 0; gq < 4; ++gq) F += (R)kPGL4W[gq] * speed(mid + half * (R)kPGL4X[gq]);
    F *= half;
    const R ue = a0 + du;
    R x1 = kx[DEG - 1], y1 = ky[DEG - 1];
#pragma unroll
    for (int j = DEG - 2; j >= 0; --j) { x1 = x1 * ue + kx[j]; y1 = y1 * ue + ky[j]; }
    R x2 = (R)(DEG - 1) * kx[DEG - 1], y2 = (R)(DEG - 1) * ky[DEG - 1];
#pragma unroll
    for (int j = DEG - 3; j >= 0; --j) { x2 = x2 * ue + (R)(j + 1) * kx[j + 1]; y2 = y2 * ue + (R)(j + 1) * ky[j + 1]; }
    const R rn = rsqrt_n1<R>(fma(y1, y1, fma(x1, x1, speed_floor<R>())));
    const R delta = (F - target) * rn;
    du = fmin(fmax(du - delta, (R)0), h);
    const R ad = fabs(delta);
    if (ad <= tolS) break;
    const R C = (R)0.5 * fabs(x1 * x2 + y1 * y2) * (rn * rn);
    if (smooth && ad <= tolB && C * ad * ad <= tolC) break;
  }
  const R u = fmin(a0 + du, (R)1);
  bx = sax[DEG]; by = say[DEG];
#pragma unroll
  for (int j = DEG - 1; j >= 0; --j) { bx = bx * u + sax[j]; by = by * u + say[j]; }
  R x1 = kx[DEG - 1], y1 = ky[DEG - 1];
#pragma unroll
  for (int j = DEG - 2; j >= 0; --j) { x1 = x1 * u + kx[j]; y1 = y1 * u + ky[j]; }
  R x2 = (R)(DEG - 1) * kx[DEG - 1], y2 = (R)(DEG - 1) * ky[DEG - 1];
#pragma unroll
  for (int j = DEG - 3; j >= 0; --j) { x2 = x2 * u + (R)(j + 1) * kx[j + 1]; y2 = y2 * u + (R)(j + 1) * ky[j + 1]; }
  const R rn = fast_rsqrt<R>(fma(y1, y1, fma(x1, x1, speed_floor<R>())));  // cubed below: full precision
  kap = (x1 * y2 - y1 * x2) * (rn * rn * rn);
}

}  // namespace scb
