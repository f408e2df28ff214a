// ORACLE — TEST INFRASTRUCTURE ONLY (see ora_spline.hpp header).
//
// PARITY UNPINNED: the stages in this file (quintic builder A5, cost + analytic gradient A6,
// RPROP optimiser A7 of SURVEY.md §8) have NO implementation and NO test in the reference.
// They are named by the north star and specified by the paper shipped with the reference
// (reference/lau09iros.pdf §III-A, §IV-B, Algorithm 1).  This file is the frozen definition
// (DESIGN.md "New stages"); it is validated by properties only: interpolation and C2 joins,
// gradient vs central finite differences, monotone best-so-far cost.
#pragma once
#include "gl32.h"
#include "ora_spline.hpp"

namespace ora {

struct LauOpt {
  int iters = 32;
  int move_waypoints = 0;
  double lambda = 0.05;
  double e_init = 1.0, e_min = 0.25, e_max = 4.0;
  double step_e = 0.1, step_w = 0.05;
  double w_radius = 0.25;
};

// Frozen per-problem data computed once from the INPUT waypoints: base tangents (A1,
// sc:343-377 — direction and magnitude 1/2 min neighbour distance) and the join weights.
template <class R> struct LauModel {
  std::vector<V2<R>> W0;     // input waypoints
  std::vector<V2<R>> Bt;     // base tangent per waypoint; T_i = e_i * Bt_i   (Lau09 §III-A)
  std::vector<R> mag;        // |Bt_i|
  std::vector<R> alpha;      // weight of the PREVIOUS segment's cubic at inner waypoint i:
                             // anti-proportional to leg length => d_next / (d_prev + d_next)
};

template <class R> inline LauModel<R> lau_model(const std::vector<V2<R>>& path, R start_angle) {
  LauModel<R> m;
  m.W0 = path;
  std::vector<V2<R>> tang;
  from_path(path, start_angle, false, EVAL_DC, (std::vector<V2<R>>*)nullptr, &tang);
  m.Bt = tang;
  const size_t W = path.size();
  m.mag.resize(W); m.alpha.assign(W, (R)0);
  for (size_t i = 0; i < W; ++i) m.mag[i] = std::sqrt(tang[i].x * tang[i].x + tang[i].y * tang[i].y);
  for (size_t i = 1; i + 1 < W; ++i) {
    const R dp = pt_dist(path[i - 1], path[i]), dn = pt_dist(path[i], path[i + 1]);
    m.alpha[i] = dn / (dp + dn);
  }
  return m;
}

// A5.  Control points of every quintic segment from (waypoints Wp, elongations e).
//   T_i = e_i Bt_i
//   A_i = alpha_i [6(W_{i-1}-W_i) + 4T_i + 2T_{i-1}] + (1-alpha_i) [6(W_{i+1}-W_i) - 4T_i - 2T_{i+1}]
//         (second derivatives at W_i of the two neighbouring cubic Beziers that have end
//          derivatives T; one-sided at the two ends)
//   P0 = W_i, P1 = W_i + T_i/5, P2 = A_i/20 + 2P1 - W_i,
//   P5 = W_{i+1}, P4 = W_{i+1} - T_{i+1}/5, P3 = A_{i+1}/20 + 2P4 - W_{i+1}
template <class R>
inline void lau_second_derivs(const LauModel<R>& m, const std::vector<V2<R>>& Wp, const std::vector<R>& e,
                              std::vector<V2<R>>& T, std::vector<V2<R>>& A) {
  const int W = (int)Wp.size();
  T.resize(W); A.resize(W);
  for (int i = 0; i < W; ++i) T[i] = e[i] * m.Bt[i];
  for (int i = 0; i < W; ++i) {
    V2<R> prev = {0, 0}, next = {0, 0};
    if (i > 0) prev = (R)6 * (Wp[i - 1] - Wp[i]) + (R)4 * T[i] + (R)2 * T[i - 1];
    if (i + 1 < W) next = (R)6 * (Wp[i + 1] - Wp[i]) - (R)4 * T[i] - (R)2 * T[i + 1];
    if (i == 0) A[i] = next;
    else if (i + 1 == W) A[i] = prev;
    else A[i] = m.alpha[i] * prev + ((R)1 - m.alpha[i]) * next;
  }
}

template <class R>
inline Spline<R> lau_build(const LauModel<R>& m, const std::vector<V2<R>>& Wp, const std::vector<R>& e) {
  const int W = (int)Wp.size();
  std::vector<V2<R>> T, A;
  lau_second_derivs(m, Wp, e, T, A);
  Spline<R> sp;
  sp.ctrl.resize(W - 1);
  for (int i = 0; i + 1 < W; ++i) {
    const V2<R> P0 = Wp[i], P5 = Wp[i + 1];
    const V2<R> P1 = P0 + ((R)1 / (R)5) * T[i];
    const V2<R> P4 = P5 - ((R)1 / (R)5) * T[i + 1];
    const V2<R> P2 = ((R)1 / (R)20) * A[i] + (R)2 * P1 - P0;
    const V2<R> P3 = ((R)1 / (R)20) * A[i + 1] + (R)2 * P4 - P5;
    sp.ctrl[i] = {P0, P1, P2, P3, P4, P5};
  }
  return sp;
}

// Basis values at the 32 Gauss–Legendre nodes mapped to [0,1]:  B'(u) = sum_j c1[j] P_j,
// B''(u) = sum_j c2[j] P_j for a quintic; wq = w/2.
struct LauNodes {
  double u[32], wq[32], c1[32][6], c2[32][6];
  LauNodes() {
    for (int q = 0; q < 32; ++q) {
      const double t = 0.5 * GL32_X[q] + 0.5, s = 1 - t;
      u[q] = t; wq[q] = 0.5 * GL32_W[q];
      const double b4[5] = {s * s * s * s, 4 * t * s * s * s, 6 * t * t * s * s, 4 * t * t * t * s, t * t * t * t};
      const double b3[4] = {s * s * s, 3 * t * s * s, 3 * t * t * s, t * t * t};
      for (int j = 0; j < 6; ++j) {
        const double lo = j > 0 ? b4[j - 1] : 0, hi = j < 5 ? b4[j] : 0;
        c1[q][j] = 5 * (lo - hi);
        const double a = j >= 2 ? b3[j - 2] : 0, b = (j >= 1 && j <= 4) ? b3[j - 1] : 0, c = j <= 3 ? b3[j] : 0;
        c2[q][j] = 20 * (a - 2 * b + c);
      }
    }
  }
};
inline const LauNodes& lau_nodes() { static LauNodes n; return n; }

// A6.  J = sum_seg int_0^1 [ kappa(u)^2 |B'(u)| + lambda |B'(u)| ] du
//        = sum_seg sum_q wq [ c^2 / n^5 + lambda n ],  c = x'y'' - y'x'',  n = |B'|
// by 32-point Gauss–Legendre per segment, and dJ/dP_j for every control point.
// Summation order (mirrored by the CUDA kernel): nodes q = 0..31 in table order within a
// segment, segments in order.
template <class R>
inline R lau_cost_grad(const Spline<R>& sp, R lambda, std::vector<std::vector<V2<R>>>* grad) {
  const LauNodes& N = lau_nodes();
  const int nseg = sp.n_segments();
  if (grad) grad->assign(nseg, std::vector<V2<R>>(6, V2<R>{0, 0}));
  R J = 0;
  for (int s = 0; s < nseg; ++s) {
    const std::vector<V2<R>>& P = sp.ctrl[s];
    R Js = 0;
    for (int q = 0; q < 32; ++q) {
      R x1 = 0, y1 = 0, x2 = 0, y2 = 0;
      for (int j = 0; j < 6; ++j) {
        x1 += (R)N.c1[q][j] * P[j].x; y1 += (R)N.c1[q][j] * P[j].y;
        x2 += (R)N.c2[q][j] * P[j].x; y2 += (R)N.c2[q][j] * P[j].y;
      }
      const R n2 = x1 * x1 + y1 * y1;
      const R n = std::sqrt(n2);
      const R c = x1 * y2 - y1 * x2;
      const R in5 = (R)1 / (n2 * n2 * n);
      const R w = (R)N.wq[q];
      Js += w * (c * c * in5 + lambda * n);
      if (grad) {
        const R k1 = (R)2 * c * in5;            // d/dc of c^2/n^5
        const R k2 = (R)5 * c * c * in5 / n2;   // 5 c^2 / n^7
        const R k3 = lambda / n;
        const R fx1 = k1 * y2 - k2 * x1 + k3 * x1;
        const R fy1 = -k1 * x2 - k2 * y1 + k3 * y1;
        const R fx2 = -k1 * y1, fy2 = k1 * x1;
        for (int j = 0; j < 6; ++j) {
          (*grad)[s][j].x += w * (fx1 * (R)N.c1[q][j] + fx2 * (R)N.c2[q][j]);
          (*grad)[s][j].y += w * (fy1 * (R)N.c1[q][j] + fy2 * (R)N.c2[q][j]);
        }
      }
    }
    J += Js;
  }
  return J;
}

// Chain rule from control-point gradients to parameters (transposes lau_build exactly).
template <class R>
inline void lau_backprop(const LauModel<R>& m, const std::vector<std::vector<V2<R>>>& g,
                         std::vector<V2<R>>& gW, std::vector<R>& ge) {
  const int W = (int)m.W0.size();
  std::vector<V2<R>> gT(W, V2<R>{0, 0}), gA(W, V2<R>{0, 0});
  gW.assign(W, V2<R>{0, 0});
  for (int i = 0; i + 1 < W; ++i) {
    const std::vector<V2<R>>& s = g[i];
    // P0 = W_i; P1 = W_i + T_i/5; P2 = A_i/20 + W_i + 2T_i/5
    gW[i] = gW[i] + (s[0] + s[1] + s[2]);
    gT[i] = gT[i] + ((R)0.2 * s[1] + (R)0.4 * s[2]);
    gA[i] = gA[i] + (R)0.05 * s[2];
    // P5 = W_{i+1}; P4 = W_{i+1} - T_{i+1}/5; P3 = A_{i+1}/20 + W_{i+1} - 2T_{i+1}/5
    gW[i + 1] = gW[i + 1] + (s[3] + s[4] + s[5]);
    gT[i + 1] = gT[i + 1] - ((R)0.4 * s[3] + (R)0.2 * s[4]);
    gA[i + 1] = gA[i + 1] + (R)0.05 * s[3];
  }
  for (int i = 0; i < W; ++i) {
    const R ap = (i == 0) ? (R)0 : (i + 1 == W) ? (R)1 : m.alpha[i];  // weight of "prev"
    const R an = (i + 1 == W) ? (R)0 : (i == 0) ? (R)1 : ((R)1 - m.alpha[i]);
    const V2<R> a = gA[i];
    if (i > 0) { gW[i - 1] = gW[i - 1] + ((R)6 * ap) * a; gT[i - 1] = gT[i - 1] + ((R)2 * ap) * a; }
    if (i + 1 < W) { gW[i + 1] = gW[i + 1] + ((R)6 * an) * a; gT[i + 1] = gT[i + 1] - ((R)2 * an) * a; }
    gW[i] = gW[i] - ((R)6 * (ap + an)) * a;
    gT[i] = gT[i] + ((R)4 * (ap - an)) * a;
  }
  ge.resize(W);
  for (int i = 0; i < W; ++i) ge[i] = gT[i].x * m.Bt[i].x + gT[i].y * m.Bt[i].y;
}

template <class R> struct LauResult {
  Spline<R> spline;
  std::vector<V2<R>> Wp;
  std::vector<R> e;
  R cost;
  std::vector<R> history;  // J at every evaluation (K+1 entries); +inf where the iterate collides
  bool collision_free = true;  // false: every iterate crossed an obstacle edge (cost = +inf)
};

// Collision test of one spline (Lau09 IV-B): every segment cut into M chords B(k/M) -> B((k+1)/M), every chord
// against every obstacle edge with the reference's segment test (intersects, sc:142-178).
template <class R> inline bool lau_collides(const Spline<R>& sp, const std::vector<Edge<R>>& edges, int M) {
  for (const auto& seg : sp.ctrl) {
    V2<R> a = seg[0];
    for (int k = 1; k <= M; ++k) {
      const R u = k == M ? (R)1 : (R)k / (R)M;
      std::vector<V2<R>> t = seg;
      for (size_t r = 1; r < t.size(); ++r)
        for (size_t i = 0; i + r < t.size(); ++i) t[i] = {((R)1 - u) * t[i].x + u * t[i + 1].x, ((R)1 - u) * t[i].y + u * t[i + 1].y};
      V2<R> pt;
      for (const auto& e : edges)
        if (seg_intersects(a, t[0], e.a, e.b, pt)) return true;
      a = t[0];
    }
  }
  return false;
}

// A7.  iRPROP- with a fixed budget of K gradient iterations (Lau09 Alg. 1 adapted to use the
// sign of the analytic gradient): step *1.2 while the sign holds, *0.5 and skip the move on
// a sign change; parameters clamped to their boxes; best-so-far parameters are returned.
// A gradient component within 1e-9 of the largest one of its class counts as zero (so that
// rounding noise on exactly-stationary parameters cannot decide a move).
//
// Collision term (Lau09 IV-B: "if the trajectory is not collision-free according to the obstacle map, its cost is set to
// infinity"; there the optimiser is derivative-free and simply rejects such a trial step).  With an edge list and
// M = collision_chords > 0, every iterate is tested (lau_collides) and a colliding iterate has J = +inf: it can never
// become the best-so-far, and it is REJECTED - the parameters return to the best collision-free iterate found so far,
// every step size is halved and the gradient memory cleared, so the next trial is a shorter step from there.  While no
// collision-free iterate exists yet (the initial spline collides) the gradient steps go on.  The returned spline is the
// best collision-free iterate; if there is none, collision_free = false and cost = +inf.
template <class R> inline LauResult<R> lau_optimise(const LauModel<R>& m, const LauOpt& o,
                                                    const std::vector<Edge<R>>* edges = nullptr, int chords = 0) {
  const int W = (int)m.W0.size();
  std::vector<V2<R>> Wp = m.W0;
  std::vector<R> e(W, (R)o.e_init);
  std::vector<R> se(W, (R)o.step_e), pe(W, (R)0);
  std::vector<V2<R>> sw(W), pw(W, V2<R>{0, 0});
  for (int i = 0; i < W; ++i) sw[i] = {(R)o.step_w * m.mag[i], (R)o.step_w * m.mag[i]};
  LauResult<R> best;
  auto rprop1 = [](R g, R& gprev, R& step, R smin, R smax) -> R {
    const R s = g * gprev;
    if (s > 0) step = std::min(step * (R)1.2, smax);
    else if (s < 0) { step = std::max(step * (R)0.5, smin); g = 0; }
    gprev = g;
    return g > 0 ? -step : g < 0 ? step : (R)0;
  };
  std::vector<std::vector<V2<R>>> g;
  std::vector<V2<R>> gW; std::vector<R> ge;
  for (int it = 0; it <= o.iters; ++it) {
    Spline<R> sp = lau_build(m, Wp, e);
    R J = lau_cost_grad(sp, (R)o.lambda, &g);
    const bool hit = edges && chords > 0 && !edges->empty() && lau_collides(sp, *edges, chords);
    if (hit) J = (R)INFINITY;
    best.history.push_back(J);
    if (it == 0 || J < best.cost) { best.cost = J; best.spline = sp; best.Wp = Wp; best.e = e; }
    if (it == o.iters) break;
    if (hit && best.cost < (R)INFINITY) {  // rejected trial: back to the best collision-free iterate, shorter steps
      Wp = best.Wp; e = best.e;
      for (int i = 0; i < W; ++i) {
        se[i] = std::max(se[i] * (R)0.5, (R)1e-6); pe[i] = 0;
        const R smin = (R)1e-6 * m.mag[i];
        sw[i] = {std::max(sw[i].x * (R)0.5, smin), std::max(sw[i].y * (R)0.5, smin)}; pw[i] = {0, 0};
      }
      continue;
    }
    lau_backprop(m, g, gW, ge);
    R gemax = 0, gwmax = 0;
    for (int i = 0; i < W; ++i) gemax = std::max(gemax, std::fabs(ge[i]));
    for (int i = 1; i + 1 < W; ++i) gwmax = std::max(gwmax, std::max(std::fabs(gW[i].x), std::fabs(gW[i].y)));
    const R te = (R)1e-9 * gemax, tw = (R)1e-9 * gwmax;
    for (int i = 0; i < W; ++i) {
      R gi = std::fabs(ge[i]) <= te ? (R)0 : ge[i];
      e[i] += rprop1(gi, pe[i], se[i], (R)1e-6, (R)1.0);
      e[i] = std::min((R)o.e_max, std::max((R)o.e_min, e[i]));
    }
    if (o.move_waypoints) {
      for (int i = 1; i + 1 < W; ++i) {
        R gx = std::fabs(gW[i].x) <= tw ? (R)0 : gW[i].x, gy = std::fabs(gW[i].y) <= tw ? (R)0 : gW[i].y;
        const R smin = (R)1e-6 * m.mag[i], smax = (R)0.5 * m.mag[i];
        Wp[i].x += rprop1(gx, pw[i].x, sw[i].x, smin, smax);
        Wp[i].y += rprop1(gy, pw[i].y, sw[i].y, smin, smax);
        const R r = (R)o.w_radius;
        Wp[i].x = std::min(m.W0[i].x + r, std::max(m.W0[i].x - r, Wp[i].x));
        Wp[i].y = std::min(m.W0[i].y + r, std::max(m.W0[i].y - r, Wp[i].y));
      }
    }
  }
  best.collision_free = best.cost < (R)INFINITY;
  return best;
}

}  // namespace ora
