// ORACLE — TEST INFRASTRUCTURE ONLY.  Not shipped, not linked by the product library.
//
// CPU restatement of the spline side of sea-current's planning hot path
// (reference: /root/reference/sea_current.hpp, cited below as "sc:LINE").
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may compile, link, import or execute anything under oracle/.
//
// Template parameter R is the scalar the reference's `float` is mapped to:
//   R = float   reference-faithful arithmetic (the reference computes the spline side in
//               float32 with double Gauss–Legendre tables, sc:768-880)
//   R = double  the FP64 parity oracle for the CUDA FP64 mode.
// Places where the reference mixes float and double are kept as written (casts through
// double that are identities when R = double).
//
// Pinned by: examples/output.json arclength block (sc:343-377, 599-683, 767-896) — see
// tests/test_oracle_golden.py.  chebfit's float QR (Eigen HouseholderQR, not in the tree)
// is restated from the published algorithm; pinned only loosely by output.json pos_x/pos_y.
#pragma once
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstddef>
#include <stdexcept>
#include <vector>

namespace ora {

template <class R> struct V2 { R x, y; };
template <class R> inline V2<R> operator+(V2<R> a, V2<R> b) { return {a.x + b.x, a.y + b.y}; }
template <class R> inline V2<R> operator-(V2<R> a, V2<R> b) { return {a.x - b.x, a.y - b.y}; }
template <class R> inline V2<R> operator*(R k, V2<R> a) { return {k * a.x, k * a.y}; }
template <class R> inline V2<R> neg(V2<R> a) { return {-a.x, -a.y}; }

// sc:86-88  pt_dist: sqrt(pow(dx,2)+pow(dy,2)); std::pow(float,int) promotes to double.
template <class R> inline R pt_dist(V2<R> u, V2<R> v = {0, 0}) {
  double dx = (double)(v.x - u.x), dy = (double)(v.y - u.y);
  return (R)std::sqrt(dx * dx + dy * dy);
}

// Eigen's Vector2f::normalized(): v / sqrt(squaredNorm) when the norm is > 0.
template <class R> inline V2<R> normalized(V2<R> v) {
  R n2 = v.x * v.x + v.y * v.y;
  if (n2 > (R)0) { R n = std::sqrt(n2); return {v.x / n, v.y / n}; }
  return v;
}

// ---------------------------------------------------------------- A1: tangent heuristics
// sc:343-345
template <class R> inline R tangent_magnitude(V2<R> W0, V2<R> W1, V2<R> W2) {
  return (R)(0.5 * (double)std::min(pt_dist(W0, W1), pt_dist(W1, W2)));
}

// sc:347-369.  `clamp_acos` is this repo's one conscious fix (the
// reference feeds an unclamped quotient to acos and can produce NaN for collinear legs).
template <class R> inline V2<R> calc_tangent(V2<R> W0, V2<R> W1, V2<R> W2, bool clamp_acos = true) {
  const V2<R> u = W0 - W1, v = W2 - W1;
  R u_dot_v = u.x * v.x + u.y * v.y;
  const R denom = pt_dist(u) * pt_dist(v);
  R arg = u_dot_v / denom;
  if (clamp_acos) arg = std::max((R)-1, std::min((R)1, arg));
  const R theta = std::acos(arg) / 2;
  const R offset = std::atan2(u.y, u.x);
  const R test_offset = std::atan2(v.y, v.x);
  int mult = 1;
  if (test_offset - offset < 0) mult = -1;  // no angle wrapping in the reference (sc:357-358)
  V2<R> l90 = {std::sin(offset + (mult * theta)), -std::cos(offset + (mult * theta))};
  l90 = normalized(l90);
  int mult2 = -1;
  if (pt_dist(W1 + l90, W2) < pt_dist(W1 + neg(l90), W2)) mult2 = 1;
  return tangent_magnitude(W0, W1, W2) * ((R)mult2 * l90);
}

// sc:371-373
template <class R> inline V2<R> calc_start_tangent(V2<R> W0, V2<R> W1, R theta) {
  return tangent_magnitude(W0, W1, W0) * V2<R>{std::cos(theta), std::sin(theta)};
}
// sc:375-377
template <class R> inline V2<R> calc_end_tangent(V2<R> W1, V2<R> W2) {
  return tangent_magnitude(W1, W2, W1) * normalized(W2 - W1);
}

// ---------------------------------------------------------------- A3: Bézier evaluation
// sc:1096-1106  omega_table: omega^k, omega = exp(-2*pi*i/(degree+1)); the reference's
// pow(complex<float>, size_t) promotes to complex<double>, so build in double and narrow.
template <class R> inline std::vector<std::complex<R>> omega_table(int degree) {
  std::vector<std::complex<R>> om(degree + 1);
  om[0] = std::complex<R>(1, 0);
  // exp((complex<float>(pi) * -2i) / complex<float>(degree+1)) is evaluated in float
  const std::complex<R> base = std::exp(std::complex<R>(0, (R)-2 * (R)3.14159265358979323846) /
                                        std::complex<R>((R)(degree + 1), 0));
  for (int i = 1; i <= degree; ++i) {
    std::complex<double> b((double)base.real(), (double)base.imag());
    std::complex<double> p = std::pow(b, (double)i);
    om[i] = std::complex<R>((R)p.real(), (R)p.imag());
  }
  return om;
}

// sc:700-716  Q = ifft(ctrl) (Eigen::FFT<float>::inv, scaled by 1/n): plain inverse DFT.
template <class R>
inline void bf_coefficients(const std::vector<V2<R>>& ctrl, std::vector<std::complex<R>>& Qx,
                            std::vector<std::complex<R>>& Qy) {
  const int n = (int)ctrl.size();
  Qx.assign(n, {}); Qy.assign(n, {});
  for (int k = 0; k < n; ++k) {
    std::complex<double> sx = 0, sy = 0;
    for (int j = 0; j < n; ++j) {
      double ang = 2.0 * 3.14159265358979323846 * (double)((j * k) % n) / (double)n;
      std::complex<double> w(std::cos(ang), std::sin(ang));
      sx += (double)ctrl[j].x * w; sy += (double)ctrl[j].y * w;
    }
    Qx[k] = std::complex<R>((R)(sx.real() / n), (R)(sx.imag() / n));
    Qy[k] = std::complex<R>((R)(sy.real() / n), (R)(sy.imag() / n));
  }
}

// sc:721-748  Bernstein–Fourier evaluation, B(s) = Re sum_k Q_k (1 + s(omega^k - 1))^deg.
template <class R>
inline V2<R> bezier_point_bf(const std::vector<std::complex<R>>& Qx, const std::vector<std::complex<R>>& Qy,
                             const std::vector<std::complex<R>>& om, int degree, R s) {
  const std::complex<R> one(1, 0);
  R bx = 0, by = 0;
  for (int k = 0; k <= degree; ++k) {
    std::complex<R> tmp = std::pow(one + s * (om[k] - one), degree);
    bx += (Qx[k] * tmp).real();
    by += (Qy[k] * tmp).real();
  }
  return {bx, by};
}

// The same polynomial in Bernstein form by de Casteljau (what a sane implementation does;
// used to cross-check the BF restatement and as the "restated-fast" CPU variant).
template <class R> inline V2<R> bezier_point_dc(const V2<R>* ctrl, int degree, R s) {
  V2<R> tmp[8];
  for (int i = 0; i <= degree; ++i) tmp[i] = ctrl[i];
  for (int r = 1; r <= degree; ++r)
    for (int i = 0; i <= degree - r; ++i)
      tmp[i] = {tmp[i].x + s * (tmp[i + 1].x - tmp[i].x), tmp[i].y + s * (tmp[i + 1].y - tmp[i].y)};
  return tmp[0];
}

// sc:1041-1053 hodograph control points: degree * (P[j+1] - P[j]).
template <class R> inline std::vector<V2<R>> hodograph_ctrl(const std::vector<V2<R>>& c) {
  const int deg = (int)c.size() - 1;
  std::vector<V2<R>> d(deg);
  for (int j = 0; j < deg; ++j) d[j] = (R)deg * (c[j + 1] - c[j]);
  return d;
}

// One spline = list of segments, each a control polygon (sc:388-393 `ctrl_pts`).
template <class R> struct Spline {
  std::vector<std::vector<V2<R>>> ctrl;  // [segment][degree+1]
  int degree() const { return (int)ctrl[0].size() - 1; }
  int n_segments() const { return (int)ctrl.size(); }
};

enum EvalMode { EVAL_BF = 0, EVAL_DC = 1, EVAL_POW = 2 };

// Power-basis form of the same polynomial, a_k = C(n,k) * (k-th forward difference of P at 0),
// evaluated by Horner.  This is the arithmetic the CUDA kernels use.
template <class R> inline void bezier_power_coeffs(const V2<R>* ctrl, int degree, V2<R>* a) {
  V2<R> d[8];
  for (int i = 0; i <= degree; ++i) d[i] = ctrl[i];
  R binom = 1;
  for (int k = 0; k <= degree; ++k) {
    a[k] = binom * d[0];
    for (int j = 0; j + k < degree; ++j) d[j] = d[j + 1] - d[j];
    binom = binom * (R)(degree - k) / (R)(k + 1);
  }
}
template <class R> inline V2<R> horner(const V2<R>* a, int degree, R u) {
  V2<R> v = a[degree];
  for (int k = degree - 1; k >= 0; --k) v = {v.x * u + a[k].x, v.y * u + a[k].y};
  return v;
}

template <class R>
inline void eval_curve(const std::vector<V2<R>>& ctrl, const std::vector<R>& pos, int mode,
                       std::vector<V2<R>>& out) {
  const int deg = (int)ctrl.size() - 1;
  out.resize(pos.size());
  if (mode == EVAL_BF) {
    std::vector<std::complex<R>> Qx, Qy;
    bf_coefficients(ctrl, Qx, Qy);
    auto om = omega_table<R>(deg);
    for (size_t i = 0; i < pos.size(); ++i) out[i] = bezier_point_bf(Qx, Qy, om, deg, pos[i]);
  } else if (mode == EVAL_POW) {
    V2<R> a[8];
    bezier_power_coeffs(ctrl.data(), deg, a);
    for (size_t i = 0; i < pos.size(); ++i) out[i] = horner(a, deg, pos[i]);
  } else {
    for (size_t i = 0; i < pos.size(); ++i) out[i] = bezier_point_dc(ctrl.data(), deg, pos[i]);
  }
}

// ---------------------------------------------------------------- N3: tangent clipping against obstacle edges
// One obstacle edge (x0, y0) -> (x1, y1).  The reference keeps them as obstacle.lines of every obstacle of the
// planning_space (sc:193-284); shrink_tangent visits obstacles and their lines in order (sc:580-581), which is the
// order of this flat list.
template <class R> struct Edge { V2<R> a, b; };

// sc:137-178: intersection of segment l = (l0, l1) with segment k = (k0, k1); collinear and parallel cases are
// "no intersection" exactly as in the reference.
template <class R> inline bool seg_intersects(V2<R> l0, V2<R> l1, V2<R> k0, V2<R> k1, V2<R>& pt) {
  auto cross = [](V2<R> u, V2<R> v) { return u.x * v.y - u.y * v.x; };  // sc:137-139
  const R a = cross(k0 - l0, l1 - l0);
  const R b = cross(l1 - l0, k1 - k0);
  if (b == 0) return false;  // sc:145-162: a == 0 (collinear) and a != 0 (parallel) both return false
  const R u = a / b;
  const R c = cross(k0 - l0, k1 - k0);
  const R t = c / b;         // sc:166-169: d = b
  if (0 <= u && u <= 1 && 0 <= t && t <= 1) { pt = t * (l1 - l0) + l0; return true; }
  return false;
}

// sc:575-596 with k = 1 (sc:612): the tangent is cut back, edge after edge, where the segments W -> W + T or
// W -> W - T cross an edge, so that both control points W + T (this segment) and W - T (the previous one) stay on
// W's side of every edge.
template <class R> inline V2<R> shrink_tangent(V2<R> T, V2<R> Wp, const std::vector<Edge<R>>& edges) {
  V2<R> ret = T;
  for (const Edge<R>& e : edges) {
    V2<R> pt;
    if (seg_intersects(Wp + ret, Wp, e.a, e.b, pt)) ret = pt - Wp;
    if (seg_intersects(Wp - ret, Wp, e.a, e.b, pt)) ret = Wp - pt;
  }
  return ret;
}

// ---------------------------------------------------------------- A2: from_path
// sc:599-683 (obstacle list empty => shrink_tangent, sc:575-596, returns k*T with k = 1).
// `preview` reproduces the 10,001-point dense sampling of every segment (sc:679, 750-763)
// that nothing downstream consumes; it exists so the timed CPU baseline does the
// reference's work.  start_angle NaN => atan2 of the first leg (sc:606-609).
template <class R>
inline Spline<R> from_path(const std::vector<V2<R>>& path, R start_angle, bool preview = false,
                           int eval_mode = EVAL_BF, std::vector<V2<R>>* preview_pts = nullptr,
                           std::vector<V2<R>>* tangents_out = nullptr, const std::vector<Edge<R>>* edges = nullptr) {
  if (path.size() < 2) throw std::invalid_argument("Not enough points for a path");
  const size_t W = path.size();
  std::vector<V2<R>> tang; tang.reserve(W);
  if (std::isnan(start_angle)) {
    V2<R> d = path[1] - path[0];
    start_angle = std::atan2(d.y, d.x);
  }
  tang.push_back(calc_start_tangent(path[0], path[1], start_angle));
  V2<R> Te = calc_end_tangent(path[W - 2], path[W - 1]);
  for (size_t i = 1; i + 1 < W; ++i) tang.push_back(calc_tangent(path[i - 1], path[i], path[i + 1]));
  tang.push_back(Te);
  if (edges && !edges->empty())  // sc:615, 620, 637
    for (size_t i = 0; i < W; ++i) tang[i] = shrink_tangent(tang[i], path[i], *edges);
  Spline<R> sp;
  sp.ctrl.resize(W - 1);
  for (size_t i = 0; i + 1 < W; ++i)
    sp.ctrl[i] = {path[i], path[i] + tang[i], path[i + 1] - tang[i + 1], path[i + 1]};
  if (preview) {
    const R precision = (R)0.0001;
    const int n = (int)std::round(1.0f / precision);
    std::vector<R> pos(n + 1);
    for (int i = 0; i <= n; ++i) pos[i] = std::max((R)0, std::min((R)i * precision, (R)1));
    std::vector<V2<R>> pts;
    for (size_t i = 0; i + 1 < W; ++i) {
      eval_curve(sp.ctrl[i], pos, eval_mode, pts);
      if (preview_pts) preview_pts->insert(preview_pts->end(), pts.begin(), pts.end());
    }
  }
  if (tangents_out) *tangents_out = tang;
  return sp;
}

}  // namespace ora
