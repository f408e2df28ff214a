// ORACLE — TEST INFRASTRUCTURE ONLY (see ora_spline.hpp header).
//
// One planning problem end to end in FP64, in the shape of the product's batch record
// (include/seacurrent_b200.h): spline (reference cubic sc:599-683, or Lau quintic +
// optimiser, ora_lau.hpp) -> arc-length tables (sc:767-896) -> toppra on the 1-DOF Hermite
// path (sc:1191-1265 + the tp: chain in ora_algo.hpp) -> Spline parametrizer -> per-gridpoint
// record and the uniform-dt stream (sc:1235-1262).
//
// Two steps deviate from the reference ON PURPOSE (DESIGN.md "Deviations"):
//   * arc length -> curve parameter: the reference fits a degree-9 Chebyshev polynomial per
//     segment in float (resample/chebfit, sc:898-1005, 1109-1170; ~1e-4 L error).  The
//     product inverts the table exactly: table search + Newton steps to convergence on a 4-point
//     Gauss–Legendre integral of |B'| (arclen_invert below).  The reference variant is kept
//     in ora_arclen.hpp (resample) and compared against the fixture separately.
//   * the clamped cubic spline of the parametrizer is solved by the Thomas algorithm instead
//     of a dense column-pivoted QR (tp:geometric_path/piecewise_poly_path.cpp:162); solve
//     mode SOLVE_DENSE_QR restates the reference and the two agree to ~1e-13.
#pragma once
#include "ora_algo.hpp"
#include "ora_arclen.hpp"
#include "ora_lau.hpp"

namespace ora {

enum { ST_BAD_PATH = 16, ST_BAD_LIMITS = 17, ST_TIME_KNOTS = 18, ST_COLLISION = 19 };
static const int INV_NEWTON_ITERS = 40;       // cap; the iteration stops when it has converged (|step| <= 1e-12 h)
static const double GL4_X[4] = {-0.8611363115940525752, -0.3399810435848562648, 0.3399810435848562648, 0.8611363115940525752};
static const double GL4_W[4] = {0.3478548451374538574, 0.6521451548625461426, 0.6521451548625461426, 0.3478548451374538574};

struct PlanIn {
  std::vector<V2<double>> path;
  double start_angle = NAN;
  double vel_lo = -1, vel_hi = 1;
  const double* vel_lo_grid = nullptr;  // [G] or null
  const double* vel_hi_grid = nullptr;
  double acc_lo = -1, acc_hi = 1;
  double vel_start = 0, vel_end = 0;
  int G = 101;
  double x_cap = 100.0;
  double dt = (double)0.02f;
  double omega_max = 0, acc_lat_max = 0;  // curvature-aware velocity caps (0 = off), see scb_batch_in
  double a_brake = 0, t_react = 0;        // obstacle-distance speed limit, Lau09 III-B eq. 2 (a_brake = 0: off)
  int collision_chords = 0;               // M > 0: collision verdict of the finished spline (Lau09 IV-B), see scb_batch_in
  std::vector<Edge<double>> edges;        // obstacle edges for the tangent clipping (REF_CUBIC), see scb_batch_in
  int spline_mode = 0;  // 0 REF_CUBIC, 1 LAU_QUINTIC
  LauOpt opt;
  int solve_mode = tp::SOLVE_THOMAS;
  int param_mode = 0;   // 0 parametrizer::Spline (sc:1231), 1 parametrizer::ConstAccel (tp:parametrizer/const_accel.cpp)
  int S_cap = 0;
  bool timing_only = false;  // SCB_STAGE_TIMING alone: gen_vel_prof on a path of length path_len, no curve
  double path_len = 0;
};

struct PlanOut {
  int status = tp::ERR_UNINITIALIZED, fail_stage = -1;
  int degree = 3;
  std::vector<double> ctrl_x, ctrl_y;      // (W-1)*(d+1)
  double L = 0, T = 0, cost = 0;
  std::vector<double> table;               // (W-1)*101
  std::vector<double> K_lo, K_hi, x, t, px, py, v, a, omega;  // G each
  int n_samples = 0;
  std::vector<double> s_time, s_pos, s_vel, s_acc, s_px, s_py, s_omega;
};

// Per-segment power-basis coefficients of B, B', B'' (what the CUDA kernels keep).
struct SegPoly {
  int deg;
  V2<double> a[6], d1[5], d2[4];
  void set(const std::vector<V2<double>>& ctrl) {
    deg = (int)ctrl.size() - 1;
    bezier_power_coeffs(ctrl.data(), deg, a);
    for (int k = 0; k < deg; ++k) d1[k] = (double)(k + 1) * a[k + 1];
    for (int k = 0; k + 1 < deg; ++k) d2[k] = (double)(k + 1) * d1[k + 1];
  }
  double speed(double u) const { V2<double> d = horner(d1, deg - 1, u); return std::sqrt(d.x * d.x + d.y * d.y); }
};

// arc length q in [0, L]  ->  (segment, u).  seg_end[j] = cumulative length at the end of
// segment j; table = per-segment cumulative tables (101 entries each).
inline void arclen_invert(const std::vector<SegPoly>& polys, const std::vector<double>& table,
                          const std::vector<double>& seg_end, double q, int& seg, double& u) {
  const int nseg = (int)polys.size();
  q = std::min(std::max(q, 0.0), seg_end[nseg - 1]);
  seg = nseg - 1;
  for (int j = 0; j < nseg; ++j) if (q <= seg_end[j]) { seg = j; break; }
  const double r = q - (seg > 0 ? seg_end[seg - 1] : 0.0);
  const double* c = &table[(size_t)seg * 101];
  int lo = 0, hi = 100;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (c[mid] <= r) lo = mid; else hi = mid; }
  const double h = 0.01, a = std::min(lo * h, 1.0);
  const double target = r - c[lo];
  double du = h * target / (c[lo + 1] - c[lo]);
  du = std::min(std::max(du, 0.0), h);
  const SegPoly& sp = polys[seg];
  for (int it = 0; it < INV_NEWTON_ITERS; ++it) {
    const double half = 0.5 * du, mid = a + half;
    double F = 0;
    for (int g = 0; g < 4; ++g) F += GL4_W[g] * sp.speed(mid + half * GL4_X[g]);
    F *= half;
    const double delta = (F - target) / sp.speed(a + du);
    du = std::min(std::max(du - delta, 0.0), h);
    if (std::fabs(delta) <= 1e-12 * h) break;  // converged: the next correction is ~delta^2
  }
  u = std::min(a + du, 1.0);
}

inline void curve_at(const SegPoly& sp, double u, double& px, double& py, double& kappa) {
  const V2<double> p = horner(sp.a, sp.deg, u), d = horner(sp.d1, sp.deg - 1, u), dd = horner(sp.d2, sp.deg - 2, u);
  px = p.x; py = p.y;
  const double n2 = d.x * d.x + d.y * d.y;
  kappa = (d.x * dd.y - d.y * dd.x) / (n2 * std::sqrt(n2));  // sc:1029
}

inline PlanOut plan_one(const PlanIn& in) {
  PlanOut o;
  const int W = (int)in.path.size(), G = in.G;
  // ---- spline
  Spline<double> sp;
  bool bad = false;
  int nseg = 0;
  std::vector<double> seg_end;
  std::vector<SegPoly> polys;
  if (in.timing_only) { o.L = in.path_len; } else {
  for (int i = 0; i < W; ++i) bad = bad || !std::isfinite(in.path[i].x) || !std::isfinite(in.path[i].y);
  for (int i = 0; i + 1 < W && !bad; ++i) bad = !(pt_dist(in.path[i], in.path[i + 1]) > 0.0);
  if (bad) { o.status = ST_BAD_PATH; return o; }
  if (in.spline_mode == 0) {
    sp = from_path(in.path, in.start_angle, false, EVAL_BF, (std::vector<V2<double>>*)nullptr, (std::vector<V2<double>>*)nullptr, &in.edges);
    o.cost = 0;
  } else {
    LauModel<double> m = lau_model(in.path, in.start_angle);
    // the collision term lives INSIDE the optimiser loop in quintic mode (ora_lau.hpp: rejected trials, best
    // collision-free iterate); the verdict on the finished spline below is then the same test on that iterate
    LauResult<double> r = lau_optimise(m, in.opt, &in.edges, in.collision_chords);
    sp = r.spline; o.cost = r.cost;
  }
  o.degree = sp.degree();
  for (auto& seg : sp.ctrl) for (auto& p : seg) { o.ctrl_x.push_back(p.x); o.ctrl_y.push_back(p.y); }
  // ---- arc length
  ArclengthData<double> ad = arclength(sp, 0.01, EVAL_POW);
  o.L = ad.arclength;
  for (auto& seg : ad.segments) o.table.insert(o.table.end(), seg.begin(), seg.end());
  if (!(o.L > 0) || !std::isfinite(o.L)) { o.status = ST_BAD_PATH; return o; }
  nseg = sp.n_segments();
  // ---- collision verdict (Lau09 IV-B: a trajectory that is not collision-free has infinite cost): every segment cut
  //      into M chords (de Casteljau), every chord against every obstacle edge with intersects (sc:142-178)
  if (in.collision_chords > 0 && !in.edges.empty()) {
    const int M = in.collision_chords;
    for (int j = 0; j < nseg; ++j) {
      V2<double> a = sp.ctrl[j][0];
      for (int k = 1; k <= M; ++k) {
        const double u = k == M ? 1.0 : (double)k / (double)M;
        std::vector<V2<double>> t = sp.ctrl[j];
        for (size_t r = 1; r < t.size(); ++r)
          for (size_t i = 0; i + r < t.size(); ++i) t[i] = {(1.0 - u) * t[i].x + u * t[i + 1].x, (1.0 - u) * t[i].y + u * t[i + 1].y};
        V2<double> pt;
        for (const auto& e : in.edges)
          if (seg_intersects(a, t[0], e.a, e.b, pt)) { o.status = ST_COLLISION; o.cost = INFINITY; return o; }
        a = t[0];
      }
    }
  }
  seg_end.resize(nseg);
  double acc = 0;
  for (int j = 0; j < nseg; ++j) { acc += ad.segments[j].back(); seg_end[j] = acc; }
  polys.resize(nseg);
  for (int j = 0; j < nseg; ++j) polys[j].set(sp.ctrl[j]);
  }
  // ---- toppra (sc:1191-1233)
  tp::Problem pb;
  pb.path = tp::PPoly::hermite({{0.0}, {o.L}}, {{in.vel_start}, {in.vel_end}}, {0, 1});
  pb.gridpoints = tp::linspaced(G, 0.0, 1.0);
  const double *vlg = in.vel_lo_grid, *vhg = in.vel_hi_grid;
  const double vlo = in.vel_lo, vhi = in.vel_hi;
  const tp::Vec& gp = pb.gridpoints;
  // Curvature-aware limits (Lau09 III-B; the reference's hook is the position-dependent vel_lim_func,
  // sc:1175-1188): the box of gridpoint i is intersected with [-c_i, c_i],
  // c_i = min(omega_max / |kappa_i|, sqrt(acc_lat_max / |kappa_i|)), kappa_i = curvature at arc length q(s_i).
  // Obstacle-distance limit (Lau09 III-B eq. 2): c_i = -a_brake t_react + sqrt(a_brake^2 t_react^2 + 2 a_brake d_i),
  // d_i = distance from the curve point of gridpoint i to the closest obstacle edge (point to segment).
  // LinearJointAcceleration::check throws when lower > upper (tp:constraint/linear_joint_acceleration.cpp:18-22);
  // a NaN limit, which every comparison of the reference lets through, is reported the same way
  if (!(in.acc_lo <= in.acc_hi)) { o.status = ST_BAD_LIMITS; return o; }
  if (vlg && vhg) {
    for (int i = 0; i < G; ++i) if (std::isnan(vlg[i]) || std::isnan(vhg[i])) { o.status = ST_BAD_LIMITS; return o; }
  } else if (std::isnan(vlo) || std::isnan(vhi)) { o.status = ST_BAD_LIMITS; return o; }
  const bool obst = in.a_brake > 0 && !in.edges.empty();
  std::vector<double> cl, ch;
  if (!in.timing_only && (in.omega_max > 0 || in.acc_lat_max > 0 || obst)) {
    cl.resize(G); ch.resize(G);
    for (int i = 0; i < G; ++i) {
      const double q = pb.path.eval_single(gp[i], 0)[0];
      int seg; double u, bx, by, kap;
      arclen_invert(polys, o.table, seg_end, q, seg, u);
      curve_at(polys[seg], u, bx, by, kap);
      double lo = vlg && vhg ? vlg[i] : vlo, hi = vlg && vhg ? vhg[i] : vhi;
      const double k = std::fabs(kap);
      if (k > 0) {
        double c = INFINITY;
        if (in.omega_max > 0) c = in.omega_max / k;
        if (in.acc_lat_max > 0) c = std::min(c, std::sqrt(in.acc_lat_max / k));
        hi = std::min(hi, c); lo = std::max(lo, -c);
      }
      if (obst) {
        double d2 = INFINITY;
        for (const auto& e : in.edges) {
          const double ex = e.b.x - e.a.x, ey = e.b.y - e.a.y, rx = bx - e.a.x, ry = by - e.a.y, ee = ex * ex + ey * ey;
          double t = ee > 0 ? (rx * ex + ry * ey) / ee : 0.0;
          t = std::min(std::max(t, 0.0), 1.0);
          const double qx = rx - t * ex, qy = ry - t * ey;
          d2 = std::min(d2, qx * qx + qy * qy);
        }
        const double at = in.a_brake * in.t_react, c = -at + std::sqrt(at * at + 2.0 * in.a_brake * std::sqrt(d2));
        hi = std::min(hi, c); lo = std::max(lo, -c);
      }
      cl[i] = lo; ch[i] = hi;
    }
    vlg = cl.data(); vhg = ch.data();
  }
  pb.vel_lim = [&](double s, tp::Vec& lo, tp::Vec& hi) {
    if (vlg && vhg) {
      size_t i = std::lower_bound(gp.begin(), gp.end(), s) - gp.begin();
      if (i >= gp.size()) i = gp.size() - 1;
      lo[0] = vlg[i]; hi[0] = vhg[i];
    } else { lo[0] = vlo; hi[0] = vhi; }
  };
  pb.acc_lo = {in.acc_lo}; pb.acc_hi = {in.acc_hi};
  pb.interpolation = true;
  pb.x_init[1] = in.x_cap;
  tp::ParamData pd;
  try {
    pd = tp::compute_path_parametrization(pb, 0, 0);  // sc:1225
  } catch (const std::runtime_error&) { o.status = ST_BAD_LIMITS; return o; }
  o.status = pd.ret_code; o.fail_stage = pd.fail_stage;
  o.K_lo.assign(G, 0); o.K_hi.assign(G, 0); o.x.assign(G, 0);
  for (size_t i = 0; i < pd.controllable_sets.size(); ++i) { o.K_lo[i] = pd.controllable_sets[i][0]; o.K_hi[i] = pd.controllable_sets[i][1]; }
  for (size_t i = 0; i < pd.parametrization.size(); ++i) o.x[i] = pd.parametrization[i];
  if (pd.ret_code != tp::OK) return o;
  // ---- parametrizer: traj(t, order) is Spline::eval_impl (tp:parametrizer/spline.cpp:47-49) or ConstAccel::eval_impl
  tp::SplineParam spp;
  tp::ConstAccelParam cap;
  std::function<double(double, int)> traj;
  if (in.param_mode == 1) {
    cap = tp::const_accel_parametrizer(pb.path, pd.gridpoints, pd.parametrization);
    if (!std::isfinite(cap.ts.back())) { o.status = ST_TIME_KNOTS; return o; }  // two neighbouring gridpoints at rest
    o.t = cap.ts;
    traj = [&cap](double t, int order) { return cap.eval_single(t, order)[0]; };
  } else {
    spp = tp::spline_parametrizer(pb.path, pd.gridpoints, pd.parametrization, in.solve_mode);
    if ((int)spp.ts.size() != G) { o.status = ST_TIME_KNOTS; return o; }
    o.t = spp.ts;
    traj = [&spp](double t, int order) { return spp.traj.eval_single(t, order)[0]; };
  }
  o.T = o.t.back() - o.t.front();
  o.px.resize(G); o.py.resize(G); o.v.resize(G); o.a.resize(G); o.omega.resize(G);
  for (int i = 0; i < G; ++i) {
    o.v[i] = traj(o.t[i], 1);
    o.a[i] = traj(o.t[i], 2);
    if (in.timing_only) continue;
    const double q = pb.path.eval_single(gp[i], 0)[0];
    int seg; double u, kap;
    arclen_invert(polys, o.table, seg_end, q, seg, u);
    curve_at(polys[seg], u, o.px[i], o.py[i], kap);
    o.omega[i] = o.v[i] * kap;
  }
  // ---- uniform-dt stream (sc:1235-1262 + resample + angular_velocity)
  if (in.S_cap > 0) {
    const int S = (int)std::ceil(o.T / in.dt);
    o.n_samples = S;
    const int n = std::min(S, in.S_cap);
    tp::Vec times = tp::linspaced(S, o.t.front(), o.t.back());
    o.s_time.resize(n); o.s_pos.resize(n); o.s_vel.resize(n); o.s_acc.resize(n);
    o.s_px.resize(n); o.s_py.resize(n); o.s_omega.resize(n);
    for (int i = 0; i < n; ++i) {
      o.s_time[i] = times[i];
      o.s_pos[i] = traj(times[i], 0);
      o.s_vel[i] = traj(times[i], 1);
      o.s_acc[i] = traj(times[i], 2);
      if (in.timing_only) continue;
      int seg; double u, kap;
      arclen_invert(polys, o.table, seg_end, o.s_pos[i], seg, u);
      curve_at(polys[seg], u, o.s_px[i], o.s_py[i], kap);
      o.s_omega[i] = o.s_vel[i] * kap;
    }
  }
  return o;
}

}  // namespace ora
