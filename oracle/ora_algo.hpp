// ORACLE — TEST INFRASTRUCTURE ONLY (see ora_spline.hpp header).
// toppra constraints, Seidel stage set-up, backward/forward passes, Spline parametrizer and
// sea-current's gen_vel_prof glue restated.  Citations: "tp:" = third-party/toppra/cpp/src/
// toppra/, "sc:" = sea_current.hpp.
#pragma once
#include "ora_seidel.hpp"

namespace ora {
namespace tp {

typedef std::function<void(double s, Vec& lower, Vec& upper)> VelLimFn;  // sc:1175 vel_lim_func

enum ReturnCode { OK = 0, ERR_UNKNOWN = 1, ERR_FAIL_CONTROLLABLE = 2, ERR_FAIL_FORWARD_PASS = 3,
                  ERR_UNINITIALIZED = 4, ERR_FAIL_FEASIBLE = 5 };  // tp:algorithm.hpp:13-32

struct Problem {
  PPoly path;
  Vec gridpoints;
  // velocity constraint (tp:constraint/linear_joint_velocity.{hpp,cpp}); absent if !has_vel
  bool has_vel = true;
  VelLimFn vel_lim;
  double maxsd = 1e8;  // linear_joint_velocity.hpp:33
  // acceleration constraint (tp:constraint/linear_joint_acceleration.cpp)
  Vec acc_lo, acc_hi;
  bool interpolation = true;  // sc:1210
  double x_init[2] = {0, 100};  // tp:algorithm.hpp:152 m_initXBound
};

struct Stages {
  int N = 0, nC = 0;
  Vec deltas;
  std::vector<std::vector<std::array<double, 3>>> A;  // [N+1][nC] (tp:solver/seidel.cpp:204-263 m_A)
  std::vector<std::array<double, 2>> low, high;       // [N+1] (u, x)
};

// tp:constraint/linear_joint_velocity.cpp:24-61
inline void velocity_xbounds(const Problem& pb, std::vector<std::array<double, 2>>& xbound) {
  const int G = (int)pb.gridpoints.size(), nd = pb.path.dof;
  xbound.resize(G);
  Vec lo(nd), hi(nd);
  for (int i = 0; i < G; ++i) {
    pb.vel_lim(pb.gridpoints[i], lo, hi);
    Vec v = pb.path.eval_single(pb.gridpoints[i], 1);
    double sdmin = -pb.maxsd, sdmax = pb.maxsd;
    for (int k = 0; k < nd; ++k) {
      const double vinv = 1.0 / v[k];
      if (v[k] > 0) { sdmax = std::min(hi[k] * vinv, sdmax); sdmin = std::max(lo[k] * vinv, sdmin); }
      else if (v[k] < 0) { sdmax = std::min(lo[k] * vinv, sdmax); sdmin = std::max(hi[k] * vinv, sdmin); }
      else if (hi[k] < 0 || lo[k] > 0) throw std::runtime_error("BoxConstraint is infeasible");
    }
    xbound[i][0] = 0;
    if (sdmin > 0) xbound[i][0] = sdmin * sdmin;
    xbound[i][1] = sdmax * sdmax;
  }
}

// tp:solver.cpp:26-63 + tp:solver/seidel.cpp:204-263 + tp:constraint.cpp:100-162,185-204 +
// tp:constraint/linear_joint_acceleration.cpp:22-49
inline Stages build_stages(const Problem& pb) {
  Stages st;
  const int G = (int)pb.gridpoints.size(), N = G - 1, m = pb.path.dof;
  st.N = N;
  st.deltas.resize(N);
  for (int i = 0; i < N; ++i) {
    st.deltas[i] = pb.gridpoints[i + 1] - pb.gridpoints[i];
    if (st.deltas[i] <= 0) throw std::invalid_argument("Invalid times.");
  }
  // collocation parameters: a = q', b = q'', c = 0; F = [I; -I], g = [acc_hi; -acc_lo]
  std::vector<Vec> a(G), b(G);
  for (int i = 0; i < G; ++i) { a[i] = pb.path.eval_single(pb.gridpoints[i], 1); b[i] = pb.path.eval_single(pb.gridpoints[i], 2); }
  const int mm = pb.interpolation ? 2 * m : m;  // variables after interpolation doubling
  const int k = 2 * mm;                         // rows of F
  st.nC = 2 + k;
  st.A.assign(G, std::vector<std::array<double, 3>>(st.nC, {0, 0, 0}));
  for (int i = 0; i < G; ++i) {
    Vec ai(mm), bi(mm), ci(mm, 0.0);
    for (int d = 0; d < m; ++d) { ai[d] = a[i][d]; bi[d] = b[i][d]; }
    if (pb.interpolation) {
      for (int d = 0; d < m; ++d) {
        if (i < N) ai[m + d] = a[i + 1][d] + 2 * st.deltas[i] * b[i + 1][d];
        else ai[m + d] = a[N][d];
        bi[m + d] = b[std::min(i + 1, N)][d];
      }
    }
    // F (2*mm x mm): interpolation => blkdiag([I;-I],[I;-I]); rows ordered
    // [+I block0; -I block0; +I block1; -I block1], g = [hi; -lo; hi; -lo].
    int row = 2;
    const int nblocks = pb.interpolation ? 2 : 1;
    for (int blk = 0; blk < nblocks; ++blk) {
      for (int sgn = 0; sgn < 2; ++sgn)
        for (int d = 0; d < m; ++d) {
          const double F = sgn == 0 ? 1.0 : -1.0;
          const double g = sgn == 0 ? pb.acc_hi[d] : -pb.acc_lo[d];
          st.A[i][row++] = {F * ai[blk * m + d], F * bi[blk * m + d], F * ci[blk * m + d] - g};
        }
    }
  }
  st.low.assign(G, {-INF, -INF}); st.high.assign(G, {INF, INF});
  if (pb.has_vel) {
    std::vector<std::array<double, 2>> xb;
    velocity_xbounds(pb, xb);
    for (int i = 0; i < G; ++i) { st.low[i][1] = std::max(st.low[i][1], xb[i][0]); st.high[i][1] = std::min(st.high[i][1], xb[i][1]); }
  }
  return st;
}

// tp:solver/seidel.cpp:265-347 (warm-start reordering is dead code there: solve_lp2d is
// called with the unordered m_A[i], l.335 — so it is not restated).
inline bool solve_stagewise(Stages& st, int i, const double g[2], const double x[2], const double xNext[2], double sol[2]) {
  double low[2] = {st.low[i][0], st.low[i][1]}, high[2] = {st.high[i][0], st.high[i][1]};
  low[1] = std::max(low[1], x[0]);
  high[1] = std::min(high[1], x[1]);
  if (i < st.N) {
    const double delta = st.deltas[i];
    if (xNext[0] <= -INF) st.A[i][0] = {0, 0, -1}; else st.A[i][0] = {-2 * delta, -1, xNext[0]};
    if (xNext[1] >= INF) st.A[i][1] = {0, 0, -1}; else st.A[i][1] = {2 * delta, 1, -xNext[1]};
  } else {
    st.A[i][0] = {0, 0, -1}; st.A[i][1] = {0, 0, -1};
  }
  const double v[2] = {-g[0], -g[1]};
  LpSol s = solve_lp2d(v, st.A[i], low, high);
  if (s.feasible) { sol[0] = s.optvar[0]; sol[1] = s.optvar[1]; return true; }
  return false;
}

struct ParamData {
  Vec gridpoints, parametrization;
  std::vector<std::array<double, 2>> controllable_sets, feasible_sets;
  int ret_code = ERR_UNINITIALIZED;
  int fail_stage = -1;
};

// tp:algorithm.cpp:28-68
inline int compute_controllable_sets(const Problem& pb, Stages& st, ParamData& d, double vel_end) {
  const int N = st.N;
  const double g_upper[2] = {1e-9, -1}, g_lower[2] = {-1e-9, 1};
  d.controllable_sets.assign(N + 1, {0, 0});
  d.controllable_sets[N] = {vel_end * vel_end, vel_end * vel_end};
  double sol[2];
  for (int i = N - 1; i >= 0; --i) {
    const double xn[2] = {d.controllable_sets[i + 1][0], d.controllable_sets[i + 1][1]};
    if (!solve_stagewise(st, i, g_upper, pb.x_init, xn, sol)) { d.fail_stage = i; return ERR_FAIL_CONTROLLABLE; }
    d.controllable_sets[i][1] = sol[1];
    if (!solve_stagewise(st, i, g_lower, pb.x_init, xn, sol)) { d.fail_stage = i; return ERR_FAIL_CONTROLLABLE; }
    d.controllable_sets[i][0] = std::max(0.0, sol[1]);
  }
  return OK;
}

// tp:algorithm/toppra.cpp:15-46.  NB parametrization(0) = vel_start, not squared (l.23).
inline int compute_forward_pass(Stages& st, ParamData& d, double vel_start) {
  const int N = st.N;
  d.parametrization.assign(N + 1, 0.0);
  d.parametrization[0] = vel_start;
  double sol[2];
  for (int i = 0; i < N; ++i) {
    const double g[2] = {-2 * st.deltas[i], -1};
    const double x[2] = {d.parametrization[i], d.parametrization[i]};
    const double xn[2] = {d.controllable_sets[i + 1][0], d.controllable_sets[i + 1][1]};
    if (!solve_stagewise(st, i, g, x, xn, sol)) { d.fail_stage = i; return ERR_FAIL_FORWARD_PASS; }
    d.parametrization[i + 1] = std::min(d.controllable_sets[i + 1][1],
                                        std::max(d.controllable_sets[i + 1][0], d.parametrization[i] + 2 * st.deltas[i] * sol[0]));
  }
  return OK;
}

// tp:algorithm.cpp:70-103
inline int compute_feasible_sets(Stages& st, ParamData& d) {
  const double g_upper[2] = {1e-9, -1}, g_lower[2] = {-1e-9, 1};
  const double x[2] = {0, 100}, xn[2] = {0, 100};
  d.feasible_sets.assign(st.N + 1, {0, 0});
  double sol[2];
  for (int i = 0; i < st.N + 1; ++i) {
    if (!solve_stagewise(st, i, g_upper, x, xn, sol)) return ERR_FAIL_FEASIBLE;
    d.feasible_sets[i][1] = sol[1];
    if (!solve_stagewise(st, i, g_lower, x, xn, sol)) return ERR_FAIL_FEASIBLE;
    d.feasible_sets[i][0] = sol[1];
  }
  return OK;
}

// tp:algorithm.cpp:14-26
inline ParamData compute_path_parametrization(const Problem& pb, double vel_start, double vel_end) {
  ParamData d;
  d.gridpoints = pb.gridpoints;
  Stages st = build_stages(pb);
  d.ret_code = compute_controllable_sets(pb, st, d, vel_end);
  d.parametrization.assign(st.N + 1, 0.0);
  if (d.ret_code > 0) return d;
  d.ret_code = compute_forward_pass(st, d, vel_start);
  return d;
}

// ---------------------------------------------------------------- parametrizer::Spline
// tp:parametrizer.cpp:6-21 + tp:parametrizer/spline.cpp:10-45
struct SplineParam {
  PPoly traj;  // q(t)
  Vec ts;      // knot times (after dropping dt < 1e-8 knots)
  Vec ss;      // path positions kept
  double duration() const { return ts.back() - ts.front(); }
};

inline SplineParam spline_parametrizer(const PPoly& path, const Vec& gridpoints, const Vec& vsquared, int solve_mode) {
  const int G = (int)gridpoints.size();
  Vec vs(G);
  for (int i = 0; i < G; ++i) vs[i] = std::sqrt(vsquared[i]);
  Vec t(G, 0.0), sg(gridpoints);
  std::vector<int> skip;
  for (int i = 1; i < G; ++i) {
    const double sd_average = (vs[i - 1] + vs[i]) / 2;
    double delta_t;
    if (sd_average > 1e-8) delta_t = (gridpoints[i] - gridpoints[i - 1]) / sd_average;
    else delta_t = 5;
    t[i] = t[i - 1] + delta_t;
    if (delta_t < 1e-8) skip.push_back(i);
  }
  for (int i = (int)skip.size() - 1; i > -1; --i) { t.erase(t.begin() + skip[i]); sg.erase(sg.begin() + skip[i]); }
  std::vector<Vec> q(sg.size());
  for (size_t i = 0; i < sg.size(); ++i) q[i] = path.eval_single(sg[i], 0);
  auto I = path.interval();
  Vec bc0 = path.eval_single(I[0], 1), bc1 = path.eval_single(I[1], 1);
  for (double& v : bc0) v *= vs[0];
  for (double& v : bc1) v *= vs[G - 1];
  SplineParam sp;
  sp.traj = cubic_spline_clamped(q, t, bc0, bc1, solve_mode);
  sp.ts = t; sp.ss = sg;
  return sp;
}

// ---------------------------------------------------------------- parametrizer::ConstAccel
// tp:parametrizer/const_accel.cpp:19-33 (process_internals), 35-58 (evalParams), 60-115 (eval_impl): the path
// acceleration is constant on every grid interval, so s(t) is a quadratic per interval and nothing is solved.
struct ConstAccelParam {
  PPoly path;
  Vec gridpoints, vs, ts, us;
  Vec eval_single(double t, int order) const {
    // first knot of m_ts[1 .. G-2] that is >= t; clamps to the first / last interval (const_accel.cpp:42-48)
    const int G = (int)ts.size();
    const int k = (int)(std::lower_bound(ts.begin() + 1, ts.begin() + (G - 1), t) - ts.begin()) - 1;
    const double dt = t - ts[k];
    const double u = us[k];
    const double v = vs[k] + dt * u;
    double ss = gridpoints[k] + dt * vs[k] + 0.5 * dt * dt * us[k];
    auto I = path.interval();
    ss = std::min(std::max(ss, I[0]), I[1]);  // const_accel.cpp:80-81
    if (order == 0) return path.eval_single(ss, 0);
    Vec ps = path.eval_single(ss, 1);
    if (order == 1) { for (double& x : ps) x *= v; return ps; }
    Vec pss = path.eval_single(ss, 2);
    for (size_t j = 0; j < ps.size(); ++j) pss[j] = pss[j] * v * v + ps[j] * u;
    return pss;
  }
};

inline ConstAccelParam const_accel_parametrizer(const PPoly& path, const Vec& gridpoints, const Vec& vsquared) {
  const int G = (int)gridpoints.size();
  ConstAccelParam c;
  c.path = path; c.gridpoints = gridpoints;
  c.vs.resize(G); c.ts.assign(G, 0.0); c.us.assign(G - 1, 0.0);
  for (int i = 0; i < G; ++i) c.vs[i] = std::sqrt(vsquared[i]);
  for (int i = 0; i + 1 < G; ++i) {
    c.us[i] = 0.5 * (c.vs[i + 1] * c.vs[i + 1] - c.vs[i] * c.vs[i]) / (gridpoints[i + 1] - gridpoints[i]);
    c.ts[i + 1] = c.ts[i] + 2 * (gridpoints[i + 1] - gridpoints[i]) / (c.vs[i + 1] + c.vs[i]);
  }
  return c;
}

// ---------------------------------------------------------------- sc:1191-1265 gen_vel_prof<1>
struct VelocityProfile {
  std::vector<float> pos, vel, acc;  // sc:1245-1262 (double -> float)
  Vec time;
  ParamData pd;
  SplineParam sp;
};

inline VelocityProfile gen_vel_prof(double pos_end, double pos_start, double vel_end, double vel_start,
                                    const VelLimFn& vel_lim, double acc_min, double acc_max, float dt = 0.02f,
                                    int N = 100, int solve_mode = SOLVE_DENSE_QR, double x_cap = 100.0) {
  Problem pb;
  pb.path = PPoly::hermite({{pos_start}, {pos_end}}, {{vel_start}, {vel_end}}, {0, 1});
  pb.gridpoints = linspaced(N + 1, 0.0, 1.0);  // tp:algorithm.cpp:137 with m_N = 100 by default
  pb.vel_lim = vel_lim;
  pb.acc_lo = {acc_min}; pb.acc_hi = {acc_max};
  pb.interpolation = true;
  pb.x_init[1] = x_cap;
  VelocityProfile out;
  out.pd = compute_path_parametrization(pb, 0, 0);  // sc:1225: (0, 0), vel_start/vel_end unused here
  if (out.pd.ret_code != OK) return out;
  out.sp = spline_parametrizer(pb.path, out.pd.gridpoints, out.pd.parametrization, solve_mode);
  const double t0 = out.sp.ts.front(), t1 = out.sp.ts.back();
  const int length = (int)std::ceil((t1 - t0) / dt);
  out.time = linspaced(length, t0, t1);
  out.pos.resize(length); out.vel.resize(length); out.acc.resize(length);
  for (int i = 0; i < length; ++i) {
    out.pos[i] = (float)out.sp.traj.eval_single(out.time[i], 0)[0];
    out.vel[i] = (float)out.sp.traj.eval_single(out.time[i], 1)[0];
    out.acc[i] = (float)out.sp.traj.eval_single(out.time[i], 2)[0];
  }
  return out;
}

}  // namespace tp
}  // namespace ora
