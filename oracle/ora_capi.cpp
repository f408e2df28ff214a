// ORACLE — TEST INFRASTRUCTURE ONLY.  C entry points over the header-only oracle so that
// tests/ (ctypes), __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// can call it.  Never linked into libseacurrent_b200.so.
#include <chrono>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/seacurrent_b200.h"
#include "ora_plan.hpp"
#include "ora_fmt.hpp"

using namespace ora;

extern "C" {

int ora_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// ---------------------------------------------------------------- Seidel (test_solver.cpp)
int ora_lp1d(double v0, double v1, const double* A, int n, double* optvar) {
  std::vector<std::array<double, 2>> a(n);
  for (int i = 0; i < n; ++i) a[i] = {A[2 * i], A[2 * i + 1]};
  tp::LpSol1d s = tp::solve_lp1d(v0, v1, a, n);
  *optvar = s.optvar;
  return s.feasible ? 1 : 0;
}

int ora_lp2d(const double* v, const double* A, int n, const double* low, const double* high, double* optvar,
             int* active) {
  std::vector<std::array<double, 3>> a(n);
  for (int i = 0; i < n; ++i) a[i] = {A[3 * i], A[3 * i + 1], A[3 * i + 2]};
  tp::LpSol s = tp::solve_lp2d(v, a, low, high);
  optvar[0] = s.optvar[0]; optvar[1] = s.optvar[1];
  if (active) { active[0] = s.active_c[0]; active[1] = s.active_c[1]; }
  return s.feasible ? 1 : 0;
}

// ---------------------------------------------------------------- generic toppra problem
// coeffs: [nseg][4][dof] (row 0 multiplies ds^3), breaks [nseg+1].  Constant velocity and
// acceleration boxes.  Outputs K [G][2], x [G], feas [G][2] (any may be null).
// Returns the ReturnCode of computePathParametrization; *feas_rc that of computeFeasibleSets.
int ora_toppra_ppoly(int nseg, int dof, const double* coeffs, const double* breaks, int G, const double* vel_lo,
                     const double* vel_hi, const double* acc_lo, const double* acc_hi, int interpolation,
                     double sd_start, double sd_end, double* K, double* x, double* feas, int* feas_rc) {
  tp::Problem pb;
  pb.path.breaks.assign(breaks, breaks + nseg + 1);
  for (int s = 0; s < nseg; ++s) {
    std::vector<tp::Vec> c(4, tp::Vec(dof));
    for (int r = 0; r < 4; ++r) for (int k = 0; k < dof; ++k) c[r][k] = coeffs[((size_t)s * 4 + r) * dof + k];
    pb.path.c0.push_back(c);
  }
  pb.path.finish();
  pb.gridpoints = tp::linspaced(G, breaks[0], breaks[nseg]);
  tp::Vec vl(vel_lo, vel_lo + dof), vh(vel_hi, vel_hi + dof);
  pb.vel_lim = [vl, vh](double, tp::Vec& lo, tp::Vec& hi) { lo = vl; hi = vh; };
  pb.acc_lo.assign(acc_lo, acc_lo + dof); pb.acc_hi.assign(acc_hi, acc_hi + dof);
  pb.interpolation = interpolation != 0;
  tp::ParamData d = tp::compute_path_parametrization(pb, sd_start, sd_end);
  for (int i = 0; i < G; ++i) {
    if (K && i < (int)d.controllable_sets.size()) { K[2 * i] = d.controllable_sets[i][0]; K[2 * i + 1] = d.controllable_sets[i][1]; }
    if (x && i < (int)d.parametrization.size()) x[i] = d.parametrization[i];
  }
  if (feas) {
    tp::Stages st = tp::build_stages(pb);
    tp::ParamData f;
    const int rc = tp::compute_feasible_sets(st, f);
    if (feas_rc) *feas_rc = rc;
    for (int i = 0; i < G; ++i) { feas[2 * i] = f.feasible_sets[i][0]; feas[2 * i + 1] = f.feasible_sets[i][1]; }
  }
  return d.ret_code;
}

// The same with velocity limits that vary along the path (LinearJointVelocityVarying, sc:1177-1188): vel_lo_grid /
// vel_hi_grid [G][dof] are the limits AT the gridpoints (the constraint evaluates vel_lim_func there and nowhere else).
int ora_toppra_ppoly_grid(int nseg, int dof, const double* coeffs, const double* breaks, int G, const double* vel_lo_grid,
                          const double* vel_hi_grid, const double* acc_lo, const double* acc_hi, int interpolation,
                          double sd_start, double sd_end, double* K, double* x, int* fail_stage) {
  tp::Problem pb;
  pb.path.breaks.assign(breaks, breaks + nseg + 1);
  for (int s = 0; s < nseg; ++s) {
    std::vector<tp::Vec> c(4, tp::Vec(dof));
    for (int r = 0; r < 4; ++r) for (int k = 0; k < dof; ++k) c[r][k] = coeffs[((size_t)s * 4 + r) * dof + k];
    pb.path.c0.push_back(c);
  }
  pb.path.finish();
  pb.gridpoints = tp::linspaced(G, breaks[0], breaks[nseg]);
  const tp::Vec gp = pb.gridpoints;
  pb.vel_lim = [=](double s, tp::Vec& lo, tp::Vec& hi) {
    int i = 0;  // the gridpoint this position is
    double best = INFINITY;
    for (int j = 0; j < G; ++j) { const double d = std::fabs(gp[j] - s); if (d < best) { best = d; i = j; } }
    lo.assign(vel_lo_grid + (size_t)i * dof, vel_lo_grid + (size_t)(i + 1) * dof);
    hi.assign(vel_hi_grid + (size_t)i * dof, vel_hi_grid + (size_t)(i + 1) * dof);
  };
  pb.acc_lo.assign(acc_lo, acc_lo + dof); pb.acc_hi.assign(acc_hi, acc_hi + dof);
  pb.interpolation = interpolation != 0;
  tp::ParamData d = tp::compute_path_parametrization(pb, sd_start, sd_end);
  for (int i = 0; i < G; ++i) {
    if (K && i < (int)d.controllable_sets.size()) { K[2 * i] = d.controllable_sets[i][0]; K[2 * i + 1] = d.controllable_sets[i][1]; }
    if (x && i < (int)d.parametrization.size()) x[i] = d.parametrization[i];
  }
  if (fail_stage) *fail_stage = d.fail_stage;
  return d.ret_code;
}

// Spline parametrizer over a generic ppoly path (test_spline_parametrizer.cpp): returns the
// knot times ts [G] and evaluates q(t), q'(t) at `n_eval` times.  solve_mode 0 dense QR, 1 Thomas.
int ora_spline_param(int nseg, int dof, const double* coeffs, const double* breaks, int G, const double* gridpoints,
                     const double* vsq, int solve_mode, double* ts, int n_eval, const double* t_eval, double* q,
                     double* qd, double* qdd) {
  tp::PPoly path;
  path.breaks.assign(breaks, breaks + nseg + 1);
  for (int s = 0; s < nseg; ++s) {
    std::vector<tp::Vec> c(4, tp::Vec(dof));
    for (int r = 0; r < 4; ++r) for (int k = 0; k < dof; ++k) c[r][k] = coeffs[((size_t)s * 4 + r) * dof + k];
    path.c0.push_back(c);
  }
  path.finish();
  tp::Vec gp(gridpoints, gridpoints + G), v(vsq, vsq + G);
  tp::SplineParam sp = tp::spline_parametrizer(path, gp, v, solve_mode);
  for (size_t i = 0; i < sp.ts.size(); ++i) ts[i] = sp.ts[i];
  for (int i = 0; i < n_eval; ++i) {
    tp::Vec a = sp.traj.eval_single(t_eval[i], 0), b = sp.traj.eval_single(t_eval[i], 1), c = sp.traj.eval_single(t_eval[i], 2);
    for (int k = 0; k < dof; ++k) { q[i * dof + k] = a[k]; qd[i * dof + k] = b[k]; if (qdd) qdd[i * dof + k] = c[k]; }
  }
  return (int)sp.ts.size();
}

// ConstAccel parametrizer over the 1-DOF Hermite path of gen_vel_prof (0 -> L on s in [0, 1], end slopes vs / ve):
// knot times ts [G] and q, q', q'' at `n_eval` times (tp:parametrizer/const_accel.cpp).
void ora_const_accel_param(double L, double vs, double ve, int G, const double* gridpoints, const double* vsq, double* ts,
                           int n_eval, const double* t_eval, double* q, double* qd, double* qdd) {
  tp::PPoly path = tp::PPoly::hermite({{0.0}, {L}}, {{vs}, {ve}}, {0, 1});
  tp::ConstAccelParam c = tp::const_accel_parametrizer(path, tp::Vec(gridpoints, gridpoints + G), tp::Vec(vsq, vsq + G));
  for (int i = 0; i < G; ++i) ts[i] = c.ts[i];
  for (int i = 0; i < n_eval; ++i) {
    q[i] = c.eval_single(t_eval[i], 0)[0]; qd[i] = c.eval_single(t_eval[i], 1)[0]; qdd[i] = c.eval_single(t_eval[i], 2)[0];
  }
}

// Clamped cubic spline through (t_i, y_i) with end slopes (test_cubic_spline*.cpp vs scipy).
void ora_cubic_spline(int n, const double* t, const double* y, double bc0, double bc1, int solve_mode, int n_eval,
                      const double* t_eval, double* out) {
  std::vector<tp::Vec> pos(n);
  for (int i = 0; i < n; ++i) pos[i] = {y[i]};
  tp::PPoly p = tp::cubic_spline_clamped(pos, tp::Vec(t, t + n), {bc0}, {bc1}, solve_mode);
  for (int i = 0; i < n_eval; ++i) out[i] = p.eval_single(t_eval[i], 0)[0];
}

// Hermite path eval (test_hermite_spline.cpp)
void ora_hermite_eval(int n, int dof, const double* pos, const double* vel, const double* times, int n_eval,
                      const double* s_eval, int order, double* out) {
  std::vector<tp::Vec> P(n), V(n);
  for (int i = 0; i < n; ++i) { P[i].assign(pos + i * dof, pos + (i + 1) * dof); V[i].assign(vel + i * dof, vel + (i + 1) * dof); }
  tp::PPoly p = tp::PPoly::hermite(P, V, tp::Vec(times, times + n));
  for (int i = 0; i < n_eval; ++i) { tp::Vec r = p.eval_single(s_eval[i], order); for (int k = 0; k < dof; ++k) out[i * dof + k] = r[k]; }
}

// ---------------------------------------------------------------- reference-faithful E1 (float32 spline side)
// examples/json_test.cpp:28-58 / zmq_test.cpp:64-92 body: from_path -> arclength ->
// gen_vel_prof<1> -> resample -> angular_velocity, with the reference's arithmetic types.
// mode bit 0: dense 10,001-point preview sampling of every segment (sc:679);
// mode bit 1: Bernstein-Fourier evaluation with complex pow (sc:721-748), else de Casteljau;
// mode bit 2: dense column-pivoted QR in the parametrizer (tp:...piecewise_poly_path.cpp:162),
//             else Thomas ("restated-fast").
struct E1Out {
  Spline<float> sp; ArclengthData<float> ad; tp::VelocityProfile prof; Resampled<float> rs; std::vector<float> angvel;
};
static E1Out e1_reference(const float* wx, const float* wy, int W, float start_angle, double vel_lo, double vel_hi,
                          double acc_lo, double acc_hi, int N, float dt, int mode) {
  E1Out o;
  std::vector<V2<float>> path(W);
  for (int i = 0; i < W; ++i) path[i] = {wx[i], wy[i]};
  const int ev = (mode & 2) ? EVAL_BF : EVAL_DC;
  o.sp = from_path(path, start_angle, (mode & 1) != 0, ev);
  o.ad = arclength(o.sp, 0.01f, ev);
  tp::VelLimFn lim = [vel_lo, vel_hi](double, tp::Vec& lo, tp::Vec& hi) { lo[0] = vel_lo; hi[0] = vel_hi; };
  o.prof = tp::gen_vel_prof((double)o.ad.arclength, 0.0, 0.0, 0.0, lim, acc_lo, acc_hi, dt, N,
                            (mode & 4) ? tp::SOLVE_DENSE_QR : tp::SOLVE_THOMAS);
  if (o.prof.pd.ret_code != tp::OK) return o;
  std::vector<float> pos = o.prof.pos;
  o.rs = resample(o.sp, pos, o.ad, true, ev);
  std::vector<float> curv = curvature(o.sp, o.rs, ev);
  std::vector<float> vel(o.prof.vel.begin(), o.prof.vel.begin() + std::min(o.prof.vel.size(), curv.size()));
  curv.resize(vel.size());
  o.angvel = angular_velocity(curv, vel);
  return o;
}

// Single problem, all outputs (fixture test).  Buffers sized by the caller: ctrl 8*(W-1)
// floats (x,y interleaved per control point), tables (W-1)*101, streams cap entries.
int ora_e1_single(const float* wx, const float* wy, int W, float start_angle, double vel_lo, double vel_hi,
                  double acc_lo, double acc_hi, int N, float dt, int mode, float* ctrl, float* tables, float* total,
                  int cap, int* n_samples, double* time, float* pos, float* vel, float* acc, float* px, float* py,
                  float* angvel, int* n_resampled) {
  E1Out o = e1_reference(wx, wy, W, start_angle, vel_lo, vel_hi, acc_lo, acc_hi, N, dt, mode);
  for (int s = 0; s + 1 < W; ++s) for (int j = 0; j < 4; ++j) { ctrl[(s * 4 + j) * 2] = o.sp.ctrl[s][j].x; ctrl[(s * 4 + j) * 2 + 1] = o.sp.ctrl[s][j].y; }
  for (int s = 0; s + 1 < W; ++s) for (int k = 0; k < 101; ++k) tables[s * 101 + k] = o.ad.segments[s][k];
  *total = o.ad.arclength;
  const int S = (int)o.prof.time.size();
  *n_samples = S;
  for (int i = 0; i < std::min(S, cap); ++i) { time[i] = o.prof.time[i]; pos[i] = o.prof.pos[i]; vel[i] = o.prof.vel[i]; acc[i] = o.prof.acc[i]; }
  const int Rn = (int)o.rs.pts.size();
  *n_resampled = Rn;
  for (int i = 0; i < std::min(Rn, cap); ++i) { px[i] = o.rs.pts[i].x; py[i] = o.rs.pts[i].y; angvel[i] = i < (int)o.angvel.size() ? o.angvel[i] : 0.f; }
  return o.prof.pd.ret_code;
}

// Timed CPU baseline ("port" of the reference's E1 body), OpenMP over problems.
// wx/wy [P][W] float; limits [P] double.  Returns elapsed seconds; *checksum defeats DCE.
double ora_e1_batch_time(int P, int W, const float* wx, const float* wy, const double* vel_lo, const double* vel_hi,
                         const double* acc_lo, const double* acc_hi, int N, float dt, int mode, int nthreads,
                         double* checksum) {
  double sum = 0;
  auto t0 = std::chrono::steady_clock::now();
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic) reduction(+ : sum)
  for (int p = 0; p < P; ++p) {
    E1Out o = e1_reference(wx + (size_t)p * W, wy + (size_t)p * W, W, NAN, vel_lo[p], vel_hi[p], acc_lo[p], acc_hi[p], N, dt, mode);
    sum += o.ad.arclength + (o.prof.time.empty() ? 0.0 : o.prof.time.back()) + (o.angvel.empty() ? 0.f : o.angvel[o.angvel.size() / 2]);
  }
  auto t1 = std::chrono::steady_clock::now();
  if (checksum) *checksum = sum;
  return std::chrono::duration<double>(t1 - t0).count();
}

// ---------------------------------------------------------------- Lau stages (property tests)
// Evaluate the quintic built from (Wp, e) over the model frozen at `path`: control points
// [(W-1)*6*2], cost J, gradients ge [W], gW [W*2].
double ora_lau_eval(const double* path_xy, int W, double start_angle, const double* Wp_xy, const double* e, double lambda,
                    double* ctrl, double* ge, double* gW) {
  std::vector<V2<double>> path(W), Wp(W);
  for (int i = 0; i < W; ++i) { path[i] = {path_xy[2 * i], path_xy[2 * i + 1]}; Wp[i] = {Wp_xy[2 * i], Wp_xy[2 * i + 1]}; }
  LauModel<double> m = lau_model(path, start_angle);
  std::vector<double> ev(e, e + W);
  Spline<double> sp = lau_build(m, Wp, ev);
  std::vector<std::vector<V2<double>>> g;
  const double J = lau_cost_grad(sp, lambda, &g);
  std::vector<V2<double>> gw; std::vector<double> gee;
  lau_backprop(m, g, gw, gee);
  for (int s = 0; s + 1 < W; ++s) for (int j = 0; j < 6; ++j) { if (ctrl) { ctrl[(s * 6 + j) * 2] = sp.ctrl[s][j].x; ctrl[(s * 6 + j) * 2 + 1] = sp.ctrl[s][j].y; } }
  for (int i = 0; i < W; ++i) { if (ge) ge[i] = gee[i]; if (gW) { gW[2 * i] = gw[i].x; gW[2 * i + 1] = gw[i].y; } }
  return J;
}

// Run the optimiser; history gets iters+1 costs.  Returns the best cost.
double ora_lau_optimise(const double* path_xy, int W, double start_angle, const scb_opt* opt, double* ctrl, double* e_out,
                        double* history) {
  std::vector<V2<double>> path(W);
  for (int i = 0; i < W; ++i) path[i] = {path_xy[2 * i], path_xy[2 * i + 1]};
  LauOpt o;
  o.iters = opt->iters; o.move_waypoints = opt->move_waypoints; o.lambda = opt->lambda; o.e_init = opt->e_init;
  o.e_min = opt->e_min; o.e_max = opt->e_max; o.step_e = opt->step_e; o.step_w = opt->step_w; o.w_radius = opt->w_radius;
  LauResult<double> r = lau_optimise(lau_model(path, start_angle), o);
  for (int s = 0; s + 1 < W; ++s) for (int j = 0; j < 6; ++j) { ctrl[(s * 6 + j) * 2] = r.spline.ctrl[s][j].x; ctrl[(s * 6 + j) * 2 + 1] = r.spline.ctrl[s][j].y; }
  for (int i = 0; i < W; ++i) if (e_out) e_out[i] = r.e[i];
  for (size_t i = 0; i < r.history.size(); ++i) if (history) history[i] = r.history[i];
  return r.cost;
}

// ---------------------------------------------------------------- the product's batch, on the CPU
// Same structs and layouts as scb_plan_batch (FP64 host buffers).  solve_mode: 0 dense QR
// (reference), 1 Thomas.  Returns elapsed seconds through *seconds.
int ora_plan_batch(const scb_batch_in* in, const scb_batch_out* out, int spline_mode, int solve_mode, int nthreads,
                   double* seconds) {
  const int P = in->P, W = in->W, G = in->G;
  const double *wx = (const double*)in->wx, *wy = (const double*)in->wy, *sa = (const double*)in->start_angle;
  const double *vlo = (const double*)in->vel_lo, *vhi = (const double*)in->vel_hi;
  const double *vlg = (const double*)in->vel_lo_grid, *vhg = (const double*)in->vel_hi_grid;
  const double *alo = (const double*)in->acc_lo, *ahi = (const double*)in->acc_hi;
  const double *vs = (const double*)in->vel_start, *ve = (const double*)in->vel_end;
  const int d = spline_mode == 0 ? 3 : 5, nc = (W - 1) * (d + 1), S_cap = out->S_cap;
  auto t0 = std::chrono::steady_clock::now();
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
#pragma omp parallel for schedule(dynamic)
  for (int p = 0; p < P; ++p) {
    PlanIn pi;
    if (in->stages == SCB_STAGE_TIMING) { pi.timing_only = true; pi.path_len = ((const double*)in->path_len)[p]; }
    else {
      pi.path.resize(W);
      for (int i = 0; i < W; ++i) pi.path[i] = {wx[(size_t)p * W + i], wy[(size_t)p * W + i]};
    }
    pi.start_angle = sa ? sa[p] : NAN;
    std::vector<double> gl, gh;
    if (vlg && vhg) {
      gl.resize(G); gh.resize(G);
      for (int i = 0; i < G; ++i) { gl[i] = vlg[(size_t)i * P + p]; gh[i] = vhg[(size_t)i * P + p]; }
      pi.vel_lo_grid = gl.data(); pi.vel_hi_grid = gh.data();
    } else { pi.vel_lo = vlo[p]; pi.vel_hi = vhi[p]; }
    pi.acc_lo = alo[p]; pi.acc_hi = ahi[p];
    pi.vel_start = vs ? vs[p] : 0.0; pi.vel_end = ve ? ve[p] : 0.0;
    pi.G = G;
    pi.x_cap = in->x_cap > 0 ? in->x_cap : 100.0;
    pi.dt = in->dt > 0 ? in->dt : (double)0.02f;
    pi.omega_max = in->omega_max; pi.acc_lat_max = in->acc_lat_max;
    pi.a_brake = in->a_brake; pi.t_react = in->t_react;
    pi.collision_chords = in->collision_chords;
    pi.param_mode = in->param_mode;
    if (in->obstacle_edges && in->n_obstacle_edges > 0) {
      const double* ed = (const double*)in->obstacle_edges;
      for (int k = 0; k < in->n_obstacle_edges; ++k) pi.edges.push_back({{ed[4 * k], ed[4 * k + 1]}, {ed[4 * k + 2], ed[4 * k + 3]}});
    }
    pi.spline_mode = spline_mode;
    pi.opt.iters = in->opt.iters; pi.opt.move_waypoints = in->opt.move_waypoints; pi.opt.lambda = in->opt.lambda;
    pi.opt.e_init = in->opt.e_init; pi.opt.e_min = in->opt.e_min; pi.opt.e_max = in->opt.e_max;
    pi.opt.step_e = in->opt.step_e; pi.opt.step_w = in->opt.step_w; pi.opt.w_radius = in->opt.w_radius;
    pi.solve_mode = solve_mode;
    pi.S_cap = S_cap;
    PlanOut o = plan_one(pi);
    if (out->status) out->status[p] = o.status;
    if (out->fail_stage) out->fail_stage[p] = o.fail_stage;
    auto put = [&](void* dst, size_t idx, double v) { if (dst) ((double*)dst)[idx] = v; };
    for (int k = 0; k < nc; ++k) {
      put(out->ctrl_x, (size_t)p * nc + k, k < (int)o.ctrl_x.size() ? o.ctrl_x[k] : 0.0);
      put(out->ctrl_y, (size_t)p * nc + k, k < (int)o.ctrl_y.size() ? o.ctrl_y[k] : 0.0);
    }
    put(out->arclen_total, p, o.L); put(out->T_total, p, o.T); put(out->cost, p, o.cost);
    const int nt = (W - 1) * SCB_ARCLEN_TABLE;
    for (int k = 0; k < nt; ++k) put(out->arclen_table, (size_t)p * nt + k, k < (int)o.table.size() ? o.table[k] : 0.0);
    auto grid = [&](void* dst, const std::vector<double>& v) { if (dst) for (int i = 0; i < G; ++i) ((double*)dst)[(size_t)i * P + p] = i < (int)v.size() ? v[i] : 0.0; };
    grid(out->K_lo, o.K_lo); grid(out->K_hi, o.K_hi); grid(out->x, o.x); grid(out->t, o.t);
    grid(out->px, o.px); grid(out->py, o.py); grid(out->v, o.v); grid(out->a, o.a); grid(out->omega, o.omega);
    if (S_cap > 0) {
      if (out->n_samples) out->n_samples[p] = o.n_samples;
      auto strm = [&](void* dst, const std::vector<double>& v) { if (dst) for (int i = 0; i < S_cap; ++i) ((double*)dst)[(size_t)i * P + p] = i < (int)v.size() ? v[i] : 0.0; };
      strm(out->s_time, o.s_time); strm(out->s_pos, o.s_pos); strm(out->s_vel, o.s_vel); strm(out->s_acc, o.s_acc);
      strm(out->s_px, o.s_px); strm(out->s_py, o.s_py); strm(out->s_omega, o.s_omega);
    }
  }
  auto t1 = std::chrono::steady_clock::now();
  if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
  return 0;
}

// ---- global planner front end (ora_fmt.hpp)
void ora_halton(int b, int n, int* state_fi, float* out) {
  ora::HaltonState st; st.f = state_fi[0]; st.i = state_fi[1];
  const std::vector<float> v = ora::halton(b, n, st);
  for (int j = 0; j < n; ++j) out[j] = v[j];
  state_fi[0] = st.f; state_fi[1] = st.i;
}

static std::vector<ora::FmtObstacle> unpack_obstacles(int n_obs, const int* nv, const float* verts, const int* nl, const float* lines,
                                                      const float* rects, const int* closed) {
  std::vector<ora::FmtObstacle> obs(n_obs);
  size_t vo = 0, lo = 0;
  for (int k = 0; k < n_obs; ++k) {
    for (int i = 0; i < nv[k]; ++i, ++vo) obs[k].vertices.push_back(ora::P2{verts[2 * vo], verts[2 * vo + 1]});
    obs[k].lines.assign(lines + 4 * lo, lines + 4 * (lo + nl[k])); lo += nl[k];
    obs[k].x_max = rects[4 * k]; obs[k].x_min = rects[4 * k + 1]; obs[k].y_max = rects[4 * k + 2]; obs[k].y_min = rects[4 * k + 3];
    obs[k].closed = closed[k] != 0;
  }
  return obs;
}

int ora_obstacle_contains(int nv, const float* verts, int nl, const float* lines, const float* rect, int closed, float px, float py) {
  const std::vector<ora::FmtObstacle> o = unpack_obstacles(1, &nv, verts, &nl, lines, rect, &closed);
  return ora::obstacle_contains(o[0], ora::P2{px, py}) ? 1 : 0;
}

// rect = (x_min, x_max, y_min, y_max) of the planning space; state = (x.f, x.i, y.f, y.i); returns the point count
int ora_sample_free(int n, const float* rect, int n_obs, const int* nv, const float* verts, const int* nl, const float* lines,
                    const float* obs_rects, const int* closed, int n_free, const float* free_rects, int* state, float* out_xy,
                    int max_rounds) {
  const std::vector<ora::FmtObstacle> obs = unpack_obstacles(n_obs, nv, verts, nl, lines, obs_rects, closed);
  ora::HaltonState xs, ys; xs.f = state[0]; xs.i = state[1]; ys.f = state[2]; ys.i = state[3];
  const std::vector<float> fr(free_rects, free_rects + 4 * (size_t)n_free);
  const std::vector<ora::P2> pts = ora::sample_free(n, rect, obs, fr, xs, ys, max_rounds);
  for (size_t i = 0; i < pts.size(); ++i) { out_xy[2 * i] = pts[i].x; out_xy[2 * i + 1] = pts[i].y; }
  state[0] = xs.f; state[1] = xs.i; state[2] = ys.f; state[3] = ys.i;
  return (int)pts.size();
}

// Q queries over one node set; path_idx [Q][max_path] node indices (0 = init, 1..n samples, n+1 = goal)
void ora_fmt_batch(int Q, int n, const float* nx, const float* ny, int E, const float* edges, const float* ix, const float* iy,
                   const float* gx, const float* gy, float rn, int max_path, int* status, int* path_len, int* closed,
                   int* path_idx, float* path_x, float* path_y, float* cost) {
  std::vector<ora::P2> samples(n);
  for (int i = 0; i < n; ++i) samples[i] = ora::P2{nx[i], ny[i]};
  const std::vector<float> ed(edges, edges + 4 * (size_t)E);
#pragma omp parallel for schedule(dynamic)
  for (int q = 0; q < Q; ++q) {
    const ora::P2 a{ix[q], iy[q]}, b{gx[q], gy[q]};
    const ora::FmtResult r = ora::fmt_star(samples, ed, a, b, rn);
    status[q] = r.status; closed[q] = r.closed; path_len[q] = (int)r.path.size(); cost[q] = r.status == 0 ? r.cost : FLT_MAX;
    if (r.status == 0 && (int)r.path.size() > max_path) { status[q] = 2; continue; }
    for (size_t k = 0; k < r.path.size(); ++k) {
      const int id = r.path[k];
      const ora::P2 p = id == 0 ? a : (id == n + 1 ? b : samples[id - 1]);
      if (path_idx) path_idx[(size_t)q * max_path + k] = id;
      path_x[(size_t)q * max_path + k] = p.x; path_y[(size_t)q * max_path + k] = p.y;
    }
  }
}

}  // extern "C"
