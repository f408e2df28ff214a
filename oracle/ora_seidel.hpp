// ORACLE — TEST INFRASTRUCTURE ONLY (see ora_spline.hpp header).
// toppra's Seidel LP solver restated: tp:solver/seidel-internal.hpp:54-57,73-133 (solve_lp1d)
// and tp:solver/seidel.cpp:14-25,30-39,50-198 (value, project_linear_constraint, solve_lp2d).
#pragma once
#include "ora_toppra.hpp"

namespace ora {
namespace tp {

static const double TINY = 1e-10;           // seidel-internal.hpp:54
static const double REL_TOLERANCE = 1e-10;  // :56
static const double ABS_TOLERANCE = 1e-13;  // :57
enum { LOW_0 = -1, HIGH_0 = -2, LOW_1 = -3, HIGH_1 = -4 };
inline int LOWc(int i) { return i == 0 ? LOW_0 : LOW_1; }
inline int HIGHc(int i) { return i == 0 ? HIGH_0 : HIGH_1; }

struct LpSol1d { bool feasible; double optvar; int active_c; };
struct LpSol { bool feasible = false; double optval = 0; double optvar[2] = {0, 0}; int active_c[2] = {0, 0}; };

// max v0*x + v1  s.t.  a[i]*x + b[i] <= 0.  A1d is rows of (a, b).
inline LpSol1d solve_lp1d(double v0, double /*v1*/, const std::vector<std::array<double, 2>>& A, int nrows) {
  double cur_min = -INF, cur_max = INF;
  int active_c_min = -1, active_c_max = -2;
  const bool maximize = v0 > 0;
  for (int i = 0; i < nrows; ++i) {
    const double a = A[i][0], b = A[i][1];
    if (std::fabs(a) < ABS_TOLERANCE) {
      if (b > ABS_TOLERANCE) return {false, 0, 0};
      continue;
    }
    if (a > 0) {
      if (a * cur_max + b > 0) { cur_max = std::min(-b / a, cur_max); active_c_max = i; }
    } else if (a < 0 && a * cur_min + b > 0) {
      cur_min = std::max(-b / a, cur_min); active_c_min = i;
    }
  }
  if (cur_min - cur_max > std::max(std::fabs(cur_min), std::fabs(cur_max)) * REL_TOLERANCE ||
      cur_min == INF || cur_max == -INF)
    return {false, 0, 0};
  if (v0 == 0) {
    if (cur_min != -INF) {
      if (cur_max != INF) return {true, (cur_max / 2 + cur_min / 2), active_c_min};
      return {true, cur_min, active_c_min};
    }
    return {true, cur_max, active_c_max};
  }
  if (maximize) return {true, cur_max, active_c_max};
  return {true, cur_min, active_c_min};
}

// seidel.cpp:14-25: coeffs . [vars 1] with 0 * inf := 0
inline double value3(const double* row, const double* vars) {
  double res = row[2];
  for (int i = 0; i < 2; ++i) res += (row[i] == 0 && !std::isfinite(vars[i]) ? 0 : row[i] * vars[i]);
  return res;
}
// seidel.cpp:30-39
inline std::array<double, 2> project(const double* Aj, const double* d_tan, const double* zero_prj) {
  return {Aj[0] * d_tan[0] + Aj[1] * d_tan[1], value3(Aj, zero_prj)};
}

// seidel.cpp:50-198.  A: nrows x 3 row-major.
inline LpSol solve_lp2d(const double v[2], const std::vector<std::array<double, 3>>& A, const double low[2],
                        const double high[2]) {
  const int nrows = (int)A.size();
  std::vector<std::array<double, 2>> A_1d(nrows + 4);
  double cur[2];
  LpSol sol;
  for (int i = 0; i < 2; ++i) {
    if (low[i] > high[i]) {
      if (low[i] - high[i] > std::max(std::fabs(low[i]), std::fabs(high[i])) * REL_TOLERANCE ||
          low[i] == INF || high[i] == -INF)
        return LpSol{};
      const int j = 1 - i;
      sol.optvar[i] = (low[i] + high[i]) / 2;
      for (int r = 0; r < nrows; ++r) A_1d[r] = {A[r][j], A[r][2] + sol.optvar[i] * A[r][i]};
      A_1d[nrows] = {-1, low[j]};
      A_1d[nrows + 1] = {1, -high[j]};
      LpSol1d s1 = solve_lp1d(v[j], 0., A_1d, nrows + 2);
      if (!s1.feasible) return LpSol{};
      sol.optvar[j] = s1.optvar;
      sol.feasible = true;
      sol.optval = v[0] * sol.optvar[0] + v[1] * sol.optvar[1];
      sol.active_c[0] = v[i] > 0 ? HIGHc(i) : LOWc(i);
      switch (s1.active_c - nrows) {
        case 0: sol.active_c[1] = LOWc(j); break;
        case 1: sol.active_c[1] = HIGHc(j); break;
        default: sol.active_c[1] = s1.active_c;
      }
      return sol;
    }
    cur[i] = v[i] > 0 ? high[i] : low[i];
    sol.active_c[i] = v[i] > 0 ? HIGHc(i) : LOWc(i);
  }
  for (int i = 0; i < nrows; ++i) {
    if (value3(A[i].data(), cur) < ABS_TOLERANCE) continue;
    sol.active_c[0] = i;
    double zero_prj[2] = {A[i][0], A[i][1]};
    const double sq = zero_prj[0] * zero_prj[0] + zero_prj[1] * zero_prj[1];
    const double f = -A[i][2] / sq;
    zero_prj[0] *= f; zero_prj[1] *= f;
    const double d_tan[2] = {-A[i][1], A[i][0]};
    const double v1d = v[0] * d_tan[0] + v[1] * d_tan[1];
    for (int j = 0; j < i; ++j) A_1d[j] = project(A[j].data(), d_tan, zero_prj);
    const double r0[3] = {-1., 0., low[0]}, r1[3] = {1., 0., -high[0]}, r2[3] = {0., -1., low[1]}, r3[3] = {0., 1., -high[1]};
    A_1d[i] = project(r0, d_tan, zero_prj);
    A_1d[i + 1] = project(r1, d_tan, zero_prj);
    A_1d[i + 2] = project(r2, d_tan, zero_prj);
    A_1d[i + 3] = project(r3, d_tan, zero_prj);
    LpSol1d s1 = solve_lp1d(v1d, 0., A_1d, 4 + i);
    if (!s1.feasible) return LpSol{};
    cur[0] = zero_prj[0] + s1.optvar * d_tan[0];
    cur[1] = zero_prj[1] + s1.optvar * d_tan[1];
    if (s1.active_c < 0) { if (i < 2) sol.active_c[i] = s1.active_c; }  // reference writes past a 2-array for i >= 2 (UB)
    else if (s1.active_c >= i) sol.active_c[1] = i - s1.active_c - 1;
    else sol.active_c[1] = s1.active_c;
  }
  sol.feasible = true;
  sol.optvar[0] = cur[0]; sol.optvar[1] = cur[1];
  sol.optval = v[0] * cur[0] + v[1] * cur[1];
  return sol;
}

}  // namespace tp
}  // namespace ora
