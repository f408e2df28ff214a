// ORACLE — TEST INFRASTRUCTURE ONLY (see ora_spline.hpp header).
// Arc-length tables (A4), Chebyshev fit/eval (A16), resample (A16), curvature / angular
// velocity (A17) restated from /root/reference/sea_current.hpp ("sc:LINE").
#pragma once
#include "gl32.h"
#include "ora_spline.hpp"

namespace ora {

// sc:322-326
template <class R> struct ArclengthData {
  R arclength = 0;
  std::vector<std::vector<R>> segments;   // cumulative length per segment, n+1 entries
  std::vector<std::vector<R>> positions;  // curve parameter of each entry
};

// sc:767-896.  precision = 0.01 => 100 sub-intervals x 32 Gauss–Legendre nodes per segment.
// Mixed arithmetic kept as in the reference: node position and weight*norm products go
// through double, every accumulator is R.
template <class R>
inline ArclengthData<R> arclength(const Spline<R>& sp, R precision = (R)0.01, int eval_mode = EVAL_BF) {
  ArclengthData<R> ad;
  const int nseg = sp.n_segments();
  ad.segments.resize(nseg); ad.positions.resize(nseg);
  R total = 0;
  const int n = (int)std::round(1.0f / (float)precision);
  for (int s = 0; s < nseg; ++s) {
    std::vector<R> pos(n + 1);
    for (int i = 0; i <= n; ++i) pos[i] = std::min((R)i * precision, (R)1);
    std::vector<R> deriv_pos((size_t)32 * n);
    for (int k = 0; k < n; ++k) {
      const R b = pos[k + 1], a = pos[k];
      for (int i = 0; i < 32; ++i)
        deriv_pos[(size_t)k * 32 + i] = (R)((double)((b - a) / 2) * GL32_X[i] + (double)((b + a) / 2));
    }
    std::vector<V2<R>> d;
    eval_curve(hodograph_ctrl(sp.ctrl[s]), deriv_pos, eval_mode, d);
    std::vector<R> seg(n + 1, (R)0), segpos(n + 1, (R)0);
    for (int k = 0; k < n; ++k) {
      const R b = pos[k + 1], a = pos[k];
      for (int i = 0; i < 32; ++i) {
        const V2<R> p = d[(size_t)k * 32 + i];
        seg[k + 1] = (R)((double)seg[k + 1] + GL32_W[i] * (double)std::sqrt((p.x * p.x) + (p.y * p.y)));
      }
      seg[k + 1] *= ((b - a) / (R)2);
      segpos[k + 1] = b;
      total += seg[k + 1];
    }
    R sum = 0;
    for (R& arc : seg) { sum += arc; arc = sum; }
    ad.segments[s] = seg; ad.positions[s] = segpos;
  }
  ad.arclength = total;
  return ad;
}

// ---------------------------------------------------------------- chebfit / chebeval
// sc:329-335
template <class R> struct ChebPoly { std::vector<R> coeffs; R xmin, xmax; };

// Least squares min ||T c - y|| by Householder QR (Eigen::HouseholderQR<MatrixXf>::solve,
// sc:1134-1137; Eigen is not in the tree — standard algorithm restated).
template <class R>
inline std::vector<R> householder_lstsq(std::vector<R> A, int m, int n, std::vector<R> y) {
  // A row-major m x n
  for (int k = 0; k < n; ++k) {
    R tail = 0;
    for (int i = k + 1; i < m; ++i) tail += A[(size_t)i * n + k] * A[(size_t)i * n + k];
    const R c0 = A[(size_t)k * n + k];
    if (tail == (R)0) continue;
    R beta = std::sqrt(c0 * c0 + tail);
    if (c0 >= 0) beta = -beta;
    const R tau = (beta - c0) / beta;
    const R inv = (R)1 / (c0 - beta);
    for (int i = k + 1; i < m; ++i) A[(size_t)i * n + k] *= inv;  // essential part, v0 = 1
    A[(size_t)k * n + k] = beta;
    for (int j = k + 1; j < n; ++j) {
      R dot = A[(size_t)k * n + j];
      for (int i = k + 1; i < m; ++i) dot += A[(size_t)i * n + k] * A[(size_t)i * n + j];
      dot *= tau;
      A[(size_t)k * n + j] -= dot;
      for (int i = k + 1; i < m; ++i) A[(size_t)i * n + j] -= dot * A[(size_t)i * n + k];
    }
    R dot = y[k];
    for (int i = k + 1; i < m; ++i) dot += A[(size_t)i * n + k] * y[i];
    dot *= tau;
    y[k] -= dot;
    for (int i = k + 1; i < m; ++i) y[i] -= dot * A[(size_t)i * n + k];
  }
  std::vector<R> c(n);
  for (int k = n - 1; k >= 0; --k) {
    R s = y[k];
    for (int j = k + 1; j < n; ++j) s -= A[(size_t)k * n + j] * c[j];
    c[k] = s / A[(size_t)k * n + k];
  }
  return c;
}

// sc:1109-1138.  NB `degree` is the number of basis functions n (polynomial degree n-1).
template <class R>
inline ChebPoly<R> chebfit(const std::vector<R>& x, const std::vector<R>& y, int degree) {
  const int n = degree, m = (int)x.size();
  const R xmax = *std::max_element(x.begin(), x.end());
  const R xmin = *std::min_element(x.begin(), x.end());
  std::vector<R> xn(m);
  for (int i = 0; i < m; ++i) xn[i] = ((R)2 * x[i] - (xmax + xmin)) / (xmax - xmin);
  std::vector<R> T((size_t)m * n, (R)0);
  for (int i = 0; i < m; ++i) {
    T[(size_t)i * n + 0] = 1;
    if (n >= 1) T[(size_t)i * n + 1] = xn[i];
    for (int j = 2; j < n; ++j)
      T[(size_t)i * n + j] = ((R)2 * xn[i]) * T[(size_t)i * n + j - 1] - T[(size_t)i * n + j - 2];
  }
  return {householder_lstsq(T, m, n, y), xmin, xmax};
}

// sc:1140-1170
template <class R>
inline std::vector<R> chebeval(const std::vector<R>& x, const ChebPoly<R>& b, int degree) {
  const int n = degree, m = (int)x.size();
  std::vector<R> y(m, (R)0);
  for (int i = 0; i < m; ++i) {
    const R xn = ((R)2 * x[i] - (b.xmax + b.xmin)) / (b.xmax - b.xmin);
    R t0 = 1, t1 = xn;
    R acc = (R)0;
    acc += b.coeffs[0] * t0;
    if (n >= 1) acc += b.coeffs[1] * t1;
    for (int j = 2; j < n; ++j) {
      const R t2 = ((R)2 * xn) * t1 - t0;
      acc += b.coeffs[j] * t2;
      t0 = t1; t1 = t2;
    }
    y[i] = acc;
  }
  return y;
}

// ---------------------------------------------------------------- A16: resample
// sc:898-1005.  Returns, for every output row, the segment index, the curve parameter and
// the evaluated point.  Quirks kept: (1) block offsets are profile samples,
// not segment lengths (sc:942-944); (2) the last sample of block i is also the first of
// block i+1, the final profile sample is never used and surplus rows are trimmed from the
// end of the last block (sc:937-939, 986-996).
template <class R> struct Resampled {
  std::vector<int> seg;
  std::vector<R> u;
  std::vector<V2<R>> pts;
};

template <class R>
inline Resampled<R> resample(const Spline<R>& sp, std::vector<R>& profile_pos, const ArclengthData<R>& ad,
                             bool nudge_positions, int eval_mode = EVAL_BF) {
  const int rows = (int)profile_pos.size();
  if (nudge_positions) {
    profile_pos[0] = 0;
    profile_pos[rows - 1] = ad.arclength;
    for (int i = 1; i < rows - 1; ++i) {
      if (profile_pos[i] < profile_pos[i - 1] || profile_pos[i] > profile_pos[i + 1])
        profile_pos[i] = (profile_pos[i - 1] + profile_pos[i + 1]) / 2;
      if (profile_pos[i] < 0) profile_pos[i] = 0;
      if (profile_pos[i] > ad.arclength) profile_pos[i] = ad.arclength;
    }
  }
  const int nseg = sp.n_segments();
  Resampled<R> out;
  int j = 0;
  R offset = 0;
  size_t block_size_sum = 0;
  std::vector<size_t> block_rows(nseg);
  for (int i = 0; i < nseg; ++i) {
    const std::vector<R>& seg = ad.segments[i];
    const R last = seg.back();
    const int start = j;
    for (; j < rows && (profile_pos[j] - offset <= last); ++j) {}
    if (i + 1 == nseg && i == 0) j = rows;
    else if (i + 1 == nseg) j = rows - 1;
    j -= 1;
    offset = profile_pos[j];
    std::vector<R> block((size_t)std::max(0, (j + 1) - start));
    for (size_t k = 0; k < block.size(); ++k) block[k] = profile_pos[start + k] - profile_pos[start];
    block_size_sum += block.size();
    block_rows[i] = block.size();
    const int degree = (int)std::min<long>(10L, (long)seg.size());
    const ChebPoly<R> poly = chebfit(seg, ad.positions[i], degree);
    std::vector<R> fixed = chebeval(block, poly, degree);
    for (R& f : fixed) { if (f < 0) f = 0; if (f > 1) f = 1; }
    std::vector<V2<R>> pts;
    eval_curve(sp.ctrl[i], fixed, eval_mode, pts);
    for (size_t k = 0; k < fixed.size(); ++k) { out.seg.push_back(i); out.u.push_back(fixed[k]); out.pts.push_back(pts[k]); }
  }
  if (block_size_sum > (size_t)rows) {
    const size_t diff = block_size_sum - rows;
    out.seg.resize(out.seg.size() - diff); out.u.resize(out.u.size() - diff); out.pts.resize(out.pts.size() - diff);
  }
  return out;
}

// ---------------------------------------------------------------- A17
// sc:1017-1039: kappa = (x'y'' - y'x'') / (x'^2 + y'^2)^1.5, via two hodographs evaluated
// at the same parameters.  std::pow(float, 1.5) runs in double.
template <class R>
inline std::vector<R> curvature(const Spline<R>& sp, const Resampled<R>& rs, int eval_mode = EVAL_BF) {
  std::vector<R> k(rs.u.size());
  size_t i0 = 0;
  while (i0 < rs.u.size()) {  // one hodograph evaluation per run of samples on the same segment
    size_t i1 = i0;
    while (i1 < rs.u.size() && rs.seg[i1] == rs.seg[i0]) ++i1;
    auto d1c = hodograph_ctrl(sp.ctrl[rs.seg[i0]]);
    auto d2c = hodograph_ctrl(d1c);
    std::vector<R> p(rs.u.begin() + i0, rs.u.begin() + i1);
    std::vector<V2<R>> d1, d2;
    eval_curve(d1c, p, eval_mode, d1); eval_curve(d2c, p, eval_mode, d2);
    for (size_t i = i0; i < i1; ++i) {
      const V2<R> a = d1[i - i0], b = d2[i - i0];
      k[i] = (R)((double)(a.x * b.y - a.y * b.x) / std::pow((double)(a.x * a.x + a.y * a.y), 1.5));
    }
    i0 = i1;
  }
  return k;
}

// sc:1055-1067
template <class R>
inline std::vector<R> angular_velocity(const std::vector<R>& curv, const std::vector<R>& vel) {
  std::vector<R> w(curv.size());
  for (size_t i = 0; i < w.size(); ++i) w[i] = vel[i] * curv[i];
  return w;
}

}  // namespace ora
