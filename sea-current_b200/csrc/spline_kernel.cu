// Spline stage on the GPU, one warp per path (sm_100a):
//   tangent heuristics           sea_current.hpp:343-377
//   segment builder from_path    sea_current.hpp:599-683 (REF_CUBIC, control points only: the
//                                reference's 10,001-point preview sampling feeds nothing downstream)
//   Lau09 quintic builder + curvature/arc-length cost + analytic gradient + iRPROP- loop
//                                (LAU_QUINTIC; no reference code — DESIGN.md "New stages")
//   arc-length tables            sea_current.hpp:767-896 (100 sub-intervals x 32-point
//                                Gauss-Legendre per segment, cumulative table of 101 entries)
// Per-path state (waypoints, tangents, control points, optimiser state) is staged in shared
// memory; lanes map to waypoints / segments / quadrature nodes / sub-intervals depending on the
// phase; reductions and scans are warp shuffles.  Per-path arrays are [P][n] so a warp's loads
// and stores are contiguous.
#include "gl32_device.h"
#include "scb_common.cuh"

namespace scb {

constexpr unsigned kFull = 0xffffffffu;

template <class R> __device__ __forceinline__ R warp_sum(R v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(kFull, v, off);
  return v;
}
template <class R> __device__ __forceinline__ R warp_max(R v) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v = fmax(v, __shfl_xor_sync(kFull, v, off));
  return v;
}

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
  const R mult = (test_offset - offset < 0) ? (R)-1 : (R)1;
  R sn, cs;
  sincos(offset + mult * theta, &sn, &cs);
  R lx = sn, ly = -cs;
  const R n2 = lx * lx + ly * ly;
  if (n2 > 0) { const R n = sqrt(n2); lx = lx / n; ly = ly / n; }
  const R m2 = dist2<R>(w1x + lx, w1y + ly, w2x, w2y) < dist2<R>(w1x - lx, w1y - ly, w2x, w2y) ? (R)1 : (R)-1;
  const R mag = (R)0.5 * fmin(du, dv);  // sc:343-345: |W0W1| = |u|, |W1W2| = |v|
  tx = mag * (m2 * lx); ty = mag * (m2 * ly);
}

// Per-warp shared-memory carve-up.
template <class R> struct Smem {
  R* base; int off;
  __device__ __forceinline__ Smem(R* b) : base(b), off(0) {}
  __device__ __forceinline__ R* take(int n) { R* p = base + off; off += n; return p; }
};

__host__ __device__ inline int spline_smem_reals(int W, int deg) {
  const int nseg = W - 1;
  int n = 4 * W + 2 * nseg * (deg + 1) + 2 * nseg * deg;  // waypoints, tangents, ctrl, hodograph
  if (deg == 5) n += 25 * W + 2 * nseg * 6;                 // optimiser state + control-point gradient
  return n;
}

template <class R, int DEG>
__global__ void __launch_bounds__(128) spline_kernel(SplineArgs g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int p = blockIdx.x * wpb + warp;
  if (p >= g.P) return;
  const int W = g.W, nseg = W - 1, NC = DEG + 1;
  Smem<R> sm((R*)smem_raw + (size_t)warp * spline_smem_reals(W, DEG));
  R* wx = sm.take(W); R* wy = sm.take(W); R* tx = sm.take(W); R* ty = sm.take(W);
  R* cx = sm.take(nseg * NC); R* cy = sm.take(nseg * NC);
  R* hx = sm.take(nseg * DEG); R* hy = sm.take(nseg * DEG);

  // ---- load waypoints, validate
  bool bad = false;
  for (int i = lane; i < W; i += 32) {
    const R x = ((const R*)g.wx)[(size_t)p * W + i], y = ((const R*)g.wy)[(size_t)p * W + i];
    wx[i] = x; wy[i] = y;
    bad = bad || !isfinite(x) || !isfinite(y);
  }
  __syncwarp();
  for (int i = lane; i < nseg; i += 32) bad = bad || !(dist2<R>(wx[i], wy[i], wx[i + 1], wy[i + 1]) > 0);
  bad = __any_sync(kFull, bad);
  const int nc = nseg * NC, nt = nseg * kTable;
  R* o_cx = g.ctrl_x ? (R*)g.ctrl_x + (size_t)p * nc : nullptr;
  R* o_cy = g.ctrl_y ? (R*)g.ctrl_y + (size_t)p * nc : nullptr;
  R* o_tab = (R*)g.table + (size_t)p * nt;
  if (bad) {
    for (int i = lane; i < nc; i += 32) { if (o_cx) o_cx[i] = 0; if (o_cy) o_cy[i] = 0; }
    for (int i = lane; i < nt; i += 32) o_tab[i] = 0;
    if (lane == 0) { ((R*)g.L)[p] = 0; if (g.cost) ((R*)g.cost)[p] = 0; g.status[p] = SCB_ST_ERR_BAD_PATH; }
    return;
  }

  // ---- A1: tangents (lanes <-> waypoints)
  for (int i = lane; i < W; i += 32) {
    R ax, ay;
    if (i == 0) {  // sc:371-373, 606-609
      R ang = g.start_angle ? ((const R*)g.start_angle)[p] : Num<R>::nan();
      if (isnan(ang)) ang = atan2(wy[1] - wy[0], wx[1] - wx[0]);
      const R mag = (R)0.5 * dist2<R>(wx[0], wy[0], wx[1], wy[1]);
      R sn, cs; sincos(ang, &sn, &cs);
      ax = mag * cs; ay = mag * sn;
    } else if (i == W - 1) {  // sc:375-377
      const R dx = wx[i] - wx[i - 1], dy = wy[i] - wy[i - 1];
      const R mag = (R)0.5 * dist2<R>(wx[i - 1], wy[i - 1], wx[i], wy[i]);
      const R n2 = dx * dx + dy * dy, n = sqrt(n2);
      ax = mag * (dx / n); ay = mag * (dy / n);
    } else {
      calc_tangent<R>(wx[i - 1], wy[i - 1], wx[i], wy[i], wx[i + 1], wy[i + 1], ax, ay);
    }
    tx[i] = ax; ty[i] = ay;
  }
  __syncwarp();

  R cost = 0;
  if constexpr (DEG == 3) {
    // ---- A2: cubic control points [W_i, W_i + T_i, W_{i+1} - T_{i+1}, W_{i+1}]  (sc:661-681)
    for (int s = lane; s < nseg; s += 32) {
      cx[4 * s] = wx[s]; cy[4 * s] = wy[s];
      cx[4 * s + 1] = wx[s] + tx[s]; cy[4 * s + 1] = wy[s] + ty[s];
      cx[4 * s + 2] = wx[s + 1] - tx[s + 1]; cy[4 * s + 2] = wy[s + 1] - ty[s + 1];
      cx[4 * s + 3] = wx[s + 1]; cy[4 * s + 3] = wy[s + 1];
    }
  } else {
    // ---- A5-A7: Lau quintic + optimiser.  tx/ty hold the frozen base tangents Bt.
    R* w0x = sm.take(W); R* w0y = sm.take(W); R* mag = sm.take(W); R* alpha = sm.take(W);
    R* e = sm.take(W); R* se = sm.take(W); R* pe = sm.take(W);
    R* swx = sm.take(W); R* swy = sm.take(W); R* pwx = sm.take(W); R* pwy = sm.take(W);
    R* be = sm.take(W); R* bwx = sm.take(W); R* bwy = sm.take(W);
    R* Tx = sm.take(W); R* Ty = sm.take(W); R* Ax = sm.take(W); R* Ay = sm.take(W);
    R* gTx = sm.take(W); R* gTy = sm.take(W); R* gAx = sm.take(W); R* gAy = sm.take(W);
    R* gWx = sm.take(W); R* gWy = sm.take(W); R* ge = sm.take(W);
    R* gx = sm.take(nseg * 6); R* gy = sm.take(nseg * 6);
    const scb_opt& o = g.opt;
    for (int i = lane; i < W; i += 32) {
      w0x[i] = wx[i]; w0y[i] = wy[i];
      mag[i] = sqrt(tx[i] * tx[i] + ty[i] * ty[i]);
      R al = 0;
      if (i > 0 && i < W - 1) {
        const R dp = dist2<R>(wx[i - 1], wy[i - 1], wx[i], wy[i]), dn = dist2<R>(wx[i], wy[i], wx[i + 1], wy[i + 1]);
        al = dn / (dp + dn);
      }
      alpha[i] = al;
      e[i] = (R)o.e_init; se[i] = (R)o.step_e; pe[i] = 0;
      swx[i] = swy[i] = (R)o.step_w * mag[i]; pwx[i] = pwy[i] = 0;
    }
    // basis values of B', B'' at this lane's Gauss-Legendre node on [0,1]
    R c1[6], c2[6], wq;
    {
      const double t = 0.5 * GL32_X[lane] + 0.5, s = 1 - t;
      wq = (R)(0.5 * GL32_W[lane]);
      const double b4[5] = {s * s * s * s, 4 * t * s * s * s, 6 * t * t * s * s, 4 * t * t * t * s, t * t * t * t};
      const double b3[4] = {s * s * s, 3 * t * s * s, 3 * t * t * s, t * t * t};
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        const double lo = j > 0 ? b4[j - 1] : 0, hi = j < 5 ? b4[j] : 0;
        c1[j] = (R)(5 * (lo - hi));
        const double a = j >= 2 ? b3[j - 2] : 0, b = (j >= 1 && j <= 4) ? b3[j - 1] : 0, c = j <= 3 ? b3[j] : 0;
        c2[j] = (R)(20 * (a - 2 * b + c));
      }
    }
    __syncwarp();
    const R lambda = (R)o.lambda;
    R bestJ = 0;
    auto build = [&]() {
      for (int i = lane; i < W; i += 32) { Tx[i] = e[i] * tx[i]; Ty[i] = e[i] * ty[i]; }
      __syncwarp();
      for (int i = lane; i < W; i += 32) {
        R px = 0, py = 0, nx = 0, ny = 0;
        if (i > 0) {
          px = (R)6 * (wx[i - 1] - wx[i]) + (R)4 * Tx[i] + (R)2 * Tx[i - 1];
          py = (R)6 * (wy[i - 1] - wy[i]) + (R)4 * Ty[i] + (R)2 * Ty[i - 1];
        }
        if (i + 1 < W) {
          nx = (R)6 * (wx[i + 1] - wx[i]) - (R)4 * Tx[i] - (R)2 * Tx[i + 1];
          ny = (R)6 * (wy[i + 1] - wy[i]) - (R)4 * Ty[i] - (R)2 * Ty[i + 1];
        }
        if (i == 0) { Ax[i] = nx; Ay[i] = ny; }
        else if (i + 1 == W) { Ax[i] = px; Ay[i] = py; }
        else { Ax[i] = alpha[i] * px + ((R)1 - alpha[i]) * nx; Ay[i] = alpha[i] * py + ((R)1 - alpha[i]) * ny; }
      }
      __syncwarp();
      for (int s = lane; s < nseg; s += 32) {
        const R fifth = (R)1 / (R)5, twentieth = (R)1 / (R)20;
        const R p0x = wx[s], p0y = wy[s], p5x = wx[s + 1], p5y = wy[s + 1];
        const R p1x = p0x + fifth * Tx[s], p1y = p0y + fifth * Ty[s];
        const R p4x = p5x - fifth * Tx[s + 1], p4y = p5y - fifth * Ty[s + 1];
        cx[6 * s] = p0x; cy[6 * s] = p0y; cx[6 * s + 1] = p1x; cy[6 * s + 1] = p1y;
        cx[6 * s + 2] = twentieth * Ax[s] + (R)2 * p1x - p0x; cy[6 * s + 2] = twentieth * Ay[s] + (R)2 * p1y - p0y;
        cx[6 * s + 3] = twentieth * Ax[s + 1] + (R)2 * p4x - p5x; cy[6 * s + 3] = twentieth * Ay[s + 1] + (R)2 * p4y - p5y;
        cx[6 * s + 4] = p4x; cy[6 * s + 4] = p4y; cx[6 * s + 5] = p5x; cy[6 * s + 5] = p5y;
      }
      __syncwarp();
    };
    const int K = o.iters;
    for (int it = 0; it <= K; ++it) {
      build();
      // ---- A6: cost and control-point gradient; lanes <-> the 32 nodes of one segment
      R J = 0;
      for (int s = 0; s < nseg; ++s) {
        R x1 = 0, y1 = 0, x2 = 0, y2 = 0;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          const R px = cx[6 * s + j], py = cy[6 * s + j];
          x1 += c1[j] * px; y1 += c1[j] * py; x2 += c2[j] * px; y2 += c2[j] * py;
        }
        const R n2 = x1 * x1 + y1 * y1, n = sqrt(n2);
        const R c = x1 * y2 - y1 * x2;
        const R in5 = (R)1 / (n2 * n2 * n);
        const R Js = warp_sum<R>(wq * (c * c * in5 + lambda * n));
        J += Js;
        if (it < K) {
          const R k1 = (R)2 * c * in5, k2 = (R)5 * c * c * in5 / n2, k3 = lambda / n;
          const R fx1 = k1 * y2 - k2 * x1 + k3 * x1, fy1 = -k1 * x2 - k2 * y1 + k3 * y1;
          const R fx2 = -k1 * y1, fy2 = k1 * x1;
#pragma unroll
          for (int j = 0; j < 6; ++j) {
            const R vx = warp_sum<R>(wq * (fx1 * c1[j] + fx2 * c2[j]));
            const R vy = warp_sum<R>(wq * (fy1 * c1[j] + fy2 * c2[j]));
            if (lane == j) { gx[6 * s + j] = vx; gy[6 * s + j] = vy; }
          }
        }
      }
      if (it == 0 || J < bestJ) {
        bestJ = J;
        for (int i = lane; i < W; i += 32) { be[i] = e[i]; bwx[i] = wx[i]; bwy[i] = wy[i]; }
      }
      if (it == K) break;
      __syncwarp();
      // ---- chain rule to (T, A, W) per waypoint, gathering from the two adjacent segments
      for (int i = lane; i < W; i += 32) {
        R wxs = 0, wys = 0, txs = 0, tys = 0, axs = 0, ays = 0;
        if (i > 0) {
          const R* sx = gx + 6 * (i - 1); const R* sy = gy + 6 * (i - 1);
          wxs += (sx[3] + sx[4] + sx[5]); wys += (sy[3] + sy[4] + sy[5]);
          txs -= ((R)0.4 * sx[3] + (R)0.2 * sx[4]); tys -= ((R)0.4 * sy[3] + (R)0.2 * sy[4]);
          axs += (R)0.05 * sx[3]; ays += (R)0.05 * sy[3];
        }
        if (i < nseg) {
          const R* sx = gx + 6 * i; const R* sy = gy + 6 * i;
          wxs += (sx[0] + sx[1] + sx[2]); wys += (sy[0] + sy[1] + sy[2]);
          txs += ((R)0.2 * sx[1] + (R)0.4 * sx[2]); tys += ((R)0.2 * sy[1] + (R)0.4 * sy[2]);
          axs += (R)0.05 * sx[2]; ays += (R)0.05 * sy[2];
        }
        gWx[i] = wxs; gWy[i] = wys; gTx[i] = txs; gTy[i] = tys; gAx[i] = axs; gAy[i] = ays;
      }
      __syncwarp();
      R gemax = 0, gwmax = 0;
      for (int i = lane; i < W; i += 32) {
        auto ap = [&](int k) { return k == 0 ? (R)0 : (k == W - 1 ? (R)1 : alpha[k]); };
        auto an = [&](int k) { return k == W - 1 ? (R)0 : (k == 0 ? (R)1 : (R)1 - alpha[k]); };
        R wxs = gWx[i], wys = gWy[i], txs = gTx[i], tys = gTy[i];
        if (i > 0) {  // waypoint i is the "next" of i-1
          const R k6 = (R)6 * an(i - 1), k2 = (R)2 * an(i - 1);
          wxs += k6 * gAx[i - 1]; wys += k6 * gAy[i - 1]; txs -= k2 * gAx[i - 1]; tys -= k2 * gAy[i - 1];
        }
        {
          const R k6 = (R)6 * (ap(i) + an(i)), k4 = (R)4 * (ap(i) - an(i));
          wxs -= k6 * gAx[i]; wys -= k6 * gAy[i]; txs += k4 * gAx[i]; tys += k4 * gAy[i];
        }
        if (i + 1 < W) {  // waypoint i is the "prev" of i+1
          const R k6 = (R)6 * ap(i + 1), k2 = (R)2 * ap(i + 1);
          wxs += k6 * gAx[i + 1]; wys += k6 * gAy[i + 1]; txs += k2 * gAx[i + 1]; tys += k2 * gAy[i + 1];
        }
        const R gei = txs * tx[i] + tys * ty[i];
        ge[i] = gei;
        gemax = fmax(gemax, fabs(gei));
        if (i > 0 && i + 1 < W) gwmax = fmax(gwmax, fmax(fabs(wxs), fabs(wys)));
        // gW is final only now; stash in registers until every lane has read gWx/gWy/gTx/gTy
        Tx[i] = wxs; Ty[i] = wys;  // Tx/Ty are free until the next build()
      }
      gemax = warp_max<R>(gemax); gwmax = warp_max<R>(gwmax);
      __syncwarp();
      // ---- A7: iRPROP-
      const R te = (R)1e-9 * gemax, tw = (R)1e-9 * gwmax;
      auto rprop1 = [](R gv, R& gprev, R& step, R smin, R smax) -> R {
        const R s = gv * gprev;
        if (s > 0) step = fmin(step * (R)1.2, smax);
        else if (s < 0) { step = fmax(step * (R)0.5, smin); gv = 0; }
        gprev = gv;
        return gv > 0 ? -step : (gv < 0 ? step : (R)0);
      };
      for (int i = lane; i < W; i += 32) {
        const R gi = fabs(ge[i]) <= te ? (R)0 : ge[i];
        R ei = e[i] + rprop1(gi, pe[i], se[i], (R)1e-6, (R)1.0);
        e[i] = fmin((R)o.e_max, fmax((R)o.e_min, ei));
        if (o.move_waypoints && i > 0 && i + 1 < W) {
          const R gvx = fabs(Tx[i]) <= tw ? (R)0 : Tx[i], gvy = fabs(Ty[i]) <= tw ? (R)0 : Ty[i];
          const R smin = (R)1e-6 * mag[i], smax = (R)0.5 * mag[i], r = (R)o.w_radius;
          R nx = wx[i] + rprop1(gvx, pwx[i], swx[i], smin, smax);
          R ny = wy[i] + rprop1(gvy, pwy[i], swy[i], smin, smax);
          wx[i] = fmin(w0x[i] + r, fmax(w0x[i] - r, nx));
          wy[i] = fmin(w0y[i] + r, fmax(w0y[i] - r, ny));
        }
      }
      __syncwarp();
    }
    __syncwarp();
    for (int i = lane; i < W; i += 32) { e[i] = be[i]; wx[i] = bwx[i]; wy[i] = bwy[i]; }
    __syncwarp();
    build();
    cost = bestJ;
  }
  __syncwarp();

  // ---- control points out (coalesced), hodograph in power basis per segment
  for (int i = lane; i < nc; i += 32) { if (o_cx) o_cx[i] = cx[i]; if (o_cy) o_cy[i] = cy[i]; }
  for (int s = lane; s < nseg; s += 32) {
    R dx[DEG], dy[DEG];  // hodograph control points DEG * (P[j+1] - P[j])  (sc:1041-1053)
#pragma unroll
    for (int j = 0; j < DEG; ++j) { dx[j] = (R)DEG * (cx[NC * s + j + 1] - cx[NC * s + j]); dy[j] = (R)DEG * (cy[NC * s + j + 1] - cy[NC * s + j]); }
    R binom = 1;
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      hx[DEG * s + k] = binom * dx[0]; hy[DEG * s + k] = binom * dy[0];
#pragma unroll
      for (int j = 0; j + k < DEG - 1; ++j) { dx[j] = dx[j + 1] - dx[j]; dy[j] = dy[j + 1] - dy[j]; }
      binom = binom * (R)(DEG - 1 - k) / (R)(k + 1);
    }
  }
  __syncwarp();

  // ---- A4: arc-length pieces; lanes <-> sub-intervals, 32 nodes sequentially (sc:854-880),
  // then a segmented warp scan gives the cumulative tables (sc:882-886) without re-reading HBM.
  const int items = nseg * kSub;
  R Lpart = 0, carry = 0;
  int carry_seg = -1;
  for (int base = 0; base < items; base += 32) {
    const int it = base + lane;
    const bool live = it < items;
    const int s = live ? it / kSub : nseg - 1, k = live ? it % kSub : kSub - 1;
    const R a = fmin((R)k * (R)0.01, (R)1), b = fmin((R)(k + 1) * (R)0.01, (R)1);
    const R half = (b - a) / 2, mid = (b + a) / 2;
    R acc = 0;
    if (live) {
      R kx[DEG], ky[DEG];
#pragma unroll
      for (int j = 0; j < DEG; ++j) { kx[j] = hx[DEG * s + j]; ky[j] = hy[DEG * s + j]; }
#pragma unroll 4
      for (int i = 0; i < 32; ++i) {
        const R u = half * (R)GL32_X[i] + mid;
        R dx = kx[DEG - 1], dy = ky[DEG - 1];
#pragma unroll
        for (int j = DEG - 2; j >= 0; --j) { dx = dx * u + kx[j]; dy = dy * u + ky[j]; }
        acc += (R)GL32_W[i] * sqrt(dx * dx + dy * dy);
      }
      acc *= half;
    }
    // inclusive scan within the segment
    R v = acc;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const R t = __shfl_up_sync(kFull, v, off);
      const int ts = __shfl_up_sync(kFull, s, off);
      if (lane >= off && ts == s) v += t;
    }
    if (s == carry_seg) v += carry;
    if (live) {
      o_tab[s * kTable + k + 1] = v;
      if (k == 0) o_tab[s * kTable] = 0;
      if (k == kSub - 1) Lpart += v;
    }
    carry = __shfl_sync(kFull, v, 31);
    carry_seg = __shfl_sync(kFull, s, 31);
  }
  const R L = warp_sum<R>(Lpart);
  if (lane == 0) {
    ((R*)g.L)[p] = L;
    if (g.cost) ((R*)g.cost)[p] = cost;
    g.status[p] = (L > 0 && isfinite(L)) ? SCB_ST_OK : SCB_ST_ERR_BAD_PATH;
  }
}

size_t spline_smem_bytes(int W, int spline_mode, int real_size, int warps) {
  return (size_t)spline_smem_reals(W, spline_mode == SCB_REF_CUBIC ? 3 : 5) * real_size * warps;
}

template <class R> void launch_spline(const SplineArgs& a, int spline_mode, cudaStream_t st) {
  const int deg = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  int warps = 4;
  while (warps > 1 && (size_t)spline_smem_reals(a.W, deg) * sizeof(R) * warps > 200 * 1024) warps >>= 1;
  const size_t smem = (size_t)spline_smem_reals(a.W, deg) * sizeof(R) * warps;
  const int grid = (a.P + warps - 1) / warps;
  if (deg == 3) {
    cudaFuncSetAttribute(spline_kernel<R, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    spline_kernel<R, 3><<<grid, warps * 32, smem, st>>>(a);
  } else {
    cudaFuncSetAttribute(spline_kernel<R, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    spline_kernel<R, 5><<<grid, warps * 32, smem, st>>>(a);
  }
}
template void launch_spline<double>(const SplineArgs&, int, cudaStream_t);
template void launch_spline<float>(const SplineArgs&, int, cudaStream_t);

}  // namespace scb
