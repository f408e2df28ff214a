// Path stage, warp-team variant (sm_100a): one WARP owns PB consecutive paths from the waypoints to
// the per-gridpoint geometry, and never meets another warp at a barrier.  Same stages, same
// arithmetic definitions and same outputs as path_kernel.cu (which remains the variant for paths
// whose state does not fit a warp's slice of shared memory):
//   tangent heuristics sc:343-377, segment builder sc:599-683, Lau09 quintic + cost/gradient +
//   iRPROP- (DESIGN.md "New stages"), arc-length tables sc:767-896, curve point and curvature at the
//   arc length q(s_i) of every toppra gridpoint (resample sc:898-1005 + curvature sc:1017-1039).
//
// Why warps and not CTAs (profiles/r1b_*): the CTA-team kernel spent 18 % of its stall samples at
// __syncthreads (the optimiser has ~9 short dependent phases per iteration) and held the FP64 pipe
// at 46 %.  Here the phases of one path are ordered by __syncwarp only, the 16 resident warps of an
// SM are in different phases at any time, and the FP64 pipe sees a steady mix.
//
// Work mapping inside a warp (lanes are re-mapped phase by phase):
//   (path, waypoint) / (path, segment) items           tangents, builder, chain rule, iRPROP-
//   (path, segment, 4-node group) items, 8 per segment  cost + control-point gradient; the 13 partial
//                                                       sums of a lane go through a [13][32] slab of
//                                                       shared memory and are added in lane order
//   (path, segment, sub-interval) items                arc-length pieces
//   (gridpoint, path) items, path fastest              geometry at the toppra gridpoints
//
// Arc-length pieces: the reference integrates |B'| over every 0.01-wide sub-interval with a 32-point
// Gauss-Legendre rule.  On such an interval the integrand is analytic and the 6-point rule already
// agrees with the 32-point one to ~1e-15 unless the speed has a (near-)zero within a few interval
// widths.  Each lane therefore evaluates the 4- and the 6-point rules (10 nodes) and accepts the
// 6-point value when the two agree to 1e-11 relative (FP32: 2e-6); otherwise it runs the reference's
// 32 nodes in the reference's order.  Measured on C2/C3: 0.1 % of the pieces take the long way and
// accepted pieces differ from the 32-point value by < 1e-11 relative (DESIGN.md "Arc length").
#include "gl32_device.h"
#include "path_device.cuh"

namespace scb {

constexpr int kWarpsPerCta = 4;
constexpr int kLanesPerSeg = 8;                  // cost/gradient: lanes per segment ...
constexpr int kNodesPerItem = 32 / kLanesPerSeg; // ... and quadrature nodes per lane
constexpr int kPart = 13;                        // J + 6 x-gradients + 6 y-gradients
constexpr int kPartStride = 33;                  // row stride of the [13][32] partial slab (bank-conflict-free reads)

// Shared-memory layout of one warp's slice (host and device agree through this struct).
struct WarpLayout {
  int W, nseg, PB, deg;
  size_t ints_bytes, segend_off, reals_off;  // byte offsets inside the slice
  // offsets in reals from reals_off
  size_t scal, wx, wy, tx, ty, cx, cy, uni;  // uni: start of the optimiser / tail union
  size_t total_reals;
  __host__ __device__ WarpLayout(int W_, int PB_, int deg_) : W(W_), nseg(W_ - 1), PB(PB_), deg(deg_) {
    ints_bytes = 64;                                  // bad[8], improve[8]
    segend_off = ints_bytes;
    reals_off = ints_bytes;
    size_t o = 0;
    const size_t pw = (size_t)PB * W, ps = (size_t)PB * nseg, nc = deg + 1;
    scal = o; o += (size_t)PB * 8;
    wx = o; o += pw; wy = o; o += pw; tx = o; o += pw; ty = o; o += pw;
    cx = o; o += ps * nc; cy = o; o += ps * nc;
    uni = o;
    o += deg == 5 ? 25 * pw + 13 * ps + (size_t)kPart * kPartStride : 0;  // optimiser state
    total_reals = o;
  }
  __host__ __device__ size_t bytes(int real_size) const { return (reals_off + total_reals * real_size + 15) / 16 * 16; }
};

template <class R, int DEG>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 4) path_warp_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NC = DEG + 1;
  constexpr unsigned kFullMask = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int W = g.W, nseg = W - 1, PB = g.PB, P = g.P;
  // ---- CTA-wide, once: basis values of B', B'' at the 32 cost nodes (quintic only)
  R* nb = (R*)smem_raw;  // [32][13]: c1[6], c2[6], wq
  size_t nb_bytes = 0;
  if constexpr (DEG == 5) {
    nb_bytes = (sizeof(R) * 32 * kPart + 15) / 16 * 16;
    if (threadIdx.x < 32) lau_node_basis<R>(nb + threadIdx.x * kPart, threadIdx.x);
    __syncthreads();  // the only CTA barrier; every thread reaches it
  }
  const WarpLayout lay(W, PB, DEG);
  unsigned char* slice = smem_raw + nb_bytes + (size_t)warp * lay.bytes((int)sizeof(R));
  const int team = blockIdx.x * kWarpsPerCta + warp;
  const int p0 = team * PB;
  const int PBa = min(PB, P - p0);
  if (PBa <= 0) return;
  int* bad = (int*)slice;
  int* improve = bad + 8;
  R* rb = (R*)(slice + lay.reals_off);
  R* scal = rb + lay.scal;
  R* wx = rb + lay.wx; R* wy = rb + lay.wy; R* tx = rb + lay.tx; R* ty = rb + lay.ty;
  R* cx = rb + lay.cx; R* cy = rb + lay.cy;
  R* uni = rb + lay.uni;
  const int nW = PBa * W, nS = PBa * nseg;

  // ---- P0: waypoints in (the warp's slice of [P][W] is contiguous), validation
  if (lane < 8) { bad[lane] = 0; improve[lane] = 0; }
  __syncwarp();
  for (int it = lane; it < nW; it += 32) {
    const R x = ((const R*)g.wx)[(size_t)p0 * W + it], y = ((const R*)g.wy)[(size_t)p0 * W + it];
    wx[it] = x; wy[it] = y;
    if (!isfinite(x) || !isfinite(y)) bad[it / W] = 1;
  }
  __syncwarp();
  for (int it = lane; it < nS; it += 32) {
    const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
    if (!(dist2<R>(wx[i], wy[i], wx[i + 1], wy[i + 1]) > 0)) bad[pl] = 1;
  }
  // ---- P1: tangents (sc:343-377, 606-638)
  for (int it = lane; it < nW; it += 32) {
    const int pl = it / W, i = it - pl * W;
    R ox, oy;
    waypoint_tangent<R>(wx + pl * W, wy + pl * W, W, i, g.start_angle ? ((const R*)g.start_angle)[p0 + pl] : Num<R>::nan(), ox, oy);
    tx[it] = ox; ty[it] = oy;
  }
  __syncwarp();

  if constexpr (DEG == 3) {
    // ---- A2: cubic control points [W_i, W_i + T_i, W_{i+1} - T_{i+1}, W_{i+1}]  (sc:661-681)
    for (int it = lane; it < nS; it += 32) {
      const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
      R* ox = cx + 4 * it; R* oy = cy + 4 * it;
      ox[0] = wx[i]; oy[0] = wy[i];
      ox[1] = wx[i] + tx[i]; oy[1] = wy[i] + ty[i];
      ox[2] = wx[i + 1] - tx[i + 1]; oy[2] = wy[i + 1] - ty[i + 1];
      ox[3] = wx[i + 1]; oy[3] = wy[i + 1];
    }
    if (lane < PBa) scal[lane * 8 + 1] = 0;
  } else {
    // ================= Lau09 quintic builder (A5), cost + gradient (A6), iRPROP- (A7) =================
    // definitions: DESIGN.md "New stages"; tx/ty hold the frozen base tangents Bt
    Carve<R> sm(uni);
    const size_t pw = (size_t)PB * W, ps = (size_t)PB * nseg;
    R* w0x = sm.take(pw); R* w0y = sm.take(pw); R* mag = sm.take(pw); R* alpha = sm.take(pw);
    R* e = sm.take(pw); R* se = sm.take(pw); R* pe = sm.take(pw);
    R* swx = sm.take(pw); R* swy = sm.take(pw); R* pwx = sm.take(pw); R* pwy = sm.take(pw);
    R* be = sm.take(pw); R* bwx = sm.take(pw); R* bwy = sm.take(pw);
    R* Tx = sm.take(pw); R* Ty = sm.take(pw); R* Ax = sm.take(pw); R* Ay = sm.take(pw);
    R* gTx = sm.take(pw); R* gTy = sm.take(pw); R* gAx = sm.take(pw); R* gAy = sm.take(pw);
    R* gWx = sm.take(pw); R* gWy = sm.take(pw); R* ge = sm.take(pw);
    R* gx = sm.take(ps * 6); R* gy = sm.take(ps * 6); R* Js = sm.take(ps);
    R* part = sm.take((size_t)kPart * kPartStride);
    const scb_opt& o = g.opt;
    for (int it = lane; it < nW; it += 32) {
      const int pl = it / W, i = it - pl * W;
      w0x[it] = wx[it]; w0y[it] = wy[it];
      mag[it] = sqrt(tx[it] * tx[it] + ty[it] * ty[it]);
      R al = 0;
      if (i > 0 && i < W - 1) {
        const R dp = dist2<R>(wx[it - 1], wy[it - 1], wx[it], wy[it]), dn = dist2<R>(wx[it], wy[it], wx[it + 1], wy[it + 1]);
        al = dn / (dp + dn);
      }
      alpha[it] = al;
      e[it] = (R)o.e_init; se[it] = (R)o.step_e; pe[it] = 0;
      swx[it] = swy[it] = (R)o.step_w * mag[it]; pwx[it] = pwy[it] = 0;
    }
    __syncwarp();
    const R lambda = (R)o.lambda;
    auto build = [&]() {
      for (int it = lane; it < nW; it += 32) { Tx[it] = e[it] * tx[it]; Ty[it] = e[it] * ty[it]; }
      __syncwarp();
      for (int it = lane; it < nW; it += 32) {
        const int pl = it / W, i = it - pl * W;
        R px = 0, py = 0, nx = 0, ny = 0;
        if (i > 0) {
          px = (R)6 * (wx[it - 1] - wx[it]) + (R)4 * Tx[it] + (R)2 * Tx[it - 1];
          py = (R)6 * (wy[it - 1] - wy[it]) + (R)4 * Ty[it] + (R)2 * Ty[it - 1];
        }
        if (i + 1 < W) {
          nx = (R)6 * (wx[it + 1] - wx[it]) - (R)4 * Tx[it] - (R)2 * Tx[it + 1];
          ny = (R)6 * (wy[it + 1] - wy[it]) - (R)4 * Ty[it] - (R)2 * Ty[it + 1];
        }
        if (i == 0) { Ax[it] = nx; Ay[it] = ny; }
        else if (i + 1 == W) { Ax[it] = px; Ay[it] = py; }
        else { Ax[it] = alpha[it] * px + ((R)1 - alpha[it]) * nx; Ay[it] = alpha[it] * py + ((R)1 - alpha[it]) * ny; }
      }
      __syncwarp();
      for (int it = lane; it < nS; it += 32) {
        const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
        const R fifth = (R)1 / (R)5, twentieth = (R)1 / (R)20;
        const R p0x = wx[i], p0y = wy[i], p5x = wx[i + 1], p5y = wy[i + 1];
        const R p1x = p0x + fifth * Tx[i], p1y = p0y + fifth * Ty[i];
        const R p4x = p5x - fifth * Tx[i + 1], p4y = p5y - fifth * Ty[i + 1];
        R* ox = cx + 6 * it; R* oy = cy + 6 * it;
        ox[0] = p0x; oy[0] = p0y; ox[1] = p1x; oy[1] = p1y;
        ox[2] = twentieth * Ax[i] + (R)2 * p1x - p0x; oy[2] = twentieth * Ay[i] + (R)2 * p1y - p0y;
        ox[3] = twentieth * Ax[i + 1] + (R)2 * p4x - p5x; oy[3] = twentieth * Ay[i + 1] + (R)2 * p4y - p5y;
        ox[4] = p4x; oy[4] = p4y; ox[5] = p5x; oy[5] = p5y;
      }
      __syncwarp();
    };
    const int K = o.iters;
    const int n_items = nS * kLanesPerSeg, n_rounds = (n_items + 31) / 32;
    constexpr int kSegPerRound = 32 / kLanesPerSeg;
    for (int iter = 0; iter <= K; ++iter) {
      build();
      const bool want_grad = iter < K;
      // ---- A6: cost and control-point gradient
      for (int r = 0; r < n_rounds; ++r) {
        const int it = r * 32 + lane;
        const bool live = it < n_items;
        const int sg = live ? it / kLanesPerSeg : 0, grp = it % kLanesPerSeg;
        R Px[6], Py[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) { Px[j] = cx[6 * sg + j]; Py[j] = cy[6 * sg + j]; }
        R J = 0, ggx[6], ggy[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) { ggx[j] = 0; ggy[j] = 0; }
#pragma unroll
        for (int qn = 0; qn < kNodesPerItem; ++qn) {
          const R* b = nb + (grp * kNodesPerItem + qn) * kPart;
          R x1 = 0, y1 = 0, x2 = 0, y2 = 0;
#pragma unroll
          for (int j = 0; j < 6; ++j) {
            x1 += b[j] * Px[j]; y1 += b[j] * Py[j]; x2 += b[6 + j] * Px[j]; y2 += b[6 + j] * Py[j];
          }
          // n = |B'|; every power of 1/n comes from one rsqrt (no divisions in the node loop)
          const R n2 = x1 * x1 + y1 * y1, rn = fast_rsqrt<R>(n2), n = n2 * rn, rn2 = rn * rn;
          const R c = x1 * y2 - y1 * x2;
          const R in5 = (rn2 * rn2) * rn;
          const R wq = b[12];
          J += wq * (c * c * in5 + lambda * n);
          if (want_grad) {
            const R k1 = (R)2 * c * in5, k2 = ((R)5 * c * c * in5) * rn2, k3 = lambda * rn;
            const R fx1 = wq * (k1 * y2 - k2 * x1 + k3 * x1), fy1 = wq * (-k1 * x2 - k2 * y1 + k3 * y1);
            const R fx2 = wq * (-k1 * y1), fy2 = wq * (k1 * x1);
#pragma unroll
            for (int j = 0; j < 6; ++j) { ggx[j] += fx1 * b[j] + fx2 * b[6 + j]; ggy[j] += fy1 * b[j] + fy2 * b[6 + j]; }
          }
        }
        // partial sums -> [13][32] slab; then one lane per (segment, component) adds 8 partials in lane order
        part[lane] = live ? J : (R)0;
        if (want_grad) {
#pragma unroll
          for (int j = 0; j < 6; ++j) { part[(1 + j) * kPartStride + lane] = live ? ggx[j] : (R)0; part[(7 + j) * kPartStride + lane] = live ? ggy[j] : (R)0; }
        }
        __syncwarp();
        const int n_out = kSegPerRound * (want_grad ? kPart : 1);
        for (int oi = lane; oi < n_out; oi += 32) {
          const int sl = oi % kSegPerRound, comp = oi / kSegPerRound;
          const int sgo = r * kSegPerRound + sl;
          if (sgo < nS) {
            const R* pp = part + comp * kPartStride + sl * kLanesPerSeg;
            R acc = pp[0];
#pragma unroll
            for (int k = 1; k < kLanesPerSeg; ++k) acc += pp[k];
            if (comp == 0) Js[sgo] = acc;
            else if (comp <= 6) gx[6 * sgo + comp - 1] = acc;
            else gy[6 * sgo + comp - 7] = acc;
          }
        }
        __syncwarp();
      }
      if (lane < PBa) {  // J of each path: segments in order; best-so-far bookkeeping
        R J = 0;
        for (int s = 0; s < nseg; ++s) J += Js[lane * nseg + s];
        const bool better = iter == 0 || J < scal[lane * 8 + 1];
        if (better) scal[lane * 8 + 1] = J;
        improve[lane] = better ? 1 : 0;
      }
      __syncwarp();
      // ---- chain rule to (T, A, W) per waypoint, gathering from the two adjacent segments
      for (int it = lane; it < nW; it += 32) {
        const int pl = it / W, i = it - pl * W;
        if (improve[pl]) { be[it] = e[it]; bwx[it] = wx[it]; bwy[it] = wy[it]; }
        if (iter == K) continue;
        R wxs = 0, wys = 0, txs = 0, tys = 0, axs = 0, ays = 0;
        if (i > 0) {
          const R* sx = gx + 6 * (pl * nseg + i - 1); const R* sy = gy + 6 * (pl * nseg + i - 1);
          wxs += (sx[3] + sx[4] + sx[5]); wys += (sy[3] + sy[4] + sy[5]);
          txs -= ((R)0.4 * sx[3] + (R)0.2 * sx[4]); tys -= ((R)0.4 * sy[3] + (R)0.2 * sy[4]);
          axs += (R)0.05 * sx[3]; ays += (R)0.05 * sy[3];
        }
        if (i < nseg) {
          const R* sx = gx + 6 * (pl * nseg + i); const R* sy = gy + 6 * (pl * nseg + i);
          wxs += (sx[0] + sx[1] + sx[2]); wys += (sy[0] + sy[1] + sy[2]);
          txs += ((R)0.2 * sx[1] + (R)0.4 * sx[2]); tys += ((R)0.2 * sy[1] + (R)0.4 * sy[2]);
          axs += (R)0.05 * sx[2]; ays += (R)0.05 * sy[2];
        }
        gWx[it] = wxs; gWy[it] = wys; gTx[it] = txs; gTy[it] = tys; gAx[it] = axs; gAy[it] = ays;
      }
      if (iter == K) break;
      __syncwarp();
      for (int it = lane; it < nW; it += 32) {
        const int pl = it / W, i = it - pl * W;
        auto ap = [&](int k) { return k == 0 ? (R)0 : (k == W - 1 ? (R)1 : alpha[pl * W + k]); };
        auto an = [&](int k) { return k == W - 1 ? (R)0 : (k == 0 ? (R)1 : (R)1 - alpha[pl * W + k]); };
        R wxs = gWx[it], wys = gWy[it], txs = gTx[it], tys = gTy[it];
        if (i > 0) {  // waypoint i is the "next" of i-1
          const R k6 = (R)6 * an(i - 1), k2 = (R)2 * an(i - 1);
          wxs += k6 * gAx[it - 1]; wys += k6 * gAy[it - 1]; txs -= k2 * gAx[it - 1]; tys -= k2 * gAy[it - 1];
        }
        {
          const R k6 = (R)6 * (ap(i) + an(i)), k4 = (R)4 * (ap(i) - an(i));
          wxs -= k6 * gAx[it]; wys -= k6 * gAy[it]; txs += k4 * gAx[it]; tys += k4 * gAy[it];
        }
        if (i + 1 < W) {  // waypoint i is the "prev" of i+1
          const R k6 = (R)6 * ap(i + 1), k2 = (R)2 * ap(i + 1);
          wxs += k6 * gAx[it + 1]; wys += k6 * gAy[it + 1]; txs += k2 * gAx[it + 1]; tys += k2 * gAy[it + 1];
        }
        ge[it] = txs * tx[it] + tys * ty[it];
        Tx[it] = wxs; Ty[it] = wys;  // final dJ/dW; Tx/Ty are free until the next build()
      }
      __syncwarp();
      if (lane < PBa) {
        R gemax = 0, gwmax = 0;
        for (int i = 0; i < W; ++i) {
          gemax = fmax(gemax, fabs(ge[lane * W + i]));
          if (i > 0 && i + 1 < W) gwmax = fmax(gwmax, fmax(fabs(Tx[lane * W + i]), fabs(Ty[lane * W + i])));
        }
        scal[lane * 8 + 2] = gemax; scal[lane * 8 + 3] = gwmax;
      }
      __syncwarp();
      // ---- A7: iRPROP-
      for (int it = lane; it < nW; it += 32) {
        const int pl = it / W, i = it - pl * W;
        const R te = (R)1e-9 * scal[pl * 8 + 2], tw = (R)1e-9 * scal[pl * 8 + 3];
        const R gi = fabs(ge[it]) <= te ? (R)0 : ge[it];
        const R ei = e[it] + rprop_step<R>(gi, pe[it], se[it], (R)1e-6, (R)1.0);
        e[it] = fmin((R)o.e_max, fmax((R)o.e_min, ei));
        if (o.move_waypoints && i > 0 && i + 1 < W) {
          const R gvx = fabs(Tx[it]) <= tw ? (R)0 : Tx[it], gvy = fabs(Ty[it]) <= tw ? (R)0 : Ty[it];
          const R smin = (R)1e-6 * mag[it], smax = (R)0.5 * mag[it], rr = (R)o.w_radius;
          const R nx = wx[it] + rprop_step<R>(gvx, pwx[it], swx[it], smin, smax);
          const R ny = wy[it] + rprop_step<R>(gvy, pwy[it], swy[it], smin, smax);
          wx[it] = fmin(w0x[it] + rr, fmax(w0x[it] - rr, nx));
          wy[it] = fmin(w0y[it] + rr, fmax(w0y[it] - rr, ny));
        }
      }
      __syncwarp();
    }
    __syncwarp();
    for (int it = lane; it < nW; it += 32) { e[it] = be[it]; wx[it] = bwx[it]; wy[it] = bwy[it]; }
    __syncwarp();
    build();
  }
  __syncwarp();

  // ---- control points, cost and the path verdict so far -> global; arc_kernel.cu takes over from here
  const int nc = nseg * NC;
  for (int it = lane; it < PBa * nc; it += 32) {
    const bool b = bad[it / nc] != 0;
    ((R*)g.ctrl_x)[(size_t)p0 * nc + it] = b ? (R)0 : cx[it];
    ((R*)g.ctrl_y)[(size_t)p0 * nc + it] = b ? (R)0 : cy[it];
  }
  if (lane < PBa) {
    const int p = p0 + lane;
    if (g.cost) ((R*)g.cost)[p] = bad[lane] ? (R)0 : scal[lane * 8 + 1];
    g.status[p] = bad[lane] ? SCB_ST_ERR_BAD_PATH : SCB_ST_OK;
  }
}

// Launch geometry of the warp-team kernel: paths per warp so that the short per-waypoint phases fill
// the lanes, and a slice small enough for 16 resident warps per SM.  False = use the CTA-team kernel.
bool path_warp_shape(int W, int spline_mode, int real_size, int& PB, size_t& smem) {
  const int deg = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  const size_t nb = deg == 5 ? ((size_t)real_size * 32 * kPart + 15) / 16 * 16 : 0;
  const size_t budget = 56 * 1024;  // per CTA of 4 warps: 4 CTAs = 16 warps per SM
  int best = 0; double best_eff = 0;
  for (int pb = 1; pb <= 8; ++pb) {
    const size_t need = nb + kWarpsPerCta * WarpLayout(W, pb, deg).bytes(real_size);
    if (need > budget) break;
    const int nw = pb * W, ns = pb * (W - 1);
    const double e_small = (double)nw / (32.0 * ((nw + 31) / 32));
    const double e_cost = (double)(ns * kLanesPerSeg) / (32.0 * ((ns * kLanesPerSeg + 31) / 32));
    const double eff = (deg == 5 ? 0.8 * e_cost + 0.2 * e_small : e_small);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = pb; }
  }
  if (!best) return false;
  PB = best;
  smem = nb + kWarpsPerCta * WarpLayout(W, best, deg).bytes(real_size);
  return true;
}

template <class R> void launch_path_warp(const PathArgs& a0, int spline_mode, int PB, size_t smem, cudaStream_t st) {
  PathArgs a = a0;
  a.PB = PB; a.table_in_smem = 0;
  const int per_cta = PB * kWarpsPerCta, grid = (a.P + per_cta - 1) / per_cta;
  if (spline_mode == SCB_REF_CUBIC) {
    cudaFuncSetAttribute(path_warp_kernel<R, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(path_warp_kernel<R, 3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    path_warp_kernel<R, 3><<<grid, kWarpsPerCta * 32, smem, st>>>(a);
  } else {
    cudaFuncSetAttribute(path_warp_kernel<R, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(path_warp_kernel<R, 5>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    path_warp_kernel<R, 5><<<grid, kWarpsPerCta * 32, smem, st>>>(a);
  }
}
template void launch_path_warp<double>(const PathArgs&, int, int, size_t, cudaStream_t);
template void launch_path_warp<float>(const PathArgs&, int, int, size_t, cudaStream_t);

}  // namespace scb
