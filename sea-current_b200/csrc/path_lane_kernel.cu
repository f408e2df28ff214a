// Path stage, lane-per-waypoint variant (sm_100a): the default for paths of up to 32 waypoints.
//   tangent heuristics sc:343-377, segment builder sc:599-683 (cubic), Lau09 quintic builder + cost /
//   analytic gradient + iRPROP- (DESIGN.md "New stages" is the frozen definition).
//
// Mapping.  A warp owns PB = floor(32 / W) consecutive paths.  Lane l owns waypoint i = l % W of path
// l / W for the whole kernel AND segment i (waypoints i, i+1) when i < W-1.  Everything a waypoint needs
// from its neighbours travels by __shfl_up/down(1): there is no shared-memory hand-over, no __syncwarp
// and no lane re-mapping between the phases of an optimiser iteration (the predecessor,
// path_warp_kernel.cu, spent half of its issue slots and stall samples in those phases:
// profiles/r1f_path_warp_kernel_f64.lines.txt).  Shared memory only holds per-lane private optimiser
// state (a register spill area the owner alone touches).
//
// Cost + gradient of a segment run on ONE lane, all 32 Gauss-Legendre nodes in sequence, so node
// constants are warp-uniform constant-bank operands (LAU_PAIR) and partial sums never leave
// registers.  The hodograph is evaluated in the power basis,
//     B'(u) = sum_m c_m u^m,  c_m = C(4,m) Delta^m D_0,  D_j = 5 (P_{j+1} - P_j),
// by Horner (7 FMAs per axis for B', B'' instead of 12 in the Bernstein form), and the gradient is
// accumulated against the powers of u (G_m = sum f1 u^m, H_m = sum f2 u^m: 9 per axis instead of 12)
// and mapped back to (T, A, W) once per iteration.  D_j are formed from (W, T, A) directly,
//     D0 = T_i, D1 = T_i + A_i/4, D2 = 5 (W_{i+1} - W_i) - 2 (T_i + T_{i+1}) + (A_{i+1} - A_i)/4,
//     D3 = T_{i+1} - A_{i+1}/4, D4 = T_{i+1},
// which is the frozen control-point construction with the O(|W|) terms cancelled analytically.
// Nodes are taken in symmetric pairs about u = 1/2 (segment_cost below): ~53 FP64 instructions per node
// (the Bernstein-form predecessor needed ~95).
//
// Arithmetic is FP64 for both storage types S: iRPROP- acts on gradient SIGNS, so its iterates are
// only reproducible where the gradient is resolved far below the decision threshold (1e-9 of the
// largest component) -- like toppra's feasibility tolerances this does not exist in FP32
// (DESIGN.md "Precision modes").  S = float only changes what is loaded and stored.
#include <cstdlib>

#include "lau_nodes_device.h"
#include "path_device.cuh"

namespace scb {

namespace {

constexpr unsigned kAll = 0xffffffffu;
constexpr int kLaneWarps = 4;     // warps per CTA
constexpr int kLaneSlots = 15;    // private FP64 slots per lane in shared memory:
enum { kSe, kPe, kSwx, kSwy, kPwx, kPwy, kW0x, kW0y, kMag, kBe, kBwx, kBwy, kAlpha, kAnPrev, kApNext,
       // SHADOW mapping (path_pair_kernel): state of the path's LAST waypoint, carried by the lane of its last segment
       kSeS, kPeS, kBeS, kBtSx, kBtSy, kSx, kSy, kAnPrevS };
constexpr int kPairSlots = 23;


// What one cost node contributes, given B' = (x1, y1) and B'' = (x2, y2) there: the weighted integrand (returned)
// and, if GRAD, the weighted forces f1 = dJ/dB', f2 = dJ/dB''.
//   d/dB' = k1 (y2, -x2) + (k3 - k2) B',  d/dB'' = k1 (-y1, x1);  k1 = 2c/n^5, k2 = 5c^2/n^7, k3 = lambda/n
template <bool GRAD>
__device__ __forceinline__ double node_terms(double x1, double y1, double x2, double y2, double lambda, double wq, double w2,
                                             double& fx1, double& fy1, double& fx2, double& fy2) {
  // n = |B'|; every power of 1/n comes from one rsqrt
  const double n2 = fma(x1, x1, y1 * y1), rn = fast_rsqrt<double>(n2), n = n2 * rn, rn2 = rn * rn;
  const double c = fma(x1, y2, -(y1 * x2));
  const double in5 = (rn2 * rn2) * rn;
  const double t = c * in5, cc5 = t * c;  // c / n^5, c^2 / n^5
  if (GRAD) {
    const double wk1 = w2 * t;
    const double wk32 = wq * fma(-5.0, cc5 * rn2, lambda * rn);
    fx1 = fma(wk1, y2, wk32 * x1); fy1 = fma(-wk1, x2, wk32 * y1);
    fx2 = -(wk1 * y1); fy2 = wk1 * x1;
  }
  return fma(lambda, n, cc5);
}

// Cost of one quintic segment and, if GRAD, its gradient with respect to the power-basis coefficients of
// the hodograph (gc?[m] = dJ/dc_m).  cx, cy: c_0..c_4.  The 32 nodes are 16 pairs u = 1/2 -+ h with equal
// weights: with B'(1/2 + v) = sum_m e_m v^m the two values of a pair are E(h^2) -+ h O(h^2) (even and odd
// parts: 11 instructions per axis for B' and B'' at BOTH nodes instead of 14), and the moments
// G_m = sum f1 v^m, H_m = sum f2 v^m take the pair's sum (even m) or difference (odd m): 13 instead of 18.
template <bool GRAD>
__device__ __forceinline__ double segment_cost(const double (&cx)[5], const double (&cy)[5], double lambda,
                                               double (&gcx)[5], double (&gcy)[5]) {
  // Taylor shift to u = 1/2: e_m = sum_k c_k C(k,m) 2^(m-k)
  const double ex0 = cx[0] + 0.5 * cx[1] + 0.25 * cx[2] + 0.125 * cx[3] + 0.0625 * cx[4];
  const double ex1 = cx[1] + cx[2] + 0.75 * cx[3] + 0.5 * cx[4];
  const double ex2 = cx[2] + 1.5 * cx[3] + 1.5 * cx[4], ex3 = cx[3] + 2.0 * cx[4], ex4 = cx[4];
  const double ey0 = cy[0] + 0.5 * cy[1] + 0.25 * cy[2] + 0.125 * cy[3] + 0.0625 * cy[4];
  const double ey1 = cy[1] + cy[2] + 0.75 * cy[3] + 0.5 * cy[4];
  const double ey2 = cy[2] + 1.5 * cy[3] + 1.5 * cy[4], ey3 = cy[3] + 2.0 * cy[4], ey4 = cy[4];
  // B''(1/2 + v) = e1 + 2 e2 v + 3 e3 v^2 + 4 e4 v^3
  const double tx2 = 2.0 * ex2, tx3 = 3.0 * ex3, tx4 = 4.0 * ex4, ty2 = 2.0 * ey2, ty3 = 3.0 * ey3, ty4 = 4.0 * ey4;
  double J = 0;
  double Gx[5] = {0, 0, 0, 0, 0}, Gy[5] = {0, 0, 0, 0, 0}, Hx[4] = {0, 0, 0, 0}, Hy[4] = {0, 0, 0, 0};
#pragma unroll 2
  for (int q = 0; q < 16; ++q) {
    const double h = LAU_PAIR[q][0], w = LAU_PAIR[q][1], ww = LAU_PAIR[q][2], h3 = LAU_PAIR[q][3];
    const double wq = LAU_PAIR[q][4], w2 = LAU_PAIR[q][5];
    const double Ex = fma(fma(ex4, w, ex2), w, ex0), Ox = h * fma(ex3, w, ex1);
    const double Ey = fma(fma(ey4, w, ey2), w, ey0), Oy = h * fma(ey3, w, ey1);
    const double Ax = fma(tx3, w, ex1), Bx = h * fma(tx4, w, tx2);
    const double Ay = fma(ty3, w, ey1), By = h * fma(ty4, w, ty2);
    double fx1p, fy1p, fx2p, fy2p, fx1m, fy1m, fx2m, fy2m;
    const double jm = node_terms<GRAD>(Ex - Ox, Ey - Oy, Ax - Bx, Ay - By, lambda, wq, w2, fx1m, fy1m, fx2m, fy2m);
    const double jp = node_terms<GRAD>(Ex + Ox, Ey + Oy, Ax + Bx, Ay + By, lambda, wq, w2, fx1p, fy1p, fx2p, fy2p);
    J = fma(wq, jm + jp, J);
    if (GRAD) {
      const double sx1 = fx1p + fx1m, dx1 = fx1p - fx1m, sy1 = fy1p + fy1m, dy1 = fy1p - fy1m;
      const double sx2 = fx2p + fx2m, dx2 = fx2p - fx2m, sy2 = fy2p + fy2m, dy2 = fy2p - fy2m;
      Gx[0] += sx1; Gx[2] = fma(sx1, w, Gx[2]); Gx[4] = fma(sx1, ww, Gx[4]); Gx[1] = fma(dx1, h, Gx[1]); Gx[3] = fma(dx1, h3, Gx[3]);
      Gy[0] += sy1; Gy[2] = fma(sy1, w, Gy[2]); Gy[4] = fma(sy1, ww, Gy[4]); Gy[1] = fma(dy1, h, Gy[1]); Gy[3] = fma(dy1, h3, Gy[3]);
      Hx[0] += sx2; Hx[2] = fma(sx2, w, Hx[2]); Hx[1] = fma(dx2, h, Hx[1]); Hx[3] = fma(dx2, h3, Hx[3]);
      Hy[0] += sy2; Hy[2] = fma(sy2, w, Hy[2]); Hy[1] = fma(dy2, h, Hy[1]); Hy[3] = fma(dy2, h3, Hy[3]);
    }
  }
  if (GRAD) {
    // gradient with respect to e_m (B'' = sum_m m e_m v^(m-1)), then the transposed Taylor shift
    double gex[5], gey[5];
    gex[0] = Gx[0]; gey[0] = Gy[0];
#pragma unroll
    for (int m = 1; m < 5; ++m) { gex[m] = fma((double)m, Hx[m - 1], Gx[m]); gey[m] = fma((double)m, Hy[m - 1], Gy[m]); }
    gcx[0] = gex[0]; gcy[0] = gey[0];
    gcx[1] = 0.5 * gex[0] + gex[1]; gcy[1] = 0.5 * gey[0] + gey[1];
    gcx[2] = 0.25 * gex[0] + gex[1] + gex[2]; gcy[2] = 0.25 * gey[0] + gey[1] + gey[2];
    gcx[3] = 0.125 * gex[0] + 0.75 * gex[1] + 1.5 * gex[2] + gex[3]; gcy[3] = 0.125 * gey[0] + 0.75 * gey[1] + 1.5 * gey[2] + gey[3];
    gcx[4] = 0.0625 * gex[0] + 0.5 * gex[1] + 1.5 * gex[2] + 2.0 * gex[3] + gex[4];
    gcy[4] = 0.0625 * gey[0] + 0.5 * gey[1] + 1.5 * gey[2] + 2.0 * gey[3] + gey[4];
  }
  return J;
}

// c_m = C(4,m) Delta^m D_0
__device__ __forceinline__ void hodograph_power(double d0, double d1, double d2, double d3, double d4, double (&c)[5]) {
  const double a0 = d1 - d0, a1 = d2 - d1, a2 = d3 - d2, a3 = d4 - d3;
  const double b0 = a1 - a0, b1 = a2 - a1, b2 = a3 - a2;
  const double e0 = b1 - b0, e1 = b2 - b1;
  c[0] = d0; c[1] = 4.0 * a0; c[2] = 6.0 * b0; c[3] = 4.0 * e0; c[4] = e1 - e0;
}
// transpose of hodograph_power: gradient with respect to D_0..D_4
__device__ __forceinline__ void hodograph_power_T(const double (&g)[5], double (&gd)[5]) {
  gd[0] = g[0] - 4.0 * g[1] + 6.0 * g[2] - 4.0 * g[3] + g[4];
  gd[1] = 4.0 * g[1] - 12.0 * g[2] + 12.0 * g[3] - 4.0 * g[4];
  gd[2] = 6.0 * g[2] - 12.0 * g[3] + 6.0 * g[4];
  gd[3] = 4.0 * g[3] - 4.0 * g[4];
  gd[4] = g[4];
}

// ---- collision term inside the optimiser loop (Lau09 IV-B; DESIGN.md "New stages", A7) ------------------------------
// Does the lane's segment, cut into M chords B(k/M) -> B((k+1)/M), cross an obstacle edge (the reference's segment
// test, sc:142-178)?  The segment is given by its start point and hodograph control points D_0..D_4 (control points
// P_0 = W, P_{j+1} = P_j + D_j / 5).  Two culls keep the common case cheap: the float bounding box of the control
// polygon (the curve lies in its convex hull) against the float bounding box of every edge - rounded outwards, so a
// cull never hides a crossing - and, for the few edges that survive, the chord's box against the edge's.  Edge data
// of the first kCollSmem edges sits in shared memory (one broadcast read per edge), the rest comes from global memory.
constexpr int kCollSmem = 128;
struct CollEdges {
  const double* e;   // [n][4] in shared memory (first kCollSmem edges)
  const float* box;  // [n][4] = (xlo, xhi, ylo, yhi), rounded outwards
  const float* ef;   // [n][4] = the end points as floats (oriented-box cull)
  int ns, n;         // staged / all
  const void* g;     // all edges in global memory, storage type S
};

// The staged edge data is read through explicit shared-memory loads: the pointers reach segment_hits through a struct
// and a call, where the compiler no longer knows their address space and emits generic loads (long-scoreboard stalls on
// the cull loop's first line were 6 % of the kernel's samples, profiles/r2n per-line profile).
__device__ __forceinline__ float4 lds_f4(const float* base, int idx) {
  float4 v;
  const unsigned a = (unsigned)__cvta_generic_to_shared(base) + 16u * (unsigned)idx;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void lds_edge(const double* base, int idx, double& x0, double& y0, double& x1, double& y1) {
  const unsigned a = (unsigned)__cvta_generic_to_shared(base) + 32u * (unsigned)idx;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x0), "=d"(y0) : "r"(a));
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x1), "=d"(y1) : "r"(a + 16u));
}

template <class S>
__device__ __forceinline__ void stage_coll_edges(const PathArgs& g, double* se, float* sb, float* sf, int tid, int T) {
  const int ns = g.n_coll_edges < kCollSmem ? g.n_coll_edges : kCollSmem;
  const S* ge = (const S*)g.coll_edges;
  for (int k = tid; k < ns; k += T) {
    const double x0 = (double)ge[4 * k], y0 = (double)ge[4 * k + 1], x1 = (double)ge[4 * k + 2], y1 = (double)ge[4 * k + 3];
    se[4 * k] = x0; se[4 * k + 1] = y0; se[4 * k + 2] = x1; se[4 * k + 3] = y1;
    sb[4 * k] = __double2float_rd(fmin(x0, x1)); sb[4 * k + 1] = __double2float_ru(fmax(x0, x1));
    sb[4 * k + 2] = __double2float_rd(fmin(y0, y1)); sb[4 * k + 3] = __double2float_ru(fmax(y0, y1));
    sf[4 * k] = (float)x0; sf[4 * k + 1] = (float)y0; sf[4 * k + 2] = (float)x1; sf[4 * k + 3] = (float)y1;
  }
}

// NW: 32-edge words of the candidate mask (1 when the staged list has at most 32 edges: the common case walks one word)
template <class S, int NW>
__device__ __noinline__ bool segment_hits_nw(double wx, double wy, const double (&dx)[5], const double (&dy)[5], const CollEdges& ce, int M) {
  // control polygon -> float box of the hull, and the hull's box in the frame of the segment's chord W_i -> W_i+1:
  // s = n . (P - W) across the chord (n = the chord turned by 90 degrees, not normalised), a = d . (P - W) along it.
  // A wall beside a diagonal leg lies inside the axis-aligned box but outside the oriented one.
  float xlo, xhi, ylo, yhi, slo, shi, alo, ahi;
  float fwx, fwy, fdx, fdy;
  {
    double rx[5], ry[5];
    double px = 0, py = 0, ax = 0, bx = 0, ay = 0, by = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      px = fma(0.2, dx[j], px); py = fma(0.2, dy[j], py);
      rx[j] = px; ry[j] = py;
      ax = fmin(ax, px); bx = fmax(bx, px); ay = fmin(ay, py); by = fmax(by, py);
    }
    // the chord points below come from the power basis: a hair outside the hull by rounding; widen by 1e-6 of its size
    const float mx = 1e-6f * (float)(bx - ax) + 1e-30f, my = 1e-6f * (float)(by - ay) + 1e-30f;
    xlo = __double2float_rd(wx + ax) - mx; xhi = __double2float_ru(wx + bx) + mx;
    ylo = __double2float_rd(wy + ay) - my; yhi = __double2float_ru(wy + by) + my;
    const double cdx = rx[4], cdy = ry[4];  // the chord
    double s0 = 0, s1 = 0, a0 = 0, a1 = 0;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const double sj = fma(cdx, ry[j], -(cdy * rx[j])), aj = fma(cdx, rx[j], cdy * ry[j]);
      s0 = fmin(s0, sj); s1 = fmax(s1, sj); a0 = fmin(a0, aj); a1 = fmax(a1, aj);
    }
    fwx = (float)wx; fwy = (float)wy; fdx = (float)cdx; fdy = (float)cdy;
    // float evaluation of the edge side: coordinates rounded to float (rel. 6e-8 each), two products summed; the margin
    // is 16 float ulps of the largest term that can occur, |d| (|W| + |edge|) - far more than the rounding, far less
    // than a millimetre for paths of a few hundred metres
    const float big = (fabsf(fdx) + fabsf(fdy)) * (fabsf(fwx) + fabsf(fwy) + (float)(bx - ax) + (float)(by - ay) + 1.0f);
    const float mg = 1e-6f * big + 1e-30f;
    slo = __double2float_rd(s0) - mg; shi = __double2float_ru(s1) + mg; alo = __double2float_rd(a0) - mg; ahi = __double2float_ru(a1) + mg;
  }
  // candidate edges: those whose box meets the hull's (bit mask over the staged edges)
  unsigned cand[NW];
  unsigned any = 0;
#pragma unroll
  for (int w = 0; w < NW; ++w) {
    unsigned m = 0;
    const int e0 = 32 * w, e1 = ce.ns < e0 + 32 ? ce.ns : e0 + 32;
    for (int e = e0; e < e1; ++e) {
      const float4 b = lds_f4(ce.box, e);
      if (b.x > xhi || b.y < xlo || b.z > yhi || b.w < ylo) continue;
      const float4 q = lds_f4(ce.ef, e);
      const float r0x = q.x - fwx, r0y = q.y - fwy, r1x = q.z - fwx, r1y = q.w - fwy;
      const float mgs = 1e-6f * (fabsf(fdx) + fabsf(fdy)) * (fabsf(q.x) + fabsf(q.y) + fabsf(q.z) + fabsf(q.w));  // the edge's own size
      const float n0 = fdx * r0y - fdy * r0x, n1 = fdx * r1y - fdy * r1x;
      if (fmaxf(n0, n1) < slo - mgs || fminf(n0, n1) > shi + mgs) continue;  // the whole edge on one side of the hull
      const float t0 = fdx * r0x + fdy * r0y, t1 = fdx * r1x + fdy * r1y;
      if (fmaxf(t0, t1) < alo - mgs || fminf(t0, t1) > ahi + mgs) continue;  // ... or before / beyond it
      m |= 1u << (e - e0);
    }
    cand[w] = m; any |= m;
  }
  const bool tail = ce.n > ce.ns;  // an edge list longer than the stage: the rest is walked from global memory
  if (!any && !tail) return false;
  // power basis of the segment: B(u) = W + u (c0 + u (c1/2 + u (c2/3 + u (c3/4 + u c4/5))))
  double cx[5], cy[5];
  hodograph_power(dx[0], dx[1], dx[2], dx[3], dx[4], cx);
  hodograph_power(dy[0], dy[1], dy[2], dy[3], dy[4], cy);
  const double qx1 = 0.5 * cx[1], qx2 = cx[2] * (1.0 / 3.0), qx3 = 0.25 * cx[3], qx4 = 0.2 * cx[4];
  const double qy1 = 0.5 * cy[1], qy2 = cy[2] * (1.0 / 3.0), qy3 = 0.25 * cy[3], qy4 = 0.2 * cy[4];
  const double invM = 1.0 / (double)M;
  const S* ge = (const S*)ce.g;
  double ax = wx, ay = wy;
  float fax = (float)wx, fay = (float)wy;
  for (int k = 1; k <= M; ++k) {  // every chord point is formed once; the candidates are tested chord by chord
    const double u = k == M ? 1.0 : (double)k * invM;
    const double qx = fma(u, fma(u, fma(u, fma(u, fma(u, qx4, qx3), qx2), qx1), cx[0]), wx);
    const double qy = fma(u, fma(u, fma(u, fma(u, fma(u, qy4, qy3), qy2), qy1), cy[0]), wy);
    const float fqx = (float)qx, fqy = (float)qy;
    // float box of the chord, widened by two float ulps of its coordinates (the conversions round to nearest)
    const float gx = 2.4e-7f * fmaxf(fabsf(fax), fabsf(fqx)) + 1e-30f, gy = 2.4e-7f * fmaxf(fabsf(fay), fabsf(fqy)) + 1e-30f;
    const float cxl = fminf(fax, fqx) - gx, cxh = fmaxf(fax, fqx) + gx, cyl = fminf(fay, fqy) - gy, cyh = fmaxf(fay, fqy) + gy;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      unsigned m = cand[w];
      while (m) {
        const int e = 32 * w + __ffs(m) - 1;
        m &= m - 1;
        const float4 b = lds_f4(ce.box, e);
        if (b.x > cxh || b.y < cxl || b.z > cyh || b.w < cyl) continue;
        double e0x, e0y, e1x, e1y, ix, iy;
        lds_edge(ce.e, e, e0x, e0y, e1x, e1y);
        if (seg_intersects<double>(ax, ay, qx, qy, e0x, e0y, e1x, e1y, ix, iy)) return true;
      }
    }
    if (tail) {
      for (int e = ce.ns; e < ce.n; ++e) {
        const double e0x = (double)ge[4 * e], e0y = (double)ge[4 * e + 1], e1x = (double)ge[4 * e + 2], e1y = (double)ge[4 * e + 3];
        if (fmin(e0x, e1x) > fmax(ax, qx) || fmax(e0x, e1x) < fmin(ax, qx) || fmin(e0y, e1y) > fmax(ay, qy) || fmax(e0y, e1y) < fmin(ay, qy)) continue;
        double ix, iy;
        if (seg_intersects<double>(ax, ay, qx, qy, e0x, e0y, e1x, e1y, ix, iy)) return true;
      }
    }
    ax = qx; ay = qy; fax = fqx; fay = fqy;
  }
  return false;
}

}  // namespace


template <class S>
__device__ __forceinline__ bool segment_hits(double wx, double wy, const double (&dx)[5], const double (&dy)[5], const CollEdges& ce, int M) {
  return ce.ns <= 32 ? segment_hits_nw<S, 1>(wx, wy, dx, dy, ce, M) : segment_hits_nw<S, kCollSmem / 32>(wx, wy, dx, dy, ce, M);
}

// ---- how a lane talks to the other lanes of its path -------------------------------------------------
// WarpComm: the path lives inside one warp (W <= 32, several paths per warp): shuffles only.
// CtaComm:  one path per CTA (33 <= W <= 512 waypoints, C5's long-horizon paths): neighbours through a
//           double-buffered exchange area in shared memory, ONE __syncthreads per exchange (a thread can be
//           at most one exchange ahead of the slowest, and the buffer it then writes is not the one being read).
struct WarpComm {
  int base, W;        // first lane and length of this lane's path
  double* st;         // private slots: st[k * sstride]
  int sstride;
  template <int N> __device__ __forceinline__ void prev(const double (&v)[N], double (&pv)[N]) {
#pragma unroll
    for (int k = 0; k < N; ++k) pv[k] = __shfl_up_sync(kAll, v[k], 1);
  }
  template <int N> __device__ __forceinline__ void next(const double (&v)[N], double (&nx)[N]) {
#pragma unroll
    for (int k = 0; k < N; ++k) nx[k] = __shfl_down_sync(kAll, v[k], 1);
  }
  template <int N> __device__ __forceinline__ void both(const double (&v)[N], double (&pv)[N], double (&nx)[N]) { prev(v, pv); next(v, nx); }
  // sum over the path's first n lanes, in lane order (DESIGN.md "New stages", A6: segments in order)
  __device__ __forceinline__ double path_sum(double v, int n) {
    double acc = 0;
    for (int s = 0; s < n; ++s) acc += __shfl_sync(kAll, v, (base + s) & 31);
    return acc;
  }
  __device__ __forceinline__ void path_max2(double& a, double& b) {
    double ma = 0, mb = 0;
    for (int k = 0; k < W; ++k) { ma = fmax(ma, __shfl_sync(kAll, a, (base + k) & 31)); mb = fmax(mb, __shfl_sync(kAll, b, (base + k) & 31)); }
    a = ma; b = mb;
  }
  __device__ __forceinline__ bool path_any(bool f) {
    const unsigned m = __ballot_sync(kAll, f);
    const unsigned pm = (W >= 32 ? kAll : ((1u << W) - 1u)) << (base & 31);
    return (m & pm) != 0;
  }
};

struct CtaComm {
  static constexpr int kMaxVals = 6;
  int T, tid;          // threads of the CTA (a multiple of 32), this thread
  double* xbuf;        // [2][kMaxVals][T] exchange area
  double* red;         // [2][2][32] per-warp partials
  double* st;          // private slots: st[k * sstride]
  int sstride;
  int flip = 0, rflip = 0;
  template <int N> __device__ __forceinline__ double* stage(const double (&v)[N]) {
    static_assert(N <= kMaxVals, "exchange area too small");
    double* b = xbuf + (size_t)flip * kMaxVals * T;
    flip ^= 1;
#pragma unroll
    for (int k = 0; k < N; ++k) b[k * T + tid] = v[k];
    __syncthreads();
    return b;
  }
  template <int N> __device__ __forceinline__ void prev(const double (&v)[N], double (&pv)[N]) {
    const double* b = stage(v);
    const int j = max(tid - 1, 0);
#pragma unroll
    for (int k = 0; k < N; ++k) pv[k] = b[k * T + j];
  }
  template <int N> __device__ __forceinline__ void next(const double (&v)[N], double (&nx)[N]) {
    const double* b = stage(v);
    const int j = min(tid + 1, T - 1);
#pragma unroll
    for (int k = 0; k < N; ++k) nx[k] = b[k * T + j];
  }
  template <int N> __device__ __forceinline__ void both(const double (&v)[N], double (&pv)[N], double (&nx)[N]) {
    const double* b = stage(v);
    const int jm = max(tid - 1, 0), jp = min(tid + 1, T - 1);
#pragma unroll
    for (int k = 0; k < N; ++k) { pv[k] = b[k * T + jm]; nx[k] = b[k * T + jp]; }
  }
  // deterministic tree sum: xor butterfly inside each warp, warp partials added in warp order by every thread
  __device__ __forceinline__ double path_sum(double v, int n) {
    if (tid >= n) v = 0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(kAll, v, off);
    double* r = red + (size_t)rflip * 64; rflip ^= 1;
    if ((tid & 31) == 0) r[tid >> 5] = v;
    __syncthreads();
    double acc = 0;
    for (int w = 0; w < (T >> 5); ++w) acc += r[w];
    return acc;
  }
  __device__ __forceinline__ void path_max2(double& a, double& b) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) { a = fmax(a, __shfl_xor_sync(kAll, a, off)); b = fmax(b, __shfl_xor_sync(kAll, b, off)); }
    double* r = red + (size_t)rflip * 64; rflip ^= 1;
    if ((tid & 31) == 0) { r[tid >> 5] = a; r[32 + (tid >> 5)] = b; }
    __syncthreads();
    double ma = 0, mb = 0;
    for (int w = 0; w < (T >> 5); ++w) { ma = fmax(ma, r[w]); mb = fmax(mb, r[32 + w]); }
    a = ma; b = mb;
  }
  __device__ __forceinline__ bool path_any(bool f) { return __syncthreads_or(f ? 1 : 0) != 0; }
};

// PairComm: paths packed over a PAIR of warps (64 lanes, one lane per SEGMENT: path_pair_kernel), so a path may straddle
// the two warps.  Everything goes through a double-buffered exchange area of the pair and ONE named barrier of 64
// threads per exchange (the same one-exchange-ahead argument as CtaComm).  Sums run over the path's lanes in lane
// order, like WarpComm's: the bits of the cost do not depend on the mapping.
struct PairComm {
  static constexpr int kMaxVals = 6, T = 64;
  int tid, base, n, bar;  // lane within the pair, first lane and number of lanes of this lane's path, barrier id
  double* xbuf;           // [2][kMaxVals][64]
  double* st;
  int sstride;
  int flip = 0;
  template <int N> __device__ __forceinline__ const double* stage(const double (&v)[N]) {
    static_assert(N <= kMaxVals, "exchange area too small");
    double* b = xbuf + (size_t)flip * kMaxVals * T;
    flip ^= 1;
#pragma unroll
    for (int k = 0; k < N; ++k) b[k * T + tid] = v[k];
    asm volatile("bar.sync %0, %1;" ::"r"(bar), "r"(T) : "memory");
    return b;
  }
  template <int N> __device__ __forceinline__ void prev(const double (&v)[N], double (&pv)[N]) {
    const double* b = stage(v);
    const int j = max(tid - 1, 0);
#pragma unroll
    for (int k = 0; k < N; ++k) pv[k] = b[k * T + j];
  }
  template <int N> __device__ __forceinline__ void next(const double (&v)[N], double (&nx)[N]) {
    const double* b = stage(v);
    const int j = min(tid + 1, T - 1);
#pragma unroll
    for (int k = 0; k < N; ++k) nx[k] = b[k * T + j];
  }
  template <int N> __device__ __forceinline__ void both(const double (&v)[N], double (&pv)[N], double (&nx)[N]) {
    const double* b = stage(v);
    const int jm = max(tid - 1, 0), jp = min(tid + 1, T - 1);
#pragma unroll
    for (int k = 0; k < N; ++k) { pv[k] = b[k * T + jm]; nx[k] = b[k * T + jp]; }
  }
  __device__ __forceinline__ double path_sum(double v, int cnt) {
    const double a[1] = {v};
    const double* b = stage(a);
    const double* q = b + base;
    const int m = min(cnt, T - base);  // (lanes without a path may start near the end of the area)
    double acc = 0;
    for (int s = 0; s < m; ++s) acc += q[s];
    return acc;
  }
  __device__ __forceinline__ void path_max2(double& a, double& c) {
    const double v[2] = {a, c};
    const double* b = stage(v);
    const double* q = b + base;
    const int m = min(n, T - base);
    // compare-and-select instead of fmax: ma / mb start at 0 and can never become NaN this way, and a NaN operand fails
    // the comparison and is skipped - what fmax does - at half the instructions (this loop was 6 % of the kernel's)
    double ma = 0, mb = 0;
#pragma unroll 3
    for (int k = 0; k < m; ++k) { const double x = q[k], y = q[T + k]; ma = x > ma ? x : ma; mb = y > mb ? y : mb; }
    a = ma; c = mb;
  }
  __device__ __forceinline__ bool path_any(bool f) {
    const double v[1] = {f ? 1.0 : 0.0};
    const double* b = stage(v);
    double m = 0;
    for (int k = 0; k < n; ++k) m = fmax(m, b[min(base + k, T - 1)]);
    return m != 0;
  }
};

// One waypoint lane from the waypoints to the control points of its segment.  p: problem, i: waypoint index,
// own: the lane has a waypoint.  Every lane of the warp / CTA runs this to the end (collectives inside).
// COLL: the collision term inside the optimiser loop is compiled in (a separate instantiation, so that the default
// kernel stays the register-exact code it was: 127 registers, no stack traffic in the loop)
// SHADOW (path_pair_kernel): one lane per SEGMENT - waypoint i and segment i for i = 0 .. nseg-1 - and the lane of the
// last segment also carries the path's last waypoint (its elongation is a parameter of the optimiser like any other):
// what the lane of waypoint nseg would have computed and handed to its neighbour is computed here, with the same
// expressions in the same order, and substituted for the neighbour's values.  Every lane executes the shadow code
// (a handful of instructions per iteration against ~1,600 for the segment's cost); only the last-segment lane keeps it.
template <class S, int DEG, class Comm, bool COLL, bool SHADOW = false>
__device__ __forceinline__ void lane_path(const PathArgs& g, Comm& cm, int p, int i, bool own, const CollEdges& ce) {
  static_assert(!(SHADOW && (COLL || DEG != 5)), "the shadow mapping is built for the plain quintic optimiser");
  constexpr int NC = DEG + 1;
  const int W = g.W, nseg = W - 1, ss = cm.sstride;
  double* st = cm.st;
  const bool first = i == 0, last = i == nseg;
  const bool hasS = SHADOW && own && i == nseg - 1;
  const bool seg = own && !last;          // this lane also owns segment i
  const bool inner = own && !first && !last;

  // ---- waypoints in (consecutive lanes read consecutive addresses), neighbours, validation
  double wx = own ? (double)((const S*)g.wx)[(size_t)p * W + i] : (double)i;
  double wy = own ? (double)((const S*)g.wy)[(size_t)p * W + i] : 0.0;
  double xm, ym, xp, yp;
  {
    const double v[2] = {wx, wy}; double a[2], b[2];
    cm.both(v, a, b);
    xm = a[0]; ym = a[1]; xp = b[0]; yp = b[1];
  }
  double sx = 0, sy = 0;  // SHADOW: the last waypoint
  if constexpr (SHADOW) {
    if (hasS) { sx = (double)((const S*)g.wx)[(size_t)p * W + nseg]; sy = (double)((const S*)g.wy)[(size_t)p * W + nseg]; xp = sx; yp = sy; }
  }
  const double dprev = dist2<double>(xm, ym, wx, wy), dnext = dist2<double>(wx, wy, xp, yp);
  const bool bad = cm.path_any((own && (!isfinite(wx) || !isfinite(wy) || (seg && !(dnext > 0)))) ||
                               (hasS && (!isfinite(sx) || !isfinite(sy))));

  // ---- tangents (sc:343-377, 606-638)
  double btx, bty;
  if (first) {
    double ang = g.start_angle && own ? (double)((const S*)g.start_angle)[p] : CUDART_NAN;
    if (isnan(ang)) ang = atan2(yp - wy, xp - wx);
    double sn, cs; sincos(ang, &sn, &cs);
    btx = (0.5 * dnext) * cs; bty = (0.5 * dnext) * sn;
  } else if (last) {
    const double dx = wx - xm, dy = wy - ym, n = sqrt(dx * dx + dy * dy);
    btx = (0.5 * dprev) * (dx / n); bty = (0.5 * dprev) * (dy / n);
  } else {
    calc_tangent<double>(xm, ym, wx, wy, xp, yp, btx, bty);
  }
  if (g.n_edges > 0) shrink_tangent<double, S>(wx, wy, (const S*)g.edges, g.n_edges, btx, bty);  // sc:615, 620, 637
  double bsx = 0, bsy = 0;  // SHADOW: base tangent of the last waypoint (the `last` branch above, seen from its neighbour)
  if constexpr (SHADOW) {
    const double dps = dist2<double>(wx, wy, sx, sy);
    const double dx = sx - wx, dy = sy - wy, n = sqrt(dx * dx + dy * dy);
    bsx = (0.5 * dps) * (dx / n); bsy = (0.5 * dps) * (dy / n);
  }

  // T and A of this waypoint and the next one: what both the optimiser and the final control points need
  double alpha = 0;
  double eS = 1.0;  // SHADOW: elongation of the last waypoint
  auto waypoint_TA = [&](double e, double& Tx, double& Ty, double& Txp, double& Typ, double& wxp, double& wyp,
                         double& Ax, double& Ay, double& Axp, double& Ayp) {
    Tx = e * btx; Ty = e * bty;
    const double v[4] = {Tx, Ty, wx, wy}; double pv[4], nx[4];
    cm.both(v, pv, nx);
    if constexpr (SHADOW) {
      if (hasS) { nx[0] = eS * st[kBtSx * ss]; nx[1] = eS * st[kBtSy * ss]; nx[2] = st[kSx * ss]; nx[3] = st[kSy * ss]; }
    }
    Txp = nx[0]; Typ = nx[1]; wxp = nx[2]; wyp = nx[3];
    // (written out with the contraction the compiler chose for `6 (Wp - W) + 4 T + 2 Tp`, so that the SHADOW copy
    // below is the same arithmetic by construction)
    const double pvx = fma(2.0, pv[0], fma(6.0, pv[2] - wx, 4.0 * Tx)), pvy = fma(2.0, pv[1], fma(6.0, pv[3] - wy, 4.0 * Ty));
    const double nxx = 6.0 * (wxp - wx) - 4.0 * Tx - 2.0 * Txp, nxy = 6.0 * (wyp - wy) - 4.0 * Ty - 2.0 * Typ;
    Ax = first ? nxx : (last ? pvx : alpha * pvx + (1.0 - alpha) * nxx);
    Ay = first ? nxy : (last ? pvy : alpha * pvy + (1.0 - alpha) * nxy);
    const double a2[2] = {Ax, Ay}; double an2[2];
    cm.next(a2, an2);
    if constexpr (SHADOW) {
      if (hasS) {  // A of the last waypoint: the "previous" cubic alone (the `last` case of the lane it replaces)
        an2[0] = fma(2.0, Tx, fma(6.0, wx - wxp, 4.0 * Txp)); an2[1] = fma(2.0, Ty, fma(6.0, wy - wyp, 4.0 * Typ));
      }
    }
    Axp = an2[0]; Ayp = an2[1];
  };

  double cost = 0;
  double e = 1.0;
  if constexpr (DEG == 5) {
    // ================= Lau09 quintic builder (A5), cost + gradient (A6), iRPROP- (A7) =================
    const scb_opt& o = g.opt;
    const double lambda = o.lambda;
    double an_prev, ap_next;
    {
      const double mag = sqrt(btx * btx + bty * bty);
      alpha = inner ? dnext / (dprev + dnext) : 0.0;
      // weights of the "previous" / "next" cubic in A_i (one-sided at the ends)
      const double ap = first ? 0.0 : (last ? 1.0 : alpha), an = last ? 0.0 : (first ? 1.0 : 1.0 - alpha);
      st[kSe * ss] = o.step_e; st[kPe * ss] = 0;
      st[kSwx * ss] = o.step_w * mag; st[kSwy * ss] = o.step_w * mag; st[kPwx * ss] = 0; st[kPwy * ss] = 0;
      st[kW0x * ss] = wx; st[kW0y * ss] = wy; st[kMag * ss] = mag;
      st[kBe * ss] = o.e_init; st[kBwx * ss] = wx; st[kBwy * ss] = wy;
      const double v[2] = {ap, an}; double pv[2], nx[2];
      cm.both(v, pv, nx);
      an_prev = pv[1]; ap_next = nx[0];
      if constexpr (SHADOW) {
        if (hasS) ap_next = 1.0;  // the last waypoint's weight of its "previous" cubic
        st[kSeS * ss] = o.step_e; st[kPeS * ss] = 0; st[kBeS * ss] = o.e_init;
        st[kBtSx * ss] = bsx; st[kBtSy * ss] = bsy; st[kSx * ss] = sx; st[kSy * ss] = sy;
        st[kAnPrevS * ss] = an;   // ... and the weight of this lane's waypoint's "next" cubic, which it sees as an_prev
        eS = o.e_init;
      }
      st[kAlpha * ss] = alpha; st[kAnPrev * ss] = an_prev; st[kApNext * ss] = ap_next;
      e = o.e_init;
    }
    const int K = o.iters;
    for (int iter = 0; iter <= K; ++iter) {
      // ---- A5: T, A per waypoint; hodograph control points D of the lane's segment
      alpha = st[kAlpha * ss];
      double Tx, Ty, Txp, Typ, wxp, wyp, Ax, Ay, Axp, Ayp;
      waypoint_TA(e, Tx, Ty, Txp, Typ, wxp, wyp, Ax, Ay, Axp, Ayp);
      double cx[5], cy[5];
      const double hdx[5] = {Tx, fma(0.25, Ax, Tx), fma(0.25, Axp - Ax, fma(5.0, wxp - wx, -2.0 * (Tx + Txp))), fma(-0.25, Axp, Txp), Txp};
      const double hdy[5] = {Ty, fma(0.25, Ay, Ty), fma(0.25, Ayp - Ay, fma(5.0, wyp - wy, -2.0 * (Ty + Typ))), fma(-0.25, Ayp, Typ), Typ};
      hodograph_power(hdx[0], hdx[1], hdx[2], hdx[3], hdx[4], cx);
      hodograph_power(hdy[0], hdy[1], hdy[2], hdy[3], hdy[4], cy);
      // ---- A6
      double gcx[5], gcy[5], Js;
      if (iter < K) Js = segment_cost<true>(cx, cy, lambda, gcx, gcy);
      else Js = segment_cost<false>(cx, cy, lambda, gcx, gcy);
      if (!seg) Js = 0;
      double J = cm.path_sum(Js, nseg);
      // collision term (Lau09 IV-B): a colliding iterate has infinite cost; it is rejected below
      bool hit = false;
      if constexpr (COLL) {
        hit = cm.path_any(seg && segment_hits<S>(wx, wy, hdx, hdy, ce, g.coll_chords));
        if (hit) J = CUDART_INF;
      }
      const bool better = iter == 0 || J < cost;
      if (better) { cost = J; st[kBe * ss] = e; st[kBwx * ss] = wx; st[kBwy * ss] = wy; }
      if constexpr (SHADOW) { if (better) st[kBeS * ss] = eS; }
      if (iter == K) break;
      // ---- chain rule: D -> (T, A, W) of the segment's two waypoints, then A -> (T, W) of the neighbours
      double gdx[5], gdy[5];
      hodograph_power_T(gcx, gdx); hodograph_power_T(gcy, gdy);
      if (!seg) {
#pragma unroll
        for (int m = 0; m < 5; ++m) { gdx[m] = 0; gdy[m] = 0; }
      }
      // end side of the segment (waypoint i+1) is picked up by the next lane
      double ein[6];
      double gTSx = 0, gTSy = 0, gASx = 0, gASy = 0;  // SHADOW: gradient w.r.t. T and A of the last waypoint
      {
        const double eout[6] = {gdx[3] + gdx[4] - 2.0 * gdx[2], gdy[3] + gdy[4] - 2.0 * gdy[2], 0.25 * (gdx[2] - gdx[3]),
                                0.25 * (gdy[2] - gdy[3]), 5.0 * gdx[2], 5.0 * gdy[2]};
        cm.prev(eout, ein);
        if constexpr (SHADOW) {  // the lane it replaces adds what it receives to its own (zero) segment part
          gTSx = 0.0 + eout[0]; gTSy = 0.0 + eout[1]; gASx = 0.0 + eout[2]; gASy = 0.0 + eout[3];
        }
      }
      double gTx = gdx[0] + gdx[1] - 2.0 * gdx[2], gTy = gdy[0] + gdy[1] - 2.0 * gdy[2];
      double gAx = 0.25 * (gdx[1] - gdx[2]), gAy = 0.25 * (gdy[1] - gdy[2]);
      double gWx = -5.0 * gdx[2], gWy = -5.0 * gdy[2];
      if (!first) { gTx += ein[0]; gTy += ein[1]; gAx += ein[2]; gAy += ein[3]; gWx += ein[4]; gWy += ein[5]; }
      {
        const double ap = first ? 0.0 : (last ? 1.0 : alpha), an = last ? 0.0 : (first ? 1.0 : 1.0 - alpha);
        const double v[2] = {gAx, gAy}; double gm[2], gp[2];
        cm.both(v, gm, gp);
        if constexpr (SHADOW) {
          if (hasS) { gp[0] = gASx; gp[1] = gASy; }
          // the last waypoint: ap = 1, an = 0, so k4 = 4; then its `!first` part with gm = this lane's gA
          const double aS = st[kAnPrevS * ss];
          gTSx = fma(4.0, gASx, gTSx); gTSy = fma(4.0, gASy, gTSy);
          gTSx = fma(-2.0 * aS, gAx, gTSx); gTSy = fma(-2.0 * aS, gAy, gTSy);
        }
        const double k6 = 6.0 * (ap + an), k4 = 4.0 * (ap - an);
        gWx = fma(-k6, gAx, gWx); gWy = fma(-k6, gAy, gWy); gTx = fma(k4, gAx, gTx); gTy = fma(k4, gAy, gTy);
        if (!first) {  // waypoint i is the "next" of i-1
          const double a = st[kAnPrev * ss];
          gWx = fma(6.0 * a, gm[0], gWx); gWy = fma(6.0 * a, gm[1], gWy); gTx = fma(-2.0 * a, gm[0], gTx); gTy = fma(-2.0 * a, gm[1], gTy);
        }
        if (!last) {   // waypoint i is the "previous" of i+1
          const double a = st[kApNext * ss];
          gWx = fma(6.0 * a, gp[0], gWx); gWy = fma(6.0 * a, gp[1], gWy); gTx = fma(2.0 * a, gp[0], gTx); gTy = fma(2.0 * a, gp[1], gTy);
        }
      }
      const double ge = gTx * btx + gTy * bty;
      // largest gradient component of each class over the path (components below 1e-9 of it count as zero)
      double gemax = own ? fabs(ge) : 0.0, gwmax = inner ? fmax(fabs(gWx), fabs(gWy)) : 0.0;
      double geS = 0;
      if constexpr (SHADOW) {
        const double bx = st[kBtSx * ss], by = st[kBtSy * ss];
        geS = gTSx * bx + gTSy * by;
        if (hasS) gemax = fmax(gemax, fabs(geS));
      }
      cm.path_max2(gemax, gwmax);
      // ---- A7: iRPROP-.  A rejected (colliding) trial, once a collision-free iterate exists: back to the best such
      // iterate, every step halved, gradient memory cleared.  The lanes of a rejected path still walked the chain rule
      // above with their warp (its shuffles need every lane); only what they keep differs.
      {
        const bool reject = COLL && hit && cost < CUDART_INF;
        const double te = 1e-9 * gemax, tw = 1e-9 * gwmax;
        const double se0 = st[kSe * ss];
        double se = se0, pe = st[kPe * ss];
        const double gi = fabs(ge) <= te ? 0.0 : ge;
        const double ei = e + rprop_step<double>(gi, pe, se, 1e-6, 1.0);
        e = reject ? st[kBe * ss] : fmin((double)o.e_max, fmax((double)o.e_min, ei));
        st[kSe * ss] = reject ? fmax(0.5 * se0, 1e-6) : se; st[kPe * ss] = reject ? 0.0 : pe;
        if constexpr (SHADOW) {  // the same step for the last waypoint's elongation
          double seS = st[kSeS * ss], peS = st[kPeS * ss];
          const double giS = fabs(geS) <= te ? 0.0 : geS;
          const double eiS = eS + rprop_step<double>(giS, peS, seS, 1e-6, 1.0);
          eS = fmin((double)o.e_max, fmax((double)o.e_min, eiS));
          st[kSeS * ss] = seS; st[kPeS * ss] = peS;
        }
        if (o.move_waypoints && inner) {
          const double mag = st[kMag * ss], w0x = st[kW0x * ss], w0y = st[kW0y * ss];
          const double swx0 = st[kSwx * ss], swy0 = st[kSwy * ss];
          double swx = swx0, swy = swy0, pwx = st[kPwx * ss], pwy = st[kPwy * ss];
          const double gvx = fabs(gWx) <= tw ? 0.0 : gWx, gvy = fabs(gWy) <= tw ? 0.0 : gWy;
          const double smin = 1e-6 * mag, smax = 0.5 * mag, rr = o.w_radius;
          const double nx = wx + rprop_step<double>(gvx, pwx, swx, smin, smax);
          const double ny = wy + rprop_step<double>(gvy, pwy, swy, smin, smax);
          wx = reject ? st[kBwx * ss] : fmin(w0x + rr, fmax(w0x - rr, nx));
          wy = reject ? st[kBwy * ss] : fmin(w0y + rr, fmax(w0y - rr, ny));
          st[kSwx * ss] = reject ? fmax(0.5 * swx0, smin) : swx; st[kSwy * ss] = reject ? fmax(0.5 * swy0, smin) : swy;
          st[kPwx * ss] = reject ? 0.0 : pwx; st[kPwy * ss] = reject ? 0.0 : pwy;
        }
      }
    }
    e = st[kBe * ss]; wx = st[kBwx * ss]; wy = st[kBwy * ss];
    if constexpr (SHADOW) eS = st[kBeS * ss];
    alpha = st[kAlpha * ss];
  }

  // ---- control points of the lane's segment (the frozen construction), cost and the path verdict -> global
  {
    double ox[NC], oy[NC];
    if constexpr (DEG == 3) {  // sc:661-681
      const double Tx = e * btx, Ty = e * bty;
      const double v[4] = {Tx, Ty, wx, wy}; double nx[4];
      cm.next(v, nx);
      ox[0] = wx; oy[0] = wy; ox[1] = wx + Tx; oy[1] = wy + Ty;
      ox[2] = nx[2] - nx[0]; oy[2] = nx[3] - nx[1]; ox[3] = nx[2]; oy[3] = nx[3];
    } else {
      double Tx, Ty, Txp, Typ, wxp, wyp, Ax, Ay, Axp, Ayp;
      waypoint_TA(e, Tx, Ty, Txp, Typ, wxp, wyp, Ax, Ay, Axp, Ayp);
      const double fifth = 1.0 / 5.0, twentieth = 1.0 / 20.0;
      const double p1x = wx + fifth * Tx, p1y = wy + fifth * Ty, p4x = wxp - fifth * Txp, p4y = wyp - fifth * Typ;
      ox[0] = wx; oy[0] = wy; ox[1] = p1x; oy[1] = p1y;
      ox[2] = twentieth * Ax + 2.0 * p1x - wx; oy[2] = twentieth * Ay + 2.0 * p1y - wy;
      ox[3] = twentieth * Axp + 2.0 * p4x - wxp; oy[3] = twentieth * Ayp + 2.0 * p4y - wyp;
      ox[4] = p4x; oy[4] = p4y; ox[5] = wxp; oy[5] = wyp;
    }
    if (seg) {
      S* dx = (S*)g.ctrl_x + ((size_t)p * nseg + i) * NC; S* dy = (S*)g.ctrl_y + ((size_t)p * nseg + i) * NC;
#pragma unroll
      for (int j = 0; j < NC; ++j) { dx[j] = bad ? (S)0 : (S)ox[j]; dy[j] = bad ? (S)0 : (S)oy[j]; }
    }
    if (own && first) {
      if (g.cost) ((S*)g.cost)[p] = bad ? (S)0 : (S)cost;
      g.status[p] = bad ? SCB_ST_ERR_BAD_PATH : SCB_ST_OK;
    }
  }
}

// W <= 32: a warp owns PB = floor(32 / W) consecutive paths
template <class S, int DEG, bool COLL>
__global__ void __launch_bounds__(kLaneWarps * 32, 4) path_lane_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char lane_smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int W = g.W, PB = g.PB, P = g.P;
  const int p0 = (blockIdx.x * kLaneWarps + warp) * PB;
  const int pl = lane / W, i = lane - pl * W, p = p0 + pl;
  WarpComm cm;
  cm.base = pl * W; cm.W = W; cm.sstride = 32;
  cm.st = (double*)lane_smem_raw + (size_t)warp * kLaneSlots * 32 + lane;
  CollEdges ce{nullptr, nullptr, nullptr, 0, 0, nullptr};
  if constexpr (COLL) {  // every thread reaches the barrier (no early return above)
    double* se = (double*)lane_smem_raw + (size_t)kLaneWarps * kLaneSlots * 32;
    float* sb = (float*)(se + 4 * kCollSmem);
    float* sf = sb + 4 * kCollSmem;
    stage_coll_edges<S>(g, se, sb, sf, threadIdx.x, kLaneWarps * 32);
    __syncthreads();
    ce = CollEdges{se, sb, sf, g.n_coll_edges < kCollSmem ? g.n_coll_edges : kCollSmem, g.n_coll_edges, g.coll_edges};
  }
  if (p0 >= P) return;  // the whole warp leaves together
  lane_path<S, DEG, WarpComm, COLL>(g, cm, p, i, pl < PB && p < P, ce);
}

// 3 <= W <= 32, plain quintic optimiser: one lane per SEGMENT, floor(64 / (W - 1)) paths packed over each PAIR of warps
// (W = 10: 7 paths on 63 of 64 lanes, against 3 paths per warp = 54 of 64 above; W = 20: 3 paths per pair against 2).
// Same arithmetic, same summation order: bit-identical control points and cost (tests/test_gpu_parity.py).
template <class S>
__global__ void __launch_bounds__(kLaneWarps * 32, 4) path_pair_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char lane_smem_raw[];
  constexpr int kPairs = kLaneWarps / 2;
  const int pair = threadIdx.x >> 6, t = threadIdx.x & 63;
  const int nseg = g.W - 1, PBp = 64 / nseg, P = g.P;
  const int p0 = (blockIdx.x * kPairs + pair) * PBp;
  const int pl = t / nseg, i = t - pl * nseg, p = p0 + pl;
  PairComm cm;
  cm.tid = t; cm.base = pl * nseg; cm.n = nseg; cm.bar = 1 + pair; cm.sstride = 64;
  cm.xbuf = (double*)lane_smem_raw + (size_t)pair * 2 * PairComm::kMaxVals * 64;
  cm.st = (double*)lane_smem_raw + (size_t)kPairs * 2 * PairComm::kMaxVals * 64 + (size_t)pair * kPairSlots * 64 + t;
  const CollEdges ce{nullptr, nullptr, nullptr, 0, 0, nullptr};
  if (p0 >= P) return;  // the whole pair leaves together (its barrier is its own)
  lane_path<S, 5, PairComm, false, true>(g, cm, p, i, pl < PBp && p < P, ce);
}
constexpr size_t kPairSmem = ((size_t)(kLaneWarps / 2) * 2 * PairComm::kMaxVals * 64 + (size_t)(kLaneWarps / 2) * kPairSlots * 64) * sizeof(double);

// 33 <= W <= 512: one path per CTA of ceil(W / 32) warps (128 registers per thread: 512 threads at most)
template <class S, int DEG, bool COLL>
__global__ void __launch_bounds__(512, 1) path_lane_cta_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char lane_smem_raw[];
  const int T = blockDim.x, tid = threadIdx.x;
  CtaComm cm;
  cm.T = T; cm.tid = tid; cm.sstride = T;
  cm.xbuf = (double*)lane_smem_raw;
  cm.red = cm.xbuf + (size_t)2 * CtaComm::kMaxVals * T;
  cm.st = cm.red + 128 + tid;
  CollEdges ce{nullptr, nullptr, nullptr, 0, 0, nullptr};
  if constexpr (COLL) {
    double* se = cm.red + 128 + (size_t)kLaneSlots * T;
    float* sb = (float*)(se + 4 * kCollSmem);
    float* sf = sb + 4 * kCollSmem;
    stage_coll_edges<S>(g, se, sb, sf, tid, T);
    __syncthreads();
    ce = CollEdges{se, sb, sf, g.n_coll_edges < kCollSmem ? g.n_coll_edges : kCollSmem, g.n_coll_edges, g.coll_edges};
  }
  lane_path<S, DEG, CtaComm, COLL>(g, cm, blockIdx.x, tid, tid < g.W, ce);
}

bool path_lane_shape(int W, int& PB, size_t& smem, int spline_mode) {
  if (W < 2 || W > 512) return false;
  const size_t coll = spline_mode == SCB_REF_CUBIC ? 0 : (size_t)kCollSmem * (4 * sizeof(double) + 8 * sizeof(float));  // staged edges
  if (W <= 32) {
    PB = 32 / W;
    smem = spline_mode == SCB_REF_CUBIC ? 0 : (size_t)kLaneWarps * kLaneSlots * 32 * sizeof(double) + coll;
  } else {
    PB = 1;
    const size_t T = (size_t)(W + 31) / 32 * 32;
    smem = ((size_t)2 * CtaComm::kMaxVals * T + 128 + (size_t)kLaneSlots * T) * sizeof(double) + coll;
  }
  return true;
}

template <class S> void launch_path_lane(const PathArgs& a0, int spline_mode, int PB, size_t smem, cudaStream_t st) {
  PathArgs a = a0;
  a.PB = PB; a.table_in_smem = 0;
  if (a.W <= 32) {
    const int per_cta = PB * kLaneWarps, grid = (a.P + per_cta - 1) / per_cta;
    if (spline_mode == SCB_REF_CUBIC) path_lane_kernel<S, 3, false><<<grid, kLaneWarps * 32, smem, st>>>(a);
    else if (a.n_coll_edges > 0) path_lane_kernel<S, 5, true><<<grid, kLaneWarps * 32, smem, st>>>(a);
    else {
      // the pair mapping where it packs more paths onto the lanes (SCB_PATH_PAIR = 0 / 1: never / wherever it can run)
      const char* e = getenv("SCB_PATH_PAIR");
      const int nseg = a.W - 1, per_pair = a.W >= 3 ? 64 / nseg : 0;
      const bool can = a.W >= 3, wins = can && per_pair > 2 * PB;
      if (e ? (atoi(e) != 0 && can) : wins) {
        const int per_cta2 = per_pair * (kLaneWarps / 2), grid2 = (a.P + per_cta2 - 1) / per_cta2;
        path_pair_kernel<S><<<grid2, kLaneWarps * 32, kPairSmem, st>>>(a);
      } else {
        path_lane_kernel<S, 5, false><<<grid, kLaneWarps * 32, smem, st>>>(a);
      }
    }
  } else {
    const int T = (a.W + 31) / 32 * 32;
    if (spline_mode == SCB_REF_CUBIC) {
      cudaFuncSetAttribute(path_lane_cta_kernel<S, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      path_lane_cta_kernel<S, 3, false><<<a.P, T, smem, st>>>(a);
    } else if (a.n_coll_edges > 0) {
      cudaFuncSetAttribute(path_lane_cta_kernel<S, 5, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      path_lane_cta_kernel<S, 5, true><<<a.P, T, smem, st>>>(a);
    } else {
      cudaFuncSetAttribute(path_lane_cta_kernel<S, 5, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      path_lane_cta_kernel<S, 5, false><<<a.P, T, smem, st>>>(a);
    }
  }
}
template void launch_path_lane<double>(const PathArgs&, int, int, size_t, cudaStream_t);
template void launch_path_lane<float>(const PathArgs&, int, int, size_t, cudaStream_t);

}  // namespace scb
