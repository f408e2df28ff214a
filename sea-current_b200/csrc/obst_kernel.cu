// Obstacle-distance speed limit (Lau09 III-B eq. 2; scb_batch_in.a_brake / t_react), one thread per (gridpoint,
// problem) element (sm_100a).  Reads the curve points px, py [G][P] the arc kernel wrote, finds the distance d to the
// closest obstacle edge (point to segment; the edge list is staged in shared memory and every thread walks it),
// and writes the velocity box of the gridpoint intersected with [-c, c],
//   c = -a_brake t_react + sqrt(a_brake^2 t_react^2 + 2 a_brake d),
// into the per-gridpoint limit arrays the toppra kernels read (vel_lo_grid / vel_hi_grid; the curvature caps, if any,
// are applied on top by those kernels).  Elementwise and coalesced: 2 reads + 2 writes of a real per element.
// The reference's hook for a position-dependent limit is vel_lim_func (sea_current.hpp:1175-1188).
//
// collide_kernel: collision verdict of the finished spline (Lau09 IV-B: "if the trajectory is not collision-free
// according to the obstacle map, its cost is set to infinity"; scb_batch_in.collision_chords).  One thread per
// (problem, segment): the segment is cut into M chords B(k/M) -> B((k+1)/M) (de Casteljau in FP64) and every chord is
// tested against every obstacle edge with the reference's segment test (intersects, sea_current.hpp:142-178).  A hit
// turns status[p] from OK into SCB_ST_ERR_COLLISION and cost[p] into +inf; the timing kernels skip such a problem.
#include "scb_common.cuh"
#include "path_device.cuh"

namespace scb {

namespace {

template <class S, int DEG>
__global__ void __launch_bounds__(256) collide_kernel(CollideArgs g) {
  __shared__ double se[1024 * 4];
  const int ns = g.E < 1024 ? g.E : 1024;
  const S* __restrict__ ge = (const S*)g.edges;
  for (int i = threadIdx.x; i < 4 * ns; i += 256) se[i] = (double)ge[i];
  __syncthreads();
  const int nseg = g.W - 1;
  const size_t item = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (item >= (size_t)g.P * nseg) return;
  const int p = (int)(item / nseg);
  if (g.status[p] != SCB_ST_OK) return;
  const S* __restrict__ cx = (const S*)g.ctrl_x + item * (DEG + 1);
  const S* __restrict__ cy = (const S*)g.ctrl_y + item * (DEG + 1);
  double bx[DEG + 1], by[DEG + 1];
#pragma unroll
  for (int j = 0; j <= DEG; ++j) { bx[j] = (double)cx[j]; by[j] = (double)cy[j]; }
  double ax = bx[0], ay = by[0];
  bool hit = false;
  for (int k = 1; k <= g.M && !hit; ++k) {
    const double u = k == g.M ? 1.0 : (double)k / (double)g.M;
    double tx[DEG + 1], ty[DEG + 1];
#pragma unroll
    for (int j = 0; j <= DEG; ++j) { tx[j] = bx[j]; ty[j] = by[j]; }
#pragma unroll
    for (int r = 1; r <= DEG; ++r)
#pragma unroll
      for (int j = 0; j <= DEG - r; ++j) { tx[j] = (1.0 - u) * tx[j] + u * tx[j + 1]; ty[j] = (1.0 - u) * ty[j] + u * ty[j + 1]; }
    const double qx = tx[0], qy = ty[0];
    // chord and edge with disjoint bounding boxes cannot cross (the segment test's u and t are then outside [0, 1] by
    // far more than its rounding): four compares skip the two divisions for almost every pair
    const double cxl = fmin(ax, qx), cxh = fmax(ax, qx), cyl = fmin(ay, qy), cyh = fmax(ay, qy);
    for (int e = 0; e < g.E && !hit; ++e) {
      double e0x, e0y, e1x, e1y, ix, iy;
      if (e < ns) { e0x = se[4 * e]; e0y = se[4 * e + 1]; e1x = se[4 * e + 2]; e1y = se[4 * e + 3]; }
      else { e0x = (double)ge[4 * e]; e0y = (double)ge[4 * e + 1]; e1x = (double)ge[4 * e + 2]; e1y = (double)ge[4 * e + 3]; }
      if (fmin(e0x, e1x) > cxh || fmax(e0x, e1x) < cxl || fmin(e0y, e1y) > cyh || fmax(e0y, e1y) < cyl) continue;
      hit = seg_intersects<double>(ax, ay, qx, qy, e0x, e0y, e1x, e1y, ix, iy);
    }
    ax = qx; ay = qy;
  }
  if (hit) {  // several segments of one problem may write: the same values
    g.status[p] = SCB_ST_ERR_COLLISION;
    if (g.cost) ((S*)g.cost)[p] = (S)CUDART_INF;
  }
}

constexpr int kObstThreads = 256, kObstEdgesSmem = 512, kObstPer = 4;

template <class S>
__global__ void __launch_bounds__(kObstThreads) obst_cap_kernel(ObstArgs g) {
  // per edge: start point, direction, 1 / |direction|^2 (0 for a degenerate edge: the distance to its point), in double
  // for the exact distance and in float for the screen in front of it
  __shared__ double se[kObstEdgesSmem * 5];
  __shared__ float sf[kObstEdgesSmem * 6];  // the same five + the edge's share of the screen's error bound
  const int ns = g.E < kObstEdgesSmem ? g.E : kObstEdgesSmem;
  const S* __restrict__ ge = (const S*)g.edges;
  for (int k = threadIdx.x; k < ns; k += kObstThreads) {
    const double ax = (double)ge[4 * k], ay = (double)ge[4 * k + 1], ex = (double)ge[4 * k + 2] - ax, ey = (double)ge[4 * k + 3] - ay;
    const double ee = ex * ex + ey * ey;
    se[5 * k] = ax; se[5 * k + 1] = ay; se[5 * k + 2] = ex; se[5 * k + 3] = ey; se[5 * k + 4] = ee > 0 ? 1.0 / ee : 0.0;
    sf[6 * k] = (float)ax; sf[6 * k + 1] = (float)ay; sf[6 * k + 2] = (float)ex; sf[6 * k + 3] = (float)ey;
    sf[6 * k + 4] = ee > 0 ? (float)(1.0 / ee) : 0.0f;
    sf[6 * k + 5] = (float)(fabs(ax) + fabs(ay) + fabs(ex) + fabs(ey));
  }
  __syncthreads();
  const size_t n = (size_t)g.G * g.P;
  const double ab = g.a_brake, at = g.a_brake * g.t_react, at2 = at * at;
  // kObstPer elements per thread, a grid stride apart (coalesced): one broadcast read of an edge serves all of them --
  // with one element per thread the kernel was bound by the shared-memory reads of the edge list, not by the FP64 pipe
  const size_t stride = (size_t)gridDim.x * kObstThreads;
  for (size_t o0 = (size_t)blockIdx.x * kObstThreads + threadIdx.x; o0 < n; o0 += stride * kObstPer) {
    double bx[kObstPer], by[kObstPer], d2[kObstPer];
#pragma unroll
    for (int j = 0; j < kObstPer; ++j) {
      const size_t o = o0 + j * stride;
      const bool in = o < n;
      bx[j] = in ? (double)((const S*)g.px)[o] : 0.0; by[j] = in ? (double)((const S*)g.py)[o] : 0.0; d2[j] = CUDART_INF;
    }
    // Float screen: the distance to every edge in float (FP32 pipe, twice the FP64 rate and otherwise idle), then the
    // exact FP64 distance only to the edges that can still be the closest one.  A float distance is off by at most a few
    // float ulps of the coordinates involved: err <= 4e-6 (|point| + |edge|) is twenty times that.  Every edge with
    // d_float <= d_float_min + 2 err_max is evaluated exactly, the true closest edge is among them, and the minimum of
    // exact values does not depend on which others are - the result has the bits of the all-FP64 loop it replaces
    // (65.5 M points x 24 edges at ~12 FP64 instructions each were 2.1 ms of the C3 obstacle step).
    float fbx[kObstPer], fby[kObstPer], fmin2[kObstPer], fpt[kObstPer], ferr[kObstPer];
#pragma unroll
    for (int j = 0; j < kObstPer; ++j) {
      fbx[j] = (float)bx[j]; fby[j] = (float)by[j]; fmin2[j] = CUDART_INF_F; fpt[j] = fabsf(fbx[j]) + fabsf(fby[j]); ferr[j] = 0.0f;
    }
    for (int k = 0; k < ns; ++k) {
      const float ax = sf[6 * k], ay = sf[6 * k + 1], ex = sf[6 * k + 2], ey = sf[6 * k + 3], ie = sf[6 * k + 4], sz = sf[6 * k + 5];
#pragma unroll
      for (int j = 0; j < kObstPer; ++j) {
        const float rx = fbx[j] - ax, ry = fby[j] - ay;
        const float t = fminf(fmaxf((rx * ex + ry * ey) * ie, 0.0f), 1.0f);
        const float qx = fmaf(-t, ex, rx), qy = fmaf(-t, ey, ry);
        fmin2[j] = fminf(fmin2[j], fmaf(qx, qx, qy * qy));
        ferr[j] = fmaxf(ferr[j], sz);
      }
    }
    float thr2[kObstPer];
#pragma unroll
    for (int j = 0; j < kObstPer; ++j) {
      const float slack = 8e-6f * (fpt[j] + ferr[j]) + 1e-30f;  // 2 err_max
      const float lim = sqrtf(fmin2[j]) + slack;
      thr2[j] = lim * lim * 1.000001f;
    }
    for (int k = 0; k < ns; ++k) {
      const float fax = sf[6 * k], fay = sf[6 * k + 1], fex = sf[6 * k + 2], fey = sf[6 * k + 3], fie = sf[6 * k + 4];
      bool need = false, nd[kObstPer];
#pragma unroll
      for (int j = 0; j < kObstPer; ++j) {
        const float rx = fbx[j] - fax, ry = fby[j] - fay;
        const float t = fminf(fmaxf((rx * fex + ry * fey) * fie, 0.0f), 1.0f);
        const float qx = fmaf(-t, fex, rx), qy = fmaf(-t, fey, ry);
        nd[j] = fmaf(qx, qx, qy * qy) <= thr2[j];
        need = need || nd[j];
      }
      if (!need) continue;
      const double ax = se[5 * k], ay = se[5 * k + 1], ex = se[5 * k + 2], ey = se[5 * k + 3], ie = se[5 * k + 4];
#pragma unroll
      for (int j = 0; j < kObstPer; ++j) {
        if (!nd[j]) continue;
        const double rx = bx[j] - ax, ry = by[j] - ay;
        const double t = fmin(fmax((rx * ex + ry * ey) * ie, 0.0), 1.0);
        const double qx = fma(-t, ex, rx), qy = fma(-t, ey, ry);
        d2[j] = fmin(d2[j], fma(qx, qx, qy * qy));
      }
    }
    for (int k = ns; k < g.E; ++k) {  // an edge list longer than the shared-memory stage: the rest from global memory
      const double ax = (double)ge[4 * k], ay = (double)ge[4 * k + 1], ex = (double)ge[4 * k + 2] - ax, ey = (double)ge[4 * k + 3] - ay;
      const double ee = ex * ex + ey * ey;
#pragma unroll
      for (int j = 0; j < kObstPer; ++j) {
        const double rx = bx[j] - ax, ry = by[j] - ay;
        const double t = fmin(fmax(ee > 0 ? (rx * ex + ry * ey) / ee : 0.0, 0.0), 1.0);
        const double qx = rx - t * ex, qy = ry - t * ey;
        d2[j] = fmin(d2[j], qx * qx + qy * qy);
      }
    }
#pragma unroll
    for (int j = 0; j < kObstPer; ++j) {
      const size_t o = o0 + j * stride;
      if (o >= n) break;
      const int p = (int)(o % (size_t)g.P);
      const double c = -at + sqrt(at2 + 2.0 * ab * sqrt(d2[j]));
      double lo, hi;
      if (g.vel_lo_grid && g.vel_hi_grid) { lo = (double)((const S*)g.vel_lo_grid)[o]; hi = (double)((const S*)g.vel_hi_grid)[o]; }
      else { lo = (double)((const S*)g.vel_lo)[p]; hi = (double)((const S*)g.vel_hi)[p]; }
      ((S*)g.out_lo)[o] = (S)fmax(lo, -c);
      ((S*)g.out_hi)[o] = (S)fmin(hi, c);
    }
  }
}

}  // namespace

template <class R> void launch_obst_cap(const ObstArgs& a, cudaStream_t st) {
  const size_t n = (size_t)a.G * a.P;
  const size_t want = (n + (size_t)kObstThreads * kObstPer - 1) / ((size_t)kObstThreads * kObstPer);
  const int grid = (int)(want < (size_t)148 * 16 ? (want ? want : 1) : (size_t)148 * 16);
  obst_cap_kernel<R><<<grid, kObstThreads, 0, st>>>(a);
}
template void launch_obst_cap<double>(const ObstArgs&, cudaStream_t);
template void launch_obst_cap<float>(const ObstArgs&, cudaStream_t);

template <class R> void launch_collide(const CollideArgs& a, int degree, cudaStream_t st) {
  const size_t n = (size_t)a.P * (a.W - 1);
  const int grid = (int)((n + 255) / 256);
  if (degree == 3) collide_kernel<R, 3><<<grid, 256, 0, st>>>(a);
  else collide_kernel<R, 5><<<grid, 256, 0, st>>>(a);
}
template void launch_collide<double>(const CollideArgs&, int, cudaStream_t);
template void launch_collide<float>(const CollideArgs&, int, cudaStream_t);

}  // namespace scb
