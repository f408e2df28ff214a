// Path stage on the GPU (sm_100a): everything that depends only on the waypoints.
//   tangent heuristics           sea_current.hpp:343-377
//   segment builder from_path    sea_current.hpp:599-683 (REF_CUBIC: control points only — the
//                                reference's 10,001-point preview sampling feeds nothing downstream)
//   Lau09 quintic builder + curvature/arc-length cost + analytic gradient + iRPROP- loop
//                                (LAU_QUINTIC; no reference code — DESIGN.md "New stages")
//   arc-length tables            sea_current.hpp:767-896 (100 sub-intervals x 32-point
//                                Gauss-Legendre per segment, cumulative table of 101 entries)
//   curve point and curvature at the arc length q(s_i) of every toppra gridpoint
//                                (resample sc:898-1005 + curvature sc:1017-1039; exact table
//                                inversion instead of the float Chebyshev fit, DESIGN.md "Deviations")
//
// One CTA owns PB consecutive paths.  All per-path state (waypoints, optimiser state, control
// points, gradients, power-basis coefficients and — when it fits — the arc-length tables) lives in
// shared memory; the CTA's threads are re-mapped phase by phase onto the work items that exist in
// that phase: (path, waypoint), (path, segment), (path, segment, node group), (path, segment,
// sub-interval), (gridpoint, path).  No per-item warp reductions except a 4-lane butterfly in the
// cost/gradient phase.  Per-path arrays are [P][n] in HBM (a CTA's slice is contiguous); gridpoint
// arrays are [G][P] and are written with the path index fastest.
#include "gl32_device.h"
#include "scb_common.cuh"

namespace scb {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kNG = 4;          // lanes per segment in the cost/gradient phase
constexpr int kNodesPerLane = 32 / kNG;

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
  const R mult = (test_offset - offset < 0) ? (R)-1 : (R)1;  // no angle wrapping, as the reference
  R sn, cs;
  sincos(offset + mult * theta, &sn, &cs);
  R lx = sn, ly = -cs;
  const R n2 = lx * lx + ly * ly;
  if (n2 > 0) { const R n = sqrt(n2); lx = lx / n; ly = ly / n; }
  const R m2 = dist2<R>(w1x + lx, w1y + ly, w2x, w2y) < dist2<R>(w1x - lx, w1y - ly, w2x, w2y) ? (R)1 : (R)-1;
  const R mag = (R)0.5 * fmin(du, dv);  // sc:343-345: |W0W1| = |u|, |W1W2| = |v|
  tx = mag * (m2 * lx); ty = mag * (m2 * ly);
}

// Shared-memory budget, in reals, of a CTA that owns PB paths (host and device agree on it).
__host__ __device__ inline size_t path_smem_reals(int W, int deg, int PB, bool table_in_smem) {
  const size_t nseg = W - 1;
  size_t per = (size_t)(deg == 5 ? 29 : 4) * W                       // per-waypoint arrays
               + nseg * (size_t)(2 * (deg + 1) + 2 * deg + 2 * (deg + 1) + 3 + (deg == 5 ? 13 : 0))
               + 8;                                                   // per-path scalars
  if (table_in_smem) per += nseg * kTable;
  return per * PB + (deg == 5 ? 32 * 13 : 0);
}
constexpr int kPathSmemIntBytes = 256;  // flags in front of the reals

template <class R> struct Carve {
  R* base; size_t off;
  __device__ __forceinline__ Carve(R* b) : base(b), off(0) {}
  __device__ __forceinline__ R* take(size_t n) { R* p = base + off; off += n; return p; }
};

template <class R> struct HermiteQ {  // q(s) of gen_vel_prof's 1-DOF path (piecewise_poly_path.cpp:364-375)
  R c0, c1, c2;
  __device__ __forceinline__ void init(R L, R vs, R ve) { c2 = vs; c0 = (ve - (R)2 * L + vs); c1 = (ve - vs - (R)3 * c0) / (R)2; }
  __device__ __forceinline__ R q(R s) const { return c0 * (s * s * s) + c1 * (s * s) + c2 * s; }
};

__device__ __constant__ double kPGL4X[4] = {-0.8611363115940525752, -0.3399810435848562648, 0.3399810435848562648,
                                            0.8611363115940525752};
__device__ __constant__ double kPGL4W[4] = {0.3478548451374538574, 0.6521451548625461426, 0.6521451548625461426,
                                            0.3478548451374538574};

template <class R, int DEG>
__global__ void __launch_bounds__(512, 1) path_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NC = DEG + 1;
  const int tid = threadIdx.x, T = blockDim.x;
  const int W = g.W, nseg = W - 1, PB = g.PB, P = g.P;
  const int p0 = blockIdx.x * PB;
  const int PBa = min(PB, P - p0);  // paths of this CTA that exist
  int* bad = (int*)smem_raw;        // [PB]
  int* improve = bad + 32;          // [PB]
  Carve<R> sm((R*)(smem_raw + kPathSmemIntBytes));
  // cumulative segment ends are FP64 in both precision modes (first carve: 8-byte aligned): the arc
  // position of a gridpoint is resolved against them before it is narrowed to a segment-local R
  double* segend = (double*)sm.take((size_t)PB * nseg * (sizeof(double) / sizeof(R)));
  R* wx = sm.take(PB * W); R* wy = sm.take(PB * W); R* tx = sm.take(PB * W); R* ty = sm.take(PB * W);
  R* cx = sm.take(PB * nseg * NC); R* cy = sm.take(PB * nseg * NC);
  R* hx = sm.take(PB * nseg * DEG); R* hy = sm.take(PB * nseg * DEG);
  R* ax = sm.take(PB * nseg * NC); R* ay = sm.take(PB * nseg * NC);
  R* segtot = sm.take(PB * nseg);
  R* scal = sm.take(PB * 8);  // per path: 0 J, 1 bestJ, 2 gemax, 3 gwmax, 4 L, 5 vel_start, 6 vel_end

  // ---- P0: load waypoints (coalesced: the CTA's slice of [P][W] is contiguous), validate
  if (tid < 32) { bad[tid] = 0; improve[tid] = 0; }
  __syncthreads();
  for (int it = tid; it < PBa * W; it += T) {
    const R x = ((const R*)g.wx)[(size_t)p0 * W + it], y = ((const R*)g.wy)[(size_t)p0 * W + it];
    wx[it] = x; wy[it] = y;
    if (!isfinite(x) || !isfinite(y)) bad[it / W] = 1;
  }
  __syncthreads();
  for (int it = tid; it < PBa * nseg; it += T) {
    const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
    if (!(dist2<R>(wx[i], wy[i], wx[i + 1], wy[i + 1]) > 0)) bad[pl] = 1;
  }
  // ---- P1: tangents, one item per (path, waypoint)
  for (int it = tid; it < PBa * W; it += T) {
    const int pl = it / W, i = it - pl * W;
    const R* x = wx + pl * W; const R* y = wy + pl * W;
    R ox, oy;
    if (i == 0) {  // sc:371-373, 606-609
      R ang = g.start_angle ? ((const R*)g.start_angle)[p0 + pl] : Num<R>::nan();
      if (isnan(ang)) ang = atan2(y[1] - y[0], x[1] - x[0]);
      const R mag = (R)0.5 * dist2<R>(x[0], y[0], x[1], y[1]);
      R sn, cs; sincos(ang, &sn, &cs);
      ox = mag * cs; oy = mag * sn;
    } else if (i == W - 1) {  // sc:375-377
      const R dx = x[i] - x[i - 1], dy = y[i] - y[i - 1];
      const R mag = (R)0.5 * dist2<R>(x[i - 1], y[i - 1], x[i], y[i]);
      const R n = sqrt(dx * dx + dy * dy);
      ox = mag * (dx / n); oy = mag * (dy / n);
    } else {
      calc_tangent<R>(x[i - 1], y[i - 1], x[i], y[i], x[i + 1], y[i + 1], ox, oy);
    }
    tx[it] = ox; ty[it] = oy;
  }
  __syncthreads();

  if constexpr (DEG == 3) {
    // ---- A2: cubic control points [W_i, W_i + T_i, W_{i+1} - T_{i+1}, W_{i+1}]  (sc:661-681)
    for (int it = tid; it < PBa * nseg; it += T) {
      const int pl = it / nseg, s = it - pl * nseg, i = pl * W + s;
      R* ox = cx + 4 * it; R* oy = cy + 4 * it;
      ox[0] = wx[i]; oy[0] = wy[i];
      ox[1] = wx[i] + tx[i]; oy[1] = wy[i] + ty[i];
      ox[2] = wx[i + 1] - tx[i + 1]; oy[2] = wy[i + 1] - ty[i + 1];
      ox[3] = wx[i + 1]; oy[3] = wy[i + 1];
    }
    if (tid < PBa) scal[tid * 8 + 1] = 0;
  } else {
#include "path_kernel_lau.inl"
  }
  __syncthreads();
#include "path_kernel_tail.inl"
}

size_t path_smem_bytes(int W, int spline_mode, int real_size, int PB, bool table_in_smem) {
  return kPathSmemIntBytes + path_smem_reals(W, spline_mode == SCB_REF_CUBIC ? 3 : 5, PB, table_in_smem) * real_size;
}

// Launch geometry: PB paths per CTA so that two CTAs fit one SM when possible; the arc-length
// tables stay in shared memory when one path's state fits next to them.
bool path_launch_shape(int W, int spline_mode, int real_size, int& PB, int& threads, bool& table_in_smem, size_t& smem) {
  const size_t budget2 = 111 * 1024, budget1 = 224 * 1024;
  const int deg = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  auto bytes = [&](int pb, bool tab) { return path_smem_bytes(W, spline_mode, real_size, pb, tab); };
  table_in_smem = bytes(1, true) <= budget1;
  if (bytes(1, table_in_smem) > budget1) return false;
  const int lanes_per_path = (W - 1) * (deg == 5 ? kNG : 8);  // items of the widest compute phase
  if (bytes(1, table_in_smem) <= budget2) {
    // two CTAs of <= 256 threads per SM (128 registers each): as many paths as fill 256 lanes and fit
    PB = 1;
    while (PB < 32 && bytes(PB + 1, table_in_smem) <= budget2 && (PB + 1) * lanes_per_path <= 256) ++PB;
    threads = 256;
  } else {
    PB = 1;  // one big CTA per SM
    threads = 512;
  }
  smem = bytes(PB, table_in_smem);
  return true;
}

template <class R> void launch_path(const PathArgs& a0, int spline_mode, cudaStream_t st) {
  PathArgs a = a0;
  int PB, threads; bool tab; size_t smem;
  if (!path_launch_shape(a.W, spline_mode, (int)sizeof(R), PB, threads, tab, smem)) return;
  a.PB = PB; a.table_in_smem = tab ? 1 : 0;
  const int grid = (a.P + PB - 1) / PB;
  if (spline_mode == SCB_REF_CUBIC) {
    cudaFuncSetAttribute(path_kernel<R, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(path_kernel<R, 3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    path_kernel<R, 3><<<grid, threads, smem, st>>>(a);
  } else {
    cudaFuncSetAttribute(path_kernel<R, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(path_kernel<R, 5>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    path_kernel<R, 5><<<grid, threads, smem, st>>>(a);
  }
}
template void launch_path<double>(const PathArgs&, int, cudaStream_t);
template void launch_path<float>(const PathArgs&, int, cudaStream_t);

}  // namespace scb
