// Path stage, CTA-team variant (sm_100a): what depends only on the waypoints, up to the control points.
//   tangent heuristics           sea_current.hpp:343-377
//   segment builder from_path    sea_current.hpp:599-683 (REF_CUBIC: control points only — the
//                                reference's 10,001-point preview sampling feeds nothing downstream)
//   Lau09 quintic builder + curvature/arc-length cost + analytic gradient + iRPROP- loop
//                                (LAU_QUINTIC; no reference code — DESIGN.md "New stages")
// Arc-length tables and the geometry at the toppra gridpoints follow in arc_kernel.cu.
//
// This is the variant for paths whose optimiser state does not fit a warp's slice of shared memory
// (long paths, e.g. C5's 200 waypoints); paths of up to 32 waypoints run path_lane_kernel.cu.  One CTA owns PB
// consecutive paths, all per-path state lives in shared memory, and the CTA's threads are re-mapped
// phase by phase onto the items of that phase: (path, waypoint), (path, segment), (path, segment,
// node group).  Per-path arrays are [P][n] in HBM (a CTA's slice is contiguous).
#include <cstdlib>

#include "path_device.cuh"

namespace scb {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kNG = 4;          // lanes per segment in the cost/gradient phase
constexpr int kNodesPerLane = 32 / kNG;

// Shared-memory budget, in reals, of a CTA that owns PB paths (host and device agree on it).
__host__ __device__ inline size_t path_smem_reals(int W, int deg, int PB, bool table_in_smem) {
  const size_t nseg = W - 1;
  size_t per = (size_t)(deg == 5 ? 29 : 4) * W                       // per-waypoint arrays
               + nseg * (size_t)(2 * (deg + 1) + (deg == 5 ? 13 : 0))
               + 8;                                                   // per-path scalars
  (void)table_in_smem;
  return per * PB + (deg == 5 ? 32 * 13 : 0);
}
constexpr int kPathSmemIntBytes = 256;  // flags in front of the reals

// R: arithmetic and shared-memory type (always double, see path_lane_kernel.cu "Arithmetic"); S: storage type in HBM
template <class R, int DEG, class S>
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
  R* wx = sm.take(PB * W); R* wy = sm.take(PB * W); R* tx = sm.take(PB * W); R* ty = sm.take(PB * W);
  R* cx = sm.take(PB * nseg * NC); R* cy = sm.take(PB * nseg * NC);
  R* scal = sm.take(PB * 8);  // per path: 0 J, 1 bestJ, 2 gemax, 3 gwmax, 4 L, 5 vel_start, 6 vel_end

  // ---- P0: load waypoints (coalesced: the CTA's slice of [P][W] is contiguous), validate
  if (tid < 32) { bad[tid] = 0; improve[tid] = 0; }
  __syncthreads();
  for (int it = tid; it < PBa * W; it += T) {
    const R x = (R)((const S*)g.wx)[(size_t)p0 * W + it], y = (R)((const S*)g.wy)[(size_t)p0 * W + it];
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
      R ang = g.start_angle ? (R)((const S*)g.start_angle)[p0 + pl] : Num<R>::nan();
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
    if (g.n_edges > 0) shrink_tangent<R, S>(x[i], y[i], (const S*)g.edges, g.n_edges, ox, oy);  // sc:615, 620, 637
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
  // ---- control points, cost and the path verdict so far -> global; arc_kernel.cu takes over from here
  {
    const int nc = nseg * NC;
    for (int it = tid; it < PBa * nc; it += T) {
      const bool b = bad[it / nc] != 0;
      ((S*)g.ctrl_x)[(size_t)p0 * nc + it] = b ? (S)0 : (S)cx[it];
      ((S*)g.ctrl_y)[(size_t)p0 * nc + it] = b ? (S)0 : (S)cy[it];
    }
    if (tid < PBa) {
      const int p = p0 + tid;
      if (g.cost) ((S*)g.cost)[p] = bad[tid] ? (S)0 : (S)scal[tid * 8 + 1];
      g.status[p] = bad[tid] ? SCB_ST_ERR_BAD_PATH : SCB_ST_OK;
    }
  }
}

size_t path_smem_bytes(int W, int spline_mode, int real_size, int PB) {
  return kPathSmemIntBytes + path_smem_reals(W, spline_mode == SCB_REF_CUBIC ? 3 : 5, PB, false) * real_size;
}

static bool use_lane_kernel() {  // SCB_PATH_TEAM=cta forces the CTA-team kernel (A/B measurements, tests)
  const char* e = getenv("SCB_PATH_TEAM");
  return !(e && e[0] == 'c');
}

// Launch geometry of the path stage: the lane-per-waypoint kernels (threads = 128 as the marker) for W <= 512,
// else the CTA-team kernel: PB paths per CTA so that two CTAs fit one SM when possible.  False = the path is too
// long for the CTA kernel or for arc_kernel's shared-memory power bases.  `real_size` is the STORAGE size
// (arc kernel); the path kernels compute in FP64 in both modes.  The arc-length tables always live in
// global memory (table_in_smem is kept for the ABI of this helper only).
bool path_launch_shape(int W, int spline_mode, int real_size, int& PB, int& threads, bool& table_in_smem, size_t& smem) {
  table_in_smem = false;
  size_t arc_smem;
  if (!arc_paths_per_cta(W, spline_mode, real_size, arc_smem)) return false;
  if (use_lane_kernel() && path_lane_shape(W, PB, smem, spline_mode)) { threads = 128; return true; }
  const size_t budget2 = 111 * 1024, budget1 = 224 * 1024;
  const int deg = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  auto bytes = [&](int pb) { return path_smem_bytes(W, spline_mode, 8, pb); };
  if (bytes(1) > budget1) return false;
  const int lanes_per_path = (W - 1) * (deg == 5 ? kNG : 8);  // items of the widest compute phase
  if (bytes(1) <= budget2) {
    // two CTAs of <= 256 threads per SM (128 registers each): as many paths as fill 256 lanes and fit
    PB = 1;
    while (PB < 32 && bytes(PB + 1) <= budget2 && (PB + 1) * lanes_per_path <= 256) ++PB;
    threads = 256;
  } else {
    PB = 1;  // one big CTA per SM
    threads = 512;
  }
  smem = bytes(PB);
  return true;
}

// path stage = control points (this file or path_lane_kernel.cu) + arc stage (arc_kernel.cu): 2 launches.
// R is the storage type of the mode; the control-point kernels compute in FP64.
template <class R> void launch_path(const PathArgs& a0, int spline_mode, cudaStream_t st, cudaEvent_t between) {
  PathArgs a = a0;
  int PB, threads; bool tab; size_t smem;
  if (!path_launch_shape(a.W, spline_mode, (int)sizeof(R), PB, threads, tab, smem)) return;
  if (threads == 128) {
    launch_path_lane<R>(a0, spline_mode, PB, smem, st);
  } else {
    a.PB = PB; a.table_in_smem = 0;
    const int grid = (a.P + PB - 1) / PB;
    if (spline_mode == SCB_REF_CUBIC) {
      cudaFuncSetAttribute(path_kernel<double, 3, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(path_kernel<double, 3, R>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
      path_kernel<double, 3, R><<<grid, threads, smem, st>>>(a);
    } else {
      cudaFuncSetAttribute(path_kernel<double, 5, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaFuncSetAttribute(path_kernel<double, 5, R>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
      path_kernel<double, 5, R><<<grid, threads, smem, st>>>(a);
    }
  }
  if (between) cudaEventRecord(between, st);  // splits spline_ms into path_ms + arc_ms
  launch_arc<R>(a0, spline_mode, st);
}
template void launch_path<double>(const PathArgs&, int, cudaStream_t, cudaEvent_t);
template void launch_path<float>(const PathArgs&, int, cudaStream_t, cudaEvent_t);

}  // namespace scb
