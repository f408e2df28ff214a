// Arc stage on the GPU (sm_100a): everything between the control points and the toppra kernels.
//   arc-length tables    sea_current.hpp:767-896 (100 sub-intervals per segment, cumulative table of
//                        101 entries), path length, final path status
//   geometry             curve point and curvature at the arc length q(s_i) of every toppra gridpoint
//                        (resample sc:898-1005 + curvature sc:1017-1039; exact table inversion instead
//                        of the float Chebyshev fit, DESIGN.md "Deviations")
// Input: the control points the path kernels wrote ([P][(W-1)(d+1)], one path contiguous).
//
// Split from the path kernels because the two halves want different machines (profiles/r1c_*, r1d_*):
// the optimiser needs 128 registers per thread (16 warps per SM); these phases are embarrassingly
// parallel, need ~80, and run here with 24 warps per SM and no per-path serial sections.
//
// One CTA owns PB consecutive paths; its threads are re-mapped per phase:
//   (path, segment)                 power basis of the segment and of its hodograph -> shared memory
//   (path, segment, piece)          one 0.01-wide piece per thread and turn -> raw piece lengths in the table
//   warp per (path, segment)        lanes 0..24 re-read 4 consecutive pieces each; a warp scan of the lane
//                                   totals gives the cumulative table, stored 32 B per lane
//   (gridpoint, path), path fastest geometry; stores to the [G][P] arrays are PB-contiguous
//
// Arc-length pieces: the reference integrates |B'| over every 0.01-wide sub-interval with a 32-point
// Gauss-Legendre rule.  The integrand is analytic there; a thread evaluates the 7-point Kronrod extension
// of the 3-point Gauss rule (7 nodes, exact to degree 11 like the 6-point Gauss rule) and keeps it when
// the embedded Gauss value agrees to 1e-9 relative (FP32: 1e-5) -- K7's own error is then ~the square of
// that -- and otherwise runs the reference's 32 nodes in the reference's order.  Measured on the C2/C3
// control points (numpy): 0.27 % / 0.01 % of the pieces take the long way, accepted pieces are within
// 2.4e-14 relative of the 32-point value (DESIGN.md "Arc length").
#include "path_device.cuh"

namespace scb {

constexpr int kArcThreads = 256;
constexpr int kPiecesPerLane = 4, kScanLanes = kSub / kPiecesPerLane;  // 25 lanes x 4 pieces = 100
// Coarse inverse of a segment's table: for the 33 arc lengths j / 32 of the segment's length the piece that holds
// them (one byte each).  The lookup of a gridpoint starts from the linear interpolation between two entries instead of
// from the proportional guess r / length, which assumes a constant speed along the segment and was off by more than
// the 4-wide prefetch window on most gridpoints of a quintic segment: the walk that repaired it was 13 % of this
// kernel's stall samples and 8.5 % of its instructions (profiles/r2j arc_kernel lines, path_device.cuh:195-196).
constexpr int kCoarse = 32, kCoarseStride = 36;

__host__ __device__ inline size_t arc_smem_bytes(int W, int deg, int PB, int real_size, bool table_in_smem) {
  const size_t nseg = W - 1;
  const size_t reals = (size_t)PB * nseg * (2 * (deg + 1) + 2 * deg + 2 + (table_in_smem ? kTable : 0));
  const size_t coarse = (size_t)PB * nseg * kCoarseStride;  // inverse-table guesses (bytes), whichever memory the tables are in
  return 64 + sizeof(double) * PB * (nseg + 4) + (reals * real_size + 15) / 16 * 16 + coarse;
}

// TS: the CTA's arc-length tables also live in shared memory (the geometry phase reads them from there; the
// table lookups out of L1/L2 were a third of this kernel's stall samples, profiles/r1i_arc_kernel_f64.lines.txt)
template <class R, int DEG, bool TS>
__global__ void __launch_bounds__(kArcThreads, 3) arc_kernel(PathArgs g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NC = DEG + 1;
  constexpr unsigned kFullMask = 0xffffffffu;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarps = T >> 5;
  const int W = g.W, nseg = W - 1, PB = g.PB, P = g.P;
  const int p0 = blockIdx.x * PB;
  const int PBa = min(PB, P - p0);
  int* bad = (int*)smem_raw;                         // [PB] (PB <= 16)
  double* segend = (double*)(smem_raw + 64);         // [PB * nseg] cumulative segment ends, FP64 in both modes
  double* hqc = segend + (size_t)PB * nseg;          // [PB][4]: c0, c1, c2 of q(s) (gen_vel_prof's Hermite path)
  Carve<R> sm((R*)(smem_raw + 64 + sizeof(double) * (size_t)PB * (nseg + 4)));
  R* ax = sm.take((size_t)PB * nseg * NC); R* ay = sm.take((size_t)PB * nseg * NC);
  R* hx = sm.take((size_t)PB * nseg * DEG); R* hy = sm.take((size_t)PB * nseg * DEG);
  R* segtot = sm.take((size_t)PB * nseg);
  R* invlen = sm.take((size_t)PB * nseg);            // 100 / segment length
  R* stab = TS ? sm.take((size_t)PB * nseg * kTable) : nullptr;
  unsigned char* coarse = smem_raw + arc_smem_bytes(W, DEG, PB, (int)sizeof(R), TS) - (size_t)PB * nseg * kCoarseStride;
  const int nS = PBa * nseg, nc = nseg * NC, nt = nseg * kTable;

  // ---- status of the path kernel, power basis of every segment
  if (tid < PBa) bad[tid] = g.status[p0 + tid] != SCB_ST_OK;
  for (int it = tid; it < nS; it += T) {
    const R* cx = (const R*)g.ctrl_x + (size_t)p0 * nc + (size_t)it * NC;
    const R* cy = (const R*)g.ctrl_y + (size_t)p0 * nc + (size_t)it * NC;
    power_basis<R, DEG>(cx, cy, ax + NC * it, ay + NC * it, hx + DEG * it, hy + DEG * it);
  }
  __syncthreads();

  // ---- A4, step 1: every 0.01-wide piece of every segment, one piece per thread and turn (all lanes busy;
  // the warp-per-segment form kept 25 of 32).  Raw piece lengths go to the table's own slots [k + 1].
  R* tab = (R*)g.table + (size_t)p0 * nt;
  R* wtab = TS ? stab : tab;  // where the pieces and the lookups live
  for (int it = tid; it < nS * kSub; it += T) {
    const int sg = it / kSub, k = it - sg * kSub;
    R kx[DEG], ky[DEG];
#pragma unroll
    for (int j = 0; j < DEG; ++j) { kx[j] = hx[DEG * sg + j]; ky[j] = hy[DEG * sg + j]; }
    auto speed = [&](R u) {
      R dx = kx[DEG - 1], dy = ky[DEG - 1];
#pragma unroll
      for (int j = DEG - 2; j >= 0; --j) { dx = dx * u + kx[j]; dy = dy * u + ky[j]; }
      // full-precision rsqrt here: a one-sided 1e-13 error in every sample would bias the path length, and the
      // timing stage differentiates what it derives from it (see rsqrt_n1 in scb_common.cuh)
      const R n2 = fma(dy, dy, fma(dx, dx, speed_floor<R>()));  // > 0 also at a cusp: no guard around the rsqrt
      return n2 * fast_rsqrt<R>(n2);
    };
    const R a = fmin((R)k * (R)0.01, (R)1), b = fmin((R)(k + 1) * (R)0.01, (R)1);  // sc:851
    const R half = (b - a) / 2, mid = (b + a) / 2;
    // Gauss 3 inside Kronrod 7 (7 nodes, K7 exact to degree 11 like the 6-point Gauss rule)
    const R f0 = speed(mid);
    const R f1 = speed(mid - half * (R)kK7X[0]) + speed(mid + half * (R)kK7X[0]);
    const R f2 = speed(mid - half * (R)kK7X[1]) + speed(mid + half * (R)kK7X[1]);
    const R f3 = speed(mid - half * (R)kK7X[2]) + speed(mid + half * (R)kK7X[2]);
    const R s3 = (R)(8.0 / 9.0) * f0 + (R)(5.0 / 9.0) * f2;
    const R s7 = (R)kK7W[0] * f0 + (R)kK7W[1] * f1 + (R)kK7W[2] * f2 + (R)kK7W[3] * f3;
    R piece = s7 * half;
    const R tol = sizeof(R) == 8 ? (R)1e-9 : (R)1e-5;
    if (!(fabs(s3 - s7) <= tol * s7)) {  // the reference's 32 nodes in table order (sc:854-880); rare
      R acc = 0;
#pragma unroll 4
      for (int i = 0; i < 32; ++i) acc += (R)GL32_W[i] * speed(half * (R)GL32_X[i] + mid);
      piece = acc * half;
    }
    wtab[(size_t)sg * kTable + k + 1] = piece;
  }
  __syncthreads();  // pieces are read back by another thread of this CTA

  // ---- A4, step 2: cumulative table, one warp per (path, segment): lanes 0..24 own 4 consecutive pieces
  for (int sg = warp; sg < nS; sg += nwarps) {
    R* c = tab + (size_t)sg * kTable;
    R* cs = wtab + (size_t)sg * kTable;
    const bool live = lane < kScanLanes;
    R run[kPiecesPerLane];
#pragma unroll
    for (int j = 0; j < kPiecesPerLane; ++j) run[j] = live ? cs[lane * kPiecesPerLane + j + 1] : (R)0;
    // lane-local running sums, then an inclusive warp scan of the lane totals (sc:882-886 adds the
    // pieces one after the other; this order differs from it by rounding only)
#pragma unroll
    for (int j = 1; j < kPiecesPerLane; ++j) run[j] = run[j - 1] + run[j];
    R tot = live ? run[kPiecesPerLane - 1] : (R)0, incl = tot;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const R up = __shfl_up_sync(kFullMask, incl, off);
      if (lane >= off) incl += up;
    }
    const R base = incl - tot;
    if (live) {
#pragma unroll
      for (int j = 0; j < kPiecesPerLane; ++j) {
        const R v = base + run[j];
        c[lane * kPiecesPerLane + j + 1] = v;
        if (TS) cs[lane * kPiecesPerLane + j + 1] = v;
      }
    }
    if (lane == 0) { c[0] = 0; if (TS) cs[0] = 0; }
    if (lane == kScanLanes - 1) segtot[sg] = incl;
    {  // coarse inverse: lane j finds the piece that holds the arc length (j / 32) of the segment
      __syncwarp();  // the table entries above were written by other lanes of this warp (shared or global memory)
      const R tot = __shfl_sync(kFullMask, incl, kScanLanes - 1);
      unsigned char* cz = coarse + (size_t)sg * kCoarseStride;
      const R target = tot * ((R)lane * (R)(1.0 / kCoarse));
      int lo = 0, hi = kSub;  // largest k in [0, 99] with cs[k] <= target
      while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (cs[mid] <= target) lo = mid; else hi = mid; }
      cz[lane] = (unsigned char)lo;
      if (lane == 0) cz[kCoarse] = (unsigned char)(kSub - 1);
    }
  }
  __syncthreads();

  // ---- path length, final status
  if (tid < PBa) {
    const int p = p0 + tid;
    double acc = 0;
    for (int s = 0; s < nseg; ++s) {
      const R len = segtot[tid * nseg + s];
      acc += (double)len; segend[tid * nseg + s] = acc; invlen[tid * nseg + s] = (R)kSub / len;
    }
    const bool ok = !bad[tid] && acc > 0 && isfinite(acc);
    if (!ok) bad[tid] = 1;
    {
      HermiteQ<double> hq;
      hq.init(acc, g.vel_start ? (double)((const R*)g.vel_start)[p] : 0.0, g.vel_end ? (double)((const R*)g.vel_end)[p] : 0.0);
      hqc[tid * 4] = hq.c0; hqc[tid * 4 + 1] = hq.c1; hqc[tid * 4 + 2] = hq.c2;
    }
    ((R*)g.L)[p] = ok ? (R)acc : (R)0;
    if (!ok && g.cost) ((R*)g.cost)[p] = 0;
    g.status[p] = ok ? SCB_ST_OK : SCB_ST_ERR_BAD_PATH;
  }
  __syncthreads();
  for (int it = tid; it < PBa * nt; it += T) if (bad[it / nt]) tab[it] = 0;
  if (g.ctrl_x && g.ctrl_y) {
    for (int it = tid; it < PBa * nc; it += T) {
      if (bad[it / nc]) { ((R*)g.ctrl_x)[(size_t)p0 * nc + it] = 0; ((R*)g.ctrl_y)[(size_t)p0 * nc + it] = 0; }
    }
  }
  __syncthreads();  // the tables written above are read below by other threads of this CTA

  // ---- geometry at the toppra gridpoints: q(s_i) -> (segment, u) -> point, curvature
  if (g.px || g.py || g.kappa || g.px32 || g.py32) {
    const int G = g.G, N = G - 1;
    const double step = 1.0 / (double)N;
    R* o_px = (R*)g.px; R* o_py = (R*)g.py; R* o_k = (R*)g.kappa;
    float* o_px32 = g.px32; float* o_py32 = g.py32;  // float records beside (or instead of) the R arrays
    int sg = 0;
    for (int it = tid; it < PBa * G; it += T) {
      const int i = it / PBa, pl = it - i * PBa;
      const size_t oo = (size_t)i * P + p0 + pl;
      if (bad[pl]) {
        if (o_px) o_px[oo] = 0; if (o_py) o_py[oo] = 0; if (o_k) o_k[oo] = 0;
        if (o_px32) o_px32[oo] = 0; if (o_py32) o_py32[oo] = 0;
        continue;
      }
      R bx, by, kap;
      gridpoint_geometry<R, DEG>(i, N, step, segend + pl * nseg, nseg, hqc + pl * 4, invlen + pl * nseg, sg,
                                 coarse + (size_t)pl * nseg * kCoarseStride, kCoarseStride, wtab + (size_t)pl * nt,
                                 ax + (size_t)pl * nseg * NC, ay + (size_t)pl * nseg * NC, hx + (size_t)pl * nseg * DEG,
                                 hy + (size_t)pl * nseg * DEG, bx, by, kap);
      if (o_px) o_px[oo] = bx;
      if (o_py) o_py[oo] = by;
      if (o_k) o_k[oo] = kap;
      if (o_px32) o_px32[oo] = (float)bx;
      if (o_py32) o_py32[oo] = (float)by;
    }
  }
}

// paths per CTA: as many (up to 8) as keep three CTAs on an SM with the tables in shared memory; long paths
// keep their tables in global memory (8, 4, 2 or 1 paths per CTA by the size of the power bases); 0 = too long
int arc_paths_per_cta(int W, int spline_mode, int real_size, size_t& smem, bool* table_in_smem) {
  const int deg = spline_mode == SCB_REF_CUBIC ? 3 : 5;
  for (int pb = 8; pb >= 1; --pb) {
    smem = arc_smem_bytes(W, deg, pb, real_size, true);
    if (smem <= 74 * 1024) { if (table_in_smem) *table_in_smem = true; return pb; }
  }
  if (table_in_smem) *table_in_smem = false;
  for (int pb = 8; pb >= 1; pb >>= 1) {
    smem = arc_smem_bytes(W, deg, pb, real_size, false);
    if (smem <= (pb > 1 ? 40 * 1024 : 200 * 1024)) return pb;
  }
  return 0;
}

template <class R, int DEG, bool TS> static void launch_arc_t(const PathArgs& a, int grid, size_t smem, cudaStream_t st) {
  cudaFuncSetAttribute(arc_kernel<R, DEG, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (TS) cudaFuncSetAttribute(arc_kernel<R, DEG, TS>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  arc_kernel<R, DEG, TS><<<grid, kArcThreads, smem, st>>>(a);
}

template <class R> void launch_arc(const PathArgs& a0, int spline_mode, cudaStream_t st) {
  PathArgs a = a0;
  size_t smem; bool ts;
  const int PB = arc_paths_per_cta(a.W, spline_mode, (int)sizeof(R), smem, &ts);
  if (!PB) return;
  a.PB = PB;
  const int grid = (a.P + PB - 1) / PB;
  if (spline_mode == SCB_REF_CUBIC) { if (ts) launch_arc_t<R, 3, true>(a, grid, smem, st); else launch_arc_t<R, 3, false>(a, grid, smem, st); }
  else { if (ts) launch_arc_t<R, 5, true>(a, grid, smem, st); else launch_arc_t<R, 5, false>(a, grid, smem, st); }
}
template void launch_arc<double>(const PathArgs&, int, cudaStream_t);
template void launch_arc<float>(const PathArgs&, int, cudaStream_t);

}  // namespace scb
