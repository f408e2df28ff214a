// FMT* over a shared set of sampled nodes, one CTA per planning query (sm_100a).  SURVEY 8f N4.
//
// Reference: planning_space::fast_marching_trees (sea_current.hpp:1339-1407) with cost (sc:1315-1326), near
// (sc:1328-1337), intersects (sc:142-178) and pt_dist (sc:87-89).  The reference keeps V_open / V_unvisited /
// V_closed as std::unordered_set<Vector2f> and the costs and parents in hash maps keyed by the point; here a
// node is an index (0 = x_init, 1..n = the shared samples, n+1 = x_goal) and the three sets are one byte of state
// per node in shared memory, next to the node's coordinates, cost-to-come and parent, plus V_open as a compact
// index list, so one query's whole search state lives in one SM (21 bytes per node, up to 8,192 samples) and a
// batch of queries over the same map (MPC goals, multi-robot, replans) fills the GPU with one CTA per query.
//
// One expansion of the lowest-cost open node z:
//   (1) X_near = unvisited nodes within the radius of z         -> compacted into a list (block-wide)
//   (2) per x in X_near, ONE WARP scans the open nodes within the radius of x, lanes striding over the open list
//       (the frontier: far fewer entries than nodes).  Per 32 entries the lanes cost their own candidate
//       cost[y] + |xy|, then the warp tests the candidates that would beat its best so far cheapest first, one
//       obstacle edge per lane; the first collision-free one ends the round (the reference costs every candidate,
//       collision test included: a colliding one evaluates to FLT_MAX and can never win, so the argmin is the
//       same).  Ties: lowest index -> parent[x], cost[x]
//   (3) the newly connected nodes become open only after the sweep (V_open.merge, sc:1381), z is closed,
//       block-wide argmin over the open set picks the next z.
// Float arithmetic is the reference's, operation for operation: this file is compiled with --fmad=false
// (Makefile) so the cross products and the double-precision distance are not contracted, and results are
// bit-identical to the CPU restatement the tests hold.  Exact cost ties go to the lowest node index (the reference: hash order).
#include <cfloat>

#include "scb_common.cuh"

namespace scb {

namespace {

constexpr int kFmtThreads = 256;
constexpr int kFmtCoopMax = 4;  // up to this many candidates per 32 open nodes are collision-tested cooperatively, cheapest first
enum : unsigned char { FMT_UNVISITED = 0, FMT_OPEN = 1, FMT_CLOSED = 2, FMT_OPEN_NEW = 3, FMT_ABSENT = 4 };

__device__ __forceinline__ float fmt_dist(float ax, float ay, float bx, float by) {  // sc:87-89 (pow(float, int) is double)
  const double dx = (double)(bx - ax), dy = (double)(by - ay);
  return (float)sqrt(dx * dx + dy * dy);
}

// Float screen in front of the exact test `(float)sqrt(dx^2 + dy^2 in double) <= rn^2`: with s = fl(dx^2 + dy^2)
// (relative error <= 2.4e-7) and r = rn^2, s > r^2 (1 + 1e-5) is certainly far and s < r^2 (1 - 1e-5) certainly
// near; only the sliver between goes through the double-precision square root, so the decision is the exact one.
// 0 = far, 2 = near, 1 = undecided.
__device__ __forceinline__ int fmt_near_screen(float ax, float ay, float bx, float by, float lo, float hi) {
  const float dx = bx - ax, dy = by - ay, s = dx * dx + dy * dy;
  return s > hi ? 0 : (s < lo ? 2 : 1);
}

// sc:142-178 with l = (x, y) and k = the obstacle edge; colinear / parallel -> false
__device__ __forceinline__ bool fmt_hits(float l0x, float l0y, float l1x, float l1y, const float* __restrict__ e) {
  const float k0x = e[0], k0y = e[1], k1x = e[2], k1y = e[3];
  const float ex = k0x - l0x, ey = k0y - l0y, dlx = l1x - l0x, dly = l1y - l0y, dkx = k1x - k0x, dky = k1y - k0y;
  const float a = ex * dly - ey * dlx;
  const float b = dlx * dky - dly * dkx;
  if (b == 0) return false;
  const float u = a / b;
  const float c = ex * dky - ey * dkx;
  const float t = c / b;
  return 0 <= u && u <= 1 && 0 <= t && t <= 1;
}

__global__ void __launch_bounds__(kFmtThreads) fmt_kernel(FmtArgs g) {
  extern __shared__ unsigned char smem_raw[];
  const int n = g.n, nn = n + 2, GOAL = n + 1, q = blockIdx.x;
  float* X = (float*)smem_raw;
  float* Y = X + nn;
  float* cost = Y + nn;
  typedef unsigned short idx_t;  // nn <= 8,194
  idx_t* parent = (idx_t*)(cost + nn);
  idx_t* xlist = parent + nn;    // X_near of this expansion
  idx_t* ol = xlist + nn;        // V_open as a compact list (ping-pong with nl)
  idx_t* nl = ol + nn;
  unsigned char* state = (unsigned char*)(nl + nn);
  const float* edges = g.edges;  // the obstacle edge list: shared memory when it fits (launch_fmt), else global / L1
  if (g.edges_in_smem) {
    float* se = (float*)(smem_raw + (((size_t)nn * 21 + 15) & ~(size_t)15));
    for (int i = threadIdx.x; i < 4 * g.E; i += kFmtThreads) se[i] = g.edges[i];
    edges = se;
  }
  __shared__ int s_count, s_ocount, s_ncount, s_z;
  __shared__ float s_red_c[kFmtThreads / 32];
  __shared__ int s_red_i[kFmtThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = kFmtThreads / 32;
  const float ix = g.ix[q], iy = g.iy[q], gx = g.gx[q], gy = g.gy[q];
  for (int i = tid; i < nn; i += kFmtThreads) {
    float x, y;
    if (i == 0) { x = ix; y = iy; } else if (i == GOAL) { x = gx; y = gy; } else { x = g.nx[i - 1]; y = g.ny[i - 1]; }
    X[i] = x; Y[i] = y; cost[i] = FLT_MAX; parent[i] = 0;
    unsigned char s = FMT_UNVISITED;
    if (i == 0) s = FMT_OPEN;
    else if (i != GOAL && ((x == gx && y == gy) || (x == ix && y == iy))) s = FMT_ABSENT;  // set semantics
    state[i] = s;
  }
  if (tid == 0) { s_z = 0; s_count = 0; s_ocount = 1; s_ncount = 0; ol[0] = 0; }
  __syncthreads();
  if (tid == 0) cost[0] = 0.f;
  const double rn2 = (double)g.rn * (double)g.rn;  // near(): pt_dist <= std::pow(dist, 2), sc:1331
  const float near_lo = (float)(rn2 * rn2 * (1.0 - 1e-5)), near_hi = (float)(rn2 * rn2 * (1.0 + 1e-5));
  int closed = 0, status = 1;
  if (ix == gx && iy == gy) {  // while (z != x_goal) never runs: the path is {x_init}
    if (tid == 0) {
      g.status[q] = 0; g.path_len[q] = 1; g.cost[q] = 0.f;
      if (g.closed) g.closed[q] = 0;
      if (g.max_path >= 1 && g.path_x && g.path_y) { g.path_x[(size_t)q * g.max_path] = ix; g.path_y[(size_t)q * g.max_path] = iy; }
    }
    return;
  }
  __syncthreads();
  int z = 0, ocount = 1;
  while (true) {
    // (1) X_near
    const float zx = X[z], zy = Y[z];
    for (int i = tid; i < nn; i += kFmtThreads) {
      if (state[i] != FMT_UNVISITED) continue;
      const int cls = fmt_near_screen(X[i], Y[i], zx, zy, near_lo, near_hi);
      if (cls == 2 || (cls == 1 && (double)fmt_dist(X[i], Y[i], zx, zy) <= rn2)) xlist[atomicAdd(&s_count, 1)] = (idx_t)i;
    }
    __syncthreads();
    const int count = s_count;
    // (2) one warp per x, lanes over the open list (its order is arbitrary: ties are resolved by index explicitly).
    //     (wbest, wy) = the cheapest collision-free candidate so far, warp-uniform.  Per 32 list entries the lanes
    //     cost their own entry, then the warp tests the candidates that would beat (wbest, wy) cheapest first,
    //     the obstacle edges spread over the lanes; the first free one ends the round (the rest are dearer).
    //     With more than kFmtCoopMax candidates the lanes test their own edges in parallel instead.
    for (int li = warp; li < count; li += nwarps) {
      const int x = xlist[li];
      const float xx = X[x], xy = Y[x];
      float wbest = FLT_MAX;
      int wy = -1;
      for (int k0 = 0; k0 < ocount; k0 += 32) {
        const int k = k0 + lane;
        float c = FLT_MAX, yx = 0.f, yy = 0.f;
        int y = -1;
        if (k < ocount) {
          y = ol[k]; yx = X[y]; yy = Y[y];
          const int cls = fmt_near_screen(xx, xy, yx, yy, near_lo, near_hi);
          if (cls == 0) y = -1;
          else {
            const float d = fmt_dist(xx, xy, yx, yy);  // pt_dist is symmetric bit for bit ((-d)^2 == d^2)
            if (cls == 1 && !((double)d <= rn2)) y = -1; else c = cost[y] + d;
          }
        }
        bool cand = y >= 0 && (c < wbest || (c == wbest && y < wy));
        if (__popc(__ballot_sync(0xffffffffu, cand)) > kFmtCoopMax) {
          // many candidates (a dense neighbourhood): every lane tests its own edge against the obstacles, then one
          // argmin over the free ones
          if (cand) {
            bool hit = false;
            for (int e = 0; e < g.E && !hit; ++e) hit = fmt_hits(xx, xy, yx, yy, edges + 4 * e);
            cand = !hit;
          }
          float mc = cand ? c : FLT_MAX;
          int my = cand ? y : 0x7fffffff;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(0xffffffffu, mc, o);
            const int oy = __shfl_xor_sync(0xffffffffu, my, o);
            if (oc < mc || (oc == mc && oy < my)) { mc = oc; my = oy; }
          }
          if (my != 0x7fffffff) { wbest = mc; wy = my; }
          continue;
        }
        while (__any_sync(0xffffffffu, cand)) {
          float mc = cand ? c : FLT_MAX;
          int my = cand ? y : 0x7fffffff;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const float oc = __shfl_xor_sync(0xffffffffu, mc, o);
            const int oy = __shfl_xor_sync(0xffffffffu, my, o);
            if (oc < mc || (oc == mc && oy < my)) { mc = oc; my = oy; }
          }
          const int src = __ffs(__ballot_sync(0xffffffffu, cand && y == my)) - 1;
          const float tx = __shfl_sync(0xffffffffu, yx, src), ty = __shfl_sync(0xffffffffu, yy, src);
          bool hit = false;
          for (int e = lane; e < g.E; e += 32) hit = hit || fmt_hits(xx, xy, tx, ty, edges + 4 * e);
          hit = __any_sync(0xffffffffu, hit);
          if (lane == src) cand = false;
          if (!hit) { wbest = mc; wy = my; break; }
        }
      }
      if (lane == 0 && wy >= 0) { parent[x] = (idx_t)wy; cost[x] = wbest; state[x] = FMT_OPEN_NEW; }
    }
    __syncthreads();
    // (3) V_open <- V_open + newly connected - z (rebuilt into the other list), next z = argmin cost over it
    float bc = FLT_MAX;
    int bi = -1;
    for (int k = tid; k < ocount; k += kFmtThreads) {
      const int y = ol[k];
      if (y == z) continue;
      nl[atomicAdd(&s_ncount, 1)] = (idx_t)y;
      const float c = cost[y];
      if (bi < 0 || c < bc || (c == bc && y < bi)) { bc = c; bi = y; }
    }
    for (int k = tid; k < count; k += kFmtThreads) {
      const int x = xlist[k];
      if (state[x] != FMT_OPEN_NEW) continue;
      state[x] = FMT_OPEN;
      nl[atomicAdd(&s_ncount, 1)] = (idx_t)x;
      const float c = cost[x];
      if (bi < 0 || c < bc || (c == bc && x < bi)) { bc = c; bi = x; }
    }
    if (tid == 0) state[z] = FMT_CLOSED;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float oc = __shfl_xor_sync(0xffffffffu, bc, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi >= 0 && (bi < 0 || oc < bc || (oc == bc && oi < bi))) { bc = oc; bi = oi; }
    }
    if (lane == 0) { s_red_c[warp] = bc; s_red_i[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      float c = s_red_c[0]; int i = s_red_i[0];
      for (int w = 1; w < nwarps; ++w) {
        const float oc = s_red_c[w]; const int oi = s_red_i[w];
        if (oi >= 0 && (i < 0 || oc < c || (oc == c && oi < i))) { c = oc; i = oi; }
      }
      s_z = i; s_count = 0; s_ocount = s_ncount; s_ncount = 0;
    }
    __syncthreads();
    ++closed;
    z = s_z; ocount = s_ocount;
    { idx_t* t = ol; ol = nl; nl = t; }
    if (z < 0) break;                      // V_open is empty: std::nullopt
    if (z == GOAL) { status = 0; break; }
  }
  if (tid != 0) return;
  if (g.closed) g.closed[q] = closed;
  if (status != 0) { g.status[q] = 1; g.path_len[q] = 0; g.cost[q] = FLT_MAX; return; }
  int len = 1;
  for (int p = GOAL; p != 0; p = parent[p]) ++len;
  g.path_len[q] = len; g.cost[q] = cost[GOAL];
  if (len > g.max_path) { g.status[q] = 2; return; }  // the caller's path buffer is too short; nothing written
  g.status[q] = 0;
  if (!g.path_x || !g.path_y) return;
  float* px = g.path_x + (size_t)q * g.max_path; float* py = g.path_y + (size_t)q * g.max_path;
  int k = len - 1;
  for (int p = GOAL; p != 0; p = parent[p], --k) { px[k] = X[p]; py[k] = Y[p]; }
  px[0] = X[0]; py[0] = Y[0];
}

}  // namespace

size_t fmt_smem_bytes(int n) { return (((size_t)(n + 2) * 21 + 15) & ~(size_t)15); }  // 3 floats, 4 16-bit indices, 1 state byte per node

cudaError_t launch_fmt(const FmtArgs& a_, cudaStream_t st) {
  FmtArgs a = a_;
  size_t smem = fmt_smem_bytes(a.n);
  a.edges_in_smem = a.E > 0 && smem + (size_t)a.E * 16 <= 200 * 1024;
  if (a.edges_in_smem) smem += (size_t)a.E * 16;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(fmt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  fmt_kernel<<<a.Q, kFmtThreads, smem, st>>>(a);
  return cudaGetLastError();
}

}  // namespace scb
