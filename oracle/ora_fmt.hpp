// ORACLE (test infrastructure, never shipped) — CPU restatement of the reference's global planner front end:
// Halton sampling of the free space and FMT* over the sampled nodes.
//   halton                      sea_current.hpp:100-132
//   intersects                  sea_current.hpp:142-178
//   obstacle::contains          sea_current.hpp:200-250
//   planning_space::sample_free sea_current.hpp:1294-1313
//   planning_space::cost / near sea_current.hpp:1315-1337
//   fast_marching_trees         sea_current.hpp:1339-1407
// Arithmetic follows the reference's types: float cross products and sums, and pt_dist (sc:87-89) in the
// double precision `std::pow(float, int)` promotes it to, rounded to float on return.  Built with
// -ffp-contract=off (oracle/Makefile).
//
// Parity unpinned on one point: the reference walks std::unordered_set<Vector2f> in hash order, so when two
// candidates have EXACTLY equal float cost the one it keeps depends on libstdc++'s bucket layout.  Here (and in
// the CUDA kernel) the lowest node index wins.  Nodes are indexed 0 = x_init, 1..n = samples, n+1 = x_goal.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <vector>

namespace ora {

struct HaltonState { int f = 0, i = 0; };

// sc:100-132.  `std::floor(i / b)` there is an integer division followed by a no-op floor.
inline std::vector<float> halton(int b, int n, HaltonState& st) {
  std::vector<float> nums(n > 0 ? n : 0);
  int f = 0, i = 1;
  if (st.i != 0 && st.f != 0) { i = st.i; f = st.f; }
  for (int j = 0; j < n; ++j) {
    const int x = i - f;
    if (x == 1) { f = 1; i *= b; }
    else {
      int y = i / b;
      while (x <= y) y = y / b;
      f = (b + 1) * y - x;
    }
    nums[j] = static_cast<float>(f) / i;
  }
  st.f = f; st.i = i;
  return nums;
}

struct P2 { float x, y; };

inline float fmt_pt_dist(P2 u, P2 v) {  // sc:87-89
  const double dx = (double)(v.x - u.x), dy = (double)(v.y - u.y);
  return (float)std::sqrt(dx * dx + dy * dy);
}

// sc:142-178; the colinear and parallel branches both answer "no".
inline bool fmt_intersects(P2 l0, P2 l1, P2 k0, P2 k1, P2* at = nullptr) {
  const float ex = k0.x - l0.x, ey = k0.y - l0.y, dlx = l1.x - l0.x, dly = l1.y - l0.y, dkx = k1.x - k0.x, dky = k1.y - k0.y;
  const float a = ex * dly - ey * dlx;
  const float b = dlx * dky - dly * dkx;
  if (b == 0) return false;
  const float u = a / b;
  const float c = ex * dky - ey * dkx;
  const float t = c / b;
  if (0 <= u && u <= 1 && 0 <= t && t <= 1) {
    if (at) { at->x = t * dlx + l0.x; at->y = t * dly + l0.y; }
    return true;
  }
  return false;
}

struct FmtObstacle {  // sc:193-284: what contains() reads
  std::vector<P2> vertices;
  std::vector<float> lines;  // [n][4]
  float x_max, x_min, y_max, y_min;
  bool closed;
};

inline bool obstacle_contains(const FmtObstacle& o, P2 p) {  // sc:200-250
  if (!(p.x <= o.x_max && p.x >= o.x_min && p.y <= o.y_max && p.y >= o.y_min)) return false;
  const size_t nl = o.lines.size() / 4;
  if (o.closed) {
    const P2 outside{o.x_max + 1, o.y_max + 1};
    int count = 0;
    std::vector<bool> corners(o.vertices.size());
    for (size_t l = 0; l < nl; ++l) {
      P2 ip;
      if (!fmt_intersects(outside, p, P2{o.lines[4 * l], o.lines[4 * l + 1]}, P2{o.lines[4 * l + 2], o.lines[4 * l + 3]}, &ip)) continue;
      bool counted = false;
      for (size_t i = 0; i < o.vertices.size(); ++i) {
        // the reference compares float differences with the double literal 0.00001
        if ((double)std::fabs(ip.x - o.vertices[i].x) <= 0.00001 && (double)std::fabs(ip.y - o.vertices[i].y) <= 0.00001) {
          if (corners[i]) { counted = true; continue; }
          corners[i] = true; count++; counted = true;
          break;
        }
      }
      if (!counted) count++;
    }
    return count % 2 == 1;
  }
  for (size_t l = 0; l < nl; ++l) {  // open obstacle: within 0.01 of a line (dist_pt_line, sc:181-190)
    const float x0 = p.x, y0 = p.y, x1 = o.lines[4 * l], y1 = o.lines[4 * l + 1], x2 = o.lines[4 * l + 2], y2 = o.lines[4 * l + 3];
    const float nrm = std::sqrt((x2 - x1) * (x2 - x1) + (y2 - y1) * (y2 - y1));
    const float dist = std::fabs((x2 - x1) * (y1 - y0) - (x1 - x0) * (y2 - y1)) / nrm;
    if (dist < 0.01f) return true;
  }
  return false;
}

// sc:1294-1313 with the free-space allocations given as rectangles (x_min, x_max, y_min, y_max); the point set starts
// as {(0, 0)} like the reference's.  Returns points in insertion order.  max_rounds guards the reference's
// `while (n > pts.size())`, which never ends when nothing is free.
inline std::vector<P2> sample_free(int n, const float rect[4], const std::vector<FmtObstacle>& obs,
                                   const std::vector<float>& free_rects, HaltonState& xs, HaltonState& ys, int max_rounds = 64) {
  std::vector<P2> pts{P2{0, 0}};
  for (int round = 0; (int)pts.size() < n && round < max_rounds; ++round) {
    const int m = n - (int)pts.size();
    const std::vector<float> xt = halton(2, m, xs), yt = halton(3, m, ys);
    for (int i = 0; i < m; ++i) {
      const P2 t{(rect[1] - rect[0]) * xt[i] + rect[0], (rect[3] - rect[2]) * yt[i] + rect[2]};
      bool in_obs = false, allocated = false;
      for (const auto& o : obs) if (obstacle_contains(o, t)) { in_obs = true; break; }
      for (size_t k = 0; k + 3 < free_rects.size(); k += 4)
        if (t.x >= free_rects[k] && t.x <= free_rects[k + 1] && t.y >= free_rects[k + 2] && t.y <= free_rects[k + 3]) { allocated = true; break; }
      if (in_obs || !allocated) continue;
      bool dup = false;
      for (const auto& q : pts) if (q.x == t.x && q.y == t.y) { dup = true; break; }
      if (!dup) pts.push_back(t);
    }
  }
  return pts;
}

struct FmtResult { int status; std::vector<int> path; float cost; int closed; };  // status 0 = path, 1 = std::nullopt

// sc:1339-1407, loop for loop: every candidate parent is costed INCLUDING its collision test (sc:1365-1369), strict
// `<` keeps the earlier candidate, a node is connected only if the edge to the chosen parent is collision-free.
// The CUDA kernel prunes the collision tests (a candidate that cannot beat the best so far is never tested); this
// literal form is what it is checked against.
inline FmtResult fmt_star(const std::vector<P2>& samples, const std::vector<float>& edges, P2 init, P2 goal, float rn) {
  const float inf = FLT_MAX;
  const int n = (int)samples.size(), nn = n + 2, GOAL = n + 1;
  std::vector<P2> node(nn);
  node[0] = init; node[GOAL] = goal;
  for (int i = 0; i < n; ++i) node[i + 1] = samples[i];
  enum { UNVISITED = 0, OPEN = 1, CLOSED = 2, ABSENT = 4 };
  std::vector<uint8_t> state(nn, UNVISITED);
  state[0] = OPEN;
  for (int i = 1; i <= n; ++i)  // set semantics: a sample equal to x_goal is the same element; one equal to x_init is dropped
    if ((node[i].x == goal.x && node[i].y == goal.y) || (node[i].x == init.x && node[i].y == init.y)) state[i] = ABSENT;
  std::vector<float> cost(nn, inf);
  std::vector<int> parent(nn, -1);
  cost[0] = 0; parent[0] = 0;
  FmtResult r{1, {}, 0.f, 0};
  if (init.x == goal.x && init.y == goal.y) { r.status = 0; r.path = {0}; return r; }
  const double rn2 = std::pow((double)rn, 2);  // near(): pt_dist <= std::pow(dist, 2), sc:1331
  auto near_ = [&](int a, int b) { return (double)fmt_pt_dist(node[a], node[b]) <= rn2; };
  auto edge_cost = [&](int a, int b) -> float {  // sc:1315-1326
    for (size_t k = 0; k + 3 < edges.size(); k += 4)
      if (fmt_intersects(node[a], node[b], P2{edges[k], edges[k + 1]}, P2{edges[k + 2], edges[k + 3]})) return inf;
    return fmt_pt_dist(node[a], node[b]);
  };
  int z = 0;
  while (z != GOAL) {
    std::vector<int> open_new;
    for (int x = 0; x < nn; ++x) {
      if (state[x] != UNVISITED || !near_(x, z)) continue;
      int y_min = -1;
      float c_min = 0;
      for (int y = 0; y < nn; ++y) {
        if (state[y] != OPEN || !near_(y, x)) continue;
        const float c = cost[y] + edge_cost(x, y);
        if (y_min < 0 || c < c_min) { y_min = y; c_min = c; }
      }
      if (y_min < 0) continue;
      const float e = edge_cost(x, y_min);
      if (e != inf) { parent[x] = y_min; cost[x] = cost[y_min] + e; open_new.push_back(x); }
    }
    for (int x : open_new) state[x] = OPEN;  // V_open.merge(V_open_new): after the sweep, sc:1381
    state[z] = CLOSED; ++r.closed;
    int zn = -1;
    for (int y = 0; y < nn; ++y)
      if (state[y] == OPEN && (zn < 0 || cost[y] < cost[zn])) zn = y;
    if (zn < 0) return r;
    z = zn;
  }
  r.status = 0; r.cost = cost[GOAL];
  for (int p = GOAL; p != 0; p = parent[p]) r.path.push_back(p);
  r.path.push_back(0);
  for (size_t a = 0, b = r.path.size() - 1; a < b; ++a, --b) { int t = r.path[a]; r.path[a] = r.path[b]; r.path[b] = t; }
  return r;
}

}  // namespace ora
