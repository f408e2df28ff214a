// BASELINE.json config 4 from C++: P randomized planning problems (W = 10, G = 500) planned on every GPU of the box
// through include/sea_current_b200.hpp (b200::multi_gpu_planner -> scb_multi_*), the way an MPC outer loop uses it:
// summaries of all problems, then the trajectories of the k best.  Prints one JSON line.
//   multi_gpu_c4 <n_devices> <P> [repeats]
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <vector>

#include "sea_current_b200.hpp"

namespace b200 = turtle::sc::b200;

static double u01(uint64_t& s) {  // SplitMix64
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return (double)((z ^ (z >> 31)) >> 11) * (1.0 / 9007199254740992.0);
}

template <class T> static T* pinned(size_t n) {  // pinned host memory: the D2H / H2D copies run at PCIe speed
  void* p = scb_alloc_pinned(n * sizeof(T));
  if (!p) { std::fprintf(stderr, "scb_alloc_pinned failed\n"); std::exit(2); }
  return (T*)p;
}

int main(int argc, char** argv) {
  const int R = argc > 1 ? std::atoi(argv[1]) : 1, P = argc > 2 ? std::atoi(argv[2]) : 1 << 16, reps = argc > 3 ? std::atoi(argv[3]) : 3;
  const int W = 10, G = 500, K = 16;
  double *wx = pinned<double>((size_t)P * W), *wy = pinned<double>((size_t)P * W);
  double *vlo = pinned<double>(P), *vhi = pinned<double>(P), *alo = pinned<double>(P), *ahi = pinned<double>(P);
  uint64_t seed = 0x5EAC0DE4ull;
  const double pi = 3.14159265358979323846;
  for (int p = 0; p < P; ++p) {  // legs U(2, 10) m, heading change U(-100, 100) deg, vel U(0.5, 3), acc U(0.25, 4)
    double x = 0, y = 0, th = (2 * u01(seed) - 1) * pi;
    wx[(size_t)p * W] = 0; wy[(size_t)p * W] = 0;
    for (int i = 1; i < W; ++i) {
      if (i > 1) th += (2 * u01(seed) - 1) * (100.0 * pi / 180.0);
      const double leg = 2 + 8 * u01(seed);
      x += leg * std::cos(th); y += leg * std::sin(th);
      wx[(size_t)p * W + i] = x; wy[(size_t)p * W + i] = y;
    }
    vhi[p] = 0.5 + 2.5 * u01(seed); vlo[p] = -vhi[p];
    ahi[p] = 0.25 + 3.75 * u01(seed); alo[p] = -ahi[p];
    if (u01(seed) < 0.01) alo[p] = 0.1 * ahi[p];  // cannot brake: ERR_FAIL_CONTROLLABLE
  }
  scb_batch_in in{};
  in.P = P; in.W = W; in.G = G; in.wx = wx; in.wy = wy; in.vel_lo = vlo; in.vel_hi = vhi; in.acc_lo = alo; in.acc_hi = ahi;
  scb_default_opt(&in.opt);
  in.opt.move_waypoints = 1;
  scb_summary* summ = pinned<scb_summary>(P);
  int32_t lo, hi;
  scb_shard_range(P, 0, R, &lo, &hi);
  b200::multi_gpu_planner pl(R, SCB_FP64, SCB_LAU_QUINTIC, hi - lo, W, G);
  std::vector<int32_t> ids(P), st(K);
  std::vector<float> t((size_t)G * K), px((size_t)G * K), py((size_t)G * K), v((size_t)G * K), om((size_t)G * K);
  std::vector<double> T(K);
  double best = 1e300;
  int failed = 0;
  for (int rep = 0; rep < reps + 1; ++rep) {  // first pass = warm-up (allocations)
    const auto t0 = std::chrono::steady_clock::now();
    pl.plan_summary(in, true, summ);
    std::iota(ids.begin(), ids.end(), 0);
    auto key = [&](int32_t i) { return summ[i].status == 0 ? summ[i].T_total : 1e300; };
    auto less = [&](int32_t a, int32_t b) { return key(a) < key(b); };
    std::nth_element(ids.begin(), ids.begin() + K, ids.end(), less);   // O(P): the K fastest, then those in order
    std::sort(ids.begin(), ids.begin() + K, less);
    scb_batch_out out{};
    out.status = st.data(); out.T_total = T.data(); out.t = t.data(); out.px = px.data(); out.py = py.data(); out.v = v.data(); out.omega = om.data();
    pl.fetch(K, ids.data(), out);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (rep > 0 && sec < best) best = sec;
  }
  for (int p = 0; p < P; ++p) failed += summ[p].status != 0;
  bool ok = true;
  for (int k = 0; k < K; ++k) ok = ok && st[k] == 0 && T[k] == summ[ids[k]].T_total && std::fabs((double)t[(size_t)(G - 1) * K + k] - T[k]) <= 1e-6 * T[k];
  std::printf("{\"devices\": %d, \"P\": %d, \"W\": %d, \"G\": %d, \"seconds\": %.6f, \"trajectories_per_s\": %.1f, \"failed\": %d, \"best_T\": %.9g, "
              "\"fetched_consistent\": %s}\n", pl.devices(), P, W, G, best, P / best, failed, T[0], ok ? "true" : "false");
  return ok ? 0 : 1;
}
