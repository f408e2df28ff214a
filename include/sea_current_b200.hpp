// sea_current_b200.hpp — host-side drop-in for the planning hot path of sea_current.hpp.
//
// Keeps namespace turtle::sc and the call sequence a sea-current caller writes
// (examples/json_test.cpp:28-65, examples/zmq_test.cpp:64-92):
//
//     bezier_spline pad = bezier_spline::from_path(path, space);      // sc:402, sc:599-683
//     arclength_data ad = pad.arclength();                            // sc:413, sc:767-896
//     velocity_profile prof = gen_vel_prof<1>(pos_end, pos_start, vel_end, vel_start,
//                                             vel_lim, acc_min, acc_max);   // sc:1191-1265
//     bezier_spline re = pad.resample(prof.pos[0], ad, true);         // sc:419, sc:898-1005
//     std::vector<float> w = re.angular_velocity(prof);               // sc:416, sc:1055-1067
//     json j = serialize_path_to_json(re, prof, ad, w);               // sc:1423-1457
//
// Every numeric step is a batch-of-1 call into the C ABI of include/seacurrent_b200.h, i.e. a
// CUDA kernel launch on a B200; nothing is computed on the host and there is no CPU fallback:
// without libseacurrent_b200.so and an sm_100 device every entry point throws.  Batched callers
// should use turtle::sc::b200::batch_planner (or the C ABI directly) instead: one call plans
// thousands of paths.
//
// The reference's value types come from Eigen (Vector2f, VectorXf, Eigen::Vector<T, N>) and
// toppra (toppra::Vector).  Eigen is not part of this image, so this header carries small
// stand-ins with the members the call sequence uses; when <Eigen/Core> has been included first
// they are not defined and the real ones are used.
//
// Deliberate differences from the reference (DESIGN.md "Deviations"):
//   * bezier_spline::pts / positions are empty after from_path: the 10,001-point-per-segment
//     preview (sc:679) feeds nothing downstream.  resample() fills pts as in the reference.
//   * resample() inverts the arc-length table exactly instead of through the float Chebyshev
//     least-squares fit (sc:1109-1170); it does not print (sc:946-947, 999-1002).
//   * gen_vel_prof throws planning_error when toppra's verdict is not OK; the reference fetches
//     the ReturnCode and ignores it in Release builds (sc:1225-1227).
//   * serialize_path_to_json writes velocity[i] / acceleration[i] of sample i; the reference's
//     format_vec_vecx (sc:1410-1420) repeats element 0 (index slip).
//   * the header's free functions are inline (the reference's are not: ODR).
#ifndef SEA_CURRENT_B200_HPP
#define SEA_CURRENT_B200_HPP

#include <algorithm>
#include <array>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <initializer_list>
#include <limits>
#include <memory>
#include <optional>
#include <ostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <unordered_set>
#include <vector>

#include "seacurrent_b200.h"

// ------------------------------------------------------------------ Eigen / toppra stand-ins
#ifndef EIGEN_CORE_H
namespace Eigen {
struct Vector2f {
  float v[2];
  Vector2f(float x = 0, float y = 0) : v{x, y} {}
  float x() const { return v[0]; }
  float y() const { return v[1]; }
  float& operator()(int i) { return v[i]; }
  float operator()(int i) const { return v[i]; }
  float& operator[](int i) { return v[i]; }
  float operator[](int i) const { return v[i]; }
  bool operator==(const Vector2f& o) const { return v[0] == o.v[0] && v[1] == o.v[1]; }
  bool operator!=(const Vector2f& o) const { return !(*this == o); }
};
template <class T> struct VectorX {
  std::vector<T> d;
  VectorX() = default;
  explicit VectorX(long n) : d((size_t)n, T(0)) {}
  VectorX(const T* b, const T* e) : d(b, e) {}
  static VectorX Zero(long n) { return VectorX(n); }
  long rows() const { return (long)d.size(); }
  long size() const { return (long)d.size(); }
  T* data() { return d.data(); }
  const T* data() const { return d.data(); }
  T& operator()(long i, long = 0) { return d[(size_t)i]; }
  T operator()(long i, long = 0) const { return d[(size_t)i]; }
  T& operator[](long i) { return d[(size_t)i]; }
  T operator[](long i) const { return d[(size_t)i]; }
};
using VectorXf = VectorX<float>;
using VectorXd = VectorX<double>;
template <class T, int N> struct Vector {
  std::array<T, N> v{};
  Vector() = default;
  Vector(std::initializer_list<T> l) { int i = 0; for (T x : l) if (i < N) v[i++] = x; }
  int rows() const { return N; }
  T& operator()(int i) { return v[i]; }
  T operator()(int i) const { return v[i]; }
};
struct MatrixX2f {  // Matrix<float, Dynamic, 2>
  std::vector<float> cx, cy;
  long rows() const { return (long)cx.size(); }
  float operator()(long i, int j) const { return j == 0 ? cx[(size_t)i] : cy[(size_t)i]; }
};
}  // namespace Eigen
namespace toppra {
typedef double value_type;
typedef Eigen::VectorXd Vector;
}  // namespace toppra
#define SEA_CURRENT_B200_OWN_EIGEN 1
#endif

namespace turtle::sc {

using Eigen::Vector2f;
using Eigen::VectorXf;
using toppra::value_type;
#ifdef SEA_CURRENT_B200_OWN_EIGEN
using Eigen::MatrixX2f;
#else
struct MatrixX2f {
  std::vector<float> cx, cy;
  long rows() const { return (long)cx.size(); }
  float operator()(long i, int j) const { return j == 0 ? cx[(size_t)i] : cy[(size_t)i]; }
};
#endif

// toppra::ReturnCode of the failed plan (tp:algorithm.hpp:13-32) + the gridpoint it failed at
struct planning_error : std::runtime_error {
  int status, fail_stage;
  planning_error(int st, int fs, const std::string& what) : std::runtime_error(what), status(st), fail_stage(fs) {}
};

struct bounding_rect {  // sc:61-84
  float x_max, x_min, y_max, y_min;
  bounding_rect(float x_max, float x_min, float y_max, float y_min) : x_max(x_max), x_min(x_min), y_max(y_max), y_min(y_min) {}
  inline bool contains(const Vector2f& p) const { return p.x() <= x_max && p.x() >= x_min && p.y() <= y_max && p.y() >= y_min; }
  inline void enclose_point(const Vector2f& p) {
    if (p.x() > x_max) x_max = p.x();
    if (p.x() < x_min) x_min = p.x();
    if (p.y() > y_max) y_max = p.y();
    if (p.y() < y_min) y_min = p.y();
  }
};

// sc:87-89: std::pow(float, int) promotes to double, so the distance is formed in double and rounded on return
inline float pt_dist(const Vector2f& u, const Vector2f& v = Vector2f(0, 0)) {
  const double dx = (double)(v.x() - u.x()), dy = (double)(v.y() - u.y());
  return (float)std::sqrt(dx * dx + dy * dy);
}

struct halton_state {  // sc:93-97
  int f = 0, i = 0;
  halton_state(int f, int i) : f(f), i(i) {}
  halton_state() = default;
};
// sc:100-132: the next n elements of the base-b Halton (van der Corput) sequence; advances `state`
inline std::vector<float> halton(const int b, const int n, halton_state& state) {
  std::vector<float> nums((size_t)(n > 0 ? n : 0));
  int f = 0, i = 1;
  if (state.i != 0 && state.f != 0) { i = state.i; f = state.f; }
  for (int j = 0; j < n; ++j) {
    const int x = i - f;
    if (x == 1) { f = 1; i *= b; }
    else {
      int y = i / b;
      while (x <= y) y = y / b;
      f = (b + 1) * y - x;
    }
    nums[(size_t)j] = static_cast<float>(f) / i;
  }
  state.f = f; state.i = i;
  return nums;
}

inline float cross2d(const Vector2f& u, const Vector2f& v) { return u.x() * v.y() - u.y() * v.x(); }  // sc:137-139
// sc:142-178: do segments l and k cross, and where.  Colinear and parallel segments answer "no", as in the reference.
inline std::tuple<bool, Vector2f> intersects(const std::tuple<Vector2f, Vector2f>& l, const std::tuple<Vector2f, Vector2f>& k) {
  const Vector2f &l0 = std::get<0>(l), &l1 = std::get<1>(l), &k0 = std::get<0>(k), &k1 = std::get<1>(k);
  const Vector2f e(k0.x() - l0.x(), k0.y() - l0.y()), dl(l1.x() - l0.x(), l1.y() - l0.y()), dk(k1.x() - k0.x(), k1.y() - k0.y());
  const float a = cross2d(e, dl), b = cross2d(dl, dk);
  if (b == 0) return {false, Vector2f(0, 0)};
  const float u = a / b, c = cross2d(e, dk), t = c / b;
  if (0 <= u && u <= 1 && 0 <= t && t <= 1) return {true, Vector2f(t * dl.x() + l0.x(), t * dl.y() + l0.y())};
  return {false, Vector2f(0, 0)};
}
inline float dist_pt_line(const Vector2f& p1, const Vector2f& p2, const Vector2f a) {  // sc:181-190 (the infinite line)
  const float x0 = a.x(), y0 = a.y(), x1 = p1.x(), y1 = p1.y(), x2 = p2.x(), y2 = p2.y();
  return std::abs((x2 - x1) * (y1 - y0) - (x1 - x0) * (y2 - y1)) / std::sqrt((x2 - x1) * (x2 - x1) + (y2 - y1) * (y2 - y1));
}

// sc:193-284.  The path reads obstacle.lines (shrink_tangent, sc:575-596); `contains` and the bounding box serve
// sample_free.  `closed` is never set by the reference's constructors (an uninitialised read in contains); here it
// defaults to true (the even-odd test) and the caller clears it for a polyline obstacle.
struct obstacle {
  std::vector<std::tuple<Vector2f, Vector2f>> lines;
  std::vector<Vector2f> vertices;
  bounding_rect bound_rect = {0, 0, 0, 0};
  bool closed = true;
  int64_t id = 0;
  obstacle() = default;
  // sc:251-267: a polyline through the vertices.  Quirks kept: with an EVEN vertex count the reference appends the
  // first vertex to its constructor argument (not to the member), i.e. the edge list is closed, `vertices` is not;
  // the bounding box encloses every vertex of that local list but the last.
  obstacle(std::vector<Vector2f> vs) : vertices(vs) {
    if (vs.empty()) return;
    bound_rect = {vs[0].x(), vs[0].x(), vs[0].y(), vs[0].y()};
    if (vs.size() % 2 == 0) vs.push_back(vs[0]);
    for (size_t i = 0; i + 1 < vs.size(); ++i) { lines.emplace_back(vs[i], vs[i + 1]); bound_rect.enclose_point(vs[i]); }
  }
  // sc:269-281: explicit edge list over the vertices
  obstacle(std::vector<Vector2f> vs, const std::vector<std::tuple<int, int>>& edges) : vertices(std::move(vs)) {
    if (!vertices.empty()) bound_rect = {vertices[0].x(), vertices[0].x(), vertices[0].y(), vertices[0].y()};
    for (const auto& e : edges) {
      if (std::get<0>(e) < 0 || std::get<1>(e) < 0 || (size_t)std::get<0>(e) >= vertices.size() || (size_t)std::get<1>(e) >= vertices.size())
        throw std::invalid_argument("edge must reference indices within the vertex list");
      lines.emplace_back(vertices[(size_t)std::get<0>(e)], vertices[(size_t)std::get<1>(e)]);
      bound_rect.enclose_point(vertices[(size_t)std::get<0>(e)]);
      bound_rect.enclose_point(vertices[(size_t)std::get<1>(e)]);
    }
  }
  // sc:200-250: even-odd crossing count of the segment from outside the bounding box (a crossing that lands on a
  // vertex counts once per vertex), or, for an open obstacle, "within 0.01 of one of its lines"
  inline bool contains(const Vector2f& point) const {
    if (!bound_rect.contains(point)) return false;
    if (closed) {
      const Vector2f outside_pt(bound_rect.x_max + 1, bound_rect.y_max + 1);
      int count = 0;
      std::vector<bool> corners(vertices.size());
      for (const auto& line : lines) {
        const auto [ints, int_pt] = intersects({outside_pt, point}, line);
        if (!ints) continue;
        bool counted = false;
        for (size_t i = 0; i < vertices.size(); ++i) {
          if (std::abs(int_pt.x() - vertices[i].x()) <= 0.00001 && std::abs(int_pt.y() - vertices[i].y()) <= 0.00001) {
            if (corners[i]) { counted = true; continue; }
            corners[i] = true; count++; counted = true;
            break;
          }
        }
        if (!counted) count++;
      }
      return count % 2 == 1;
    }
    for (const auto& line : lines)
      if (dist_pt_line(std::get<0>(line), std::get<1>(line), point) < 0.01f) return true;
    return false;
  }
};

struct hash_vector2f {  // sc:286-295
  size_t operator()(const Vector2f v) const {
    float fa = v.x(), fb = v.y();
    int32_t a, b;
    std::memcpy(&a, &fa, 4); std::memcpy(&b, &fb, 4);
    return std::hash<int32_t>()(a) ^ std::hash<int32_t>()(b);
  }
};
using point_set = std::unordered_set<Vector2f, hash_vector2f>;

// sc:298-319.  from_path reads `.obstacles`; sample_free / fast_marching_trees are the global planner front end
// (SURVEY 8f N4): sampling stays on the host (the free-space allocations are std::function callbacks), the tree
// search runs on the GPU (scb_fmt_batch), one query per call here, many per call through b200::fmt_batch.
struct planning_space {
  std::vector<obstacle> obstacles;
  std::vector<std::function<bool(Vector2f)>> free_space_allocations;
  bounding_rect bound_rect;
  halton_state x_state, y_state;
  planning_space(const bounding_rect& br) : bound_rect(br) {}

  inline std::tuple<bool, obstacle> is_obstacle(const Vector2f& p) const {  // sc:1270-1276
    for (const auto& o : obstacles) if (o.contains(p)) return {true, o};
    return {false, obstacle()};
  }
  inline bool is_free_space_allocated(const Vector2f& p) const {  // sc:1278-1283
    for (const auto& f : free_space_allocations) if (f(p)) return true;
    return false;
  }
  inline bool is_free(const Vector2f& p) const {  // sc:1285-1292
    for (const auto& o : obstacles) if (o.contains(p)) return false;
    return is_free_space_allocated(p);
  }
  // sc:1294-1313 in insertion order ((0, 0) first, as the reference's set starts).  The reference's
  // `while (n > pts.size())` never ends when no candidate is free (e.g. nothing allocated); this one throws after
  // 64 rounds that added nothing.  It does not print (sc:1306).
  inline std::vector<Vector2f> sample_free_ordered(const int n) {
    std::vector<Vector2f> pts{Vector2f(0, 0)};
    point_set seen(pts.begin(), pts.end());
    int idle = 0;
    while (n > (int)pts.size()) {
      const int m = n - (int)pts.size();
      const std::vector<float> x_test = halton(2, m, x_state), y_test = halton(3, m, y_state);
      bool grew = false;
      for (int i = 0; i < m; ++i) {
        const Vector2f test((bound_rect.x_max - bound_rect.x_min) * x_test[(size_t)i] + bound_rect.x_min,
                            (bound_rect.y_max - bound_rect.y_min) * y_test[(size_t)i] + bound_rect.y_min);
        if (is_free(test) && seen.insert(test).second) { pts.push_back(test); grew = true; }
      }
      idle = grew ? 0 : idle + 1;
      if (idle >= 64) throw std::runtime_error("sample_free: no free point found (is any free space allocated?)");
    }
    return pts;
  }
  inline point_set sample_free(const int n) {
    const std::vector<Vector2f> pts = sample_free_ordered(n);
    return point_set(pts.begin(), pts.end());
  }
  inline float cost(const Vector2f a, const Vector2f b) const {  // sc:1315-1326
    for (const auto& o : obstacles)
      for (const auto& l : o.lines)
        if (std::get<0>(intersects({a, b}, l))) return std::numeric_limits<float>::max();
    return pt_dist(a, b);
  }
  inline point_set near(const Vector2f b, const point_set& nodes, const float dist) const {  // sc:1328-1337
    point_set out;
    for (const auto& a : nodes)
      if (pt_dist(a, b) <= std::pow(dist, 2) && a != b) out.insert(a);
    return out;
  }
  // sc:1339-1407: samples n free points (advancing the Halton states, like the reference) and searches on the GPU
  inline std::optional<std::vector<Vector2f>> fast_marching_trees(const Vector2f& x_init, const Vector2f& x_goal, const int n, const float rn);
};

struct arclength_data {  // sc:322-326
  float arclength = 0;
  std::vector<VectorXf> segments;
  std::vector<VectorXf> positions;
};

struct velocity_profile {  // sc:379-386
  std::vector<VectorXf> pos, vel, acc;
  toppra::Vector time;
  velocity_profile(std::vector<VectorXf> pos, std::vector<VectorXf> vel, std::vector<VectorXf> acc, toppra::Vector time)
      : pos(std::move(pos)), vel(std::move(vel)), acc(std::move(acc)), time(std::move(time)) {}
};

using vel_lim_func = std::function<std::tuple<toppra::Vector, toppra::Vector>(value_type time)>;  // sc:1175

namespace b200 {

// Spline construction used by the drop-in calls: the reference's cubic (default) or the Lau09
// quintic + optimiser.  Also settable with SEA_CURRENT_B200_SPLINE=lau in the environment.
inline int& spline_mode() {
  static int m = [] { const char* e = std::getenv("SEA_CURRENT_B200_SPLINE"); return (e && (e[0] == 'l' || e[0] == 'L')) ? SCB_LAU_QUINTIC : SCB_REF_CUBIC; }();
  return m;
}
// Parametrizer gen_vel_prof builds: toppra::parametrizer::Spline as in the reference (sc:1231), or ConstAccel
// (tp:parametrizer/const_accel.cpp).  Also settable with SEA_CURRENT_B200_PARAM=const_accel in the environment.
inline int& param_mode() {
  static int m = [] { const char* e = std::getenv("SEA_CURRENT_B200_PARAM"); return (e && (e[0] == 'c' || e[0] == 'C')) ? SCB_PARAM_CONST_ACCEL : SCB_PARAM_SPLINE; }();
  return m;
}
inline int& device() { static int d = 0; return d; }

inline void check(int rc, scb_ctx* ctx, const char* what) {
  if (rc == SCB_OK) return;
  std::string msg = std::string(what) + ": " + scb_strerror(rc);
  if (ctx && rc == SCB_E_CUDA) msg += std::string(" [") + scb_last_cuda_error(ctx) + "]";
  throw std::runtime_error(msg);
}

// RAII owner of one scb_ctx (FP64).  One per host thread and GPU.
class batch_planner {
 public:
  batch_planner(int device, int precision, int spline_mode, int max_problems, int max_waypoints, int max_gridpoints, int max_samples)
      : precision_(precision), mode_(spline_mode) {
    scb_config cfg{device, precision, spline_mode, max_problems, max_waypoints, max_gridpoints, max_samples};
    check(scb_create(&ctx_, &cfg), nullptr, "scb_create");
  }
  ~batch_planner() { if (ctx_) scb_destroy(ctx_); }
  batch_planner(const batch_planner&) = delete;
  batch_planner& operator=(const batch_planner&) = delete;
  scb_ctx* ctx() const { return ctx_; }
  int mode() const { return mode_; }
  int degree() const { return mode_ == SCB_REF_CUBIC ? 3 : 5; }
  // host buffers in, host buffers out; see seacurrent_b200.h for the layouts
  void plan(const scb_batch_in& in, const scb_batch_out& out) { check(scb_plan_batch(ctx_, &in, &out), ctx_, "scb_plan_batch"); }
  // rank, then fetch: the batch stays in HBM, the 32-byte summaries come back; fetch() then reads the
  // records of the few problems the caller keeps (host arrays laid out as scb_batch_out with P := n)
  void plan_summary(const scb_batch_in& in, bool record_f32, scb_summary* summaries) {
    check(scb_plan_batch_summary(ctx_, &in, record_f32 ? 1 : 0, summaries), ctx_, "scb_plan_batch_summary");
  }
  void fetch(int n, const int32_t* problem_ids, const scb_batch_out& out) { check(scb_fetch(ctx_, n, problem_ids, &out), ctx_, "scb_fetch"); }
  scb_timing timing() const { scb_timing t{}; scb_get_timing(ctx_, &t); return t; }

 private:
  scb_ctx* ctx_ = nullptr;
  int precision_, mode_;
};

// One process, several GPUs: contiguous blocks of ceil(P / R) problems, block r on device r with its own scb_ctx and
// host thread (scb_multi_*).  Problems share nothing, so there is no collective; the summaries land in the caller's
// array in problem order.  max_problems is the capacity PER DEVICE.
class multi_gpu_planner {
 public:
  multi_gpu_planner(int n_devices, int precision, int spline_mode, int max_problems, int max_waypoints, int max_gridpoints,
                    const int32_t* devices = nullptr) {
    scb_config cfg{0, precision, spline_mode, max_problems, max_waypoints, max_gridpoints, 0};
    check(scb_multi_create(&m_, &cfg, n_devices, devices), nullptr, "scb_multi_create");
  }
  ~multi_gpu_planner() { if (m_) scb_multi_destroy(m_); }
  multi_gpu_planner(const multi_gpu_planner&) = delete;
  multi_gpu_planner& operator=(const multi_gpu_planner&) = delete;
  int devices() const { return scb_multi_devices(m_); }
  void plan(const scb_batch_in& in, const scb_batch_out& out) { check(scb_multi_plan_batch(m_, &in, &out), nullptr, "scb_multi_plan_batch"); }
  void plan_summary(const scb_batch_in& in, bool record_f32, scb_summary* summaries) {
    check(scb_multi_plan_summary(m_, &in, record_f32 ? 1 : 0, summaries), nullptr, "scb_multi_plan_summary");
  }
  void fetch(int n, const int32_t* problem_ids, const scb_batch_out& out) { check(scb_multi_fetch(m_, n, problem_ids, &out), nullptr, "scb_multi_fetch"); }
  scb_timing timing(int r) const { scb_timing t{}; scb_multi_get_timing(m_, r, &t); return t; }

 private:
  scb_multi* m_ = nullptr;
};

constexpr int kDropInMaxWaypoints = 4096, kDropInGrid = 101 /* tp:algorithm.hpp:140 */, kDropInMaxSamples = 1 << 17;

// Per-thread context behind the reference-shaped calls.  spline_mode() and device() are mutable globals: the context
// is rebuilt when either differs from what it was created with, so the mode a call checks is the mode its kernels run.
inline batch_planner& context_for(std::unique_ptr<batch_planner>& slot, int& dev_of_slot, int max_problems) {
  if (!slot || slot->mode() != spline_mode() || dev_of_slot != device()) {
    slot.reset();
    slot = std::make_unique<batch_planner>(device(), SCB_FP64, spline_mode(), max_problems, kDropInMaxWaypoints, kDropInGrid, kDropInMaxSamples);
    dev_of_slot = device();
  }
  return *slot;
}
inline batch_planner& drop_in_planner() {  // batch-of-1
  static thread_local std::unique_ptr<batch_planner> pl;
  static thread_local int dev = -1;
  return context_for(pl, dev, 1);
}

// FMT* for many (x_init, x_goal) queries over ONE sampled map (scb_fmt_batch: one CTA per query).  An empty
// optional = the reference's std::nullopt (the open set ran empty).
inline std::vector<std::optional<std::vector<Vector2f>>> fmt_batch(const planning_space& space, const std::vector<Vector2f>& samples,
                                                                   const std::vector<std::tuple<Vector2f, Vector2f>>& queries, const float rn,
                                                                   std::vector<float>* costs = nullptr) {
  const int Q = (int)queries.size(), n = (int)samples.size();
  std::vector<float> nx((size_t)n), ny((size_t)n), edges, ix((size_t)Q), iy((size_t)Q), gx((size_t)Q), gy((size_t)Q);
  for (int i = 0; i < n; ++i) { nx[(size_t)i] = samples[(size_t)i].x(); ny[(size_t)i] = samples[(size_t)i].y(); }
  for (const auto& o : space.obstacles)
    for (const auto& l : o.lines) { edges.push_back(std::get<0>(l).x()); edges.push_back(std::get<0>(l).y()); edges.push_back(std::get<1>(l).x()); edges.push_back(std::get<1>(l).y()); }
  for (int q = 0; q < Q; ++q) {
    ix[(size_t)q] = std::get<0>(queries[(size_t)q]).x(); iy[(size_t)q] = std::get<0>(queries[(size_t)q]).y();
    gx[(size_t)q] = std::get<1>(queries[(size_t)q]).x(); gy[(size_t)q] = std::get<1>(queries[(size_t)q]).y();
  }
  const int M = n + 2;  // a path visits a node at most once
  std::vector<int32_t> status((size_t)Q), len((size_t)Q);
  std::vector<float> px((size_t)Q * M), py((size_t)Q * M), cost((size_t)Q);
  scb_fmt_in in{};
  in.Q = Q; in.n_nodes = n; in.n_edges = (int)(edges.size() / 4); in.max_path = M;
  in.node_x = nx.data(); in.node_y = ny.data(); in.edges = edges.empty() ? nullptr : edges.data();
  in.init_x = ix.data(); in.init_y = iy.data(); in.goal_x = gx.data(); in.goal_y = gy.data(); in.rn = rn;
  scb_fmt_out out{};
  out.status = status.data(); out.path_len = len.data(); out.path_x = px.data(); out.path_y = py.data(); out.cost = cost.data();
  batch_planner& pl = drop_in_planner();
  check(scb_fmt_batch(pl.ctx(), &in, &out), pl.ctx(), "scb_fmt_batch");
  std::vector<std::optional<std::vector<Vector2f>>> paths((size_t)Q);
  for (int q = 0; q < Q; ++q) {
    if (status[(size_t)q] != SCB_FMT_OK) continue;
    std::vector<Vector2f> p;
    for (int k = 0; k < len[(size_t)q]; ++k) p.emplace_back(px[(size_t)q * M + k], py[(size_t)q * M + k]);
    paths[(size_t)q] = std::move(p);
  }
  if (costs) *costs = cost;
  return paths;
}

}  // namespace b200

inline std::optional<std::vector<Vector2f>> planning_space::fast_marching_trees(const Vector2f& x_init, const Vector2f& x_goal, const int n,
                                                                                  const float rn) {
  const std::vector<Vector2f> samples = sample_free_ordered(n);
  return b200::fmt_batch(*this, samples, {{x_init, x_goal}}, rn)[0];
}

class bezier_spline {  // sc:388-431
 public:
  std::vector<std::vector<Vector2f>> ctrl_pts;
  MatrixX2f pts;
  std::vector<VectorXf> positions;

  bezier_spline() = default;

  inline int n_pts() const { return (int)pts.rows(); }
  inline int n_segments() const { return (int)ctrl_pts.size(); }
  inline int degree() const { return ctrl_pts.empty() ? 0 : (int)ctrl_pts[0].size() - 1; }

  // sc:599-683 (tangent heuristics sc:343-377, tangent clipping against ps.obstacles sc:575-596).
  static bezier_spline from_path(const std::vector<Vector2f>& path, const planning_space& ps,
                                 float start_angle = std::numeric_limits<float>::quiet_NaN()) {
    if (!ps.obstacles.empty() && b200::drop_in_planner().mode() != SCB_REF_CUBIC)
      throw std::invalid_argument("sea_current_b200: obstacle clipping is defined for the reference's cubic spline only");
    if (path.size() < 2) throw std::invalid_argument("sea_current_b200: a path needs at least two waypoints");
    bezier_spline s;
    s.wx_.resize(path.size()); s.wy_.resize(path.size());
    for (size_t i = 0; i < path.size(); ++i) { s.wx_[i] = path[i].x(); s.wy_[i] = path[i].y(); }
    s.start_angle_ = start_angle;
    for (const obstacle& ob : ps.obstacles)
      for (const auto& ln : ob.lines) {
        s.edges_.push_back(std::get<0>(ln).x()); s.edges_.push_back(std::get<0>(ln).y());
        s.edges_.push_back(std::get<1>(ln).x()); s.edges_.push_back(std::get<1>(ln).y());
      }
    s.build();
    return s;
  }

  // sc:767-896.  Only the reference's default precision exists on the device (100 sub-intervals).
  arclength_data arclength(const float precision = 0.01f) const {
    if (std::fabs(precision - 0.01f) > 1e-9f) throw std::invalid_argument("sea_current_b200: arclength precision is fixed at 0.01");
    return ad_;
  }

  // sc:898-1005.  Mutates profile_pos when nudge_positions is set, as the reference does.
  bezier_spline resample(VectorXf& profile_pos, arclength_data /*ad*/, bool nudge_positions = false) const {
    if (profile_pos.rows() <= 0) throw std::invalid_argument("The vector of positions to be sampled must not be empty");
    bezier_spline r = *this;
    r.arc_.assign(profile_pos.data(), profile_pos.data() + profile_pos.rows());
    r.geometry(nullptr, nudge_positions);
    for (long i = 0; i < profile_pos.rows(); ++i) profile_pos(i) = (float)r.arc_[(size_t)i];
    return r;
  }

  // sc:1017-1039: curvature at every resampled point
  std::vector<float> curvature() const { return std::vector<float>(kappa_.begin(), kappa_.end()); }

  // sc:1055-1067: omega_i = v_i * kappa_i
  std::vector<float> angular_velocity(const velocity_profile& vel_prof) const {
    if (vel_prof.vel.empty() || (size_t)vel_prof.vel[0].size() != arc_.size())
      throw std::invalid_argument("curvature and velocity vectors must be the same size");
    bezier_spline r = *this;
    std::vector<double> v(arc_.size());
    for (size_t i = 0; i < v.size(); ++i) v[i] = vel_prof.vel[0][(long)i];
    r.geometry(v.data(), false);
    return std::vector<float>(r.kappa_.begin(), r.kappa_.end());
  }

 private:
  std::vector<double> wx_, wy_, arc_, kappa_, edges_;  // edges_: [n][4] obstacle edges of the planning_space
  std::vector<double> cx_, cy_, tab_;                  // FP64 control points and arc-length tables of build(): what
                                                       // resample / angular_velocity evaluate (no second from_path)
  float start_angle_ = std::numeric_limits<float>::quiet_NaN();
  arclength_data ad_;

  void build() {  // SCB_STAGE_SPLINE: control points + arc-length tables
    b200::batch_planner& pl = b200::drop_in_planner();
    const int W = (int)wx_.size(), nseg = W - 1, nc = pl.degree() + 1;
    std::vector<double> cx((size_t)nseg * nc), cy((size_t)nseg * nc), tab((size_t)nseg * SCB_ARCLEN_TABLE);
    double L = 0, sa = start_angle_;
    int32_t status = 0;
    scb_batch_in in{};
    in.P = 1; in.W = W; in.G = b200::kDropInGrid; in.wx = wx_.data(); in.wy = wy_.data(); in.start_angle = &sa;
    in.obstacle_edges = edges_.empty() ? nullptr : edges_.data(); in.n_obstacle_edges = (int32_t)(edges_.size() / 4);
    in.stages = SCB_STAGE_SPLINE;
    scb_default_opt(&in.opt);
    scb_batch_out out{};
    out.status = &status; out.ctrl_x = cx.data(); out.ctrl_y = cy.data(); out.arclen_total = &L; out.arclen_table = tab.data();
    pl.plan(in, out);
    if (status != SCB_ST_OK) throw planning_error(status, -1, "sea_current_b200: bad path (non-finite waypoint or zero-length leg)");
    ctrl_pts.assign((size_t)nseg, std::vector<Vector2f>((size_t)nc));
    cx_ = cx; cy_ = cy; tab_ = tab;
    ad_ = arclength_data();
    ad_.arclength = (float)L;
    for (int s = 0; s < nseg; ++s) {
      for (int j = 0; j < nc; ++j) ctrl_pts[(size_t)s][(size_t)j] = Vector2f((float)cx[(size_t)s * nc + j], (float)cy[(size_t)s * nc + j]);
      VectorXf seg(SCB_ARCLEN_TABLE), pos(SCB_ARCLEN_TABLE);
      for (int k = 0; k < SCB_ARCLEN_TABLE; ++k) { seg(k) = (float)tab[(size_t)s * SCB_ARCLEN_TABLE + k]; pos(k) = std::fmin((float)k * 0.01f, 1.0f); }
      ad_.segments.push_back(seg); ad_.positions.push_back(pos);
    }
  }

  void geometry(const double* vel, bool nudge) {  // SCB_STAGE_GEOMETRY alone at arc_: the spline of build() is an input
    b200::batch_planner& pl = b200::drop_in_planner();
    const int S = (int)arc_.size();
    if (S > b200::kDropInMaxSamples) throw std::invalid_argument("sea_current_b200: too many samples for the drop-in context");
    if ((int)cx_.size() != (int)(wx_.size() - 1) * (pl.degree() + 1))
      throw std::logic_error("sea_current_b200: b200::spline_mode() changed between from_path and resample / angular_velocity");
    std::vector<double> px((size_t)S), py((size_t)S), om((size_t)S), vv;
    int32_t status = 0, ns = S;
    scb_batch_in in{};
    in.P = 1; in.W = (int)wx_.size(); in.G = b200::kDropInGrid;
    in.stages = SCB_STAGE_GEOMETRY; in.nudge_positions = nudge ? 1 : 0;
    scb_default_opt(&in.opt);
    scb_batch_out out{};
    out.ctrl_x = cx_.data(); out.ctrl_y = cy_.data(); out.arclen_table = tab_.data();
    out.status = &status; out.S_cap = S; out.n_samples = &ns; out.s_pos = arc_.data();
    if (vel) { vv.assign(vel, vel + S); out.s_vel = vv.data(); }
    out.s_px = px.data(); out.s_py = py.data(); out.s_omega = om.data();
    pl.plan(in, out);
    if (status != SCB_ST_OK) throw planning_error(status, -1, "sea_current_b200: bad path");
    pts.cx.assign(px.begin(), px.end()); pts.cy.assign(py.begin(), py.end());
    kappa_ = om;  // s_vel absent: v = 1, i.e. curvature; present: angular velocity
  }
};

// sc:1191-1265.  N = 1 is the instantiation sea-current's own call sequence uses (arc length is the one DOF) and runs
// the timing stage of scb_plan_batch; N = 2 .. SCB_MAX_DOF run scb_gen_vel_prof (the generic-row LP kernel, then the same
// parametrizer and sampler once per degree of freedom).
template <int N>
inline velocity_profile gen_vel_prof(const Eigen::Vector<value_type, N>& pos_end, const Eigen::Vector<value_type, N>& pos_start,
                                     const Eigen::Vector<value_type, N>& vel_end, const Eigen::Vector<value_type, N>& vel_start,
                                     const vel_lim_func& vel_lim, const Eigen::Vector<value_type, N>& acc_min,
                                     const Eigen::Vector<value_type, N>& acc_max, const float dt = 0.02f) {
  static_assert(N >= 1 && N <= SCB_MAX_DOF, "sea_current_b200: gen_vel_prof<N> is built for 1 <= N <= SCB_MAX_DOF degrees of freedom");
  b200::batch_planner& pl = b200::drop_in_planner();
  const int G = b200::kDropInGrid, S_cap = b200::kDropInMaxSamples;
  if constexpr (N > 1) {
    // vel_lim_func tabulated at toppra's gridpoints, [G][N] (one problem: P = 1)
    std::vector<double> vlo((size_t)G * N), vhi((size_t)G * N);
    for (int i = 0; i < G; ++i) {
      const double s = i == G - 1 ? 1.0 : (double)i * (1.0 / (double)(G - 1));
      auto lim = vel_lim(s);
      for (int k = 0; k < N; ++k) { vlo[(size_t)i * N + k] = std::get<0>(lim)(k); vhi[(size_t)i * N + k] = std::get<1>(lim)(k); }
    }
    double p0[N], p1[N], v0[N], v1[N], alo[N], ahi[N];
    for (int k = 0; k < N; ++k) { p0[k] = pos_start(k); p1[k] = pos_end(k); v0[k] = vel_start(k); v1[k] = vel_end(k); alo[k] = acc_min(k); ahi[k] = acc_max(k); }
    int32_t status = 0, fail = -1, ns = 0;
    std::vector<double> st((size_t)S_cap), sp((size_t)S_cap * N), sv((size_t)S_cap * N), sa((size_t)S_cap * N);
    scb_ndof_in in{};
    in.P = 1; in.dof = N; in.G = G; in.interpolation = 1;
    in.pos_start = p0; in.pos_end = p1; in.vel_start = v0; in.vel_end = v1;
    in.vel_lo_grid = vlo.data(); in.vel_hi_grid = vhi.data(); in.acc_lo = alo; in.acc_hi = ahi;
    in.dt = (double)dt; in.param_mode = b200::param_mode();
    scb_ndof_out out{};
    out.status = &status; out.fail_stage = &fail; out.S_cap = S_cap; out.n_samples = &ns;
    out.s_time = st.data(); out.s_pos = sp.data(); out.s_vel = sv.data(); out.s_acc = sa.data();
    b200::check(scb_gen_vel_prof(pl.ctx(), &in, &out), pl.ctx(), "scb_gen_vel_prof");
    if (status != SCB_ST_OK) {
      std::ostringstream m;
      m << "sea_current_b200: toppra verdict " << status << " at gridpoint " << fail;
      throw planning_error(status, fail, m.str());
    }
    if (ns > S_cap) throw std::runtime_error("sea_current_b200: trajectory longer than the drop-in sample capacity");
    std::vector<VectorXf> pos, vel, acc;
    for (int k = 0; k < N; ++k) {
      VectorXf a(ns), b(ns), c(ns);
      for (int i = 0; i < ns; ++i) {
        a(i) = (float)sp[(size_t)k * S_cap + i]; b(i) = (float)sv[(size_t)k * S_cap + i]; c(i) = (float)sa[(size_t)k * S_cap + i];
      }
      pos.push_back(a); vel.push_back(b); acc.push_back(c);
    }
    toppra::Vector time(ns);
    for (int i = 0; i < ns; ++i) time(i) = st[(size_t)i];
    return velocity_profile(pos, vel, acc, time);
  }
  // vel_lim_func of the path position s (sc:1177-1188), tabulated at toppra's gridpoints
  std::vector<double> vlo((size_t)G), vhi((size_t)G);
  for (int i = 0; i < G; ++i) {
    const double s = i == G - 1 ? 1.0 : (double)i * (1.0 / (double)(G - 1));
    auto lim = vel_lim(s);
    vlo[(size_t)i] = std::get<0>(lim)(0); vhi[(size_t)i] = std::get<1>(lim)(0);
  }
  const double L = pos_end(0) - pos_start(0), vs = vel_start(0), ve = vel_end(0), alo = acc_min(0), ahi = acc_max(0);
  int32_t status = 0, fail = -1, ns = 0;
  std::vector<double> st((size_t)S_cap), sp((size_t)S_cap), sv((size_t)S_cap), sa((size_t)S_cap);
  scb_batch_in in{};
  in.P = 1; in.W = 2; in.G = G; in.path_len = &L; in.vel_lo_grid = vlo.data(); in.vel_hi_grid = vhi.data();
  in.acc_lo = &alo; in.acc_hi = &ahi; in.vel_start = &vs; in.vel_end = &ve;
  in.stages = SCB_STAGE_TIMING; in.lp_mode = SCB_LP_SEIDEL; in.dt = (double)dt; in.param_mode = b200::param_mode();
  scb_default_opt(&in.opt);
  scb_batch_out out{};
  out.status = &status; out.fail_stage = &fail; out.S_cap = S_cap; out.n_samples = &ns;
  out.s_time = st.data(); out.s_pos = sp.data(); out.s_vel = sv.data(); out.s_acc = sa.data();
  pl.plan(in, out);
  if (status != SCB_ST_OK) {
    std::ostringstream m;
    m << "sea_current_b200: toppra verdict " << status << " at gridpoint " << fail;
    throw planning_error(status, fail, m.str());
  }
  if (ns > S_cap) throw std::runtime_error("sea_current_b200: trajectory longer than the drop-in sample capacity");
  const double p0 = pos_start(0);
  VectorXf pos(ns), vel(ns), acc(ns);
  toppra::Vector time(ns);
  for (int i = 0; i < ns; ++i) { pos(i) = (float)(sp[(size_t)i] + p0); vel(i) = (float)sv[(size_t)i]; acc(i) = (float)sa[(size_t)i]; time(i) = st[(size_t)i]; }
  return velocity_profile({pos}, {vel}, {acc}, time);
}

// Minimal JSON value for the one document the path emits (an array of flat sample objects):
// `o << std::setw(4) << j` and j.dump(4) print what nlohmann::json prints for sc:1439-1454
// (keys in alphabetical order, 4-space indent).
class json {
 public:
  struct sample { double time, velocity, acceleration, angularVelocity, x, y; };
  std::vector<sample> samples;
  size_t size() const { return samples.size(); }
  std::string dump(int indent = -1) const {
    std::string o;
    const bool pretty = indent >= 0;
    auto nl = [&](int depth) { if (pretty) { o += '\n'; o.append((size_t)(depth * indent), ' '); } };
    auto num = [&](double v) {  // shortest round-trip digits, laid out as nlohmann's dtoa (fixed for 1e-4 <= |v| < 1e15)
      if (!std::isfinite(v)) { o += "null"; return; }
      if (v == 0) { o += std::signbit(v) ? "-0.0" : "0.0"; return; }
      char buf[40];
      auto r = std::to_chars(buf, buf + sizeof buf, std::fabs(v), std::chars_format::scientific);
      std::string sci(buf, r.ptr), dig;
      const size_t epos = sci.find('e');
      for (size_t i = 0; i < epos; ++i) if (sci[i] != '.') dig += sci[i];
      const int k = (int)dig.size(), n = std::atoi(sci.c_str() + epos + 1) + 1;  // value = 0.dig * 10^n
      if (v < 0) o += '-';
      if (k <= n && n <= 15) { o += dig; o.append((size_t)(n - k), '0'); o += ".0"; }
      else if (0 < n && n <= 15) { o.append(dig, 0, (size_t)n); o += '.'; o.append(dig, (size_t)n, std::string::npos); }
      else if (-4 < n && n <= 0) { o += "0."; o.append((size_t)(-n), '0'); o += dig; }
      else {
        o += dig[0];
        if (k > 1) { o += '.'; o.append(dig, 1, std::string::npos); }
        const int e = n - 1;
        o += 'e'; o += e < 0 ? '-' : '+';
        const int ae = e < 0 ? -e : e;
        if (ae < 10) o += '0';
        o += std::to_string(ae);
      }
    };
    auto key = [&](const char* k, int depth) { nl(depth); o += '"'; o += k; o += pretty ? "\": " : "\":"; };
    if (samples.empty()) return "[]";
    o += '[';
    for (size_t i = 0; i < samples.size(); ++i) {
      const sample& s = samples[i];
      nl(1); o += '{';
      key("acceleration", 2); num(s.acceleration); o += ',';
      key("angularVelocity", 2); num(s.angularVelocity); o += ',';
      key("holonomicAngularVelocity", 2); num(0.0); o += ',';
      key("holonomicRotation", 2); num(0.0); o += ',';
      key("pose", 2); o += '{';
      key("translation", 3); o += '{';
      key("x", 4); num(s.x); o += ',';
      key("y", 4); num(s.y);
      nl(3); o += '}';
      nl(2); o += "},";
      key("time", 2); num(s.time); o += ',';
      key("velocity", 2); num(s.velocity);
      nl(1); o += '}';
      if (i + 1 < samples.size()) o += ',';
    }
    nl(0); o += ']';
    return o;
  }
  friend std::ostream& operator<<(std::ostream& os, const json& j) {
    const int w = (int)os.width();
    os.width(0);
    return os << j.dump(w > 0 ? w : -1);
  }
};

// sc:1423-1457
inline json serialize_path_to_json(const bezier_spline& spline, const velocity_profile& vel_prof, const arclength_data& /*arclens*/,
                                   const std::vector<float>& ang_vel) {
  json j;
  const long n = spline.pts.rows();
  if (vel_prof.vel.empty() || vel_prof.vel[0].size() < n || vel_prof.time.size() < n || (long)ang_vel.size() < n)
    throw std::invalid_argument("serialize_path_to_json: profile shorter than the resampled spline");
  j.samples.resize((size_t)n);
  for (long i = 0; i < n; ++i)
    j.samples[(size_t)i] = {(double)(float)vel_prof.time(i), (double)vel_prof.vel[0](i), (double)vel_prof.acc[0](i), (double)ang_vel[(size_t)i],
                            (double)spline.pts(i, 0), (double)spline.pts(i, 1)};
  return j;
}

namespace b200 {

// The request/response body of examples/zmq_test.cpp:26-101 without the socket: text
//   "acc_min acc_max vel_min vel_max\n x0 y0\n x1 y1 ..."  ->  pretty-printed JSON (setw(4)).
inline std::string handle_text_request(const std::string& req) {
  std::istringstream ss(req);
  float acc_min_val, acc_max_val, vel_min_val, vel_max_val;
  if (!(ss >> acc_min_val >> acc_max_val >> vel_min_val >> vel_max_val)) throw std::invalid_argument("request: four limits expected");
  std::vector<Vector2f> path;
  float x, y, max_x = 0, max_y = 0;
  while (ss >> x >> y) { path.emplace_back(x, y); max_x = std::fmax(max_x, std::fabs(x)); max_y = std::fmax(max_y, std::fabs(y)); }
  planning_space space(bounding_rect(max_x, -max_x, max_y, -max_y));
  bezier_spline pad = bezier_spline::from_path(path, space);
  const arclength_data ad = pad.arclength();
  const Eigen::Vector<value_type, 1> pos_end{ad.arclength}, zero{0.0}, amin{acc_min_val}, amax{acc_max_val};
  auto vel_lim = [&](value_type) {
    toppra::Vector lo(1), hi(1);
    lo(0) = vel_min_val; hi(0) = vel_max_val;
    return std::make_tuple(lo, hi);
  };
  velocity_profile prof = gen_vel_prof<1>(pos_end, zero, zero, zero, vel_lim, amin, amax);
  bezier_spline re = pad.resample(prof.pos[0], ad, true);
  const std::vector<float> ang_vel = re.angular_velocity(prof);
  return serialize_path_to_json(re, prof, ad, ang_vel).dump(4) + "\n";
}

// Batched variant of the request above: several requests in one message, separated by blank lines;
// the response is a JSON array with one element per request, in order: the array handle_text_request would return
// for it, or {"error": ..., "status": toppra ReturnCode / SCB_ST_*, "fail_stage": gridpoint} where that call would
// throw.  Requests with the same waypoint count are planned together: three scb_plan_batch calls per group
// (from_path + arclength; gen_vel_prof; resample + angular_velocity) in place of four per request, with the same
// float hand-overs between them as the one-request sequence, so every element is identical to the one-request
// response.
inline std::string handle_text_request_batch(const std::string& req) {
  struct request { float amin, amax, vmin, vmax; std::vector<double> x, y; std::string error; int status = 0, fail = -1; json out; };
  std::vector<request> rq;
  {  // split at blank lines
    std::istringstream all(req);
    std::string line, block;
    auto flush = [&] {
      if (block.find_first_not_of(" \t\r\n") == std::string::npos) { block.clear(); return; }
      request r{};
      std::istringstream ss(block);
      if (!(ss >> r.amin >> r.amax >> r.vmin >> r.vmax)) r.error = "request: four limits expected";
      float px, py;
      while (ss >> px >> py) { r.x.push_back(px); r.y.push_back(py); }
      if (r.error.empty() && r.x.size() < 2) r.error = "sea_current_b200: a path needs at least two waypoints";
      rq.push_back(std::move(r));
      block.clear();
    };
    while (std::getline(all, line)) {
      if (line.find_first_not_of(" \t\r") == std::string::npos) flush(); else { block += line; block += '\n'; }
    }
    flush();
  }
  static thread_local std::unique_ptr<batch_planner> pl_slot;
  static thread_local int pl_dev = -1;
  batch_planner& pl = context_for(pl_slot, pl_dev, 1 << 16);
  const int G = kDropInGrid;
  const double dt = (double)0.02f, nan = std::numeric_limits<double>::quiet_NaN();
  std::vector<int> order;  // requests grouped by waypoint count
  for (size_t i = 0; i < rq.size(); ++i) if (rq[i].error.empty()) order.push_back((int)i);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return rq[(size_t)a].x.size() < rq[(size_t)b].x.size(); });
  for (size_t g0 = 0; g0 < order.size();) {
    size_t g1 = g0;
    const int W = (int)rq[(size_t)order[g0]].x.size();
    while (g1 < order.size() && (int)rq[(size_t)order[g1]].x.size() == W && g1 - g0 < (size_t)(1 << 16)) ++g1;
    const int P = (int)(g1 - g0);
    auto R = [&](int p) -> request& { return rq[(size_t)order[g0 + (size_t)p]]; };
    // (1) from_path + arclength
    std::vector<double> wx((size_t)P * W), wy((size_t)P * W), sa((size_t)P, nan), L((size_t)P);
    std::vector<int32_t> st1((size_t)P), st2((size_t)P), fail((size_t)P), ns((size_t)P);
    for (int p = 0; p < P; ++p)
      for (int w = 0; w < W; ++w) { wx[(size_t)p * W + w] = R(p).x[(size_t)w]; wy[(size_t)p * W + w] = R(p).y[(size_t)w]; }
    scb_batch_in in{};
    in.P = P; in.W = W; in.G = G; in.wx = wx.data(); in.wy = wy.data(); in.start_angle = sa.data(); in.stages = SCB_STAGE_SPLINE;
    scb_default_opt(&in.opt);
    scb_batch_out out{};
    out.status = st1.data(); out.arclen_total = L.data();
    pl.plan(in, out);
    // (2) gen_vel_prof on the float arc length (arclength_data holds a float), limits tabulated per gridpoint
    std::vector<double> Lf((size_t)P), vlo((size_t)G * P), vhi((size_t)G * P), alo((size_t)P), ahi((size_t)P), zero((size_t)P, 0.0), T((size_t)P);
    for (int p = 0; p < P; ++p) {
      Lf[(size_t)p] = st1[(size_t)p] == SCB_ST_OK ? (double)(float)L[(size_t)p] : 1.0;
      alo[(size_t)p] = R(p).amin; ahi[(size_t)p] = R(p).amax;
      for (int i = 0; i < G; ++i) { vlo[(size_t)i * P + p] = R(p).vmin; vhi[(size_t)i * P + p] = R(p).vmax; }
    }
    scb_batch_in tin{};
    tin.P = P; tin.W = 2; tin.G = G; tin.path_len = Lf.data(); tin.vel_lo_grid = vlo.data(); tin.vel_hi_grid = vhi.data();
    tin.acc_lo = alo.data(); tin.acc_hi = ahi.data(); tin.vel_start = zero.data(); tin.vel_end = zero.data();
    tin.stages = SCB_STAGE_TIMING; tin.lp_mode = SCB_LP_SEIDEL; tin.dt = dt; tin.param_mode = param_mode();
    scb_default_opt(&tin.opt);
    scb_batch_out tout{};
    tout.status = st2.data(); tout.fail_stage = fail.data(); tout.T_total = T.data();
    pl.plan(tin, tout);  // durations first: they size the sample arrays
    int S = 1;
    for (int p = 0; p < P; ++p)
      if (st1[(size_t)p] == SCB_ST_OK && st2[(size_t)p] == SCB_ST_OK) S = std::max(S, (int)std::ceil(T[(size_t)p] / dt) + 2);
    if (S > kDropInMaxSamples) throw std::runtime_error("sea_current_b200: trajectory longer than the drop-in sample capacity");
    const size_t SP = (size_t)S * P;
    std::vector<double> s_t(SP), s_p(SP), s_v(SP), s_a(SP), s_x(SP), s_y(SP), s_o(SP);
    tout.S_cap = S; tout.n_samples = ns.data(); tout.T_total = nullptr;
    tout.s_time = s_t.data(); tout.s_pos = s_p.data(); tout.s_vel = s_v.data(); tout.s_acc = s_a.data();
    pl.plan(tin, tout);
    // (3) resample (nudged) + angular_velocity at the float positions / velocities of the profile
    for (int p = 0; p < P; ++p) {
      const bool ok = st1[(size_t)p] == SCB_ST_OK && st2[(size_t)p] == SCB_ST_OK && ns[(size_t)p] <= S;
      if (!ok) ns[(size_t)p] = 0;
      for (int i = 0; i < ns[(size_t)p]; ++i) {
        const size_t o = (size_t)i * P + p;
        s_p[o] = (double)(float)s_p[o]; s_v[o] = (double)(float)s_v[o];
      }
    }
    std::vector<int32_t> st3((size_t)P);
    in.stages = SCB_STAGE_SPLINE | SCB_STAGE_GEOMETRY; in.nudge_positions = 1;
    scb_batch_out gout{};
    gout.status = st3.data(); gout.S_cap = S; gout.n_samples = ns.data(); gout.s_pos = s_p.data(); gout.s_vel = s_v.data();
    gout.s_px = s_x.data(); gout.s_py = s_y.data(); gout.s_omega = s_o.data();
    pl.plan(in, gout);
    for (int p = 0; p < P; ++p) {
      request& r = R(p);
      if (st1[(size_t)p] != SCB_ST_OK) { r.status = st1[(size_t)p]; r.error = "sea_current_b200: bad path (non-finite waypoint or zero-length leg)"; continue; }
      if (st2[(size_t)p] != SCB_ST_OK) {
        r.status = st2[(size_t)p]; r.fail = fail[(size_t)p];
        r.error = "sea_current_b200: toppra verdict " + std::to_string(r.status) + " at gridpoint " + std::to_string(r.fail);
        continue;
      }
      r.out.samples.resize((size_t)ns[(size_t)p]);
      for (int i = 0; i < ns[(size_t)p]; ++i) {
        const size_t o = (size_t)i * P + p;
        r.out.samples[(size_t)i] = {(double)(float)s_t[o], (double)(float)s_v[o], (double)(float)s_a[o], (double)(float)s_o[o],
                                    (double)(float)s_x[o], (double)(float)s_y[o]};
      }
    }
    g0 = g1;
  }
  std::string resp = "[";
  for (size_t i = 0; i < rq.size(); ++i) {
    if (i) resp += ",";
    if (rq[i].error.empty()) { resp += rq[i].out.dump(-1); continue; }
    std::string msg;
    for (char c : rq[i].error) { if (c == '"' || c == '\\') msg += '\\'; msg += c; }
    resp += "{\"error\":\"" + msg + "\",\"fail_stage\":" + std::to_string(rq[i].fail) + ",\"status\":" + std::to_string(rq[i].status) + "}";
  }
  return resp + "]\n";
}

}  // namespace b200
}  // namespace turtle::sc

#endif  // SEA_CURRENT_B200_HPP
