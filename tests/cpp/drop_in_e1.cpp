// Drop-in check: the call sequence a sea-current user writes (examples/json_test.cpp:28-65 and the
// request body of examples/zmq_test.cpp:26-101), compiled against include/sea_current_b200.hpp
// instead of sea_current.hpp.  tests/test_drop_in.py builds it with g++, runs it on the GPU box and
// compares what it prints with the oracle.
//
//   drop_in_e1 json <out.json>      the six-waypoint case of examples/json_test.cpp
//   drop_in_e1 text <out.json>      the request of examples/zmq_test.py:7-10 through handle_text_request
//   drop_in_e1 lines <unused>       edge lists of the two obstacle constructors (host only, sc:251-281)
//   drop_in_e1 obst <unused>        from_path over a planning_space with obstacles (shrink_tangent, sc:575-596)
//   drop_in_e1 textbatch <out.json> five requests in one message through handle_text_request_batch; each one also through
//                                   handle_text_request into <out.json>.<k> (or its exception text)
//   drop_in_e1 sample <n>           planning_space::sample_free over the map of examples/test.cpp:250-259 (host only)
//   drop_in_e1 ndof <unused>        gen_vel_prof<3>: three degrees of freedom, a velocity limit that drops mid-path (sc:1191-1265)
//   drop_in_e1 fmt <n>              fast_marching_trees(( -0.5, 1), (1, -1), n, 1) as examples/test.cpp:284 calls it, then
//                                   b200::fmt_batch with three queries over a fresh sample set
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>

#include "sea_current_b200.hpp"

using namespace Eigen;
using namespace turtle::sc;

static int run_json_case(const char* out_path) {
  planning_space space(bounding_rect{1, -1, 1, -1});
  const std::vector<Vector2f> path = {Vector2f(0, 0), Vector2f(10, 0), Vector2f(10, 10), Vector2f(10, 20), Vector2f(30, 50), Vector2f(0, 0)};
  bezier_spline pad = bezier_spline::from_path(path, space);
  const arclength_data ad = pad.arclength();
  const Eigen::Vector<value_type, 1> pos_end{ad.arclength}, pos_start{0}, vel_end{0}, vel_start{0}, acc_min{-0.25}, acc_max{0.25};
  auto vel_lim = [](toppra::value_type) {
    toppra::Vector lower{1}, upper{1};
    lower(0, 0) = 0;
    upper(0, 0) = 1;
    return std::make_tuple(lower, upper);
  };
  velocity_profile prof = gen_vel_prof<1>(pos_end, pos_start, vel_end, vel_start, vel_lim, acc_min, acc_max);
  bezier_spline re = pad.resample(prof.pos[0], ad, true);
  const std::vector<float> ang_vel = re.angular_velocity(prof);
  json j = serialize_path_to_json(re, prof, ad, ang_vel);
  std::ofstream o(out_path);
  o << std::setw(4) << j << std::endl;
  // one summary line for the test: arclength, sample count, first table entries, a control point
  std::cout << std::setprecision(9) << "{\"arclength\": " << ad.arclength << ", \"n_samples\": " << re.n_pts()
            << ", \"n_segments\": " << pad.n_segments() << ", \"degree\": " << pad.degree() << ", \"seg0_1\": " << ad.segments[0](1)
            << ", \"ctrl_1_x\": " << pad.ctrl_pts[0][1].x() << ", \"ctrl_1_y\": " << pad.ctrl_pts[0][1].y() << "}" << std::endl;
  return 0;
}

// Two obstacles near the corner (10, 0) of an L-shaped path: a 4-vertex one (even count: the reference closes its edge
// list) and a 3-vertex polyline (open).
static std::vector<obstacle> corner_obstacles() {
  obstacle a(std::vector<Vector2f>{Vector2f(8, -2.5f), Vector2f(8, -0.5f), Vector2f(9, -0.5f), Vector2f(9, -2.5f)});
  obstacle b(std::vector<Vector2f>{Vector2f(11, 2), Vector2f(13, 2), Vector2f(13, 3)});
  obstacle c(std::vector<Vector2f>{Vector2f(20, 20), Vector2f(21, 20), Vector2f(21, 21)}, std::vector<std::tuple<int, int>>{{0, 2}, {2, 1}});
  return {a, b, c};
}

static void print_lines(const std::vector<obstacle>& obs) {
  std::cout << "{\"lines\": [";
  bool first = true;
  for (const auto& o : obs)
    for (const auto& l : o.lines) {
      std::cout << (first ? "" : ", ") << "[" << std::get<0>(l).x() << ", " << std::get<0>(l).y() << ", " << std::get<1>(l).x() << ", " << std::get<1>(l).y() << "]";
      first = false;
    }
  std::cout << "], \"n_vertices\": [";
  for (size_t i = 0; i < obs.size(); ++i) std::cout << (i ? ", " : "") << obs[i].vertices.size();
  std::cout << "]";
}

static int run_obstacle_case() {
  planning_space space(bounding_rect{30, -30, 30, -30});
  space.obstacles = corner_obstacles();
  const std::vector<Vector2f> path = {Vector2f(0, 0), Vector2f(10, 0), Vector2f(10, 10), Vector2f(0, 10)};
  bezier_spline sp = bezier_spline::from_path(path, space);
  const arclength_data ad = sp.arclength();
  print_lines(space.obstacles);
  std::cout << std::setprecision(9) << ", \"arclength\": " << ad.arclength << ", \"ctrl\": [";
  for (int s = 0; s < sp.n_segments(); ++s)
    for (int k = 0; k < 4; ++k) std::cout << (s + k ? ", " : "") << "[" << sp.ctrl_pts[s][k].x() << ", " << sp.ctrl_pts[s][k].y() << "]";
  std::cout << "]}" << std::endl;
  return 0;
}

// examples/test.cpp:250-259, with all of the rectangle allocated as free space (the example allocates none, so its
// sample_free cannot return)
static planning_space example_space() {
  planning_space space(bounding_rect{1, -1, 1, -1});
  space.obstacles.push_back(obstacle({Vector2f(-0.5, 0), Vector2f(1, 0), Vector2f(1, 1), Vector2f(0, 1)}));
  space.obstacles.push_back(obstacle({Vector2f(0, -0.5), Vector2f(1, 0), Vector2f(1, 1), Vector2f(0, 1)}));
  space.obstacles.push_back(obstacle({Vector2f(-0.6f, 0.148f), Vector2f(-1, 0.148f), Vector2f(-1, 0), Vector2f(-0.6f, 0)}));
  space.free_space_allocations.push_back([](Vector2f) { return true; });
  return space;
}

static void print_pts(const char* key, const std::vector<Vector2f>& pts) {
  std::cout << "\"" << key << "\": [";
  for (size_t i = 0; i < pts.size(); ++i) std::cout << (i ? ", " : "") << "[" << pts[i].x() << ", " << pts[i].y() << "]";
  std::cout << "]";
}

static int run_sample_case(int n) {
  planning_space space = example_space();
  const std::vector<Vector2f> pts = space.sample_free_ordered(n);
  std::cout << std::setprecision(9) << "{";
  print_pts("pts", pts);
  const point_set again = space.sample_free(n);  // the Halton states carry on (sc:1302-1303)
  std::cout << ", \"state\": [" << space.x_state.f << ", " << space.x_state.i << ", " << space.y_state.f << ", " << space.y_state.i << "]"
            << ", \"second_set_size\": " << again.size() << ", \"contains\": [" << space.obstacles[0].contains(Vector2f(0.5, 0.5)) << ", "
            << space.obstacles[0].contains(Vector2f(-0.4f, 0.9f)) << ", " << space.is_free(Vector2f(0.5, 0.5)) << ", " << space.is_free(Vector2f(-0.9f, -0.9f))
            << "], \"cost\": [" << space.cost(Vector2f(-0.9f, -0.9f), Vector2f(-0.5f, -0.6f)) << ", "
            << (space.cost(Vector2f(0.5f, -0.9f), Vector2f(0.5f, 0.9f)) == std::numeric_limits<float>::max()) << "]"
            << ", \"near\": " << space.near(Vector2f(0, 0), point_set(pts.begin(), pts.end()), 0.5f).size() << "}" << std::endl;
  return 0;
}

static int run_fmt_case(int n) {
  planning_space space = example_space();
  const auto path = space.fast_marching_trees(Vector2f(-0.5, 1), Vector2f(1, -1), n, 1);
  std::cout << std::setprecision(9) << "{\"found\": " << path.has_value() << ", ";
  print_pts("path", path.value_or(std::vector<Vector2f>{}));
  const std::vector<Vector2f> samples = space.sample_free_ordered(n);  // second draw: the states have advanced
  std::vector<float> costs;
  const auto paths = b200::fmt_batch(space, samples, {{Vector2f(-0.9f, -0.9f), Vector2f(-0.9f, 0.9f)}, {Vector2f(-0.5, 1), Vector2f(1, -1)},
                                                       {Vector2f(0.5f, 0.5f), Vector2f(-0.9f, -0.9f)}}, 0.6f, &costs);
  std::cout << ", ";
  print_pts("samples", samples);
  std::cout << ", \"batch\": [";
  for (size_t q = 0; q < paths.size(); ++q) {
    std::cout << (q ? ", " : "") << "{\"found\": " << paths[q].has_value() << ", \"cost\": " << (paths[q] ? costs[q] : -1.f) << ", ";
    print_pts("path", paths[q].value_or(std::vector<Vector2f>{}));
    std::cout << "}";
  }
  std::cout << "]}" << std::endl;
  return 0;
}

static int run_text_batch(const char* out_path) {
  const std::vector<std::string> reqs = {
      "-0.5 0.5 -0.25 0.25\n0 0\n10 0\n10 10",                          // examples/zmq_test.py:7-10
      "-0.25 0.25 0 1\n0 0\n10 0\n10 10\n10 20\n30 50\n0 0",            // examples/json_test.cpp:19-26
      "-1.5 1.5 -2 2\n0 0\n4 1\n6 5",                                    // same waypoint count as the first: one group
      "-0.5 0.5 -0.25 0.25\n0 0\n10 0\n10 0\n20 5",                      // zero-length leg: bad path
      "-0.5 0.5 0.1 0.25\n0 0\n10 0\n10 10"};                            // velocity box excludes rest: toppra refuses
  std::string all;
  for (size_t k = 0; k < reqs.size(); ++k) all += reqs[k] + (k + 1 < reqs.size() ? "\n\n" : "\n");
  { std::ofstream o(out_path); o << b200::handle_text_request_batch(all); }
  for (size_t k = 0; k < reqs.size(); ++k) {
    std::ofstream o(std::string(out_path) + "." + std::to_string(k));
    try { o << b200::handle_text_request(reqs[k]); }
    catch (const std::exception& e) { o << "\"" << e.what() << "\"" << std::endl; }
  }
  std::cout << "{\"ok\": true, \"n\": " << reqs.size() << "}" << std::endl;
  return 0;
}

// gen_vel_prof<N> with N = 3 (sc:1191, `requires (N >= 1)`): the reference's signature, end before start
static int run_ndof_case() {
  const Eigen::Vector<value_type, 3> pos_end{4.0, -1.5, 2.0}, pos_start{0.5, 1.0, 2.5}, vel_end{0.2, 0.0, 0.0}, vel_start{0.0, -0.1, 0.0},
      acc_min{-0.5, -0.4, -0.3}, acc_max{0.5, 0.6, 0.3};
  auto vel_lim = [](toppra::value_type s) {
    toppra::Vector lower(3), upper(3);
    const double f = (s > 0.405 && s < 0.595) ? 0.25 : 1.0;
    for (int k = 0; k < 3; ++k) { upper(k) = f * (1.0 + 0.5 * k); lower(k) = -upper(k); }
    return std::make_tuple(lower, upper);
  };
  velocity_profile prof = gen_vel_prof<3>(pos_end, pos_start, vel_end, vel_start, vel_lim, acc_min, acc_max, 0.05f);
  std::cout << std::setprecision(9) << "{\"n\": " << prof.time.rows() << ", \"dof\": " << prof.pos.size() << ", \"T\": " << prof.time(prof.time.rows() - 1);
  for (const char* name : {"pos", "vel", "acc"}) {
    const std::vector<VectorXf>& v = !std::strcmp(name, "pos") ? prof.pos : (!std::strcmp(name, "vel") ? prof.vel : prof.acc);
    std::cout << ", \"" << name << "\": [";
    for (size_t k = 0; k < v.size(); ++k) {
      std::cout << (k ? ", [" : "[");
      for (long i = 0; i < v[k].rows(); ++i) std::cout << (i ? ", " : "") << v[k](i);
      std::cout << "]";
    }
    std::cout << "]";
  }
  std::cout << "}" << std::endl;
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 3) { std::cerr << "usage: drop_in_e1 json|text <out.json>" << std::endl; return 2; }
  try {
    if (!std::strcmp(argv[1], "json")) return run_json_case(argv[2]);
    if (!std::strcmp(argv[1], "lines")) { print_lines(corner_obstacles()); std::cout << "}" << std::endl; return 0; }
    if (!std::strcmp(argv[1], "obst")) return run_obstacle_case();
    if (!std::strcmp(argv[1], "textbatch")) return run_text_batch(argv[2]);
    if (!std::strcmp(argv[1], "sample")) return run_sample_case(std::atoi(argv[2]));
    if (!std::strcmp(argv[1], "fmt")) return run_fmt_case(std::atoi(argv[2]));
    if (!std::strcmp(argv[1], "ndof")) return run_ndof_case();
    const std::string req = "-0.5 0.5 -0.25 0.25\n0 0\n10 0\n10 10";
    std::ofstream o(argv[2]);
    o << b200::handle_text_request(req);
    std::cout << "{\"ok\": true}" << std::endl;
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "drop_in_e1: " << e.what() << std::endl;
    return 3;
  }
}
