// Header-only C++ view of graphslam_cuda.h for the reference's graph node
// (reference src/graph/src/graph.cpp).  gs::PoseGraph stands where the node keeps
// `gtsam::NonlinearFactorGraph graph; gtsam::Values initial;` (graph.cpp:6-7); gs::GraphNode is the
// node's own bookkeeping (prior_factor / new_factor / loop_factor / solve, graph.cpp:27-132) with ROS stripped.
#pragma once
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "graphslam_cuda.h"

namespace gs {

struct Error : std::runtime_error {
  gs_status status;
  Error(gs_status s, const std::string& m) : std::runtime_error(m), status(s) {}
};

class PoseGraph {
 public:
  explicit PoseGraph(const gs_params* p = nullptr) {
    gs_status s = gs_create(&g_, p);
    if (s != GS_OK) {
      std::string m = g_ ? gs_last_error(g_) : "gs_create failed";
      if (g_) gs_destroy(g_);
      g_ = nullptr;
      throw Error(s, m);
    }
  }
  ~PoseGraph() {
    if (g_) gs_destroy(g_);
  }
  PoseGraph(const PoseGraph&) = delete;
  PoseGraph& operator=(const PoseGraph&) = delete;

  void add_vertex(uint64_t key, const double xyt[3]) { ck(gs_add_vertex(g_, key, xyt)); }                 // graph.cpp:58,86
  void add_prior(uint64_t key, const double xyt[3], const double cov9[9]) { ck(gs_add_prior(g_, key, xyt, cov9)); }  // :57
  void add_between(uint64_t k1, uint64_t k2, const double d[3], const double cov9[9]) {                   // :87-92,106-111
    ck(gs_add_between(g_, k1, k2, d, cov9));
  }
  gs_stats optimize() {                                                                                    // :119
    gs_stats st;
    gs_status s = gs_optimize(g_, &st);
    if (s != GS_OK && s != GS_ENOTCONVERGED) throw Error(s, gs_last_error(g_));
    return st;
  }
  void pose(uint64_t key, double xyt[3]) { ck(gs_get_pose(g_, key, xyt)); }                                // :123-125
  double chi2() {
    double v = 0;
    ck(gs_chi2(g_, &v));
    return v;
  }
  void marginal_covariance(uint64_t key, double cov9[9]) { ck(gs_marginal_covariance(g_, key, cov9)); }     // :120,126-127
  gs_g2o_counts load_g2o(const char* path) {               // g2o / TORO text; the reference has no persistence (:5-10)
    gs_g2o_counts c;
    ck(gs_load_g2o(g_, path, &c));
    return c;
  }
  void save_g2o(const char* path) { ck(gs_save_g2o(g_, path)); }
  size_t num_vertices() const {
    size_t n = 0;
    gs_num_vertices(g_, &n);
    return n;
  }
  size_t num_factors() const {                                                                              // graph.nrFactors()
    size_t n = 0;
    gs_num_factors(g_, &n);
    return n;
  }
  bool closest(uint64_t key, int skip, uint64_t* out, double* dist = nullptr) {                            // :146-178
    gs_status s = gs_closest_keyframe(g_, key, skip, out, dist);
    if (s == GS_ENOKEY) return false;
    ck(s);
    return true;
  }
  gs_graph* handle() { return g_; }

 private:
  void ck(gs_status s) {
    if (s != GS_OK) throw Error(s, gs_last_error(g_));
  }
  gs_graph* g_ = nullptr;
};

// Plain-struct mirrors of the wire types the node reads (src/common/msg/*.msg), ROS-free.
struct Pose2DWithCovariance {           // Pose2DWithCovariance.msg
  double x = 0, y = 0, theta = 0;
  std::array<double, 9> covariance{};   // row-major
};
struct Factor {                         // Factor.msg (ids widened from int8)
  uint64_t id_1 = 0, id_2 = 0;
  Pose2DWithCovariance delta;
};
struct Keyframe {                       // Keyframe.msg without the point cloud
  uint64_t id = 0;
  Pose2DWithCovariance pose_opti;
};
struct Keyframes {                      // Keyframes.msg: what publish_keyframes sends (graph.cpp:18-25)
  std::vector<Keyframe> keyframes;
};
struct Registration {                   // Registration.msg
  bool first_frame_flag = false, keyframe_flag = false, loop_closure_flag = false;
  Keyframe keyframe_new, keyframe_last;
  Factor factor_new, factor_loop;
};

// One entry of a recorded registration stream (graphslam_b200/regstream.py writes the same text format):
//   REG <first> <keyframe> <loop> [NEW id_1 dx dy dtheta c0..c8] [LOOP id_1 id_2 dx dy dtheta c0..c8]
//   SOLVE        where the node's solve() runs (the call the reference has commented out, graph.cpp:195)
struct StreamEvent {
  bool solve = false;
  Registration reg;
};

inline std::vector<StreamEvent> read_registration_stream(const char* path) {
  std::vector<StreamEvent> out;
  std::FILE* f = std::fopen(path, "r");
  if (!f) throw Error(GS_EINVAL, std::string("cannot open ") + path);
  std::vector<char> buf(1 << 16);
  int lineno = 0;
  auto bad = [&](const char* why) {
    std::fclose(f);
    throw Error(GS_EINVAL, std::string(path) + ":" + std::to_string(lineno) + ": " + why);
  };
  while (std::fgets(buf.data(), (int)buf.size(), f)) {
    ++lineno;
    if (char* hash = std::strchr(buf.data(), '#')) *hash = 0;
    std::vector<std::string> t;
    for (char* tok = std::strtok(buf.data(), " \t\r\n"); tok; tok = std::strtok(nullptr, " \t\r\n")) t.emplace_back(tok);
    if (t.empty()) continue;
    StreamEvent ev;
    if (t[0] == "SOLVE") {
      ev.solve = true;
      out.push_back(ev);
      continue;
    }
    if (t[0] != "REG" || t.size() < 4) bad("not a registration line");
    ev.reg.first_frame_flag = t[1] == "1";
    ev.reg.keyframe_flag = t[2] == "1";
    ev.reg.loop_closure_flag = t[3] == "1";
    auto pose = [&](size_t k, Pose2DWithCovariance& d) {   // 12 numbers from token k
      d.x = std::strtod(t[k].c_str(), nullptr);
      d.y = std::strtod(t[k + 1].c_str(), nullptr);
      d.theta = std::strtod(t[k + 2].c_str(), nullptr);
      for (int q = 0; q < 9; ++q) d.covariance[q] = std::strtod(t[k + 3 + q].c_str(), nullptr);
    };
    size_t k = 4;
    while (k < t.size()) {
      if (t[k] == "NEW" && k + 14 <= t.size()) {
        ev.reg.factor_new.id_1 = std::strtoull(t[k + 1].c_str(), nullptr, 10);
        pose(k + 2, ev.reg.factor_new.delta);
        k += 14;
      } else if (t[k] == "LOOP" && k + 15 <= t.size()) {
        ev.reg.factor_loop.id_1 = std::strtoull(t[k + 1].c_str(), nullptr, 10);
        ev.reg.factor_loop.id_2 = std::strtoull(t[k + 2].c_str(), nullptr, 10);
        pose(k + 3, ev.reg.factor_loop.delta);
        k += 15;
      } else {
        bad("unknown or truncated field");
      }
    }
    out.push_back(ev);
  }
  std::fclose(f);
  return out;
}

class GraphNode {
 public:
  double sigma_xy_prior = 0.1, sigma_th_prior = 0.1;   // graph.cpp:13-14
  int keyframes_to_skip_in_loop_closing = 10;          // graph.cpp:15
  bool faithful_compose = false;                       // true: seed with the reference's own compose (graph.hpp:30-43)
  std::vector<Keyframe> keyframes;
  PoseGraph graph;

  void prior_factor(Registration input) {              // graph.cpp:27-61
    ++keyframe_ids_;
    const double q[9] = {sigma_xy_prior * sigma_xy_prior, 0, 0, 0, sigma_xy_prior * sigma_xy_prior, 0, 0, 0,
                         sigma_th_prior * sigma_th_prior};
    const double pose[3] = {0, 0, 0};
    input.keyframe_new.id = keyframe_ids_;
    input.keyframe_new.pose_opti.x = input.keyframe_new.pose_opti.y = input.keyframe_new.pose_opti.theta = 0;
    keyframes.push_back(input.keyframe_new);
    graph.add_vertex(keyframe_ids_, pose);
    graph.add_prior(keyframe_ids_, pose, q);
  }
  void new_factor(Registration input) {                // graph.cpp:63-95
    ++keyframe_ids_;
    // graph.cpp:72 composes from input.keyframe_last.pose_opti: what the scanner got from the last_keyframe service
    // (graph.cpp:134-144) before it published.  A message without it (id 0) falls back to the node's own last keyframe.
    const Pose2DWithCovariance& last = input.keyframe_last.id ? input.keyframe_last.pose_opti : keyframes.back().pose_opti;
    const double a[3] = {last.x, last.y, last.theta};
    const double d[3] = {input.factor_new.delta.x, input.factor_new.delta.y, input.factor_new.delta.theta};
    double pn[3];
    if (faithful_compose) gs_compose_reference(a, d, pn);
    else gs_compose(a, d, pn);
    input.keyframe_new.id = keyframe_ids_;
    input.keyframe_new.pose_opti.x = pn[0];
    input.keyframe_new.pose_opti.y = pn[1];
    input.keyframe_new.pose_opti.theta = pn[2];
    keyframes.push_back(input.keyframe_new);
    input.factor_new.id_2 = keyframe_ids_;
    graph.add_vertex(keyframe_ids_, pn);
    graph.add_between(input.factor_new.id_1, input.factor_new.id_2, d, input.factor_new.delta.covariance.data());
  }
  void loop_factor(const Registration& input) {        // graph.cpp:97-113
    const double d[3] = {input.factor_loop.delta.x, input.factor_loop.delta.y, input.factor_loop.delta.theta};
    graph.add_between(input.factor_loop.id_1, input.factor_loop.id_2, d, input.factor_loop.delta.covariance.data());
  }
  gs_stats solve() {                                   // graph.cpp:115-132
    gs_stats st = graph.optimize();
    for (auto& kf : keyframes) {
      double p[3];
      graph.pose(kf.id, p);
      kf.pose_opti.x = p[0];
      kf.pose_opti.y = p[1];
      kf.pose_opti.theta = p[2];
    }
    return st;  // warm start: the library keeps the optimum as the next initial values (graph.cpp:130)
  }
  bool last_keyframe(Keyframe* out) const {            // LastKeyframe service, graph.cpp:134-144
    if (keyframes.empty()) return false;
    *out = keyframes.back();
    return true;
  }
  Keyframes publish_keyframes() const {                // graph.cpp:18-25 (the message, not the transport)
    Keyframes out;
    out.keyframes = keyframes;
    return out;
  }
  // replay of a recorded stream: every REG goes through registration_callback with keyframe_last filled from the
  // last_keyframe service, as the scanner does; SOLVE runs solve().  Returns the stats of the last solve.
  gs_stats replay(const std::vector<StreamEvent>& events) {
    gs_stats st{};
    for (const StreamEvent& ev : events) {
      if (ev.solve) {
        st = solve();
        continue;
      }
      Registration reg = ev.reg;
      if (reg.keyframe_flag) last_keyframe(&reg.keyframe_last);
      registration_callback(reg, false);
    }
    return st;
  }
  void registration_callback(const Registration& input, bool do_solve) {   // graph.cpp:180-200
    if (input.first_frame_flag) {
      prior_factor(input);
    } else if (input.keyframe_flag) {
      new_factor(input);
      if (input.loop_closure_flag) loop_factor(input);
      if (do_solve) solve();  // the call the reference has commented out (graph.cpp:195)
    }
  }

 private:
  uint64_t keyframe_ids_ = 0;                          // graph.cpp:207
};

}  // namespace gs
