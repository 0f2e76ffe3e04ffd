// TEST INFRASTRUCTURE ONLY: the SI-RRT planner of host/SIRRTPlanner.hpp instantiated on the CPU oracle
// (oracle/sirrt_oracle.h).  Used by tests/test_planner.py to check that the B200 backend yields IDENTICAL plans
// (same seed, same host logic, two checkers) and to time the CPU side of the plan-time metric.
//   planner_oracle <scene.pack> <seed> [--max-iters N] [--max-time S] [--fixed] [--baseline]      -> result JSON on stdout
// --baseline: the TIMED CPU arm (oracle_set_fast: boolean early exits + the reference's AABB tree over constant-pose
// windows, built before the clock starts like the reference's scene initialisation); results are bit-identical.
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <iterator>
#include <stdexcept>

#include "../include/sirrt_pack.h"
#include "../manipulationplanning-si-rrt_b200/host/ResultsLog.hpp"
#include "../manipulationplanning-si-rrt_b200/host/SIRRTPlanner.hpp"
#include "../oracle/sirrt_oracle.h"

struct OracleBackend {
  Oracle* o = nullptr;
  int dof_ = 0, n_links = 0, T = 0, fps_ = 0;
  std::vector<float> lim;
  std::vector<double> start, goal;
  std::vector<std::vector<double>> trees;  // flat point stores
  OracleBackend(const std::vector<unsigned char>& pack, int mode) {
    o = oracle_create(pack.data(), pack.size(), mode);
    if (!o) throw std::runtime_error(std::string("oracle: ") + oracle_last_error());
    oracle_info(o, &dof_, &n_links, &T, &fps_, nullptr);
    lim.resize(2 * dof_);
    oracle_info(o, nullptr, nullptr, nullptr, nullptr, lim.data());
    const double* p = (const double*)(pack.data() + sizeof(SirrtPackHeader));
    start.assign(p, p + dof_);
    goal.assign(p + dof_, p + 2 * dof_);
  }
  ~OracleBackend() { if (o) oracle_destroy(o); }
  int dof() const { return dof_; }
  int frame_count() const { return T; }
  int fps() const { return fps_; }
  std::vector<std::pair<float, float>> get_planned_robot_limits() const {
    std::vector<std::pair<float, float>> r;
    for (int j = 0; j < dof_; j++) r.emplace_back(lim[2 * j], lim[2 * j + 1]);
    return r;
  }
  std::vector<double> start_configuration() const { return start; }
  std::vector<double> end_configuration() const { return goal; }
  std::vector<std::pair<int, int>> get_safe_intervals(const std::vector<double>& q) {
    std::vector<int32_t> buf(2 * 2048);
    const int n = oracle_safe_intervals(o, q.data(), buf.data(), 2048);
    std::vector<std::pair<int, int>> r;
    for (int i = 0; i < n; i++) r.emplace_back(buf[2 * i], buf[2 * i + 1]);
    return r;
  }
  int edge_first_collision(const double* q0, const double* dir, int steps, double t0, double t1) {
    std::vector<double> q(dof_);
    for (int s = 0; s <= steps; s++) {
      for (int j = 0; j < dof_; j++) q[j] = q0[j] + dir[j] * (double)s / (double)steps;
      const int frame = (int)(t0 + (t1 - t0) * (double)s / (double)steps);
      if (oracle_check_frame(o, q.data(), frame)) return s;
    }
    return -1;
  }
  int knn_new() { trees.emplace_back(); return (int)trees.size() - 1; }
  void knn_add(int t, const double* p) { trees[t].insert(trees[t].end(), p, p + dof_); }
  size_t knn_nearest(int t, const double* q) {
    int64_t idx = 0;
    double d2 = 0;
    oracle_knn_nearest(trees[t].data(), (int64_t)(trees[t].size() / dof_), dof_, q, 1, &idx, &d2);
    return (size_t)idx;
  }
  void knn_radius(int t, const double* q, double r2, std::vector<size_t>& out) {
    const int64_t n = (int64_t)(trees[t].size() / dof_);
    std::vector<int64_t> idx((size_t)n + 1);
    const int64_t m = oracle_knn_radius(trees[t].data(), n, dof_, q, r2, idx.data(), n);
    out.assign(idx.begin(), idx.begin() + m);
  }
};

// --check-prefetch: the same backend WITH the optional knn_radius_prefetch hook of the planner template (what B200Backend
// uses to launch the radius query of set_parent before the safe-interval call).  Nothing is prefetched here; the hook only
// records (tree, q, r2, tree size) and knn_radius counts how many of its queries were announced exactly -- the planner-side
// half of that optimisation, checked on the CPU.  Plans must not depend on the hook.
struct OraclePrefetchBackend : OracleBackend {
  using OracleBackend::OracleBackend;
  bool have = false;
  int h_tree = -1;
  double h_r2 = 0;
  size_t h_size = 0;
  std::vector<double> h_q;
  uint64_t hints = 0, queries = 0, matched = 0, mismatched = 0, overwritten = 0;
  void knn_radius_prefetch(int t, const double* q, double r2) {
    if (have) overwritten++;
    have = true; h_tree = t; h_r2 = r2; h_size = trees[t].size(); h_q.assign(q, q + dof_);
    hints++;
  }
  void knn_add(int t, const double* p) { OracleBackend::knn_add(t, p); }
  void knn_radius(int t, const double* q, double r2, std::vector<size_t>& out) {
    queries++;
    if (have) {
      have = false;
      if (t == h_tree && r2 == h_r2 && trees[t].size() == h_size && !memcmp(h_q.data(), q, sizeof(double) * dof_)) matched++;
      else mismatched++;
    }
    OracleBackend::knn_radius(t, q, r2, out);
  }
};

template <typename BE> static int run_planner(BE& be, const char* scene_path, int seed, uint64_t max_iters, double max_time, bool baseline);

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s <scene.pack> <seed> [--max-iters N] [--max-time S] [--fixed]\n", argv[0]); return 2; }
  std::ifstream f(argv[1], std::ios::binary);
  if (!f) { fprintf(stderr, "cannot open %s\n", argv[1]); return 1; }
  std::vector<unsigned char> pack((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  const int seed = atoi(argv[2]);
  uint64_t max_iters = 0;
  double max_time = -1;
  int mode = ORACLE_REFERENCE_EXACT;
  bool baseline = false, check_prefetch = false;
  for (int i = 3; i < argc; i++) {
    if (!strcmp(argv[i], "--max-iters") && i + 1 < argc) max_iters = strtoull(argv[++i], nullptr, 10);
    else if (!strcmp(argv[i], "--max-time") && i + 1 < argc) max_time = atof(argv[++i]);
    else if (!strcmp(argv[i], "--fixed")) mode = ORACLE_FIXED;
    else if (!strcmp(argv[i], "--baseline")) baseline = true;
    else if (!strcmp(argv[i], "--check-prefetch")) check_prefetch = true;
  }
  try {
    if (check_prefetch) { OraclePrefetchBackend be(pack, mode); return run_planner(be, argv[1], seed, max_iters, max_time, baseline); }
    OracleBackend be(pack, mode);
    return run_planner(be, argv[1], seed, max_iters, max_time, baseline);
  } catch (const std::exception& e) {
    fprintf(stderr, "planner_oracle: %s\n", e.what());
    return 1;
  }
}

static void prefetch_extras(OracleBackend&, MDP::ResultsLog&) {}
static void prefetch_extras(OraclePrefetchBackend& be, MDP::ResultsLog& log) {
  log.extra.push_back({"radius_hints", (double)be.hints});
  log.extra.push_back({"radius_queries", (double)be.queries});
  log.extra.push_back({"radius_queries_announced_exactly", (double)be.matched});
  log.extra.push_back({"radius_hint_mismatches", (double)be.mismatched});
  log.extra.push_back({"radius_hints_overwritten", (double)be.overwritten});
}

template <typename BE> static int run_planner(BE& be, const char* scene_path, int seed, uint64_t max_iters, double max_time, bool baseline) {
  {
    if (baseline) { oracle_set_fast(1); oracle_bvh_build(be.o); }
    MDP::MSIRRT::PlannerConnect<BE> planner(be, seed);
    if (!planner.valid()) { fprintf(stderr, "planner_oracle: %s\n", planner.error().c_str()); return 1; }
    if (max_iters) planner.set_max_iterations(max_iters);
    if (max_time > 0) planner.set_max_planning_time(max_time);
    const bool result = planner.solve();
    MDP::ResultsLog log;
    log.path_to_scene_json = scene_path;
    const auto& st = planner.stats();
    log.set_counters(st, be.n_links);
    log.has_result = result;
    log.start_tree_vertices = planner.get_start_tree_vertices_count();
    log.goal_tree_vertices = planner.get_goal_tree_vertices_count();
    if (result) {
      auto path = planner.get_final_path();
      log.extra.push_back({"check_path", planner.check_path(path) ? 1.0 : 0.0});
      for (const auto& s : planner.get_result_path(0.0)) log.final_path.push_back({s.q, s.time});
      log.path_cost = log.final_path.back().time;
    }
    log.extra.push_back({"iterations", (double)st.iterations});
    log.extra.push_back({"motion_checks", (double)st.n_motion_checks});
    prefetch_extras(be, log);
    std::cout << log.to_json() << std::endl;
    return 0;
  }
}
