// sirrt_plan -- SI-RRT planner executable on the B200 backend (counterpart of MSIRRT/src/main.cpp:41-118).
//   sirrt_plan <scene.pack> <result_dir> <seed> [--max-iters N] [--max-time S] [--no-edge-table] [--fixed] [--device D] [--stdout]
// The scene arrives as a scene pack (tools/scene_pack.py) instead of scene JSON + planner config JSON (the
// reference ignores the config values anyway: PlannerConnect.cpp:67-70).
#include <chrono>
#include <cstdio>
#include <cstring>
#include <iostream>

#include "B200Backend.hpp"
#include "ResultsLog.hpp"
#include "SIRRTPlanner.hpp"

int main(int argc, char** argv) {
  if (argc < 4) {
    fprintf(stderr, "usage: %s <scene.pack> <result_dir> <seed> [--max-iters N] [--max-time S] [--no-edge-table] [--fixed] [--device D] [--stdout]\n", argv[0]);
    return 2;
  }
  const std::string pack_path = argv[1], out_dir = argv[2];
  const int seed = atoi(argv[3]);
  uint64_t max_iters = 0;
  double max_time = -1;
  bool edge_table = true, to_stdout = false;
  int parity = SIRRT_PARITY_REFERENCE_EXACT, device = 0;
  for (int i = 4; i < argc; i++) {
    if (!strcmp(argv[i], "--max-iters") && i + 1 < argc) max_iters = strtoull(argv[++i], nullptr, 10);
    else if (!strcmp(argv[i], "--max-time") && i + 1 < argc) max_time = atof(argv[++i]);
    else if (!strcmp(argv[i], "--no-edge-table")) edge_table = false;
    else if (!strcmp(argv[i], "--fixed")) parity = SIRRT_PARITY_FIXED;
    else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--stdout")) to_stdout = true;
  }
  using clk = std::chrono::steady_clock;
  auto ns = [](clk::time_point a, clk::time_point b) { return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(b - a).count(); };
  try {
    const auto t_start = clk::now();
    MDP::CollisionManager cm = MDP::CollisionManager::from_pack_file(pack_path, device, parity);
    const auto t_scene = clk::now();
    MDP::B200Backend be(cm, device, edge_table);
    MDP::MSIRRT::PlannerConnect<MDP::B200Backend> planner(be, seed);
    if (!planner.valid()) { fprintf(stderr, "sirrt_plan: %s\n", planner.error().c_str()); return 1; }
    if (max_iters) planner.set_max_iterations(max_iters);
    if (max_time > 0) planner.set_max_planning_time(max_time);
    const auto t_cfg = clk::now();
    const bool result = planner.solve();
    const auto t_end = clk::now();

    MDP::ResultsLog log;
    log.path_to_scene_json = pack_path; log.path_to_result_folder = out_dir;
    log.full_execution_time_ns = ns(t_start, t_end); log.planner_config_time_ns = ns(t_start, t_cfg);
    log.scene_init_time_ns = ns(t_start, t_scene);
    const MDP::MSIRRT::PlannerStats& st = planner.stats();
    log.set_counters(st, cm.link_count());
    log.algorithm_solving_time_ns = ns(t_cfg, t_end);
    log.has_result = result;
    log.start_tree_vertices = planner.get_start_tree_vertices_count();
    log.goal_tree_vertices = planner.get_goal_tree_vertices_count();
    if (result) {
      auto path = planner.get_final_path();
      if (!planner.check_path(path)) fprintf(stderr, "sirrt_plan: warning: check_path failed (PlannerConnect.cpp:934-954)\n");
      for (const auto& s : planner.get_result_path(0.0)) log.final_path.push_back({s.q, s.time});
      log.path_cost = log.final_path.back().time;
    }
    log.extra = {{"iterations", (double)st.iterations}, {"motion_checks", (double)st.n_motion_checks}, {"edge_tables", (double)be.n_tables()},
                 {"edge_cell_batches", (double)be.n_cell_batches()}, {"edge_table_lookups", (double)be.n_lookups()}};
    {
      const uint64_t* h = be.uses_histogram();
      const char* names[6] = {"edges_used_1", "edges_used_2", "edges_used_3", "edges_used_4_7", "edges_used_8_15", "edges_used_16_up"};
      for (int i = 0; i < 6; i++) log.extra.push_back({names[i], (double)h[i]});
    }
    if (to_stdout) std::cout << log.to_json() << std::endl;
    else {
      const std::string fn = log.save(std::to_string(std::chrono::system_clock::to_time_t(std::chrono::system_clock::now())), seed);
      printf("%s\n", fn.c_str());
    }
    return 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "sirrt_plan: %s\n", e.what());
    return 1;
  }
}
