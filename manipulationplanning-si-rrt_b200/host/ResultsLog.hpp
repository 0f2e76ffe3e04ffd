// ResultsLog.hpp -- planner result log in the reference's JSON schema:
// STRRT_Planner/src/config_read_writer/ResultsWriter.cpp:52-124 (top-level keys, file name
// "<dir>/MSIRRT_jump_planner_logs_<unix time>_<seed>.json") and MSIRRT/src/main.cpp:70-116
// ("final_planner_data": has_result, start/goal_tree_vertices, final_path [{time, robot_angles}], path_cost),
// so mass_test/final_analysis.ipynb and multiagent_eval/do_mass_test_multithread.py:167-222 keep working.
// Counters: those the reference derives from the planner's own calls and results are reproduced exactly
// (number_of_collision_checks, _get_safe_interval, _detected_collisions :374, _safe_intervals and
// _sum_of_safe_interval_frames :511-545, _forward_kinematics = one per state check + one per interval query :350,:492,
// _broadphase_get_safe_interval = links per interval query :495-497).  number_of_broadphase_collision_checks is written
// as links x state checks: the reference stops at the first colliding link (:357-377), the GPU evaluates every link of a
// state in one launch.  The two narrow-phase counters count coal's BVH-traversal callbacks (:604, :650), which have no
// counterpart here; the backend's own exact pair tests are reported there when the caller supplies them, else 0.
// The FK timers are part of the fused kernels and stay 0.
#pragma once
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

namespace MDP {

struct ResultsLog {
  std::string path_to_scene_json, path_to_config_json, path_to_result_folder, planner_type = "MSIRRT_jump";
  uint64_t full_execution_time_ns = 0, planner_config_time_ns = 0, algorithm_solving_time_ns = 0;
  uint64_t collision_check_time_ns = 0, get_safe_interval_time_ns = 0, scene_init_time_ns = 0;
  uint64_t number_of_collision_checks = 0, number_of_get_safe_interval = 0;
  uint64_t number_of_broadphase_collision_checks = 0, number_of_narrowphase_collision_checks = 0;
  uint64_t number_of_broadphase_get_safe_interval = 0, number_of_narrowphase_get_safe_interval = 0;
  uint64_t number_of_distances_checks = 0, number_of_forward_kinematics = 0;
  uint64_t number_of_detected_collisions = 0, number_of_safe_intervals = 0, number_of_sum_of_safe_interval_frames = 0;
  bool has_result = false;
  int start_tree_vertices = 0, goal_tree_vertices = 0;
  struct State { std::vector<double> q; double time; };
  std::vector<State> final_path;
  double path_cost = 0;
  std::vector<std::pair<std::string, double>> extra;  // backend counters (not in the reference schema), under "b200"

  // PlannerStats (SIRRTPlanner.hpp) -> the reference's counters; `links` = collision links of the planned robot
  template <class Stats>
  void set_counters(const Stats& st, int links) {
    algorithm_solving_time_ns = st.solve_ns; collision_check_time_ns = st.motion_ns; get_safe_interval_time_ns = st.safe_interval_ns;
    number_of_collision_checks = st.n_state_checks; number_of_get_safe_interval = st.n_safe_interval_calls;
    number_of_detected_collisions = st.n_detected_collisions;
    number_of_safe_intervals = st.n_safe_intervals; number_of_sum_of_safe_interval_frames = st.sum_safe_interval_frames;
    number_of_forward_kinematics = st.n_state_checks + st.n_safe_interval_calls;
    number_of_broadphase_collision_checks = (uint64_t)links * st.n_state_checks;
    number_of_broadphase_get_safe_interval = (uint64_t)links * st.n_safe_interval_calls;
  }

  static std::string num(double v) { char b[40]; snprintf(b, sizeof b, "%.17g", v); return b; }
  static std::string esc(const std::string& s) {
    std::string o;
    for (char c : s) { if (c == '"' || c == '\\') o += '\\'; o += c; }
    return o;
  }

  std::string to_json() const {
    std::ostringstream o;
    o << "{\"planner_type\":\"" << esc(planner_type) << "\",\"path_to_scene_json\":\"" << esc(path_to_scene_json) << "\",\"path_to_config_json\":\""
      << esc(path_to_config_json) << "\",\"full_execution_time_ns\":" << full_execution_time_ns << ",\"planner_config_time_ns\":" << planner_config_time_ns
      << ",\"algorithm_solving_time_ns\":" << algorithm_solving_time_ns << ",\"collision_check_time_ns\":" << collision_check_time_ns
      << ",\"distance_check_time_ns\":0,\"forward_kinematics_in_collision_check_time_ns\":0,\"forward_kinematics_in_distance_check_time_ns\":0"
      << ",\"total_forward_kinematics_time_ns\":0,\"get_safe_interval_time_ns\":" << get_safe_interval_time_ns
      << ",\"collision_check_init_time_ns\":" << scene_init_time_ns << ",\"safe_intervals_init_time_ns\":" << scene_init_time_ns
      << ",\"number_of_collision_checks\":" << number_of_collision_checks << ",\"number_of_get_safe_interval\":" << number_of_get_safe_interval
      << ",\"number_of_broadphase_collision_checks\":" << number_of_broadphase_collision_checks
      << ",\"number_of_narrowphase_collision_checks\":" << number_of_narrowphase_collision_checks
      << ",\"number_of_broadphase_get_safe_interval\":" << number_of_broadphase_get_safe_interval
      << ",\"number_of_narrowphase_get_safe_interval\":" << number_of_narrowphase_get_safe_interval
      << ",\"number_of_distances_checks\":" << number_of_distances_checks << ",\"number_of_forward_kinematics\":" << number_of_forward_kinematics
      << ",\"number_of_detected_collisions\":" << number_of_detected_collisions << ",\"number_of_safe_intervals\":" << number_of_safe_intervals
      << ",\"number_of_sum_of_safe_interval_frames\":" << number_of_sum_of_safe_interval_frames;
    o << ",\"final_planner_data\":{\"has_result\":" << (has_result ? "true" : "false") << ",\"start_tree_vertices\":" << start_tree_vertices
      << ",\"goal_tree_vertices\":" << goal_tree_vertices;
    if (has_result) {
      o << ",\"final_path\":[";
      for (size_t i = 0; i < final_path.size(); i++) {
        o << (i ? "," : "") << "{\"time\":" << num(final_path[i].time) << ",\"robot_angles\":[";
        for (size_t j = 0; j < final_path[i].q.size(); j++) o << (j ? "," : "") << num(final_path[i].q[j]);
        o << "]}";
      }
      o << "],\"path_cost\":" << num(path_cost);
    }
    o << "},\"path_results\":[]";
    if (!extra.empty()) {
      o << ",\"b200\":{";
      for (size_t i = 0; i < extra.size(); i++) o << (i ? "," : "") << "\"" << esc(extra[i].first) << "\":" << num(extra[i].second);
      o << "}";
    }
    o << "}";
    return o.str();
  }

  // ResultsWriter.cpp:119: <dir>/<planner>_planner_logs_<file_name>_<seed>.json
  std::string save(const std::string& file_name, int random_seed) const {
    const std::string fn = path_to_result_folder + "/" + planner_type + "_planner_logs_" + file_name + "_" + std::to_string(random_seed) + ".json";
    std::ofstream f(fn);
    f << to_json();
    return fn;
  }
};

}  // namespace MDP
