// SIRRTPlanner.hpp -- SI-RRT (bidirectional RRT-Connect over (configuration, safe interval) vertices) re-stated
// above the collision-checker boundary, following MSIRRT/src/PlannerConnect.cpp of the reference line by line
// in behaviour (RNG consumption order, set_parent departure-time scan and "jump" logic, warm-up, connect,
// prune) but without Eigen / nanoflann / rapidjson.  SURVEY.md §8 row f1.
//
// The planner is a template over a Backend so that the SAME host logic runs
//   * on libsirrt_b200 (B200Backend below: MDP::CollisionManager shim + sirrt_knn_* + the §7.3 edge table), and
//   * on the CPU checker in tests/ (test infrastructure only; the main() for it lives under tests/),
// which is how "identical plans" is checked: same seed, two backends, waypoint-by-waypoint equality.
//
// Backend concept:
//   int dof(); int frame_count(); int fps();
//   std::vector<std::pair<float,float>> get_planned_robot_limits();
//   std::vector<double> start_configuration(), end_configuration();
//   std::vector<std::pair<int,int>> get_safe_intervals(const std::vector<double>& q);     CollisionManager.cpp:481
//   int edge_first_collision(const double* q0, const double* dir, int steps, double t0, double t1);
//        first s in 0..steps with check_collision_frame(q0 + dir*s/steps, (int)(t0 + (t1-t0)*s/steps)), else -1
//   int  knn_new(); void knn_add(int tree, const double* p);
//   size_t knn_nearest(int tree, const double* q); void knn_radius(int tree, const double* q, double r2, std::vector<size_t>& out);
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <deque>
#include <random>
#include <string>
#include <utility>
#include <vector>

namespace MDP {
namespace MSIRRT {

using Coord = std::vector<double>;

inline double coord_dist(const Coord& a, const Coord& b) {
  double s = 0;
  for (size_t i = 0; i < a.size(); i++) { const double d = a[i] - b[i]; s += d * d; }
  return std::sqrt(s);
}

struct Vertex {  // MSIRRT/include/Vertex.hpp:8-32
  Coord coords;
  std::pair<int, int> safe_interval;
  Vertex* parent = nullptr;
  double arrival_time = -1;
  double departure_from_parent_time = -1;
  int tree_id = -1;
};

struct Tree {  // MSIRRT/include/Tree.hpp, Tree.cpp:18-36 (the kd-tree lives in the backend)
  std::string tree_name;
  int tree_idx = -1;
  int knn = -1;
  std::deque<Vertex> array_of_vertices;  // deque: stable addresses like the reference's heap vertices
};

struct PathState { Coord q; double time; };  // RobotObstacleJsonInfo.hpp:17-22

struct PlannerStats {
  uint64_t n_safe_interval_calls = 0, n_motion_checks = 0, n_state_checks = 0, iterations = 0;
  // ResultsWriter counters the planner can reproduce exactly (CollisionManager.cpp:374, :511-545)
  uint64_t n_detected_collisions = 0, n_safe_intervals = 0, sum_safe_interval_frames = 0;
  uint64_t safe_interval_ns = 0, motion_ns = 0, solve_ns = 0;
};

template <class Backend>
class PlannerConnect {
 public:
  // PlannerConnect.cpp:8-96
  PlannerConnect(Backend& be, int random_seed) : be_(be), probability_gen_(0.0, 1.0), gen_(random_seed) {
    dof_ = be.dof();
    frame_count_ = be.frame_count();
    fps_ = be.fps();
    start_tree_.tree_name = "start_tree"; start_tree_.tree_idx = 0; start_tree_.knn = be.knn_new();
    goal_tree_.tree_name = "goal_tree"; goal_tree_.tree_idx = 1; goal_tree_.knn = be.knn_new();
    robot_limits_ = be.get_planned_robot_limits();
    const Coord start = be.start_configuration(), goal = be.end_configuration();
    auto start_si = safe_intervals(start);
    if (start_si.empty() || start_si[0].first != 0) { valid_ = false; error_ = "start configuration is not collision-free at frame 0"; return; }
    add_vertex(start_tree_, start, start_si[0], nullptr, -1, 0);
    root_node_ = &start_tree_.array_of_vertices[0];
    root_node_->arrival_time = 0;
    goal_coords_ = goal;
    goal_safe_intervals_ = safe_intervals(goal);
    if (goal_safe_intervals_.empty()) { valid_ = false; error_ = "goal configuration has no safe interval"; return; }
    for (const auto& si : goal_safe_intervals_) add_vertex(goal_tree_, goal_coords_, si, nullptr, -1, si.second);
    current_tree_ = &goal_tree_;
    other_tree_ = &start_tree_;
    max_planning_time_ = 20;  // :67-70 (hard-coded in the reference)
    planner_range_ = 1;
    vmax_ = 3.1415;
    if (const char* env = std::getenv("MSIRRT_PLANNER_RANGE")) { const double v = std::atof(env); if (v > 0) planner_range_ = v; }
    if (const char* env = std::getenv("MSIRRT_RADIUS_FACTOR")) { const double v = std::atof(env); if (v > 0) radius_factor_ = v; }
    if (const char* env = std::getenv("MSIRRT_MAX_DEP_TIME")) { const int v = std::atoi(env); if (v >= 0) max_dep_time_for_safe_int_check_ = v; }
  }

  bool valid() const { return valid_; }
  const std::string& error() const { return error_; }
  // extension for reproducible tests: stop after this many main-loop iterations (0 = wall clock only)
  void set_max_iterations(uint64_t n) { max_iterations_ = n; }
  void set_max_planning_time(double s) { max_planning_time_ = s; }
  const PlannerStats& stats() const { return stats_; }
  int get_start_tree_vertices_count() const { return (int)start_tree_.array_of_vertices.size(); }
  int get_goal_tree_vertices_count() const { return (int)goal_tree_.array_of_vertices.size(); }

  // PlannerConnect.cpp:105-150
  bool solve() {
    if (!valid_) return false;
    solver_start_time_ = std::chrono::steady_clock::now();
    if (root_node_->safe_interval.second < frame_count_ / 5) warmup_start_tree();
    while (check_planner_termination_condition() && !goal_reached_) {
      stats_.iterations++;
      Coord coord_rand = get_random_configuration();
      if (!extend(coord_rand)) continue;
      hint_radius(coord_rand, *current_tree_);
      auto si = safe_intervals(coord_rand);
      std::vector<Vertex*> new_nodes = set_parent(coord_rand, si);
      if (new_nodes.empty()) { swap_trees(); continue; }
      connect_trees(coord_rand, new_nodes);
      swap_trees();
    }
    stats_.solve_ns = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - solver_start_time_).count();
    return goal_reached_;
  }

  // PlannerConnect.cpp:302-318
  std::vector<Vertex*> get_final_path() const {
    std::vector<Vertex*> result;
    if (!goal_reached_) return result;
    for (Vertex* cur = finish_node_; cur; cur = cur->parent) result.push_back(cur);
    std::reverse(result.begin(), result.end());
    return result;
  }

  // MSIRRT/src/main.cpp:81-110: waypoints with waiting states made explicit, times in seconds
  std::vector<PathState> get_result_path(double start_time_s) const {
    std::vector<PathState> out;
    Vertex* prev = nullptr;
    for (Vertex* st : get_final_path()) {
      if (prev && st->parent && st->departure_from_parent_time != st->parent->arrival_time)
        out.push_back({prev->coords, st->departure_from_parent_time / (double)fps_ + start_time_s});
      out.push_back({st->coords, st->arrival_time / (double)fps_ + start_time_s});
      prev = st;
    }
    return out;
  }

  // PlannerConnect.cpp:934-954
  bool check_path(const std::vector<Vertex*>& path) {
    for (size_t i = 0; i + 1 < path.size(); i++) {
      Coord cc;
      double a = path[i]->arrival_time, d = path[i + 1]->departure_from_parent_time, b = path[i + 1]->arrival_time;
      if (is_collision_motion(path[i]->coords, path[i]->coords, a, d, nullptr) &&
          is_collision_motion(path[i]->coords, path[i + 1]->coords, d, b, nullptr)) return false;
      if (!(path[i]->safe_interval.first <= a && path[i]->safe_interval.second >= a)) return false;
      if (!(path[i]->safe_interval.first <= d && path[i]->safe_interval.second >= d)) return false;
    }
    return true;
  }

 private:
  std::vector<std::pair<int, int>> safe_intervals(const Coord& q) {
    const auto t0 = std::chrono::steady_clock::now();
    auto r = be_.get_safe_intervals(q);
    stats_.safe_interval_ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    stats_.n_safe_interval_calls++;
    stats_.n_safe_intervals += r.size();
    for (const auto& iv : r) stats_.sum_safe_interval_frames += (uint64_t)(iv.second - iv.first + 1);
    return r;
  }

  // Tree.cpp:18-36
  Vertex* add_vertex(Tree& tree, const Coord& q, std::pair<int, int> si, Vertex* parent, double departure_time, double arrival_time) {
    Vertex v;
    v.coords = q; v.safe_interval = si; v.parent = parent; v.arrival_time = arrival_time;
    v.departure_from_parent_time = departure_time; v.tree_id = tree.tree_idx;
    tree.array_of_vertices.push_back(std::move(v));
    be_.knn_add(tree.knn, tree.array_of_vertices.back().coords.data());
    return &tree.array_of_vertices.back();
  }

  // PlannerConnect.cpp:320-327 (whole seconds, like the reference's duration_cast<seconds>)
  bool check_planner_termination_condition() const {
    if (max_iterations_ && stats_.iterations >= max_iterations_) return false;
    const auto el = std::chrono::duration_cast<std::chrono::seconds>(std::chrono::steady_clock::now() - solver_start_time_).count();
    return !goal_reached_ && (double)el < max_planning_time_;
  }

  // PlannerConnect.cpp:346-359: one uniform draw per joint, in joint order
  Coord get_random_configuration() {
    Coord r(dof_);
    for (int j = 0; j < dof_; j++)
      r[j] = probability_gen_(gen_) * (robot_limits_[j].second - robot_limits_[j].first) + robot_limits_[j].first;
    return r;
  }

  // PlannerConnect.cpp:361-380.  QUIRK 11: the guard `delta <= (1 / 100000000)` is integer division = 0.
  bool extend(Coord& coords_of_new) {
    const Vertex* q_nearest = &current_tree_->array_of_vertices[be_.knn_nearest(current_tree_->knn, coords_of_new.data())];
    Coord delta_vector(dof_);
    double s = 0;
    for (int j = 0; j < dof_; j++) { delta_vector[j] = coords_of_new[j] - q_nearest->coords[j]; s += delta_vector[j] * delta_vector[j]; }
    const double delta = std::sqrt(s);
    if (delta <= 0) return false;
    if (delta < planner_range_) return true;
    for (int j = 0; j < dof_; j++) coords_of_new[j] = q_nearest->coords[j] + (delta_vector[j] / delta) * planner_range_;
    return true;
  }

  void swap_trees() { std::swap(current_tree_, other_tree_); }

  // PlannerConnect.cpp:152-189
  void warmup_start_tree() {
    current_tree_ = &start_tree_;
    other_tree_ = &goal_tree_;
    const double old_range = planner_range_;
    planner_range_ = 1.5 * root_node_->safe_interval.second / fps_ * vmax_;  // :158, evaluated left to right in double
    for (int i = 0; i < 25; i++) {
      Coord coord_rand = get_random_configuration();
      if (!extend(coord_rand)) continue;
      hint_radius(coord_rand, *current_tree_);
      auto si = safe_intervals(coord_rand);
      std::vector<Vertex*> new_nodes = set_parent(coord_rand, si);
      if (new_nodes.empty() && was_static_obstacle_) {
        coord_rand = last_valid_coord_;
        si = safe_intervals(coord_rand);
        new_nodes = set_parent(coord_rand, si);
      }
    }
    planner_range_ = old_range;
  }

  // PlannerConnect.cpp:197-281
  bool connect_trees(Coord& coord_rand, const std::vector<Vertex*>& another_tree_new_nodes) {
    const Coord original_coord_rand = coord_rand;
    while (check_planner_termination_condition()) {
      coord_rand = original_coord_rand;
      swap_trees();
      const bool is_ok = extend(coord_rand);
      swap_trees();
      if (!is_ok) break;
      hint_radius(coord_rand, *other_tree_);  // set_parent below runs on the swapped tree
      auto si = safe_intervals(coord_rand);
      swap_trees();
      std::vector<Vertex*> new_nodes = set_parent(coord_rand, si);
      swap_trees();
      if (new_nodes.empty()) break;
      if (original_coord_rand == coord_rand) {  // reached the first tree's new configuration
        for (Vertex* node : new_nodes) {
          for (Vertex* other : another_tree_new_nodes) {
            if (node->coords == other->coords && node->safe_interval == other->safe_interval) {
              if (node->tree_id == 1) {  // goal tree
                if (node->arrival_time >= other->arrival_time) { goal_reached_ = true; goal_nodes_ = {other, node}; prune_goal_tree(); }
              } else {
                if (node->arrival_time <= other->arrival_time) { goal_reached_ = true; goal_nodes_ = {node, other}; prune_goal_tree(); }
              }
              return true;
            }
          }
          if (goal_reached_) break;
        }
        break;
      }
      coord_rand = original_coord_rand;
    }
    return false;
  }

  // Backends that can overlap the radius query of the coming set_parent (same q, same r2, same tree: :416) with the
  // safe-interval call in between offer knn_radius_prefetch; for the others this is a no-op.  Answers are unaffected.
  template <typename B> static auto hint_radius_impl(B& be, int knn, const double* q, double r2, int) -> decltype(be.knn_radius_prefetch(knn, q, r2), void()) {
    be.knn_radius_prefetch(knn, q, r2);
  }
  template <typename B> static void hint_radius_impl(B&, int, const double*, double, long) {}
  void hint_radius(const Coord& q, Tree& tree) { hint_radius_impl(be_, tree.knn, q.data(), radius_factor_ * planner_range_ * planner_range_, 0); }

  // PlannerConnect.cpp:394-412: strict dist^2 < r2, then the caller sorts
  std::vector<Vertex*> get_nearest_node_by_radius(const Coord& q, double r2, Tree& tree) {
    radius_scratch_.clear();
    be_.knn_radius(tree.knn, q.data(), r2, radius_scratch_);
    std::vector<Vertex*> out;
    out.reserve(radius_scratch_.size());
    for (size_t i : radius_scratch_) out.push_back(&tree.array_of_vertices[i]);
    return out;
  }

  double time_between(const Coord& a, const Coord& b) const { return coord_dist(a, b) * (double)fps_ / vmax_; }

  // PlannerConnect.cpp:862-897 (collision_coord != nullptr) and :899-932 (nullptr).
  // Returns true on collision; *collision_coord = first colliding interpolated state.
  bool is_collision_motion(const Coord& start_coords, const Coord& end_coords, double start_time, double end_time, Coord* collision_coord) {
    const auto t0 = std::chrono::steady_clock::now();
    stats_.n_motion_checks++;
    bool res = false;
    if (collision_coord) last_valid_coord_ = start_coords;
    if (start_time >= end_time) res = true;
    else {
      Coord dir(dof_);
      double s = 0;
      for (int j = 0; j < dof_; j++) { dir[j] = end_coords[j] - start_coords[j]; s += dir[j] * dir[j]; }
      const double norm = std::sqrt(s);
      if ((norm * (double)fps_ / (end_time - start_time)) > (vmax_ + 0.001)) res = true;  // speed gate :872
      else {
        const int steps = std::max({(int)(norm / 0.1), (int)(end_time - start_time + 0.5), 1});  // :879
        // states q_s = start + dir*s/steps at frames (int)(t0 + (t1-t0)*s/steps), s = 0..steps (:883-886); the backend
        // answers with the first colliding step (per-state early exit on the CPU, one batched launch / the
        // cached edge table on the GPU)
        const int hit = be_.edge_first_collision(start_coords.data(), dir.data(), steps, start_time, end_time);
        stats_.n_state_checks += (uint64_t)(hit < 0 ? steps + 1 : hit + 1);
        auto state = [&](int step, Coord& out) {
          out.resize(dof_);
          for (int j = 0; j < dof_; j++) out[j] = start_coords[j] + dir[j] * (double)step / (double)steps;
        };
        if (hit >= 0) {
          res = true;
          stats_.n_detected_collisions++;  // the one state check that returned true (CollisionManager.cpp:372-376)
          if (collision_coord) { state(hit, *collision_coord); if (hit > 0) state(hit - 1, last_valid_coord_); }
        } else if (collision_coord) state(steps, last_valid_coord_);
      }
    }
    stats_.motion_ns += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    return res;
  }

  // PlannerConnect.cpp:414-843
  std::vector<Vertex*> set_parent(Coord& coord_rand, const std::vector<std::pair<int, int>>& safe_intervals_of_coord_rand) {
    was_static_obstacle_ = false;
    std::vector<Vertex*> nearest_nodes = get_nearest_node_by_radius(coord_rand, radius_factor_ * planner_range_ * planner_range_, *current_tree_);
    std::vector<Vertex*> added_vertices;
    const int mdt = max_dep_time_for_safe_int_check_;

    if (current_tree_ == &start_tree_) {
      const double time_to_goal = time_between(coord_rand, goal_coords_);
      // earliest possible reach time first (:433-443)
      std::sort(nearest_nodes.begin(), nearest_nodes.end(), [&](const Vertex* a, const Vertex* b) {
        return a->arrival_time + time_between(coord_rand, a->coords) < b->arrival_time + time_between(coord_rand, b->coords);
      });
      for (const std::pair<int, int>& safe_int : safe_intervals_of_coord_rand) {
        if ((safe_int.first + time_to_goal) >= frame_count_) continue;  // the goal cannot be reached from that interval (:456)
        bool found_parent = false;
        for (Vertex* cand : nearest_nodes) {
          const double time_to_node = time_between(coord_rand, cand->coords);
          int col_first = (int)(std::max(cand->arrival_time, (double)safe_int.first - time_to_node) + mdt);  // :470-471
          int col_second = col_first - 1;
          if (time_to_node < 1) { found_parent = true; break; }                       // duplicate (:475)
          if (cand->arrival_time + time_to_node > safe_int.second) break;              // sorted: nothing later fits (:480)
          if (cand->safe_interval.second + time_to_node < safe_int.first) continue;    // no overlap (:486)
          const Coord& start_coords = cand->coords;
          for (double departure_time = std::max(cand->arrival_time, (double)safe_int.first - time_to_node);
               departure_time <= std::min((double)cand->safe_interval.second, (double)safe_int.second - time_to_node); departure_time += 1) {
            double arrival_time = departure_time + time_to_node;
            if (arrival_time > safe_int.second) {  // rounding fix (:500-509)
              departure_time -= arrival_time - safe_int.second;
              if (departure_time < cand->arrival_time) continue;
              arrival_time = safe_int.second;
            }
            if (arrival_time < safe_int.first) arrival_time = safe_int.first;  // slowing down (:511-516)
            Coord collision_coord;
            if (!is_collision_motion(start_coords, coord_rand, departure_time, arrival_time, &collision_coord)) {
              added_vertices.push_back(add_vertex(*current_tree_, coord_rand, safe_int, cand, departure_time, arrival_time));
              found_parent = true;
              break;
            } else if ((departure_time > col_second || departure_time > col_first + mdt) && departure_time < arrival_time) {  // :540
              if (collision_coord.empty()) continue;  // a gate (:865,:872) rejected the motion: the reference reads an unset vector here; we try the next departure
              const auto col_si = safe_intervals(collision_coord);  // :544
              if (col_si.empty()) { was_static_obstacle_ = true; break; }
              bool no_overlaps_found = true;
              for (const std::pair<int, int>& csi : col_si) {
                const double t_cand_col = time_between(cand->coords, collision_coord);
                const double t_col_rand = time_between(coord_rand, collision_coord);
                const double pr_par_low = departure_time + t_cand_col;
                const double pr_par_high = cand->safe_interval.second + t_cand_col;
                const double pr_ran_low = safe_int.first - t_col_rand;
                const double pr_ran_high = safe_int.second - t_col_rand;
                if (csi.second < pr_par_low) continue;
                if (csi.first > std::min(pr_par_high, pr_ran_high)) break;
                const double ov_low = std::max(pr_par_low, (double)csi.first), ov_high = std::min(pr_par_high, (double)csi.second);
                if (ov_low > ov_high) continue;
                const double res_low = std::max(ov_low, pr_ran_low), res_high = std::min(ov_high, pr_ran_high);
                if (res_low > res_high) continue;
                no_overlaps_found = false;
                if ((res_low - t_cand_col - departure_time) >= 1) {  // jump (:607-612)
                  departure_time = res_low - t_cand_col - 1;
                  col_first = (int)(departure_time + 1);
                  col_second = (int)(res_high - t_cand_col);
                }
              }
              if (no_overlaps_found) { was_static_obstacle_ = true; break; }
            }
          }
          if (found_parent) { was_static_obstacle_ = false; break; }
        }
      }
    } else {
      const double time_to_start = time_between(coord_rand, root_node_->coords);
      // latest possible departure first (:651-660)
      std::sort(nearest_nodes.begin(), nearest_nodes.end(), [&](const Vertex* a, const Vertex* b) {
        return a->arrival_time - time_between(coord_rand, a->coords) > b->arrival_time - time_between(coord_rand, b->coords);
      });
      for (const std::pair<int, int>& safe_int : safe_intervals_of_coord_rand) {
        if ((safe_int.second - time_to_start) < 0) continue;  // :664
        bool found_parent = false;
        for (Vertex* cand : nearest_nodes) {
          const double time_to_node = time_between(coord_rand, cand->coords);
          int col_second = (int)(std::min(cand->arrival_time, (double)safe_int.second + time_to_node) - mdt);  // :675-676
          int col_first = col_second - 1;
          if (time_to_node < 1) { found_parent = true; break; }
          if (cand->arrival_time - time_to_node < safe_int.first) break;
          if (cand->safe_interval.first - time_to_node > safe_int.second) continue;
          const Coord& start_coords = cand->coords;
          for (double departure_time = std::min(cand->arrival_time, (double)safe_int.second + time_to_node);
               departure_time >= std::max((double)cand->safe_interval.first, (double)safe_int.first + time_to_node); departure_time -= 1) {
            double arrival_time = departure_time - time_to_node;  // travelling into the past
            if (arrival_time > safe_int.second) arrival_time = safe_int.second;
            if (arrival_time < safe_int.first) {
              departure_time += safe_int.first - arrival_time;
              if (departure_time < cand->arrival_time) continue;
              arrival_time = departure_time - time_to_node;
            }
            Coord collision_coord;
            if (!is_collision_motion(coord_rand, start_coords, arrival_time, departure_time, &collision_coord)) {
              added_vertices.push_back(add_vertex(*current_tree_, coord_rand, safe_int, cand, departure_time, arrival_time));
              found_parent = true;
              break;
            } else if ((departure_time < col_first || departure_time < col_second - mdt) && departure_time > arrival_time) {  // :748
              if (collision_coord.empty()) continue;
              auto col_si = safe_intervals(collision_coord);  // :751
              if (col_si.empty()) { was_static_obstacle_ = true; break; }
              bool no_overlaps_found = true;
              std::reverse(col_si.begin(), col_si.end());
              for (const std::pair<int, int>& csi : col_si) {
                const double t_cand_col = time_between(cand->coords, collision_coord);
                const double t_col_rand = time_between(coord_rand, collision_coord);
                const double pr_par_low = cand->safe_interval.first - t_cand_col;
                const double pr_par_high = departure_time - t_cand_col;
                const double pr_ran_low = safe_int.first + t_col_rand;
                const double pr_ran_high = safe_int.second + t_col_rand;
                if (csi.first > pr_par_high) continue;
                if (csi.second < std::min(pr_par_low, pr_ran_low)) break;
                const double ov_low = std::max(pr_par_low, (double)csi.first), ov_high = std::min(pr_par_high, (double)csi.second);
                if (ov_low > ov_high) continue;
                const double res_low = std::max(ov_low, pr_ran_low), res_high = std::min(ov_high, pr_ran_high);
                if (res_low > res_high) continue;
                no_overlaps_found = false;
                if ((departure_time - (res_high + t_cand_col)) > 1) {  // jump (:806-811)
                  departure_time = res_high + t_cand_col + 1;
                  col_first = (int)(res_low + t_cand_col);
                  col_second = (int)(departure_time - 1);
                }
              }
              if (no_overlaps_found) { was_static_obstacle_ = true; break; }
            }
          }
          if (found_parent) break;
        }
      }
    }
    return added_vertices;
  }

  // PlannerConnect.cpp:956-995: re-validate the goal branch forward in time and append it to the start tree
  void prune_goal_tree() {
    Vertex* start_tree_node = goal_nodes_.first;
    Vertex* goal_tree_node_child = goal_nodes_.second;
    Vertex* goal_tree_node = goal_tree_node_child->parent;
    while (goal_tree_node) {
      const double time_to_node = time_between(start_tree_node->coords, goal_tree_node->coords);
      const Coord start_coords = start_tree_node->coords;
      bool found_dep_time = false;
      for (double departure_time = std::max(start_tree_node->arrival_time, (double)goal_tree_node->safe_interval.first - time_to_node);
           departure_time <= goal_tree_node_child->arrival_time; departure_time += 1) {
        const double arrival_time = departure_time + time_to_node;
        if (!is_collision_motion(start_coords, goal_tree_node->coords, departure_time, arrival_time, nullptr)) {
          start_tree_node = add_vertex(start_tree_, goal_tree_node->coords, goal_tree_node->safe_interval, start_tree_node, departure_time, arrival_time);
          found_dep_time = true;
          break;
        }
      }
      if (!found_dep_time)  // :984-988 keeps the goal-tree timing
        start_tree_node = add_vertex(start_tree_, goal_tree_node->coords, goal_tree_node->safe_interval, start_tree_node,
                                     goal_tree_node_child->arrival_time, goal_tree_node_child->departure_from_parent_time);
      goal_tree_node_child = goal_tree_node;
      goal_tree_node = goal_tree_node->parent;
    }
    finish_node_ = start_tree_node;
  }

  Backend& be_;
  std::uniform_real_distribution<> probability_gen_;
  std::mt19937 gen_;
  int dof_ = 0, frame_count_ = 0, fps_ = 0;
  Tree start_tree_, goal_tree_;
  Tree* current_tree_ = nullptr;
  Tree* other_tree_ = nullptr;
  Vertex* root_node_ = nullptr;
  Vertex* finish_node_ = nullptr;
  Coord goal_coords_;
  std::vector<std::pair<int, int>> goal_safe_intervals_;
  std::vector<std::pair<float, float>> robot_limits_;
  std::pair<Vertex*, Vertex*> goal_nodes_{nullptr, nullptr};
  double max_planning_time_ = 20, vmax_ = 3.1415, planner_range_ = 1, radius_factor_ = 4.0;
  int max_dep_time_for_safe_int_check_ = 0;
  uint64_t max_iterations_ = 0;
  bool goal_reached_ = false, was_static_obstacle_ = false, valid_ = true;
  std::string error_;
  Coord last_valid_coord_;
  std::vector<size_t> radius_scratch_;
  std::chrono::time_point<std::chrono::steady_clock> solver_start_time_;
  PlannerStats stats_;
};

}  // namespace MSIRRT
}  // namespace MDP
