// B200Backend.hpp -- the planner-facing adapter over libsirrt_b200: MDP::CollisionManager (the drop-in shim of
// the reference class) for safe intervals, sirrt_knn_* for the trees' nearest / radius queries
// (MSIRRT/src/Tree.cpp:18-36, PlannerConnect.cpp:382-412), and the §7.3 edge table for motion validation:
//
//   PlannerConnect::set_parent scans departure times for ONE candidate edge (PlannerConnect.cpp:494-537,
//   709-745); every is_collision_motion of that scan interpolates the same steps+1 configurations and only the
//   frames move.  The first query of an edge is answered with one batched (state, frame) launch
//   (sirrt_check_frames_batch_f64); when the same edge comes back with another departure time the backend
//   computes the collision bit rows of its states over ALL frames once (sirrt_collision_bits_f64, per-frame
//   semantics) and answers this and every later departure time by table look-ups on the host.
#pragma once
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "CollisionManager.hpp"

namespace MDP {

class B200Backend {
 public:
  // knn_pool (optional, owned by the caller, one per host thread and device): vertex stores are taken from / returned to it
  // instead of being allocated and freed per planner -- a 2 x 50 MB cudaMalloc / cudaFree pair per task, the latter a
  // device-wide synchronisation that stalls every other planner thread sharing the GPU (mass test)
  explicit B200Backend(CollisionManager& cm, int device = 0, bool use_edge_table = true, std::vector<sirrt_knn*>* knn_pool = nullptr)
      : cm_(cm), device_(device), use_table_(use_edge_table), pool_(knn_pool) {}
  ~B200Backend() {
    for (sirrt_knn* k : knn_) {
      if (!k) continue;
      if (pool_ && sirrt_knn_clear(k) == SIRRT_OK) pool_->push_back(k);
      else sirrt_knn_destroy(k);
    }
  }
  static void destroy_pool(std::vector<sirrt_knn*>& pool) { for (sirrt_knn* k : pool) sirrt_knn_destroy(k); pool.clear(); }
  B200Backend(const B200Backend&) = delete;
  B200Backend& operator=(const B200Backend&) = delete;

  int dof() const { return cm_.dof(); }
  int frame_count() const { return cm_.frame_count(); }
  int fps() const { return cm_.fps(); }
  std::vector<std::pair<float, float>> get_planned_robot_limits() const { return cm_.get_planned_robot_limits(); }
  std::vector<double> start_configuration() const { return cm_.start_configuration(); }
  std::vector<double> end_configuration() const { return cm_.end_configuration(); }
  std::vector<std::pair<int, int>> get_safe_intervals(const std::vector<double>& q) { return cm_.get_safe_intervals(q); }

  int edge_first_collision(const double* q0, const double* dir, int steps, double t0, double t1) {
    const int d = dof(), n = steps + 1, T = frame_count();
    const bool same = steps == e_steps_ && e_q0_.size() == (size_t)d && !memcmp(e_q0_.data(), q0, d * sizeof(double)) &&
                      !memcmp(e_dir_.data(), dir, d * sizeof(double));
    if (!same) {
      e_q0_.assign(q0, q0 + d); e_dir_.assign(dir, dir + d); e_steps_ = steps; e_uses_ = 0; e_have_table_ = false;
      e_states_.resize((size_t)n * d);
      for (int s = 0; s < n; s++)
        for (int j = 0; j < d; j++) e_states_[(size_t)s * d + j] = q0[j] + dir[j] * (double)s / (double)steps;  // PlannerConnect.cpp:885
    }
    e_uses_++;
    frames_.resize(n);
    for (int s = 0; s < n; s++) frames_[s] = (int32_t)(t0 + (t1 - t0) * (double)s / (double)steps);  // :886
    if (use_table_ && e_uses_ >= 2 && !e_have_table_) {
      cm_.collision_bits(e_states_, SIRRT_SEM_FRAME, e_bits_);
      e_have_table_ = true;
      n_tables_++;
    }
    if (e_have_table_) {
      const int W = cm_.words();
      for (int s = 0; s < n; s++) {
        const int f = frames_[s];
        if (f < 0 || f >= T) return s;  // the reference indexes unchecked (CollisionManager.cpp:369); we answer "collided"
        if ((e_bits_[(size_t)s * W + (f >> 5)] >> (f & 31)) & 1u) return s;
      }
      n_lookups_++;
      return -1;
    }
    cells_.assign(n, 0);
    if (sirrt_check_frames_batch_f64(cm_.handle(), e_states_.data(), frames_.data(), n, cells_.data()) != SIRRT_OK)
      throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
    n_cell_batches_++;
    for (int s = 0; s < n; s++) if (cells_[s]) return s;
    return -1;
  }

  int knn_new() {
    sirrt_knn* k = nullptr;
    if (pool_ && !pool_->empty()) {
      k = pool_->back(); pool_->pop_back();
      knn_.push_back(k);
      return (int)knn_.size() - 1;
    }
    if (sirrt_knn_create(dof(), 1 << 20, device_, &k)  /* the reference reserves 1e6 vertices per tree (PlannerConnect.cpp:16-18) */ != SIRRT_OK) throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
    knn_.push_back(k);
    return (int)knn_.size() - 1;
  }
  void knn_add(int tree, const double* p) { ck(sirrt_knn_add(knn_[tree], p, 1)); }
  size_t knn_nearest(int tree, const double* q) {
    int64_t idx = -1;
    double d2 = 0;
    ck(sirrt_knn_nearest(knn_[tree], q, 1, &idx, &d2));
    return (size_t)idx;
  }
  void knn_radius(int tree, const double* q, double r2, std::vector<size_t>& out) {
    int64_t n = 0;
    rad_.resize(std::max<size_t>(rad_.size(), 256));
    for (;;) {
      const int rc = sirrt_knn_radius(knn_[tree], q, r2, rad_.data(), (int64_t)rad_.size(), &n);
      if (rc == SIRRT_E_OVERFLOW || n > (int64_t)rad_.size()) { rad_.resize((size_t)n); continue; }
      ck(rc);
      break;
    }
    out.assign(rad_.begin(), rad_.begin() + n);
  }

  uint64_t n_tables() const { return n_tables_; }
  uint64_t n_cell_batches() const { return n_cell_batches_; }
  uint64_t n_lookups() const { return n_lookups_; }

 private:
  void ck(int rc) const { if (rc != SIRRT_OK) throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error()); }
  CollisionManager& cm_;
  int device_;
  bool use_table_;
  std::vector<sirrt_knn*>* pool_ = nullptr;
  std::vector<sirrt_knn*> knn_;
  std::vector<int64_t> rad_;
  std::vector<double> e_q0_, e_dir_, e_states_;
  int e_steps_ = -1, e_uses_ = 0;
  bool e_have_table_ = false;
  std::vector<uint32_t> e_bits_;
  std::vector<int32_t> frames_;
  std::vector<uint8_t> cells_;
  uint64_t n_tables_ = 0, n_cell_batches_ = 0, n_lookups_ = 0;
};

}  // namespace MDP
