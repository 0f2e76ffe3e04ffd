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
      note_edge_done();
      e_q0_.assign(q0, q0 + d); e_dir_.assign(dir, dir + d); e_steps_ = steps; e_uses_ = 0; e_have_table_ = false;
      e_states_.resize((size_t)n * d);
      for (int s = 0; s < n; s++)
        for (int j = 0; j < d; j++) e_states_[(size_t)s * d + j] = q0[j] + dir[j] * (double)s / (double)steps;  // PlannerConnect.cpp:885
    }
    e_uses_++;
    frames_.resize(n);
    for (int s = 0; s < n; s++) frames_[s] = (int32_t)(t0 + (t1 - t0) * (double)s / (double)steps);  // :886
    if (use_table_ && e_uses_ >= 2) {
      // windowed table: the scan moves every state's frame by about one per departure time, so the rows are only
      // computed for kWinTiles 32-frame tiles from the current frames on, in the direction the scan moves; the window
      // is rebuilt when a query leaves it
      const int W = cm_.words();
      bool covered = e_have_table_;
      for (int s = 0; s < n && covered; s++) {
        const int f = frames_[s];
        if (f >= 0 && f < T && (f < 32 * e_tlo_[s] || f >= 32 * e_thi_[s])) covered = false;
      }
      if (!covered) {
        const bool forward = frames_[0] >= e_prev_f0_;
        e_tlo_.resize(n); e_thi_.resize(n);
        for (int s = 0; s < n; s++) {
          const int tile = std::min(std::max(frames_[s], 0), T - 1) >> 5;
          const int lo = std::max(0, forward ? tile : tile - (kWinTiles - 1));
          e_tlo_[s] = lo; e_thi_[s] = std::min(W, lo + kWinTiles);
        }
        e_bits_.assign((size_t)n * W, 0u);
        if (W <= kWinTiles + 1) {
          cm_.collision_bits(e_states_, SIRRT_SEM_FRAME, e_bits_);
          for (int s = 0; s < n; s++) { e_tlo_[s] = 0; e_thi_[s] = W; }
        } else if (sirrt_collision_bits_window_f64(cm_.handle(), e_states_.data(), n, SIRRT_SEM_FRAME, e_tlo_.data(), kWinTiles, e_bits_.data()) != SIRRT_OK) {
          throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
        }
        e_have_table_ = true;
        n_tables_++;
      }
      e_prev_f0_ = frames_[0];
      for (int s = 0; s < n; s++) {
        const int f = frames_[s];
        if (f < 0 || f >= T) return s;  // the reference indexes unchecked (CollisionManager.cpp:369); we answer "collided"
        if ((e_bits_[(size_t)s * W + (f >> 5)] >> (f & 31)) & 1u) return s;
      }
      n_lookups_++;
      return -1;
    }
    e_prev_f0_ = frames_[0];
    cells_.assign(n, 0);
    if (sirrt_check_frames_batch_f64(cm_.handle(), e_states_.data(), frames_.data(), n, cells_.data()) != SIRRT_OK)
      throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
    n_cell_batches_++;
    for (int s = 0; s < n; s++) if (cells_[s]) return s;
    return -1;
  }

  int knn_new() {
    sirrt_knn* k = nullptr;
    while (pool_ && !pool_->empty()) {
      k = pool_->back(); pool_->pop_back();
      int dim = 0, dev = -1;
      if (sirrt_knn_info(k, &dim, nullptr, &dev) == SIRRT_OK && dim == dof() && dev == device_) {  // a pooled store of another robot / device is of no use
        sirrt_knn_host_buffers(k, 1);
        knn_.push_back(k);
        return (int)knn_.size() - 1;
      }
      sirrt_knn_destroy(k);
      k = nullptr;
    }
    // the reference reserves 1e6 vertices per tree (PlannerConnect.cpp:16-18); the store grows on demand instead
    if (sirrt_knn_create(dof(), 1 << 14, device_, &k) != SIRRT_OK) throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
    sirrt_knn_host_buffers(k, 1);  // the planner passes host vectors only
    knn_.push_back(k);
    return (int)knn_.size() - 1;
  }
  void knn_add(int tree, const double* p) { pf_valid_ = false; ck(sirrt_knn_add(knn_[tree], p, 1)); }
  // Optional hook of the planner template: the radius query that opens set_parent (PlannerConnect.cpp:414-416) depends on
  // the steered configuration only, so it is LAUNCHED here, before the planner's get_safe_intervals call (another object,
  // another stream: the two run side by side on the GPU), and collected by knn_radius below.  Same answers, one round
  // trip per iteration hidden.
  void knn_radius_prefetch(int tree, const double* q, double r2) {
    pf_valid_ = false;
    if (sirrt_knn_radius_begin(knn_[tree], q, r2) != SIRRT_OK) return;  // knn_radius asks synchronously
    pf_tree_ = tree; pf_r2_ = r2; pf_q_.assign(q, q + dof());
    pf_valid_ = true;
  }
  size_t knn_nearest(int tree, const double* q) {
    pf_valid_ = false;
    int64_t idx = -1;
    double d2 = 0;
    ck(sirrt_knn_nearest(knn_[tree], q, 1, &idx, &d2));
    return (size_t)idx;
  }
  void knn_radius(int tree, const double* q, double r2, std::vector<size_t>& out) {
    int64_t n = 0;
    rad_.resize(std::max<size_t>(rad_.size(), 256));
    if (pf_valid_) {
      pf_valid_ = false;
      if (pf_tree_ == tree && pf_r2_ == r2 && !memcmp(pf_q_.data(), q, sizeof(double) * dof())) {
        const int rc = sirrt_knn_radius_end(knn_[tree], rad_.data(), (int64_t)rad_.size(), &n);
        if (rc == SIRRT_OK && n <= (int64_t)rad_.size()) { out.assign(rad_.begin(), rad_.begin() + n); n_prefetched_++; return; }
        if (rc != SIRRT_OK && rc != SIRRT_E_OVERFLOW) ck(rc);
        rad_.resize(std::max<size_t>((size_t)n, rad_.size()));  // more hits than fit: ask again below
      }  // (a prefetch for another query is settled and dropped by the library)
    }
    for (;;) {
      const int rc = sirrt_knn_radius(knn_[tree], q, r2, rad_.data(), (int64_t)rad_.size(), &n);
      if (rc == SIRRT_E_OVERFLOW || n > (int64_t)rad_.size()) { rad_.resize((size_t)n); continue; }
      ck(rc);
      break;
    }
    out.assign(rad_.begin(), rad_.begin() + n);
  }

  // histogram of queries per distinct edge: 1, 2, 3, 4-7, 8-15, 16+ (what the "table on the 2nd use" rule is judged by)
  const uint64_t* uses_histogram() { note_edge_done(); e_uses_ = 0; return uses_hist_; }
  uint64_t n_tables() const { return n_tables_; }
  uint64_t n_cell_batches() const { return n_cell_batches_; }
  uint64_t n_lookups() const { return n_lookups_; }
  uint64_t n_prefetched_radius() const { return n_prefetched_; }

 private:
  void note_edge_done() {
    if (e_uses_ <= 0) return;
    const int b = e_uses_ == 1 ? 0 : e_uses_ == 2 ? 1 : e_uses_ == 3 ? 2 : e_uses_ < 8 ? 3 : e_uses_ < 16 ? 4 : 5;
    uses_hist_[b]++;
  }
  void ck(int rc) const { if (rc != SIRRT_OK) throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error()); }
  CollisionManager& cm_;
  int device_;
  bool use_table_;
  std::vector<sirrt_knn*>* pool_ = nullptr;
  std::vector<sirrt_knn*> knn_;
  std::vector<int64_t> rad_;
  bool pf_valid_ = false;   // a radius query launched by knn_radius_prefetch waits to be collected
  int pf_tree_ = -1;
  double pf_r2_ = 0;
  std::vector<double> pf_q_;
  uint64_t n_prefetched_ = 0;
  std::vector<double> e_q0_, e_dir_, e_states_;
  int e_steps_ = -1, e_uses_ = 0;
  bool e_have_table_ = false;
  std::vector<uint32_t> e_bits_;
  std::vector<int32_t> frames_, e_tlo_, e_thi_;  // e_tlo_/e_thi_: tiles [lo, hi) each state's table row is valid for
  int e_prev_f0_ = 0;
  static constexpr int kWinTiles = 3;            // >= 64 frames ahead of the query that built the window
  std::vector<uint8_t> cells_;
  uint64_t n_tables_ = 0, n_cell_batches_ = 0, n_lookups_ = 0, uses_hist_[6] = {0, 0, 0, 0, 0, 0};
};

}  // namespace MDP
