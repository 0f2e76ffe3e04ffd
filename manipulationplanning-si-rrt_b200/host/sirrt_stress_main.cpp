// sirrt_stress -- determinism under concurrency: many host threads, each with its own scene on the same GPU, issue the
// planner's call mix (get_safe_intervals, per-edge cell batches, edge tables, kNN add / nearest / radius with stores
// created and destroyed as the mass test does) on IDENTICAL inputs; every answer must equal the one a single thread
// computed alone.  Prints one JSON line with the number of mismatches per entry point.
//   sirrt_stress <scene.pack> <threads> <rounds> [device]
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

#include "CollisionManager.hpp"

struct Answers {
  std::vector<std::vector<std::pair<int, int>>> si;
  std::vector<std::vector<uint8_t>> cells;
  std::vector<std::vector<uint32_t>> tables;
  std::vector<int64_t> nearest;
  std::vector<std::vector<int64_t>> radius;
};

static void run(MDP::CollisionManager& cm, int device, int rounds, Answers& out) {
  const int dof = cm.dof(), T = cm.frame_count();
  const auto lim = cm.get_planned_robot_limits();
  std::mt19937_64 rng(12345);
  auto rnd_q = [&](std::vector<double>& q) {
    q.resize(dof);
    for (int j = 0; j < dof; j++) q[j] = std::uniform_real_distribution<double>(lim[j].first, lim[j].second)(rng);
  };
  for (int r = 0; r < rounds; r++) {
    sirrt_knn* k = nullptr;
    if (sirrt_knn_create(dof, 1 << 20, device, &k) != SIRRT_OK) throw std::runtime_error(sirrt_last_error());
    std::vector<double> q, q2;
    for (int it = 0; it < 12; it++) {
      rnd_q(q);
      sirrt_knn_add(k, q.data(), 1);
      rnd_q(q2);
      int64_t idx = -1; double d2 = 0;
      sirrt_knn_nearest(k, q2.data(), 1, &idx, &d2);
      out.nearest.push_back(idx);
      std::vector<int64_t> hits(256); int64_t n = 0;
      sirrt_knn_radius(k, q2.data(), 9.0, hits.data(), 256, &n);
      hits.resize((size_t)std::min<int64_t>(n, 256));
      out.radius.push_back(hits);
      out.si.push_back(cm.get_safe_intervals(q2));
      // one edge: cell batch, then the edge table
      const int steps = 8 + (it % 9);
      std::vector<double> states((size_t)(steps + 1) * dof);
      std::vector<int32_t> frames(steps + 1);
      for (int s = 0; s <= steps; s++) {
        for (int j = 0; j < dof; j++) states[(size_t)s * dof + j] = q[j] + (q2[j] - q[j]) * 0.3 * s / steps;
        frames[s] = (int)((it * 37 + s * 3) % T);
      }
      std::vector<uint8_t> cells(steps + 1);
      if (sirrt_check_frames_batch_f64(cm.handle(), states.data(), frames.data(), steps + 1, cells.data()) != SIRRT_OK) throw std::runtime_error(sirrt_last_error());
      out.cells.push_back(cells);
      std::vector<uint32_t> bits;
      cm.collision_bits(states, SIRRT_SEM_FRAME, bits);
      out.tables.push_back(bits);
    }
    sirrt_knn_destroy(k);
  }
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s <scene.pack> <threads> <rounds> [device]\n", argv[0]); return 2; }
  const int n_threads = atoi(argv[2]), rounds = atoi(argv[3]), device = argc > 4 ? atoi(argv[4]) : 0;
  try {
    Answers ref;
    {
      MDP::CollisionManager cm = MDP::CollisionManager::from_pack_file(argv[1], device);
      run(cm, device, rounds, ref);
    }
    std::atomic<long long> bad_si{0}, bad_cells{0}, bad_tables{0}, bad_nn{0}, bad_rad{0}, failed{0};
    auto worker = [&]() {
      try {
        MDP::CollisionManager cm = MDP::CollisionManager::from_pack_file(argv[1], device);
        for (int rep = 0; rep < 2; rep++) {  // the second pass reuses the scene, as consecutive tasks of a pack do
          Answers a;
          run(cm, device, rounds, a);
          for (size_t i = 0; i < ref.si.size(); i++) bad_si += a.si[i] != ref.si[i];
          for (size_t i = 0; i < ref.cells.size(); i++) bad_cells += a.cells[i] != ref.cells[i];
          for (size_t i = 0; i < ref.tables.size(); i++) bad_tables += a.tables[i] != ref.tables[i];
          for (size_t i = 0; i < ref.nearest.size(); i++) bad_nn += a.nearest[i] != ref.nearest[i];
          for (size_t i = 0; i < ref.radius.size(); i++) bad_rad += a.radius[i] != ref.radius[i];
        }
      } catch (const std::exception& e) { failed++; fprintf(stderr, "sirrt_stress: %s\n", e.what()); }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < n_threads; t++) th.emplace_back(worker);
    for (auto& x : th) x.join();
    printf("{\"threads\": %d, \"rounds\": %d, \"calls_per_thread\": %zu, \"mismatch\": {\"safe_intervals\": %lld, \"cell_batches\": %lld, \"edge_tables\": %lld, "
           "\"knn_nearest\": %lld, \"knn_radius\": %lld}, \"failed_threads\": %lld}\n",
           n_threads, rounds, 2 * ref.si.size(), bad_si.load(), bad_cells.load(), bad_tables.load(), bad_nn.load(), bad_rad.load(), failed.load());
    return (bad_si + bad_cells + bad_tables + bad_nn + bad_rad + failed) ? 1 : 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "sirrt_stress: %s\n", e.what());
    return 1;
  }
}
