// sirrt_latency_main.cpp -- per-call latency of the planner-in-the-loop entry points, measured from C++ over the
// C ABI (no Python in the loop): the numbers the SI-RRT plan time is made of.
//   sirrt_latency <scene.pack> [n_calls]
// Prints one JSON object (microseconds per call, mean over n_calls after warm-up).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "CollisionManager.hpp"

static double now_us() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
template <typename F> static double bench(F f, int n) {
  for (int i = 0; i < 20; i++) f(i);
  const double t = now_us();
  for (int i = 0; i < n; i++) f(i + 20);
  return (now_us() - t) / n;
}

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s <scene.pack> [n_calls]\n", argv[0]); return 2; }
  const int n = argc > 2 ? atoi(argv[2]) : 500;
  try {
    MDP::CollisionManager cm = MDP::CollisionManager::from_pack_file(argv[1]);
    const int dof = cm.dof(), T = cm.frame_count();
    const auto lim = cm.get_planned_robot_limits();
    std::mt19937_64 rng(7);
    std::vector<std::vector<double>> qs(64, std::vector<double>(dof));
    for (auto& q : qs)
      for (int j = 0; j < dof; j++) q[j] = std::uniform_real_distribution<double>(lim[j].first, lim[j].second)(rng);
    const double si = bench([&](int i) { cm.get_safe_intervals(qs[i & 63]); }, n);
    const double cf = bench([&](int i) { int f = (i * 37) % T; cm.check_collision_frame(qs[i & 63], f); }, n);
    std::vector<double> states(32 * (size_t)dof);
    std::vector<int32_t> frames(32);
    std::vector<uint8_t> cells(32);
    for (int s = 0; s < 32; s++) { for (int j = 0; j < dof; j++) states[(size_t)s * dof + j] = qs[s][j]; frames[s] = (s * 53) % T; }
    const double cb = bench([&](int) { sirrt_check_frames_batch_f64(cm.handle(), states.data(), frames.data(), 32, cells.data()); }, n);
    std::vector<uint32_t> bits;
    const double et = bench([&](int) { cm.collision_bits(states, SIRRT_SEM_FRAME, bits); }, n / 2);
    sirrt_knn* k = nullptr;
    if (sirrt_knn_create(dof, 1 << 20, 0, &k) != SIRRT_OK) throw std::runtime_error(sirrt_last_error());
    std::vector<double> pts(5000 * (size_t)dof);
    for (double& x : pts) x = std::uniform_real_distribution<double>(-3, 3)(rng);
    sirrt_knn_add(k, pts.data(), 5000);
    int64_t idx = 0, cnt = 0;
    double d2 = 0;
    std::vector<int64_t> hits(8192);
    const double kn = bench([&](int i) { sirrt_knn_nearest(k, qs[i & 63].data(), 1, &idx, &d2); }, n);
    const double kr = bench([&](int i) { sirrt_knn_radius(k, qs[i & 63].data(), 4.0, hits.data(), (int64_t)hits.size(), &cnt); }, n);
    const double ka = bench([&](int i) { sirrt_knn_add(k, qs[i & 63].data(), 1); }, n);
    const double kan = bench([&](int i) { sirrt_knn_add(k, qs[i & 63].data(), 1); sirrt_knn_nearest(k, qs[(i + 1) & 63].data(), 1, &idx, &d2); }, n);
    sirrt_knn_destroy(k);
    printf("{\"scene\": \"%s\", \"calls\": %d, \"us\": {\"get_safe_intervals\": %.2f, \"check_collision_frame\": %.2f, \"check_frames_batch_32\": %.2f, "
           "\"edge_table_32_rows\": %.2f, \"knn_nearest_5000\": %.2f, \"knn_radius_5000\": %.2f, \"knn_add\": %.2f, \"knn_add_then_nearest\": %.2f}}\n",
           argv[1], n, si, cf, cb, et, kn, kr, ka, kan);
  } catch (const std::exception& e) {
    fprintf(stderr, "sirrt_latency: %s\n", e.what());
    return 1;
  }
  return 0;
}
