// gjk_hist.cpp -- diagnostics: FP32 GJK iterations per deferred task (k_gjk_tasks), by verdict, on one scene pack.
// Needs the library built with -DSIRRT_GJK_STATS (scripts/micro/build_gjk_hist.sh); not part of the product.
//   gjk_hist <scene.pack> [n_configs]
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "CollisionManager.hpp"

extern "C" int sirrt_debug_gjk_hist(unsigned int* out, int reset);
extern "C" int sirrt_debug_phases(unsigned long long* out, int reset);
extern "C" int sirrt_debug_fk_mode(int mode);

static void dump(const char* what, int calls) {
  unsigned int h[4][64];
  sirrt_debug_gjk_hist(&h[0][0], 1);
  const char* names[4] = {"free", "hit", "ambiguous", "ambiguous_thin_tetra"};
  printf("{\"call\": \"%s\", \"calls\": %d", what, calls);
  for (int v = 0; v < 4; v++) {
    unsigned long long n = 0, it = 0; int mx = 0;
    for (int i = 0; i < 64; i++) { n += h[v][i]; it += (unsigned long long)h[v][i] * i; if (h[v][i]) mx = i; }
    printf(", \"%s\": {\"tasks\": %llu, \"mean_iter\": %.2f, \"max_iter\": %d, \"hist\": [", names[v], n, n ? (double)it / n : 0.0, mx);
    for (int i = 0; i <= mx; i++) printf("%s%u", i ? "," : "", h[v][i]);
    printf("]}");
  }
  unsigned long long ph[32];
  sirrt_debug_phases(ph, 1);
  const double cw = ph[0] ? (double)ph[0] : 1.0, tk = ph[8] ? (double)ph[8] : 1.0;
  printf(", \"collide_bits_cycles_per_warp\": {\"warps\": %llu, \"fk_and_link_table\": %.0f, \"static\": %.0f, \"dynamic\": %.0f, \"item\": %.0f, \"sincos\": %.0f, \"chain\": %.0f, \"table\": %.0f}",
         ph[0], ph[1] / cw, ph[2] / cw, ph[3] / cw, ph[4] / cw, ph[5] / cw, ph[6] / cw, ph[7] / cw);
  printf(", \"gjk_task_cycles\": {\"tasks\": %llu, \"prologue\": %.0f, \"gjk\": %.0f, \"iters\": %.2f, \"support_A\": %.0f, \"support_B\": %.0f, \"simplex\": %.0f}", ph[8], ph[9] / tk,
         ph[10] / tk, ph[12] / tk, ph[13] / tk, ph[14] / tk, ph[15] / tk);
  printf(", \"max_cycles_over_all_calls\": {\"static\": %llu, \"dynamic\": %llu, \"item\": %llu, \"gjk\": %llu, \"gjk_iters\": %llu}", ph[16 + 2], ph[16 + 3], ph[16 + 4], ph[16 + 10], ph[16 + 12]);
  printf("}\n");
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  const int n = argc > 2 ? atoi(argv[2]) : 64;
  MDP::CollisionManager cm = MDP::CollisionManager::from_pack_file(argv[1]);
  const int dof = cm.dof(), T = cm.frame_count();
  const auto lim = cm.get_planned_robot_limits();
  std::mt19937_64 rng(7);
  std::vector<std::vector<double>> qs(n, std::vector<double>(dof));
  for (auto& q : qs)
    for (int j = 0; j < dof; j++) q[j] = std::uniform_real_distribution<double>(lim[j].first, lim[j].second)(rng);
  sirrt_debug_gjk_hist(nullptr, 1);
  sirrt_debug_phases(nullptr, 1);
  for (int i = 0; i < n; i++) cm.get_safe_intervals(qs[i]);
  dump("get_safe_intervals", n);
  sirrt_debug_fk_mode(1);
  for (int i = 0; i < n; i++) cm.get_safe_intervals(qs[i]);
  dump("get_safe_intervals_fk_serial_chain", n);
  sirrt_debug_fk_mode(0);
  std::vector<double> states(32 * (size_t)dof);
  std::vector<int32_t> frames(32);
  std::vector<uint8_t> cells(32);
  for (int r = 0; r < n / 8 + 1; r++) {
    for (int s = 0; s < 32; s++) { for (int j = 0; j < dof; j++) states[(size_t)s * dof + j] = qs[(s + r) % n][j]; frames[s] = (s * 53 + r * 17) % T; }
    sirrt_check_frames_batch_f64(cm.handle(), states.data(), frames.data(), 32, cells.data());
  }
  dump("check_frames_batch_32", n / 8 + 1);
  return 0;
}
