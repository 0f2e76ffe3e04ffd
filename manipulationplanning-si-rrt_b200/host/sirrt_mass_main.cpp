// sirrt_mass_main.cpp -- the mass test (BASELINE.json config 5; reference driver mass_test/do_mass_test.py:47-60, one
// MSIRRT_Planner process per (scene, seed) in a ProcessPoolExecutor) on one B200: ONE process and CUDA context per GPU,
// `threads` host planner threads sharing the device through their own scenes / streams.  Tasks are independent, so a
// box shards them across GPUs by launching one of these per device (scripts/mass_test.py); no inter-GPU traffic.
//   sirrt_mass <tasks.txt> <device> <threads> [--max-time s] [--max-iters n] [--fixed] [--queue counter.bin]
// tasks.txt: one "<scene.pack> <seed>" per line.  Output: one JSON line per task (task index = line number), then a
// summary line.  A task is the unit of work: a thread claims the next unclaimed line from a counter and keeps the scene
// it built for its previous task when the pack is the same (the list keeps the seeds of a pack together).  With --queue
// the counter lives in a shared file (8 bytes, mmap'ed, atomic fetch-add), so the per-GPU processes of one box, all
// given the SAME task list, drain one queue: a GPU whose tasks finish early takes the next task instead of idling (the
// list is ordered heaviest first by the driver).
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <mutex>
#include <sstream>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include "B200Backend.hpp"
#include "SIRRTPlanner.hpp"

struct Task { std::string pack; int seed; int index; };

static uint64_t fnv(const void* p, size_t n, uint64_t h) {
  const unsigned char* b = (const unsigned char*)p;
  for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
  return h;
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: %s <tasks.txt> <device> <threads> [--max-time s] [--max-iters n] [--fixed]\n", argv[0]); return 2; }
  const int device = atoi(argv[2]), n_threads = std::max(1, atoi(argv[3]));
  double max_time = -1;
  uint64_t max_iters = 0;
  int parity = SIRRT_PARITY_REFERENCE_EXACT;
  const char* queue_path = nullptr;
  for (int i = 4; i < argc; i++) {
    if (!strcmp(argv[i], "--max-time") && i + 1 < argc) max_time = atof(argv[++i]);
    else if (!strcmp(argv[i], "--max-iters") && i + 1 < argc) max_iters = strtoull(argv[++i], nullptr, 10);
    else if (!strcmp(argv[i], "--fixed")) parity = SIRRT_PARITY_FIXED;
    else if (!strcmp(argv[i], "--queue") && i + 1 < argc) queue_path = argv[++i];
  }
  static unsigned long long local_counter = 0;
  unsigned long long* counter = &local_counter;
  if (queue_path) {
    const int fd = open(queue_path, O_RDWR);
    void* m = fd >= 0 ? mmap(nullptr, 8, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0) : MAP_FAILED;
    if (m == MAP_FAILED) { fprintf(stderr, "sirrt_mass: cannot map queue counter %s\n", queue_path); return 2; }
    counter = (unsigned long long*)m;
  }
  std::vector<Task> tasks;
  {
    std::ifstream f(argv[1]);
    if (!f) { fprintf(stderr, "sirrt_mass: cannot open %s\n", argv[1]); return 2; }
    std::string line;
    while (std::getline(f, line)) {
      std::istringstream is(line);
      Task t;
      if (is >> t.pack >> t.seed) { t.index = (int)tasks.size(); tasks.push_back(t); }
    }
  }
  using clk = std::chrono::steady_clock;
  std::mutex out_mu;
  std::atomic<int> n_solved{0}, n_failed{0}, n_done{0};
  std::atomic<uint64_t> solve_ns_total{0};
  const auto t0 = clk::now();
  auto worker = [&](int tid) {
    (void)tid;
    std::unique_ptr<MDP::CollisionManager> cm;
    std::string cm_pack;
    std::vector<sirrt_knn*> knn_pool;  // this thread's vertex stores, reused by its consecutive tasks
    for (;;) {
      const size_t ti = (size_t)__atomic_fetch_add(counter, 1ull, __ATOMIC_RELAXED);
      if (ti >= tasks.size()) break;
      const Task& t = tasks[ti];
      try {
        if (!cm || cm_pack != t.pack) {
          cm.reset();
          cm.reset(new MDP::CollisionManager(MDP::CollisionManager::from_pack_file(t.pack, device, parity)));
          cm_pack = t.pack;
        }
        MDP::B200Backend be(*cm, device, true, &knn_pool);
        MDP::MSIRRT::PlannerConnect<MDP::B200Backend> planner(be, t.seed);
        if (!planner.valid()) throw std::runtime_error(planner.error());
        if (max_iters) planner.set_max_iterations(max_iters);
        if (max_time > 0) planner.set_max_planning_time(max_time);
        const auto a = clk::now();
        const bool ok = planner.solve();
        const uint64_t ns = (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(clk::now() - a).count();
        uint64_t h = 1469598103934665603ull;
        double cost = -1;
        size_t n_way = 0;
        if (ok) {
          const auto path = planner.get_result_path(0.0);
          for (const auto& s : path) { h = fnv(s.q.data(), s.q.size() * sizeof(double), h); h = fnv(&s.time, sizeof(double), h); }
          cost = path.back().time; n_way = path.size();
          n_solved++;
        }
        solve_ns_total += ns;
        n_done++;
        const MDP::MSIRRT::PlannerStats& st = planner.stats();
        std::lock_guard<std::mutex> lk(out_mu);
        printf("{\"task\": %d, \"pack\": \"%s\", \"seed\": %d, \"solved\": %s, \"solve_ms\": %.4f, \"cost\": %.17g, \"waypoints\": %zu, "
               "\"path_hash\": \"%016llx\", \"iterations\": %llu, \"safe_interval_calls\": %llu, \"state_checks\": %llu, "
               "\"detected_collisions\": %llu, \"safe_intervals\": %llu, \"safe_interval_frames\": %llu}\n",
               t.index, t.pack.c_str(), t.seed, ok ? "true" : "false", ns / 1e6, cost, n_way, (unsigned long long)h,
               (unsigned long long)st.iterations, (unsigned long long)st.n_safe_interval_calls, (unsigned long long)st.n_state_checks,
               (unsigned long long)st.n_detected_collisions, (unsigned long long)st.n_safe_intervals, (unsigned long long)st.sum_safe_interval_frames);
      } catch (const std::exception& e) {
        n_failed++;
        cm.reset();
        std::lock_guard<std::mutex> lk(out_mu);
        fprintf(stderr, "sirrt_mass: %s seed %d: %s\n", t.pack.c_str(), t.seed, e.what());
      }
    }
    MDP::B200Backend::destroy_pool(knn_pool);
  };
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; t++) th.emplace_back(worker, t);
  for (auto& x : th) x.join();
  const double wall = std::chrono::duration<double>(clk::now() - t0).count();
  printf("{\"summary\": true, \"device\": %d, \"threads\": %d, \"tasks\": %zu, \"solved\": %d, \"failed\": %d, \"wall_s\": %.4f, \"sum_solve_s\": %.4f}\n",
         device, n_threads, (size_t)n_done.load(), n_solved.load(), n_failed.load(), wall, solve_ns_total.load() / 1e9);
  return n_failed.load() ? 1 : 0;
}
