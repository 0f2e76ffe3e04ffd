// sirrt_api.cu -- C ABI of libsirrt_b200.so (include/sirrt_b200.h): scene life cycle, the batched
// collision / safe-interval / edge entry points and the single-call drop-ins.  Everything that
// computes runs in the hand-written sm_100a kernels of sirrt_kernels.cuh / sirrt_sweep.cuh /
// sirrt_knn.cuh; there is no CPU fallback: without a CUDA device every compute call fails.
#include <sched.h>
#include <sys/prctl.h>
#include <time.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#include <algorithm>
#include <atomic>
#include <utility>
#include <mutex>
#include <new>

#include "../../include/sirrt_b200.h"
#include "sirrt_kernels.cuh"
#include "sirrt_sweep.cuh"
#include "sirrt_tables.hpp"

using namespace sirrt;

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, const char* a = "") {
  snprintf(g_err, sizeof g_err, fmt, a);
  return code;
}
#define CU(x)                                                                                          \
  do {                                                                                                 \
    cudaError_t e_ = (x);                                                                              \
    if (e_ != cudaSuccess) { cudaGetLastError(); return fail(SIRRT_E_CUDA, "CUDA: %s (" #x ")", cudaGetErrorString(e_)); } \
  } while (0)

// Every host wait of the library goes through here.  Default: cudaStreamSynchronize (spins: lowest latency when each
// planner thread has a core).  SIRRT_SYNC=yield in the environment: poll the stream and give the core away between
// polls -- for boxes that run more planner threads than cores (the mass test shares 32 cores between 8 GPUs' workers;
// a planner thread spends most of its time waiting for its GPU).  SIRRT_SYNC=block: cudaStreamSynchronize on a device
// whose scheduling flag is cudaDeviceScheduleBlockingSync (set at the first scene / store creation): the thread sleeps
// until the GPU's interrupt -- slower per call, but a waiting thread costs no core at all.
// SIRRT_SYNC=nap: like yield, but the thread first SLEEPS for half of what this kind of wait took recently (a running mean
// kept by the caller's object) -- a planner thread that waits 66 us for its kernels then burns no core for the first 33 us,
// and boxes with far more planner threads than cores (8 GPUs x 32 threads on 32 cores) keep their cores for the threads
// that have launches to issue.  The timer slack of the thread is lowered to 1 us once (default 50 us would oversleep).
static int sirrt_sync_mode() {
  static const int mode = [] { const char* e = getenv("SIRRT_SYNC"); return !e ? 0 : !strcmp(e, "yield") ? 1 : !strcmp(e, "block") ? 2 : !strcmp(e, "nap") ? 3 : 0; }();
  return mode;
}
static inline double sirrt_now_us() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static inline void sirrt_nap(float mean_us) {
  static thread_local bool slack_set = false;
  if (!slack_set) { prctl(PR_SET_TIMERSLACK, 1000ul, 0ul, 0ul, 0ul); slack_set = true; }
  const long ns = (long)(mean_us * 500.f);  // half of the recent mean
  if (ns < 8000) return;                    // short waits (a kNN query): not worth a timer
  timespec ts{0, ns > 2000000 ? 2000000 : ns};
  nanosleep(&ts, nullptr);
}
static void sirrt_apply_sync_mode(int device) {
  if (sirrt_sync_mode() != 2) return;
  static std::mutex mu;
  static bool done[64] = {false};
  std::lock_guard<std::mutex> lk(mu);
  if (device < 0 || device >= 64 || done[device]) return;
  done[device] = true;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(device);
  if (cudaSetDeviceFlags(cudaDeviceScheduleBlockingSync) != cudaSuccess) cudaGetLastError();  // best effort: plain synchronisation still works
  cudaSetDevice(cur);
}
static cudaError_t sirrt_sync(cudaStream_t st) {
  const int mode = sirrt_sync_mode();
  if (mode != 1 && mode != 3) return cudaStreamSynchronize(st);
  static thread_local float mean_us = 0.f;  // nap mode: recent duration of this thread's stream waits
  const double t0 = mode == 3 ? sirrt_now_us() : 0.0;
  if (mode == 3) sirrt_nap(mean_us);
  for (;;) {
    const cudaError_t e = cudaStreamQuery(st);
    if (e != cudaErrorNotReady) {
      if (mode == 3) mean_us += 0.25f * ((float)(sirrt_now_us() - t0) - mean_us);
      return e;
    }
    sched_yield();
  }
}

// Completion through a flag in the object's pinned host block (latency path): the last block of the call's last kernel
// writes `seq` behind everything the call produced (RecheckParams::done, k_knn_nearest1, k_knn_radius1), and the host polls
// its own memory (the top bit, SIRRT_FLAG_CONTINUE, is the caller's business: scene_wait) -- no driver call, so 16 planner threads waiting on one GPU do not queue on the driver's lock, and the
// wake-up does not pay cudaStreamSynchronize's hand-shake.  The stream is queried now and then only to surface a failed
// launch (the flag would never arrive).
static cudaError_t sirrt_wait_flag(cudaStream_t st, const volatile unsigned int* flag, unsigned int seq, float* mean_us = nullptr, bool or_continue = false) {
  const int mode = sirrt_sync_mode();
  const unsigned int mask = or_continue ? 0x7fffffffu : 0xffffffffu;  // or_continue: seq | SIRRT_FLAG_CONTINUE also ends the wait
  if (mode == 3 && mean_us) {  // sleep through the first half of the expected wait, then poll
    const double t0 = sirrt_now_us();
    sirrt_nap(*mean_us);
    for (unsigned spin = 1;; spin++) {
      if ((*flag & mask) == seq) { *mean_us += 0.25f * ((float)(sirrt_now_us() - t0) - *mean_us); return cudaSuccess; }
      if ((spin & 0xffu) == 0) {
        const cudaError_t e = cudaStreamQuery(st);
        if (e == cudaSuccess) return (*flag & mask) == seq ? cudaSuccess : cudaErrorUnknown;
        if (e != cudaErrorNotReady) return e;
      }
      sched_yield();
    }
  }
  if (mode == 2) {  // sleeping wait: the stream's completion implies the flag
    const cudaError_t e = cudaStreamSynchronize(st);
    return e != cudaSuccess ? e : ((*flag & mask) == seq ? cudaSuccess : cudaErrorUnknown);
  }
  for (unsigned spin = 1;; spin++) {
    if ((*flag & mask) == seq) return cudaSuccess;
    if ((spin & (mode ? 0xffu : 0xffffu)) == 0) {
      const cudaError_t e = cudaStreamQuery(st);
      if (e == cudaSuccess) return (*flag & mask) == seq ? cudaSuccess : cudaErrorUnknown;
      if (e != cudaErrorNotReady) return e;
    }
    if (mode) sched_yield();
    else __builtin_ia32_pause();  // mode 3 without a mean falls through here as yield
  }
}

extern "C" const char* sirrt_last_error(void) { return g_err; }
extern "C" int sirrt_version(void) { return 100; }
extern "C" int sirrt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

// Planner-sized calls in flight per device, over all scenes of the process.  Splitting a row's slots / a cell's pairs over
// extra warps (SIRRT_OPT_SLOT_SPLIT) shortens a call on an idle GPU (k_collide_bits 23 -> 17 us) but multiplies the warps --
// each repeating the FP64 FK -- and under load that is paid by the other callers: with 16 planner threads per GPU the
// mass test's p50 solve time was 32-42 ms with the split against 27-29 ms without (profiles/r2e_mass_knobs_1gpu.json).  So a
// call splits only while fewer than SIRRT_SPLIT_MAX_LOAD other calls are in flight on its device.
#define SIRRT_SPLIT_MAX_LOAD 2
static std::atomic<int> g_inflight[64];
struct InFlight {
  int dev, load;
  explicit InFlight(int d) : dev(d >= 0 && d < 64 ? d : 0) { load = g_inflight[dev].fetch_add(1, std::memory_order_relaxed); }
  ~InFlight() { g_inflight[dev].fetch_sub(1, std::memory_order_relaxed); }
  InFlight(const InFlight&) = delete;
  InFlight& operator=(const InFlight&) = delete;
};

// growable device scratch buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    // small buffers grow geometrically: the planner's stores grow by one point per call and a cudaMalloc per
    // call would dominate the latency path
    size_t want = bytes < ((size_t)64 << 20) ? std::max(2 * bytes, (size_t)4096) : bytes;
    if (cudaMalloc(&p, want) != cudaSuccess) { cudaGetLastError(); return -1; }
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() { return (T*)p; }
};

struct sirrt_scene {
  int device = 0, parity_mode = 0, n_sm = 148;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaStream_t copy_in = nullptr, copy_out = nullptr;  // pipelined host-buffer batches: H2D / D2H beside the compute stream
  std::vector<cudaEvent_t> pipe_ev;                     // per block: upload done, results ready
  int pipeline = 1;
  HScene hs;
  HostTables tb;
  DScene d{};
  const float* d_aabb_f = nullptr;
  const double* d_aabb_d = nullptr;
  const unsigned int* d_hull_cells = nullptr;  // kept while SIRRT_OPT_HULL_CELLS = 0 hides it from the kernels
  SweepTables sweep{};
  std::vector<void*> owned;  // device allocations of the scene tables
  double band = 0;
  long long chunk_configs = 1 << 18;
  unsigned int wl_cap = 1u << 22;
  int use_sweep = 1;
  long long sweep_min_configs = 8192;
  long long defer_max_configs = 4096;  // hull scenes: batches up to this size hand their GJK runs to k_gjk_tasks
  DevBuf q, bits, table, frames, cells, work, tasks, counts, misc, misc2, xf, row1, seeds;
  int seed_recheck = 1;  // latency path: the FP64 re-check of a deferred hull test starts from the FP32 run's simplex (SIRRT_OPT_SEED_RECHECK)
  // latency path: the kernels of a call chained by programmatic dependent launch (SIRRT_OPT_PDL).  OFF by default: it saves
  // ~1 us of a 50 us call on an idle GPU, but the pre-launched dependents hold SM slots while they wait, and 48 planner
  // threads sharing one GPU lost 30 % of their throughput (mass test, profiles/r2e_mass_knobs_1gpu.json)
  int pdl = 0;
  int inline_recheck = 9;  // flag-armed hull calls: k_recheck is not queued, the task kernel's last block re-checks up to inline_recheck - 1 items -- one per warp of the block, a second round would take longer than handing them back (SIRRT_OPT_INLINE_RECHECK; 0 = always queue k_recheck)
  unsigned pred_tasks[3] = {1024, 256, 4096};  // deferred pair tests the last single-row / cell / table call left (task_grid)
  int pred_kind = -1;
  bool k3_deferred = false;  // the pending call was launched without k_recheck; scene_wait launches it when the flag says CONTINUE
  RecheckParams k3{};
  unsigned k3_grid = 16;
  int obb_cull = 1;      // hull pair tests preceded by the oriented-box separating-axis cull (SIRRT_OPT_OBB_CULL)
  int host_buffers = 0;  // the caller promises host memory for every buffer of the single-call / cell / edge-table entry points (SIRRT_OPT_HOST_BUFFERS)
  int cur_load = 0;    // other planner-sized calls in flight on the device when the current call started (InFlight)
  int slot_split = 1;  // planner-sized batches: dynamic slots of a row split over several warps (SIRRT_OPT_SLOT_SPLIT)
  // Latency path (planner in the loop): one pinned host block that the GPU reads and writes in place
  // (cudaMallocHost memory is device-addressable under UVA).  Small host inputs are memcpy'd into it and read
  // by the kernels over the bus; small results are written into it by the kernels and memcpy'd out after the
  // single stream synchronisation of the call -- no cudaMemcpy on that path (measured on the B200 box: every
  // copy operation adds 3-4 us to the 9 us launch + synchronise floor; scripts/micro/lat_probe.cu).
  char* pinned = nullptr;
  unsigned int* lazy_cnt = nullptr;  // device: 4 launch counters + k_recheck ticket [4] + k_gjk_tasks ticket [6], zero between calls (finish_call clears them); [7] = sequence number of the call the task kernel completed
  bool lazy_timed = false;    // ... and its sweep launch was bracketed by ev[0] / ev[1]
  bool lazy_pending = false;  // counters of the last run_bits launch are in flight to PIN_CNT
  unsigned int fin_seq = 0;   // sequence number of latency-path calls (lazy_cnt[7]: which call the task kernel has already completed)
  unsigned int done_seq = 0;  // sequence number of the completion flag (PIN_FLAG) the pending call will write
  bool done_armed = false;    // the pending call ends with a flag write: wait with sirrt_wait_flag instead of a stream sync
  float wait_mean_us = 0.f;   // SIRRT_SYNC=nap: running mean of this scene's flag waits
  std::vector<cudaEvent_t> ev;  // start/stop pairs around the sweep launches of one call
  uint64_t stats[5] = {0, 0, 0, 0, 0};
};

enum : size_t { PIN_Q = 0, PIN_Q_CAP = 256 << 10, PIN_FR = PIN_Q + PIN_Q_CAP, PIN_FR_CAP = 64 << 10, PIN_OUT = PIN_FR + PIN_FR_CAP,
                PIN_OUT_CAP = 512 << 10, PIN_CNT = PIN_OUT + PIN_OUT_CAP, PIN_CNT_CAP = 4 << 10, PIN_FLAG = PIN_CNT + 256, PIN_BYTES = PIN_CNT + PIN_CNT_CAP };

template <typename T> static int upload(sirrt_scene* s, const std::vector<T>& v, const T** out) {
  void* p = nullptr;
  size_t bytes = std::max(v.size(), (size_t)1) * sizeof(T);
  CU(cudaMalloc(&p, bytes));
  s->owned.push_back(p);
  if (!v.empty()) CU(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *out = (const T*)p;
  return 0;
}

static bool is_device_ptr(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" int sirrt_scene_create_from_pack(const void* pack, size_t bytes, int device, int parity_mode, sirrt_scene** out) {
  if (!pack || !out || (parity_mode != SIRRT_PARITY_REFERENCE_EXACT && parity_mode != SIRRT_PARITY_FIXED))
    return fail(SIRRT_E_ARG, "sirrt_scene_create_from_pack: bad argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return fail(SIRRT_E_CUDA, "no CUDA device (libsirrt_b200 has no CPU fallback)"); }
  if (device < 0 || device >= ndev) return fail(SIRRT_E_ARG, "device index out of range");
  sirrt_scene* s = new (std::nothrow) sirrt_scene;
  if (!s) return fail(SIRRT_E_ARG, "out of memory");
  std::string err;
  if (!parse_pack(pack, bytes, s->hs, err)) { delete s; return fail(SIRRT_E_PACK, "%s", err.c_str()); }
  int rc = build_tables(s->hs, parity_mode, s->tb, err);
  if (rc) { delete s; return fail(rc, "%s", err.c_str()); }
  s->device = device; s->parity_mode = parity_mode;
  // environment overrides of the latency-path knobs (A/B measurements of whole programs, e.g. the mass test)
  if (const char* e = getenv("SIRRT_OPT_PDL")) s->pdl = atoi(e) != 0;
  if (const char* e = getenv("SIRRT_OPT_SLOT_SPLIT")) s->slot_split = atoi(e) != 0;
  if (const char* e = getenv("SIRRT_OPT_OBB_CULL")) s->obb_cull = atoi(e) != 0;
  if (const char* e = getenv("SIRRT_OPT_SEED_RECHECK")) s->seed_recheck = atoi(e) != 0;
  if (const char* e = getenv("SIRRT_OPT_INLINE_RECHECK")) s->inline_recheck = std::max(0, atoi(e));
  auto bail = [&](int code) { sirrt_scene_destroy(s); return code; };
  sirrt_apply_sync_mode(device);
  if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return bail(fail(SIRRT_E_CUDA, "cudaSetDevice failed")); }
  cudaDeviceGetAttribute(&s->n_sm, cudaDevAttrMultiProcessorCount, device);
  if (cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking) != cudaSuccess) { cudaGetLastError(); return bail(fail(SIRRT_E_CUDA, "cudaStreamCreate failed")); }
  s->stream = s->own_stream;
  HostTables& tb = s->tb;
  DScene& d = s->d;
  d.T = tb.T; d.Tp = tb.Tp; d.n_tiles = tb.n_tiles; d.n_slots = tb.n_slots;
  d.n_static = tb.n_static; d.n_dynamic = tb.n_dynamic; d.parity_mode = parity_mode; d.has_hull = tb.has_hull;
#define UP(field, vec) if ((rc = upload(s, vec, &field))) return bail(rc)
  UP(d.shapes, tb.shapes); UP(d.slots, tb.slots); UP(d.static_slots, tb.static_slots); UP(d.dynamic_slots, tb.dynamic_slots);
  UP(d.pose_f, tb.pose_f); UP(d.bcw_f, tb.bcw_f); UP(d.tight_f, tb.tight_f); UP(d.pose_d, tb.pose_d); UP(d.tile_bounds, tb.tile_bounds);
  UP(d.verts_f, tb.verts_f); UP(d.verts_d, tb.verts_d); UP(d.hull_cells, tb.hull_cells); UP(d.hull_cand, tb.hull_cand); UP(d.hull_rec, tb.hull_rec); UP(s->d_aabb_f, tb.aabb_f); UP(s->d_aabb_d, tb.aabb_d);
#undef UP
  // FP32 uncertainty band: the fast path decides a pair only when |signed distance| exceeds it;
  // everything inside is re-evaluated in FP64 (k_recheck).  Scales with the coordinate magnitudes
  // the FP32 arithmetic meets (arm reach around the base), never below 2e-5 m.
  s->band = 2e-5 * std::max(1.0, tb.extent / 1.5);
  if ((rc = sweep_build(s->hs, tb, s->d, &s->sweep, s->owned))) return bail(fail(rc, "sweep table build failed"));
  if (cudaMallocHost((void**)&s->pinned, PIN_BYTES) != cudaSuccess) { cudaGetLastError(); return bail(fail(SIRRT_E_CUDA, "cudaMallocHost failed")); }
  memset(s->pinned + PIN_CNT, 0, PIN_CNT_CAP);
  // cleared ON THE SCENE'S STREAM and waited for: a plain cudaMemset runs on the legacy default stream, is asynchronous
  // to the host for device memory, and is not ordered against this non-blocking stream -- under load it could land after
  // the first kernels had started counting
  if (cudaMalloc((void**)&s->lazy_cnt, 64) != cudaSuccess || cudaMemsetAsync(s->lazy_cnt, 0, 64, s->own_stream) != cudaSuccess ||
      cudaStreamSynchronize(s->own_stream) != cudaSuccess) { cudaGetLastError(); return bail(fail(SIRRT_E_CUDA, "cudaMalloc failed")); }
  for (int i = 0; i < 64; i++) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) { cudaGetLastError(); break; }
    s->ev.push_back(e);
  }
  // the table uploads above are plain cudaMemcpy calls from pageable memory: for small copies the call may return once
  // the data sits in the driver's staging buffer, with the DMA still queued on the legacy default stream -- which this
  // scene's non-blocking stream is not ordered against.  Wait for it before the first kernel can be launched.
  if (cudaStreamSynchronize(cudaStreamLegacy) != cudaSuccess) { cudaGetLastError(); return bail(fail(SIRRT_E_CUDA, "scene upload failed")); }
  *out = s;
  return SIRRT_OK;
}

extern "C" void sirrt_scene_destroy(sirrt_scene* s) {
  if (!s) return;
  cudaSetDevice(s->device);
  if (s->own_stream) { cudaStreamSynchronize(s->own_stream); }
  for (void* p : s->owned) cudaFree(p);
  for (DevBuf* b : {&s->q, &s->bits, &s->table, &s->frames, &s->cells, &s->work, &s->tasks, &s->counts, &s->misc, &s->misc2, &s->xf, &s->row1, &s->seeds}) b->release();
  if (s->pinned) cudaFreeHost(s->pinned);
  if (s->lazy_cnt) cudaFree(s->lazy_cnt);
  for (cudaEvent_t e : s->ev) cudaEventDestroy(e);
  for (cudaEvent_t e : s->pipe_ev) cudaEventDestroy(e);
  if (s->copy_in) cudaStreamDestroy(s->copy_in);
  if (s->copy_out) cudaStreamDestroy(s->copy_out);
  if (s->own_stream) cudaStreamDestroy(s->own_stream);
  cudaGetLastError();
  delete s;
}

extern "C" int sirrt_scene_info(const sirrt_scene* s, int* dof, int* n_links, int* n_frames, int* fps, float* lim) {
  if (!s) return fail(SIRRT_E_ARG, "null scene");
  if (dof) *dof = s->hs.dof;
  if (n_links) *n_links = s->hs.robot.n_links;
  if (n_frames) *n_frames = s->hs.T;
  if (fps) *fps = s->hs.fps;
  if (lim) memcpy(lim, s->hs.robot.limits.data(), sizeof(float) * s->hs.robot.limits.size());
  return SIRRT_OK;
}
extern "C" int sirrt_scene_task(const sirrt_scene* s, double* a, double* b, double* v) {
  if (!s) return fail(SIRRT_E_ARG, "null scene");
  const size_t n = sizeof(double) * s->hs.dof;
  if (a) memcpy(a, s->hs.start_cfg.data(), n);
  if (b) memcpy(b, s->hs.end_cfg.data(), n);
  if (v) memcpy(v, s->hs.max_vel.data(), n);
  return SIRRT_OK;
}
extern "C" int sirrt_scene_set_stream(sirrt_scene* s, void* st) {
  if (!s) return fail(SIRRT_E_ARG, "null scene");
  s->stream = st ? (cudaStream_t)st : s->own_stream;
  return SIRRT_OK;
}
extern "C" int sirrt_scene_set_band(sirrt_scene* s, double band) {
  if (!s || !(band >= 0) || band > 1e-2) return fail(SIRRT_E_ARG, "band out of range");
  s->band = band;
  return SIRRT_OK;
}
extern "C" int sirrt_scene_set_option(sirrt_scene* s, int key, long long value) {
  if (!s) return fail(SIRRT_E_ARG, "null scene");
  switch (key) {
    case SIRRT_OPT_USE_SWEEP: s->use_sweep = value != 0; break;
    case SIRRT_OPT_HULL_CELLS:
      if (value != 0 && !s->d_hull_cells) s->d_hull_cells = s->d.hull_cells;
      if (value == 0 && s->d.hull_cells) { s->d_hull_cells = s->d.hull_cells; s->d.hull_cells = nullptr; }
      else if (value != 0) s->d.hull_cells = s->d_hull_cells;
      break;
    case SIRRT_OPT_CHUNK_CONFIGS: if (value < 1) return fail(SIRRT_E_ARG, "bad chunk"); s->chunk_configs = value; break;
    case SIRRT_OPT_SWEEP_MIN_CONFIGS: if (value < 1) return fail(SIRRT_E_ARG, "bad threshold"); s->sweep_min_configs = value; break;
    case SIRRT_OPT_DEFER_MAX_CONFIGS: if (value < 0) return fail(SIRRT_E_ARG, "bad threshold"); s->defer_max_configs = value; break;
    case SIRRT_OPT_PIPELINE: s->pipeline = value != 0; break;
    case SIRRT_OPT_SLOT_SPLIT: s->slot_split = value != 0; break;
    case SIRRT_OPT_PDL: s->pdl = value != 0; break;
    case SIRRT_OPT_SEED_RECHECK: s->seed_recheck = value != 0; break;
    case SIRRT_OPT_HOST_BUFFERS: s->host_buffers = value != 0; break;
    case SIRRT_OPT_OBB_CULL: s->obb_cull = value != 0; break;
    case SIRRT_OPT_INLINE_RECHECK: if (value < 0 || value > 4096) return fail(SIRRT_E_ARG, "bad inline re-check limit"); s->inline_recheck = (int)value; break;
    case SIRRT_OPT_WORKLIST_CAP: if (value < 64 || value > (1ll << 28)) return fail(SIRRT_E_ARG, "bad worklist cap"); s->wl_cap = (unsigned)value; break;
    default: return fail(SIRRT_E_ARG, "unknown option");
  }
  return SIRRT_OK;
}
extern "C" int sirrt_scene_last_stats(const sirrt_scene* s, uint64_t st[5]) {
  if (!s || !st) return fail(SIRRT_E_ARG, "null argument");
  memcpy(st, s->stats, sizeof s->stats);
  return SIRRT_OK;
}

#ifdef SIRRT_GJK_STATS
extern "C" int sirrt_debug_gjk_hist(unsigned int* out, int reset) {  // diagnostics build only (scripts/micro/gjk_hist.cpp)
  if (cudaDeviceSynchronize() != cudaSuccess) return SIRRT_E_CUDA;
  if (out && cudaMemcpyFromSymbol(out, g_gjk_hist, sizeof(unsigned int) * 4 * 64) != cudaSuccess) return SIRRT_E_CUDA;
  if (reset) { static unsigned int z[4 * 64]; if (cudaMemcpyToSymbol(g_gjk_hist, z, sizeof z) != cudaSuccess) return SIRRT_E_CUDA; }
  return SIRRT_OK;
}
extern "C" int sirrt_debug_fk_mode(int mode) { return cudaMemcpyToSymbol(g_fk_mode, &mode, sizeof mode) == cudaSuccess ? SIRRT_OK : SIRRT_E_CUDA; }
extern "C" int sirrt_debug_phases(unsigned long long* out, int reset) {
  if (cudaDeviceSynchronize() != cudaSuccess) return SIRRT_E_CUDA;
  if (out && cudaMemcpyFromSymbol(out, g_phase, sizeof(unsigned long long) * 16) != cudaSuccess) return SIRRT_E_CUDA;
  if (out && cudaMemcpyFromSymbol(out + 16, g_phase_max, sizeof(unsigned long long) * 16) != cudaSuccess) return SIRRT_E_CUDA;
  if (reset) { static unsigned long long z[16]; if (cudaMemcpyToSymbol(g_phase, z, sizeof z) != cudaSuccess || cudaMemcpyToSymbol(g_phase_max, z, sizeof z) != cudaSuccess) return SIRRT_E_CUDA; }
  return SIRRT_OK;
}
#endif

#include "sirrt_api_batch.inl"
#include "sirrt_api_knn.inl"
#include "sirrt_api_noalloc.inl"
