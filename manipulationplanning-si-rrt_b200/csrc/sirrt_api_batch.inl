// sirrt_api_batch.inl -- batched collision bits / safe intervals / cell checks / edge validation.
// Included by sirrt_api.cu.

// Grid of the deferred-task kernel on the latency path (tasks beyond the grid are looped over, so its size costs time,
// never results).  Task counts vary from 0 to several hundred between consecutive planner calls, and a call on an idle
// GPU lasts as long as its slowest task, so an idle GPU gets at least one block per SM (a warp per task for 1184 tasks).
// With other calls in flight on the device every resident block is an SM slot another planner thread waits for (a fixed
// 592-block grid doubled the mass test's p50 solve time at 16 threads per GPU): the grid then follows a decaying maximum of
// the task counts this kind of call left on the scene.
static unsigned task_grid(const sirrt_scene* s, int kind, long long rows) {
  const unsigned full = (unsigned)s->n_sm * 4;
  if (rows > 64) return full;
  const unsigned want = (2u * s->pred_tasks[kind] + 64u + SIRRT_WARPS_PER_BLOCK - 1) / SIRRT_WARPS_PER_BLOCK;
  const unsigned floor_blocks = s->cur_load < SIRRT_SPLIT_MAX_LOAD ? (unsigned)s->n_sm : 8u;
  return std::max(floor_blocks, std::min(full, want));
}
static void note_tasks(sirrt_scene* s, int kind, unsigned tasks) {
  s->pred_tasks[kind] = std::max(tasks, s->pred_tasks[kind] - s->pred_tasks[kind] / 8);
}

static int grid_for(long long warps_wanted, int n_sm, int blocks_per_sm) {
  long long blocks = (warps_wanted + SIRRT_WARPS_PER_BLOCK - 1) / SIRRT_WARPS_PER_BLOCK;
  long long cap = (long long)n_sm * blocks_per_sm;
  return (int)std::max(1ll, std::min(blocks, cap));
}

// pointer classification: a caller that has promised host buffers for the whole scene (SIRRT_OPT_HOST_BUFFERS; the C++ shim
// passes std::vector storage only) is spared the driver query per pointer and call on the planner-in-the-loop entry points
static inline bool dev_ptr(const sirrt_scene* s, const void* p) { return s->host_buffers ? false : is_device_ptr(p); }

struct Range { long long n0, cnt; };
#define SIRRT_SEED_CAP (64u << 10)  // work-list entries that can carry a re-check seed (planner-sized calls leave a handful)

// Launch of a kernel that follows another kernel of the same latency-path call: with pdl the launch carries the
// programmatic-stream-serialization attribute (the kernel's blocks may become resident while its predecessor still runs;
// it calls pdl_wait() before it touches the predecessor's results)
template <typename... KArgs, typename... Args>
static cudaError_t launch_after(bool pdl, void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

// the call's last kernel (k_recheck on the latency path) will write the next sequence number to PIN_FLAG
static void arm_done(sirrt_scene* s, RecheckParams& R) {
  R.done = (unsigned int*)(s->pinned + PIN_FLAG);
  if (++s->done_seq >= 0x40000000u) s->done_seq = 1;  // the top bit of the flag is SIRRT_FLAG_CONTINUE
  R.done_seq = s->done_seq;
  s->done_armed = true;
  s->k3_deferred = false;
}
// wait for the call in flight: by the pinned flag when its last kernel carries one, else by the stream
static cudaError_t scene_wait(sirrt_scene* s) {
  if (!s->done_armed) return sirrt_sync(s->stream);
  s->done_armed = false;
  const volatile unsigned int* flag = (const volatile unsigned int*)(s->pinned + PIN_FLAG);
  cudaError_t e = sirrt_wait_flag(s->stream, flag, s->done_seq, &s->wait_mean_us, s->k3_deferred);
  if (e != cudaSuccess || !s->k3_deferred) return e;
  s->k3_deferred = false;
  if (*flag != (s->done_seq | SIRRT_FLAG_CONTINUE)) return cudaSuccess;
  // the task kernel's last block found more FP64 items than it re-checks itself: k_recheck after all (it completes the call)
  k_recheck<<<s->k3_grid, SIRRT_WARPS_PER_BLOCK * 32, 0, s->stream>>>(s->tb.robot, s->k3);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  s->stats[0]++;
  return sirrt_wait_flag(s->stream, flag, s->done_seq, &s->wait_mean_us);
}

// Collision bit rows of configurations [0, N) (global indexing into q / d_bits).  Work is cut into
// launches of at most chunk_configs configurations so that the FP64 re-check work list stays
// bounded; a launch whose list overflowed is repeated in halves.
// lazy = true (single-launch batches of the latency path): nothing is synchronised here; the launch counters
// travel to PIN_CNT behind the kernels and the caller settles them with finish_lazy() after its one
// synchronisation (a work-list overflow then repeats the call on the checked path).
// fuse_table (latency path, at most 64 rows): k_recheck's last block also reduces the rows to intervals; *fused tells the
// caller that no k_bits_to_intervals launch is needed.
// row_tiles / win_tiles: every row only needs the win_tiles tiles starting at row_tiles[i].x (windowed edge tables).
// bits_clean: the caller guarantees that the rows read zero (slot-split launches OR their words in; otherwise the rows are
// cleared here first).  fuse_clear: the fused interval reduction leaves the rows zeroed again (internal row buffers).
static int run_bits(sirrt_scene* s, QView q, long long N, int semantics, uint32_t* d_bits, bool lazy = false, int32_t* fuse_table = nullptr,
                    int fuse_kmax = 0, bool* fused = nullptr, const int2* row_tiles = nullptr, int win_tiles = 0, bool bits_clean = false,
                    bool fuse_clear = false, uint32_t* rows_out = nullptr) {
  if (fused) *fused = false;
  s->done_armed = false;
  const int W = s->tb.n_tiles;
  cudaStream_t st = s->stream;
  if (s->work.ensure((size_t)s->wl_cap * sizeof(WorkItem))) return fail(SIRRT_E_CUDA, "cudaMalloc(work list) failed");
  std::vector<Range> todo;
  for (long long n0 = 0; n0 < N; n0 += s->chunk_configs) todo.push_back({n0, std::min(s->chunk_configs, N - n0)});
  uint64_t launches = 0, rechecked = 0, near = 0, exact = 0, sweep_ns = 0;
  const bool sweep = !row_tiles && s->use_sweep && s->sweep.supported && N >= s->sweep_min_configs;
  const bool may_defer = !sweep && s->tb.has_hull && s->defer_max_configs > 0;
  if (may_defer && s->tasks.ensure((size_t)s->wl_cap * sizeof(WorkItem))) return fail(SIRRT_E_CUDA, "cudaMalloc(task list) failed");
  while (!todo.empty()) {
    const size_t nr = todo.size();
    const bool self_clearing = lazy && nr == 1;  // counters cleared and published by k_recheck's last block: no memset, no copy
    if (!self_clearing) {
      if (s->counts.ensure(nr * 4 * sizeof(unsigned))) return fail(SIRRT_E_CUDA, "cudaMalloc(counters) failed");
      CU(cudaMemsetAsync(s->counts.p, 0, nr * 4 * sizeof(unsigned), st));
    }
    for (size_t i = 0; i < nr; i++) {
      QView qv = q;
      qv.n0 = todo[i].n0; qv.cnt = todo[i].cnt;
      WorkList wl{s->work.as<WorkItem>(), self_clearing ? s->lazy_cnt : s->counts.as<unsigned>() + 4 * i, s->wl_cap};
      if (self_clearing && may_defer && s->seed_recheck) {  // latency path on hull scenes: the FP64 re-check starts from the FP32 simplex
        if (s->seeds.ensure((size_t)SIRRT_SEED_CAP * sizeof(int4))) return fail(SIRRT_E_CUDA, "cudaMalloc(re-check seeds) failed");
        wl.seeds = s->seeds.as<int4>(); wl.seed_cap = SIRRT_SEED_CAP;
      }
      // (the single-launch latency path is not timed -- two event records per planner call -- unless it runs the sweep kernel,
      // which planner-sized calls never do by default: sweep_kernel_ms must then still say that k_sweep ran)
      const bool timed = (sweep || !(lazy && nr == 1)) && i < s->ev.size() / 2;
      RecheckParams R{};
      R.sc = s->d; R.q = qv; R.items = wl.items; R.count = wl.count; R.cap = wl.cap;
      R.aabb_d = s->d_aabb_d; R.bits = d_bits; R.n_words = W; R.cells = nullptr; R.counters = wl.count;
      R.seeds = wl.seeds; R.seed_cap = wl.seed_cap;
      if (self_clearing) {
        R.ticket = s->lazy_cnt + 4; R.publish = (unsigned int*)(s->pinned + PIN_CNT);
        R.fin = s->lazy_cnt + 7; R.fin_seq = ++s->fin_seq;
      }
      if (self_clearing && fuse_table && N <= 64) {
        R.iv = IvParams{d_bits, N, W, s->tb.T, fuse_kmax, fuse_table};
        R.iv.clear_rows = fuse_clear ? 1 : 0;
        R.iv.no_fill = fuse_clear ? 1 : 0;  // sirrt_safe_intervals reads the first `count` pairs only
        if (fused) *fused = true;
        arm_done(s, R);  // the table is the pinned host block and nothing follows this kernel: the caller waits on the flag (scene_wait)
      } else if (self_clearing && rows_out) {  // edge tables: the call's last block copies the raw rows to the pinned block, then the flag
        R.iv = IvParams{d_bits + (size_t)qv.n0 * W, qv.cnt, W, s->tb.T, 0, nullptr};
        R.iv.rows_out = rows_out;
        if (fused) *fused = true;
        arm_done(s, R);
      }
      if (timed) cudaEventRecord(s->ev[2 * i], st);
      bool skip_k3 = false;
      if (sweep) {
        launches += sweep_launch(s->sweep, s->tb.robot, s->d, qv, wl, d_bits, W, (float)s->band, s->n_sm, st);
      } else {
        BitsParams P{};
        P.sc = s->d; P.q = qv; P.wl = wl; P.aabb_f = s->d_aabb_f; P.bits = d_bits; P.n_words = W;
        P.semantics = semantics; P.band = (float)s->band; P.obb_cull = s->obb_cull;
        P.row_tiles = row_tiles ? row_tiles + qv.n0 : nullptr;
        const bool defer = may_defer && qv.cnt <= s->defer_max_configs && nr == 1;  // one task list: one range at a time
        if (defer) {
          P.tasks = WorkList{s->tasks.as<WorkItem>(), wl.count + 3, s->wl_cap};
          if (s->xf.ensure((size_t)qv.cnt * s->tb.robot.n_links * 12 * sizeof(double))) return fail(SIRRT_E_CUDA, "cudaMalloc(link transforms) failed");
          P.xf = s->xf.as<double>();
        }
        // small batches: split each row over several warps (tile chunks) to fill the machine
        long long target = (long long)s->n_sm * 32;
        const int Wr = row_tiles ? std::max(1, std::min(win_tiles, W)) : W;  // tiles a row needs
        int n_chunks = (int)std::max(1ll, std::min((long long)Wr, target / std::max(1ll, qv.cnt)));
        P.tiles_per_chunk = (Wr + n_chunks - 1) / n_chunks;
        P.n_chunks = (Wr + P.tiles_per_chunk - 1) / P.tiles_per_chunk;
        // planner-sized batches leave most of the machine idle with one warp per (row, tile): split the dynamic slots of each
        // over several warps as well (the static slots of a row become items of 32 pairs), words OR-ed into zeroed rows
        P.n_schunks = 1; P.slots_per_chunk = s->tb.n_dynamic;
        const long long row_warps = qv.cnt * P.n_chunks;
        if (s->slot_split && s->cur_load < SIRRT_SPLIT_MAX_LOAD && s->tb.n_dynamic >= 2 && row_warps * 2 <= target) {
          // down to ONE slot per warp: the launch lasts as long as its slowest warp, and a warp's (slot, link) pairs are serial
          const int want = (int)std::min<long long>(s->tb.n_dynamic, target / row_warps);
          if (want > 1) {
            P.slots_per_chunk = (s->tb.n_dynamic + want - 1) / want;
            P.n_schunks = (s->tb.n_dynamic + P.slots_per_chunk - 1) / P.slots_per_chunk;
          }
        }
        if (P.n_schunks > 1 && !bits_clean) CU(cudaMemsetAsync(d_bits + (size_t)qv.n0 * W, 0, (size_t)qv.cnt * W * sizeof(uint32_t), st));
        const long long static_items = (P.n_schunks > 1 && s->tb.n_static > 0) ? (long long)(s->tb.robot.n_links * s->tb.n_static + 31) / 32 : 0;
        const int grid = grid_for(row_warps * P.n_schunks + static_items * qv.cnt, s->n_sm, 8);
        if (s->tb.has_hull) k_collide_bits<true><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
        else k_collide_bits<false><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
        launches++;
        if (defer) {
          GjkTaskParams G{};
          G.sc = s->d; G.q = qv; G.items = P.tasks.items; G.count = P.tasks.count; G.cap = P.tasks.cap; G.wl = wl; G.xf = P.xf;
          G.bits = d_bits; G.n_words = W; G.cells = nullptr; G.band = (float)s->band;
          if (self_clearing) {  // the task kernel's last block completes the call when nothing is left for the FP64 re-check
            G.ticket = s->lazy_cnt + 6; G.publish = R.publish; G.done = R.done; G.done_seq = R.done_seq; G.fin = R.fin; G.fin_seq = R.fin_seq; G.iv = R.iv;
            // flag-armed calls: k_recheck is not queued at all (the last block re-checks a handful of items itself; with more
            // it raises SIRRT_FLAG_CONTINUE and scene_wait launches k_recheck)
            if (R.done && s->inline_recheck) { G.inline_recheck = s->inline_recheck; G.aabb_d = s->d_aabb_d; skip_k3 = true; }
          }
          // a single row rarely defers more than a few hundred pair tests: a grid of one block per SM (1184 warps) leaves fewer
          // idle blocks to launch, fence and count out
          s->pred_kind = N == 1 ? 0 : 2;
          CU(launch_after(self_clearing && s->pdl, k_gjk_tasks, self_clearing ? task_grid(s, s->pred_kind, N) : (unsigned)s->n_sm * 4, SIRRT_WARPS_PER_BLOCK * 32, st,
                          s->tb.robot, G));
          launches++;
        }
      }
      if (timed) cudaEventRecord(s->ev[2 * i + 1], st);
      if (skip_k3) { s->k3 = R; s->k3_grid = (unsigned)s->n_sm; s->k3_deferred = true; }
      else
      // a handful of rows leaves a handful of ambiguous pair tests: a small grid shortens the last-block hand-over
      CU(launch_after(self_clearing && s->pdl && !sweep, k_recheck, (self_clearing && N <= 64) ? 16u : (unsigned)s->n_sm * 2, SIRRT_WARPS_PER_BLOCK * 32, st, s->tb.robot, R));
      if (!skip_k3) launches++;
      CU(cudaGetLastError());
    }
    if (self_clearing) {
      s->lazy_timed = sweep && !s->ev.empty();
      s->lazy_pending = true;
      s->stats[0] += launches;
      return SIRRT_OK;
    }
    std::vector<unsigned> host(nr * 4);
    CU(cudaMemcpyAsync(host.data(), s->counts.p, nr * 4 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CU(sirrt_sync(st));
    for (size_t i = 0; i < nr && i < s->ev.size() / 2; i++) {
      float ms = 0;
      if (cudaEventElapsedTime(&ms, s->ev[2 * i], s->ev[2 * i + 1]) == cudaSuccess) sweep_ns += (uint64_t)(ms * 1e6);
      else cudaGetLastError();
    }
    std::vector<Range> redo;
    for (size_t i = 0; i < nr; i++) {
      if (host[4 * i] > s->wl_cap || host[4 * i + 3] > s->wl_cap) {
        if (todo[i].cnt <= 1) return fail(SIRRT_E_OVERFLOW, "FP64 re-check work list overflow on a single configuration; raise SIRRT_OPT_WORKLIST_CAP");
        const long long h = todo[i].cnt / 2;
        redo.push_back({todo[i].n0, h});
        redo.push_back({todo[i].n0 + h, todo[i].cnt - h});
        // rows of the overflowed range are recomputed from scratch (bits are overwritten by the producer)
      } else {
        rechecked += host[4 * i]; exact += host[4 * i + 1] + host[4 * i + 3]; near += host[4 * i + 2];
      }
    }
    todo.swap(redo);
  }
  s->stats[0] += launches; s->stats[1] += rechecked; s->stats[2] += near; s->stats[3] += exact; s->stats[4] += sweep_ns;
  return SIRRT_OK;
}

// after the caller's synchronisation: 0 = settled, 1 = the work list overflowed (repeat on the checked path)
static int finish_lazy(sirrt_scene* s) {
  if (!s->lazy_pending) return 0;
  s->lazy_pending = false;
  if (s->lazy_timed) {
    s->lazy_timed = false;
    float ms = 0;
    if (cudaEventElapsedTime(&ms, s->ev[0], s->ev[1]) == cudaSuccess) s->stats[4] += (uint64_t)(ms * 1e6);
    else cudaGetLastError();
  }
  const unsigned* host = (const unsigned*)(s->pinned + PIN_CNT);
  if (host[0] > s->wl_cap || host[3] > s->wl_cap) return 1;
  s->stats[1] += host[0]; s->stats[3] += host[1] + host[3]; s->stats[2] += host[2];
  if (s->pred_kind >= 0) { note_tasks(s, s->pred_kind, host[3]); s->pred_kind = -1; }  // sizes the next task grid of this kind of call
  return 0;
}

static int run_intervals(sirrt_scene* s, const uint32_t* d_bits, long long N, int32_t* d_table, int Kmax) {
  IvParams P{d_bits, N, s->tb.n_tiles, s->tb.T, Kmax, d_table};
  k_bits_to_intervals<<<(unsigned)((N + 7) / 8), 256, 0, s->stream>>>(P);
  s->stats[0]++;
  CU(cudaGetLastError());
  return SIRRT_OK;
}

// bring a caller buffer onto the device (no-op for device pointers)
static int stage_in(sirrt_scene* s, DevBuf& buf, const void* src, size_t bytes, const void** dev, size_t pin_off = 0, size_t pin_cap = 0) {
  if (dev_ptr(s, src)) { *dev = src; return 0; }
  if (bytes <= pin_cap) {  // latency path: the kernels read the pinned block in place
    memcpy(s->pinned + pin_off, src, bytes);
    *dev = s->pinned + pin_off;
    return 0;
  }
  if (buf.ensure(bytes)) return fail(SIRRT_E_CUDA, "cudaMalloc(staging) failed");
  CU(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, s->stream));
  *dev = buf.p;
  return 0;
}

static int bits_and_table(sirrt_scene* s, QView q, long long N, int semantics, int32_t* table, int Kmax, uint32_t* bits,
                          uint64_t* near_contact_count, const int2* row_tiles = nullptr, int win_tiles = 0) {
  const int W = s->tb.n_tiles;
  InFlight in_flight(s->device);
  s->cur_load = in_flight.load;
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  uint32_t* d_bits = nullptr;
  const bool bits_dev = bits && dev_ptr(s, bits);
  if (bits_dev) d_bits = bits;
  else {
    if (s->bits.ensure((size_t)N * W * sizeof(uint32_t))) return fail(SIRRT_E_CUDA, "cudaMalloc(bits) failed");
    d_bits = s->bits.as<uint32_t>();
  }
  const size_t tbytes = table ? (size_t)N * (1 + 2 * Kmax) * sizeof(int32_t) : 0, bbytes = (size_t)N * W * sizeof(uint32_t);
  const bool tdev = table && dev_ptr(s, table);
  // latency path: one launch set, results through the pinned block, one synchronisation
  const bool small = N <= s->chunk_configs && (!table || tdev || tbytes <= PIN_OUT_CAP / 2) && (!bits || bits_dev || bbytes <= PIN_OUT_CAP / 2);
  for (int attempt = 0; attempt < 2; attempt++) {
    const bool lazy = small && attempt == 0;
    if (row_tiles) CU(cudaMemsetAsync(d_bits, 0, bbytes, s->stream));  // defined content outside the rows' windows (callers only look inside)
    // planner-sized edge tables (host rows, no interval table): no copy operation and no stream synchronisation -- the
    // call's last block leaves the rows in the pinned block and raises the completion flag
    const bool rows_by_flag = lazy && !table && bits && !bits_dev && bbytes <= (16u << 10);
    bool armed = false;
    int rc = run_bits(s, q, N, semantics, d_bits, lazy, nullptr, 0, &armed, row_tiles, win_tiles, row_tiles != nullptr, false,
                      rows_by_flag ? (uint32_t*)(s->pinned + PIN_OUT + PIN_OUT_CAP / 2) : nullptr);
    if (rc) return rc;
    if (armed) {
      CU(scene_wait(s));
      if (finish_lazy(s)) continue;
      memcpy(bits, s->pinned + PIN_OUT + PIN_OUT_CAP / 2, bbytes);
      break;
    }
    int32_t* t_host = nullptr;
    uint32_t* b_host = nullptr;
    if (table) {
      int32_t* d_table = table;
      if (!tdev) {
        if (small) d_table = t_host = (int32_t*)(s->pinned + PIN_OUT);  // k_bits_to_intervals writes the pinned block in place
        else {
          if (s->table.ensure(tbytes)) return fail(SIRRT_E_CUDA, "cudaMalloc(table) failed");
          d_table = s->table.as<int32_t>();
        }
      }
      if ((rc = run_intervals(s, d_bits, N, d_table, Kmax))) return rc;
      if (!tdev && !small) CU(cudaMemcpyAsync(table, d_table, tbytes, cudaMemcpyDeviceToHost, s->stream));
    }
    if (bits && !bits_dev) {
      if (small) { b_host = (uint32_t*)(s->pinned + PIN_OUT + PIN_OUT_CAP / 2); CU(cudaMemcpyAsync(b_host, d_bits, bbytes, cudaMemcpyDeviceToHost, s->stream)); }
      else CU(cudaMemcpyAsync(bits, d_bits, bbytes, cudaMemcpyDeviceToHost, s->stream));
    }
    CU(sirrt_sync(s->stream));
    if (lazy && finish_lazy(s)) continue;  // FP64 work list overflowed: repeat with the checked, splitting loop
    if (t_host) memcpy(table, t_host, tbytes);
    if (b_host) memcpy(bits, b_host, bbytes);
    break;
  }
  if (near_contact_count) *near_contact_count = s->stats[2];
  return SIRRT_OK;
}

// Pipelined form of the batched path for HOST buffers (the end-to-end path of a caller that keeps its configurations and
// tables in host memory): the batch is cut into blocks of at most SIRRT_PIPE_CONFIGS configurations and three streams run
//   copy-in: H2D of block k+1   ||   compute: sweep + FP64 re-check + intervals of block k   ||   copy-out: D2H of block k-1
// so that only the first block's upload and the last block's download are exposed.  Results are those of the plain path
// (same kernels, same per-block work lists); a block whose FP64 work list overflowed sends the call to the plain path.
// Batches of one to two blocks (a rank's share of a 2^20 batch split over 8 GPUs) are cut in halves of at least
// SIRRT_PIPE_MIN_CONFIGS (one full wave of k_sweep warps on 148 SMs) so that their copies overlap as well.
#define SIRRT_PIPE_CONFIGS (128 << 10)
#define SIRRT_PIPE_MIN_CONFIGS (64 << 10)
static long long pipe_block_configs(const sirrt_scene* s, long long N) {
  long long per = std::min<long long>(SIRRT_PIPE_CONFIGS, s->chunk_configs);
  if (N < 2 * per) per = std::min(per, std::max<long long>(SIRRT_PIPE_MIN_CONFIGS, ((N + 1) / 2 + 1023) / 1024 * 1024));
  return per;
}
static int bits_and_table_pipelined(sirrt_scene* s, const float* hq, long long N, int32_t* table, int Kmax, uint32_t* bits, uint64_t* near_contact_count,
                                    bool* fell_back, int16_t* table16 = nullptr) {
  *fell_back = false;
  const int W = s->tb.n_tiles, dof = s->hs.dof, cols = 1 + 2 * Kmax;
  const size_t esz = table16 ? sizeof(int16_t) : sizeof(int32_t);  // compact rows: int16 entries
  if (table16) table = (int32_t*)table16;
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  if (!s->copy_in) {
    CU(cudaStreamCreateWithFlags(&s->copy_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&s->copy_out, cudaStreamNonBlocking));
  }
  const long long per = pipe_block_configs(s, N);
  const size_t nb = (size_t)((N + per - 1) / per);
  while (s->pipe_ev.size() < 2 * nb) {
    cudaEvent_t e;
    CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    s->pipe_ev.push_back(e);
  }
  if (s->q.ensure((size_t)N * dof * sizeof(float)) || s->bits.ensure((size_t)N * W * sizeof(uint32_t)) || (table && s->table.ensure((size_t)N * cols * esz)) ||
      s->work.ensure((size_t)s->wl_cap * sizeof(WorkItem)) || s->counts.ensure(nb * 4 * sizeof(unsigned)))
    return fail(SIRRT_E_CUDA, "cudaMalloc(pipeline buffers) failed");
  cudaStream_t st = s->stream;
  float* dq = s->q.as<float>();
  uint32_t* d_bits = s->bits.as<uint32_t>();
  int32_t* d_table = table ? s->table.as<int32_t>() : nullptr;
  const bool sweep = s->use_sweep && s->sweep.supported && N >= s->sweep_min_configs;
  CU(cudaMemsetAsync(s->counts.p, 0, nb * 4 * sizeof(unsigned), st));
  uint64_t launches = 0;
  for (size_t i = 0; i < nb; i++) {
    const long long n0 = (long long)i * per, cnt = std::min(per, N - n0);
    // joint rows of the SoA batch are N floats apart on both sides: one strided copy per block
    CU(cudaMemcpy2DAsync(dq + n0, (size_t)N * sizeof(float), hq + n0, (size_t)N * sizeof(float), (size_t)cnt * sizeof(float), dof, cudaMemcpyHostToDevice, s->copy_in));
    CU(cudaEventRecord(s->pipe_ev[2 * i], s->copy_in));
    CU(cudaStreamWaitEvent(st, s->pipe_ev[2 * i], 0));
    QView qv{dq, nullptr, N, n0, cnt, dof};
    WorkList wl{s->work.as<WorkItem>(), s->counts.as<unsigned>() + 4 * i, s->wl_cap};
    const bool timed = i < s->ev.size() / 2;
    if (timed) cudaEventRecord(s->ev[2 * i], st);
    if (sweep) launches += sweep_launch(s->sweep, s->tb.robot, s->d, qv, wl, d_bits, W, (float)s->band, s->n_sm, st);
    else {
      BitsParams P{};
      P.sc = s->d; P.q = qv; P.wl = wl; P.aabb_f = s->d_aabb_f; P.bits = d_bits; P.n_words = W; P.semantics = SIRRT_SEM_INTERVAL; P.band = (float)s->band; P.obb_cull = s->obb_cull;
      const long long target = (long long)s->n_sm * 32;
      int n_chunks = (int)std::max(1ll, std::min((long long)W, target / std::max(1ll, cnt)));
      P.tiles_per_chunk = (W + n_chunks - 1) / n_chunks;
      P.n_chunks = (W + P.tiles_per_chunk - 1) / P.tiles_per_chunk;
      const int grid = grid_for(cnt * P.n_chunks, s->n_sm, 8);
      if (s->tb.has_hull) k_collide_bits<true><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
      else k_collide_bits<false><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
      launches++;
    }
    if (timed) cudaEventRecord(s->ev[2 * i + 1], st);
    RecheckParams R{};
    R.sc = s->d; R.q = qv; R.items = wl.items; R.count = wl.count; R.cap = wl.cap;
    R.aabb_d = s->d_aabb_d; R.bits = d_bits; R.n_words = W; R.cells = nullptr; R.counters = wl.count;
    k_recheck<<<s->n_sm * 2, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, R);
    launches++;
    if (table) {
      IvParams I{d_bits + (size_t)n0 * W, cnt, W, s->tb.T, Kmax, table16 ? nullptr : d_table + (size_t)n0 * cols};
      if (table16) I.table16 = (int16_t*)d_table + (size_t)n0 * cols;
      k_bits_to_intervals<<<(unsigned)((cnt + 7) / 8), 256, 0, st>>>(I);
      launches++;
    }
    CU(cudaGetLastError());
    CU(cudaEventRecord(s->pipe_ev[2 * i + 1], st));
    CU(cudaStreamWaitEvent(s->copy_out, s->pipe_ev[2 * i + 1], 0));
    if (table) CU(cudaMemcpyAsync((char*)table + (size_t)n0 * cols * esz, (char*)d_table + (size_t)n0 * cols * esz, (size_t)cnt * cols * esz, cudaMemcpyDeviceToHost, s->copy_out));
    if (bits) CU(cudaMemcpyAsync(bits + (size_t)n0 * W, d_bits + (size_t)n0 * W, (size_t)cnt * W * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->copy_out));
  }
  std::vector<unsigned> host(nb * 4);
  CU(cudaMemcpyAsync(host.data(), s->counts.p, nb * 4 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CU(sirrt_sync(st));
  CU(sirrt_sync(s->copy_out));
  uint64_t rechecked = 0, near = 0, exact = 0, sweep_ns = 0;
  for (size_t i = 0; i < nb; i++) {
    if (host[4 * i] > s->wl_cap || host[4 * i + 3] > s->wl_cap) { *fell_back = true; return SIRRT_OK; }  // repeat on the plain (splitting) path
    rechecked += host[4 * i]; exact += host[4 * i + 1] + host[4 * i + 3]; near += host[4 * i + 2];
    float ms = 0;
    if (i < s->ev.size() / 2 && cudaEventElapsedTime(&ms, s->ev[2 * i], s->ev[2 * i + 1]) == cudaSuccess) sweep_ns += (uint64_t)(ms * 1e6);
    else cudaGetLastError();
  }
  s->stats[0] = launches; s->stats[1] = rechecked; s->stats[2] = near; s->stats[3] = exact; s->stats[4] = sweep_ns;
  if (near_contact_count) *near_contact_count = near;
  return SIRRT_OK;
}

extern "C" int sirrt_safe_intervals_batch(sirrt_scene* s, const float* q_soa, int64_t N, int32_t* table, int Kmax,
                                          uint32_t* bits, uint64_t* near_contact_count) {
  if (!s || !q_soa || N < 0 || (table && Kmax < 1) || (!table && !bits)) return fail(SIRRT_E_ARG, "sirrt_safe_intervals_batch: bad argument");
  if (N == 0) { if (near_contact_count) *near_contact_count = 0; return SIRRT_OK; }
  if (N > 0x7fffffffll) return fail(SIRRT_E_ARG, "N exceeds 2^31-1 configurations per call");
  CU(cudaSetDevice(s->device));
  int rc;
  if (s->pipeline && N >= 2 * SIRRT_PIPE_MIN_CONFIGS && !is_device_ptr(q_soa) && !(table && is_device_ptr(table)) && !(bits && is_device_ptr(bits))) {
    bool fell_back = false;  // host buffers, more than one block: copies overlap the kernels
    rc = bits_and_table_pipelined(s, q_soa, N, table, Kmax, bits, near_contact_count, &fell_back);
    if (rc || !fell_back) return rc;
  }
  const void* dq = nullptr;
  rc = stage_in(s, s->q, q_soa, (size_t)N * s->hs.dof * sizeof(float), &dq, PIN_Q, 64 << 10);
  if (rc) return rc;
  QView q{(const float*)dq, nullptr, N, 0, N, s->hs.dof};
  return bits_and_table(s, q, N, SIRRT_SEM_INTERVAL, table, Kmax, bits, near_contact_count);
}

// Compact form of the interval table: int16 entries (row = {count, lo0, hi0, ...}, unused = -1), half the bytes of the
// default int32 table on the way back to the host.  Host buffers; scenes with at most 32767 frames.
extern "C" int sirrt_safe_intervals_batch_i16(sirrt_scene* s, const float* q_soa, int64_t N, int16_t* table16, int Kmax, uint64_t* near_contact_count) {
  if (!s || !q_soa || !table16 || N < 0 || Kmax < 1) return fail(SIRRT_E_ARG, "sirrt_safe_intervals_batch_i16: bad argument");
  if (s->tb.T > 32767) return fail(SIRRT_E_LIMIT, "sirrt_safe_intervals_batch_i16: more than 32767 frames; use the int32 table");
  if (is_device_ptr(q_soa) || is_device_ptr(table16)) return fail(SIRRT_E_ARG, "sirrt_safe_intervals_batch_i16 takes host buffers");
  if (N == 0) { if (near_contact_count) *near_contact_count = 0; return SIRRT_OK; }
  if (N > 0x7fffffffll) return fail(SIRRT_E_ARG, "N exceeds 2^31-1 configurations per call");
  bool fell_back = false;
  int rc = bits_and_table_pipelined(s, q_soa, N, nullptr, Kmax, nullptr, near_contact_count, &fell_back, table16);
  if (rc || !fell_back) return rc;
  // a block's FP64 work list overflowed: the plain path (which splits such blocks), then narrow on the host
  std::vector<int32_t> wide((size_t)N * (1 + 2 * Kmax));
  if ((rc = sirrt_safe_intervals_batch(s, q_soa, N, wide.data(), Kmax, nullptr, near_contact_count))) return rc;
  for (size_t i = 0; i < wide.size(); i++) table16[i] = (int16_t)wide[i];
  return SIRRT_OK;
}

extern "C" int sirrt_collision_bits_f64(sirrt_scene* s, const double* q_aos, int64_t N, int semantics, uint32_t* bits) {
  if (!s || !q_aos || !bits || N < 0 || (semantics != SIRRT_SEM_INTERVAL && semantics != SIRRT_SEM_FRAME)) return fail(SIRRT_E_ARG, "sirrt_collision_bits_f64: bad argument");
  if (N == 0) return SIRRT_OK;
  CU(cudaSetDevice(s->device));
  const void* dq = nullptr;
  int rc = stage_in(s, s->q, q_aos, (size_t)N * s->hs.dof * sizeof(double), &dq, PIN_Q, 64 << 10);
  if (rc) return rc;
  QView q{nullptr, (const double*)dq, N, 0, N, s->hs.dof};
  return bits_and_table(s, q, N, semantics, nullptr, 0, bits, nullptr);
}

// §7.3 edge table, windowed: the departure-time scan of one edge (PlannerConnect.cpp:494-537) moves every state's frame
// by one per step, so the table only needs, per state, the frames around the current query.
extern "C" int sirrt_collision_bits_window_f64(sirrt_scene* s, const double* q_aos, int64_t N, int semantics, const int32_t* first_tile, int n_tiles,
                                               uint32_t* bits) {
  if (!s || !q_aos || !bits || !first_tile || N < 0 || n_tiles < 1 || (semantics != SIRRT_SEM_INTERVAL && semantics != SIRRT_SEM_FRAME))
    return fail(SIRRT_E_ARG, "sirrt_collision_bits_window_f64: bad argument");
  if (N == 0) return SIRRT_OK;
  if (is_device_ptr(first_tile)) return fail(SIRRT_E_ARG, "sirrt_collision_bits_window_f64: first_tile is a host array");
  CU(cudaSetDevice(s->device));
  const void* dq = nullptr;
  int rc = stage_in(s, s->q, q_aos, (size_t)N * s->hs.dof * sizeof(double), &dq, PIN_Q, 64 << 10);
  if (rc) return rc;
  const int W = s->tb.n_tiles;
  std::vector<int2> rt((size_t)N);
  for (int64_t i = 0; i < N; i++) {
    const int lo = std::max(0, std::min(first_tile[i], W));
    rt[i] = make_int2(lo, std::min(W, lo + n_tiles));
  }
  const int2* d_rt = nullptr;
  const size_t rb = (size_t)N * sizeof(int2);
  if (rb <= PIN_FR_CAP) { memcpy(s->pinned + PIN_FR, rt.data(), rb); d_rt = (const int2*)(s->pinned + PIN_FR); }  // read in place
  else {
    if (s->frames.ensure(rb)) return fail(SIRRT_E_CUDA, "cudaMalloc(row windows) failed");
    CU(cudaMemcpyAsync(s->frames.p, rt.data(), rb, cudaMemcpyHostToDevice, s->stream));
    CU(sirrt_sync(s->stream));  // rt is a local
    d_rt = s->frames.as<int2>();
  }
  QView q{nullptr, (const double*)dq, N, 0, N, s->hs.dof};
  return bits_and_table(s, q, N, semantics, nullptr, 0, bits, nullptr, d_rt, n_tiles);
}

// ---- a4: independent (configuration, frame) cells ------------------------------------------------
// flag_when_done: the results land in the pinned host block and nothing follows the last k_recheck, which then writes the
// completion flag (the caller waits with scene_wait)
// out_zeroed: d_out reads zero (planner-sized batches then split the pairs of a cell over several warps that only write hits)
static int run_cells(sirrt_scene* s, QView q, long long N, const int32_t* d_frames, uint8_t* d_out, int semantics, int frame_one = 0,
                     bool flag_when_done = false, bool out_zeroed = false) {
  cudaStream_t st = s->stream;
  s->done_armed = false;
  if (s->work.ensure((size_t)s->wl_cap * sizeof(WorkItem))) return fail(SIRRT_E_CUDA, "cudaMalloc(work list) failed");
  if (s->counts.ensure(4 * sizeof(unsigned))) return fail(SIRRT_E_CUDA, "cudaMalloc(counters) failed");
  const bool self_clearing = N <= (long long)(16 << 10);  // latency path: k_recheck's last block clears the counters
  // cells are cheap (one frame each): the work list cannot overflow for chunk sizes below the cap / (L*slots)
  const long long per = std::max(1ll, (long long)s->wl_cap / std::max(1, 2 * s->tb.robot.n_links * s->tb.n_slots));
  for (long long n0 = 0; n0 < N; n0 += per) {
    QView qv = q;
    qv.n0 = n0; qv.cnt = std::min(per, N - n0);
    if (!self_clearing) CU(cudaMemsetAsync(s->counts.p, 0, 4 * sizeof(unsigned), st));
    WorkList wl{s->work.as<WorkItem>(), self_clearing ? s->lazy_cnt : s->counts.as<unsigned>(), s->wl_cap};
    if (self_clearing && s->tb.has_hull && s->defer_max_configs > 0 && s->seed_recheck) {
      if (s->seeds.ensure((size_t)SIRRT_SEED_CAP * sizeof(int4))) return fail(SIRRT_E_CUDA, "cudaMalloc(re-check seeds) failed");
      wl.seeds = s->seeds.as<int4>(); wl.seed_cap = SIRRT_SEED_CAP;
    }
    CellParams P{};
    P.sc = s->d; P.q = qv; P.wl = wl; P.aabb_f = s->d_aabb_f; P.frames = d_frames; P.frame_one = frame_one; P.out = d_out; P.semantics = semantics; P.band = (float)s->band; P.obb_cull = s->obb_cull;
    // hull scenes: every cell appends at most 2*L*slots tasks, the same bound `per` keeps under the list capacity
    const bool defer = s->tb.has_hull && s->defer_max_configs > 0 && qv.cnt <= 64 * s->defer_max_configs;
    if (defer) {
      if (s->tasks.ensure((size_t)s->wl_cap * sizeof(WorkItem))) return fail(SIRRT_E_CUDA, "cudaMalloc(task list) failed");
      P.tasks = WorkList{s->tasks.as<WorkItem>(), wl.count + 3, s->wl_cap};
      if (qv.cnt <= (16 << 10)) {  // FP64 link transforms handed to the task kernel (larger batches redo FK per task)
        if (s->xf.ensure((size_t)qv.cnt * s->tb.robot.n_links * 12 * sizeof(double))) return fail(SIRRT_E_CUDA, "cudaMalloc(link transforms) failed");
        P.xf = s->xf.as<double>();
      }
    }
    P.n_pchunks = 1; P.pairs_per_chunk = 0;
    const long long target = (long long)s->n_sm * 32;
    const int rounds = (s->tb.robot.n_links * s->tb.n_slots + 31) / 32;
    if (out_zeroed && s->slot_split && s->cur_load < SIRRT_SPLIT_MAX_LOAD && rounds >= 2 && qv.cnt * 2 <= target) {
      const int want = (int)std::min<long long>(rounds, target / qv.cnt);
      const int rpc = (rounds + want - 1) / want;  // 32-pair rounds per warp
      P.pairs_per_chunk = 32 * rpc;
      P.n_pchunks = (rounds + rpc - 1) / rpc;
    }
    RecheckParams R{};
    R.sc = s->d; R.q = qv; R.items = wl.items; R.count = wl.count; R.cap = wl.cap;
    R.aabb_d = s->d_aabb_d; R.bits = nullptr; R.n_words = s->tb.n_tiles; R.cells = d_out; R.counters = wl.count;
    R.seeds = wl.seeds; R.seed_cap = wl.seed_cap;
    if (self_clearing) { R.ticket = s->lazy_cnt + 4; R.fin = s->lazy_cnt + 7; R.fin_seq = ++s->fin_seq; }
    if (self_clearing && flag_when_done && n0 + per >= N) { arm_done(s, R); R.publish = (unsigned int*)(s->pinned + PIN_CNT); }
    bool skip_k3 = false;
    const int grid = grid_for(qv.cnt * P.n_pchunks, s->n_sm, 8);
    if (s->tb.has_hull) k_check_cells<true><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
    else k_check_cells<false><<<grid, SIRRT_WARPS_PER_BLOCK * 32, 0, st>>>(s->tb.robot, P);
    if (defer) {
      GjkTaskParams G{};
      G.sc = s->d; G.q = qv; G.items = P.tasks.items; G.count = P.tasks.count; G.cap = P.tasks.cap; G.wl = wl; G.xf = P.xf;
      G.bits = nullptr; G.n_words = s->tb.n_tiles; G.cells = d_out; G.band = (float)s->band;
      if (self_clearing) { G.ticket = s->lazy_cnt + 6; G.publish = R.publish; G.done = R.done; G.done_seq = R.done_seq; G.fin = R.fin; G.fin_seq = R.fin_seq; }
      if (self_clearing && R.done && s->inline_recheck) { G.inline_recheck = s->inline_recheck; G.aabb_d = s->d_aabb_d; skip_k3 = true; }
      s->pred_kind = 1;
      CU(launch_after(self_clearing && s->pdl, k_gjk_tasks, self_clearing ? task_grid(s, 1, qv.cnt) : (unsigned)s->n_sm * 4, SIRRT_WARPS_PER_BLOCK * 32, st, s->tb.robot, G));
      s->stats[0]++;
    }
    if (skip_k3) { s->k3 = R; s->k3_grid = (unsigned)s->n_sm; s->k3_deferred = true; s->stats[0] += 1; }
    else {
      CU(launch_after(self_clearing && s->pdl, k_recheck, (self_clearing && qv.cnt <= 256) ? 16u : (unsigned)s->n_sm * 2, SIRRT_WARPS_PER_BLOCK * 32, st, s->tb.robot, R));
      s->stats[0] += 2;
    }
    CU(cudaGetLastError());
  }
  return SIRRT_OK;
}

static int check_frames_common(sirrt_scene* s, const float* q_soa, const double* q_aos, const int32_t* frames, int64_t N, uint8_t* collided) {
  if (!s || (!q_soa && !q_aos) || !frames || !collided || N < 0) return fail(SIRRT_E_ARG, "sirrt_check_frames_batch: bad argument");
  if (N == 0) return SIRRT_OK;
  if (N > 0x7fffffffll) return fail(SIRRT_E_ARG, "N exceeds 2^31-1 cells per call");
  InFlight in_flight(s->device);
  s->cur_load = in_flight.load;
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  if (s->tb.n_slots == 0 && !dev_ptr(s, frames) && !dev_ptr(s, collided)) {
    // EMPTY WORKLOAD, not a fallback: a scene without obstacles (STRRT/scene_task.json ships "obstacles": []) has
    // no (link, obstacle) pair to test, so no kernel has anything to compute; only the frame range rule remains
    for (int64_t i = 0; i < N; i++) collided[i] = (frames[i] < 0 || frames[i] >= s->tb.T) ? 1 : 0;
    return SIRRT_OK;
  }
  const void *dq = nullptr, *df = nullptr;
  int rc;
  if (N == 1 && q_aos && !dev_ptr(s, q_aos) && !dev_ptr(s, frames) && !dev_ptr(s, collided) && s->hs.dof <= SIRRT_MAXJ) {
    // latency path, one cell: angles and frame travel in the kernel parameters, the answer lands in the pinned block
    QView q1{nullptr, nullptr, 1, 0, 1, s->hs.dof, 1, {}};
    for (int j = 0; j < s->hs.dof; j++) q1.inl[j] = q_aos[j];
    uint8_t* d_out1 = (uint8_t*)(s->pinned + PIN_OUT);
    d_out1[0] = 0;
    if ((rc = run_cells(s, q1, 1, nullptr, d_out1, SIRRT_SEM_FRAME, frames[0], true, true))) return rc;
    CU(scene_wait(s));
    if (s->tb.has_hull) note_tasks(s, 1, ((const unsigned*)(s->pinned + PIN_CNT))[3]);
    collided[0] = d_out1[0];
    return SIRRT_OK;
  }
  if (q_soa) rc = stage_in(s, s->q, q_soa, (size_t)N * s->hs.dof * sizeof(float), &dq, PIN_Q, 64 << 10);
  else rc = stage_in(s, s->q, q_aos, (size_t)N * s->hs.dof * sizeof(double), &dq, PIN_Q, 64 << 10);
  if (rc) return rc;
  if ((rc = stage_in(s, s->frames, frames, (size_t)N * sizeof(int32_t), &df, PIN_FR, PIN_FR_CAP))) return rc;
  const bool odev = dev_ptr(s, collided);
  const bool opin = !odev && (size_t)N <= (size_t)(16 << 10);  // latency path: the kernels write the pinned block in place
  uint8_t* d_out = collided;
  if (opin) { d_out = (uint8_t*)(s->pinned + PIN_OUT); memset(d_out, 0, (size_t)N); }
  else if (!odev) {
    if (s->cells.ensure((size_t)N)) return fail(SIRRT_E_CUDA, "cudaMalloc(cells) failed");
    d_out = s->cells.as<uint8_t>();
  }
  QView q{q_soa ? (const float*)dq : nullptr, q_soa ? nullptr : (const double*)dq, N, 0, N, s->hs.dof};
  if ((rc = run_cells(s, q, N, (const int32_t*)df, d_out, SIRRT_SEM_FRAME, 0, opin, opin))) return rc;
  if (!odev && !opin) CU(cudaMemcpyAsync(collided, d_out, (size_t)N, cudaMemcpyDeviceToHost, s->stream));
  CU(scene_wait(s));
  if (opin && s->tb.has_hull) note_tasks(s, 1, ((const unsigned*)(s->pinned + PIN_CNT))[3]);
  if (opin) memcpy(collided, d_out, (size_t)N);
  return SIRRT_OK;
}
extern "C" int sirrt_check_frames_batch(sirrt_scene* s, const float* q_soa, const int32_t* frames, int64_t N, uint8_t* collided) {
  return check_frames_common(s, q_soa, nullptr, frames, N, collided);
}
extern "C" int sirrt_check_frames_batch_f64(sirrt_scene* s, const double* q_aos, const int32_t* frames, int64_t N, uint8_t* collided) {
  return check_frames_common(s, nullptr, q_aos, frames, N, collided);
}

// ---- single-call drop-ins ------------------------------------------------------------------------
extern "C" int sirrt_safe_intervals(sirrt_scene* s, const double* q, int32_t* lo_hi, int cap, int* n_out) {
  if (!s || !q || !n_out || cap < 0 || (cap > 0 && !lo_hi)) return fail(SIRRT_E_ARG, "sirrt_safe_intervals: bad argument");
  InFlight in_flight(s->device);
  s->cur_load = in_flight.load;
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  const int dof = s->hs.dof, W = s->tb.n_tiles, T = s->tb.T;
  const int Kmax = std::max(1, std::min(cap, (T + 1) / 2));
  const size_t tbytes = (size_t)(1 + 2 * Kmax) * sizeof(int32_t);
  if (tbytes > PIN_OUT_CAP) return fail(SIRRT_E_ARG, "sirrt_safe_intervals: cap too large");
  if (s->tb.n_slots == 0) {  // EMPTY WORKLOAD (no obstacle, no pair test, nothing to launch): one interval [0, T-1], CollisionManager.cpp:509-517
    *n_out = 1;
    if (cap < 1) return fail(SIRRT_E_OVERFLOW, "sirrt_safe_intervals: more intervals than cap");
    lo_hi[0] = 0; lo_hi[1] = T - 1;
    return SIRRT_OK;
  }
  // the row of this call lives in an internal buffer that reads zero between calls (slot-split launches OR into it; the
  // fused interval reduction clears it again)
  if (!s->row1.p) {
    if (s->row1.ensure((size_t)W * 4 + 64)) return fail(SIRRT_E_CUDA, "cudaMalloc failed");
    CU(cudaMemsetAsync(s->row1.p, 0, (size_t)W * 4 + 64, s->stream));
  }
  // latency path: q is read from, and the interval table written to, the pinned block in place; one synchronisation
  double* dq = (double*)(s->pinned + PIN_Q);
  if (dev_ptr(s, q)) { CU(cudaMemcpyAsync(dq, q, sizeof(double) * dof, cudaMemcpyDeviceToHost, s->stream)); CU(sirrt_sync(s->stream)); }
  else memcpy(dq, q, sizeof(double) * dof);
  uint32_t* d_bits = s->row1.as<uint32_t>();
  const int32_t* host_table = (const int32_t*)(s->pinned + PIN_OUT);
  QView qv{nullptr, dq, 1, 0, 1, dof, 0, {}};
  if (dof <= SIRRT_MAXJ) { qv.inl_on = 1; for (int j = 0; j < dof; j++) qv.inl[j] = dq[j]; }  // angles travel in the kernel parameters
  for (int attempt = 0; attempt < 2; attempt++) {
    bool fused = false;
    int rc = run_bits(s, qv, 1, SIRRT_SEM_INTERVAL, d_bits, attempt == 0, (int32_t*)host_table, Kmax, &fused, nullptr, 0, true, true);
    if (rc) return rc;
    if (!fused) {
      if ((rc = run_intervals(s, d_bits, 1, (int32_t*)host_table, Kmax))) return rc;
      CU(cudaMemsetAsync(d_bits, 0, (size_t)W * 4, s->stream));
    }
    CU(scene_wait(s));
    if (attempt == 0 && finish_lazy(s)) continue;
    break;
  }
  *n_out = host_table[0];
  const int n = std::min(host_table[0], std::min(cap, Kmax));
  for (int i = 0; i < 2 * n; i++) lo_hi[i] = host_table[1 + i];
  if (host_table[0] > cap) return fail(SIRRT_E_OVERFLOW, "sirrt_safe_intervals: more intervals than cap");
  return SIRRT_OK;
}

extern "C" int sirrt_check_frame(sirrt_scene* s, const double* q, int frame, int* collided) {
  if (!s || !q || !collided) return fail(SIRRT_E_ARG, "sirrt_check_frame: bad argument");
  int32_t f = frame;
  uint8_t out = 0;
  int rc = check_frames_common(s, nullptr, q, &f, 1, &out);
  if (rc) return rc;
  *collided = out;
  return SIRRT_OK;
}

// ---- a5: edge validation (PlannerConnect::is_collision_motion, PlannerConnect.cpp:862-897) -------
struct EdgeParams {
  const double *q0, *q1, *t0, *t1;
  const long long* offs;  // [n_edges+1] first cell of each edge; steps = offs[e+1]-offs[e]-1, 0 cells = gated edge
  long long n_edges, n_cells;
  int dof;
  double* q_cells;   // [n_cells][dof]
  int32_t* frames;   // [n_cells]
};

// one thread per cell: q = q0 + dir*s/steps (:885), frame = (int)(t0 + (t1-t0)*s/steps) (:886).
// Explicit _rn intrinsics keep nvcc from contracting into FMA so the values equal the host's.
__global__ void k_edge_cells(const __grid_constant__ EdgeParams P) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= P.n_cells) return;
  long long lo = 0, hi = P.n_edges;  // largest e with offs[e] <= c
  while (hi - lo > 1) { const long long mid = (lo + hi) >> 1; if (P.offs[mid] <= c) lo = mid; else hi = mid; }
  const long long e = lo;
  const double s = (double)(c - P.offs[e]), steps = (double)(P.offs[e + 1] - P.offs[e] - 1);
  for (int i = 0; i < P.dof; i++) {
    const double a = P.q0[e * P.dof + i], dir = __dsub_rn(P.q1[e * P.dof + i], a);
    P.q_cells[c * P.dof + i] = __dadd_rn(a, __ddiv_rn(__dmul_rn(dir, s), steps));
  }
  const double t0 = P.t0[e], dt = __dsub_rn(P.t1[e], t0);
  const double tf = __dadd_rn(t0, __ddiv_rn(__dmul_rn(dt, s), steps));
  P.frames[c] = (tf >= 2147483647.0 || tf <= -2147483648.0 || tf != tf) ? -1 : (int)tf;
}

struct EdgeFirstParams {
  const long long* offs;
  const uint8_t* cells;
  const double* q_cells;
  long long n_edges;
  int dof;
  int32_t* first;     // in: -1 (to be evaluated) or the host's gate code (-2 / -3)
  double* coord;      // [n_edges][dof] or null
};
__global__ void k_edge_first(const __grid_constant__ EdgeFirstParams P) {
  const int lane = threadIdx.x & 31;
  const long long e = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (e >= P.n_edges) return;
  const long long c0 = P.offs[e], c1 = P.offs[e + 1];
  int found = -1;
  for (long long b = c0; b < c1 && found < 0; b += 32) {
    const long long c = b + lane;
    const unsigned m = __ballot_sync(SIRRT_FULL, c < c1 && P.cells[c]);
    if (m) found = (int)(b - c0) + __ffs(m) - 1;
  }
  if (c1 > c0 && lane == 0) P.first[e] = found;
  if (P.coord && found >= 0)
    for (int i = lane; i < P.dof; i += 32) P.coord[e * P.dof + i] = P.q_cells[(c0 + found) * P.dof + i];
}

extern "C" int sirrt_edge_check_batch(sirrt_scene* s, const double* q0, const double* q1, const double* t0, const double* t1,
                                      int64_t n_edges, double vmax, int32_t* first, double* coord) {
  if (!s || !q0 || !q1 || !t0 || !t1 || !first || n_edges < 0) return fail(SIRRT_E_ARG, "sirrt_edge_check_batch: bad argument");
  if (n_edges == 0) return SIRRT_OK;
  if (is_device_ptr(q0) || is_device_ptr(q1) || is_device_ptr(t0) || is_device_ptr(t1) || is_device_ptr(first) || (coord && is_device_ptr(coord)))
    return fail(SIRRT_E_ARG, "sirrt_edge_check_batch takes host buffers (edge step counts are planned on the host)");
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  const int dof = s->hs.dof;
  std::vector<long long> offs(n_edges + 1);
  long long cells = 0;
  for (int64_t e = 0; e < n_edges; e++) {
    offs[e] = cells;
    first[e] = -1;
    const double a = t0[e], b = t1[e];
    if (a >= b) { first[e] = -2; continue; }  // :865
    double n2 = 0;
    for (int i = 0; i < dof; i++) { const double d = q1[e * dof + i] - q0[e * dof + i]; n2 += d * d; }
    const double norm = sqrt(n2);
    if (vmax > 0 && (norm * ((double)s->hs.fps) / (b - a)) > (vmax + 0.001)) { first[e] = -3; continue; }  // :872
    const double x1 = norm / 0.1, x2 = b - a + 0.5;
    long long steps = x1 < 2e9 ? (long long)(int)x1 : 2000000000ll, s2 = x2 < 2e9 ? (long long)(int)x2 : 2000000000ll;  // :879
    steps = std::max(std::max(steps, s2), 1ll);
    cells += steps + 1;
    if (cells > 0x7fffffffll) return fail(SIRRT_E_ARG, "edge batch expands to more than 2^31-1 cells");
  }
  offs[n_edges] = cells;
  if (cells == 0) return SIRRT_OK;
  cudaStream_t st = s->stream;
  const size_t eb = (size_t)n_edges * dof * sizeof(double);
  size_t need = 2 * eb + 2 * (size_t)n_edges * sizeof(double) + (size_t)(n_edges + 1) * sizeof(long long) + (size_t)n_edges * sizeof(int32_t) + eb + 256;
  if (s->misc.ensure(need)) return fail(SIRRT_E_CUDA, "cudaMalloc(edges) failed");
  char* p = (char*)s->misc.p;
  double* d_q0 = (double*)p; p += eb;
  double* d_q1 = (double*)p; p += eb;
  double* d_t0 = (double*)p; p += n_edges * sizeof(double);
  double* d_t1 = (double*)p; p += n_edges * sizeof(double);
  long long* d_offs = (long long*)p; p += (n_edges + 1) * sizeof(long long);
  double* d_coord = (double*)p; p += eb;
  int32_t* d_first = (int32_t*)p;
  if (s->misc2.ensure((size_t)cells * dof * sizeof(double))) return fail(SIRRT_E_CUDA, "cudaMalloc(edge cells) failed");
  if (s->frames.ensure((size_t)cells * sizeof(int32_t))) return fail(SIRRT_E_CUDA, "cudaMalloc(edge frames) failed");
  if (s->cells.ensure((size_t)cells)) return fail(SIRRT_E_CUDA, "cudaMalloc(cells) failed");
  CU(cudaMemcpyAsync(d_q0, q0, eb, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_q1, q1, eb, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_t0, t0, n_edges * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_t1, t1, n_edges * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_offs, offs.data(), (n_edges + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_first, first, n_edges * sizeof(int32_t), cudaMemcpyHostToDevice, st));
  if (coord) CU(cudaMemsetAsync(d_coord, 0, eb, st));
  EdgeParams E{d_q0, d_q1, d_t0, d_t1, d_offs, n_edges, cells, dof, s->misc2.as<double>(), s->frames.as<int32_t>()};
  k_edge_cells<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(E);
  QView qv{nullptr, s->misc2.as<double>(), cells, 0, cells, dof};
  int rc = run_cells(s, qv, cells, s->frames.as<int32_t>(), s->cells.as<uint8_t>(), SIRRT_SEM_FRAME);
  if (rc) return rc;
  EdgeFirstParams F{d_offs, s->cells.as<uint8_t>(), s->misc2.as<double>(), n_edges, dof, d_first, coord ? d_coord : nullptr};
  k_edge_first<<<(unsigned)((n_edges + 7) / 8), 256, 0, st>>>(F);
  s->stats[0] += 2;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(first, d_first, n_edges * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  if (coord) CU(cudaMemcpyAsync(coord, d_coord, eb, cudaMemcpyDeviceToHost, st));
  CU(sirrt_sync(st));
  return SIRRT_OK;
}

// ---- f4: ST-RRT* motion validation (SpaceTimeMotionValidator::checkMotion,
// STRRT_Planner/src/OMPLValidators/SpaceTimeMotionValidator.cpp:22-74 and :76-127; the state test is
// StateValidityCheckerFunctor::isValid, StateValidityCheckerFunctor.cpp:29-47 -> CollisionManager.cpp:330-336) ---
// Cell layout per motion: cell 0 = s2 (tested first, :25 / :82), cells 1..steps = interpolated states i = 0..steps-1
// at fraction i/steps (:54-57): q = from + (to - from) * f (ompl RealVectorStateSpace::interpolate), time likewise
// (TimeStateSpace::interpolate), frame = (int)((float)time * fps) in float32 (CollisionManager.cpp:333).
__global__ void k_motion_cells(const __grid_constant__ EdgeParams P, const float fps) {
  const long long c = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= P.n_cells) return;
  long long lo = 0, hi = P.n_edges;
  while (hi - lo > 1) { const long long mid = (lo + hi) >> 1; if (P.offs[mid] <= c) lo = mid; else hi = mid; }
  const long long e = lo;
  const long long k = c - P.offs[e];
  const double steps = (double)(P.offs[e + 1] - P.offs[e] - 1);
  double tm;
  if (k == 0) {
    for (int i = 0; i < P.dof; i++) P.q_cells[c * P.dof + i] = P.q1[e * P.dof + i];
    tm = P.t1[e];
  } else {
    const double f = __ddiv_rn((double)(k - 1), steps);
    for (int i = 0; i < P.dof; i++) {
      const double a = P.q0[e * P.dof + i];
      P.q_cells[c * P.dof + i] = __dadd_rn(a, __dmul_rn(__dsub_rn(P.q1[e * P.dof + i], a), f));
    }
    tm = __dadd_rn(P.t0[e], __dmul_rn(__dsub_rn(P.t1[e], P.t0[e]), f));
  }
  const float ff = __fmul_rn((float)tm, fps);
  P.frames[c] = (ff >= 2147483647.0f || ff <= -2147483648.0f || ff != ff) ? -1 : (int)ff;
}

extern "C" int sirrt_motion_check_batch(sirrt_scene* s, const double* q0, const double* q1, const double* t0, const double* t1,
                                        int64_t n, double vmax, int variant, int32_t* result, double* last_valid_fraction) {
  if (!s || !q0 || !q1 || !t0 || !t1 || !result || n < 0 || (variant != 0 && variant != 1))
    return fail(SIRRT_E_ARG, "sirrt_motion_check_batch: bad argument");
  if (n == 0) return SIRRT_OK;
  if (is_device_ptr(q0) || is_device_ptr(q1) || is_device_ptr(t0) || is_device_ptr(t1) || is_device_ptr(result) ||
      (last_valid_fraction && is_device_ptr(last_valid_fraction)))
    return fail(SIRRT_E_ARG, "sirrt_motion_check_batch takes host buffers (step counts are planned on the host)");
  CU(cudaSetDevice(s->device));
  memset(s->stats, 0, sizeof s->stats);
  const int dof = s->hs.dof;
  const double fps = (double)s->hs.fps;
  std::vector<long long> offs(n + 1);
  std::vector<int32_t> gate(n, 0);
  long long cells = 0;
  for (int64_t e = 0; e < n; e++) {
    offs[e] = cells;
    cells += 1;  // s2
    const double dt = t1[e] - t0[e];
    if (dt <= 0) { gate[e] = -2; continue; }  // :41 / :98
    double n2 = 0;
    for (int i = 0; i < dof; i++) { const double d = fabs(q1[e * dof + i] - q0[e * dof + i]); n2 += d * d; }
    const double norm = sqrt(n2);
    if (vmax > 0 && norm / dt > vmax) { gate[e] = -3; continue; }  // :46 / :103
    const double x1 = norm / 0.1, x2 = (variant == 0 ? 2.0 : 1.0) * dt * fps + 0.5;  // :51 / :108
    const long long a = x1 < 2e9 ? (long long)(int)x1 : 2000000000ll, b = x2 < 2e9 ? (long long)(int)x2 : 2000000000ll;
    cells += std::max(a, b) + 2;
    if (cells > 0x7fffffffll) return fail(SIRRT_E_ARG, "motion batch expands to more than 2^31-1 cells");
  }
  offs[n] = cells;
  cudaStream_t st = s->stream;
  const size_t eb = (size_t)n * dof * sizeof(double);
  size_t need = 2 * eb + 2 * (size_t)n * sizeof(double) + (size_t)(n + 1) * sizeof(long long) + (size_t)n * sizeof(int32_t) + 256;
  if (s->misc.ensure(need)) return fail(SIRRT_E_CUDA, "cudaMalloc(motions) failed");
  char* p = (char*)s->misc.p;
  double* d_q0 = (double*)p; p += eb;
  double* d_q1 = (double*)p; p += eb;
  double* d_t0 = (double*)p; p += n * sizeof(double);
  double* d_t1 = (double*)p; p += n * sizeof(double);
  long long* d_offs = (long long*)p; p += (n + 1) * sizeof(long long);
  int32_t* d_first = (int32_t*)p;
  if (s->misc2.ensure((size_t)cells * dof * sizeof(double))) return fail(SIRRT_E_CUDA, "cudaMalloc(motion cells) failed");
  if (s->frames.ensure((size_t)cells * sizeof(int32_t))) return fail(SIRRT_E_CUDA, "cudaMalloc(motion frames) failed");
  if (s->cells.ensure((size_t)cells)) return fail(SIRRT_E_CUDA, "cudaMalloc(cells) failed");
  CU(cudaMemcpyAsync(d_q0, q0, eb, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_q1, q1, eb, cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_t0, t0, n * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_t1, t1, n * sizeof(double), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_offs, offs.data(), (n + 1) * sizeof(long long), cudaMemcpyHostToDevice, st));
  EdgeParams E{d_q0, d_q1, d_t0, d_t1, d_offs, n, cells, dof, s->misc2.as<double>(), s->frames.as<int32_t>()};
  k_motion_cells<<<(unsigned)((cells + 255) / 256), 256, 0, st>>>(E, (float)s->hs.fps);
  QView qv{nullptr, s->misc2.as<double>(), cells, 0, cells, dof};
  int rc = run_cells(s, qv, cells, s->frames.as<int32_t>(), s->cells.as<uint8_t>(), SIRRT_SEM_FRAME);
  if (rc) return rc;
  EdgeFirstParams F{d_offs, s->cells.as<uint8_t>(), s->misc2.as<double>(), n, dof, d_first, nullptr};
  k_edge_first<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(F);
  s->stats[0] += 2;
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(result, d_first, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  CU(sirrt_sync(st));
  for (int64_t e = 0; e < n; e++) {
    const int32_t raw = result[e];  // index of the first colliding cell of this motion, -1 = none
    const long long steps = offs[e + 1] - offs[e] - 1;
    double frac = 0.0;  // lastValid.second (:79, :121)
    int32_t r;
    if (raw == 0) r = -4;                 // s2 invalid (:25 / :82)
    else if (gate[e]) r = gate[e];
    else if (raw < 0) { r = -1; frac = (double)(steps - 1) / (double)steps; }
    else { r = raw - 1; if (r >= 1) frac = (double)(r - 1) / (double)steps; }
    result[e] = r;
    if (last_valid_fraction) last_valid_fraction[e] = frac;
  }
  return SIRRT_OK;
}
