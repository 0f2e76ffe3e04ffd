// sirrt_kernels.cuh -- the collision kernels of libsirrt_b200 (sm_100a).
//
//   k_collide_bits   a3/a5: one warp per (configuration, chunk of 32-frame tiles).  FK once per
//                    configuration (FP64), then per tile: level-0 cull of link bounding spheres
//                    against the swept tile bounds of the dynamic slots (lanes = slots), and for the
//                    survivors the exact FP32 narrowphase with lanes = 32 consecutive frames, so that
//                    one __ballot_sync yields one word of the collision bitmask.
//   k_check_cells    a4: one warp per (configuration, frame) cell, lanes = (link, slot) pairs.
//   k_recheck        FP64 re-evaluation of every pair test that fell inside the FP32 uncertainty
//                    band (same formulas as the CPU oracle) -> sets bits / cell flags.
//   k_bits_to_intervals  warp ballot/scan: bit rows -> inclusive safe intervals (CollisionManager.cpp:507-546)
#pragma once
#include "sirrt_device.cuh"

#define SIRRT_WARPS_PER_BLOCK 8
#define SIRRT_QCAP 96

struct QView {           // configuration batch: float SoA [dof][N] or double AoS [N][dof]
  const float* f_soa;
  const double* d_aos;
  long long N;      // row stride of the SoA layout (= total configurations of the caller's batch)
  long long n0, cnt; // the kernel processes configurations n0 .. n0+cnt-1
  int dof;
  int inl_on;        // latency path, single configuration: the angles travel in the kernel parameters (no bus read)
  double inl[SIRRT_MAXJ];
};
__device__ __forceinline__ double load_q(const QView& q, long long n, int j) {
  if (q.inl_on) return q.inl[j];
  return q.d_aos ? q.d_aos[n * q.dof + j] : (double)q.f_soa[(long long)j * q.N + n];
}

struct WorkList {
  WorkItem* items;
  unsigned int* count;   // [0] entries pushed (may exceed cap), [1] exact tests, [2] near contacts
  unsigned int cap;
  // latency path: seeds[i] = support-point ids of the FP32 GJK simplex item i ended on (x = -1: none); the FP64 re-check
  // starts from it (warp_gjk seed_ids).  Every producer writes its entry, so a slot never carries a previous call's ids.
  int4* seeds;
  unsigned int seed_cap;
};
__device__ __forceinline__ void push_work(const WorkList& wl, int config, int frame, int link, int slot, int shifted, const int* seed = nullptr) {
  unsigned int i = atomicAdd(wl.count, 1u);
  if (i < wl.cap) {
    WorkItem w;
    w.config = config; w.frame = frame; w.link = (short)link; w.shifted = (short)shifted; w.slot = slot;
    wl.items[i] = w;
    if (wl.seeds && i < wl.seed_cap) wl.seeds[i] = seed ? make_int4(seed[0], seed[1], seed[2], seed[3]) : make_int4(-1, -1, -1, -1);
  }
}

// Programmatic dependent launch (latency path): the kernels of one call are launched back to back with the
// programmatic-stream-serialization attribute; a producer lets its dependents start launching at once (their blocks
// become resident and run their prologue while the producer still works) and a consumer waits here until the producer
// grid has completed and its writes are visible.  Both are no-ops for a kernel launched without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Warp-aggregated append of deferred tasks: every lane passes the links (bit l of mask0: the true geometry, bit l of mask2:
// the shifted functor geometry) whose pair test (link l, slot, this lane's frame) it defers; ONE atomicAdd per warp
// reserves the entries.  All 32 lanes must call.  (A lane-by-lane push_work is 32 same-address atomics and one L2 round
// trip per (slot, link) on the warp's critical path.)
__device__ __forceinline__ void push_tasks_warp(const WorkList& wl, unsigned mask0, unsigned mask2, int config, int frame, int slot, int lane) {
  const int mine = __popc(mask0) + __popc(mask2);
  int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(SIRRT_FULL, incl, o); if (lane >= o) incl += u; }
  const int total = __shfl_sync(SIRRT_FULL, incl, 31);
  if (total == 0) return;
  unsigned int base = 0;
  if (lane == 0) base = atomicAdd(wl.count, (unsigned int)total);
  unsigned int i = __shfl_sync(SIRRT_FULL, base, 0) + (unsigned int)(incl - mine);
  for (int pass = 0; pass < 2; pass++) {
    unsigned m = pass ? mask2 : mask0;
    while (m) {
      const int l = __ffs(m) - 1;
      m &= m - 1;
      if (i < wl.cap) {
        WorkItem w;
        w.config = config; w.frame = frame; w.link = (short)l; w.shifted = (short)(pass ? 2 : 0); w.slot = slot;
        wl.items[i] = w;
      }
      i++;
    }
  }
}

struct WarpCtx {   // per-warp shared memory: the planned robot's links in world space (FP32)
  float a[SIRRT_MAXL][3], b[SIRRT_MAXL][3], r[SIRRT_MAXL];   // capsule core end points + radius
  float bc[SIRRT_MAXL][3], br[SIRRT_MAXL];                   // bounding sphere
  float R[SIRRT_MAXL][9], t[SIRRT_MAXL][3];                  // link transform
  float alo[SIRRT_MAXL][3], ahi[SIRRT_MAXL][3];              // coal AABB (QUIRK 7 broadphase gate)
  float tlo[SIRRT_MAXL][3], thi[SIRRT_MAXL][3];              // tight world AABB (cull before the narrowphase)
  float oc[SIRRT_MAXL][3], oh[SIRRT_MAXL][3];                // oriented box of the link: world centre of its local AABB, local half extents (axes = columns of R)
  int type[SIRRT_MAXL];
  int qn;
  unsigned long long q[SIRRT_QCAP];                          // cooperative (hull) task queue
};

// FK in FP64 (warp form: lane j takes the sincos of joint j, lane l ends up with link l), then lane l narrows link l to
// the FP32 world-space link table.
__device__ inline void setup_links(const DRobot& rb, const DScene& sc, const QView& qv, long long n, WarpCtx& W, int lane, double* xf_out = nullptr) {
  const double my_q = (lane >= 1 && lane <= rb.n_joints && lane <= qv.dof) ? load_q(qv, n, lane - 1) : 0.0;  // lane e: chain element e = joint e - 1
  LinkXfD X;
#ifdef SIRRT_GJK_STATS
  const long long t_s0 = clock64() + (long long)(my_q == 1e300);
  long long t_s1 = 0;
  if (g_fk_mode == 1) fk_link_of_lane_serial(rb, __shfl_down_sync(SIRRT_FULL, my_q, 1), lane, X, &t_s1);
  else fk_link_of_lane(rb, my_q, lane, X, &t_s1);
  const long long t_s2 = clock64() + (long long)(X.t[0] == 1e300);
  PHASE_ADD(5, t_s1 - t_s0); PHASE_ADD(6, t_s2 - t_s1);
#else
  fk_link_of_lane(rb, my_q, lane, X);
#endif
  if (lane < rb.n_links) {
    {
      const int l = lane;
      const DShape& sh = sc.shapes[rb.link_shape[l]];
      const double* R = X.R;
      const double* t = X.t;
      for (int k = 0; k < 9; k++) W.R[l][k] = (float)R[k];
      for (int k = 0; k < 3; k++) W.t[l][k] = (float)t[k];
      if (xf_out) {  // FP64 link transforms for the deferred GJK tasks of this configuration (k_gjk_tasks)
        double* o = xf_out + ((size_t)(n - qv.n0) * rb.n_links + l) * 12;
        for (int k = 0; k < 9; k++) o[k] = R[k];
        for (int k = 0; k < 3; k++) o[9 + k] = t[k];
      }
      W.type[l] = sh.type;
      W.r[l] = sh.radius;
      W.br[l] = sh.br;
      // Everything below is derived from the NARROWED transform in FP32 (FP64 instructions are what a single warp waits for
      // on this path): the values feed culls with explicit margins and FP32 tests that only decide outside the band, so
      // the ~1e-7 m this adds is inside their budget; the boxes get an outward margin instead of directed rounding.
      float Rf[9], tf[3];
#pragma unroll
      for (int k = 0; k < 9; k++) Rf[k] = W.R[l][k];
#pragma unroll
      for (int k = 0; k < 3; k++) tf[k] = W.t[l][k];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        W.bc[l][k] = Rf[3 * k] * sh.bc[0] + Rf[3 * k + 1] * sh.bc[1] + Rf[3 * k + 2] * sh.bc[2] + tf[k];
        W.a[l][k] = tf[k] - Rf[3 * k + 2] * sh.half_len;  // capsule core: local (0,0,+-half_len)
        W.b[l][k] = tf[k] + Rf[3 * k + 2] * sh.half_len;
      }
      float cw[3];  // world position of the local AABB centre
#pragma unroll
      for (int k = 0; k < 3; k++) cw[k] = Rf[3 * k] * sh.ac[0] + Rf[3 * k + 1] * sh.ac[1] + Rf[3 * k + 2] * sh.ac[2] + tf[k];
#pragma unroll
      for (int k = 0; k < 3; k++) { W.oc[l][k] = cw[k]; W.oh[l][k] = 0.5f * (sh.amax[k] - sh.amin[k]); }
#pragma unroll
      for (int k = 0; k < 3; k++) {  // tight world AABB: centre R*ac + t, half extents sum_j |R_kj| (amax_j - amin_j)/2, widened
        float e = 0.f;
#pragma unroll
        for (int j = 0; j < 3; j++) e += fabsf(Rf[3 * k + j]) * 0.5f * (sh.amax[j] - sh.amin[j]);
        const float m = 2e-6f * (1.f + fabsf(cw[k]) + e);
        W.tlo[l][k] = cw[k] - e - m; W.thi[l][k] = cw[k] + e + m;
      }
      // coal CollisionObject::computeAABB (identity rotation, |R - I| <= 1e-12 elementwise -> translated local AABB); the
      // exact FP64 comparison only runs when the narrowed matrix looks like the identity at all
      bool ident = true;
#pragma unroll
      for (int k = 0; k < 9; k++) ident &= fabsf(Rf[k] - ((k % 4) == 0 ? 1.f : 0.f)) <= 1e-6f;
      if (ident)
        for (int k = 0; k < 9; k++) ident &= fabs(R[k] - ((k % 4) == 0 ? 1.0 : 0.0)) <= 1e-12;
      if (ident) {
#pragma unroll
        for (int k = 0; k < 3; k++) { W.alo[l][k] = sh.amin[k] + tf[k]; W.ahi[l][k] = sh.amax[k] + tf[k]; }
      } else {
#pragma unroll
        for (int k = 0; k < 3; k++) { W.alo[l][k] = cw[k] - sh.ar; W.ahi[l][k] = cw[k] + sh.ar; }
      }
    }
  }
  if (lane == 0) W.qn = 0;
  __syncwarp();
#ifdef SIRRT_GJK_STATS
  PHASE_ADD(7, clock64() - t_s2);
#endif
}

__device__ __forceinline__ int classify(float d, float band) { return d < -band ? RES_HIT : (d > band ? RES_FREE : RES_AMBIG); }

// lane-level exact test: capsule link vs box / sphere / capsule at pose (R, t)
__device__ inline int lane_exact(const WarpCtx& W, int l, const DShape& sh, const float* R, const float* t, bool ident, float band) {
  const float *a = W.a[l], *b = W.b[l];
  float d;
  if (sh.type == SH_SPHERE) d = sqrtf(pt_seg_dist2(t, a, b)) - W.r[l] - sh.radius;
  else if (sh.type == SH_BOX) {
    float da[3], db[3], la[3], lb[3];
    sub3(da, a, t); sub3(db, b, t);
    if (ident) { la[0] = da[0]; la[1] = da[1]; la[2] = da[2]; lb[0] = db[0]; lb[1] = db[1]; lb[2] = db[2]; }
    else { matTvec(la, R, da); matTvec(lb, R, db); }
    d = sqrtf(seg_box_dist2(la, lb, sh.half)) - W.r[l];
  } else {  // capsule
    float p2[3], q2[3];
    for (int k = 0; k < 3; k++) { p2[k] = t[k] - R[3 * k + 2] * sh.half_len; q2[k] = t[k] + R[3 * k + 2] * sh.half_len; }
    d = sqrtf(seg_seg_dist2(a, b, p2, q2)) - W.r[l] - sh.radius;
  }
  return classify(d, band);
}

__device__ __forceinline__ void load_pose(const DScene& sc, int slot, int f, float* R, float* t, bool need_R) {
  const float* p = sc.pose_f + (size_t)slot * 12 * sc.Tp + f;
  if (need_R) {
#pragma unroll
    for (int k = 0; k < 9; k++) R[k] = p[(size_t)k * sc.Tp];
  }
#pragma unroll
  for (int k = 0; k < 3; k++) t[k] = p[(size_t)(9 + k) * sc.Tp];
}

// warp-cooperative FP32 test of (link l, shape sh posed at slot/frame): anything involving a hull
__device__ inline int coop_test(const DScene& sc, const float* RA, const float* tA, const DShape& lsh, const DShape& sh, int slot,
                                int f, float band, int lane, int* enc_ids = nullptr, int* n_iter = nullptr, int* last_ids = nullptr) {
  float RB[9], tB[3];
  load_pose(sc, slot, f, RB, tB, true);
  GShape<float> A, B;
  A.type = lsh.type; A.verts = sc.verts_f + 3 * (size_t)lsh.vert_off; A.n = lsh.n_verts; A.cells = (sc.hull_cells && lsh.cell_off >= 0) ? sc.hull_cells + lsh.cell_off : nullptr; A.cand = sc.hull_cand; A.half_len = lsh.half_len;
  A.half[0] = lsh.half[0]; A.half[1] = lsh.half[1]; A.half[2] = lsh.half[2];
  B.type = sh.type; B.verts = sc.verts_f + 3 * (size_t)sh.vert_off; B.n = sh.n_verts; B.cells = (sc.hull_cells && sh.cell_off >= 0) ? sc.hull_cells + sh.cell_off : nullptr; B.cand = sc.hull_cand; B.half_len = sh.half_len;
  B.half[0] = sh.half[0]; B.half[1] = sh.half[1]; B.half[2] = sh.half[2];
  A.rec = (A.cells && sc.hull_rec) ? sc.hull_rec + (size_t)lsh.cell_off * SIRRT_HULL_REC : nullptr;
  B.rec = (B.cells && sc.hull_rec) ? sc.hull_rec + (size_t)sh.cell_off * SIRRT_HULL_REC : nullptr;
  float Rrel[9], trel[3], dt[3];
  matTmul3(Rrel, RA, RB);
  sub3(dt, tB, tA);
  matTvec(trel, RA, dt);
  float rsum = (lsh.type == SH_HULL ? 0.f : lsh.radius) + ((sh.type == SH_SPHERE || sh.type == SH_CAPSULE) ? sh.radius : 0.f);
  return warp_gjk<float, false>(A, B, Rrel, trel, rsum, band, lane, (float*)nullptr, enc_ids, nullptr, nullptr, n_iter, last_ids);
}

// Tight-AABB cull: a pair whose world AABBs are further apart than the FP32 uncertainty band on some axis cannot
// collide (both boxes contain their FP64 shapes: tables are rounded outwards); no decision is taken here
// other than "certainly free", so results are unchanged -- only GJK / closed-form runs are saved.
__device__ __forceinline__ bool tight_overlap(const DScene& sc, const WarpCtx& W, int l, int slot, int f, float band) {
  const float* tb = sc.tight_f + (size_t)slot * 6 * sc.Tp + f;
  bool ov = true;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const float lo = tb[(size_t)k * sc.Tp], hi = tb[(size_t)(3 + k) * sc.Tp];
    ov &= (W.tlo[l][k] - band <= hi) && (W.thi[l][k] + band >= lo);
  }
  return ov;
}

// Oriented-box cull for pair tests that would otherwise become a GJK run (anything involving a hull): the link's local
// AABB posed by the link transform against the obstacle shape's local AABB posed at (RB, tB), 15-axis separating-axis test.
// Both boxes contain their shapes, so a separating axis with more than `margin` of clearance proves the pair FREE; nothing
// else is decided here (results unchanged, GJK runs saved: 40-85 % of the deferred tasks end FREE).  The cross-product axes
// are not normalised (|a x b| <= 1), which only makes the required clearance larger; EPS guards near-parallel axes.
__host__ __device__ inline bool obb_separated(const WarpCtx& W, int l, const DShape& sb, const float* RB, const float* tB, float margin) {
  const float* RA = W.R[l];
  float hb[3], cb[3], d[3], t[3], R[3][3], AR[3][3];
#pragma unroll
  for (int k = 0; k < 3; k++) hb[k] = 0.5f * (sb.amax[k] - sb.amin[k]);
#pragma unroll
  for (int k = 0; k < 3; k++) cb[k] = RB[3 * k] * sb.ac[0] + RB[3 * k + 1] * sb.ac[1] + RB[3 * k + 2] * sb.ac[2] + tB[k];
#pragma unroll
  for (int k = 0; k < 3; k++) d[k] = cb[k] - W.oc[l][k];
#pragma unroll
  for (int i = 0; i < 3; i++) {  // axes = columns of the rotations
    t[i] = RA[i] * d[0] + RA[3 + i] * d[1] + RA[6 + i] * d[2];
#pragma unroll
    for (int j = 0; j < 3; j++) {
      R[i][j] = RA[i] * RB[j] + RA[3 + i] * RB[3 + j] + RA[6 + i] * RB[6 + j];
      AR[i][j] = fabsf(R[i][j]) + 1e-6f;
    }
  }
  const float* ha = W.oh[l];
  bool sep = false;
#pragma unroll
  for (int i = 0; i < 3; i++) sep |= fabsf(t[i]) > ha[i] + hb[0] * AR[i][0] + hb[1] * AR[i][1] + hb[2] * AR[i][2] + margin;
#pragma unroll
  for (int j = 0; j < 3; j++)
    sep |= fabsf(t[0] * R[0][j] + t[1] * R[1][j] + t[2] * R[2][j]) > ha[0] * AR[0][j] + ha[1] * AR[1][j] + ha[2] * AR[2][j] + hb[j] + margin;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      sep |= fabsf(t[i2] * R[i1][j] - t[i1] * R[i2][j]) > ha[i1] * AR[i2][j] + ha[i2] * AR[i1][j] + hb[j1] * AR[i][j2] + hb[j2] * AR[i][j1] + margin;
    }
  }
  return sep;
}

// QUIRK 7 broadphase gate: do the coal AABBs of the TRUE geometries overlap?  FP32 with margin.
__device__ __forceinline__ int aabb_gate(const WarpCtx& W, int l, const float* blo, const float* bhi) {
  const float eps = 1e-5f;
  int res = RES_HIT;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    float s1 = W.alo[l][k] - bhi[k], s2 = blo[k] - W.ahi[l][k];  // > 0: separated on this axis
    float s = fmaxf(s1, s2);
    if (s > eps) return RES_FREE;
    if (s > -eps) res = RES_AMBIG;
  }
  return res;
}

struct BitsParams {
  DScene sc;
  QView q;
  WorkList wl;
  WorkList tasks;       // items != null: hull pair tests are NOT run by the producing warp but appended here for k_gjk_tasks
  const float* aabb_f;  // robot slots: [(slot*6 + k)*Tp + f] coal AABB lo(3) hi(3); may be null
  const int2* row_tiles; // windowed rows (edge tables around a departure time): configuration i only needs tiles [x, y); null = every tile
  double* xf;           // deferred tasks: FP64 link transforms [cnt][n_links][12] written here for k_gjk_tasks; may be null
  uint32_t* bits;       // [N][W]
  int n_words;          // W
  int tiles_per_chunk;
  int n_chunks;
  // small batches: the dynamic slots of a (configuration, tile chunk) are split over n_schunks warps of slots_per_chunk
  // slots each (static slots go with slot chunk 0) and the words are OR-ed into rows the caller has zeroed
  int n_schunks, slots_per_chunk;
  int semantics;
  float band;
  int obb_cull;         // hull pair tests are preceded by the oriented-box separating-axis cull (SIRRT_OPT_OBB_CULL)
};

// process one (link, slot[, shifted]) candidate at lane-frame f, or queue it for cooperative work
#define QPACK(l, slot, sh, ln) (((unsigned long long)(unsigned)(slot) << 16) | ((unsigned long long)(l) << 8) | ((unsigned long long)(sh) << 6) | (unsigned long long)(ln))

template <bool HAS_HULL>
__global__ void __launch_bounds__(SIRRT_WARPS_PER_BLOCK * 32)
k_collide_bits(const __grid_constant__ DRobot rb, const __grid_constant__ BitsParams P) {
  __shared__ WarpCtx ctx[SIRRT_WARPS_PER_BLOCK];
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  WarpCtx& W = ctx[wib];
  const DScene& sc = P.sc;
  const int n_sch = P.n_schunks > 1 ? P.n_schunks : 1;
  // split rows: the static slots of a row are items of their own, 32 (link, slot) pairs each (beside the (tile chunk, slot
  // chunk) items): a latency-bound launch lasts as long as its slowest warp (measured: 35 000 cycles for the warp whose
  // 4 slots all came near the arm, against a mean of 18 000)
  const int n_static_items = (n_sch > 1 && sc.n_static > 0) ? (rb.n_links * sc.n_static + 31) / 32 : 0;
  const int per_row = P.n_chunks * n_sch + n_static_items;
  const long long n_items = P.q.cnt * per_row;
  const long long warp0 = (long long)blockIdx.x * SIRRT_WARPS_PER_BLOCK + wib;
  const long long n_warps = (long long)gridDim.x * SIRRT_WARPS_PER_BLOCK;
  const bool want_true_rob = (sc.parity_mode == 1) || (P.semantics == 0);
  const bool want_shift_rob = (sc.parity_mode == 0);
  const float band = P.band;
  const bool obb_cull = P.obb_cull != 0;
  unsigned int n_exact = 0;  // narrowphase pair tests evaluated by this lane (roofline accounting)

  for (long long item = warp0; item < n_items; item += n_warps) {
    const long long n = P.q.n0 + item / per_row;
    const int sub = (int)(item % per_row);
    const bool static_item = n_sch > 1 && sub >= P.n_chunks * n_sch;
    const int static_p0 = static_item ? (sub - P.n_chunks * n_sch) * 32 : 0;
    const int chunk = static_item ? 0 : sub / n_sch, sch = static_item ? 0 : sub % n_sch;
    const int j_lo = n_sch > 1 ? sch * P.slots_per_chunk : 0, j_hi = n_sch > 1 ? min(j_lo + P.slots_per_chunk, sc.n_dynamic) : sc.n_dynamic;
    int tile_lo = chunk * P.tiles_per_chunk, tile_end = sc.n_tiles;
    if (P.row_tiles) { const int2 rt = P.row_tiles[n - P.q.n0]; tile_lo += rt.x; tile_end = min(rt.y, sc.n_tiles); }
    const int tile_hi = min(tile_lo + P.tiles_per_chunk, tile_end);
    if (P.row_tiles && tile_lo >= tile_hi && !static_item) continue;  // this chunk lies past the row's window
    __syncwarp();
    PHASE_T(t_item);
    setup_links(rb, sc, P.q, n, W, lane, P.xf);
    PHASE_T(t_fk);
    PHASE_ADD(0, 1); PHASE_ADD(1, t_fk - t_item);
    const int L = rb.n_links;

    // evaluates one candidate for this lane; returns hit bit, pushes ambiguous work
    unsigned defer0 = 0, defer2 = 0;  // deferred-task mode: links whose test against the current slot this lane hands to k_gjk_tasks
    auto eval_lane = [&](int l, int slot, const DSlot& S, int f, int frame_tag, bool pass) -> bool {
      // pass: this lane's (link, slot, frame) survived its cull
      const int ltype = W.type[l];
      bool hit = false;
      float R[9], t[3];
      const bool coop_true = HAS_HULL && (ltype == SH_HULL || S.type == SH_HULL);
      bool do_true = pass && (!S.is_robot || want_true_rob);
      bool do_shift = pass && S.is_robot && want_shift_rob && S.shape_shift >= 0;
      if (do_shift) {  // broadphase gate on the true geometries' AABBs
        const float* ab = P.aabb_f + (size_t)slot * 6 * sc.Tp + f;
        float blo[3], bhi[3];
        for (int k = 0; k < 3; k++) { blo[k] = ab[(size_t)k * sc.Tp]; bhi[k] = ab[(size_t)(3 + k) * sc.Tp]; }
        int g = aabb_gate(W, l, blo, bhi);
        if (g == RES_FREE) do_shift = false;
        else if (g == RES_AMBIG) { push_work(P.wl, (int)n, frame_tag, l, slot, 1); do_shift = false; }
      }
      if (do_true && !coop_true) {
        load_pose(sc, slot, f, R, t, S.type != SH_SPHERE && !S.rot_identity);
        int r = lane_exact(W, l, sc.shapes[S.shape], R, t, S.rot_identity != 0, band);
        n_exact++;
        if (r == RES_HIT) hit = true;
        else if (r == RES_AMBIG) push_work(P.wl, (int)n, frame_tag, l, slot, 0);
      }
      bool q_true = do_true && coop_true;
      if (HAS_HULL && q_true && obb_cull) {  // oriented boxes of the link and the obstacle clear of each other: no GJK run
        load_pose(sc, slot, f, R, t, true);
        if (obb_separated(W, l, sc.shapes[S.shape], R, t, 1e-5f + band)) q_true = false;
      }
      bool coop_shift = false;
      if (do_shift) {
        coop_shift = HAS_HULL && (ltype == SH_HULL || S.type_shift == SH_HULL);
        if (coop_shift && obb_cull) {
          load_pose(sc, slot, f, R, t, true);
          if (obb_separated(W, l, sc.shapes[S.shape_shift], R, t, 1e-5f + band)) { coop_shift = false; do_shift = false; }
        }
        if (!coop_shift && do_shift) {
          load_pose(sc, slot, f, R, t, true);
          int r = lane_exact(W, l, sc.shapes[S.shape_shift], R, t, false, band);
          if (r == RES_HIT) hit = true;
          else if (r == RES_AMBIG) push_work(P.wl, (int)n, frame_tag, l, slot, 2);  // gate already passed
        }
      }
      if (HAS_HULL && P.tasks.items) {  // small batches: spread the hull tests over the whole machine (appended by the caller, warp-aggregated)
        if (q_true) defer0 |= 1u << l;
        if (do_shift && coop_shift) defer2 |= 1u << l;
      } else if (HAS_HULL) {
        bool q_shift = do_shift && coop_shift;
        unsigned m1 = __ballot_sync(SIRRT_FULL, q_true), m2 = __ballot_sync(SIRRT_FULL, q_shift);
        int base = W.qn;
        if (q_true) W.q[base + __popc(m1 & ((1u << lane) - 1))] = QPACK(l, slot, 0, lane);
        base += __popc(m1);
        if (q_shift) W.q[base + __popc(m2 & ((1u << lane) - 1))] = QPACK(l, slot, 2, lane);
        __syncwarp();
        if (lane == 0) W.qn = base + __popc(m2);
        __syncwarp();
      }
      return hit;
    };
    // drain the cooperative queue; frame of entry = fbase + entry lane (fbase < 0: static, frame 0)
    auto drain = [&](int fbase, bool& lane_hit, bool& any_hit) {
      if (!HAS_HULL) return;
      int qn = W.qn;
      for (int e = 0; e < qn; e++) {
        unsigned long long u = W.q[e];
        int slot = (int)(u >> 16), l = (int)((u >> 8) & 0xff), shf = (int)((u >> 6) & 3), ln = (int)(u & 63);
        int f = fbase < 0 ? 0 : fbase + ln;
        const DSlot S = sc.slots[slot];
        const DShape& sh = sc.shapes[shf ? S.shape_shift : S.shape];
        int r = coop_test(sc, W.R[l], W.t[l], sc.shapes[rb.link_shape[l]], sh, slot, f, band, lane);
        if (r == RES_HIT) { any_hit = true; if (lane == ln) lane_hit = true; }
        else if (r == RES_AMBIG && lane == 0) push_work(P.wl, (int)n, fbase < 0 ? -1 : f, l, slot, shf);
      }
      __syncwarp();
      if (lane == 0) W.qn = 0;
      __syncwarp();
    };

    // ---- static slots: one test per (link, slot), valid for every frame --------------------
    bool static_hit = false;
    if (n_sch == 1 || static_item) {
      const int n_pairs = static_item ? min(L * sc.n_static, static_p0 + 32) : L * sc.n_static;
      for (int p0 = static_p0; p0 < n_pairs; p0 += 32) {
        int p = p0 + lane;
        bool pass = false;
        int l = 0, slot = 0;
        DSlot S{};
        if (p < n_pairs) {
          l = p / sc.n_static; slot = sc.static_slots[p % sc.n_static];
          S = sc.slots[slot];
          const float* c = sc.bcw_f + (size_t)slot * 3 * sc.Tp;
          float dx = W.bc[l][0] - c[0], dy = W.bc[l][1] - c[sc.Tp], dz = W.bc[l][2] - c[2 * (size_t)sc.Tp];
          float rr = (W.br[l] + S.br) * 1.000001f + 1e-5f + band;
          pass = dx * dx + dy * dy + dz * dz <= rr * rr && tight_overlap(sc, W, l, slot, 0, band);
        }
        bool h = eval_lane(l, slot, S, 0, -1, pass);
        if (HAS_HULL && P.tasks.items) {
          // unsplit rows cut into tile chunks: every chunk's warp walks the static slots (it needs the closed-form static
          // hits for its own words), but the deferred tasks apply to the whole row -- appended once, by chunk 0
          if (n_sch > 1 || chunk == 0) push_tasks_warp(P.tasks, defer0, defer2, (int)n, -1, slot, lane);
          defer0 = defer2 = 0;
        }
        bool lane_hit = false, any_hit = false;
        if (HAS_HULL && W.qn > SIRRT_QCAP - 64) drain(-1, lane_hit, any_hit);
        static_hit |= __any_sync(SIRRT_FULL, h) | any_hit;
      }
      bool lane_hit = false, any_hit = false;
      drain(-1, lane_hit, any_hit);
      static_hit |= any_hit;
    }
    PHASE_T(t_static);
    PHASE_ADD(2, t_static - t_fk);
    if (static_item) {  // split rows: a static hit colours every frame of the row (OR-ed like the dynamic words)
      if (static_hit)
        for (int w = lane; w < P.n_words; w += 32) {
          uint32_t m = 0xffffffffu;
          if (sc.T - w * 32 < 32) m = (1u << (sc.T - w * 32)) - 1u;
          atomicOr(P.bits + (size_t)n * P.n_words + w, m);
        }
      continue;
    }

    // ---- dynamic slots, tile by tile ------------------------------------------------------
    uint32_t my_word = 0;  // lane (tile & 31) keeps the word of that tile until the row is flushed
    for (int tile = tile_lo; tile < tile_hi; tile++) {
      const int f = tile * 32 + lane;
      const bool valid = f < sc.T;
      const int fc = valid ? f : sc.T - 1;
      bool hit = false;
      if (!static_hit) {
        for (int j0 = j_lo; j0 < j_hi; j0 += 32) {
          const int j = j0 + lane;
          unsigned lm = 0;
          if (j < j_hi) {  // level 0: lanes = slots, swept tile bound vs link bounding spheres
            const float4 tb = sc.tile_bounds[(size_t)tile * sc.n_dynamic + j];
            for (int l = 0; l < L; l++) {
              float dx = W.bc[l][0] - tb.x, dy = W.bc[l][1] - tb.y, dz = W.bc[l][2] - tb.z;
              float rr = W.br[l] + tb.w + band;
              if (dx * dx + dy * dy + dz * dz <= rr * rr) lm |= 1u << l;
            }
          }
          unsigned surv = __ballot_sync(SIRRT_FULL, lm != 0);
          while (surv) {  // level 1: lanes = frames of the tile
            const int src = __ffs(surv) - 1;
            surv &= surv - 1;
            unsigned m = __shfl_sync(SIRRT_FULL, lm, src);
            const int slot = sc.dynamic_slots[j0 + src];
            const float* c = sc.bcw_f + (size_t)slot * 3 * sc.Tp + fc;
            const float cx = c[0], cy = c[sc.Tp], cz = c[2 * (size_t)sc.Tp];
            const DSlot S = sc.slots[slot];
            const float sbr = S.br;
            float tlo[3], thi[3];  // the slot's tight world AABB at this lane's frame (link-independent: loaded once per slot)
            {
              const float* tb = sc.tight_f + (size_t)slot * 6 * sc.Tp + fc;
#pragma unroll
              for (int k = 0; k < 3; k++) { tlo[k] = tb[(size_t)k * sc.Tp]; thi[k] = tb[(size_t)(3 + k) * sc.Tp]; }
            }
            while (m) {
              const int l = __ffs(m) - 1;
              m &= m - 1;
              float dx = W.bc[l][0] - cx, dy = W.bc[l][1] - cy, dz = W.bc[l][2] - cz;
              float rr = (W.br[l] + sbr) * 1.000001f + 1e-5f + band;
              bool pass = valid && (dx * dx + dy * dy + dz * dz <= rr * rr);
#pragma unroll
              for (int k = 0; k < 3; k++) pass &= (W.tlo[l][k] - band <= thi[k]) && (W.thi[l][k] + band >= tlo[k]);
              if (!__any_sync(SIRRT_FULL, pass)) continue;
              hit |= eval_lane(l, slot, S, fc, fc, pass);
              if (HAS_HULL && W.qn > SIRRT_QCAP - 64) { bool any = false; drain(tile * 32, hit, any); }
            }
            if (HAS_HULL && P.tasks.items) { push_tasks_warp(P.tasks, defer0, defer2, (int)n, fc, slot, lane); defer0 = defer2 = 0; }
          }
        }
        bool any = false;
        drain(tile * 32, hit, any);
      }
      uint32_t word = static_hit ? 0xffffffffu : __ballot_sync(SIRRT_FULL, hit && valid);
      if (sc.T - tile * 32 < 32) word &= (1u << (sc.T - tile * 32)) - 1u;
      if ((tile & 31) == lane) my_word = word;
      if ((tile & 31) == 31 || tile == tile_hi - 1) {  // flush up to 32 words, coalesced
        const int w0 = tile & ~31;
        const int w = w0 + lane;
        if (w >= tile_lo && w <= tile) {
          if (n_sch > 1) { if (my_word) atomicOr(P.bits + (size_t)n * P.n_words + w, my_word); }  // rows zeroed by the caller
          else P.bits[(size_t)n * P.n_words + w] = my_word;
        }
        my_word = 0;
      }
    }
    PHASE_T(t_end);
    PHASE_ADD(3, t_end - t_static); PHASE_ADD(4, t_end - t_item); PHASE_MAX(3, t_end - t_static); PHASE_MAX(4, t_end - t_item); PHASE_MAX(2, t_static - t_fk);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(SIRRT_FULL, n_exact, o);
  if (lane == 0 && n_exact) atomicAdd(P.wl.count + 1, n_exact);
}

// ------------------------------------------------------------------------------------------------
// a4: one warp per (configuration, frame) cell
// ------------------------------------------------------------------------------------------------
struct CellParams {
  DScene sc;
  QView q;
  WorkList wl;
  WorkList tasks;       // items != null: hull pair tests are deferred to k_gjk_tasks
  const float* aabb_f;
  double* xf;             // as BitsParams::xf
  const int32_t* frames;  // [N]; null: every cell uses frame_one (latency path, single cell)
  int frame_one;
  uint8_t* out;           // [N]
  int semantics;
  float band;
  // planner-sized batches: the (link, slot) pairs of a cell are split over n_pchunks warps of pairs_per_chunk pairs (a
  // multiple of 32); a warp then only ever WRITES 1 (hit) into flags the caller has zeroed
  int n_pchunks, pairs_per_chunk;
  int obb_cull;           // as BitsParams::obb_cull
};

template <bool HAS_HULL>
__global__ void __launch_bounds__(SIRRT_WARPS_PER_BLOCK * 32)
k_check_cells(const __grid_constant__ DRobot rb, const __grid_constant__ CellParams P) {
  __shared__ WarpCtx ctx[SIRRT_WARPS_PER_BLOCK];
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  WarpCtx& W = ctx[wib];
  const DScene& sc = P.sc;
  const long long warp0 = (long long)blockIdx.x * SIRRT_WARPS_PER_BLOCK + wib;
  const long long n_warps = (long long)gridDim.x * SIRRT_WARPS_PER_BLOCK;
  const bool want_true_rob = (sc.parity_mode == 1) || (P.semantics == 0);
  const bool want_shift_rob = (sc.parity_mode == 0);
  const float band = P.band;
  const int n_pch = P.n_pchunks > 1 ? P.n_pchunks : 1;
  for (long long item = warp0; item < P.q.cnt * n_pch; item += n_warps) {
    const long long n = P.q.n0 + item / n_pch;
    const int pch = (int)(item % n_pch);
    const int f = P.frames ? P.frames[n] : P.frame_one;
    if (f < 0 || f >= sc.T) {  // the reference indexes unchecked (CollisionManager.cpp:369); we answer "collided"
      if (lane == 0 && pch == 0) P.out[n] = 1;
      continue;
    }
    __syncwarp();
    setup_links(rb, sc, P.q, n, W, lane, P.xf);
    const int L = rb.n_links, n_pairs = n_pch > 1 ? min(L * sc.n_slots, (pch + 1) * P.pairs_per_chunk) : L * sc.n_slots;
    bool hit = false;
    for (int p0 = n_pch > 1 ? pch * P.pairs_per_chunk : 0; p0 < n_pairs && !hit; p0 += 32) {
      const int p = p0 + lane;
      bool h = false;
      bool q_true = false, q_shift = false;
      int l = 0, slot = 0;
      if (p < n_pairs) {
        l = p / sc.n_slots; slot = p % sc.n_slots;
        const DSlot S = sc.slots[slot];
        const float* c = sc.bcw_f + (size_t)slot * 3 * sc.Tp + f;
        const float sbr = S.br;
        float dx = W.bc[l][0] - c[0], dy = W.bc[l][1] - c[sc.Tp], dz = W.bc[l][2] - c[2 * (size_t)sc.Tp];
        float rr = (W.br[l] + sbr) * 1.000001f + 1e-5f + band;
        if (dx * dx + dy * dy + dz * dz <= rr * rr && tight_overlap(sc, W, l, slot, f, band)) {
          const DShape& lsh = sc.shapes[rb.link_shape[l]];
          const DShape& sh = sc.shapes[S.shape];
          bool do_true = !S.is_robot || want_true_rob;
          bool do_shift = S.is_robot && want_shift_rob && S.shape_shift >= 0;
          float R[9], t[3];
          if (do_shift) {
            const float* ab = P.aabb_f + (size_t)slot * 6 * sc.Tp + f;
            float blo[3], bhi[3];
            for (int k = 0; k < 3; k++) { blo[k] = ab[(size_t)k * sc.Tp]; bhi[k] = ab[(size_t)(3 + k) * sc.Tp]; }
            int g = aabb_gate(W, l, blo, bhi);
            if (g == RES_FREE) do_shift = false;
            else if (g == RES_AMBIG) { push_work(P.wl, (int)n, f, l, slot, 1); do_shift = false; }
          }
          if (do_true) {
            if (HAS_HULL && (lsh.type == SH_HULL || sh.type == SH_HULL)) {
              q_true = true;
              if (P.obb_cull) {  // oriented boxes clear of each other: no GJK run
                load_pose(sc, slot, f, R, t, true);
                if (obb_separated(W, l, sh, R, t, 1e-5f + band)) q_true = false;
              }
            } else {
              load_pose(sc, slot, f, R, t, sh.type != SH_SPHERE && !S.rot_identity);
              int r = lane_exact(W, l, sh, R, t, S.rot_identity != 0, band);
              if (r == RES_HIT) h = true;
              else if (r == RES_AMBIG) push_work(P.wl, (int)n, f, l, slot, 0);
            }
          }
          if (do_shift) {
            const DShape& ss = sc.shapes[S.shape_shift];
            if (HAS_HULL && (lsh.type == SH_HULL || ss.type == SH_HULL)) {
              q_shift = true;
              if (P.obb_cull) {
                load_pose(sc, slot, f, R, t, true);
                if (obb_separated(W, l, ss, R, t, 1e-5f + band)) q_shift = false;
              }
            } else {
              load_pose(sc, slot, f, R, t, true);
              int r = lane_exact(W, l, ss, R, t, false, band);
              if (r == RES_HIT) h = true;
              else if (r == RES_AMBIG) push_work(P.wl, (int)n, f, l, slot, 2);
            }
          }
        }
      }
      hit = __any_sync(SIRRT_FULL, h);
      if (HAS_HULL && P.tasks.items) {
        push_tasks_warp(P.tasks, q_true ? 1u << l : 0u, q_shift ? 1u << l : 0u, (int)n, f, slot, lane);
      } else if (HAS_HULL) {
        unsigned m1 = __ballot_sync(SIRRT_FULL, q_true), m2 = __ballot_sync(SIRRT_FULL, q_shift);
        for (int pass = 0; pass < 2 && !hit; pass++) {
          unsigned m = pass ? m2 : m1;
          while (m && !hit) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const int pl = __shfl_sync(SIRRT_FULL, l, src), ps = __shfl_sync(SIRRT_FULL, slot, src);
            const DSlot S = sc.slots[ps];
            const DShape& sh = sc.shapes[pass ? S.shape_shift : S.shape];
            int r = coop_test(sc, W.R[pl], W.t[pl], sc.shapes[rb.link_shape[pl]], sh, ps, f, band, lane);
            if (r == RES_HIT) hit = true;
            else if (r == RES_AMBIG && lane == 0) push_work(P.wl, (int)n, f, pl, ps, pass ? 2 : 0);
          }
        }
      }
    }
    if (lane == 0 && (hit || n_pch == 1)) P.out[n] = hit ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------------
// FP64 re-check: one warp per ambiguous pair test; formulas identical to the CPU oracle
// ------------------------------------------------------------------------------------------------
struct IvParams {
  const uint32_t* bits;
  long long N;
  int n_words, T, Kmax;
  int32_t* table;  // [N][1+2*Kmax]
  int16_t* table16 = nullptr;  // compact form of the same rows (frames < 32768); used instead of `table` when set
  uint32_t* rows_out = nullptr;  // latency path of the edge tables: the raw rows are copied here (pinned host block) by finish_call
  int no_fill = 0;             // single-call path: the caller reads only the first `count` pairs, skip the -1 padding
  int clear_rows = 0;          // fused latency path: the rows are an internal buffer that must read zero again for the next call
};

__device__ __forceinline__ int warp_excl_scan(int v, int lane, int& total) {
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int u = __shfl_up_sync(SIRRT_FULL, incl, o);
    if (lane >= o) incl += u;
  }
  total = __shfl_sync(SIRRT_FULL, incl, 31);
  return incl - v;
}

// one warp: collision row n -> its line of the interval table.  Rows are read past L1 (__ldcg): on the fused latency path
// the bits were OR-ed in by other blocks of the same launch.
template <typename OutT> __device__ inline void row_to_intervals_t(const IvParams& P, OutT* out, long long n, int lane) {
  const uint32_t* row = P.bits + (size_t)n * P.n_words;
  int n_starts = 0, n_ends = 0;
  uint32_t prev_safe = 0;  // safe bit of the last frame of the previous group (frame -1: "collision")
  for (int w0 = 0; w0 < P.n_words; w0 += 32) {
    const int w = w0 + lane;
    uint32_t s = 0;
    if (w < P.n_words) {
      const int left = P.T - w * 32;
      s = ~__ldcg(row + w) & (left >= 32 ? 0xffffffffu : ((1u << left) - 1u));
    }
    uint32_t prev_msb = __shfl_up_sync(SIRRT_FULL, s >> 31, 1);
    if (lane == 0) prev_msb = prev_safe;
    uint32_t next_lsb = __shfl_down_sync(SIRRT_FULL, s & 1u, 1);
    if (lane == 31) {
      next_lsb = 0;
      if (w + 1 < P.n_words) next_lsb = (~__ldcg(row + w + 1)) & 1u;  // frame (w+1)*32 < T, so the bit is valid
    }
    uint32_t starts = s & ~((s << 1) | prev_msb);
    uint32_t ends = s & ~((s >> 1) | (next_lsb << 31));
    int ts, te;
    int rs = n_starts + warp_excl_scan(__popc(starts), lane, ts);
    int re = n_ends + warp_excl_scan(__popc(ends), lane, te);
    while (starts) { const int b = __ffs(starts) - 1; starts &= starts - 1; if (rs < P.Kmax) out[1 + 2 * rs] = (OutT)(w * 32 + b); rs++; }
    while (ends) { const int b = __ffs(ends) - 1; ends &= ends - 1; if (re < P.Kmax) out[2 + 2 * re] = (OutT)(w * 32 + b); re++; }
    n_starts += ts; n_ends += te;
    prev_safe = __shfl_sync(SIRRT_FULL, s >> 31, 31);
  }
  // unused pairs = -1; the count is the true number of intervals even when it exceeds Kmax
  if (!P.no_fill)
    for (int k = 2 * min(n_starts, P.Kmax) + lane; k < 2 * P.Kmax; k += 32) out[1 + k] = (OutT)-1;
  if (lane == 0) out[0] = (OutT)n_starts;
}
__device__ inline void row_to_intervals(const IvParams& P, long long n, int lane) {
  if (P.table16) row_to_intervals_t<int16_t>(P, P.table16 + (size_t)n * (1 + 2 * P.Kmax), n, lane);
  else row_to_intervals_t<int32_t>(P, P.table + (size_t)n * (1 + 2 * P.Kmax), n, lane);
}

__global__ void __launch_bounds__(256) k_bits_to_intervals(const __grid_constant__ IvParams P) {
  const int lane = threadIdx.x & 31;
  const long long n = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (n >= P.N) return;
  row_to_intervals(P, n, lane);
}

struct RecheckParams {
  DScene sc;
  QView q;
  const WorkItem* items;
  const unsigned int* count;  // [0] entries pushed by the producer kernel (clamped to cap here)
  unsigned int cap;
  const double* aabb_d;     // robot slots: [(slot*Tp + f)*6 + k]
  uint32_t* bits;           // bits mode: [N][W]
  int n_words;
  uint8_t* cells;           // cell mode: [N]
  unsigned int* counters;   // [2] near-contact count
  // latency path: the last block to finish publishes counters[0..3] to `publish` (pinned host block) and clears them
  // and the ticket, so that a call needs neither a memset before nor a copy after its kernels
  unsigned int* ticket;
  unsigned int* publish;
  // completion flag in the pinned host block: the last block writes done_seq behind everything the call produced, so the
  // host can wait by polling its own memory instead of calling into the driver (sirrt_wait_flag)
  unsigned int* done;
  unsigned int done_seq;
  IvParams iv;              // latency path with a small table: iv.table != null -> the last block also turns the rows into intervals
  // early completion: when the task kernel in front of this one found the FP64 list empty its last block has already done
  // everything below (finish_call) and left fin_seq in *fin -- this kernel then returns at once
  unsigned int* fin;
  unsigned int fin_seq;
  const int4* seeds;        // per work item: the FP32 simplex to start the FP64 GJK from (WorkList::seeds), or null
  unsigned int seed_cap;
};

// End of a latency-path call, run by ONE block (all its threads) once every result of the call is visible to it: rows ->
// intervals (fused table), launch counters published to the pinned block and cleared for the next call, completion flag.
__device__ inline void finish_call(const IvParams& iv, unsigned int* counters, unsigned int* ticket, unsigned int* publish, unsigned int* done,
                                   unsigned int done_seq, unsigned int* fin, unsigned int fin_seq) {
  const int lane = threadIdx.x & 31;
  if (iv.rows_out) {  // covered, like the table, by thread 0's system fence behind the barrier below
    for (long long i = threadIdx.x; i < iv.N * iv.n_words; i += blockDim.x) iv.rows_out[i] = __ldcg(iv.bits + i);
    if (!iv.table) __syncthreads();
  }
  if (iv.table) {
    for (long long n = threadIdx.x >> 5; n < iv.N; n += SIRRT_WARPS_PER_BLOCK) row_to_intervals(iv, n, lane);
    __syncthreads();  // thread 0's system fence below is cumulative over what the block wrote before this barrier
    if (iv.clear_rows)
      for (long long i = threadIdx.x; i < iv.N * iv.n_words; i += blockDim.x) ((uint32_t*)iv.bits)[i] = 0u;
  }
  if (threadIdx.x == 0) {
    volatile unsigned int* c = counters;
    for (int i = 0; i < 4; i++) { const unsigned int v = c[i]; if (publish) publish[i] = v; c[i] = 0u; }
    if (ticket) *ticket = 0u;
    if (fin) *fin = fin_seq;
    if (done) { __threadfence_system(); *(volatile unsigned int*)done = done_seq; }
  }
}

__device__ inline double recheck_pair(const DRobot& rb, const DScene& sc, const LinkXfD& X, int l, const DShape& sh,
                                      const double* pose, int lane, int* res, const int* enc_ids = nullptr, bool* near = nullptr,
                                      const int* seed_ids = nullptr) {
  const DShape& lsh = sc.shapes[rb.link_shape[l]];
  const double *RB = pose, *tB = pose + 9;
  double d;
  if (lsh.type == SH_CAPSULE && sh.type != SH_HULL) {
    double a[3], b[3];
    for (int k = 0; k < 3; k++) { a[k] = X.t[k] - X.R[3 * k + 2] * lsh.d_half_len; b[k] = X.t[k] + X.R[3 * k + 2] * lsh.d_half_len; }
    // the oracle builds the end points as R*(0,0,+-hl)+t; identical up to the sign of zero products
    if (sh.type == SH_SPHERE) d = sqrt(pt_seg_dist2(tB, a, b)) - lsh.d_radius - sh.d_radius;
    else if (sh.type == SH_BOX) {
      double da[3], db[3], la[3], lb[3];
      sub3(da, a, tB); sub3(db, b, tB);
      matTvec(la, RB, da); matTvec(lb, RB, db);
      d = sqrt(seg_box_dist2(la, lb, sh.d_half)) - lsh.d_radius;
    } else {
      double p2[3], q2[3];
      for (int k = 0; k < 3; k++) { p2[k] = tB[k] - RB[3 * k + 2] * sh.d_half_len; q2[k] = tB[k] + RB[3 * k + 2] * sh.d_half_len; }
      d = sqrt(seg_seg_dist2(a, b, p2, q2)) - lsh.d_radius - sh.d_radius;
    }
    *res = (d - SIRRT_SECURITY_MARGIN <= SIRRT_COLLISION_THRESHOLD) ? RES_HIT : RES_FREE;
    if (near) *near = fabs(d) < 1e-6;
    return d;
  }
  GShape<double> A, B;
  A.type = lsh.type; A.verts = sc.verts_d + 3 * (size_t)lsh.vert_off; A.n = lsh.n_verts; A.cells = (sc.hull_cells && lsh.cell_off >= 0) ? sc.hull_cells + lsh.cell_off : nullptr; A.cand = sc.hull_cand; A.half_len = lsh.d_half_len;
  A.half[0] = lsh.d_half[0]; A.half[1] = lsh.d_half[1]; A.half[2] = lsh.d_half[2];
  B.type = sh.type; B.verts = sc.verts_d + 3 * (size_t)sh.vert_off; B.n = sh.n_verts; B.cells = (sc.hull_cells && sh.cell_off >= 0) ? sc.hull_cells + sh.cell_off : nullptr; B.cand = sc.hull_cand; B.half_len = sh.d_half_len;
  B.half[0] = sh.d_half[0]; B.half[1] = sh.d_half[1]; B.half[2] = sh.d_half[2];
  double Rrel[9], trel[3], dt[3];
  matTmul3(Rrel, X.R, RB);
  sub3(dt, tB, X.t);
  matTvec(trel, X.R, dt);
  double rsum = (lsh.type == SH_HULL ? 0.0 : lsh.d_radius) + ((sh.type == SH_SPHERE || sh.type == SH_CAPSULE) ? sh.d_radius : 0.0);
  if (enc_ids && enc_ids[0] >= 0) {
    // the FP32 pass ended with an enclosing but thin tetrahedron: certify the same four points in FP64 instead of
    // running GJK again (an intersection of the cores is a collision whatever the radii: distance = -rsum <= threshold)
    double mf;
    if (tetra_encloses(A, B, Rrel, trel, enc_ids, 1e-9, &mf)) {
      *res = RES_HIT;
      if (near) *near = rsum < 1e-6 && mf < 1e-6;
      return -rsum;
    }
  }
  *res = warp_gjk<double, true>(A, B, Rrel, trel, rsum, 0.0, lane, &d, nullptr, nullptr, nullptr, nullptr, nullptr, seed_ids);
  if (near) *near = fabs(d) < 1e-6;
  return d;
}

// a9: signed distance of (link l, shape sh at pose) WITH the closest point pair in world coordinates
// (coal::DistanceResult::nearest_points as the reference's DistanceCallback collects them, CollisionManager.cpp:549-600):
// pl on the planned-robot link's surface, po on the obstacle's surface.  Closed forms give the feet directly; hull pairs
// take them from the barycentric coordinates of GJK's final simplex.  For a penetrating pair the points are the cores'
// closest points (meaningless, as in coal without EPA; the reference returns before reading them).
__device__ inline double pair_distance_witness(const DRobot& rb, const DScene& sc, const LinkXfD& X, int l, const DShape& sh,
                                               const double* pose, int lane, double* pl, double* po) {
  const DShape& lsh = sc.shapes[rb.link_shape[l]];
  const double *RB = pose, *tB = pose + 9;
  double ca[3], cb[3];  // closest points of the two CORES, world frame
  double ra = lsh.type == SH_HULL ? 0.0 : lsh.d_radius, rbb = (sh.type == SH_SPHERE || sh.type == SH_CAPSULE) ? sh.d_radius : 0.0;
  if (lsh.type == SH_CAPSULE && sh.type != SH_HULL) {
    double a[3], b[3];
    for (int k = 0; k < 3; k++) { a[k] = X.t[k] - X.R[3 * k + 2] * lsh.d_half_len; b[k] = X.t[k] + X.R[3 * k + 2] * lsh.d_half_len; }
    if (sh.type == SH_SPHERE) {
      double ab[3], ap[3];
      sub3(ab, b, a); sub3(ap, tB, a);
      const double den = dot3(ab, ab), t = den > 0.0 ? clamp01(dot3(ap, ab) / den) : 0.0;
      for (int k = 0; k < 3; k++) { ca[k] = a[k] + t * ab[k]; cb[k] = tB[k]; }
    } else if (sh.type == SH_BOX) {
      double da[3], db[3], la[3], lb[3], t;
      sub3(da, a, tB); sub3(db, b, tB);
      matTvec(la, RB, da); matTvec(lb, RB, db);
      seg_box_dist2(la, lb, sh.d_half, &t);
      double ps[3], pb[3], w[3];
      for (int k = 0; k < 3; k++) {
        ps[k] = la[k] + t * (lb[k] - la[k]);
        pb[k] = ps[k] < -sh.d_half[k] ? -sh.d_half[k] : (ps[k] > sh.d_half[k] ? sh.d_half[k] : ps[k]);
        ca[k] = a[k] + t * (b[k] - a[k]);
      }
      matvec(w, RB, pb);
      for (int k = 0; k < 3; k++) cb[k] = w[k] + tB[k];
    } else {
      double p2[3], q2[3], s, t;
      for (int k = 0; k < 3; k++) { p2[k] = tB[k] - RB[3 * k + 2] * sh.d_half_len; q2[k] = tB[k] + RB[3 * k + 2] * sh.d_half_len; }
      seg_seg_params(a, b, p2, q2, s, t);
      for (int k = 0; k < 3; k++) { ca[k] = a[k] + s * (b[k] - a[k]); cb[k] = p2[k] + t * (q2[k] - p2[k]); }
    }
  } else {
    GShape<double> A, B;
    A.type = lsh.type; A.verts = sc.verts_d + 3 * (size_t)lsh.vert_off; A.n = lsh.n_verts; A.cells = (sc.hull_cells && lsh.cell_off >= 0) ? sc.hull_cells + lsh.cell_off : nullptr; A.cand = sc.hull_cand; A.half_len = lsh.d_half_len;
    A.half[0] = lsh.d_half[0]; A.half[1] = lsh.d_half[1]; A.half[2] = lsh.d_half[2];
    B.type = sh.type; B.verts = sc.verts_d + 3 * (size_t)sh.vert_off; B.n = sh.n_verts; B.cells = (sc.hull_cells && sh.cell_off >= 0) ? sc.hull_cells + sh.cell_off : nullptr; B.cand = sc.hull_cand; B.half_len = sh.d_half_len;
    B.half[0] = sh.d_half[0]; B.half[1] = sh.d_half[1]; B.half[2] = sh.d_half[2];
    double Rrel[9], trel[3], dt[3], d, v[3] = {0, 0, 0};
    matTmul3(Rrel, X.R, RB);
    sub3(dt, tB, X.t);
    matTvec(trel, X.R, dt);
    Simplex<double> S;
    S.n = 0;
    warp_gjk<double, true>(A, B, Rrel, trel, ra + rbb, 0.0, lane, &d, nullptr, &S, v);
    // barycentric coordinates of v (the point of the final simplex closest to the origin) by least squares on its vertices
    double lam[4] = {1, 0, 0, 0};
    if (S.n == 2) {
      double e[3], r[3];
      sub3(e, S.w[1], S.w[0]); sub3(r, v, S.w[0]);
      const double den = dot3(e, e);
      lam[1] = den > 0.0 ? clamp01(dot3(r, e) / den) : 0.0; lam[0] = 1.0 - lam[1];
    } else if (S.n >= 3) {
      double e1[3], e2[3], r[3];
      sub3(e1, S.w[1], S.w[0]); sub3(e2, S.w[2], S.w[0]); sub3(r, v, S.w[0]);
      const double a11 = dot3(e1, e1), a12 = dot3(e1, e2), a22 = dot3(e2, e2), b1 = dot3(r, e1), b2 = dot3(r, e2), det = a11 * a22 - a12 * a12;
      if (det > 0.0) { lam[1] = (b1 * a22 - b2 * a12) / det; lam[2] = (b2 * a11 - b1 * a12) / det; }
      lam[0] = 1.0 - lam[1] - lam[2];
    }
    double pa[3] = {0, 0, 0}, pb[3] = {0, 0, 0};
    const int ns = S.n < 1 ? 0 : (S.n > 3 ? 3 : S.n);
    for (int i = 0; i < ns; i++) {
      double xa[3], xb[3], w[3];
      gjk_point(A, S.id[i] >> 12, xa);
      gjk_point(B, S.id[i] & 4095, xb);
      matvec(w, Rrel, xb);
      for (int k = 0; k < 3; k++) { pa[k] += lam[i] * xa[k]; pb[k] += lam[i] * (w[k] + trel[k]); }
    }
    double wa[3], wb[3];
    matvec(wa, X.R, pa); matvec(wb, X.R, pb);
    for (int k = 0; k < 3; k++) { ca[k] = wa[k] + X.t[k]; cb[k] = wb[k] + X.t[k]; }
  }
  double n[3];
  sub3(n, cb, ca);
  const double len = sqrt(dot3(n, n));
  if (len > 0.0) { n[0] /= len; n[1] /= len; n[2] /= len; } else { n[0] = n[1] = n[2] = 0.0; }
  for (int k = 0; k < 3; k++) { pl[k] = ca[k] + ra * n[k]; po[k] = cb[k] - rbb * n[k]; }
  return len - ra - rbb;
}

// FP64 certificate of an FP32 "enclosed, but thin" GJK verdict (see tetra_encloses); kept out of line so that the task
// kernel's register budget stays that of the FP32 GJK
__device__ __noinline__ bool certify_enclosure(const DRobot& rb, const DScene& sc, const LinkXfD& X, int l, const DShape& sh,
                                               const double* pose, const int* enc_ids, bool* near) {
  const DShape& lsh = sc.shapes[rb.link_shape[l]];
  GShape<double> A, B;
  A.type = lsh.type; A.verts = sc.verts_d + 3 * (size_t)lsh.vert_off; A.n = lsh.n_verts; A.cells = nullptr; A.cand = nullptr; A.half_len = lsh.d_half_len;
  A.half[0] = lsh.d_half[0]; A.half[1] = lsh.d_half[1]; A.half[2] = lsh.d_half[2];
  B.type = sh.type; B.verts = sc.verts_d + 3 * (size_t)sh.vert_off; B.n = sh.n_verts; B.cells = nullptr; B.cand = nullptr; B.half_len = sh.d_half_len;
  B.half[0] = sh.d_half[0]; B.half[1] = sh.d_half[1]; B.half[2] = sh.d_half[2];
  double Rrel[9], trel[3], dt[3], mf;
  matTmul3(Rrel, X.R, pose);
  sub3(dt, pose + 9, X.t);
  matTvec(trel, X.R, dt);
  if (!tetra_encloses(A, B, Rrel, trel, enc_ids, 1e-9, &mf)) return false;
  const double rsum = (lsh.type == SH_HULL ? 0.0 : lsh.d_radius) + ((sh.type == SH_SPHERE || sh.type == SH_CAPSULE) ? sh.d_radius : 0.0);
  *near = rsum < 1e-6 && mf < 1e-6;
  return true;
}

// FK of one link only (what a re-check item needs): the chain is walked up to link `want` and no further
__device__ inline void fk_one_link(const DRobot& rb, const double* q, int want, LinkXfD& out) {
  double t[3] = {rb.base_t[0], rb.base_t[1], rb.base_t[2]}, R[9];
#pragma unroll
  for (int i = 0; i < 9; i++) R[i] = rb.base_R[i];
  int l = 0;
  for (int k = 0; k <= rb.n_joints; k++) {
    while (l < rb.n_links && rb.link_jb[l] == k) {
      if (l == want) {
        double w[3];
        matvec(w, R, rb.lo_t[l]);  // :335
        out.t[0] = t[0] + w[0]; out.t[1] = t[1] + w[1]; out.t[2] = t[2] + w[2];  // :336
        matmul3(out.R, R, rb.lo_R[l]);  // :347
        return;
      }
      l++;
    }
    if (k == rb.n_joints) break;
    double w[3];
    matvec(w, R, rb.jt[k]);  // :319,361
    t[0] += w[0]; t[1] += w[1]; t[2] += w[2];
    matmul3(R, R, rb.jR[k]);  // :320,362
    double sn, cs;
    sincos(q[k], &sn, &cs);
    const double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};  // :321,363
    matmul3(R, R, Rz);
  }
}

// QUIRK 7 broadphase gate of a `shifted == 1` item in FP64: do the coal AABBs of the true geometries overlap?
__device__ inline bool recheck_gate(const DScene& sc, const DShape& lsh, const LinkXfD& X, const double* ab) {
  double alo[3], ahi[3];
  bool ident = true;
  for (int k = 0; k < 9; k++) ident &= fabs(X.R[k] - ((k % 4) == 0 ? 1.0 : 0.0)) <= 1e-12;
  if (ident) { for (int k = 0; k < 3; k++) { alo[k] = lsh.d_amin[k] + X.t[k]; ahi[k] = lsh.d_amax[k] + X.t[k]; } }
  else {
    double c[3]; matvec(c, X.R, lsh.d_ac);
    for (int k = 0; k < 3; k++) { alo[k] = c[k] + X.t[k] - lsh.d_ar; ahi[k] = c[k] + X.t[k] + lsh.d_ar; }
  }
  bool overlap = true;
  for (int k = 0; k < 3; k++) if (alo[k] > ab[3 + k] || ahi[k] < ab[k]) overlap = false;
  return overlap;
}

// publish one decided HIT of the re-check: the bit of (config, frame), the whole row for a static slot, or the cell flag
__device__ __forceinline__ void recheck_set(const RecheckParams& P, const WorkItem& it) {
  if (P.cells) { P.cells[it.config] = 1; return; }
  uint32_t* row = P.bits + (size_t)it.config * P.n_words;
  if (it.frame >= 0) { atomicOr(row + (it.frame >> 5), 1u << (it.frame & 31)); return; }
  for (int wd = 0; wd < P.n_words; wd++) {
    uint32_t m = 0xffffffffu;
    if (P.sc.T - wd * 32 < 32) m = (1u << (P.sc.T - wd * 32)) - 1u;
    atomicOr(row + wd, m);
  }
}

// closed-form item (capsule link vs box / sphere / capsule): is this item one?  (everything else needs the warp GJK)
__device__ __forceinline__ bool recheck_is_closed(const DRobot& rb, const DScene& sc, const WorkItem& it) {
  const DSlot S = sc.slots[it.slot];
  const int lt = sc.shapes[rb.link_shape[it.link]].type;
  if (lt != SH_CAPSULE) return false;
  if (it.shifted == 1) return S.shape_shift >= 0 && sc.shapes[S.shape_shift].type != SH_HULL;
  return sc.shapes[it.shifted ? S.shape_shift : S.shape].type != SH_HULL;
}

// Two passes over the work list.  Pass 1: one LANE per closed-form item (the synthetic sweep leaves ~4e5 of them per
// 2^20 configurations; a warp per item idled 31 lanes through an FP64 FK of the whole chain).  Pass 2: one WARP per
// item that needs the warp-cooperative FP64 GJK (hull links / hull obstacles), spread over the machine because a
// planner-sized call has a handful of them and each is a long dependent chain.
// the two passes for warps warp0, warp0 + n_warps, ... (the whole grid of k_recheck, or the 8 warps of the task kernel's
// last block when a planner-sized call left only a few items: GjkTaskParams::inline_recheck)
__device__ __noinline__ void recheck_items(const DRobot& rb, const RecheckParams& P, unsigned int n_items, unsigned int warp0, unsigned int n_warps) {
  const int lane = threadIdx.x & 31;
  const DScene& sc = P.sc;
  unsigned int n_near = 0;
  for (unsigned int base = warp0 * 32; base < n_items; base += n_warps * 32) {
    const unsigned int w = base + lane;
    if (w >= n_items) continue;
    const WorkItem it = P.items[w];
    if (!recheck_is_closed(rb, sc, it)) continue;
    double q[SIRRT_MAXJ];
    for (int j = 0; j < rb.n_joints; j++) q[j] = j < P.q.dof ? load_q(P.q, it.config, j) : 0.0;
    LinkXfD X;
    fk_one_link(rb, q, it.link, X);
    const int l = it.link, f = it.frame < 0 ? 0 : it.frame;
    const DSlot S = sc.slots[it.slot];
    const double* pose = sc.pose_d + ((size_t)it.slot * sc.Tp + f) * 12;
    int shifted = it.shifted;
    if (shifted == 1) {
      if (!recheck_gate(sc, sc.shapes[rb.link_shape[l]], X, P.aabb_d + ((size_t)it.slot * sc.Tp + f) * 6)) continue;
      shifted = 2;
    }
    int res;
    const double d = recheck_pair(rb, sc, X, l, sc.shapes[shifted ? S.shape_shift : S.shape], pose, 0, &res);  // closed form: no warp primitive inside
    if (fabs(d) < 1e-6) n_near++;
    if (res == RES_HIT) recheck_set(P, it);
  }
  if (__any_sync(SIRRT_FULL, n_near != 0)) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) n_near += __shfl_xor_sync(SIRRT_FULL, n_near, o);
    if (lane == 0) atomicAdd(P.counters + 2, n_near);
  }
  if (sc.has_hull)  // a scene without a hull (links, obstacles, robot-obstacle links) has closed-form items only
  for (unsigned int w = warp0; w < n_items; w += n_warps) {
  const WorkItem it = P.items[w];
  if (recheck_is_closed(rb, sc, it)) continue;
  double q[SIRRT_MAXJ];
  for (int j = 0; j < rb.n_joints; j++) q[j] = j < P.q.dof ? load_q(P.q, it.config, j) : 0.0;
  LinkXfD X;
  fk_one_link(rb, q, it.link, X);
  const int l = it.link, f = it.frame < 0 ? 0 : it.frame;
  const DSlot S = sc.slots[it.slot];
  const double* pose = sc.pose_d + ((size_t)it.slot * sc.Tp + f) * 12;
  int shifted = it.shifted;
  if (shifted == 1) {  // undecided broadphase gate: coal AABB overlap of the true geometries, FP64
    if (!recheck_gate(sc, sc.shapes[rb.link_shape[l]], X, P.aabb_d + ((size_t)it.slot * sc.Tp + f) * 6)) continue;
    shifted = 2;
  }
  const DShape& sh = sc.shapes[shifted ? S.shape_shift : S.shape];
  int res;
  int seed[4] = {-1, -1, -1, -1};
  if (P.seeds && w < P.seed_cap && shifted == it.shifted) {  // (a gate item re-checks another geometry than the one its slot was pushed for: no seed)
    const int4 sd = P.seeds[w];
    seed[0] = sd.x; seed[1] = sd.y; seed[2] = sd.z; seed[3] = sd.w;
  }
  double d = recheck_pair(rb, sc, X, l, sh, pose, lane, &res, nullptr, nullptr, seed);
  if (lane == 0 && fabs(d) < 1e-6) atomicAdd(P.counters + 2, 1u);
  if (res != RES_HIT) continue;
  if (P.cells) { if (lane == 0) P.cells[it.config] = 1; continue; }
  uint32_t* row = P.bits + (size_t)it.config * P.n_words;
  if (it.frame >= 0) { if (lane == 0) atomicOr(row + (it.frame >> 5), 1u << (it.frame & 31)); }
  else {
    for (int wd = lane; wd < P.n_words; wd += 32) {
      uint32_t m = 0xffffffffu;
      if (sc.T - wd * 32 < 32) m = (1u << (sc.T - wd * 32)) - 1u;
      atomicOr(row + wd, m);
    }
  }
  }
}

__global__ void __launch_bounds__(SIRRT_WARPS_PER_BLOCK * 32)
k_recheck(const __grid_constant__ DRobot rb, const __grid_constant__ RecheckParams P) {
  pdl_wait();
  if (P.fin && *(volatile unsigned int*)P.fin == P.fin_seq) return;  // the task kernel's last block has completed the call
  const unsigned int n_items = min(*P.count, P.cap);
  if (P.ticket && n_items == 0) {  // nothing to re-check (the common planner call): one block completes the call at once
    if (blockIdx.x == 0) finish_call(P.iv, P.counters, nullptr, P.publish, P.done, P.done_seq, P.fin, P.fin_seq);
    return;
  }
  recheck_items(rb, P, n_items, blockIdx.x * SIRRT_WARPS_PER_BLOCK + (threadIdx.x >> 5), gridDim.x * SIRRT_WARPS_PER_BLOCK);
  if (P.ticket) {  // latency path: last block out (fuses the interval reduction,) publishes and clears the launch counters
    __shared__ bool last;
    // this block's results: bits / counters in device memory; in cell mode on the latency path the flags are the pinned host block
    if (P.done && P.cells) __threadfence_system(); else __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(P.ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    finish_call(P.iv, P.counters, P.ticket, P.publish, P.done, P.done_seq, P.fin, P.fin_seq);
  }
}

// ------------------------------------------------------------------------------------------------
// Deferred hull pair tests (planner-in-the-loop batches: one configuration, or the few dozen states of an
// edge).  The producing kernels only cull and append (configuration, frame, link, slot) tasks; here one warp
// per task runs the FP32 warp-cooperative GJK, so a single get_safe_intervals(q) on a hull robot spreads its
// few hundred GJK runs over all SMs instead of serialising them in the <= W warps that own the row.
// Decisions are the same as in the producing warp: HIT sets the bit / cell, AMBIGUOUS goes to the FP64 list.
// ------------------------------------------------------------------------------------------------
struct GjkTaskParams {
  DScene sc;
  QView q;
  const WorkItem* items;
  const unsigned int* count;  // tasks pushed (clamped to cap here; an overflow is detected by the host)
  unsigned int cap;
  WorkList wl;                // FP64 re-check list
  const double* xf;           // FP64 link transforms [cnt][n_links][12] left by the producing kernel, or null (FK is redone here)
  uint32_t* bits;             // bits mode: [N][W]
  int n_words;
  uint8_t* cells;             // cell mode: [N]
  float band;
  // latency path (ticket != null): when the FP64 re-check list is still empty after the last task, the last block out
  // completes the call here (finish_call) and k_recheck, already queued behind this kernel, returns at once
  unsigned int* ticket;       // this kernel's own ticket word
  unsigned int* publish;
  unsigned int* done;
  unsigned int done_seq;
  unsigned int* fin;
  unsigned int fin_seq;
  IvParams iv;
  // inline_recheck: k_recheck has NOT been queued behind this kernel (one launch fewer per call: the mass test is bound by
  // the GPU's kernel-launch rate).  The last block re-checks a handful of items itself before it completes the
  // call; with more it raises SIRRT_FLAG_CONTINUE on the completion flag and the host launches k_recheck after all.
  int inline_recheck;         // 0: off; n: up to n - 1 items are re-checked by the last block
  const double* aabb_d;
};
#define SIRRT_FLAG_CONTINUE 0x80000000u

__global__ void __launch_bounds__(SIRRT_WARPS_PER_BLOCK * 32)
k_gjk_tasks(const __grid_constant__ DRobot rb, const __grid_constant__ GjkTaskParams P) {
  const int lane = threadIdx.x & 31;
  const DScene& sc = P.sc;
  pdl_launch_dependents();
  pdl_wait();
  const unsigned int n_items = min(*P.count, P.cap);
  if (P.ticket && n_items == 0 && (!P.inline_recheck || *(volatile unsigned int*)P.wl.count == 0u)) {  // no deferred pair test at all
    if (blockIdx.x == 0 && *(volatile unsigned int*)P.wl.count == 0u)  // ... and nothing for the FP64 re-check either
      finish_call(P.iv, P.wl.count, nullptr, P.publish, P.done, P.done_seq, P.fin, P.fin_seq);
    return;
  }
  // blocks without a task leave at once and take no ticket: under load (16+ planner threads per GPU) every resident block
  // of this grid is an SM slot another call is waiting for
  const unsigned int n_active = max(1u, min(gridDim.x, (n_items + SIRRT_WARPS_PER_BLOCK - 1) / SIRRT_WARPS_PER_BLOCK));
  if (P.ticket && blockIdx.x >= n_active) return;
  for (unsigned int w = blockIdx.x * SIRRT_WARPS_PER_BLOCK + (threadIdx.x >> 5); w < n_items; w += gridDim.x * SIRRT_WARPS_PER_BLOCK) {
    PHASE_T(t_task);
    const WorkItem it = P.items[w];
    const int l = it.link, f = it.frame < 0 ? 0 : it.frame;
    LinkXfD Xl;
    if (P.xf) {
      const double* x = P.xf + ((size_t)(it.config - P.q.n0) * rb.n_links + l) * 12;
#pragma unroll
      for (int k = 0; k < 9; k++) Xl.R[k] = x[k];
#pragma unroll
      for (int k = 0; k < 3; k++) Xl.t[k] = x[9 + k];
    } else {
      double q[SIRRT_MAXJ];
      for (int j = 0; j < rb.n_joints; j++) q[j] = j < P.q.dof ? load_q(P.q, it.config, j) : 0.0;
      fk_one_link(rb, q, l, Xl);
    }
    float RA[9], tA[3];
#pragma unroll
    for (int k = 0; k < 9; k++) RA[k] = (float)Xl.R[k];  // the same narrowing as setup_links
#pragma unroll
    for (int k = 0; k < 3; k++) tA[k] = (float)Xl.t[k];
    const DSlot S = sc.slots[it.slot];
    const DShape& sh = sc.shapes[it.shifted ? S.shape_shift : S.shape];
    int enc_ids[4], last_ids[4] = {-1, -1, -1, -1};
#ifdef SIRRT_GJK_STATS
    int n_iter = 0;
    PHASE_T(t_pro);
    int r = coop_test(sc, RA, tA, sc.shapes[rb.link_shape[l]], sh, it.slot, f, P.band, lane, enc_ids, &n_iter, last_ids);
    if (lane == 0) atomicAdd(&g_gjk_hist[r == RES_AMBIG ? (enc_ids[0] >= 0 ? 3 : 2) : r][n_iter < 63 ? n_iter : 63], 1u);
    PHASE_T(t_gjk);
    PHASE_ADD(8, 1); PHASE_ADD(9, t_pro - t_task); PHASE_ADD(10, t_gjk - t_pro); PHASE_ADD(12, n_iter); PHASE_MAX(10, t_gjk - t_pro); PHASE_MAX(12, n_iter);
#else
    int r = coop_test(sc, RA, tA, sc.shapes[rb.link_shape[l]], sh, it.slot, f, P.band, lane, enc_ids, nullptr, last_ids);
#endif
    if (r == RES_AMBIG) {
      // a thin enclosing tetrahedron is certified here in FP64 by the warp that found it (it holds the FP64 link
      // transform); anything else goes to the FP64 re-check list
      bool near = false;
      if (enc_ids[0] >= 0 && certify_enclosure(rb, sc, Xl, l, sh, sc.pose_d + ((size_t)it.slot * sc.Tp + f) * 12, enc_ids, &near)) {
        r = RES_HIT;
        if (lane == 0 && near) atomicAdd(P.wl.count + 2, 1u);
      } else {
        if (lane == 0) push_work(P.wl, it.config, it.frame, l, it.slot, it.shifted, last_ids);
        continue;
      }
    }
    if (r != RES_HIT) continue;
    if (P.cells) { if (lane == 0) P.cells[it.config] = 1; continue; }
    uint32_t* row = P.bits + (size_t)it.config * P.n_words;
    if (it.frame >= 0) { if (lane == 0) atomicOr(row + (it.frame >> 5), 1u << (it.frame & 31)); }
    else {
      for (int wd = lane; wd < P.n_words; wd += 32) {
        uint32_t m = 0xffffffffu;
        if (sc.T - wd * 32 < 32) m = (1u << (sc.T - wd * 32)) - 1u;
        atomicOr(row + wd, m);
      }
    }
  }
  if (P.ticket) {
    __shared__ bool last;
    if (P.done && P.cells) __threadfence_system(); else __threadfence();  // cell flags of the latency path are the pinned host block
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(P.ticket, 1u) == n_active - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    if (threadIdx.x == 0) *P.ticket = 0u;
    const unsigned int n_fp64 = *(volatile unsigned int*)P.wl.count;
    if (n_fp64 == 0u) finish_call(P.iv, P.wl.count, nullptr, P.publish, P.done, P.done_seq, P.fin, P.fin_seq);
    else if (P.inline_recheck) {
      if (n_fp64 < (unsigned int)P.inline_recheck && n_fp64 <= P.wl.cap) {  // a handful: this block's 8 warps are the re-check
        RecheckParams R{};
        R.sc = P.sc; R.q = P.q; R.items = P.wl.items; R.count = P.wl.count; R.cap = P.wl.cap; R.aabb_d = P.aabb_d;
        R.bits = P.bits; R.n_words = P.n_words; R.cells = P.cells; R.counters = P.wl.count; R.done = P.done;
        R.seeds = P.wl.seeds; R.seed_cap = P.wl.seed_cap;
        recheck_items(rb, R, n_fp64, threadIdx.x >> 5, SIRRT_WARPS_PER_BLOCK);
        if (P.done && P.cells) __threadfence_system(); else __threadfence();
        __syncthreads();
        finish_call(P.iv, P.wl.count, nullptr, P.publish, P.done, P.done_seq, P.fin, P.fin_seq);
      } else if (threadIdx.x == 0 && P.done) {  // too many for one block: the host queues k_recheck (which completes the call)
        __threadfence_system();
        *(volatile unsigned int*)P.done = P.done_seq | SIRRT_FLAG_CONTINUE;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// a3 tail: collision bit rows -> safe intervals (CollisionManager.cpp:507-546), one warp per row.
// safe = ~collision.  An interval starts at a safe frame whose predecessor collides (or is frame
// -1) and ends at a safe frame whose successor collides (or is frame T).  Starts and ends are both
// ascending, so the k-th start pairs with the k-th end; each is ranked with a warp prefix sum of
// popcounts on top of a running count carried across groups of 32 words.
// ------------------------------------------------------------------------------------------------
