// sirrt_sweep.cuh -- k_sweep: the throughput kernel of the batched safe-interval path for scenes
// whose planned-robot links are all capsules and whose obstacles are boxes / spheres (the synthetic
// sweep, the mass-test cubes with the capsule xArm6).  Hull links and robot obstacles stay on the
// generic k_collide_bits.
//
// One warp owns a group of 32 configurations and pushes them through a pipeline of cull stages.
// Every stage works on a full batch of 32 items (one per lane); survivors are compacted with
// __ballot_sync / popc prefix ranks into per-warp shared-memory stacks, so the lanes stay busy at
// every level although fewer than 1 in 1000 (link, obstacle, frame) pairs reaches the exact test:
//   FK   lane = configuration   FP64 chain (RobotObstacleFCL.cpp:287-386); capsule centre, axis and
//                               1/|axis|^2 narrowed to FP32 into shared memory [link][7][32]
//   A0   lane = configuration   whole-arm bounding sphere vs the swept bounding sphere of every
//                               (32-frame tile, dynamic slot) pair (uniform float4 loads)      -> Q0
//   A1   lane = Q0 entry        the 7 link bounding spheres vs that swept sphere; one Q1 entry per
//                               surviving link (warp prefix sum of popcounts)                   -> Q1
//   B1   lane = Q1 entry        link axis vs the slot's swept capsule over the tile (segment-segment)   -> QW
//   W    lane = QW entry        analytic frame windows (sirrt_sweep_window.cuh): the obstacle centre moves piecewise
//                               linearly inside a tile, so its squared distance to the link axis is a quadratic in
//                               the frame; one root solve yields the frames that are certainly HIT (OR-ed into the
//                               row as a mask, no per-frame work) and the frames that can be neither decided FREE nor
//                               HIT from the model; only those are enumerated, one (configuration, link, slot, frame)
//                               entry each                                                                -> Q2
//   B2   lane = Q2 entry        point-segment distance from the TRUE obstacle centre to the link axis; axis-aligned
//                               boxes: separating-plane lower bound / closest-point upper bound of the segment-box
//                               distance decide FREE / HIT, the rest is UNDECIDED                       -> Q3
//   C    lane = Q3 entry        exact segment-box distance for the UNDECIDED (config, link, slot, frame)
// Pairs inside the FP32 uncertainty band go to the FP64 re-check work list exactly as in the
// generic kernel.  Collision rows accumulate in shared memory and leave as coalesced 128-byte rows.
#pragma once
#include <mutex>
#include <vector>

#include "sirrt_kernels.cuh"
#include "sirrt_sweep_window.cuh"
#include "sirrt_tables.hpp"

#define SWEEP_MAXL 8
#define SWEEP_WARPS 4
#define SWEEP_G 16       // configurations per warp (16 keeps the per-warp tables at ~7 KB -> 28 resident warps per SM)
#define SWEEP_Q0CAP 160   // 31 left over + 32 x 4 pushed per A0 round
#define SWEEP_LMAX 7      // links the queue capacities are sized for (sweep_build rejects robots with more)
#define SWEEP_Q1CAP 256   // 31 left over + 32 x SWEEP_LMAX pushed per A1 batch
#define SWEEP_QWCAP 64    // 31 left over + 32 pushed per B1 batch
#define SWEEP_Q2CAP 96    // 31 left over + 2 x 32 pushed per enumeration round of stage W
#define SWEEP_Q3CAP 64    // 31 left over + 32 pushed per B2 batch
#define SWEEP_ROWS 7      // per link: centre xyz, axis xyz (b - a), 1 / |axis|^2
#define SWEEP_LDC (SWEEP_G + 1)  // row stride of the per-warp link table (odd: rows of one configuration fall into different banks)
#define SWEEP_PER_WARP(L, Wp) ((size_t)(L) * SWEEP_ROWS * SWEEP_LDC + (size_t)SWEEP_G * (Wp) + SWEEP_Q0CAP + SWEEP_Q1CAP + SWEEP_QWCAP + SWEEP_Q2CAP + SWEEP_Q3CAP)

struct SweepTables {
  int supported = 0;
  const float4* bounds = nullptr;  // [pair]: swept bounding sphere of (tile, slot j), pair = tile*ndp + j (ndp = n_dynamic padded to x8)
  const float4* seg = nullptr;     // [pair*2]: (p0, r) and (p1, 0): swept capsule of the slot over the tile
  const float4* dyn = nullptr;     // [j]: inscribed radius, bounding radius, sphere radius or -1, offset of the slot's translation rows in pose_f (int bits)
  const float4* dyn2 = nullptr;    // [j]: box half extents, w = 1 when every pose of the slot has R == I exactly
  const float4* mdl = nullptr;     // [pair*4]: two-piece linear model of the slot's centre over the tile (sirrt_sweep_window.cuh)
  const float4* ctr = nullptr;     // [j*Tp + f]: the slot's centre at frame f (xyz)
  int ndp = 0;
  int n_links = 0;
  float link_r[SWEEP_MAXL], link_br[SWEEP_MAXL];
  int blocks_per_sm = 0;
  size_t smem = 0;
  int Wp = 0;
};

struct SweepParams {
  DScene sc;
  QView q;
  WorkList wl;
  uint32_t* bits;
  const float4* bounds;
  const float4* seg;
  const float4* dyn;
  const float4* dyn2;
  const float4* mdl;
  const float4* ctr;
  int ndp, n_words, Wp, n_links;
  float band;
  float link_r[SWEEP_MAXL], link_br[SWEEP_MAXL];
};

// FK (a1) emitting each base-stripped link's world transform to a functor; no per-thread link array
template <typename F> __device__ __forceinline__ void fk_each_link(const DRobot& rb, const double* q, F&& emit) {
  double t[3] = {rb.base_t[0], rb.base_t[1], rb.base_t[2]}, R[9];
#pragma unroll
  for (int i = 0; i < 9; i++) R[i] = rb.base_R[i];
  int l = 0;
  for (int k = 0; k <= rb.n_joints; k++) {
    while (l < rb.n_links && rb.link_jb[l] == k) {
      double w[3], Rl[9], tl[3];
      matvec(w, R, rb.lo_t[l]);          // RobotObstacleFCL.cpp:335
      tl[0] = t[0] + w[0]; tl[1] = t[1] + w[1]; tl[2] = t[2] + w[2];  // :336
      matmul3(Rl, R, rb.lo_R[l]);        // :347
      emit(l, Rl, tl);
      l++;
    }
    if (k == rb.n_joints) break;
    double w[3];
    matvec(w, R, rb.jt[k]);              // :319,361
    t[0] += w[0]; t[1] += w[1]; t[2] += w[2];
    matmul3(R, R, rb.jR[k]);             // :320,362
    double sn, cs;
    sincos(q[k], &sn, &cs);
    const double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};  // AngleAxisd(a, UnitZ) :321,363
    matmul3(R, R, Rz);
  }
}

// exact FP32 narrowphase: capsule (axis a-b, radius rl) vs box / sphere posed at (R, t)
__device__ __forceinline__ int capsule_exact(const float* a, const float* b, float rl, const DShape& sh, const float* R, const float* t,
                                             bool ident, float band) {
  float d;
  if (sh.type == SH_SPHERE) d = sqrtf(pt_seg_dist2(t, a, b)) - rl - sh.radius;
  else {
    float da[3], db[3], la[3], lb[3];
    sub3(da, a, t); sub3(db, b, t);
    if (ident) { la[0] = da[0]; la[1] = da[1]; la[2] = da[2]; lb[0] = db[0]; lb[1] = db[1]; lb[2] = db[2]; }
    else { matTvec(la, R, da); matTvec(lb, R, db); }
    d = sqrtf(seg_box_dist2(la, lb, sh.half)) - rl;
  }
  return classify(d, band);
}

// per-warp link table: row (l*7 + k), column = configuration of the group, row stride 33 floats
#define LK(l, k, cfg) lk[((l) * SWEEP_ROWS + (k)) * SWEEP_LDC + (cfg)]
// queue entries: configuration (5 bits), link (3), frame (5), pair code = tile << 11 | j (19)
#define SWEEP_PAIR(code, ndp) ((int)((code) >> 11) * (ndp) + (int)((code) & 2047u))

// stage C: exact FP32 narrowphase for up to 32 queued (configuration, link, frame, pair) entries, one per lane
__device__ __noinline__ unsigned sweep_exact_batch(const SweepParams& P, const float* lk, uint32_t* sbits, const float* s_lr,
                                                   const uint32_t* q, int count, int lane, long long cfg0) {
  __syncwarp();
  unsigned n_exact = 0;
  if (lane < count) {
    const DScene& sc = P.sc;
    const uint32_t u = q[lane];
    const int cfg = u & 31, l = (u >> 5) & 7, fr = (u >> 8) & 31;
    const int tile = u >> 24, j = (u >> 13) & 2047;
    if (!((sbits[cfg * P.Wp + tile] >> fr) & 1u)) {  // another pair may have decided this frame already
      const int slot = sc.dynamic_slots[j];
      const DSlot S = sc.slots[slot];
      const DShape& sh = sc.shapes[S.shape];
      const int f = tile * 32 + fr;
      float a[3], b[3], R[9], t[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const float c = LK(l, k, cfg), h = 0.5f * LK(l, 3 + k, cfg);
        a[k] = c - h; b[k] = c + h;
      }
      load_pose(sc, slot, f, R, t, !S.rot_identity);
      const int r = capsule_exact(a, b, s_lr[l], sh, R, t, S.rot_identity != 0, P.band);
      n_exact = 1;
      if (r == RES_HIT) atomicOr(&sbits[cfg * P.Wp + tile], 1u << fr);
      else if (r == RES_AMBIG) push_work(P.wl, (int)(cfg0 + cfg), f, l, slot, 0);
    }
  }
  __syncwarp();
  return n_exact;
}

__global__ void __launch_bounds__(SWEEP_WARPS * 32, 7)
k_sweep(const __grid_constant__ DRobot rb, const __grid_constant__ SweepParams P) {
  extern __shared__ __align__(16) unsigned char sweep_smem[];
  __shared__ float s_lr[SWEEP_MAXL];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const DScene& sc = P.sc;
  const int L = P.n_links, Wp = P.Wp, W = P.n_words, ndp = P.ndp;
  const int T = sc.T, Tp = sc.Tp, n_tiles = ndp > 0 ? sc.n_tiles : 0;
  const float* __restrict__ posef = sc.pose_f;
  float* lk = (float*)sweep_smem + (size_t)wib * SWEEP_PER_WARP(L, Wp);
  uint32_t* sbits = (uint32_t*)(lk + (size_t)L * SWEEP_ROWS * SWEEP_LDC);  // [cfg][Wp]
  uint32_t* q0 = sbits + (size_t)SWEEP_G * Wp;
  uint32_t* q1 = q0 + SWEEP_Q0CAP;
  uint32_t* qw = q1 + SWEEP_Q1CAP;
  uint32_t* q2 = qw + SWEEP_QWCAP;
  uint32_t* q3 = q2 + SWEEP_Q2CAP;
  if (threadIdx.x < SWEEP_MAXL) s_lr[threadIdx.x] = P.link_r[threadIdx.x];
  __syncthreads();
  const float band = P.band;
  const unsigned lt_mask = (1u << lane) - 1u;
  unsigned int n_exact = 0;
  const long long n_groups = (P.q.cnt + SWEEP_G - 1) / SWEEP_G;
  const int cl = lane & (SWEEP_G - 1);  // this lane's configuration of the group (both half-warps hold all 16)

  for (long long group = (long long)blockIdx.x * SWEEP_WARPS + wib; group < n_groups; group += (long long)gridDim.x * SWEEP_WARPS) {
    const long long cfg0 = P.q.n0 + group * SWEEP_G;
    const long long n = cfg0 + cl;
    const bool have = group * SWEEP_G + cl < P.q.cnt;
    __syncwarp();
    for (int i = lane; i < SWEEP_G * Wp; i += 32) sbits[i] = 0u;

    // ---- FK: lane = configuration --------------------------------------------------------------
    {
      double q[SIRRT_MAXJ];
      for (int j = 0; j < rb.n_joints; j++) q[j] = (have && j < P.q.dof) ? load_q(P.q, n, j) : 0.0;
      fk_each_link(rb, q, [&](int l, const double* Rl, const double* tl) {
        const double len = 2.0 * sc.shapes[rb.link_shape[l]].d_half_len;
        float ab[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {  // capsule core: local (0, 0, -+half_len) around the link origin
          ab[k] = (float)(Rl[3 * k + 2] * len);
          LK(l, k, cl) = (float)tl[k];  // both half-warps store the same value
          LK(l, 3 + k, cl) = ab[k];
        }
        const float den = ab[0] * ab[0] + ab[1] * ab[1] + ab[2] * ab[2];
        LK(l, 6, cl) = den > 0.f ? 1.f / den : 0.f;
      });
    }
    __syncwarp();
    // whole-arm bounding sphere of this lane's configuration (covers every link bounding sphere)
    float acx, acy, acz, ar = 0.f;
    {
      float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
#pragma unroll
      for (int l = 0; l < SWEEP_MAXL; l++)
        if (l < L) {
#pragma unroll
          for (int k = 0; k < 3; k++) {
            const float c = LK(l, k, cl);
            lo[k] = fminf(lo[k], c - P.link_br[l]); hi[k] = fmaxf(hi[k], c + P.link_br[l]);
          }
        }
      acx = 0.5f * (lo[0] + hi[0]); acy = 0.5f * (lo[1] + hi[1]); acz = 0.5f * (lo[2] + hi[2]);
#pragma unroll
      for (int l = 0; l < SWEEP_MAXL; l++)
        if (l < L) {
          const float dx = LK(l, 0, cl) - acx, dy = LK(l, 1, cl) - acy, dz = LK(l, 2, cl) - acz;
          ar = fmaxf(ar, sqrtf(dx * dx + dy * dy + dz * dz) + P.link_br[l]);
        }
      ar = ar * 1.000001f + 1e-5f + band;
    }

    // ---- static slots: once per configuration, lane = configuration ---------------------------------
    bool static_hit = false;
    for (int s = 0; s < sc.n_static && have && lane < SWEEP_G; s++) {
      const int slot = sc.static_slots[s];
      const DSlot S = sc.slots[slot];
      const DShape& sh = sc.shapes[S.shape];
      float R[9], t[3];
      load_pose(sc, slot, 0, R, t, true);
      const float* c = sc.bcw_f + (size_t)slot * 3 * Tp;
      const float ox = c[0], oy = c[Tp], oz = c[2 * (size_t)Tp];
      for (int l = 0; l < L && !static_hit; l++) {
        const float dx = LK(l, 0, cl) - ox, dy = LK(l, 1, cl) - oy, dz = LK(l, 2, cl) - oz;
        const float rr = (P.link_br[l] + S.br) * 1.000001f + 1e-5f + band;
        if (dx * dx + dy * dy + dz * dz > rr * rr) continue;
        float a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; k++) { const float cc = LK(l, k, cl), h = 0.5f * LK(l, 3 + k, cl); a[k] = cc - h; b[k] = cc + h; }
        const int r = capsule_exact(a, b, P.link_r[l], sh, R, t, S.rot_identity != 0, band);
        n_exact++;
        if (r == RES_HIT) static_hit = true;
        else if (r == RES_AMBIG) push_work(P.wl, (int)n, -1, l, slot, 0);
      }
    }
    const unsigned shm = __ballot_sync(SIRRT_FULL, static_hit);  // bit c: configuration c of the group hits a static slot
    const bool active = have && !((shm >> cl) & 1u);
    int q0n = 0, q1n = 0, qwn = 0, q2n = 0, q3n = 0;

    // ---- dynamic slots.  Every stage has exactly ONE call site (the kernel's instruction footprint decides how often the
    // 28 resident warps, each in a different stage, miss the instruction cache): nested loops, a stage runs when a full
    // batch of 32 is queued or, once everything in front of it has drained (fN), on what is left.
    for (int tile0 = 0, j0 = 0;;) {
      const bool flush = tile0 >= n_tiles;
      if (!flush) {
        // ---- A0: lane = (configuration, half of the round); arm sphere vs 4 of the round's 8 swept spheres
        const int u0 = (lane >> 4) * 4;
        const float4* bp = P.bounds + (size_t)tile0 * ndp + j0 + u0;
        unsigned mask = 0;
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const float4 tb = __ldg(bp + u);
          const float dx = acx - tb.x, dy = acy - tb.y, dz = acz - tb.z;
          const float rr = ar + tb.w;
          if (dx * dx + dy * dy + dz * dz <= rr * rr) mask |= 1u << u;
        }
        mask <<= u0;
        if (!active) mask = 0;
        if (__any_sync(SIRRT_FULL, mask != 0)) {
          int total;
          int w = q0n + warp_excl_scan(__popc(mask), lane, total);
          const unsigned base = (unsigned)cl | ((((unsigned)tile0 << 11) | (unsigned)j0) << 5);
          while (mask) {
            const int u = __ffs(mask) - 1;
            mask &= mask - 1;
            q0[w++] = base + ((unsigned)u << 5);
          }
          q0n += total;
        }
        j0 += 8;
        if (j0 >= ndp) { j0 = 0; tile0++; }
      }
      if (q0n >= 32 || flush) do {
        if (q0n > 0) {
          // ---- A1: lane = Q0 entry; link bounding spheres vs the swept sphere of the pair
          const int cnt = min(32, q0n);
          q0n -= cnt;
          __syncwarp();
          unsigned lm = 0, e = 0;
          if (lane < cnt) {
            e = q0[q0n + lane];
            const float* lkp = lk + (e & 31);
            const float4 b4 = __ldg(P.bounds + SWEEP_PAIR(e >> 5, ndp));
            const float rb4 = b4.w + band;
#pragma unroll
            for (int l = 0; l < SWEEP_MAXL; l++)
              if (l < L) {
                const float dx = lkp[(l * SWEEP_ROWS + 0) * SWEEP_LDC] - b4.x, dy = lkp[(l * SWEEP_ROWS + 1) * SWEEP_LDC] - b4.y,
                            dz = lkp[(l * SWEEP_ROWS + 2) * SWEEP_LDC] - b4.z;
                const float rr = P.link_br[l] + rb4;
                if (dx * dx + dy * dy + dz * dz <= rr * rr) lm |= 1u << l;
              }
          }
          int total;
          int w = q1n + warp_excl_scan(__popc(lm), lane, total);
          const unsigned ebase = (e & 31u) | ((e >> 5) << 8);
          while (lm) {
            const int l = __ffs(lm) - 1;
            lm &= lm - 1;
            q1[w++] = ebase | ((unsigned)l << 5);
          }
          q1n += total;
          __syncwarp();
        }
        const bool f1 = flush && q0n == 0;
        if (q1n >= 32 || f1) do {
          if (q1n > 0) {
            // ---- B1: lane = Q1 entry; link axis vs the slot's swept capsule over the tile
            const int cnt = min(32, q1n);
            q1n -= cnt;
            bool kp = false;
            unsigned e = 0;
            if (lane < cnt) {
              e = q1[q1n + lane];
              const int l = (e >> 5) & 7;
              const float* lkp = lk + l * (SWEEP_ROWS * SWEEP_LDC) + (e & 31);
              float a[3], b[3];
#pragma unroll
              for (int k = 0; k < 3; k++) { const float cc = lkp[k * SWEEP_LDC], h = 0.5f * lkp[(3 + k) * SWEEP_LDC]; a[k] = cc - h; b[k] = cc + h; }
              const float4* sp = P.seg + 2 * SWEEP_PAIR(e >> 8, ndp);
              const float4 s0 = __ldg(sp), s1 = __ldg(sp + 1);
              const float p0[3] = {s0.x, s0.y, s0.z}, p1[3] = {s1.x, s1.y, s1.z};
              const float rr = s_lr[l] + s0.w + band;
              kp = seg_seg_dist2(a, b, p0, p1) <= rr * rr;
            }
            const unsigned keep = __ballot_sync(SIRRT_FULL, kp);
            if (keep) {
              if (kp) qw[qwn + __popc(keep & lt_mask)] = e;
              qwn += __popc(keep);
            }
          }
          const bool f2 = f1 && q1n == 0;
          if (qwn >= 32 || f2) do {
            // ---- W: lane = QW entry (configuration, link, tile, slot): analytic HIT / IN frame windows
            unsigned tm = 0, code = 0;
            if (qwn > 0) {
              const int cnt = min(32, qwn);
              qwn -= cnt;
              __syncwarp();
              if (lane < cnt) {
                const unsigned e = qw[qwn + lane];
                const int cfg = e & 31, l = (e >> 5) & 7, tile = e >> 19, j = (e >> 8) & 2047;
                const float* lkp = lk + l * (SWEEP_ROWS * SWEEP_LDC) + cfg;
                const float4* mp = P.mdl + 4 * (size_t)SWEEP_PAIR(e >> 8, ndp);
                const float4 m0 = __ldg(mp), m1 = __ldg(mp + 1), m2 = __ldg(mp + 2), m3 = __ldg(mp + 3);
                const float4 dj = __ldg(P.dyn + j);
                const float rl = s_lr[l];
                unsigned in, hit;
                sweep_windows(m0, m1, m2, m3, lkp[0], lkp[SWEEP_LDC], lkp[2 * SWEEP_LDC], lkp[3 * SWEEP_LDC], lkp[4 * SWEEP_LDC],
                              lkp[5 * SWEEP_LDC], lkp[6 * SWEEP_LDC], rl + dj.y + band, rl + dj.x - band, in, hit);
                const uint32_t cur = sbits[cfg * Wp + tile];
                if (hit & ~cur) atomicOr(&sbits[cfg * Wp + tile], hit);
                tm = in & ~hit & ~cur;
                code = (e & 0xffu) | ((e >> 8) << 13);
              }
            }
            const bool f3 = f2 && qwn == 0;
            // the frames the model could not decide are enumerated into Q2, at most two per lane and round
            for (;;) {
              const unsigned b = __ballot_sync(SIRRT_FULL, tm != 0u);
              if (b) {
                if (tm) {
                  const unsigned fr = __ffs(tm) - 1;
                  tm &= tm - 1;
                  q2[q2n + __popc(b & lt_mask)] = code | (fr << 8);
                }
                q2n += __popc(b);
                const unsigned b2 = __ballot_sync(SIRRT_FULL, tm != 0u);
                if (tm) {
                  const unsigned fr = __ffs(tm) - 1;
                  tm &= tm - 1;
                  q2[q2n + __popc(b2 & lt_mask)] = code | (fr << 8);
                }
                q2n += __popc(b2);
              }
              const bool f4 = f3 && !b;
              if (q2n >= 32 || f4) do {
                // ---- B2: lane = Q2 entry (configuration, link, slot, frame), tested against the TRUE centre
                const int cnt = min(32, q2n);  // 0 only on the final drain (f4) with an empty Q2: Q3 may still hold entries
                q2n -= cnt;
                __syncwarp();
                bool want = false;
                unsigned u = 0;
                if (lane < cnt) {
                  u = q2[q2n + lane];
                  const int cfg = u & 31, l = (u >> 5) & 7, fr = (u >> 8) & 31, tile = u >> 24, j = (u >> 13) & 2047;
                  if (!((sbits[cfg * Wp + tile] >> fr) & 1u)) {  // another pair may have decided this frame already
                    const int f = tile * 32 + fr;
                    const float4 c4 = __ldg(P.ctr + (size_t)j * Tp + f);
                    const float4 dj = __ldg(P.dyn + j);
                    const float* lkp = lk + l * (SWEEP_ROWS * SWEEP_LDC) + cfg;
                    const float apx = c4.x - lkp[0], apy = c4.y - lkp[SWEEP_LDC], apz = c4.z - lkp[2 * SWEEP_LDC];
                    const float abx = lkp[3 * SWEEP_LDC], aby = lkp[4 * SWEEP_LDC], abz = lkp[5 * SWEEP_LDC];
                    float tt = (apx * abx + apy * aby + apz * abz) * lkp[6 * SWEEP_LDC];
                    tt = fminf(fmaxf(tt, -0.5f), 0.5f);
                    const float ex = apx - tt * abx, ey = apy - tt * aby, ez = apz - tt * abz;  // closest axis point -> obstacle centre
                    const float d2 = ex * ex + ey * ey + ez * ez;
                    const float rl = s_lr[l];
                    bool hit = false;
                    if (dj.z >= 0.f) {  // sphere obstacle: the point-segment distance IS the narrowphase
                      const int r = classify(sqrtf(d2) - rl - dj.z, band);
                      n_exact++;
                      if (r == RES_HIT) hit = true;
                      else if (r == RES_AMBIG) push_work(P.wl, (int)(cfg0 + cfg), f, l, sc.dynamic_slots[j], 0);
                    } else {
                      const float4 h4 = __ldg(P.dyn2 + j);
                      bool fre;
                      if (h4.w != 0.f) {
                        // axis-aligned box (world frame == box frame up to the translation).  With p* the closest axis point:
                        //   dist(segment, box) <= dist(p*, box)                      -> HIT  when dist(p*, box) < rl - band
                        //   dist(segment, box) >= |e| - h.|e| / |e|  (the plane through p* normal to e separates: the
                        //   segment lies on or behind it, the box reaches at most its support h.|n| towards it)
                        //                                                            -> FREE when that > rl + band
                        const float ax = fabsf(ex), ay = fabsf(ey), az = fabsf(ez);
                        const float qx = fmaxf(ax - h4.x, 0.f), qy = fmaxf(ay - h4.y, 0.f), qz = fmaxf(az - h4.z, 0.f);
                        const float rlo = rl - band, rhi = rl + band;
                        hit = qx * qx + qy * qy + qz * qz < rlo * rlo;
                        fre = d2 - (h4.x * ax + h4.y * ay + h4.z * az) > rhi * sqrtf(d2);
                      } else {
                        const float hi = rl + dj.y + band;  // beyond the circumscribed sphere: free
                        const float lo = rl + dj.x - band;  // inside the inscribed sphere: hit
                        fre = d2 > hi * hi;
                        hit = d2 < lo * lo;
                      }
                      want = !hit && !fre;
                    }
                    if (hit) atomicOr(&sbits[cfg * Wp + tile], 1u << fr);
                  }
                }
                const unsigned mq = __ballot_sync(SIRRT_FULL, want);
                if (mq) {
                  if (want) q3[q3n + __popc(mq & lt_mask)] = u;
                  q3n += __popc(mq);
                }
                const bool f5 = f4 && q2n == 0;
                while (q3n >= 32 || (f5 && q3n > 0)) {
                  // ---- C: lane = Q3 entry: exact segment-box narrowphase
                  const int c3 = min(32, q3n);
                  q3n -= c3;
                  n_exact += sweep_exact_batch(P, lk, sbits, s_lr, q3 + q3n, c3, lane, cfg0);
                }
              } while (q2n >= 32 || (f4 && q2n > 0));
              if (!b) break;
            }
          } while (qwn >= 32 || (f2 && qwn > 0));
        } while (q1n >= 32 || (f1 && q1n > 0));
      } while (q0n >= 32 || (flush && q0n > 0));
      if (flush) break;
    }

    // ---- rows out: lane = word, coalesced ------------------------------------------------------------
    __syncwarp();
    for (int cfg = 0; cfg < SWEEP_G; cfg++) {
      const long long nn = group * SWEEP_G + cfg;
      if (nn >= P.q.cnt) break;
      for (int w = lane; w < W; w += 32) {
        uint32_t v = ((shm >> cfg) & 1u) ? 0xffffffffu : sbits[cfg * Wp + w];
        const int left = T - w * 32;
        if (left < 32) v &= (1u << left) - 1u;
        P.bits[(size_t)(P.q.n0 + nn) * W + w] = v;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(SIRRT_FULL, n_exact, o);
  if (lane == 0 && n_exact) atomicAdd(P.wl.count + 1, n_exact);
}

// ---------------------------------------------------------------------------------------------------
static inline int sweep_upload(const void* src, size_t bytes, std::vector<void*>& owned, const void** out) {
  void* p = nullptr;
  if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return -3; }
  owned.push_back(p);
  if (cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); return -3; }
  *out = p;
  return 0;
}

static inline int sweep_build(const sirrt::HScene& hs, const sirrt::HostTables& tb, const DScene& d, SweepTables* st, std::vector<void*>& owned) {
  st->supported = 0;
  const int L = hs.robot.n_links, nd = tb.n_dynamic;
  if (tb.has_hull || tb.n_robot_slots > 0 || L > SWEEP_LMAX || nd > 2047 || tb.n_tiles > 255) return 0;  // Q1 is sized for SWEEP_LMAX links
  for (int l = 0; l < L; l++) {
    const sirrt::HShape& s = hs.robot.shapes[l + hs.robot.n_strip];
    if (s.type != SH_CAPSULE || s.bc[0] != 0 || s.bc[1] != 0 || s.bc[2] != 0) return 0;
    st->link_r[l] = (float)s.radius;
    st->link_br[l] = tb.shapes[l].br;
  }
  for (int j = 0; j < nd; j++) {
    const int t = tb.shapes[tb.slots[tb.dynamic_slots[j]].shape].type;
    if (t != SH_BOX && t != SH_SPHERE) return 0;
  }
  st->n_links = L;
  const int ndp = (nd + 7) / 8 * 8, n_pairs = tb.n_tiles * ndp;
  // pair tables, pair = tile*ndp + j.  bounds: swept bounding sphere (copied from the scene tables, slots
  // padded to a multiple of 8 with unreachable spheres); seg: swept capsule = segment between the
  // bounding-sphere centres at the tile's first and last frame, radius = largest deviation from it +
  // the slot's radius
  std::vector<float4> bounds((size_t)std::max(n_pairs, 8), make_float4(1e18f, 1e18f, 1e18f, 0.f));
  std::vector<float4> seg((size_t)std::max(n_pairs, 8) * 2, make_float4(0, 0, 0, -1));
  const int Tp = tb.Tp, T = tb.T;
  for (int tile = 0; tile < tb.n_tiles; tile++)
    for (int j = 0; j < nd; j++) {
      const size_t it = (size_t)tile * ndp + j;
      bounds[it] = tb.tile_bounds[(size_t)tile * nd + j];
      const int s = tb.dynamic_slots[j];
      const int f0 = tile * 32, f1 = std::min(tile * 32 + 32, T) - 1;
      auto ctr = [&](int f, double* c) { for (int k = 0; k < 3; k++) c[k] = (double)tb.bcw_f[((size_t)s * 3 + k) * Tp + f]; };
      double p0[3], p1[3], ab[3], den = 0;
      ctr(f0, p0); ctr(f1, p1);
      for (int k = 0; k < 3; k++) { ab[k] = p1[k] - p0[k]; den += ab[k] * ab[k]; }
      double dev = 0;
      for (int f = f0; f <= f1; f++) {
        double c[3], ap[3], t = 0, dd = 0;
        ctr(f, c);
        for (int k = 0; k < 3; k++) { ap[k] = c[k] - p0[k]; t += ap[k] * ab[k]; }
        t = den > 0 ? std::min(1.0, std::max(0.0, t / den)) : 0.0;
        for (int k = 0; k < 3; k++) { const double e = ap[k] - t * ab[k]; dd += e * e; }
        dev = std::max(dev, sqrt(dd));
      }
      const double rad = (dev + (double)tb.slots[s].br) * (1 + 1e-6) + 1e-5;
      seg[it * 2] = make_float4((float)p0[0], (float)p0[1], (float)p0[2], (float)rad);
      seg[it * 2 + 1] = make_float4((float)p1[0], (float)p1[1], (float)p1[2], 0.f);
    }
  // per dynamic slot: inscribed / bounding radius around the pose origin, sphere radius (or -1), offset of
  // the slot's translation rows in pose_f
  std::vector<float4> dyn((size_t)std::max(nd, 1), make_float4(0, 0, -1, 0));
  for (int j = 0; j < nd; j++) {
    const int s = tb.dynamic_slots[j];
    const DShape& sh = tb.shapes[tb.slots[s].shape];
    float4 v;
    v.x = sh.type == SH_SPHERE ? sh.radius : fminf(sh.half[0], fminf(sh.half[1], sh.half[2]));
    v.y = tb.slots[s].br;
    v.z = sh.type == SH_SPHERE ? sh.radius : -1.f;
    const long long toff = ((long long)s * 12 + 9) * Tp;
    if (toff + 3ll * Tp >= (1ll << 31)) return 0;
    const int toff32 = (int)toff;
    memcpy(&v.w, &toff32, 4);
    dyn[j] = v;
  }
  // per dynamic slot: box half extents + "every pose has R == I" flag; the centre per frame as one float4; the two-piece
  // linear model of the centre per (tile, slot) for the analytic frame windows of stage W
  std::vector<float4> dyn2((size_t)std::max(nd, 1), make_float4(0, 0, 0, 0));
  std::vector<float4> ctr((size_t)std::max(nd, 1) * Tp, make_float4(0, 0, 0, 0));
  std::vector<float4> mdl((size_t)std::max(n_pairs, 8) * 4, make_float4(1e18f, 1e18f, 1e18f, 0.f));
  for (int j = 0; j < nd; j++) {
    const int s = tb.dynamic_slots[j];
    const DShape& sh = tb.shapes[tb.slots[s].shape];
    dyn2[j] = make_float4(sh.half[0], sh.half[1], sh.half[2], tb.slots[s].rot_identity ? 1.f : 0.f);
    for (int f = 0; f < Tp; f++) {
      const int ff = std::min(f, T - 1);
      ctr[(size_t)j * Tp + f] = make_float4(tb.bcw_f[((size_t)s * 3 + 0) * Tp + ff], tb.bcw_f[((size_t)s * 3 + 1) * Tp + ff], tb.bcw_f[((size_t)s * 3 + 2) * Tp + ff], 0.f);
    }
    for (int tile = 0; tile < tb.n_tiles; tile++) {
      const int f0 = tile * 32, nf = std::min(32, T - f0);
      double c[32][3];
      for (int f = 0; f < nf; f++)
        for (int k = 0; k < 3; k++) c[f][k] = (double)tb.bcw_f[((size_t)s * 3 + k) * Tp + f0 + f];
      sweep_fit_model(c, nf, &mdl[((size_t)tile * ndp + j) * 4]);
    }
  }
  for (size_t it = 0; it < mdl.size() / 4; it++)  // padding pairs (never survivors of A0): an empty model
    if (mdl[it * 4].x > 1e17f) { mdl[it * 4 + 1] = make_float4(0, 0, 0, 0); mdl[it * 4 + 2] = make_float4(1e18f, 1e18f, 1e18f, 0); mdl[it * 4 + 3] = make_float4(0, 0, 0, 0); }
  const void* p = nullptr;
  if (sweep_upload(bounds.data(), bounds.size() * sizeof(float4), owned, &p)) return -3;
  st->bounds = (const float4*)p;
  if (sweep_upload(seg.data(), seg.size() * sizeof(float4), owned, &p)) return -3;
  st->seg = (const float4*)p;
  if (sweep_upload(dyn.data(), dyn.size() * sizeof(float4), owned, &p)) return -3;
  st->dyn = (const float4*)p;
  if (sweep_upload(dyn2.data(), dyn2.size() * sizeof(float4), owned, &p)) return -3;
  st->dyn2 = (const float4*)p;
  if (sweep_upload(mdl.data(), mdl.size() * sizeof(float4), owned, &p)) return -3;
  st->mdl = (const float4*)p;
  if (sweep_upload(ctr.data(), ctr.size() * sizeof(float4), owned, &p)) return -3;
  st->ctr = (const float4*)p;
  st->ndp = ndp;
  st->Wp = tb.n_tiles | 1;
  st->smem = (size_t)SWEEP_WARPS * SWEEP_PER_WARP(L, st->Wp) * 4;
  if (st->smem > 200 * 1024) return 0;
  // the dynamic shared-memory limit of a kernel is a per-device, process-wide attribute: only ever raise it (a later,
  // smaller scene must not lower the limit of an earlier, larger one)
  {
    static std::mutex mu;
    static size_t granted[64] = {0};
    std::lock_guard<std::mutex> lk(mu);
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && st->smem > granted[dev]) {
      if (cudaFuncSetAttribute(k_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)st->smem) != cudaSuccess) { cudaGetLastError(); return 0; }
      granted[dev] = st->smem;
    }
  }
  int bps = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_sweep, SWEEP_WARPS * 32, st->smem) != cudaSuccess || bps < 1) { cudaGetLastError(); return 0; }
  st->blocks_per_sm = bps;
  st->supported = 1;
  (void)d;
  return 0;
}

static inline int sweep_launch(const SweepTables& st, const DRobot& rb, const DScene& sc, const QView& q, const WorkList& wl, uint32_t* bits,
                               int n_words, float band, int n_sm, cudaStream_t stream) {
  SweepParams P{};
  P.sc = sc; P.q = q; P.wl = wl; P.bits = bits; P.bounds = st.bounds; P.seg = st.seg; P.dyn = st.dyn; P.dyn2 = st.dyn2; P.mdl = st.mdl; P.ctr = st.ctr;
  P.ndp = st.ndp; P.n_words = n_words; P.Wp = st.Wp; P.n_links = st.n_links; P.band = band;
  for (int l = 0; l < SWEEP_MAXL; l++) { P.link_r[l] = l < st.n_links ? st.link_r[l] : 0.f; P.link_br[l] = l < st.n_links ? st.link_br[l] : 0.f; }
  const long long groups = (q.cnt + SWEEP_G - 1) / SWEEP_G;
  const long long blocks = std::min((groups + SWEEP_WARPS - 1) / SWEEP_WARPS, (long long)n_sm * st.blocks_per_sm);
  k_sweep<<<(unsigned)std::max(1ll, blocks), SWEEP_WARPS * 32, st.smem, stream>>>(rb, P);
  return 1;
}
