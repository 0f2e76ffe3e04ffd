// sirrt_sweep.cuh -- k_sweep: the throughput kernel of the batched safe-interval path for scenes
// whose planned-robot links are all capsules and whose obstacles are boxes / spheres (the synthetic
// sweep, the mass-test cubes with the capsule xArm6).  Hull links and robot obstacles stay on the
// generic k_collide_bits.
//
// One warp owns a group of 32 configurations and walks a cull hierarchy whose lane mapping changes
// per level so that lanes stay busy:
//   FK        lane = configuration   FP64 chain (RobotObstacleFCL.cpp:287-386), capsule end points
//                                    narrowed to FP32 into shared memory [link][xyz a,b][32]
//   level A   lane = configuration   link bounding spheres (registers) vs the swept bounding sphere
//                                    of every dynamic slot over every 32-frame tile (uniform float4
//                                    loads) -> per-lane link mask
//   level B   lane = frame of tile   for each surviving (configuration, tile, slot): links first
//                                    tested against the slot's swept capsule (lane = link), then per
//                                    frame the point-segment distance from the obstacle centre to the
//                                    link axis decides FREE / HIT (inscribed sphere) / UNDECIDED;
//                                    __ballot_sync of the hits is one word of the collision row
//   level C   lane = queued pair     UNDECIDED (configuration, link, slot, frame) pairs are compacted
//                                    into a shared-memory queue and evaluated 32 at a time with the
//                                    exact segment-box distance
// Pairs inside the FP32 uncertainty band go to the FP64 re-check work list exactly as in the
// generic kernel.  Collision rows accumulate in shared memory and leave as coalesced 128-byte rows.
#pragma once
#include <vector>

#include "sirrt_kernels.cuh"
#include "sirrt_tables.hpp"

#define SWEEP_MAXL 8
#define SWEEP_WARPS 4
#define SWEEP_QCAP 64

struct SweepTables {
  int supported = 0;
  const float4* seg = nullptr;  // [(tile*n_dynamic + j)*2]: (p0, r) and (p1, 0): swept capsule of slot j over the tile
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
  const float4* seg;
  int n_words, Wp, n_links;
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

#define SWEEP_PACK(cfg, l, fr, tile, j) ((unsigned)(cfg) | ((unsigned)(l) << 5) | ((unsigned)(fr) << 8) | ((unsigned)(tile) << 13) | ((unsigned)(j) << 21))

__global__ void __launch_bounds__(SWEEP_WARPS * 32)
k_sweep(const __grid_constant__ DRobot rb, const __grid_constant__ SweepParams P) {
  extern __shared__ __align__(16) unsigned char sweep_smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const DScene& sc = P.sc;
  const int L = P.n_links, Wp = P.Wp, W = P.n_words, nd = sc.n_dynamic;
  const size_t per_warp = ((size_t)L * 6 * 32 + (size_t)32 * Wp + SWEEP_QCAP) * 4;
  float* ends = (float*)(sweep_smem + (size_t)wib * per_warp);  // [l][k][32]: k 0..2 = a, 3..5 = b
  uint32_t* sbits = (uint32_t*)(ends + (size_t)L * 6 * 32);     // [cfg][Wp]
  uint32_t* q3 = sbits + (size_t)32 * Wp;                       // level-C queue
  const float band = P.band;
  const unsigned lt_mask = (1u << lane) - 1u;
  unsigned int n_exact = 0;
  const long long n_groups = (P.q.cnt + 31) / 32;

  for (long long group = (long long)blockIdx.x * SWEEP_WARPS + wib; group < n_groups; group += (long long)gridDim.x * SWEEP_WARPS) {
    const long long n = P.q.n0 + group * 32 + lane;
    const bool have = group * 32 + lane < P.q.cnt;
    __syncwarp();
    for (int i = lane; i < 32 * Wp; i += 32) sbits[i] = 0u;

    // ---- FK: lane = configuration --------------------------------------------------------------
    {
      double q[SIRRT_MAXJ];
      for (int j = 0; j < rb.n_joints; j++) q[j] = (have && j < P.q.dof) ? load_q(P.q, n, j) : 0.0;
      fk_each_link(rb, q, [&](int l, const double* Rl, const double* tl) {
        const double hl = sc.shapes[rb.link_shape[l]].d_half_len;
#pragma unroll
        for (int k = 0; k < 3; k++) {  // capsule core: local (0, 0, -+half_len)
          ends[(l * 6 + k) * 32 + lane] = (float)(tl[k] - Rl[3 * k + 2] * hl);
          ends[(l * 6 + 3 + k) * 32 + lane] = (float)(tl[k] + Rl[3 * k + 2] * hl);
        }
      });
    }
    __syncwarp();
    float cx[SWEEP_MAXL], cy[SWEEP_MAXL], cz[SWEEP_MAXL];  // link bounding-sphere centres (registers)
#pragma unroll
    for (int l = 0; l < SWEEP_MAXL; l++) {
      if (l < L) {
        cx[l] = 0.5f * (ends[(l * 6 + 0) * 32 + lane] + ends[(l * 6 + 3) * 32 + lane]);
        cy[l] = 0.5f * (ends[(l * 6 + 1) * 32 + lane] + ends[(l * 6 + 4) * 32 + lane]);
        cz[l] = 0.5f * (ends[(l * 6 + 2) * 32 + lane] + ends[(l * 6 + 5) * 32 + lane]);
      } else cx[l] = cy[l] = cz[l] = 0.f;
    }

    // ---- static slots: once per configuration, lane = configuration ---------------------------------
    bool static_hit = false;
    for (int s = 0; s < sc.n_static && have; s++) {
      const int slot = sc.static_slots[s];
      const DSlot S = sc.slots[slot];
      const DShape& sh = sc.shapes[S.shape];
      float R[9], t[3];
      load_pose(sc, slot, 0, R, t, true);
      const float* c = sc.bcw_f + (size_t)slot * 3 * sc.Tp;
      const float ox = c[0], oy = c[sc.Tp], oz = c[2 * (size_t)sc.Tp];
      for (int l = 0; l < L && !static_hit; l++) {
        float a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; k++) { a[k] = ends[(l * 6 + k) * 32 + lane]; b[k] = ends[(l * 6 + 3 + k) * 32 + lane]; }
        const float dx = 0.5f * (a[0] + b[0]) - ox, dy = 0.5f * (a[1] + b[1]) - oy, dz = 0.5f * (a[2] + b[2]) - oz;
        const float rr = (P.link_br[l] + S.br) * 1.000001f + 1e-5f + band;
        if (dx * dx + dy * dy + dz * dz > rr * rr) continue;
        const int r = capsule_exact(a, b, P.link_r[l], sh, R, t, S.rot_identity != 0, band);
        n_exact++;
        if (r == RES_HIT) static_hit = true;
        else if (r == RES_AMBIG) push_work(P.wl, (int)n, -1, l, slot, 0);
      }
    }
    const bool active = have && !static_hit;
    int q3n = 0;

    // level C: evaluate queued (configuration, link, slot, frame) pairs, one per lane
    auto drain = [&](int count) {
      __syncwarp();
      if (lane < count) {
        const unsigned u = q3[lane];
        const int cfg = u & 31, l = (u >> 5) & 7, fr = (u >> 8) & 31, tile = (u >> 13) & 0xff, j = u >> 21;
        if (!((sbits[cfg * Wp + tile] >> fr) & 1u)) {  // another pair may have decided this frame already
          const int slot = sc.dynamic_slots[j];
          const DSlot S = sc.slots[slot];
          const DShape& sh = sc.shapes[S.shape];
          const int f = tile * 32 + fr;
          float a[3], b[3], R[9], t[3];
#pragma unroll
          for (int k = 0; k < 3; k++) { a[k] = ends[(l * 6 + k) * 32 + cfg]; b[k] = ends[(l * 6 + 3 + k) * 32 + cfg]; }
          load_pose(sc, slot, f, R, t, !S.rot_identity);
          const int r = capsule_exact(a, b, P.link_r[l], sh, R, t, S.rot_identity != 0, band);
          n_exact++;
          if (r == RES_HIT) atomicOr(&sbits[cfg * Wp + tile], 1u << fr);
          else if (r == RES_AMBIG) push_work(P.wl, (int)(P.q.n0 + group * 32 + cfg), f, l, slot, 0);
        }
      }
      __syncwarp();
      const int rest = q3n - count;  // < 32 by construction
      unsigned keep = 0;
      if (lane < rest) keep = q3[count + lane];
      __syncwarp();
      if (lane < rest) q3[lane] = keep;
      q3n = rest;
      __syncwarp();
    };

    // ---- dynamic slots ---------------------------------------------------------------------------
    for (int tile = 0; tile < sc.n_tiles; tile++) {
      const int f = tile * 32 + lane;
      const bool valid = f < sc.T;
      const int fc = valid ? f : sc.T - 1;
      for (int j = 0; j < nd; j++) {
        // level A: lane = configuration
        const float4 tb = __ldg(&sc.tile_bounds[(size_t)tile * nd + j]);
        unsigned lm = 0;
#pragma unroll
        for (int l = 0; l < SWEEP_MAXL; l++) {
          if (l < L) {
            const float dx = cx[l] - tb.x, dy = cy[l] - tb.y, dz = cz[l] - tb.z;
            const float rr = P.link_br[l] + tb.w + band;
            if (dx * dx + dy * dy + dz * dz <= rr * rr) lm |= 1u << l;
          }
        }
        if (!active) lm = 0;
        unsigned m = __ballot_sync(SIRRT_FULL, lm != 0);
        if (!m) continue;
        const int slot = sc.dynamic_slots[j];
        const DSlot S = sc.slots[slot];
        const DShape& sh = sc.shapes[S.shape];
        const float4 s0 = __ldg(&P.seg[((size_t)tile * nd + j) * 2]), s1 = __ldg(&P.seg[((size_t)tile * nd + j) * 2 + 1]);
        const float* pp = sc.pose_f + ((size_t)slot * 12 + 9) * sc.Tp + fc;
        const float t[3] = {pp[0], pp[sc.Tp], pp[2 * (size_t)sc.Tp]};  // obstacle centre at this lane's frame
        const bool is_sphere = sh.type == SH_SPHERE;
        const float insc = is_sphere ? sh.radius : fminf(sh.half[0], fminf(sh.half[1], sh.half[2]));
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const unsigned lms = __shfl_sync(SIRRT_FULL, lm, src);
          // level B, part 1: lane = link; link axis vs the slot's swept capsule over the tile
          bool kp = false;
          if (lane < L && ((lms >> lane) & 1u)) {
            float a[3], b[3];
#pragma unroll
            for (int k = 0; k < 3; k++) { a[k] = ends[(lane * 6 + k) * 32 + src]; b[k] = ends[(lane * 6 + 3 + k) * 32 + src]; }
            const float p0[3] = {s0.x, s0.y, s0.z}, p1[3] = {s1.x, s1.y, s1.z};
            const float rr = P.link_r[lane] + s0.w + band;
            kp = seg_seg_dist2(a, b, p0, p1) <= rr * rr;
          }
          unsigned keep = __ballot_sync(SIRRT_FULL, kp);
          if (!keep) continue;
          // level B, part 2: lane = frame
          const uint32_t cur = sbits[src * Wp + tile];
          bool done = !valid || ((cur >> lane) & 1u);
          bool hit = false;
          while (keep) {
            const int l = __ffs(keep) - 1;
            keep &= keep - 1;
            float a[3], b[3];
#pragma unroll
            for (int k = 0; k < 3; k++) { a[k] = ends[(l * 6 + k) * 32 + src]; b[k] = ends[(l * 6 + 3 + k) * 32 + src]; }
            const float d2 = pt_seg_dist2(t, a, b);
            const float rl = P.link_r[l];
            bool want = false;
            if (!done) {
              if (is_sphere) {  // the point-segment distance IS the narrowphase
                const int r = classify(sqrtf(d2) - rl - insc, band);
                n_exact++;
                if (r == RES_HIT) hit = true;
                else if (r == RES_AMBIG) push_work(P.wl, (int)(P.q.n0 + group * 32 + src), fc, l, slot, 0);
              } else {
                const float hi = rl + S.br + band;   // beyond the circumscribed sphere: free
                const float lo = rl + insc - band;   // inside the inscribed sphere: hit
                if (d2 <= hi * hi) {
                  if (d2 < lo * lo) hit = true;
                  else want = true;
                }
              }
            }
            done |= hit;
            const unsigned mq = __ballot_sync(SIRRT_FULL, want);
            if (mq) {
              if (want) q3[q3n + __popc(mq & lt_mask)] = SWEEP_PACK(src, l, lane, tile, j);
              q3n += __popc(mq);
              if (q3n >= 32) drain(32);
            }
          }
          const unsigned word = __ballot_sync(SIRRT_FULL, hit);
          if (word && lane == 0) atomicOr(&sbits[src * Wp + tile], word);
          __syncwarp();
        }
      }
    }
    if (q3n > 0) drain(q3n);

    // ---- rows out: lane = word, coalesced ------------------------------------------------------------
    __syncwarp();
    const unsigned shm = __ballot_sync(SIRRT_FULL, static_hit);
    for (int cfg = 0; cfg < 32; cfg++) {
      const long long nn = group * 32 + cfg;
      if (nn >= P.q.cnt) break;
      for (int w = lane; w < W; w += 32) {
        uint32_t v = ((shm >> cfg) & 1u) ? 0xffffffffu : sbits[cfg * Wp + w];
        const int left = sc.T - w * 32;
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
static inline int sweep_build(const sirrt::HScene& hs, const sirrt::HostTables& tb, const DScene& d, SweepTables* st, std::vector<void*>& owned) {
  st->supported = 0;
  const int L = hs.robot.n_links, nd = tb.n_dynamic;
  if (tb.has_hull || tb.n_robot_slots > 0 || L > SWEEP_MAXL || nd > 2047 || tb.n_tiles > 255) return 0;
  for (int l = 0; l < L; l++) {
    const sirrt::HShape& s = hs.robot.shapes[l + hs.robot.n_strip];
    if (s.type != SH_CAPSULE || s.bc[0] != 0 || s.bc[1] != 0 || s.bc[2] != 0) return 0;
    st->link_r[l] = (float)s.radius;
    st->link_br[l] = tb.shapes[l].br;
  }
  st->n_links = L;
  // swept capsule of every dynamic slot over every tile: segment between the bounding-sphere centres
  // at the tile's first and last frame, radius = largest deviation from it + the slot's radius
  std::vector<float4> seg((size_t)tb.n_tiles * std::max(nd, 1) * 2, make_float4(0, 0, 0, -1));
  const int Tp = tb.Tp, T = tb.T;
  for (int tile = 0; tile < tb.n_tiles; tile++)
    for (int j = 0; j < nd; j++) {
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
      seg[((size_t)tile * nd + j) * 2] = make_float4((float)p0[0], (float)p0[1], (float)p0[2], (float)rad);
      seg[((size_t)tile * nd + j) * 2 + 1] = make_float4((float)p1[0], (float)p1[1], (float)p1[2], 0.f);
    }
  void* p = nullptr;
  if (cudaMalloc(&p, seg.size() * sizeof(float4)) != cudaSuccess) { cudaGetLastError(); return -3; }
  owned.push_back(p);
  if (cudaMemcpy(p, seg.data(), seg.size() * sizeof(float4), cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); return -3; }
  st->seg = (const float4*)p;
  st->Wp = tb.n_tiles | 1;
  st->smem = (size_t)SWEEP_WARPS * ((size_t)L * 6 * 32 + (size_t)32 * st->Wp + SWEEP_QCAP) * 4;
  if (st->smem > 200 * 1024) return 0;
  if (cudaFuncSetAttribute(k_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)st->smem) != cudaSuccess) { cudaGetLastError(); return 0; }
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
  P.sc = sc; P.q = q; P.wl = wl; P.bits = bits; P.seg = st.seg; P.n_words = n_words; P.Wp = st.Wp; P.n_links = st.n_links; P.band = band;
  for (int l = 0; l < SWEEP_MAXL; l++) { P.link_r[l] = l < st.n_links ? st.link_r[l] : 0.f; P.link_br[l] = l < st.n_links ? st.link_br[l] : 0.f; }
  const long long groups = (q.cnt + 31) / 32;
  const long long blocks = std::min((groups + SWEEP_WARPS - 1) / SWEEP_WARPS, (long long)n_sm * st.blocks_per_sm);
  k_sweep<<<(unsigned)std::max(1ll, blocks), SWEEP_WARPS * 32, st.smem, stream>>>(rb, P);
  return 1;
}
