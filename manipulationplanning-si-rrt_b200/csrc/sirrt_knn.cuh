// sirrt_knn.cuh -- exact brute-force nearest-neighbour / radius / top-k kernels in configuration
// space (FP64, squared L2 accumulated in dimension order without FMA contraction, i.e. the
// arithmetic of nanoflann's L2_Simple_Adaptor::evalMetric as used by MSIRRT/src/Tree.cpp:5,18-36
// and PlannerConnect.cpp:382-412).  Points are stored structure-of-arrays [dim][capacity] so that a
// warp reads 32 consecutive points of one dimension in one 256-byte transaction.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SIRRT_KNN_MAXDIM 16
#define SIRRT_KNN_THREADS 256

struct KnnStore {
  const double* pts;  // [dim][cap]
  long long cap, n;
  int dim;
};

__device__ __forceinline__ double knn_dist2(const KnnStore& S, const double* q, long long i) {
  double r = 0.0;
#pragma unroll 1
  for (int k = 0; k < S.dim; k++) {
    const double d = __dsub_rn(q[k], S.pts[(size_t)k * S.cap + i]);
    r = __dadd_rn(r, __dmul_rn(d, d));
  }
  return r;
}

// lexicographic (distance, index) order: exact ties resolve to the lower index, which is what a
// first-wins linear scan (and the oracle) returns
__device__ __forceinline__ bool knn_less(double d1, long long i1, double d2, long long i2) {
  return d1 < d2 || (d1 == d2 && i1 < i2);
}

__device__ __forceinline__ void knn_block_min(double& d, long long& i, double* sd, long long* si) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double od = __shfl_xor_sync(0xffffffffu, d, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (knn_less(od, oi, d, i)) { d = od; i = oi; }
  }
  if (lane == 0) { sd[w] = d; si[w] = i; }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    d = lane < nw ? sd[lane] : 1.0 / 0.0;
    i = lane < nw ? si[lane] : 0x7fffffffffffffffll;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double od = __shfl_xor_sync(0xffffffffu, d, o);
      const long long oi = __shfl_xor_sync(0xffffffffu, i, o);
      if (knn_less(od, oi, d, i)) { d = od; i = oi; }
    }
    if (lane == 0) { sd[0] = d; si[0] = i; }
  }
  __syncthreads();
  d = sd[0]; i = si[0];
  __syncthreads();
}

// one warp per query: reduce the partial minima
__global__ void k_knn_final(const double* pd, const long long* pi, int n_split, long long nq, long long* idx, double* dist2) {
  const int lane = threadIdx.x & 31;
  const long long qi = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= nq) return;
  double d = 1.0 / 0.0;
  long long i = 0x7fffffffffffffffll;
  for (int s = lane; s < n_split; s += 32) {
    const double od = pd[qi * n_split + s];
    const long long oi = pi[qi * n_split + s];
    if (knn_less(od, oi, d, i)) { d = od; i = oi; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double od = __shfl_xor_sync(0xffffffffu, d, o);
    const long long oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (knn_less(od, oi, d, i)) { d = od; i = oi; }
  }
  if (lane == 0) { idx[qi] = i == 0x7fffffffffffffffll ? -1 : i; if (dist2) dist2[qi] = d; }
}

// radius search: strict dist2 < r2 (nanoflann RadiusResultSet); hits are appended with a
// warp-aggregated atomic, the host sorts the (small) hit list ascending
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_radius(const __grid_constant__ KnnStore S, const double* __restrict__ query, double r2, int* hits, unsigned int* count) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  if (threadIdx.x < S.dim) q[threadIdx.x] = query[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n_round = (S.n + stride - 1) / stride * stride;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    const bool in = i < S.n && knn_dist2(S, q, i) < r2;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (!m) continue;
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (in) hits[base + __popc(m & ((1u << lane) - 1))] = (int)i;
  }
}

// ---- latency path: ONE launch per planner query ---------------------------------------------------------
// The planner asks one nearest / one radius query per iteration on a tree of 10^2..10^5 points: launch +
// synchronise (9 us on the B200 box) dominates, so each query is a single kernel: one point per thread with the
// coordinate loads of a point issued together (the per-dimension loop of knn_dist2 would serialise dim cache
// misses), block minima to scratch, and the LAST block to finish (atomic ticket) reduces them and writes the
// answer straight into the caller-visible pinned block.
struct KnnQuery { double v[SIRRT_KNN_MAXDIM]; };  // the query point travels in the kernel parameters (no bus read)

__device__ __forceinline__ double knn_dist2_wide(const KnnStore& S, const double* q, long long i) {
  double p[SIRRT_KNN_MAXDIM];
#pragma unroll
  for (int k = 0; k < SIRRT_KNN_MAXDIM; k++) p[k] = k < S.dim ? S.pts[(size_t)k * S.cap + i] : 0.0;
  double r = 0.0;
#pragma unroll
  for (int k = 0; k < SIRRT_KNN_MAXDIM; k++)
    if (k < S.dim) { const double d = __dsub_rn(q[k], p[k]); r = __dadd_rn(r, __dmul_rn(d, d)); }  // same order, no FMA: bit-equal to knn_dist2
  return r;
}

__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_nearest1(const __grid_constant__ KnnStore S, const __grid_constant__ KnnQuery query, double* pd, long long* pi, unsigned int* ticket,
               long long* out_idx, double* out_d2, unsigned int* done, unsigned int done_seq) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ double sd[32];
  __shared__ long long si[32];
  __shared__ bool last;
  if (threadIdx.x < S.dim) q[threadIdx.x] = query.v[threadIdx.x];
  __syncthreads();
  double best = 1.0 / 0.0;
  long long bi = 0x7fffffffffffffffll;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S.n; i += (long long)gridDim.x * blockDim.x) {
    const double d = knn_dist2_wide(S, q, i);
    if (d < best) { best = d; bi = i; }  // ascending i per thread: strict < keeps the first
  }
  knn_block_min(best, bi, sd, si);
  if (gridDim.x == 1) {  // small trees (up to one block of points): no partial results, no ticket -- the block minimum is the answer
    if (threadIdx.x == 0) {
      out_idx[0] = bi == 0x7fffffffffffffffll ? -1 : bi;
      out_d2[0] = best;
      __threadfence_system();
      *(volatile unsigned int*)done = done_seq;
    }
    return;
  }
  if (threadIdx.x == 0) {
    pd[blockIdx.x] = best; pi[blockIdx.x] = bi;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  best = 1.0 / 0.0; bi = 0x7fffffffffffffffll;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
    const double od = ((volatile double*)pd)[b];
    const long long oi = ((volatile long long*)pi)[b];
    if (knn_less(od, oi, best, bi)) { best = od; bi = oi; }
  }
  knn_block_min(best, bi, sd, si);
  if (threadIdx.x == 0) {
    out_idx[0] = bi == 0x7fffffffffffffffll ? -1 : bi;
    out_d2[0] = best;
    *ticket = 0;  // ready for the next query on this stream
    __threadfence_system();
    *(volatile unsigned int*)done = done_seq;  // completion flag in the pinned block: the host polls its own memory
  }
}

// radius query, single launch: hits appended to hits[1..] with a warp-aggregated atomic on hits[0]; the last
// block copies the count and the first `first` hits into the pinned block and clears the counter
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_radius1(const __grid_constant__ KnnStore S, const __grid_constant__ KnnQuery query, double r2, int* hits, unsigned int* ticket, int* out, int first,
              unsigned int* done, unsigned int done_seq) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ bool last;
  if (threadIdx.x < S.dim) q[threadIdx.x] = query.v[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n_round = (S.n + stride - 1) / stride * stride;
  unsigned int* count = (unsigned int*)hits;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    const bool in = i < S.n && knn_dist2_wide(S, q, i) < r2;
    const unsigned m = __ballot_sync(0xffffffffu, in);
    if (!m) continue;
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(count, (unsigned)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (in) hits[1 + base + __popc(m & ((1u << lane) - 1))] = (int)i;
  }
  if (gridDim.x > 1) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
  }
  __threadfence();  // (one block: its own atomics and stores, ordered for the reads below)
  __syncthreads();
  const unsigned cnt = *(volatile unsigned int*)count;
  const int ncopy = (int)min(cnt, (unsigned)first);
  for (int i = threadIdx.x; i < ncopy; i += blockDim.x) out[1 + i] = ((volatile int*)hits)[1 + i];
  __syncthreads();  // thread 0's system fence below is cumulative over the copies made before this barrier
  if (threadIdx.x == 0) {
    out[0] = (int)cnt; *count = 0; *ticket = 0;
    __threadfence_system();
    *(volatile unsigned int*)done = done_seq;
  }
}

// k nearest of each query in ascending (distance, index) order: one block per query, k selection
// passes, pass j takes the minimum over points strictly greater than the (j-1)-th result
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_topk(const __grid_constant__ KnnStore S, const double* __restrict__ queries, int k, long long* idx, double* dist2) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ double sd[32];
  __shared__ long long si[32];
  const long long qi = blockIdx.x;
  if (threadIdx.x < S.dim) q[threadIdx.x] = queries[qi * S.dim + threadIdx.x];
  __syncthreads();
  double pd = -1.0;
  long long pi = -1;
  for (int j = 0; j < k; j++) {
    double best = 1.0 / 0.0;
    long long bi = 0x7fffffffffffffffll;
    for (long long i = threadIdx.x; i < S.n; i += blockDim.x) {
      const double d = knn_dist2(S, q, i);
      if (knn_less(pd, pi, d, i) && d < best) { best = d; bi = i; }
    }
    knn_block_min(best, bi, sd, si);
    if (threadIdx.x == 0) {
      idx[qi * k + j] = bi == 0x7fffffffffffffffll ? -1 : bi;
      if (dist2) dist2[qi * k + j] = best;
    }
    pd = best; pi = bi;
  }
}

// AoS [n][dim] staging -> SoA store at offset n0, plus the FP32 row copy [cap][row] (zero padded) the tiled multi-query
// kernel filters with, and the largest |coordinate| seen (as float bits: non-negative floats order like unsigned ints)
__global__ void k_knn_append(const double* __restrict__ aos, long long n, int dim, double* pts, long long cap, long long n0, float* pf, int row,
                             unsigned int* maxabs) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * row) return;
  const long long i = t / row;
  const int k = (int)(t % row);
  float f = 0.f;
  if (k < dim) {
    const double v = aos[i * dim + k];
    pts[(size_t)k * cap + n0 + i] = v;
    f = (float)v;
    const float up = __double2float_ru(fabs(v));
    if (__float_as_uint(up) > *(volatile unsigned int*)maxabs) atomicMax(maxabs, __float_as_uint(up));
  }
  pf[(size_t)(n0 + i) * row + k] = f;
}

// ---- throughput path: many queries per call -------------------------------------------------------------------------
// Thread = query (32 queries per warp), block = 8 warps sharing tiles of KNN_TILE points staged once in shared memory and
// read back as broadcasts.  The warps of a block cover G groups of 32 queries and 8/G slices of every tile, so a batch of
// 32 queries still fills a block and a batch of >= 256 reuses each staged point 256 times.
// Filter + verify: the scan runs in FP32 on the row copy (13 FP32 instructions per pair instead of 18 FP64 ones, on a
// pipe four times as wide); a point goes to the exact FP64 evaluation (knn_dist2: the oracle's arithmetic) only when its
// FP32 distance is within a rigorous error bound of the running FP32 minimum:
//   q~, p~ = coordinates rounded to float: |q~ - q| <= u |q|, |p~ - p| <= u |p|, u = 2^-24; the float difference, squares and
//   the D-term sum add relative errors below (D + 3) u, so the computed d~ lies in [(d - a)(1 - r), (d + a)(1 + r)] with
//   a = sqrt(D) u (max|q| + max|p|) and r = (D + 3) u.  The true nearest j and ANY point i satisfy d_j <= d_i, hence
//   d~_j <= d~_i (1 + 2.5 r) + 2 a (1 + r): a point whose d~ exceeds  d~_min (1 + 3 r) + 2.5 a  cannot be the nearest and
//   cannot tie with it.  (a, r are doubled below.)
// Every point that could be the answer is therefore compared in FP64, in ascending index order with a strict <, exactly
// as the single-query kernels and the oracle do: the result is bit-identical (index and distance).
#define KNN_TILE 256

// exact evaluation of a point that passed the FP32 filter (rare: kept out of line so the scan loop stays small)
__device__ __noinline__ void knn_verify(const KnnStore& S, const double* __restrict__ q, long long i, float d2, float k1, float k0, double& best,
                                        long long& bi, float& best32, float& thr) {
  double r = 0.0;
  for (int k = 0; k < S.dim; k++) {
    const double e = __dsub_rn(q[k], S.pts[(size_t)k * S.cap + i]);
    r = __dadd_rn(r, __dmul_rn(e, e));
  }
  if (r < best) { best = r; bi = i; }
  if (d2 < best32) {
    best32 = d2;
    const float t = sqrtf(d2) * k1 + k0;
    thr = t * t * 1.000002f;
  }
}

template <int ROW, int DU>
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_nearest_tiled(const __grid_constant__ KnnStore S, const float* __restrict__ pf, const unsigned int* __restrict__ maxabs,
                    const double* __restrict__ queries, long long nq, int G, double* pd, long long* pi) {
  __shared__ __align__(16) float tile[KNN_TILE * ROW];
  __shared__ double sd[8][32];
  __shared__ long long si[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int qg = w & (G - 1), slice = w / G, n_slices = 8 / G, per_slice = KNN_TILE / n_slices;
  const long long qi = ((long long)blockIdx.y * G + qg) * 32 + lane;
  const bool valid = qi < nq;
  float qf[DU];
  float mq = 0.f;
#pragma unroll
  for (int k = 0; k < DU; k++) {
    const double v = (valid && k < S.dim) ? queries[qi * S.dim + k] : 0.0;
    qf[k] = (float)v;
    mq = fmaxf(mq, __double2float_ru(fabs(v)));
  }
  const float u = 5.9604645e-8f;
  const float alpha = 2.f * sqrtf((float)S.dim) * u * (mq + __uint_as_float(*maxabs)) + 1e-30f, rho = 2.f * (float)(S.dim + 3) * u;
  const float k1 = 1.f + 3.f * rho, k0 = 2.5f * alpha;
  float best32 = __int_as_float(0x7f800000), thr = best32;
  double best = 1.0 / 0.0;
  long long bi = 0x7fffffffffffffffll;
  long long per = (S.n + gridDim.x - 1) / gridDim.x;
  per = (per + KNN_TILE - 1) / KNN_TILE * KNN_TILE;
  const long long lo = per * blockIdx.x, hi = min(S.n, lo + per);
  for (long long base = lo; base < hi; base += KNN_TILE) {
    __syncthreads();
    {
      const long long i = base + threadIdx.x;
      float4* dst = reinterpret_cast<float4*>(tile + threadIdx.x * ROW);
      if (i < hi) {
        const float4* src = reinterpret_cast<const float4*>(pf + (size_t)i * ROW);
#pragma unroll
        for (int k = 0; k < ROW / 4; k++) dst[k] = __ldg(src + k);
      } else {
        const float nanv = __int_as_float(0x7fc00000);  // past the end: every comparison with the threshold is false
#pragma unroll
        for (int k = 0; k < ROW / 4; k++) dst[k] = make_float4(nanv, nanv, nanv, nanv);
      }
    }
    __syncthreads();
    if (!valid) continue;
    const int p0 = slice * per_slice;
    for (int p = p0; p < p0 + per_slice; p += 4) {
      float d2[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        float v[ROW];
#pragma unroll
        for (int k = 0; k < ROW / 4; k++) {
          const float4 t = reinterpret_cast<const float4*>(tile + (p + j) * ROW)[k];
          v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
        }
        float a = 0.f;
#pragma unroll
        for (int k = 0; k < DU; k++) { const float e = qf[k] - v[k]; a = fmaf(e, e, a); }
        d2[j] = a;
      }
      // NaN rows (past the end) fail every comparison; fminf would drop them, so compare each
      if (d2[0] <= thr || d2[1] <= thr || d2[2] <= thr || d2[3] <= thr) {
#pragma unroll
        for (int j = 0; j < 4; j++)   // ascending index; thr only shrinks, so a later point is still judged rigorously
          if (d2[j] <= thr) knn_verify(S, queries + qi * S.dim, base + p + j, d2[j], k1, k0, best, bi, best32, thr);
      }
    }
  }
  sd[w][lane] = best; si[w][lane] = bi;
  __syncthreads();
  if (slice == 0 && valid) {
    for (int s2 = 1; s2 < n_slices; s2++) {
      const double od = sd[qg + s2 * G][lane];
      const long long oi = si[qg + s2 * G][lane];
      if (knn_less(od, oi, best, bi)) { best = od; bi = oi; }
    }
    pd[qi * gridDim.x + blockIdx.x] = best;
    pi[qi * gridDim.x + blockIdx.x] = bi;
  }
}

// ---- k nearest in ONE pass (k <= 32) -------------------------------------------------------------------------------
// Each warp keeps the k best (distance, index) pairs seen so far sorted across its lanes (lane j = j-th smallest).  32
// points are evaluated per step; the few that beat the current k-th are inserted one by one: the insertion position is a
// ballot + popc, the shift a single shfl_up.  Order is lexicographic (distance, index), so the result does not depend on
// the order the points are visited in.  The 8 lists of a block are merged by warp 0 the same way.
struct KnnList { double d; long long i; };

__device__ __forceinline__ void knn_list_insert(double& ld, long long& li, double cd, long long ci, int lane) {
  const int pos = __popc(__ballot_sync(0xffffffffu, knn_less(ld, li, cd, ci)));
  const double ud = __shfl_up_sync(0xffffffffu, ld, 1);
  const long long ui = __shfl_up_sync(0xffffffffu, li, 1);
  if (lane > pos) { ld = ud; li = ui; }
  else if (lane == pos) { ld = cd; li = ci; }
}

// candidates held one per lane (flag `cand`): insert those that still beat the k-th entry
__device__ __forceinline__ void knn_list_offer(double& ld, long long& li, int k, bool cand, double d, long long i, int lane) {
  unsigned m = __ballot_sync(0xffffffffu, cand);
  while (m) {
    const int src = __ffs(m) - 1;
    m &= m - 1;
    const double cd = __shfl_sync(0xffffffffu, d, src);
    const long long ci = __shfl_sync(0xffffffffu, i, src);
    const double kd = __shfl_sync(0xffffffffu, ld, k - 1);
    const long long ki = __shfl_sync(0xffffffffu, li, k - 1);
    if (knn_less(cd, ci, kd, ki)) knn_list_insert(ld, li, cd, ci, lane);
  }
}

// grid (n_split, nq): the k best of query blockIdx.y within split blockIdx.x -> part_[d|i][(q * n_split + split) * k + j]
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_topk_pass(const __grid_constant__ KnnStore S, const double* __restrict__ queries, int k, double* part_d, long long* part_i) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ double sd[8][32];
  __shared__ long long si[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const long long qi = blockIdx.y;
  if (threadIdx.x < S.dim) q[threadIdx.x] = queries[qi * S.dim + threadIdx.x];
  __syncthreads();
  const long long per = (S.n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(S.n, lo + per);
  double ld = 1.0 / 0.0;
  long long li = 0x7fffffffffffffffll;
  double kd = ld;
  long long ki = li;
  for (long long b = lo + w * 32; b < hi; b += SIRRT_KNN_THREADS) {
    const long long i = b + lane;
    const double d = i < hi ? knn_dist2_wide(S, q, i) : 1.0 / 0.0;
    const bool cand = i < hi && knn_less(d, i, kd, ki);
    if (__any_sync(0xffffffffu, cand)) {
      knn_list_offer(ld, li, k, cand, d, i, lane);
      kd = __shfl_sync(0xffffffffu, ld, k - 1);
      ki = __shfl_sync(0xffffffffu, li, k - 1);
    }
  }
  sd[w][lane] = ld; si[w][lane] = li;
  __syncthreads();
  if (w) return;
  for (int o = 1; o < 8; o++) {
    const double d = sd[o][lane];
    const long long i = si[o][lane];
    knn_list_offer(ld, li, k, lane < k && i != 0x7fffffffffffffffll, d, i, lane);
  }
  if (lane < k) {
    const size_t o = ((size_t)qi * gridDim.x + blockIdx.x) * k + lane;
    part_d[o] = ld; part_i[o] = li;
  }
}

// one warp per query: merge the n_split sorted partial lists
__global__ void k_knn_topk_merge(const double* __restrict__ part_d, const long long* __restrict__ part_i, int n_split, long long nq, int k,
                                 long long* idx, double* dist2) {
  const int lane = threadIdx.x & 31;
  const long long qi = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (qi >= nq) return;
  double ld = 1.0 / 0.0;
  long long li = 0x7fffffffffffffffll;
  const long long tot = (long long)n_split * k;
  for (long long b = 0; b < tot; b += 32) {
    const long long j = b + lane;
    const double d = j < tot ? part_d[qi * tot + j] : 1.0 / 0.0;
    const long long i = j < tot ? part_i[qi * tot + j] : 0x7fffffffffffffffll;
    knn_list_offer(ld, li, k, i != 0x7fffffffffffffffll, d, i, lane);
  }
  if (lane < k) {
    idx[qi * k + lane] = li == 0x7fffffffffffffffll ? -1 : li;
    if (dist2) dist2[qi * k + lane] = ld;
  }
}
