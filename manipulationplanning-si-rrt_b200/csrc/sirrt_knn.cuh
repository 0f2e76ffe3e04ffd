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

// grid (n_split, nq): partial minimum of query blockIdx.y over the point range of split blockIdx.x
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_partial(const __grid_constant__ KnnStore S, const double* __restrict__ queries, double* pd, long long* pi) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ double sd[32];
  __shared__ long long si[32];
  const long long qi = blockIdx.y;
  if (threadIdx.x < S.dim) q[threadIdx.x] = queries[qi * S.dim + threadIdx.x];
  __syncthreads();
  const long long per = (S.n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(S.n, lo + per);
  double best = 1.0 / 0.0;
  long long bi = 0x7fffffffffffffffll;
  for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
    const double d = knn_dist2(S, q, i);
    if (d < best) { best = d; bi = i; }  // ascending i per thread: strict < keeps the first
  }
  knn_block_min(best, bi, sd, si);
  if (threadIdx.x == 0) { pd[qi * gridDim.x + blockIdx.x] = best; pi[qi * gridDim.x + blockIdx.x] = bi; }
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
               long long* out_idx, double* out_d2) {
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
  }
}

// radius query, single launch: hits appended to hits[1..] with a warp-aggregated atomic on hits[0]; the last
// block copies the count and the first `first` hits into the pinned block and clears the counter
__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_radius1(const __grid_constant__ KnnStore S, const __grid_constant__ KnnQuery query, double r2, int* hits, unsigned int* ticket, int* out, int first) {
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
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!last) return;
  __threadfence();
  const unsigned cnt = *(volatile unsigned int*)count;
  const int ncopy = (int)min(cnt, (unsigned)first);
  for (int i = threadIdx.x; i < ncopy; i += blockDim.x) out[1 + i] = ((volatile int*)hits)[1 + i];
  __syncthreads();
  if (threadIdx.x == 0) { out[0] = (int)cnt; *count = 0; *ticket = 0; }
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

// AoS [n][dim] staging -> SoA store at offset n0
__global__ void k_knn_append(const double* __restrict__ aos, long long n, int dim, double* pts, long long cap, long long n0) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * dim) return;
  const long long i = t / dim;
  const int k = (int)(t % dim);
  pts[(size_t)k * cap + n0 + i] = aos[t];
}
