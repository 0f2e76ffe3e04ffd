// sirrt_api_noalloc.inl -- sirrt_knn_nearest_noalloc: exact 1-NN for callers that may not allocate (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  Included by sirrt_api.cu (same translation unit as the
// scratch array below).  No counterpart in the reference.
//
// The kernel is the multi-block form of k_knn_nearest1 (sirrt_knn.cuh): every block scans a slice of the points in the
// oracle's arithmetic (FP64, dimension order, no FMA; knn_dist2_wide), leaves its minimum in SCRATCH, and the last block
// of a query to finish (ticket) reduces the block minima -- ties to the lowest index, as every other kNN entry point.
// The scratch -- block minima and tickets -- is NOT allocated: it is the file-scope device array g_noalloc_scratch, of a
// fixed size that bounds the number of queries per call.  Every GPU that loads the library holds the array until exit.
//
// ALL CALLS SHARE THE ARRAY.  It keeps the last call's contents; the tickets (which must start at zero) are cleared on
// the caller's stream in front of every launch.  Two calls that overlap on one GPU (different streams, or different host
// threads) RACE on it: the caller serialises them.
#define SIRRT_NOALLOC_BLOCKS 64  // blocks per query = block minima per query in the scratch
struct NoAllocScratch {
  double d2[SIRRT_KNN_NOALLOC_MAX_QUERIES][SIRRT_NOALLOC_BLOCKS];
  long long idx[SIRRT_KNN_NOALLOC_MAX_QUERIES][SIRRT_NOALLOC_BLOCKS];
  unsigned int ticket[SIRRT_KNN_NOALLOC_MAX_QUERIES];
};
// 256 queries x 64 blocks x 16 bytes + 1 KB of tickets = 257 KB per device
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

__global__ void __launch_bounds__(SIRRT_KNN_THREADS)
k_knn_nearest_noalloc(const __grid_constant__ KnnStore S, const double* __restrict__ queries, NoAllocScratch* scratch, long long* out_idx, double* out_d2) {
  __shared__ double q[SIRRT_KNN_MAXDIM];
  __shared__ double sd[32];
  __shared__ long long si[32];
  __shared__ bool last;
  const int qi = blockIdx.y;
  if (threadIdx.x < S.dim) q[threadIdx.x] = queries[(size_t)qi * S.dim + threadIdx.x];
  __syncthreads();
  double best = 1.0 / 0.0;
  long long bi = 0x7fffffffffffffffll;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S.n; i += (long long)gridDim.x * blockDim.x) {
    const double d = knn_dist2_wide(S, q, i);
    if (d < best) { best = d; bi = i; }  // ascending i per thread: strict < keeps the first
  }
  knn_block_min(best, bi, sd, si);
  if (threadIdx.x == 0) {
    scratch->d2[qi][blockIdx.x] = best; scratch->idx[qi][blockIdx.x] = bi;
    __threadfence();
    last = atomicAdd(&scratch->ticket[qi], 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  best = 1.0 / 0.0; bi = 0x7fffffffffffffffll;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
    const double od = ((volatile double*)scratch->d2[qi])[b];
    const long long oi = ((volatile long long*)scratch->idx[qi])[b];
    if (knn_less(od, oi, best, bi)) { best = od; bi = oi; }
  }
  knn_block_min(best, bi, sd, si);
  if (threadIdx.x == 0) {
    out_idx[qi] = bi == 0x7fffffffffffffffll ? -1 : bi;
    out_d2[qi] = best;
  }
}

// Device address of the scratch array, looked up once per device WITH THAT DEVICE CURRENT (in host code the array's name is
// a host-side stand-in, not a device address).
static int noalloc_scratch_ptr(int device, NoAllocScratch** out) {
  static std::mutex mu;
  static NoAllocScratch* ptr[64] = {nullptr};
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) { cudaGetLastError(); return fail(SIRRT_E_CUDA, "sirrt_knn_nearest_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e)); }
    ptr[device] = (NoAllocScratch*)p;
  }
  *out = ptr[device];
  return SIRRT_OK;
}

extern "C" int sirrt_knn_nearest_noalloc(int device, void* cuda_stream, const double* pts_soa, int64_t n, int64_t stride, int dim, const double* queries,
                                         int64_t nq, int64_t* idx, double* dist2) {
  if (!pts_soa || !queries || !idx || !dist2 || n < 0 || nq < 0 || stride < n || dim < 1 || dim > SIRRT_KNN_MAXDIM)
    return fail(SIRRT_E_ARG, "sirrt_knn_nearest_noalloc: bad argument");
  if (nq > SIRRT_KNN_NOALLOC_MAX_QUERIES)
    return fail(SIRRT_E_LIMIT, "sirrt_knn_nearest_noalloc: more than SIRRT_KNN_NOALLOC_MAX_QUERIES (256) queries per call; the scratch array is fixed, split the batch");
  if (n > 0x7fffffffll) return fail(SIRRT_E_LIMIT, "sirrt_knn_nearest_noalloc: more than 2^31-1 points");
  if (device < 0 || device >= 64) return fail(SIRRT_E_ARG, "sirrt_knn_nearest_noalloc: device index out of range");
  if (nq == 0) return SIRRT_OK;
  CU(cudaSetDevice(device));
  NoAllocScratch* scratch = nullptr;
  int rc = noalloc_scratch_ptr(device, &scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  // the tickets must read zero when the kernel starts; everything else in the array is written before it is read
  CU(cudaMemsetAsync(scratch->ticket, 0, sizeof(unsigned int) * (size_t)nq, st));
  KnnStore S{pts_soa, (long long)stride, (long long)n, dim};
  const unsigned blocks = (unsigned)std::max<long long>(1, std::min<long long>(SIRRT_NOALLOC_BLOCKS, (n + SIRRT_KNN_THREADS - 1) / SIRRT_KNN_THREADS));
  static_assert(sizeof(long long) == sizeof(int64_t), "index type");
  k_knn_nearest_noalloc<<<dim3(blocks, (unsigned)nq), SIRRT_KNN_THREADS, 0, st>>>(S, queries, scratch, (long long*)idx, dist2);
  CU(cudaGetLastError());
  return SIRRT_OK;
}
