// lat_probe.cu -- host<->device round-trip building blocks, measured to design the planner-in-the-loop latency path.
// nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o lat_probe lat_probe.cu
#include <chrono>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
__global__ void k_empty() {}
__global__ void k_copy(const double* in, double* out, int n) { if (threadIdx.x < n) out[threadIdx.x] = in[threadIdx.x] + 1.0; }
__global__ void k_copy_flag(const double* in, double* out, int n, volatile unsigned* flag, unsigned tag) {
  if (threadIdx.x < n) out[threadIdx.x] = in[threadIdx.x] + 1.0;
  __syncthreads();
  if (threadIdx.x == 0) { __threadfence_system(); *flag = tag; }
}
static double now() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
template <typename F> static double bench(const char* name, F f, int n = 2000) {
  for (int i = 0; i < 50; i++) f(i);
  double t = now();
  for (int i = 0; i < n; i++) f(i + 50);
  t = (now() - t) / n;
  printf("%-58s %8.2f us\n", name, t);
  return t;
}
int main() {
  cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  double *d_in, *d_out, *p_in, *p_out; double h_in[8] = {1, 2, 3, 4, 5, 6, 7, 8}, h_out[8];
  unsigned* p_flag;
  cudaMalloc(&d_in, 64); cudaMalloc(&d_out, 64);
  cudaHostAlloc(&p_in, 64, cudaHostAllocMapped); cudaHostAlloc(&p_out, 64, cudaHostAllocMapped); cudaHostAlloc(&p_flag, 64, cudaHostAllocMapped);
  memcpy(p_in, h_in, 64); *p_flag = 0;
  bench("launch empty + streamSync", [&](int) { k_empty<<<1, 32, 0, st>>>(); cudaStreamSynchronize(st); });
  bench("2 launches + streamSync", [&](int) { k_empty<<<1, 32, 0, st>>>(); k_empty<<<1, 32, 0, st>>>(); cudaStreamSynchronize(st); });
  bench("4 launches + streamSync", [&](int) { for (int j = 0; j < 4; j++) k_empty<<<1, 32, 0, st>>>(); cudaStreamSynchronize(st); });
  bench("pageable H2D 48B + kernel + pageable D2H 48B + sync", [&](int) {
    cudaMemcpyAsync(d_in, h_in, 48, cudaMemcpyHostToDevice, st); k_copy<<<1, 32, 0, st>>>(d_in, d_out, 6);
    cudaMemcpyAsync(h_out, d_out, 48, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st); });
  bench("pinned H2D 48B + kernel + pinned D2H 48B + sync", [&](int) {
    cudaMemcpyAsync(d_in, p_in, 48, cudaMemcpyHostToDevice, st); k_copy<<<1, 32, 0, st>>>(d_in, d_out, 6);
    cudaMemcpyAsync(p_out, d_out, 48, cudaMemcpyDeviceToHost, st); cudaStreamSynchronize(st); });
  bench("zero-copy in/out kernel + streamSync", [&](int) { k_copy<<<1, 32, 0, st>>>(p_in, p_out, 6); cudaStreamSynchronize(st); });
  bench("zero-copy in/out kernel + spin on mapped flag", [&](int i) {
    k_copy_flag<<<1, 32, 0, st>>>(p_in, p_out, 6, p_flag, (unsigned)i + 1);
    while (*(volatile unsigned*)p_flag != (unsigned)i + 1) {} });
  bench("3 kernels, last zero-copy + spin on mapped flag", [&](int i) {
    k_copy<<<1, 32, 0, st>>>(p_in, d_in, 6); k_copy<<<1, 32, 0, st>>>(d_in, d_out, 6);
    k_copy_flag<<<1, 32, 0, st>>>(d_out, p_out, 6, p_flag, (unsigned)i + 1);
    while (*(volatile unsigned*)p_flag != (unsigned)i + 1) {} });
  bench("memsetAsync + kernel + sync", [&](int) { cudaMemsetAsync(d_in, 0, 16, st); k_empty<<<1, 32, 0, st>>>(); cudaStreamSynchronize(st); });
  cudaPointerAttributes a;
  bench("cudaPointerGetAttributes(host ptr)", [&](int) { cudaPointerGetAttributes(&a, h_in); });
  bench("cudaSetDevice(0)", [&](int) { cudaSetDevice(0); });
  // CUDA graph of 3 kernels
  cudaGraph_t g; cudaGraphExec_t ge;
  cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
  k_copy<<<1, 32, 0, st>>>(p_in, d_in, 6); k_copy<<<1, 32, 0, st>>>(d_in, d_out, 6); k_copy<<<1, 32, 0, st>>>(d_out, p_out, 6);
  cudaStreamEndCapture(st, &g); cudaGraphInstantiate(&ge, g, 0);
  bench("graph(3 kernels) launch + streamSync", [&](int) { cudaGraphLaunch(ge, st); cudaStreamSynchronize(st); });
  printf("check %f err %s\n", p_out[0], cudaGetErrorString(cudaGetLastError()));
  return 0;
}
