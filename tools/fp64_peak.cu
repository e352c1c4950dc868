// FP64 microbenchmark for the roofline denominators (SURVEY.md section 8d asks for a MEASURED DFMA peak):
//   latency   : cycles per DFMA of one dependent chain in a single warp
//   peak      : DFMA throughput with every SM saturated (8 independent chains per thread, 1024 threads per SM x 2)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak.bin tools/fp64_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

__global__ void chain(double* out, long long* cycles, int n, double a, double b) {
  double x = threadIdx.x * 1e-3;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < n; ++i) x = fma(x, a, b);
  long long t1 = clock64();
  if (threadIdx.x == 0) *cycles = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = x;
}

template <int ILP>
__global__ void flood(double* out, int n, double a, double b) {
  double x[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) x[k] = threadIdx.x * 1e-3 + k;
  for (int i = 0; i < n; ++i) {
#pragma unroll
    for (int k = 0; k < ILP; ++k) x[k] = fma(x[k], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) s += x[k];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  double* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(double) * 4096 * 1024);
  cudaMallocManaged(&cyc, sizeof(long long));
  const int n = 1 << 16;
  chain<<<1, 32>>>(out, cyc, n, 0.999999, 1e-9);
  cudaDeviceSynchronize();
  chain<<<1, 32>>>(out, cyc, n, 0.999999, 1e-9);
  cudaDeviceSynchronize();
  const double lat = (double)*cyc / n;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int blocks = p.multiProcessorCount * 2, threads = 1024, iters = 1 << 14;
  flood<8><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    flood<8><<<blocks, threads>>>(out, iters, 0.999999, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const double fmas = (double)blocks * threads * 8.0 * iters;
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_latency_cycles\": %.2f, \"fp64_tflops\": %.2f, "
         "\"dfma_per_clk_per_sm_at_max_clock\": %.1f, \"max_clock_mhz\": %d}\n",
         p.name, p.multiProcessorCount, lat, 2.0 * fmas / (best * 1e-3) / 1e12,
         fmas / (best * 1e-3) / ((double)clk * 1e3) / p.multiProcessorCount, clk / 1000);
  return 0;
}
