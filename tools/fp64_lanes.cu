// Microbenchmark behind the lane-packing decision of the latency-bound kernels: how long does ONE warp take per
// FP64 instruction as a function of which lanes are active?  (a) dependent DFMA chain (latency), (b) 8 independent
// chains (issue / pipe throughput of a single warp), for active-lane sets {32, low 16, low 8, 1, even lanes, 20}.
// If the FP64 pipe serves a warp in passes over lane groups and skips empty groups, packing the few surviving
// trials of the straggler regime into the low lanes of a warp is worth it; if not, lane placement is irrelevant.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_lanes.bin tools/fp64_lanes.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int ILP>
__global__ void run(double* out, long long* cycles, int n, unsigned mask, double a, double b) {
  const bool on = (mask >> threadIdx.x) & 1u;
  double x[ILP];
#pragma unroll
  for (int k = 0; k < ILP; ++k) x[k] = threadIdx.x * 1e-3 + k;
  long long t0 = 0, t1 = 0;
  if (on) {
    t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
#pragma unroll
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int k = 0; k < ILP; ++k) x[k] = fma(x[k], a, b);
    }
    t1 = clock64();
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < ILP; ++k) s += x[k];
  out[threadIdx.x] = s;
  if (on && (mask & ((1u << threadIdx.x) - 1u)) == 0) *cycles = t1 - t0;  // lowest active lane
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(double) * 64);
  cudaMallocManaged(&cyc, sizeof(long long));
  const int n = 1 << 13;
  struct { const char* name; unsigned mask; } sets[] = {{"all32", 0xffffffffu}, {"low16", 0xffffu}, {"low8", 0xffu},
                                                         {"lane0", 1u}, {"even16", 0x55555555u}, {"low20", 0xfffffu},
                                                         {"high16", 0xffff0000u}};
  printf("{");
  for (int s = 0; s < 7; ++s) {
    double lat = 0, thr = 0;
    for (int rep = 0; rep < 2; ++rep) {
      run<1><<<1, 32>>>(out, cyc, n, sets[s].mask, 0.999999, 1e-9);
      cudaDeviceSynchronize();
      lat = (double)*cyc / (n * 8.0);
      run<8><<<1, 32>>>(out, cyc, n, sets[s].mask, 0.999999, 1e-9);
      cudaDeviceSynchronize();
      thr = (double)*cyc / (n * 64.0);
    }
    printf("%s\"%s\": {\"dep_chain_cycles_per_dfma\": %.2f, \"ilp8_cycles_per_dfma\": %.2f}", s ? ", " : "", sets[s].name,
           lat, thr);
  }
  printf("}\n");
  return 0;
}
