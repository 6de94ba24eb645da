// Microbenchmark: MUFU ex2 throughput, f32 vs packed bf16x2 / f16x2 (per-SM results per clock).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  float a[8]; unsigned b[8];
  for (int i = 0; i < 8; ++i) { a[i] = -0.001f * (threadIdx.x + i); b[i] = 0xBC00BC00u + threadIdx.x + i; }
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(b[i]));
      if (MODE == 2) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(b[i]));
    }
  }
  long long t1 = clock64();
  float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + __uint_as_float(b[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 4096, threads = 1024;
  const char* names[3] = {"ex2.f32", "ex2.bf16x2", "ex2.f16x2"};
  for (int m = 0; m < 3; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      if (m == 0) k<0><<<148, threads>>>(out, iters, cyc);
      if (m == 1) k<1><<<148, threads>>>(out, iters, cyc);
      if (m == 2) k<2><<<148, threads>>>(out, iters, cyc);
      cudaDeviceSynchronize();
    }
    double instr_lanes = double(iters) * 8 * threads;   // per SM
    double res = instr_lanes * (m == 0 ? 1 : 2);
    printf("%-12s %lld cycles: %.2f lane-instr/clk/SM, %.2f results/clk/SM\n", names[m], *cyc, instr_lanes / *cyc, res / *cyc);
  }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
}
