#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned* out, int iters, long long* cyc) {
  unsigned a[16];
  for (int i = 0; i < 16; ++i) a[i] = 0xbc00bc00u + threadIdx.x + i;   // two small negative bf16/f16
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(a[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(a[i]));
      if (MODE == 2) { float f = __uint_as_float(a[i]); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f)); a[i] = __float_as_uint(f); }
    }
  }
  long long t1 = clock64();
  unsigned s = 0; for (int i = 0; i < 16; ++i) s ^= a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  unsigned* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 4096;
  for (int mode = 0; mode < 3; ++mode)
    for (int threads : {256, 512, 1024}) {
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, threads>>>(out, iters, cyc);
        if (mode == 1) k<1><<<148, threads>>>(out, iters, cyc);
        if (mode == 2) k<2><<<148, threads>>>(out, iters, cyc);
        cudaDeviceSynchronize();
      }
      printf("mode %d (%s) threads %4d: %.2f instr/clk/SM = %.2f exp/clk/SM\n", mode, mode == 0 ? "bf16x2" : mode == 1 ? "f16x2" : "f32", threads,
             double(iters) * 16 * threads / *cyc, double(iters) * 16 * threads / *cyc * (mode < 2 ? 2 : 1));
    }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
}
