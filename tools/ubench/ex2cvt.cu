// Does F2FP.BF16.F32.PACK_AB (cvt.rn.bf16x2.f32) compete with MUFU.EX2 for the same pipe?
// Per thread and iteration: 16 ex2, plus (MODE 1) 8 cvt packs, (MODE 2) 8 PRMT packs, (MODE 3) 8 cvt packs alone.
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned* out, int iters, long long* cyc) {
  float a[16]; unsigned acc[8];
  for (int i = 0; i < 16; ++i) a[i] = -1.0f - 0.001f * float(threadIdx.x + i);
  for (int i = 0; i < 8; ++i) acc[i] = 0;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i)
      if (MODE != 3) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      unsigned pk = 0;
      if (MODE == 1 || MODE == 3) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(pk) : "f"(a[2 * i + 1]), "f"(a[2 * i]));
      if (MODE == 2) asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(pk) : "r"(__float_as_uint(a[2 * i])), "r"(__float_as_uint(a[2 * i + 1])));
      acc[i] ^= pk;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = -a[i] - 1.0f;    // keep the chain alive (FMA pipe)
  }
  long long t1 = clock64();
  unsigned s = 0; for (int i = 0; i < 8; ++i) s ^= acc[i];
  for (int i = 0; i < 16; ++i) s ^= __float_as_uint(a[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  unsigned* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 4096;
  const char* names[] = {"16 ex2", "16 ex2 + 8 cvt.bf16x2", "16 ex2 + 8 prmt", "8 cvt.bf16x2 only"};
  for (int mode = 0; mode < 4; ++mode)
    for (int threads : {256, 512}) {
      for (int rep = 0; rep < 2; ++rep) {
        if (mode == 0) k<0><<<148, threads>>>(out, iters, cyc);
        if (mode == 1) k<1><<<148, threads>>>(out, iters, cyc);
        if (mode == 2) k<2><<<148, threads>>>(out, iters, cyc);
        if (mode == 3) k<3><<<148, threads>>>(out, iters, cyc);
        cudaDeviceSynchronize();
      }
      printf("%-24s threads %4d: %.1f clk per iteration per SM sub-partition warp-set (%.2f thread-iterations/clk/SM)\n", names[mode], threads,
             double(*cyc) / iters, double(iters) * threads / *cyc);
    }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
}
