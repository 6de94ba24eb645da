// Microbenchmark: MUFU ex2 throughput per SM as a function of resident warps per sub-partition and ILP,
// bare and inside the softmax instruction mix (FFMA -> EX2 -> FADD, F2FP pack).
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
template <int ILP, int MIX>
__global__ void k(float* out, int iters, long long* cyc, float c, float mc) {
  float a[ILP], l[4] = {0, 0, 0, 0}; unsigned pk = 0;
  for (int i = 0; i < ILP; ++i) a[i] = -0.001f * (threadIdx.x + i);
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (MIX == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }
      else {
        float x = fmaf(a[i], c, -mc), y;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
        l[i & 3] += y;
        if (i & 1) { __nv_bfloat162 u = __floats2bfloat162_rn(y, a[i - 1]); pk ^= *reinterpret_cast<unsigned*>(&u); }
        a[i] = y - 1.5f;
      }
    }
  }
  long long t1 = clock64();
  float s = l[0] + l[1] + l[2] + l[3] + __uint_as_float(pk); for (int i = 0; i < ILP; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
  float* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMallocManaged(&cyc, 8);
  const int iters = 4096;
  for (int mix = 0; mix < 2; ++mix)
    for (int threads : {128, 256, 512, 1024}) {
      for (int ilp : {8, 32}) {
        for (int rep = 0; rep < 2; ++rep) {
          if (mix == 0 && ilp == 8) k<8, 0><<<148, threads>>>(out, iters, cyc, 0.18f, 1.f);
          if (mix == 0 && ilp == 32) k<32, 0><<<148, threads>>>(out, iters, cyc, 0.18f, 1.f);
          if (mix == 1 && ilp == 8) k<8, 1><<<148, threads>>>(out, iters, cyc, 0.18f, 1.f);
          if (mix == 1 && ilp == 32) k<32, 1><<<148, threads>>>(out, iters, cyc, 0.18f, 1.f);
          cudaDeviceSynchronize();
        }
        printf("mix %d warps/SMSP %d ilp %2d: %.2f ex2/clk/SM\n", mix, threads / 128, ilp, double(iters) * ilp * threads / *cyc);
      }
    }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
}
