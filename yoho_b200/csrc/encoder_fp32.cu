// float32 SIMT kernels of the encoder: the 1e-4 parity path (YOHO_COMPUTE_FP32) and the on-device
// yardstick the tcgen05 kernels are debugged against.  fp32 operands, fp32 FMA accumulation.
//
// One register-tiled SGEMM covers every contraction of whisper.py:117-135: the dense layers
// (y = x @ kernel), the two convolutions as implicit GEMM (row -> (clip, t), k -> (tap, channel),
// zero padding folded into the loader) and the two attention products (q k^T with b_transposed,
// softmax(q k^T) v) batched over (clip, head) in blockIdx.z.
#include "kernels.cuh"

namespace yoho {

constexpr int BM = 128, BN = 64, BK = 16;
constexpr int kSgemmThreads = 256;  // 16 x 16 threads, 8 x 4 outputs each

__device__ __forceinline__ float gelu_tanh_exact(float x) {
  // flax nn.gelu(approximate=True): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
  const float c = 0.7978845608028654f;
  return 0.5f * x * (1.0f + tanhf(c * (x + 0.044715f * x * x * x)));
}

__device__ __forceinline__ float load_a(const SgemmArgs& g, const float* __restrict__ a, int row, int k) {
  if (row >= g.m || k >= g.k) return 0.0f;
  if (!g.conv) return __ldg(a + size_t(row) * g.lda + k);
  int clip = row / g.conv_out, t = row - clip * g.conv_out;
  int tap = k / g.conv_ch, ch = k - tap * g.conv_ch;
  int ts = g.conv_stride * t + tap - 1;
  if (ts < 0 || ts >= g.conv_len) return 0.0f;
  return __ldg(a + (size_t(clip) * g.conv_len + ts) * g.conv_ch + ch);
}

__device__ __forceinline__ float load_b(const SgemmArgs& g, const float* __restrict__ b, int k, int n) {
  if (k >= g.k || n >= g.n) return 0.0f;
  return g.b_transposed ? __ldg(b + size_t(n) * g.ldb + k) : __ldg(b + size_t(k) * g.ldb + n);
}

__global__ void __launch_bounds__(kSgemmThreads)
sgemm_kernel(const SgemmArgs g) {
  __shared__ float As[2][BK][BM + 4];
  __shared__ float Bs[2][BK][BN + 4];
  const int z1 = blockIdx.z / g.nz2, z2 = blockIdx.z - z1 * g.nz2;
  const float* __restrict__ a = g.a + z1 * g.sa1 + z2 * g.sa2;
  const float* __restrict__ b = g.b + z1 * g.sb1 + z2 * g.sb2;
  float* __restrict__ c = g.c + z1 * g.sc1 + z2 * g.sc2;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;

  // loader mapping: A tile 128 x 16 -> 8 scalars per thread; B tile 16 x 64 -> 4 per thread
  const int a_r = tid >> 1, a_k = (tid & 1) * 8;           // row a_r, k a_k..a_k+7
  const int b_k = g.b_transposed ? (tid & 3) * 4 : tid >> 4;  // transposed: n = tid>>2, k 4-run
  const int b_n = g.b_transposed ? tid >> 2 : (tid & 15) * 4;
  const bool a_vec = !g.conv && (g.lda % 4 == 0) && (reinterpret_cast<uintptr_t>(a) % 16 == 0);
  const bool b_vec = (g.ldb % 4 == 0) && (reinterpret_cast<uintptr_t>(b) % 16 == 0);

  float ra[8], rb[4];
  auto fetch = [&](int k0) {
    int row = m0 + a_r, kk = k0 + a_k;
    if (a_vec && row < g.m && kk + 8 <= g.k) {
      const float4* p = reinterpret_cast<const float4*>(a + size_t(row) * g.lda + kk);
      float4 u = __ldg(p), w = __ldg(p + 1);
      ra[0] = u.x; ra[1] = u.y; ra[2] = u.z; ra[3] = u.w; ra[4] = w.x; ra[5] = w.y; ra[6] = w.z; ra[7] = w.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) ra[j] = load_a(g, a, row, kk + j);
    }
    if (g.b_transposed) {
      int n = n0 + b_n, kb = k0 + b_k;
      if (b_vec && n < g.n && kb + 4 <= g.k) {
        float4 u = __ldg(reinterpret_cast<const float4*>(b + size_t(n) * g.ldb + kb));
        rb[0] = u.x; rb[1] = u.y; rb[2] = u.z; rb[3] = u.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) rb[j] = load_b(g, b, kb + j, n);
      }
    } else {
      int kb = k0 + b_k, n = n0 + b_n;
      if (b_vec && kb < g.k && n + 4 <= g.n) {
        float4 u = __ldg(reinterpret_cast<const float4*>(b + size_t(kb) * g.ldb + n));
        rb[0] = u.x; rb[1] = u.y; rb[2] = u.z; rb[3] = u.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) rb[j] = load_b(g, b, kb, n + j);
      }
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int j = 0; j < 8; ++j) As[buf][a_k + j][a_r] = ra[j];
    if (g.b_transposed) {
#pragma unroll
      for (int j = 0; j < 4; ++j) Bs[buf][b_k + j][b_n] = rb[j];
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) Bs[buf][b_k][b_n + j] = rb[j];
    }
  };

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  const int nk = (g.k + BK - 1) / BK;
  fetch(0);
  stash(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) fetch((kt + 1) * BK);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float av[8], bv[4];
      float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8 + 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w; av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
      bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      stash(buf ^ 1);
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    int row = m0 + ty * 8 + i;
    if (row >= g.m) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int col = n0 + tx * 4 + j;
      if (col >= g.n) continue;
      float v = acc[i][j] * g.alpha;
      if (g.bias) v += __ldg(g.bias + col);
      if (g.gelu) v = gelu_tanh_exact(v);
      if (g.pos) v += __ldg(g.pos + size_t(row % g.pos_period) * g.n + col);
      float* dst = c + size_t(row) * g.ldc + col;
      *dst = g.accumulate ? *dst + v : v;
    }
  }
}

int launch_sgemm(const SgemmArgs& a, cudaStream_t s) {
  if (a.m <= 0 || a.n <= 0 || a.k <= 0) return YOHO_OK;
  long long nz = (long long)a.nz1 * a.nz2;
  if (nz > 65535) return fail(YOHO_ERR_UNSUPPORTED, "sgemm: %lld batches", nz);
  dim3 grid(ceil_div(a.n, BN), ceil_div(a.m, BM), unsigned(nz));
  if (grid.y > 65535) return fail(YOHO_ERR_UNSUPPORTED, "sgemm: m=%d too large for one launch", a.m);
  ProfScope prof(s, kProfOther, 2.0 * double(nz) * a.m * double(a.n) * a.k);
  sgemm_kernel<<<grid, kSgemmThreads, 0, s>>>(a);
  YOHO_LAUNCH_CHECK("sgemm_kernel");
  return YOHO_OK;
}

// In-place softmax over the last axis, one warp per row (whisper.py:64).
__global__ void __launch_bounds__(256)
softmax_rows_kernel(float* __restrict__ x, long long rows, int cols_all, int causal_period) {
  long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float* p = x + row * cols_all;
  int cols = cols_all;
  if (causal_period > 0) {                      // masked scores are -inf -> probability 0 (whisper.py:59-60, :157)
    cols = min(cols_all, int(row % causal_period) + 1);
    for (int i = cols + lane; i < cols_all; i += 32) p[i] = 0.0f;
  }
  float mx = -INFINITY;
  for (int i = lane; i < cols; i += 32) mx = fmaxf(mx, p[i]);
#pragma unroll
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float sum = 0.0f;
  for (int i = lane; i < cols; i += 32) {
    float e = expf(p[i] - mx);
    p[i] = e;
    sum += e;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float inv = 1.0f / sum;
  for (int i = lane; i < cols; i += 32) p[i] *= inv;
}

int launch_softmax_rows(float* x, long long rows, int cols, cudaStream_t s, int causal_period) {
  if (rows == 0) return YOHO_OK;
  long long blocks = (rows + 7) / 8;
  if (blocks > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "softmax: %lld rows", rows);
  softmax_rows_kernel<<<unsigned(blocks), 256, 0, s>>>(x, rows, cols, causal_period);
  YOHO_LAUNCH_CHECK("softmax_rows_kernel");
  return YOHO_OK;
}

}  // namespace yoho
