// Memory-bound helpers of the encoder: LayerNorm, casts, weight packing / folding.
#include "kernels.cuh"

namespace yoho {

constexpr int kLnMaxVec = 10;  // float4 per lane -> dims <= 1280

__device__ __forceinline__ void store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void store4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}

// One warp per row.  Reference whisper.py:77,82,134 (flax nn.LayerNorm defaults).
template <typename OutT>
__global__ void __launch_bounds__(256, 4)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ scale, const float* __restrict__ bias,
                 OutT* __restrict__ y, int rows, int dims) {
  pdl_launch_dependents();
  pdl_wait();
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int nvec = dims >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + size_t(row) * dims);
  float4 v[kLnMaxVec];
  float s = 0.0f, q = 0.0f;
#pragma unroll
  for (int i = 0; i < kLnMaxVec; ++i) {
    int idx = i * 32 + lane;
    if (idx < nvec) {
      v[i] = xr[idx];
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      q = fmaf(v[i].x, v[i].x, fmaf(v[i].y, v[i].y, fmaf(v[i].z, v[i].z, fmaf(v[i].w, v[i].w, q))));
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    q += __shfl_xor_sync(0xffffffffu, q, o);
  }
  const float mean = s / float(dims);
  const float var = fmaxf(0.0f, q / float(dims) - mean * mean);
  const float rstd = rsqrtf(var + 1e-6f);
  const float4* sc = reinterpret_cast<const float4*>(scale);
  const float4* bi = reinterpret_cast<const float4*>(bias);
  OutT* yr = y + size_t(row) * dims;
#pragma unroll
  for (int i = 0; i < kLnMaxVec; ++i) {
    int idx = i * 32 + lane;
    if (idx < nvec) {
      float4 g = __ldg(sc + idx), b = __ldg(bi + idx), o;
      o.x = fmaf((v[i].x - mean) * rstd, g.x, b.x);
      o.y = fmaf((v[i].y - mean) * rstd, g.y, b.y);
      o.z = fmaf((v[i].z - mean) * rstd, g.z, b.z);
      o.w = fmaf((v[i].w - mean) * rstd, g.w, b.w);
      store4(yr + idx * 4, o);
    }
  }
}

template <typename OutT>
static int launch_layernorm(const float* x, const float* scale, const float* bias, OutT* y, int rows, int dims, cudaStream_t s) {
  if (dims % 4 != 0 || dims > kLnMaxVec * 128) return fail(YOHO_ERR_UNSUPPORTED, "layernorm: dims=%d (need multiple of 4, <= %d)", dims, kLnMaxVec * 128);
  if (rows == 0) return YOHO_OK;
  ProfScope prof(s, kProfLayerNorm, double(rows) * double(dims) * double(sizeof(float) + sizeof(OutT)));
  YOHO_CUDA_CHECK(launch_pdl(layernorm_kernel<OutT>, dim3(ceil_div(rows, 8)), dim3(256), 0, s, x, scale, bias, y, rows, dims));
  YOHO_LAUNCH_CHECK("layernorm_kernel");
  return YOHO_OK;
}
int launch_layernorm_f32(const float* x, const float* sc, const float* b, float* y, int rows, int dims, cudaStream_t s) {
  return launch_layernorm<float>(x, sc, b, y, rows, dims, s);
}
int launch_layernorm_bf16(const float* x, const float* sc, const float* b, __nv_bfloat16* y, int rows, int dims, cudaStream_t s) {
  return launch_layernorm<__nv_bfloat16>(x, sc, b, y, rows, dims, s);
}

// dst[n][k] = bf16(src[k][n]) through a 32x33 shared tile; columns k..ld_dst-1 are zeroed.
__global__ void transpose_cast_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int k, int n, int ld_dst, int k_pad) {
  __shared__ float tile[32][33];
  const int k0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int kk = k0 + j, nn = n0 + threadIdx.x;
    tile[j][threadIdx.x] = (kk < k && nn < n) ? src[size_t(kk) * n + nn] : 0.0f;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    int nn = n0 + j, kk = k0 + threadIdx.x;
    if (nn < n && kk < k_pad) dst[size_t(nn) * ld_dst + kk] = __float2bfloat16_rn(tile[threadIdx.x][j]);
  }
}
int launch_transpose_cast(const float* src, __nv_bfloat16* dst, int k, int n, int ld_dst, cudaStream_t s, int k_pad) {
  if (k_pad < 0) k_pad = ld_dst;
  dim3 grid(ceil_div(n, 32), ceil_div(k_pad, 32));
  ProfScope prof(s, kProfOther, double(k) * n * 6.0);
  transpose_cast_kernel<<<grid, dim3(32, 8), 0, s>>>(src, dst, k, n, ld_dst, k_pad);
  YOHO_LAUNCH_CHECK("transpose_cast_kernel");
  return YOHO_OK;
}

// LayerNorm folded into the GEMM that consumes it (GemmBf16Args::ln_*): one thread per output column n.
//   w'[k] = scale[k] * src[k][n];  dst[n][k] = bf16(w'[k] - mean_k w');  bias_out[n] = bias_in[n] + sum_k beta[k] * src[k][n]
__global__ void __launch_bounds__(128)
ln_fold_kernel(const float* __restrict__ src, const float* __restrict__ scale, const float* __restrict__ beta,
               const float* __restrict__ bias_in, __nv_bfloat16* __restrict__ dst, float* __restrict__ bias_out, int k, int n) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= n) return;
  float c = 0.0f, b = bias_in ? bias_in[col] : 0.0f;
  for (int kk = 0; kk < k; ++kk) {
    const float w = src[size_t(kk) * n + col];
    c = fmaf(scale[kk], w, c);
    b = fmaf(beta[kk], w, b);
  }
  const float m = c / float(k);
  __nv_bfloat16* row = dst + size_t(col) * k;
  for (int kk = 0; kk < k; kk += 2) {                       // k is even (a multiple of 64 in every model)
    const float w0 = src[size_t(kk) * n + col], w1 = src[size_t(kk + 1) * n + col];
    *reinterpret_cast<__nv_bfloat162*>(row + kk) = __floats2bfloat162_rn(fmaf(scale[kk], w0, -m), fmaf(scale[kk + 1], w1, -m));
  }
  bias_out[col] = b;
}
int launch_ln_fold(const float* src, const float* scale, const float* beta, const float* bias_in, __nv_bfloat16* dst,
                   float* bias_out, int k, int n, cudaStream_t s) {
  if (k % 2 != 0) return fail(YOHO_ERR_UNSUPPORTED, "ln_fold: k=%d must be even", k);
  ln_fold_kernel<<<ceil_div(n, 128), 128, 0, s>>>(src, scale, beta, bias_in, dst, bias_out, k, n);
  YOHO_LAUNCH_CHECK("ln_fold_kernel");
  return YOHO_OK;
}

__global__ void cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, size_t n) {
  pdl_launch_dependents();
  pdl_wait();
  size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t step = size_t(gridDim.x) * blockDim.x;
  if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
    const size_t n4 = n >> 2;                                 // 128-bit loads, 64-bit stores
    for (size_t v = i; v < n4; v += step) store4(dst + 4 * v, reinterpret_cast<const float4*>(src)[v]);
    i += n4 << 2;
  }
  for (; i < n; i += step) dst[i] = __float2bfloat16_rn(src[i]);
}
__global__ void cast_f32_kernel(const __nv_bfloat16* __restrict__ src, float* __restrict__ dst, size_t n) {
  size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t step = size_t(gridDim.x) * blockDim.x;
  for (; i < n; i += step) dst[i] = __bfloat162float(src[i]);
}
static int grid_for(size_t n) { size_t g = (n + 255) / 256; return int(g > 148 * 32 ? 148 * 32 : (g ? g : 1)); }
int launch_cast_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t s) {
  ProfScope prof(s, kProfOther, double(n) * 6.0);
  YOHO_CUDA_CHECK(launch_pdl(cast_bf16_kernel, dim3(grid_for(n)), dim3(256), 0, s, src, dst, n));
  YOHO_LAUNCH_CHECK("cast_bf16_kernel");
  return YOHO_OK;
}
int launch_cast_f32(const __nv_bfloat16* src, float* dst, size_t n, cudaStream_t s) {
  ProfScope prof(s, kProfOther, double(n) * 6.0);
  cast_f32_kernel<<<grid_for(n), 256, 0, s>>>(src, dst, n);
  YOHO_LAUNCH_CHECK("cast_f32_kernel");
  return YOHO_OK;
}

int launch_fill_zero(void* p, size_t bytes, cudaStream_t s) {
  cudaError_t e = cudaMemsetAsync(p, 0, bytes, s);
  if (e != cudaSuccess) return fail(YOHO_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  return YOHO_OK;
}

}  // namespace yoho
