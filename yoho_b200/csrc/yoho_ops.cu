// Memory-bound operators of YOHO's own audio encoder (reference yoho/src/nn/model.py:16-29,
// yoho/src/nn/layers.py:7-116): RMSNorm (+SiLU, + time max-pool), RoPE, the SwiGLU gate and the weight
// preparation that folds the query groups of GroupedQueryAttention.
#include "kernels.cuh"

namespace yoho {

constexpr int kRmsMaxVec = 10;  // float4 per lane -> dims <= 1280

__device__ __forceinline__ void rms_store4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void rms_store4(__nv_bfloat16* p, float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&lo);
  u.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }

// One warp per OUTPUT row.  y = [silu](x * rsqrt(mean(x^2) + 1e-6) * scale)      (flax nn.RMSNorm defaults)
// kPool: output row (b, t) = elementwise max of the results for input rows (b, 2t) and (b, 2t+1)
//        (nn.max_pool(x, (1, 2), (1, 2)) on (B, S, D), model.py:20).
template <typename OutT, bool kSilu, bool kPool>
__global__ void __launch_bounds__(256)
rmsnorm_kernel(const float* __restrict__ x, const float* __restrict__ scale, OutT* __restrict__ y, int rows_out, int dims,
               int in_per_clip, int out_per_clip) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows_out) return;
  const int nvec = dims >> 2;
  const float4* sc = reinterpret_cast<const float4*>(scale);
  float4 acc[kRmsMaxVec];
  constexpr int kIn = kPool ? 2 : 1;
#pragma unroll
  for (int r = 0; r < kIn; ++r) {
    size_t in_row = size_t(row);
    if (kPool) { const int b = row / out_per_clip, t = row - b * out_per_clip; in_row = size_t(b) * in_per_clip + 2 * t + r; }
    const float4* xr = reinterpret_cast<const float4*>(x + in_row * dims);
    float4 v[kRmsMaxVec];
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < kRmsMaxVec; ++i) {
      int idx = i * 32 + lane;
      if (idx < nvec) {
        v[i] = xr[idx];
        q = fmaf(v[i].x, v[i].x, fmaf(v[i].y, v[i].y, fmaf(v[i].z, v[i].z, fmaf(v[i].w, v[i].w, q))));
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rinv = rsqrtf(q / float(dims) + 1e-6f);
#pragma unroll
    for (int i = 0; i < kRmsMaxVec; ++i) {
      int idx = i * 32 + lane;
      if (idx < nvec) {
        float4 g = __ldg(sc + idx), o;
        o.x = v[i].x * rinv * g.x; o.y = v[i].y * rinv * g.y; o.z = v[i].z * rinv * g.z; o.w = v[i].w * rinv * g.w;
        if (kSilu) { o.x = silu_f(o.x); o.y = silu_f(o.y); o.z = silu_f(o.z); o.w = silu_f(o.w); }
        if (r == 0) acc[i] = o;
        else { acc[i].x = fmaxf(acc[i].x, o.x); acc[i].y = fmaxf(acc[i].y, o.y); acc[i].z = fmaxf(acc[i].z, o.z); acc[i].w = fmaxf(acc[i].w, o.w); }
      }
    }
  }
  OutT* yr = y + size_t(row) * dims;
#pragma unroll
  for (int i = 0; i < kRmsMaxVec; ++i) {
    int idx = i * 32 + lane;
    if (idx < nvec) rms_store4(yr + idx * 4, acc[i]);
  }
}

template <typename OutT, bool kSilu, bool kPool>
static int launch_rms(const float* x, const float* scale, OutT* y, int rows_out, int dims, int in_per_clip, int out_per_clip, cudaStream_t s) {
  if (dims % 4 != 0 || dims > kRmsMaxVec * 128) return fail(YOHO_ERR_UNSUPPORTED, "rmsnorm: dims=%d (need multiple of 4, <= %d)", dims, kRmsMaxVec * 128);
  if (rows_out == 0) return YOHO_OK;
  ProfScope prof(s, kProfLayerNorm, double(rows_out) * double(dims) * double((kPool ? 2 : 1) * sizeof(float) + sizeof(OutT)));
  rmsnorm_kernel<OutT, kSilu, kPool><<<ceil_div(rows_out, 8), 256, 0, s>>>(x, scale, y, rows_out, dims, in_per_clip, out_per_clip);
  YOHO_LAUNCH_CHECK("rmsnorm_kernel");
  return YOHO_OK;
}
int launch_rmsnorm_f32(const float* x, const float* scale, float* y, int rows, int dims, int silu, cudaStream_t s) {
  return silu ? launch_rms<float, true, false>(x, scale, y, rows, dims, 0, 1, s) : launch_rms<float, false, false>(x, scale, y, rows, dims, 0, 1, s);
}
int launch_rmsnorm_bf16(const float* x, const float* scale, __nv_bfloat16* y, int rows, int dims, int silu, cudaStream_t s) {
  return silu ? launch_rms<__nv_bfloat16, true, false>(x, scale, y, rows, dims, 0, 1, s)
              : launch_rms<__nv_bfloat16, false, false>(x, scale, y, rows, dims, 0, 1, s);
}
int launch_rmsnorm_silu_pool_f32(const float* x, const float* scale, float* y, int batch, int in_len, int dims, cudaStream_t s) {
  return launch_rms<float, true, true>(x, scale, y, batch * (in_len / 2), dims, in_len, in_len / 2, s);
}
int launch_rmsnorm_silu_pool_bf16(const float* x, const float* scale, __nv_bfloat16* y, int batch, int in_len, int dims, cudaStream_t s) {
  return launch_rms<__nv_bfloat16, true, true>(x, scale, y, batch * (in_len / 2), dims, in_len, in_len / 2, s);
}

// RoPE of layers.py:25-43 in place on the first `n_rot_heads` 64-wide heads of every row of `qkv` (row stride ld):
// pair (x[2i], x[2i+1]) of position s is rotated by the angle (s + 1) * 10^-i; `table` holds (len, 2, 32)
// float32 {cos, sin} built by the host in the reference's float32 operation order.
__device__ __forceinline__ float2 ld2(const float* p) { return *reinterpret_cast<const float2*>(p); }
__device__ __forceinline__ float2 ld2(const __nv_bfloat16* p) { return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p)); }
__device__ __forceinline__ void st2(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
__device__ __forceinline__ void st2(__nv_bfloat16* p, float2 v) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(v.x, v.y); }

template <typename T>
__global__ void __launch_bounds__(256)
rope_kernel(T* __restrict__ qkv, const float* __restrict__ table, long long rows, int seq, int ld, int n_rot_heads) {
  const int pairs_per_row = n_rot_heads * 32;
  const long long total = rows * pairs_per_row;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / pairs_per_row;
    const int pr = int(i - row * pairs_per_row), head = pr >> 5, j = pr & 31;
    const int pos = int(row % seq);
    const float c = __ldg(table + size_t(pos) * 64 + j), sn = __ldg(table + size_t(pos) * 64 + 32 + j);
    T* p = qkv + size_t(row) * ld + head * 64 + 2 * j;
    const float2 v = ld2(p);
    st2(p, make_float2(v.x * c - v.y * sn, v.x * sn + v.y * c));
  }
}
template <typename T>
static int launch_rope(T* qkv, const float* table, long long rows, int seq, int ld, int n_rot_heads, cudaStream_t s) {
  if (rows == 0 || n_rot_heads == 0) return YOHO_OK;
  const long long total = rows * n_rot_heads * 32;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ProfScope prof(s, kProfOther, double(total) * 4.0 * sizeof(T));
  rope_kernel<T><<<int(blocks), 256, 0, s>>>(qkv, table, rows, seq, ld, n_rot_heads);
  YOHO_LAUNCH_CHECK("rope_kernel");
  return YOHO_OK;
}
int launch_rope_f32(float* qkv, const float* table, long long rows, int seq, int ld, int n_rot_heads, cudaStream_t s) {
  return launch_rope<float>(qkv, table, rows, seq, ld, n_rot_heads, s);
}
int launch_rope_bf16(__nv_bfloat16* qkv, const float* table, long long rows, int seq, int ld, int n_rot_heads, cudaStream_t s) {
  return launch_rope<__nv_bfloat16>(qkv, table, rows, seq, ld, n_rot_heads, s);
}

// SwiGLU gate (layers.py:20): g[r][c] = silu(h[r][c]) * h[r][hidden + c], h is (rows, 2 * hidden) = [w1 x | w2 x].
template <typename T>
__global__ void __launch_bounds__(256)
swiglu_gate_kernel(const T* __restrict__ h, T* __restrict__ g, long long rows, int hidden) {
  const int half = hidden >> 1;
  const long long total = rows * half;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / half;
    const int c = int(i - row * half) * 2;
    const float2 a = ld2(h + size_t(row) * 2 * hidden + c), b = ld2(h + size_t(row) * 2 * hidden + hidden + c);
    st2(g + size_t(row) * hidden + c, make_float2(silu_f(a.x) * b.x, silu_f(a.y) * b.y));
  }
}
template <typename T>
static int launch_swiglu(const T* h, T* g, long long rows, int hidden, cudaStream_t s) {
  if (hidden % 2 != 0) return fail(YOHO_ERR_UNSUPPORTED, "swiglu: hidden=%d must be even", hidden);
  if (rows == 0) return YOHO_OK;
  long long blocks = (rows * (hidden / 2) + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  ProfScope prof(s, kProfOther, double(rows) * hidden * 3.0 * sizeof(T));
  swiglu_gate_kernel<T><<<int(blocks), 256, 0, s>>>(h, g, rows, hidden);
  YOHO_LAUNCH_CHECK("swiglu_gate_kernel");
  return YOHO_OK;
}
int launch_swiglu_gate_f32(const float* h, float* g, long long rows, int hidden, cudaStream_t s) { return launch_swiglu<float>(h, g, rows, hidden, s); }
int launch_swiglu_gate_bf16(const __nv_bfloat16* h, __nv_bfloat16* g, long long rows, int hidden, cudaStream_t s) {
  return launch_swiglu<__nv_bfloat16>(h, g, rows, hidden, s);
}

// dst (d_in, kv_heads * 64; row stride ld_dst) = sum over the query group of src (d_in, kv_heads, groups, 64).  The einsum at
// layers.py:77 ("g h s d, h a d -> h s a") sums the scores over g, and RoPE is the same linear map for every
// head, so attention sees only the group sum of the query projections.
__global__ void group_sum_kernel(const float* __restrict__ src, float* __restrict__ dst, int d_in, int kv_heads, int groups, int ld_dst) {
  const int total = d_in * kv_heads * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int d = i & 63, h = (i >> 6) % kv_heads, r = i / (kv_heads * 64);
    float acc = 0.0f;
    for (int g = 0; g < groups; ++g) acc += src[((size_t(r) * kv_heads + h) * groups + g) * 64 + d];
    dst[size_t(r) * ld_dst + h * 64 + d] = acc;
  }
}
int launch_group_sum(const float* src, float* dst, int d_in, int kv_heads, int groups, int ld_dst, cudaStream_t s) {
  group_sum_kernel<<<ceil_div(d_in * kv_heads * 64, 256), 256, 0, s>>>(src, dst, d_in, kv_heads, groups, ld_dst);
  YOHO_LAUNCH_CHECK("group_sum_kernel");
  return YOHO_OK;
}

}  // namespace yoho
