// Kernels of the Whisper text decoder's token-by-token path (reference yoho/src/nn/whisper.py:138-165 run with a
// KV cache instead of the reference's full re-run per token, :27 / :274-281).  One generated token = M = batch
// activation rows against every decoder weight, so each contraction is a weight-streaming "skinny" GEMM: the job
// is to put the whole GPU's load/store units on a 1-4 MB weight matrix for a couple of microseconds, not to feed
// tensor cores.  The decode position lives in device memory so that one launch sequence (or one CUDA graph) serves
// every token.
#include "kernels.cuh"

#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace yoho {

__device__ __forceinline__ float dec_gelu_tanh(float x) {
  const float c = 0.7978845608028654f;
  return 0.5f * x * (1.0f + tanhf(c * (x + 0.044715f * x * x * x)));
}

// ---- x[b, s, :] = embedding[token[b, p0 + s]] + positional_embedding[p0 + s]            whisper.py:153-155
__global__ void __launch_bounds__(128)
embed_kernel(const int* __restrict__ tokens, int tok_ld, const int* __restrict__ pos_ptr, int s_len,
             const float* __restrict__ emb, const float* __restrict__ pos, float* __restrict__ x, int dims, int vocab) {
  const int s = blockIdx.x, b = blockIdx.y;
  const int p0 = pos_ptr ? *pos_ptr : 0;     // < tok_ld: advance_kernel saturates the counter
  int tok = tokens[size_t(b) * tok_ld + p0 + s];
  tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);
  const float4* e = reinterpret_cast<const float4*>(emb + size_t(tok) * dims);
  const float4* p = reinterpret_cast<const float4*>(pos + size_t(p0 + s) * dims);
  float4* o = reinterpret_cast<float4*>(x + (size_t(b) * s_len + s) * dims);
  for (int i = threadIdx.x; i < dims / 4; i += blockDim.x) {
    float4 a = __ldg(e + i), c = __ldg(p + i);
    o[i] = make_float4(a.x + c.x, a.y + c.y, a.z + c.z, a.w + c.w);
  }
}

int launch_embed(const int* tokens, int tok_ld, const int* pos_ptr, int batch, int s_len, const float* emb, const float* pos,
                 float* x, int dims, int vocab, cudaStream_t s) {
  if (batch <= 0 || s_len <= 0) return YOHO_OK;
  if (dims % 4 != 0 || batch > 65535) return fail(YOHO_ERR_UNSUPPORTED, "embed: dims=%d batch=%d", dims, batch);
  embed_kernel<<<dim3(s_len, batch), 128, 0, s>>>(tokens, tok_ld, pos_ptr, s_len, emb, pos, x, dims, vocab);
  YOHO_LAUNCH_CHECK("embed_kernel");
  return YOHO_OK;
}

// ---- skinny GEMM: out[b, n] (+)= act(sum_k x[b, k] w[k, n] + bias[n]), b < BT <= 64 ------------------------------
// CTA = 32 output columns x the k-range of its cluster rank; its 8 warps take interleaved weight rows (one 128-byte
// row segment per warp-load, eight loads in flight per warp), the activations of the k-range sit in shared memory
// (k-major, 128-bit broadcast reads); partial sums are reduced across warps in shared memory and across the KS CTAs of
// a thread-block cluster through distributed shared memory (deterministic: no atomics).
constexpr int kSkThreads = 256, kSkWarps = 8, kSkCols = 32, kSkChunk = 256;

template <int BT, int KS>
__global__ void __launch_bounds__(kSkThreads)
skinny_gemm_kernel(const SkinnyArgs a) {
  extern __shared__ __align__(16) float sk_smem[];
  float* xs = sk_smem;                                  // [kSkChunk][BT] (4-float groups xor-swizzled by k), later red[8][BT][32]
  float* part = sk_smem + (kSkChunk * BT > kSkWarps * BT * 32 ? kSkChunk * BT : kSkWarps * BT * 32);   // [BT][32]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rank = KS > 1 ? int(blockIdx.y) : 0;
  const int col = blockIdx.x * kSkCols + lane;
  const bool col_ok = col < a.n;
  const int k_per = (a.k + KS - 1) / KS;
  const int k_begin = rank * k_per, k_end = min(a.k, k_begin + k_per);
  constexpr int G = BT / 4;                             // 4-float groups per k row

  float acc[BT];
#pragma unroll
  for (int b = 0; b < BT; ++b) acc[b] = 0.0f;

  // the first eight weight rows of every warp do not depend on the activations: pull their lines towards the SM before
  // the LayerNorm statistics and the activation staging (a step is a chain of such latencies, not bandwidth)
  {
    const int len0 = min(kSkChunk, k_end - k_begin);
    const float* wp0 = a.w + size_t(k_begin) * a.ldw + col;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int kk = warp + u * kSkWarps;
      if (col_ok && kk < len0) asm volatile("prefetch.global.L1 [%0];" ::"l"(wp0 + size_t(kk) * a.ldw));
    }
  }

  // fused LayerNorm: every CTA recomputes the statistics of the (few) activation rows over the full k
  __shared__ float ln_mean[BT], ln_rstd[BT];
  if (a.ln_scale) {
    for (int b = warp; b < a.b; b += kSkWarps) {
      const float* row = a.x + size_t(b) * a.ldx;
      float s1 = 0.0f, s2 = 0.0f;
      for (int k = lane; k < a.k; k += 32) { const float v = __ldg(row + k); s1 += v; s2 = fmaf(v, v, s2); }
#pragma unroll
      for (int o = 16; o; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
      if (lane == 0) {
        const float mean = s1 / a.k;
        ln_mean[b] = mean;
        ln_rstd[b] = rsqrtf(fmaxf(s2 / a.k - mean * mean, 0.0f) + 1e-6f);
      }
    }
  }

  for (int k0 = k_begin; k0 < k_end; k0 += kSkChunk) {
    const int len = min(kSkChunk, k_end - k0);
    __syncthreads();
    // rows of the chunk the warps will touch: len rounded up to the 64-row step of the loop below
    const int len_pad = min(kSkChunk, (len + 8 * kSkWarps - 1) / (8 * kSkWarps) * (8 * kSkWarps));
    for (int idx = tid; idx < BT * len_pad; idx += kSkThreads) {
      const int b = idx / len_pad, kk = idx - b * len_pad;
      float v = 0.0f;
      if (b < a.b && kk < len) {
        v = __ldg(a.x + size_t(b) * a.ldx + k0 + kk);
        if (a.ln_scale) v = fmaf((v - ln_mean[b]) * ln_rstd[b], __ldg(a.ln_scale + k0 + kk), __ldg(a.ln_bias + k0 + kk));
      }
      xs[kk * BT + (((b >> 2) ^ (kk & (G - 1))) << 2) + (b & 3)] = v;
    }
    __syncthreads();
    const float* wp = a.w + size_t(k0) * a.ldw + col;
    for (int kk0 = warp; kk0 < len; kk0 += 8 * kSkWarps) {
      float wv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int kk = kk0 + u * kSkWarps;
        wv[u] = (col_ok && kk < len) ? __ldg(wp + size_t(kk) * a.ldw) : 0.0f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int kk = kk0 + u * kSkWarps;       // < kSkChunk by construction
        const float4* xr = reinterpret_cast<const float4*>(xs + kk * BT);
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const float4 xv = xr[g ^ (kk & (G - 1))];
          acc[4 * g + 0] = fmaf(xv.x, wv[u], acc[4 * g + 0]);
          acc[4 * g + 1] = fmaf(xv.y, wv[u], acc[4 * g + 1]);
          acc[4 * g + 2] = fmaf(xv.z, wv[u], acc[4 * g + 2]);
          acc[4 * g + 3] = fmaf(xv.w, wv[u], acc[4 * g + 3]);
        }
      }
    }
  }
  __syncthreads();
  float* red = xs;
#pragma unroll
  for (int b = 0; b < BT; ++b) red[(warp * BT + b) * 32 + lane] = acc[b];
  __syncthreads();
  for (int o = tid; o < BT * 32; o += kSkThreads) {
    float v = 0.0f;
#pragma unroll
    for (int w = 0; w < kSkWarps; ++w) v += red[w * BT * 32 + o];
    part[o] = v;
  }
  // every CTA of the cluster finishes its own share of the BT x 32 outputs: partials of the peers come through
  // distributed shared memory, then bias / GELU / accumulate / store
  constexpr int kOutPer = BT * 32 / KS;
  if (KS > 1) cg::this_cluster().sync();
  else __syncthreads();
  const long long row_off = a.row_off_ptr ? (long long)(*a.row_off_ptr) * a.row_off_mul : 0;
  for (int i = tid; i < kOutPer; i += kSkThreads) {
    const int o = rank * kOutPer + i;
    const int b = o >> 5, c = blockIdx.x * kSkCols + (o & 31);
    float v = 0.0f;
    if (KS > 1) {
      cg::cluster_group cluster = cg::this_cluster();
#pragma unroll
      for (int r = 0; r < KS; ++r) v += cluster.map_shared_rank(part, r)[o];
    } else {
      v = part[o];
    }
    if (b >= a.b || c >= a.n) continue;
    if (a.bias) v += __ldg(a.bias + c);
    if (a.gelu) v = dec_gelu_tanh(v);
    float* dst;
    if (a.n_split > 0) dst = c < a.n_split ? a.out + (long long)b * a.out_row_stride + c
                                           : a.out2 + (long long)b * a.out2_row_stride + row_off + (c - a.n_split);
    else dst = a.out + (long long)b * a.out_row_stride + row_off + c;
    *dst = a.accumulate ? *dst + v : v;
  }
  if (KS > 1) cg::this_cluster().sync();          // peers keep their shared memory until everyone has read it
}

template <int BT, int KS>
static int skinny_launch(const SkinnyArgs& a, cudaStream_t s) {
  const size_t main_floats = size_t(kSkChunk) * BT > size_t(kSkWarps) * BT * 32 ? size_t(kSkChunk) * BT : size_t(kSkWarps) * BT * 32;
  const size_t smem = (main_floats + size_t(BT) * 32) * sizeof(float);
  static thread_local int attr_device = -1;      // the opt-in is per device; set it once per host thread and device
  int dev = 0;
  YOHO_CUDA_CHECK(cudaGetDevice(&dev));
  if (attr_device != dev) {
    YOHO_CUDA_CHECK(cudaFuncSetAttribute(skinny_gemm_kernel<BT, KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    attr_device = dev;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ceil_div(a.n, kSkCols), KS, 1);
  cfg.blockDim = dim3(kSkThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = KS; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  // No cluster attribute for the variant without a k split: launched as 1 x 1 x 1 clusters, two of those instantiations
  // computed their shared-memory addresses from SR_CgaCtaId and faulted ("operation not supported on global/shared
  // address space") or returned garbage; as plain launches all of them pass.
  cfg.numAttrs = KS > 1 ? 1 : 0;
  YOHO_CUDA_CHECK(cudaLaunchKernelEx(&cfg, skinny_gemm_kernel<BT, KS>, a));
  YOHO_LAUNCH_CHECK("skinny_gemm_kernel");
  return YOHO_OK;
}

int launch_skinny_gemm(const SkinnyArgs& a, cudaStream_t s) {
  if (a.b <= 0 || a.n <= 0 || a.k <= 0) return YOHO_OK;
  if (a.b > 64) return fail(YOHO_ERR_UNSUPPORTED, "skinny gemm: %d rows (<= 64 per call)", a.b);
  // enough column blocks to fill the GPU on their own (the vocabulary projection): no k split
  const bool wide = ceil_div(a.n, kSkCols) >= 296;
  ProfScope prof(s, kProfDecGemm, (double(a.n) * a.k + double(a.b) * (a.n + a.k)) * sizeof(float));
#define YOHO_SK(BT)                                                          \
  return wide ? skinny_launch<BT, 1>(a, s) : skinny_launch<BT, 8>(a, s)
  if (a.b <= 4) { YOHO_SK(4); }
  if (a.b <= 8) { YOHO_SK(8); }
  if (a.b <= 16) { YOHO_SK(16); }
  if (a.b <= 32) { YOHO_SK(32); }
  YOHO_SK(64);
#undef YOHO_SK
}

// ---- single-query attention against a key/value cache --------------------------------------------------------------
// One query row per (head, clip).  scores = (q . k_j) / 8 (q and k each carry 64^-0.25, whisper.py:54-57), softmax over
// the first `len` cache rows, out = sum_j p_j v_j.  `len` comes from device memory for the growing self-attention cache
// (position + 1) or is the constant number of encoder positions for the cross attention.
// The cache rows of one (head, clip) are split over the kDaSplit CTAs of a thread-block cluster ("flash decoding"): one
// clip alone still puts heads x 8 CTAs on the memory system, every CTA keeps all of its <= 256 rows' loads in flight, and
// the partial (max, sum, out) triples are merged by rank 0 through distributed shared memory.
constexpr int kDaThreads = 256, kDaSplit = 8, kDaLocal = 256;

__global__ void __launch_bounds__(kDaThreads)
decode_attention_kernel(const DecodeAttArgs a) {
  __shared__ float sc[kDaLocal];
  __shared__ float red[kDaThreads / 32];
  __shared__ __align__(16) float part[16][64];
  __shared__ float loc[66];                      // this CTA's {out[64], max, sum}
  cg::cluster_group cluster = cg::this_cluster();
  const int h = blockIdx.x, b = blockIdx.y, rank = blockIdx.z;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int len = a.len_ptr ? *a.len_ptr + a.len_add : a.len_const;
  const int per = (len + kDaSplit - 1) / kDaSplit;
  const int j_begin = rank * per;
  const int n = max(0, min(len, j_begin + per) - j_begin);          // <= kDaLocal (launcher)
  const float* kv = a.kv + (long long)b * a.kv_clip_stride + size_t(j_begin) * a.kv_ld + h * 64;

  // scores: 8 lanes per cache row (8 dims each), 4 rows per warp-load, every load of the CTA's slice in flight at once
  const int sub = lane & 7, kq = lane >> 3;
  const float4* qp = reinterpret_cast<const float4*>(a.q + size_t(b) * a.ldq + h * 64 + sub * 8);
  const float4 q0 = __ldg(qp), q1 = __ldg(qp + 1);
  float d[kDaLocal / 32];
#pragma unroll
  for (int i = 0; i < kDaLocal / 32; ++i) {
    const int j = i * 32 + warp * 4 + kq;
    d[i] = 0.0f;
    if (j < n) {
      const float4* kp = reinterpret_cast<const float4*>(kv + size_t(j) * a.kv_ld + a.k_off + sub * 8);
      const float4 k0 = __ldg(kp), k1 = __ldg(kp + 1);
      d[i] = (q0.x * k0.x + q0.y * k0.y + q0.z * k0.z + q0.w * k0.w) + (q1.x * k1.x + q1.y * k1.y + q1.z * k1.z + q1.w * k1.w);
    }
  }
#pragma unroll
  for (int i = 0; i < kDaLocal / 32; ++i) {
    d[i] += __shfl_xor_sync(0xffffffffu, d[i], 1);
    d[i] += __shfl_xor_sync(0xffffffffu, d[i], 2);
    d[i] += __shfl_xor_sync(0xffffffffu, d[i], 4);
    const int j = i * 32 + warp * 4 + kq;
    if (sub == 0 && j < n) sc[j] = d[i] * 0.125f;
  }
  __syncthreads();
  float mx = tid < n ? sc[tid] : -INFINITY;
#pragma unroll
  for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
#pragma unroll
  for (int w = 1; w < kDaThreads / 32; ++w) mx = fmaxf(mx, red[w]);
  __syncthreads();
  float e = 0.0f;
  if (tid < n) { e = expf(sc[tid] - mx); sc[tid] = e; }
  float sum = e;
#pragma unroll
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) red[warp] = sum;
  __syncthreads();
  sum = 0.0f;
#pragma unroll
  for (int w = 0; w < kDaThreads / 32; ++w) sum += red[w];
  // out: 16 threads per cache row (one float4 of the 64 dims each), 16 rows per pass
  {
    const int slice = tid >> 4, d4 = tid & 15;
    const float* vp = kv + a.v_off + d4 * 4;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
    for (int i = 0; i < kDaLocal / 16; ++i) {
      const int j = i * 16 + slice;
      if (j < n) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(vp + size_t(j) * a.kv_ld));
        const float p = sc[j];
        acc.x = fmaf(p, v.x, acc.x); acc.y = fmaf(p, v.y, acc.y); acc.z = fmaf(p, v.z, acc.z); acc.w = fmaf(p, v.w, acc.w);
      }
    }
    *reinterpret_cast<float4*>(&part[slice][d4 * 4]) = acc;
  }
  __syncthreads();
  if (tid < 64) {
    float o = 0.0f;
#pragma unroll
    for (int s2 = 0; s2 < 16; ++s2) o += part[s2][tid];
    loc[tid] = o;
  }
  if (tid == 0) { loc[64] = mx; loc[65] = sum; }
  cluster.sync();
  if (rank == 0 && tid < 64) {
    float m = -INFINITY;
    for (int r = 0; r < kDaSplit; ++r) m = fmaxf(m, cluster.map_shared_rank(loc, r)[64]);
    float tot = 0.0f, o = 0.0f;
    for (int r = 0; r < kDaSplit; ++r) {
      const float* rl = cluster.map_shared_rank(loc, r);
      const float wgt = expf(rl[64] - m);        // 0 for an empty slice (max = -inf, sum = 0)
      tot = fmaf(rl[65], wgt, tot);
      o = fmaf(rl[tid], wgt, o);
    }
    a.out[size_t(b) * a.ldo + h * 64 + tid] = o / tot;
  }
  cluster.sync();          // peers keep their shared memory until rank 0 has merged
}

int launch_decode_attention(const DecodeAttArgs& a, int batch, int heads, int max_len, cudaStream_t s) {
  if (batch <= 0 || heads <= 0) return YOHO_OK;
  if (batch > 65535) return fail(YOHO_ERR_UNSUPPORTED, "decode attention: batch=%d", batch);
  if (max_len > kDaSplit * kDaLocal) return fail(YOHO_ERR_UNSUPPORTED, "decode attention: %d cache rows (<= %d)", max_len, kDaSplit * kDaLocal);
  if (a.kv_ld % 4 != 0 || a.k_off % 4 != 0 || a.v_off % 4 != 0 || a.ldq % 4 != 0) return fail(YOHO_ERR_UNSUPPORTED, "decode attention: strides must be multiples of 4 floats");
  ProfScope prof(s, kProfDecAttention, double(batch) * heads * max_len * 2.0 * 64 * sizeof(float));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(heads, batch, kDaSplit);
  cfg.blockDim = dim3(kDaThreads, 1, 1);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = kDaSplit;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  YOHO_CUDA_CHECK(cudaLaunchKernelEx(&cfg, decode_attention_kernel, a));
  YOHO_LAUNCH_CHECK("decode_attention_kernel");
  return YOHO_OK;
}

// ---- greedy choice: tokens[b, p + 1] = argmax_v logits[b, v] unless that position is forced ---------------------------
// First maximum wins (jnp.argmax, whisper.py:279).  A sequence that has produced `eot` keeps producing it.  The last CTA
// to finish advances the position; the counter saturates at the last slot of the token buffer, so a step past the end
// rewrites that slot instead of running off the caches.
__global__ void __launch_bounds__(256)
greedy_pick_kernel(const float* __restrict__ logits, int ld, int vocab, int* __restrict__ tokens, int tok_ld, int max_len,
                   int* __restrict__ pos_ptr, const unsigned char* __restrict__ forced, int eot) {
  __shared__ float bv[8];
  __shared__ int bi[8];
  const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int p = pos_ptr[0];
  const bool write = p + 1 < max_len && !(forced && forced[p + 1]);
  float best = -INFINITY;
  int idx = 0x7fffffff;
  if (write) {
    const float* row = logits + size_t(b) * ld;
    for (int v = tid; v < vocab; v += 256) {
      const float x = row[v];
      if (x > best) { best = x; idx = v; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
    }
    if (lane == 0) { bv[warp] = best; bi[warp] = idx; }
  }
  __syncthreads();
  if (tid == 0) {
    if (write) {
      for (int w = 1; w < 8; ++w)
        if (bv[w] > best || (bv[w] == best && bi[w] < idx)) { best = bv[w]; idx = bi[w]; }
      int* t = tokens + size_t(b) * tok_ld;
      t[p + 1] = (eot >= 0 && t[p] == eot) ? eot : idx;
    }
    __threadfence();
    if (atomicAdd(pos_ptr + 1, 1) == int(gridDim.x) - 1) {      // every CTA has read p by the time the last one gets here
      pos_ptr[1] = 0;
      pos_ptr[0] = min(p + 1, max_len - 1);
    }
  }
}

int launch_greedy_pick(const float* logits, int ld, int vocab, int* tokens, int tok_ld, int max_len, int batch, int* pos_ptr,
                       const unsigned char* forced, int eot, cudaStream_t s) {
  if (batch <= 0) return YOHO_OK;
  greedy_pick_kernel<<<batch, 256, 0, s>>>(logits, ld, vocab, tokens, tok_ld, max_len, pos_ptr, forced, eot);
  YOHO_LAUNCH_CHECK("greedy_pick_kernel");
  return YOHO_OK;
}

}  // namespace yoho
