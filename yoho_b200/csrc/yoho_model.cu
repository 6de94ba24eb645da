// YOHO's own audio encoder forward (reference yoho/src/nn/model.py:9-29 with the blocks of
// yoho/src/nn/layers.py:7-116; what Model.encode_audio runs, model.py:82-83) behind the C ABI.
//
//   x (B, S, M) -> Conv(k3, SAME) -> RMSNorm -> SiLU -> max-pool over time pairs -> Conv(k3) -> RMSNorm -> SiLU
//     -> n_blocks x [ x += GQA(RMSNorm(x)) ; x += SwiGLU(RMSNorm(x)) ]                    -> (B, S / 2, D)
//
// GroupedQueryAttention as written in the reference sums the scores over the query group (the einsum at
// layers.py:77 drops g), and RoPE is the same linear map for every head, so the attention is exactly
// kv_heads-head attention with q = rope(x (sum_g Wq[h, g])): the group sum is folded into the packed weight.
// The dense contractions reuse the tcgen05 GEMM (3-tap implicit GEMM for both convolutions, the zero padding
// is TMA out-of-bounds fill) and the tcgen05 attention kernel; RMSNorm / SiLU / pooling / RoPE / the SwiGLU gate
// are the memory-bound kernels of yoho_ops.cu.
#include "kernels.cuh"

#include <vector>

namespace yoho {

constexpr int kYmFixed = 7, kYmPerBlock = 9;

struct YmBlockOff {
  size_t n0, wq, wk, wv, wo, n1, w1, w2, w3;   // float32 originals (flax layouts)
  size_t wqkv;                                 // float32 (D, 3 kvd): [sum_g wq | wk | wv]
  size_t wqkv_h, wo_h, w12_h, w3_h;            // bf16 (out, in)
};
struct YmBlob {
  size_t conv0_w, conv0_b, n0, conv1_w, conv1_b, n1, rope;
  size_t conv0_w_h, conv1_w_h;
  std::vector<YmBlockOff> blocks;
  size_t total = 0;
};

static int ym_kvd(const yoho_model_desc& d) { return d.kv_heads * 64; }
static int ym_hidden(const yoho_model_desc& d) { return 3 * d.dims; }   // int(dims * 3), layers.py:113

static size_t ym_tensor_elems(const yoho_model_desc& d, int idx) {
  const size_t D = d.dims, M = d.n_mels, kvd = ym_kvd(d), H = ym_hidden(d);
  if (idx < kYmFixed) {
    switch (idx) {
      case 0: return 3 * M * D;
      case 3: return 3 * D * D;
      case 6: return size_t(d.seq_len / 2) * 64;
      default: return D;
    }
  }
  switch ((idx - kYmFixed) % kYmPerBlock) {
    case 1: return D * size_t(d.q_heads) * 64;
    case 2: case 3: return D * kvd;
    case 4: return kvd * D;
    case 6: case 7: return D * H;
    case 8: return H * D;
    default: return D;
  }
}

static YmBlob ym_plan_blob(const yoho_model_desc& d) {
  YmBlob L;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 256); size_t o = off; off += bytes; return o; };
  const size_t D = d.dims, kvd = ym_kvd(d), H = ym_hidden(d), f = sizeof(float), h = sizeof(__nv_bfloat16);
  L.conv0_w = take(ym_tensor_elems(d, 0) * f); L.conv0_b = take(D * f); L.n0 = take(D * f);
  L.conv1_w = take(ym_tensor_elems(d, 3) * f); L.conv1_b = take(D * f); L.n1 = take(D * f);
  L.rope = take(ym_tensor_elems(d, 6) * f);
  L.blocks.resize(d.n_blocks);
  for (auto& b : L.blocks) {
    b.n0 = take(D * f); b.wq = take(D * size_t(d.q_heads) * 64 * f); b.wk = take(D * kvd * f); b.wv = take(D * kvd * f);
    b.wo = take(kvd * D * f); b.n1 = take(D * f); b.w1 = take(D * H * f); b.w2 = take(D * H * f); b.w3 = take(H * D * f);
  }
  L.conv0_w_h = take(D * 3 * size_t(d.n_mels) * h);
  L.conv1_w_h = take(D * 3 * D * h);
  for (auto& b : L.blocks) {
    b.wqkv = take(D * 3 * kvd * f);
    b.wqkv_h = take(3 * kvd * D * h);
    b.wo_h = take(D * kvd * h);
    b.w12_h = take(2 * H * D * h);
    b.w3_h = take(D * H * h);
  }
  L.total = align_up(off, 1024);
  return L;
}

static int ym_check_desc(const yoho_model_desc* d, const char* who) {
  if (!d) return fail(YOHO_ERR_INVALID, "%s: desc is null", who);
  if (d->n_mels <= 0 || d->dims <= 0 || d->q_heads <= 0 || d->kv_heads <= 0 || d->n_blocks < 0 || d->seq_len <= 1)
    return fail(YOHO_ERR_INVALID, "%s: desc has a non-positive field", who);
  if (d->q_heads % d->kv_heads != 0)
    return fail(YOHO_ERR_INVALID, "%s: q_heads=%d is not a multiple of kv_heads=%d (layers.py:61)", who, d->q_heads, d->kv_heads);
  if (d->dims != d->q_heads * 64)
    return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d, q_heads=%d: kernels need head_dim = 64 (the reference config: 384 / 6)", who, d->dims, d->q_heads);
  if (d->n_mels % 64 != 0) return fail(YOHO_ERR_UNSUPPORTED, "%s: n_mels=%d must be a multiple of 64 (the reference config: 128)", who, d->n_mels);
  if (d->dims > 1280) return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d > 1280", who, d->dims);
  return YOHO_OK;
}

struct YmWs {
  size_t melh = 0, c = 0, p0 = 0, x = 0, xn = 0, qkv = 0, att = 0, h = 0, g = 0, scores = 0;
  int score_clips = 0;
  size_t total = 0;
};

static YmWs ym_plan_ws(const yoho_model_desc& d, int batch, int frames, int compute) {
  YmWs w;
  const size_t B = batch, S = frames, S2 = frames / 2, D = d.dims, kvd = ym_kvd(d), H = ym_hidden(d);
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 1024); size_t o = off; off += bytes; return o; };
  const size_t e = compute == YOHO_COMPUTE_BF16 ? sizeof(__nv_bfloat16) : sizeof(float);
  if (compute == YOHO_COMPUTE_BF16) w.melh = take(B * S * d.n_mels * e);
  w.c = take(B * S * D * sizeof(float));       // conv outputs (Conv_0: B*S rows, Conv_1: B*S2 rows)
  w.p0 = take(B * S2 * D * e);                 // pooled stem activations
  w.x = take(B * S2 * D * sizeof(float));      // residual stream
  w.xn = take(B * S2 * D * e);
  w.qkv = take(B * S2 * 3 * kvd * e);
  w.att = take(B * S2 * kvd * e);
  w.h = take(B * S2 * 2 * H * e);
  w.g = take(B * S2 * H * e);
  if (compute != YOHO_COMPUTE_BF16) {
    const size_t per_clip = size_t(d.kv_heads) * S2 * S2 * sizeof(float);
    size_t clips = (size_t(1) << 30) / (per_clip ? per_clip : 1);
    if (clips < 1) clips = 1;
    if (clips > B) clips = B;
    if (clips * d.kv_heads > 65535) clips = 65535 / d.kv_heads;
    w.score_clips = int(clips);
    w.scores = take(clips * per_clip);
  }
  w.total = align_up(off, 1024);
  return w;
}

#define YOHO_TRY(expr) do { int _r = (expr); if (_r != YOHO_OK) return _r; } while (0)

static int ym_forward_bf16(yoho_ctx* ctx, const yoho_model_desc& d, const YmBlob& L, const char* blob, const float* mel, int batch,
                           int frames, float* out, char* ws, const YmWs& w, cudaStream_t s) {
  using bf = __nv_bfloat16;
  const int S = frames, S2 = frames / 2, D = d.dims, M = d.n_mels, kvd = ym_kvd(d), H = ym_hidden(d);
  const int rows = batch * S2;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  auto H16 = [&](size_t off) { return reinterpret_cast<const bf*>(blob + off); };
  bf* melh = reinterpret_cast<bf*>(ws + w.melh);
  float* c = reinterpret_cast<float*>(ws + w.c);
  bf* p0 = reinterpret_cast<bf*>(ws + w.p0);
  float* x = reinterpret_cast<float*>(ws + w.x);
  bf* xn = reinterpret_cast<bf*>(ws + w.xn);
  bf* qkv = reinterpret_cast<bf*>(ws + w.qkv);
  bf* att = reinterpret_cast<bf*>(ws + w.att);
  bf* hbuf = reinterpret_cast<bf*>(ws + w.h);
  bf* gbuf = reinterpret_cast<bf*>(ws + w.g);

  // Conv_0 (model.py:17): three shifted TMA boxes over (B, S, M); rows -1 and S are out of bounds = the SAME padding
  YOHO_TRY(launch_cast_bf16(mel, melh, size_t(batch) * S * M, s));
  {
    GemmBf16Args g;
    g.a = melh; g.a_ld = M; g.a_cols = M; g.batch = batch; g.rows_per_batch = S;
    g.w = H16(L.conv0_w_h); g.n = D; g.k_per_tap = M; g.n_taps = 3;
    g.taps[0] = {0, -1}; g.taps[1] = {0, 0}; g.taps[2] = {0, 1};
    g.out_f32 = c; g.ldc = D; g.bias = F32(L.conv0_b);
    YOHO_TRY(launch_gemm_bf16(ctx, g, s));
  }
  // RMSNorm -> SiLU -> max_pool (1, 2) (model.py:18-20)
  YOHO_TRY(launch_rmsnorm_silu_pool_bf16(c, F32(L.n0), p0, batch, S, D, s));
  // Conv_1 -> RMSNorm -> SiLU (model.py:22-24)
  {
    GemmBf16Args g;
    g.a = p0; g.a_ld = D; g.a_cols = D; g.batch = batch; g.rows_per_batch = S2;
    g.w = H16(L.conv1_w_h); g.n = D; g.k_per_tap = D; g.n_taps = 3;
    g.taps[0] = {0, -1}; g.taps[1] = {0, 0}; g.taps[2] = {0, 1};
    g.out_f32 = c; g.ldc = D; g.bias = F32(L.conv1_b);
    YOHO_TRY(launch_gemm_bf16(ctx, g, s));
  }
  YOHO_TRY(launch_rmsnorm_f32(c, F32(L.n1), x, rows, D, 1, s));
  for (int i = 0; i < d.n_blocks; ++i) {
    const YmBlockOff& b = L.blocks[i];
    // x += GQA(RMSNorm(x))        layers.py:101-104, 46-92
    YOHO_TRY(launch_rmsnorm_bf16(x, F32(b.n0), xn, rows, D, 0, s));
    {
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(b.wqkv_h); g.n = 3 * kvd; g.k_per_tap = D;
      g.out_bf16 = qkv; g.ldc = 3 * kvd;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    YOHO_TRY(launch_rope_bf16(qkv, F32(L.rope), rows, S2, 3 * kvd, 2 * d.kv_heads, s));   // q and k heads (layers.py:72-74)
    YOHO_TRY(launch_attention_bf16(ctx, qkv, att, batch, S2, d.kv_heads, s));              // scores / sqrt(64) (layers.py:83-84)
    {
      GemmBf16Args g;
      g.a = att; g.a_ld = kvd; g.a_cols = kvd; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(b.wo_h); g.n = D; g.k_per_tap = kvd;
      g.out_f32 = x; g.ldc = D; g.accumulate = 1;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    // x += SwiGLU(RMSNorm(x))     layers.py:106-109, 7-20
    YOHO_TRY(launch_rmsnorm_bf16(x, F32(b.n1), xn, rows, D, 0, s));
    {
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(b.w12_h); g.n = 2 * H; g.k_per_tap = D;
      g.out_bf16 = hbuf; g.ldc = 2 * H;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    YOHO_TRY(launch_swiglu_gate_bf16(hbuf, gbuf, rows, H, s));
    {
      GemmBf16Args g;
      g.a = gbuf; g.a_ld = H; g.a_cols = H; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(b.w3_h); g.n = D; g.k_per_tap = H;
      g.out_f32 = x; g.ldc = D; g.accumulate = 1;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
  }
  YOHO_CUDA_CHECK(cudaMemcpyAsync(out, x, size_t(rows) * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return YOHO_OK;
}

static int ym_forward_fp32(const yoho_model_desc& d, const YmBlob& L, const char* blob, const float* mel, int batch, int frames,
                           float* out, char* ws, const YmWs& w, cudaStream_t s) {
  const int S = frames, S2 = frames / 2, D = d.dims, M = d.n_mels, kvd = ym_kvd(d), H = ym_hidden(d), KH = d.kv_heads;
  const int rows = batch * S2;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  float* c = reinterpret_cast<float*>(ws + w.c);
  float* p0 = reinterpret_cast<float*>(ws + w.p0);
  float* x = reinterpret_cast<float*>(ws + w.x);
  float* xn = reinterpret_cast<float*>(ws + w.xn);
  float* qkv = reinterpret_cast<float*>(ws + w.qkv);
  float* att = reinterpret_cast<float*>(ws + w.att);
  float* hbuf = reinterpret_cast<float*>(ws + w.h);
  float* gbuf = reinterpret_cast<float*>(ws + w.g);
  float* scores = reinterpret_cast<float*>(ws + w.scores);
  if ((long long)batch * S > 65535LL * 128) return fail(YOHO_ERR_UNSUPPORTED, "fp32 yoho encoder: %lld rows exceed one launch", (long long)batch * S);

  auto conv = [&](const float* a, int len, int ch, const float* wt, const float* bias) {
    SgemmArgs g;
    g.a = a; g.b = wt; g.c = c; g.m = batch * len; g.n = D; g.k = 3 * ch; g.ldb = D; g.ldc = D;
    g.conv = 1; g.conv_len = len; g.conv_out = len; g.conv_ch = ch; g.conv_stride = 1; g.bias = bias;
    return launch_sgemm(g, s);
  };
  auto dense = [&](const float* a, int lda, int k, const float* wt, int n, float* cc, int ldc, int acc) {
    SgemmArgs g;
    g.a = a; g.b = wt; g.c = cc; g.m = rows; g.n = n; g.k = k; g.lda = lda; g.ldb = n; g.ldc = ldc; g.accumulate = acc;
    return launch_sgemm(g, s);
  };
  YOHO_TRY(conv(mel, S, M, F32(L.conv0_w), F32(L.conv0_b)));
  YOHO_TRY(launch_rmsnorm_silu_pool_f32(c, F32(L.n0), p0, batch, S, D, s));
  YOHO_TRY(conv(p0, S2, D, F32(L.conv1_w), F32(L.conv1_b)));
  YOHO_TRY(launch_rmsnorm_f32(c, F32(L.n1), x, rows, D, 1, s));
  for (int i = 0; i < d.n_blocks; ++i) {
    const YmBlockOff& b = L.blocks[i];
    YOHO_TRY(launch_rmsnorm_f32(x, F32(b.n0), xn, rows, D, 0, s));
    YOHO_TRY(dense(xn, D, D, F32(b.wqkv), 3 * kvd, qkv, 3 * kvd, 0));
    YOHO_TRY(launch_rope_f32(qkv, F32(L.rope), rows, S2, 3 * kvd, 2 * KH, s));
    for (int b0 = 0; b0 < batch; b0 += w.score_clips) {
      const int nb = batch - b0 < w.score_clips ? batch - b0 : w.score_clips;
      const float* qb = qkv + size_t(b0) * S2 * 3 * kvd;
      SgemmArgs g;   // scores = q k^T / sqrt(64)
      g.a = qb; g.b = qb + kvd; g.c = scores; g.m = S2; g.n = S2; g.k = 64; g.lda = 3 * kvd; g.ldb = 3 * kvd; g.ldc = S2;
      g.b_transposed = 1; g.alpha = 0.125f; g.nz1 = nb; g.nz2 = KH;
      g.sa1 = (long long)S2 * 3 * kvd; g.sa2 = 64; g.sb1 = g.sa1; g.sb2 = 64; g.sc1 = (long long)KH * S2 * S2; g.sc2 = (long long)S2 * S2;
      YOHO_TRY(launch_sgemm(g, s));
      YOHO_TRY(launch_softmax_rows(scores, (long long)nb * KH * S2, S2, s));
      SgemmArgs g2;  // out = attention v, heads merged "h s d -> s (h d)"
      g2.a = scores; g2.b = qb + 2 * kvd; g2.c = att + size_t(b0) * S2 * kvd; g2.m = S2; g2.n = 64; g2.k = S2;
      g2.lda = S2; g2.ldb = 3 * kvd; g2.ldc = kvd; g2.nz1 = nb; g2.nz2 = KH;
      g2.sa1 = g.sc1; g2.sa2 = g.sc2; g2.sb1 = g.sa1; g2.sb2 = 64; g2.sc1 = (long long)S2 * kvd; g2.sc2 = 64;
      YOHO_TRY(launch_sgemm(g2, s));
    }
    YOHO_TRY(dense(att, kvd, kvd, F32(b.wo), D, x, D, 1));
    YOHO_TRY(launch_rmsnorm_f32(x, F32(b.n1), xn, rows, D, 0, s));
    YOHO_TRY(dense(xn, D, D, F32(b.w1), H, hbuf, 2 * H, 0));
    YOHO_TRY(dense(xn, D, D, F32(b.w2), H, hbuf + H, 2 * H, 0));
    YOHO_TRY(launch_swiglu_gate_f32(hbuf, gbuf, rows, H, s));
    YOHO_TRY(dense(gbuf, H, H, F32(b.w3), D, x, D, 1));
  }
  YOHO_CUDA_CHECK(cudaMemcpyAsync(out, x, size_t(rows) * D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return YOHO_OK;
}

// dst[r][c0 + c] = src[r][c] for a (rows, cols) float32 matrix into a wider row-major matrix
__global__ void copy_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int ld_dst, int c0) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += gridDim.x * blockDim.x) {
    const int r = i / cols, c = i - r * cols;
    dst[size_t(r) * ld_dst + c0 + c] = src[i];
  }
}

}  // namespace yoho

using namespace yoho;

extern "C" int yoho_model_num_tensors(const yoho_model_desc* desc) {
  if (!desc || desc->n_blocks < 0) return fail(YOHO_ERR_INVALID, "yoho_model_num_tensors: bad desc");
  return kYmFixed + kYmPerBlock * desc->n_blocks;
}

extern "C" long long yoho_model_tensor_elems(const yoho_model_desc* desc, int idx) {
  if (ym_check_desc(desc, "yoho_model_tensor_elems") != YOHO_OK) return YOHO_ERR_INVALID;
  if (idx < 0 || idx >= kYmFixed + kYmPerBlock * desc->n_blocks) return fail(YOHO_ERR_INVALID, "yoho_model_tensor_elems: index %d", idx);
  return (long long)ym_tensor_elems(*desc, idx);
}

extern "C" size_t yoho_model_weights_bytes(const yoho_model_desc* desc) {
  if (ym_check_desc(desc, "yoho_model_weights_bytes") != YOHO_OK) return 0;
  return ym_plan_blob(*desc).total;
}

extern "C" int yoho_model_out_len(int frames) {
  if (frames <= 1) return fail(YOHO_ERR_INVALID, "yoho_model_out_len: frames=%d", frames);
  return frames / 2;
}

extern "C" int yoho_model_weights_pack(yoho_ctx* ctx, const yoho_model_desc* desc, const float* const* tensors, int n_tensors,
                                       void* blob_, size_t blob_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !tensors || !blob_) return fail(YOHO_ERR_INVALID, "yoho_model_weights_pack: null argument");
  YOHO_TRY(ym_check_desc(desc, "yoho_model_weights_pack"));
  const yoho_model_desc& d = *desc;
  const int expected = kYmFixed + kYmPerBlock * d.n_blocks;
  if (n_tensors != expected) return fail(YOHO_ERR_INVALID, "yoho_model_weights_pack: %d tensors given, %d expected", n_tensors, expected);
  for (int i = 0; i < n_tensors; ++i)
    if (!tensors[i]) return fail(YOHO_ERR_INVALID, "yoho_model_weights_pack: tensor %d is null", i);
  YmBlob L = ym_plan_blob(d);
  if (blob_bytes < L.total) return fail(YOHO_ERR_WORKSPACE, "yoho_model_weights_pack: blob %zu < %zu bytes", blob_bytes, L.total);
  char* blob = static_cast<char*>(blob_);
  const int D = d.dims, kvd = ym_kvd(d), H = ym_hidden(d);
  auto put = [&](size_t off, int idx) -> int {
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + off, tensors[idx], ym_tensor_elems(d, idx) * sizeof(float), cudaMemcpyDefault, s));
    return YOHO_OK;
  };
  const size_t fixed[kYmFixed] = {L.conv0_w, L.conv0_b, L.n0, L.conv1_w, L.conv1_b, L.n1, L.rope};
  for (int j = 0; j < kYmFixed; ++j) YOHO_TRY(put(fixed[j], j));
  for (int i = 0; i < d.n_blocks; ++i) {
    const YmBlockOff& b = L.blocks[i];
    const size_t offs[kYmPerBlock] = {b.n0, b.wq, b.wk, b.wv, b.wo, b.n1, b.w1, b.w2, b.w3};
    for (int j = 0; j < kYmPerBlock; ++j) YOHO_TRY(put(offs[j], kYmFixed + i * kYmPerBlock + j));
  }
  using bf = __nv_bfloat16;
  auto F32 = [&](size_t off) { return reinterpret_cast<float*>(blob + off); };
  auto H16 = [&](size_t off) { return reinterpret_cast<bf*>(blob + off); };
  // flax kernels are (in, out) / (k, in, out); the tensor-core path wants (out, in) with `in` contiguous
  YOHO_TRY(launch_transpose_cast(F32(L.conv0_w), H16(L.conv0_w_h), 3 * d.n_mels, D, 3 * d.n_mels, s));
  YOHO_TRY(launch_transpose_cast(F32(L.conv1_w), H16(L.conv1_w_h), 3 * D, D, 3 * D, s));
  for (int i = 0; i < d.n_blocks; ++i) {
    const YmBlockOff& b = L.blocks[i];
    // fused float32 (D, 3 kvd) = [sum_g wq | wk | wv]
    YOHO_TRY(launch_group_sum(F32(b.wq), F32(b.wqkv), D, d.kv_heads, d.q_heads / d.kv_heads, 3 * kvd, s));
    const int blocks = ceil_div(D * kvd, 256);
    copy_cols_kernel<<<blocks, 256, 0, s>>>(F32(b.wk), F32(b.wqkv), D, kvd, 3 * kvd, kvd);
    YOHO_LAUNCH_CHECK("copy_cols_kernel");
    copy_cols_kernel<<<blocks, 256, 0, s>>>(F32(b.wv), F32(b.wqkv), D, kvd, 3 * kvd, 2 * kvd);
    YOHO_LAUNCH_CHECK("copy_cols_kernel");
    YOHO_TRY(launch_transpose_cast(F32(b.wqkv), H16(b.wqkv_h), D, 3 * kvd, D, s));
    YOHO_TRY(launch_transpose_cast(F32(b.wo), H16(b.wo_h), kvd, D, kvd, s));
    YOHO_TRY(launch_transpose_cast(F32(b.w1), H16(b.w12_h), D, H, D, s));
    YOHO_TRY(launch_transpose_cast(F32(b.w2), H16(b.w12_h) + size_t(H) * D, D, H, D, s));
    YOHO_TRY(launch_transpose_cast(F32(b.w3), H16(b.w3_h), H, D, H, s));
  }
  return YOHO_OK;
}

extern "C" size_t yoho_model_workspace_bytes(const yoho_model_desc* desc, int batch, int frames, int compute) {
  if (ym_check_desc(desc, "yoho_model_workspace_bytes") != YOHO_OK) return 0;
  if (batch <= 0 || frames <= 1) return 1024;
  return ym_plan_ws(*desc, batch, frames, compute).total;
}

extern "C" int yoho_model_encoder_fwd(yoho_ctx* ctx, const yoho_model_desc* desc, const void* blob, const float* mel, int batch,
                                      int frames, float* out, int compute, void* ws, size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !blob || !mel || !out) return fail(YOHO_ERR_INVALID, "yoho_model_encoder_fwd: null argument");
  YOHO_TRY(ym_check_desc(desc, "yoho_model_encoder_fwd"));
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "yoho_model_encoder_fwd: device is sm_%d%d, kernels are sm_100a only", ctx->cc_major, ctx->cc_minor);
  if (compute != YOHO_COMPUTE_FP32 && compute != YOHO_COMPUTE_BF16) return fail(YOHO_ERR_INVALID, "yoho_model_encoder_fwd: compute=%d", compute);
  if (batch < 0 || frames <= 1) return fail(YOHO_ERR_INVALID, "yoho_model_encoder_fwd: batch=%d frames=%d", batch, frames);
  if (frames / 2 > desc->seq_len / 2)
    return fail(YOHO_ERR_INVALID, "yoho_model_encoder_fwd: %d frames give %d positions but the RoPE table was packed for %d", frames, frames / 2, desc->seq_len / 2);
  if (batch == 0) return YOHO_OK;
  if ((long long)batch * frames > 0x7fffffffLL / 8) return fail(YOHO_ERR_UNSUPPORTED, "yoho_model_encoder_fwd: batch too large for 32-bit row indices");
  YmWs w = ym_plan_ws(*desc, batch, frames, compute);
  if (!ws || ws_bytes < w.total) return fail(YOHO_ERR_WORKSPACE, "yoho_model_encoder_fwd: workspace %zu < %zu bytes", ws_bytes, w.total);
  if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0 || reinterpret_cast<uintptr_t>(blob) % 256 != 0)
    return fail(YOHO_ERR_INVALID, "yoho_model_encoder_fwd: workspace must be 1024-byte and blob 256-byte aligned");
  YmBlob L = ym_plan_blob(*desc);
  if (compute == YOHO_COMPUTE_BF16)
    return ym_forward_bf16(ctx, *desc, L, static_cast<const char*>(blob), mel, batch, frames, out, static_cast<char*>(ws), w, s);
  return ym_forward_fp32(*desc, L, static_cast<const char*>(blob), mel, batch, frames, out, static_cast<char*>(ws), w, s);
}
