// Whisper text decoder behind the C ABI: reference yoho/src/nn/whisper.py:138-165
// (TextDecoder), :91-114 (DecoderBlock), :37-67 (MultiHeadAttention) and the greedy loop at :260-285.
//
//   yoho_decoder_fwd      teacher-forced forward == TextDecoder.__call__: tokens (B, S) + encoder states -> logits
//                         (B, S, vocab); float32 SIMT kernels (the 1e-4 parity path).
//   yoho_decoder_begin    start of a generation: cross-attention keys / values of every layer computed once per clip
//   yoho_decoder_step     (the reference recomputes them per token), position reset; then one token per call with a
//   yoho_decoder_generate self-attention KV cache.  The position is a device-side counter, so every step is the same
//                         launch sequence (CUDA-graph friendly); greedy choice = first argmax (:279).
#include "kernels.cuh"

#include <vector>

namespace yoho {

constexpr int kDecFixed = 4, kDecPerLayer = 24;

struct DecLayerOff {
  // float32 originals, in pack order
  size_t ln0_s, ln0_b, s_wq, s_bq, s_wk, s_wv, s_bv, s_wo, s_bo;
  size_t ln1_s, ln1_b, c_wq, c_bq, c_wk, c_wv, c_bv, c_wo, c_bo;
  size_t ln2_s, ln2_b, w1, b1, w2, b2;
  // derived: fused [k | v] projections and their biases [0 | bv] (key_proj has no bias, whisper.py:33)
  size_t s_wkv, s_bkv, c_wkv, c_bkv;
  size_t s_wqkv, s_bqkv;   // [q | k | v] and [bq | 0 | bv] for the one-launch self-attention projection of the cached step
};
struct DecBlob {
  size_t pos, emb, lnf_s, lnf_b;
  size_t emb_t;            // (D, vocab_ld) float32: the tied embedding transposed for the vocabulary projection (:163)
  std::vector<DecLayerOff> layers;
  size_t total = 0;
};

static int vocab_ld(const yoho_decoder_desc& d) { return int(align_up(size_t(d.vocab), 32)); }

static size_t dec_tensor_elems(const yoho_decoder_desc& d, int idx) {
  const size_t D = d.dims, A = d.audio_dims;
  if (idx < kDecFixed) {
    switch (idx) {
      case 0: return size_t(d.seq_len) * D;
      case 1: return size_t(d.vocab) * D;
      default: return D;
    }
  }
  switch ((idx - kDecFixed) % kDecPerLayer) {
    case 2: case 4: case 5: case 7: case 11: case 16: return D * D;   // self q k v o, cross q o
    case 13: case 14: return A * D;                                     // cross k v act on the encoder states
    case 20: case 22: return 4 * D * D;
    case 21: return 4 * D;
    default: return D;
  }
}

static DecBlob dec_plan_blob(const yoho_decoder_desc& d) {
  DecBlob L;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 256); size_t o = off; off += bytes; return o; };
  const size_t D = d.dims, A = d.audio_dims, f = sizeof(float);
  L.pos = take(size_t(d.seq_len) * D * f);
  L.emb = take(size_t(d.vocab) * D * f);
  L.lnf_s = take(D * f); L.lnf_b = take(D * f);
  L.layers.resize(d.n_layers);
  for (auto& l : L.layers) {
    l.ln0_s = take(D * f); l.ln0_b = take(D * f);
    l.s_wq = take(D * D * f); l.s_bq = take(D * f); l.s_wk = take(D * D * f); l.s_wv = take(D * D * f); l.s_bv = take(D * f);
    l.s_wo = take(D * D * f); l.s_bo = take(D * f);
    l.ln1_s = take(D * f); l.ln1_b = take(D * f);
    l.c_wq = take(D * D * f); l.c_bq = take(D * f); l.c_wk = take(A * D * f); l.c_wv = take(A * D * f); l.c_bv = take(D * f);
    l.c_wo = take(D * D * f); l.c_bo = take(D * f);
    l.ln2_s = take(D * f); l.ln2_b = take(D * f);
    l.w1 = take(4 * D * D * f); l.b1 = take(4 * D * f); l.w2 = take(4 * D * D * f); l.b2 = take(D * f);
  }
  L.emb_t = take(D * size_t(vocab_ld(d)) * f);
  for (auto& l : L.layers) {
    l.s_wkv = take(D * 2 * D * f); l.s_bkv = take(2 * D * f);
    l.c_wkv = take(A * 2 * D * f); l.c_bkv = take(2 * D * f);
    l.s_wqkv = take(D * 3 * D * f); l.s_bqkv = take(3 * D * f);
  }
  L.total = align_up(off, 1024);
  return L;
}

static int dec_check(const yoho_decoder_desc* d, const char* who) {
  if (!d) return fail(YOHO_ERR_INVALID, "%s: desc is null", who);
  if (d->vocab <= 0 || d->seq_len <= 0 || d->dims <= 0 || d->n_heads <= 0 || d->n_layers < 0 || d->audio_dims <= 0 || d->audio_len <= 0)
    return fail(YOHO_ERR_INVALID, "%s: desc has a non-positive field", who);
  if (d->dims != d->n_heads * 64)
    return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d, n_heads=%d: kernels need d_head = 64 (every Whisper size)", who, d->dims, d->n_heads);
  if (d->dims > 1280 || d->audio_dims % 4 != 0) return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d audio_dims=%d", who, d->dims, d->audio_dims);
  return YOHO_OK;
}

// out (rows, 2 * cols) <- [a | b] column blocks of two (rows, cols) matrices
__global__ void dec_pair_cols_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, int rows, int cols) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)rows * cols; i += (long long)gridDim.x * blockDim.x) {
    const int r = int(i / cols), c = int(i - (long long)r * cols);
    out[size_t(r) * 2 * cols + c] = a[i];
    out[size_t(r) * 2 * cols + cols + c] = b[i];
  }
}

// dst (rows, ld) columns [c0, c0 + cols) <- src (rows, cols)
__global__ void dec_place_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int ld, int c0) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)rows * cols; i += (long long)gridDim.x * blockDim.x) {
    const int r = int(i / cols), c = int(i - (long long)r * cols);
    dst[size_t(r) * ld + c0 + c] = src[i];
  }
}

// dst (cols, ld) <- transpose of src (rows = vocab, cols = D), zero padded to ld
__global__ void dec_transpose_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int ld) {
  __shared__ float tile[32][33];
  const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? src[size_t(r) * cols + c] : 0.0f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < cols && r < ld) dst[size_t(c) * ld + r] = tile[threadIdx.x][i];
  }
}

#define YOHO_TRY(expr) do { int _r = (expr); if (_r != YOHO_OK) return _r; } while (0)

// ------------------------------------------------------------------------------ generation state
// cache = [ header (1024 B: int position) | self k|v (L, B, seq_len, 2D) | cross k|v (L, B, audio_len, 2D) ]
struct DecCache {
  int* pos;
  float* self_kv;
  float* cross_kv;
  size_t self_layer, cross_layer;   // floats per layer
  size_t total;
};
static DecCache dec_plan_cache(const yoho_decoder_desc& d, int batch, void* base) {
  DecCache c;
  const size_t D = d.dims;
  c.self_layer = size_t(batch) * d.seq_len * 2 * D;
  c.cross_layer = size_t(batch) * d.audio_len * 2 * D;
  char* p = static_cast<char*>(base);
  c.pos = reinterpret_cast<int*>(p);
  size_t off = 1024;
  c.self_kv = reinterpret_cast<float*>(p + off);
  off += align_up(c.self_layer * d.n_layers * sizeof(float), 1024);
  c.cross_kv = reinterpret_cast<float*>(p + off);
  off += align_up(c.cross_layer * d.n_layers * sizeof(float), 1024);
  c.total = off;
  return c;
}

struct StepWs { size_t x, xn, q, att, h, logits, total; };
static StepWs dec_plan_step_ws(const yoho_decoder_desc& d, int batch) {
  StepWs w;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 1024); size_t o = off; off += bytes; return o; };
  const size_t B = batch, D = d.dims, f = sizeof(float);
  w.x = take(B * D * f); w.xn = take(B * D * f); w.q = take(B * D * f); w.att = take(B * D * f); w.h = take(B * 4 * D * f);
  w.logits = take(B * size_t(vocab_ld(d)) * f);
  w.total = align_up(off, 1024);
  return w;
}

struct DenseOpt {
  const float* bias = nullptr;
  const float* ln_scale = nullptr; const float* ln_bias = nullptr;     // LayerNorm of the input rows, fused into the load
  float* out2 = nullptr; long long out2_row_stride = 0; int n_split = 0;
  const int* row_off_ptr = nullptr; int row_off_mul = 0;
  int gelu = 0, accumulate = 0;
};

// out (batch rows) = act(LN?(x) @ w + bias); the skinny GEMM keeps <= 64 activation rows in registers, larger batches go
// in slices (the weights then come from L2)
static int dec_skinny(const float* x, int ldx, int batch, int k, const float* w, int ldw, int n, float* out, long long out_row_stride,
                      const DenseOpt& o, cudaStream_t s) {
  for (int b0 = 0; b0 < batch; b0 += 64) {
    SkinnyArgs a;
    a.x = x + size_t(b0) * ldx; a.ldx = ldx; a.w = w; a.ldw = ldw; a.bias = o.bias;
    a.out = out + (long long)b0 * out_row_stride; a.out_row_stride = out_row_stride;
    a.out2 = o.out2 ? o.out2 + (long long)b0 * o.out2_row_stride : nullptr; a.out2_row_stride = o.out2_row_stride; a.n_split = o.n_split;
    a.row_off_ptr = o.row_off_ptr; a.row_off_mul = o.row_off_mul;
    a.ln_scale = o.ln_scale; a.ln_bias = o.ln_bias;
    a.b = batch - b0 < 64 ? batch - b0 : 64; a.k = k; a.n = n; a.gelu = o.gelu; a.accumulate = o.accumulate;
    int r = launch_skinny_gemm(a, s);
    if (r != YOHO_OK) return r;
  }
  return YOHO_OK;
}

// One token for every clip: consumes tokens[:, p], appends to the self cache at p, leaves logits for position p + 1.
// 8 launches per layer: every LayerNorm is fused into the GEMM that consumes it (each CTA recomputes the statistics of
// the few rows; from 17 rows on a separate LayerNorm launch is cheaper), q | k | v of the self attention is one GEMM whose
// k | v columns land in the cache row of the current position.
static int dec_step(const yoho_decoder_desc& d, const DecBlob& L, const char* blob, const DecCache& c, const int* tokens, int tok_ld,
                    int batch, char* ws, const StepWs& w, float* logits, int logits_ld, cudaStream_t s) {
  const int D = d.dims, H = d.n_heads, T = d.audio_len, S = d.seq_len;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  float* x = reinterpret_cast<float*>(ws + w.x);
  float* xn = reinterpret_cast<float*>(ws + w.xn);
  float* q = reinterpret_cast<float*>(ws + w.q);
  float* att = reinterpret_cast<float*>(ws + w.att);
  float* h = reinterpret_cast<float*>(ws + w.h);
  const bool fuse_ln = batch <= 16;
  // input of a GEMM that consumes LN(x): x itself with the LayerNorm fused into the load, or a normalised copy
  auto normed = [&](size_t scale, size_t bias, DenseOpt& o, const float** in) -> int {
    if (fuse_ln) { o.ln_scale = F32(scale); o.ln_bias = F32(bias); *in = x; return YOHO_OK; }
    *in = xn;
    return launch_layernorm_f32(x, F32(scale), F32(bias), xn, batch, D, s);
  };
  const float* in = nullptr;
  YOHO_TRY(launch_embed(tokens, tok_ld, c.pos, batch, 1, F32(L.emb), F32(L.pos), x, D, d.vocab, s));
  for (int i = 0; i < d.n_layers; ++i) {
    const DecLayerOff& l = L.layers[i];
    float* self_kv = c.self_kv + c.self_layer * i;
    const float* cross_kv = c.cross_kv + c.cross_layer * i;
    // q += out_proj(attention(LN(q), cache))                         whisper.py:95-99
    {
      DenseOpt o;
      o.bias = F32(l.s_bqkv);
      YOHO_TRY(normed(l.ln0_s, l.ln0_b, o, &in));
      o.out2 = self_kv; o.out2_row_stride = (long long)S * 2 * D; o.n_split = D; o.row_off_ptr = c.pos; o.row_off_mul = 2 * D;
      YOHO_TRY(dec_skinny(in, D, batch, D, F32(l.s_wqkv), 3 * D, 3 * D, q, D, o, s));
    }
    {
      DecodeAttArgs a;
      a.q = q; a.ldq = D; a.kv = self_kv; a.kv_clip_stride = (long long)S * 2 * D; a.kv_ld = 2 * D; a.k_off = 0; a.v_off = D;
      a.len_ptr = c.pos; a.len_add = 1; a.out = att; a.ldo = D;
      YOHO_TRY(launch_decode_attention(a, batch, H, S, s));
    }
    { DenseOpt o; o.bias = F32(l.s_bo); o.accumulate = 1; YOHO_TRY(dec_skinny(att, D, batch, D, F32(l.s_wo), D, D, x, D, o, s)); }
    // x += out_proj(attention(LN(x), encoder states))                whisper.py:101-105
    {
      DenseOpt o;
      o.bias = F32(l.c_bq);
      YOHO_TRY(normed(l.ln1_s, l.ln1_b, o, &in));
      YOHO_TRY(dec_skinny(in, D, batch, D, F32(l.c_wq), D, D, q, D, o, s));
    }
    {
      DecodeAttArgs a;
      a.q = q; a.ldq = D; a.kv = cross_kv; a.kv_clip_stride = (long long)T * 2 * D; a.kv_ld = 2 * D; a.k_off = 0; a.v_off = D;
      a.len_const = T; a.out = att; a.ldo = D;
      YOHO_TRY(launch_decode_attention(a, batch, H, T, s));
    }
    { DenseOpt o; o.bias = F32(l.c_bo); o.accumulate = 1; YOHO_TRY(dec_skinny(att, D, batch, D, F32(l.c_wo), D, D, x, D, o, s)); }
    // x += Dense_1(gelu(Dense_0(LN(x))))                             whisper.py:107-112
    {
      DenseOpt o;
      o.bias = F32(l.b1); o.gelu = 1;
      YOHO_TRY(normed(l.ln2_s, l.ln2_b, o, &in));
      YOHO_TRY(dec_skinny(in, D, batch, D, F32(l.w1), 4 * D, 4 * D, h, 4 * D, o, s));
    }
    { DenseOpt o; o.bias = F32(l.b2); o.accumulate = 1; YOHO_TRY(dec_skinny(h, 4 * D, batch, 4 * D, F32(l.w2), D, D, x, D, o, s)); }
  }
  // logits = LN(x) @ embedding.T                                    whisper.py:162-163
  DenseOpt o;
  YOHO_TRY(normed(L.lnf_s, L.lnf_b, o, &in));
  return dec_skinny(in, D, batch, D, F32(L.emb_t), vocab_ld(d), d.vocab, logits, logits_ld, o, s);
}

// ------------------------------------------------------------------------------ teacher-forced forward (float32)
struct FwdWs { size_t x, xn, q, kv, ckv, att, h, scores; int score_clips; size_t total; };
static FwdWs dec_plan_fwd_ws(const yoho_decoder_desc& d, int batch, int s_len) {
  FwdWs w;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 1024); size_t o = off; off += bytes; return o; };
  const size_t B = batch, S = s_len, D = d.dims, T = d.audio_len, f = sizeof(float);
  w.x = take(B * S * D * f); w.xn = take(B * S * D * f); w.q = take(B * S * D * f); w.kv = take(B * S * 2 * D * f);
  w.ckv = take(B * T * 2 * D * f); w.att = take(B * S * D * f); w.h = take(B * S * 4 * D * f);
  const size_t per_clip = size_t(d.n_heads) * S * (T > S ? T : S) * f;
  size_t clips = (size_t(1) << 30) / (per_clip ? per_clip : 1);
  if (clips < 1) clips = 1;
  if (clips > B) clips = B;
  if (clips * d.n_heads > 65535) clips = 65535 / d.n_heads;
  w.score_clips = int(clips);
  w.scores = take(clips * per_clip);
  w.total = align_up(off, 1024);
  return w;
}

static int dec_forward(const yoho_decoder_desc& d, const DecBlob& L, const char* blob, const int* tokens, const float* states, int batch,
                       int S, float* logits, char* ws, const FwdWs& w, cudaStream_t s) {
  const int D = d.dims, H = d.n_heads, T = d.audio_len, A = d.audio_dims;
  const int rows = batch * S;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  float* x = reinterpret_cast<float*>(ws + w.x);
  float* xn = reinterpret_cast<float*>(ws + w.xn);
  float* q = reinterpret_cast<float*>(ws + w.q);
  float* kv = reinterpret_cast<float*>(ws + w.kv);
  float* ckv = reinterpret_cast<float*>(ws + w.ckv);
  float* att = reinterpret_cast<float*>(ws + w.att);
  float* h = reinterpret_cast<float*>(ws + w.h);
  float* scores = reinterpret_cast<float*>(ws + w.scores);
  auto dense = [&](const float* a, int m, int lda, int k, const float* wt, int ldb, int n, const float* bias, float* c, int ldc, int gelu, int acc) {
    SgemmArgs g;
    g.a = a; g.b = wt; g.c = c; g.m = m; g.n = n; g.k = k; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
    g.bias = bias; g.gelu = gelu; g.accumulate = acc;
    return launch_sgemm(g, s);
  };
  // softmax(q k^T / 8 [+ causal mask]) v over (clip, head), keys / values taken from `kvbuf` (rows of [k | v])
  auto attend = [&](const float* kvbuf, int n_keys, int causal) -> int {
    for (int b0 = 0; b0 < batch; b0 += w.score_clips) {
      const int nb = batch - b0 < w.score_clips ? batch - b0 : w.score_clips;
      SgemmArgs g;
      g.a = q + size_t(b0) * S * D; g.b = kvbuf + size_t(b0) * n_keys * 2 * D; g.c = scores;
      g.m = S; g.n = n_keys; g.k = 64; g.lda = D; g.ldb = 2 * D; g.ldc = n_keys; g.b_transposed = 1; g.alpha = 0.125f;
      g.nz1 = nb; g.nz2 = H;
      g.sa1 = (long long)S * D; g.sa2 = 64; g.sb1 = (long long)n_keys * 2 * D; g.sb2 = 64;
      g.sc1 = (long long)H * S * n_keys; g.sc2 = (long long)S * n_keys;
      YOHO_TRY(launch_sgemm(g, s));
      YOHO_TRY(launch_softmax_rows(scores, (long long)nb * H * S, n_keys, s, causal ? S : 0));
      SgemmArgs g2;
      g2.a = scores; g2.b = kvbuf + size_t(b0) * n_keys * 2 * D + D; g2.c = att + size_t(b0) * S * D;
      g2.m = S; g2.n = 64; g2.k = n_keys; g2.lda = n_keys; g2.ldb = 2 * D; g2.ldc = D; g2.nz1 = nb; g2.nz2 = H;
      g2.sa1 = g.sc1; g2.sa2 = g.sc2; g2.sb1 = g.sb1; g2.sb2 = 64; g2.sc1 = (long long)S * D; g2.sc2 = 64;
      YOHO_TRY(launch_sgemm(g2, s));
    }
    return YOHO_OK;
  };
  YOHO_TRY(launch_embed(tokens, S, nullptr, batch, S, F32(L.emb), F32(L.pos), x, D, d.vocab, s));
  for (int i = 0; i < d.n_layers; ++i) {
    const DecLayerOff& l = L.layers[i];
    YOHO_TRY(launch_layernorm_f32(x, F32(l.ln0_s), F32(l.ln0_b), xn, rows, D, s));
    YOHO_TRY(dense(xn, rows, D, D, F32(l.s_wq), D, D, F32(l.s_bq), q, D, 0, 0));
    YOHO_TRY(dense(xn, rows, D, D, F32(l.s_wkv), 2 * D, 2 * D, F32(l.s_bkv), kv, 2 * D, 0, 0));
    YOHO_TRY(attend(kv, S, 1));
    YOHO_TRY(dense(att, rows, D, D, F32(l.s_wo), D, D, F32(l.s_bo), x, D, 0, 1));
    YOHO_TRY(launch_layernorm_f32(x, F32(l.ln1_s), F32(l.ln1_b), xn, rows, D, s));
    YOHO_TRY(dense(xn, rows, D, D, F32(l.c_wq), D, D, F32(l.c_bq), q, D, 0, 0));
    YOHO_TRY(dense(states, batch * T, A, A, F32(l.c_wkv), 2 * D, 2 * D, F32(l.c_bkv), ckv, 2 * D, 0, 0));
    YOHO_TRY(attend(ckv, T, 0));
    YOHO_TRY(dense(att, rows, D, D, F32(l.c_wo), D, D, F32(l.c_bo), x, D, 0, 1));
    YOHO_TRY(launch_layernorm_f32(x, F32(l.ln2_s), F32(l.ln2_b), xn, rows, D, s));
    YOHO_TRY(dense(xn, rows, D, D, F32(l.w1), 4 * D, 4 * D, F32(l.b1), h, 4 * D, 1, 0));
    YOHO_TRY(dense(h, rows, 4 * D, 4 * D, F32(l.w2), D, D, F32(l.b2), x, D, 0, 1));
  }
  YOHO_TRY(launch_layernorm_f32(x, F32(L.lnf_s), F32(L.lnf_b), xn, rows, D, s));
  return dense(xn, rows, D, D, F32(L.emb_t), vocab_ld(d), d.vocab, nullptr, logits, d.vocab, 0, 0);
}

}  // namespace yoho

using namespace yoho;

extern "C" int yoho_decoder_num_tensors(const yoho_decoder_desc* desc) {
  if (!desc || desc->n_layers < 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_num_tensors: bad desc");
  return kDecFixed + kDecPerLayer * desc->n_layers;
}

extern "C" long long yoho_decoder_tensor_elems(const yoho_decoder_desc* desc, int idx) {
  if (dec_check(desc, "yoho_decoder_tensor_elems") != YOHO_OK) return YOHO_ERR_INVALID;
  if (idx < 0 || idx >= kDecFixed + kDecPerLayer * desc->n_layers) return fail(YOHO_ERR_INVALID, "yoho_decoder_tensor_elems: index %d", idx);
  return (long long)dec_tensor_elems(*desc, idx);
}

extern "C" size_t yoho_decoder_weights_bytes(const yoho_decoder_desc* desc) {
  if (dec_check(desc, "yoho_decoder_weights_bytes") != YOHO_OK) return 0;
  return dec_plan_blob(*desc).total;
}

extern "C" int yoho_decoder_weights_pack(yoho_ctx* ctx, const yoho_decoder_desc* desc, const float* const* tensors, int n_tensors,
                                         void* blob_, size_t blob_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !tensors || !blob_) return fail(YOHO_ERR_INVALID, "yoho_decoder_weights_pack: null argument");
  YOHO_TRY(dec_check(desc, "yoho_decoder_weights_pack"));
  const yoho_decoder_desc& d = *desc;
  const int want = kDecFixed + kDecPerLayer * d.n_layers;
  if (n_tensors != want) return fail(YOHO_ERR_INVALID, "yoho_decoder_weights_pack: %d tensors given, %d expected", n_tensors, want);
  for (int i = 0; i < n_tensors; ++i)
    if (!tensors[i]) return fail(YOHO_ERR_INVALID, "yoho_decoder_weights_pack: tensor %d is null", i);
  DecBlob L = dec_plan_blob(d);
  if (blob_bytes < L.total) return fail(YOHO_ERR_WORKSPACE, "yoho_decoder_weights_pack: blob %zu < %zu bytes", blob_bytes, L.total);
  if (reinterpret_cast<uintptr_t>(blob_) % 256 != 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_weights_pack: blob must be 256-byte aligned");
  char* blob = static_cast<char*>(blob_);
  auto put = [&](size_t off, int idx) -> int {
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + off, tensors[idx], dec_tensor_elems(d, idx) * sizeof(float), cudaMemcpyDefault, s));
    return YOHO_OK;
  };
  YOHO_TRY(put(L.pos, 0)); YOHO_TRY(put(L.emb, 1)); YOHO_TRY(put(L.lnf_s, 2)); YOHO_TRY(put(L.lnf_b, 3));
  const int D = d.dims, A = d.audio_dims;
  auto F32 = [&](size_t off) { return reinterpret_cast<float*>(blob + off); };
  for (int i = 0; i < d.n_layers; ++i) {
    const DecLayerOff& l = L.layers[i];
    const size_t offs[kDecPerLayer] = {l.ln0_s, l.ln0_b, l.s_wq, l.s_bq, l.s_wk, l.s_wv, l.s_bv, l.s_wo, l.s_bo,
                                       l.ln1_s, l.ln1_b, l.c_wq, l.c_bq, l.c_wk, l.c_wv, l.c_bv, l.c_wo, l.c_bo,
                                       l.ln2_s, l.ln2_b, l.w1, l.b1, l.w2, l.b2};
    for (int j = 0; j < kDecPerLayer; ++j) YOHO_TRY(put(offs[j], kDecFixed + i * kDecPerLayer + j));
    dec_pair_cols_kernel<<<ceil_div(D * D, 256), 256, 0, s>>>(F32(l.s_wk), F32(l.s_wv), F32(l.s_wkv), D, D);
    YOHO_LAUNCH_CHECK("dec_pair_cols_kernel");
    dec_pair_cols_kernel<<<ceil_div(A * D, 256), 256, 0, s>>>(F32(l.c_wk), F32(l.c_wv), F32(l.c_wkv), A, D);
    YOHO_LAUNCH_CHECK("dec_pair_cols_kernel");
    dec_place_cols_kernel<<<ceil_div(D * D, 256), 256, 0, s>>>(F32(l.s_wq), F32(l.s_wqkv), D, D, 3 * D, 0);
    YOHO_LAUNCH_CHECK("dec_place_cols_kernel");
    dec_place_cols_kernel<<<ceil_div(D * 2 * D, 256), 256, 0, s>>>(F32(l.s_wkv), F32(l.s_wqkv), D, 2 * D, 3 * D, D);
    YOHO_LAUNCH_CHECK("dec_place_cols_kernel");
    YOHO_TRY(launch_fill_zero(blob + l.s_bqkv, size_t(3) * D * sizeof(float), s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.s_bqkv, blob + l.s_bq, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.s_bqkv + size_t(2) * D * sizeof(float), blob + l.s_bv, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    for (size_t pair : {l.s_bkv, l.c_bkv}) YOHO_TRY(launch_fill_zero(blob + pair, size_t(2) * D * sizeof(float), s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.s_bkv + size_t(D) * sizeof(float), blob + l.s_bv, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.c_bkv + size_t(D) * sizeof(float), blob + l.c_bv, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
  }
  dec_transpose_kernel<<<dim3(ceil_div(vocab_ld(d), 32), ceil_div(D, 32)), dim3(32, 8), 0, s>>>(F32(L.emb), F32(L.emb_t), d.vocab, D, vocab_ld(d));
  YOHO_LAUNCH_CHECK("dec_transpose_kernel");
  return YOHO_OK;
}

static int dec_common(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, const char* who) {
  if (!ctx || !blob) return fail(YOHO_ERR_INVALID, "%s: null argument", who);
  YOHO_TRY(dec_check(desc, who));
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "%s: device is sm_%d%d, kernels are sm_100a only", who, ctx->cc_major, ctx->cc_minor);
  if (reinterpret_cast<uintptr_t>(blob) % 256 != 0) return fail(YOHO_ERR_INVALID, "%s: blob must be 256-byte aligned", who);
  return YOHO_OK;
}

extern "C" size_t yoho_decoder_workspace_bytes(const yoho_decoder_desc* desc, int batch, int s_len) {
  if (dec_check(desc, "yoho_decoder_workspace_bytes") != YOHO_OK) return 0;
  if (batch <= 0 || s_len <= 0) return 1024;
  return dec_plan_fwd_ws(*desc, batch, s_len).total;
}

extern "C" int yoho_decoder_fwd(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, const int32_t* tokens, const float* states,
                                int batch, int s_len, float* logits, void* ws, size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  YOHO_TRY(dec_common(ctx, desc, blob, "yoho_decoder_fwd"));
  if (!tokens || !states || !logits) return fail(YOHO_ERR_INVALID, "yoho_decoder_fwd: null argument");
  if (batch < 0 || s_len <= 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_fwd: batch=%d s_len=%d", batch, s_len);
  if (s_len > desc->seq_len) return fail(YOHO_ERR_INVALID, "yoho_decoder_fwd: %d tokens but positional_embedding has %d rows (whisper.py:155)", s_len, desc->seq_len);
  if (batch == 0) return YOHO_OK;
  if ((long long)batch * (desc->audio_len > s_len ? desc->audio_len : s_len) > 65535LL * 128) return fail(YOHO_ERR_UNSUPPORTED, "yoho_decoder_fwd: batch too large for one launch");
  FwdWs w = dec_plan_fwd_ws(*desc, batch, s_len);
  if (!ws || ws_bytes < w.total) return fail(YOHO_ERR_WORKSPACE, "yoho_decoder_fwd: workspace %zu < %zu bytes", ws_bytes, w.total);
  if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_fwd: workspace must be 1024-byte aligned");
  DecBlob L = dec_plan_blob(*desc);
  return dec_forward(*desc, L, static_cast<const char*>(blob), tokens, states, batch, s_len, logits, static_cast<char*>(ws), w, s);
}

extern "C" size_t yoho_decoder_cache_bytes(const yoho_decoder_desc* desc, int batch) {
  if (dec_check(desc, "yoho_decoder_cache_bytes") != YOHO_OK) return 0;
  if (batch <= 0) return 1024;
  return dec_plan_cache(*desc, batch, nullptr).total;
}

extern "C" size_t yoho_decoder_step_workspace_bytes(const yoho_decoder_desc* desc, int batch) {
  if (dec_check(desc, "yoho_decoder_step_workspace_bytes") != YOHO_OK) return 0;
  if (batch <= 0) return 1024;
  return dec_plan_step_ws(*desc, batch).total;
}

extern "C" int yoho_decoder_begin(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob_, const float* states, int batch,
                                  void* cache, size_t cache_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  YOHO_TRY(dec_common(ctx, desc, blob_, "yoho_decoder_begin"));
  if (!states || !cache) return fail(YOHO_ERR_INVALID, "yoho_decoder_begin: null argument");
  if (batch <= 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_begin: batch=%d", batch);
  if (reinterpret_cast<uintptr_t>(cache) % 1024 != 0) return fail(YOHO_ERR_INVALID, "yoho_decoder_begin: cache must be 1024-byte aligned");
  DecCache c = dec_plan_cache(*desc, batch, cache);
  if (cache_bytes < c.total) return fail(YOHO_ERR_WORKSPACE, "yoho_decoder_begin: cache %zu < %zu bytes", cache_bytes, c.total);
  const long long rows = (long long)batch * desc->audio_len;
  if (rows > 65535LL * 128) return fail(YOHO_ERR_UNSUPPORTED, "yoho_decoder_begin: %lld encoder rows exceed one launch", rows);
  const char* blob = static_cast<const char*>(blob_);
  DecBlob L = dec_plan_blob(*desc);
  YOHO_CUDA_CHECK(cudaMemsetAsync(cache, 0, 1024, s));
  const int D = desc->dims, A = desc->audio_dims;
  for (int i = 0; i < desc->n_layers; ++i) {
    SgemmArgs g;   // cross k | v of layer i: states @ [key_proj | value_proj] + [0 | bv]   (whisper.py:104 -> :45-46)
    g.a = states; g.b = reinterpret_cast<const float*>(blob + L.layers[i].c_wkv); g.c = c.cross_kv + c.cross_layer * i;
    g.m = int(rows); g.n = 2 * D; g.k = A; g.lda = A; g.ldb = 2 * D; g.ldc = 2 * D;
    g.bias = reinterpret_cast<const float*>(blob + L.layers[i].c_bkv);
    YOHO_TRY(launch_sgemm(g, s));
  }
  return YOHO_OK;
}

static int dec_step_checked(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, int32_t* tokens, int batch, void* cache,
                            size_t cache_bytes, float* logits, const unsigned char* forced, int eot, int n_steps, void* ws, size_t ws_bytes,
                            cudaStream_t s, const char* who) {
  YOHO_TRY(dec_common(ctx, desc, blob, who));
  if (!tokens || !cache) return fail(YOHO_ERR_INVALID, "%s: null argument", who);
  if (batch <= 0 || n_steps < 0) return fail(YOHO_ERR_INVALID, "%s: batch=%d n_steps=%d", who, batch, n_steps);
  DecCache c = dec_plan_cache(*desc, batch, cache);
  if (cache_bytes < c.total) return fail(YOHO_ERR_WORKSPACE, "%s: cache %zu < %zu bytes", who, cache_bytes, c.total);
  StepWs w = dec_plan_step_ws(*desc, batch);
  if (!ws || ws_bytes < w.total) return fail(YOHO_ERR_WORKSPACE, "%s: workspace %zu < %zu bytes", who, ws_bytes, w.total);
  if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0 || reinterpret_cast<uintptr_t>(cache) % 1024 != 0)
    return fail(YOHO_ERR_INVALID, "%s: workspace and cache must be 1024-byte aligned", who);
  DecBlob L = dec_plan_blob(*desc);
  float* lg = logits ? logits : reinterpret_cast<float*>(static_cast<char*>(ws) + w.logits);
  const int lg_ld = logits ? desc->vocab : vocab_ld(*desc);
  for (int i = 0; i < n_steps; ++i) {
    YOHO_TRY(dec_step(*desc, L, static_cast<const char*>(blob), c, tokens, desc->seq_len, batch, static_cast<char*>(ws), w, lg, lg_ld, s));
    YOHO_TRY(launch_greedy_pick(lg, lg_ld, desc->vocab, tokens, desc->seq_len, desc->seq_len, batch, c.pos, forced, eot, s));
  }
  return YOHO_OK;
}

extern "C" int yoho_decoder_step(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, int32_t* tokens, int batch, void* cache,
                                 size_t cache_bytes, float* logits, const unsigned char* forced, int eot, void* ws, size_t ws_bytes,
                                 yoho_stream_t stream) {
  return dec_step_checked(ctx, desc, blob, tokens, batch, cache, cache_bytes, logits, forced, eot, 1, ws, ws_bytes,
                          static_cast<cudaStream_t>(stream), "yoho_decoder_step");
}

extern "C" int yoho_decoder_generate(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, int32_t* tokens, int batch, void* cache,
                                     size_t cache_bytes, const unsigned char* forced, int eot, int n_steps, void* ws, size_t ws_bytes,
                                     yoho_stream_t stream) {
  if (desc && n_steps > desc->seq_len - 1) return fail(YOHO_ERR_INVALID, "yoho_decoder_generate: %d steps exceed the %d-token buffer", n_steps, desc->seq_len);
  return dec_step_checked(ctx, desc, blob, tokens, batch, cache, cache_bytes, nullptr, forced, eot, n_steps, ws, ws_bytes,
                          static_cast<cudaStream_t>(stream), "yoho_decoder_generate");
}

// ------------------------------------------------------------------------------ test hook
extern "C" int yoho_test_skinny_gemm(yoho_ctx* ctx, const float* x, const float* w, const float* bias, float* out, int b, int k, int n,
                                     int gelu, int accumulate, yoho_stream_t stream_) {
  if (!ctx || !x || !w || !out) return fail(YOHO_ERR_INVALID, "yoho_test_skinny_gemm: null argument");
  DenseOpt o;
  o.bias = bias; o.gelu = gelu; o.accumulate = accumulate;
  return dec_skinny(x, k, b, k, w, n, n, out, n, o, static_cast<cudaStream_t>(stream_));
}
