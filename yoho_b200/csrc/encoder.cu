// Whisper audio-encoder forward (reference yoho/src/nn/whisper.py:117-135) behind the C ABI:
// weight packing, workspace planning and the per-layer launch sequence for both compute modes.
//
//   YOHO_COMPUTE_BF16: tcgen05 GEMM / attention kernels; fp32 residual stream and LayerNorm statistics,
//                      bf16 only as tensor-core operands, fp32 accumulation.
//   YOHO_COMPUTE_FP32: SIMT fp32 kernels, the 1e-4 parity path.
#include "kernels.cuh"

#include <cstdlib>
#include <vector>

namespace yoho {

// LayerNorm_0 / LayerNorm_1 of the blocks folded into the GEMMs around them (default) or run as stand-alone kernels
// (YOHO_LN_FUSE=0: A/B measurement; read once per process, it decides what yoho_encoder_weights_pack writes)
static bool ln_fuse() {
  static const bool on = [] { const char* e = getenv("YOHO_LN_FUSE"); return !(e && e[0] == '0'); }();
  return on;
}

constexpr int kTensorsFixed = 7, kTensorsPerLayer = 15;

static int conv_out_len(int frames) { return (frames + 2 - 3) / 2 + 1; }
// Conv_0 as three taps over the bf16 mel (B, F, M): every tap's K is rounded up to the 64-wide k-block; the columns
// past M come out of the TMA as zeros (out of bounds), the matching weight columns are zero
static int conv0_tap_k(int n_mels) { return int(align_up(size_t(n_mels), 64)); }
static int conv0_k(int n_mels) { return 3 * conv0_tap_k(n_mels); }

// Offsets (in bytes) of everything inside the packed weight blob.
struct LayerOff {
  size_t ln0_s, ln0_b, wq, bq, wk, wv, bv, wo, bo, ln1_s, ln1_b, w1, b1, w2, b2;  // float32 originals
  size_t bqkv;                                                                   // float32 [bq | 0 | bv]
  size_t bqkv_ln, b1_ln;                                                         // float32: LayerNorm bias carried through the weights, beta @ W + bias
  size_t wqkv_h, wo_h, w1_h, w2_h;                                               // bf16 (out, in); fused: wqkv_h / w1_h carry the LN scale, column means removed
};
struct BlobLayout {
  size_t conv0_w, conv0_b, conv1_w, conv1_b, pos, lnf_s, lnf_b;
  size_t conv0_w_h, conv1_w_h;
  std::vector<LayerOff> layers;
  size_t total = 0;
};

static size_t tensor_elems(const yoho_enc_desc& d, int idx) {
  const size_t D = d.dims, M = d.n_mels;
  if (idx < kTensorsFixed) {
    switch (idx) {
      case 0: return 3 * M * D;
      case 2: return 3 * D * D;
      case 4: return size_t(d.seq_len) * D;
      default: return D;
    }
  }
  switch ((idx - kTensorsFixed) % kTensorsPerLayer) {
    case 2: case 4: case 5: case 7: return D * D;
    case 11: case 13: return 4 * D * D;
    case 12: return 4 * D;
    default: return D;
  }
}

static BlobLayout plan_blob(const yoho_enc_desc& d) {
  BlobLayout L;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 256); size_t o = off; off += bytes; return o; };
  const size_t D = d.dims, f = sizeof(float), h = sizeof(__nv_bfloat16);
  L.conv0_w = take(tensor_elems(d, 0) * f); L.conv0_b = take(D * f);
  L.conv1_w = take(tensor_elems(d, 2) * f); L.conv1_b = take(D * f);
  L.pos = take(tensor_elems(d, 4) * f);
  L.lnf_s = take(D * f); L.lnf_b = take(D * f);
  L.layers.resize(d.n_layers);
  for (auto& l : L.layers) {
    l.ln0_s = take(D * f); l.ln0_b = take(D * f);
    l.wq = take(D * D * f); l.bq = take(D * f);
    l.wk = take(D * D * f);
    l.wv = take(D * D * f); l.bv = take(D * f);
    l.wo = take(D * D * f); l.bo = take(D * f);
    l.ln1_s = take(D * f); l.ln1_b = take(D * f);
    l.w1 = take(4 * D * D * f); l.b1 = take(4 * D * f);
    l.w2 = take(4 * D * D * f); l.b2 = take(D * f);
  }
  L.conv0_w_h = take(D * conv0_k(d.n_mels) * h);
  L.conv1_w_h = take(D * 3 * D * h);
  for (auto& l : L.layers) {
    l.bqkv = take(3 * D * f);
    l.bqkv_ln = take(3 * D * f);
    l.b1_ln = take(4 * D * f);
    l.wqkv_h = take(3 * D * D * h);
    l.wo_h = take(D * D * h);
    l.w1_h = take(4 * D * D * h);
    l.w2_h = take(4 * D * D * h);
  }
  L.total = align_up(off, 1024);
  return L;
}

static int check_desc(const yoho_enc_desc* d, const char* who) {
  if (!d) return fail(YOHO_ERR_INVALID, "%s: desc is null", who);
  if (d->n_mels <= 0 || d->dims <= 0 || d->n_heads <= 0 || d->n_layers < 0 || d->seq_len <= 0)
    return fail(YOHO_ERR_INVALID, "%s: desc has a non-positive field", who);
  if (d->dims != d->n_heads * 64)
    return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d, n_heads=%d: kernels need d_head = 64 (every Whisper size)", who, d->dims, d->n_heads);
  if (d->dims > 1280) return fail(YOHO_ERR_UNSUPPORTED, "%s: dims=%d > 1280", who, d->dims);
  if (d->n_mels % 8 != 0) return fail(YOHO_ERR_UNSUPPORTED, "%s: n_mels=%d must be a multiple of 8 (16-byte rows for the TMA of the conv stem)", who, d->n_mels);
  return YOHO_OK;
}

// ------------------------------------------------------------------------------ workspace plan
struct WsPlan {
  // bf16 path
  size_t a0 = 0, c1 = 0, x = 0, xn = 0, qkv = 0, att = 0, h = 0, scores = 0, lnst = 0;
  int score_clips = 0;
  size_t total = 0;
};

static WsPlan plan_ws(const yoho_enc_desc& d, int batch, int frames, int compute) {
  WsPlan w;
  const size_t B = batch, T = conv_out_len(frames), D = d.dims, Fp = 2 * T;
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 1024); size_t o = off; off += bytes; return o; };
  if (compute == YOHO_COMPUTE_BF16) {
    const size_t e = sizeof(__nv_bfloat16);
    w.x = take(B * T * D * sizeof(float));
    w.xn = take(B * T * D * e);
    w.qkv = take(B * T * 3 * D * e);
    w.att = take(B * T * D * e);
    const size_t h_bytes = B * T * 4 * D * e;
    const size_t a0_bytes = B * size_t(frames) * d.n_mels * e, c1_bytes = B * Fp * D * e;   // a0: bf16 copy of the mel
    const size_t stem = align_up(a0_bytes, 1024) + c1_bytes;
    w.h = take(h_bytes > stem ? h_bytes : stem);  // stem buffers are dead before the first Dense_0
    w.a0 = w.h;
    w.c1 = w.h + align_up(a0_bytes, 1024);
    w.lnst = take(B * T * size_t(gemm_ln_parts(d.dims)) * sizeof(float2));   // per-row {sum, sum of squares} partials of the residual stream
  } else {
    const size_t e = sizeof(float);
    w.x = take(B * T * D * e);
    w.xn = take(B * T * D * e);
    w.qkv = take(B * T * 3 * D * e);
    w.att = take(B * T * D * e);
    const size_t h_bytes = B * T * 4 * D * e, c1_bytes = B * size_t(frames) * D * e;
    w.h = take(h_bytes > c1_bytes ? h_bytes : c1_bytes);
    w.c1 = w.h;
    const size_t per_clip = size_t(d.n_heads) * T * T * e;
    size_t clips = (size_t(1) << 30) / (per_clip ? per_clip : 1);
    if (clips < 1) clips = 1;
    if (clips > B) clips = B;
    if (clips * d.n_heads > 65535) clips = 65535 / d.n_heads;
    w.score_clips = int(clips);
    w.scores = take(clips * per_clip);
  }
  w.total = align_up(off, 1024);
  return w;
}

// ------------------------------------------------------------------------------ forward, bf16
#define YOHO_TRY(expr) do { int _r = (expr); if (_r != YOHO_OK) return _r; } while (0)

static int forward_bf16(yoho_ctx* ctx, const yoho_enc_desc& d, const BlobLayout& L, const char* blob, const float* mel,
                        int batch, int frames, void* out, int out_dtype, char* ws, const WsPlan& w, cudaStream_t s) {
  using bf = __nv_bfloat16;
  const int T = conv_out_len(frames), D = d.dims, Fp = 2 * T, Kt = conv0_tap_k(d.n_mels);
  const int rows = batch * T;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  auto H16 = [&](size_t off) { return reinterpret_cast<const bf*>(blob + off); };
  bf* a0 = reinterpret_cast<bf*>(ws + w.a0);
  bf* c1 = reinterpret_cast<bf*>(ws + w.c1);
  float* x = reinterpret_cast<float*>(ws + w.x);
  bf* xn = reinterpret_cast<bf*>(ws + w.xn);
  bf* qkv = reinterpret_cast<bf*>(ws + w.qkv);
  bf* att = reinterpret_cast<bf*>(ws + w.att);
  bf* hbuf = reinterpret_cast<bf*>(ws + w.h);
  float2* lnst = reinterpret_cast<float2*>(ws + w.lnst);
  const bool fuse = ln_fuse();

  // Conv_0 + GELU (whisper.py:125-126) as an implicit GEMM: three shifted TMA boxes x[t-1], x[t], x[t+1] over the bf16 mel
  // (B, F, M); padding=1 is the TMA's out-of-bounds zero fill (row -1 and row F of each clip), as is the K padding of each
  // tap (columns M..Kt-1).  Output rows F..Fp-1 (the pad row Conv_1 reads) are stored as zeros.
  YOHO_TRY(launch_cast_bf16(mel, a0, size_t(batch) * frames * d.n_mels, s));
  {
    GemmBf16Args g;
    g.a = a0; g.a_ld = d.n_mels; g.a_cols = d.n_mels; g.batch = batch; g.rows_per_batch = Fp; g.a_rows = frames;
    g.w = H16(L.conv0_w_h); g.n = D; g.k_per_tap = Kt; g.n_taps = 3; g.flop_k = 3 * d.n_mels;
    g.taps[0] = {0, -1}; g.taps[1] = {0, 0}; g.taps[2] = {0, 1};
    g.out_bf16 = c1; g.ldc = D; g.bias = F32(L.conv0_b); g.gelu = 1;
    g.mask_period = Fp; g.mask_valid = frames;
    YOHO_TRY(launch_gemm_bf16(ctx, g, s));
  }
  // Conv_1 (k3, stride 2) + GELU + sinusoid table (whisper.py:127-129) on the row-pair view (B, T, 2D)
  {
    GemmBf16Args g;
    g.a = c1; g.a_ld = 2 * D; g.a_cols = 2 * D; g.batch = batch; g.rows_per_batch = T;
    g.w = H16(L.conv1_w_h); g.n = D; g.k_per_tap = D; g.n_taps = 3;
    g.taps[0] = {D, -1};  // x[2t-1] = second half of pair t-1 (row -1 -> TMA zero fill = the padding)
    g.taps[1] = {0, 0};   // x[2t]
    g.taps[2] = {D, 0};   // x[2t+1] (row 2T-1 >= frames is the zeroed pad row)
    g.out_f32 = x; g.ldc = D; g.bias = F32(L.conv1_b); g.gelu = 1;
    g.pos = F32(L.pos); g.pos_period = T;
    if (fuse && d.n_layers > 0) { g.xb_out = xn; g.xb_ld = D; g.ln_stats_out = lnst; }   // feeds LayerNorm_0 of the first block
    YOHO_TRY(launch_gemm_bf16(ctx, g, s));
  }
  // Fused: the LayerNorms of the blocks (whisper.py:77,82) have no kernel of their own.  Every GEMM that writes the residual
  // stream x (float32) also writes xn = bf16(x) and per-row {sum, sum of squares} partials; the GEMM that consumes LN(x)
  // runs on xn with weights that carry the LN scale (column means removed) and finishes the normalisation in its epilogue.
  const int parts = gemm_ln_parts(D);
  for (int i = 0; fuse && i < d.n_layers; ++i) {
    const LayerOff& l = L.layers[i];
    {   // q | k | v of LN(x)                         whisper.py:76-78, 43-46
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.wqkv_h); g.n = 3 * D; g.k_per_tap = D;
      g.out_bf16 = qkv; g.ldc = 3 * D; g.bias = F32(l.bqkv_ln);
      g.ln_stats_in = lnst; g.ln_parts_in = parts; g.ln_dims = D;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    YOHO_TRY(launch_attention_bf16(ctx, qkv, att, batch, T, d.n_heads, s));
    {   // x += out_proj(attention), and the inputs of LayerNorm_1     whisper.py:50, 79
      GemmBf16Args g;
      g.a = att; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.wo_h); g.n = D; g.k_per_tap = D;
      g.out_f32 = x; g.ldc = D; g.bias = F32(l.bo);
      g.add_resid = 1; g.xb_out = xn; g.xb_ld = D; g.ln_stats_out = lnst;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    {   // gelu(Dense_0(LN(x)))                       whisper.py:82-84
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.w1_h); g.n = 4 * D; g.k_per_tap = D;
      g.out_bf16 = hbuf; g.ldc = 4 * D; g.bias = F32(l.b1_ln); g.gelu = 1;
      g.ln_stats_in = lnst; g.ln_parts_in = parts; g.ln_dims = D;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    {   // x += Dense_1(...)                          whisper.py:85-86
      GemmBf16Args g;
      g.a = hbuf; g.a_ld = 4 * D; g.a_cols = 4 * D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.w2_h); g.n = D; g.k_per_tap = 4 * D;
      g.out_f32 = x; g.ldc = D; g.bias = F32(l.b2);
      if (i + 1 < d.n_layers) { g.add_resid = 1; g.xb_out = xn; g.xb_ld = D; g.ln_stats_out = lnst; }   // the next block's LayerNorm_0
      else g.accumulate = 1;                                               // last block: the final LayerNorm reads x itself
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
  }
  for (int i = 0; !fuse && i < d.n_layers; ++i) {
    const LayerOff& l = L.layers[i];
    // x += out_proj(attention(LN(x)))          whisper.py:76-79, 37-67
    YOHO_TRY(launch_layernorm_bf16(x, F32(l.ln0_s), F32(l.ln0_b), xn, rows, D, s));
    {
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.wqkv_h); g.n = 3 * D; g.k_per_tap = D;
      g.out_bf16 = qkv; g.ldc = 3 * D; g.bias = F32(l.bqkv);
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    YOHO_TRY(launch_attention_bf16(ctx, qkv, att, batch, T, d.n_heads, s));
    {
      GemmBf16Args g;
      g.a = att; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.wo_h); g.n = D; g.k_per_tap = D;
      g.out_f32 = x; g.ldc = D; g.bias = F32(l.bo); g.accumulate = 1;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    // x += Dense_1(gelu(Dense_0(LN(x))))       whisper.py:81-86
    YOHO_TRY(launch_layernorm_bf16(x, F32(l.ln1_s), F32(l.ln1_b), xn, rows, D, s));
    {
      GemmBf16Args g;
      g.a = xn; g.a_ld = D; g.a_cols = D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.w1_h); g.n = 4 * D; g.k_per_tap = D;
      g.out_bf16 = hbuf; g.ldc = 4 * D; g.bias = F32(l.b1); g.gelu = 1;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    {
      GemmBf16Args g;
      g.a = hbuf; g.a_ld = 4 * D; g.a_cols = 4 * D; g.batch = 1; g.rows_per_batch = rows;
      g.w = H16(l.w2_h); g.n = D; g.k_per_tap = 4 * D;
      g.out_f32 = x; g.ldc = D; g.bias = F32(l.b2); g.accumulate = 1;
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
  }
  // final LayerNorm (whisper.py:134); bf16 states are rounded here, once
  if (out_dtype == YOHO_DTYPE_BF16) return launch_layernorm_bf16(x, F32(L.lnf_s), F32(L.lnf_b), static_cast<bf*>(out), rows, D, s);
  return launch_layernorm_f32(x, F32(L.lnf_s), F32(L.lnf_b), static_cast<float*>(out), rows, D, s);
}

// ------------------------------------------------------------------------------ forward, fp32
static int forward_fp32(const yoho_enc_desc& d, const BlobLayout& L, const char* blob, const float* mel, int batch,
                        int frames, float* out, char* ws, const WsPlan& w, cudaStream_t s) {
  const int T = conv_out_len(frames), D = d.dims, H = d.n_heads, M = d.n_mels;
  const int rows = batch * T;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  float* c1 = reinterpret_cast<float*>(ws + w.c1);
  float* x = reinterpret_cast<float*>(ws + w.x);
  float* xn = reinterpret_cast<float*>(ws + w.xn);
  float* qkv = reinterpret_cast<float*>(ws + w.qkv);
  float* att = reinterpret_cast<float*>(ws + w.att);
  float* hbuf = reinterpret_cast<float*>(ws + w.h);
  float* scores = reinterpret_cast<float*>(ws + w.scores);

  auto dense = [&](const float* a, int lda, int k, const float* wt, int n, const float* bias, float* c, int ldc, int gelu, int acc) {
    SgemmArgs g;
    g.a = a; g.b = wt; g.c = c; g.m = rows; g.n = n; g.k = k; g.lda = lda; g.ldb = n; g.ldc = ldc;
    g.bias = bias; g.gelu = gelu; g.accumulate = acc;
    return launch_sgemm(g, s);
  };
  // the launch grid's y extent caps rows per call at 65535 * 128; split the conv over clips to be safe
  const int clip_step = 2048;
  for (int b0 = 0; b0 < batch; b0 += clip_step) {
    const int nb = batch - b0 < clip_step ? batch - b0 : clip_step;
    SgemmArgs g;  // Conv_0 + GELU
    g.a = mel + size_t(b0) * frames * M; g.b = F32(L.conv0_w); g.c = c1 + size_t(b0) * frames * D;
    g.m = nb * frames; g.n = D; g.k = 3 * M; g.ldb = D; g.ldc = D;
    g.conv = 1; g.conv_len = frames; g.conv_out = frames; g.conv_ch = M; g.conv_stride = 1;
    g.bias = F32(L.conv0_b); g.gelu = 1;
    YOHO_TRY(launch_sgemm(g, s));
    SgemmArgs g2;  // Conv_1 + GELU + sinusoid table
    g2.a = c1 + size_t(b0) * frames * D; g2.b = F32(L.conv1_w); g2.c = x + size_t(b0) * T * D;
    g2.m = nb * T; g2.n = D; g2.k = 3 * D; g2.ldb = D; g2.ldc = D;
    g2.conv = 1; g2.conv_len = frames; g2.conv_out = T; g2.conv_ch = D; g2.conv_stride = 2;
    g2.bias = F32(L.conv1_b); g2.gelu = 1; g2.pos = F32(L.pos); g2.pos_period = T;
    YOHO_TRY(launch_sgemm(g2, s));
  }
  if (rows > 65535 * 128) return fail(YOHO_ERR_UNSUPPORTED, "fp32 encoder: %d rows exceed one launch", rows);
  for (int i = 0; i < d.n_layers; ++i) {
    const LayerOff& l = L.layers[i];
    YOHO_TRY(launch_layernorm_f32(x, F32(l.ln0_s), F32(l.ln0_b), xn, rows, D, s));
    YOHO_TRY(dense(xn, D, D, F32(l.wq), D, F32(l.bq), qkv, 3 * D, 0, 0));
    YOHO_TRY(dense(xn, D, D, F32(l.wk), D, nullptr, qkv + D, 3 * D, 0, 0));
    YOHO_TRY(dense(xn, D, D, F32(l.wv), D, F32(l.bv), qkv + 2 * D, 3 * D, 0, 0));
    for (int b0 = 0; b0 < batch; b0 += w.score_clips) {
      const int nb = batch - b0 < w.score_clips ? batch - b0 : w.score_clips;
      const float* qb = qkv + size_t(b0) * T * 3 * D;
      SgemmArgs g;  // scores = (q k^T) / 8   (whisper.py:54-60: both operands scaled by 64^-0.25)
      g.a = qb; g.b = qb + D; g.c = scores; g.m = T; g.n = T; g.k = 64; g.lda = 3 * D; g.ldb = 3 * D; g.ldc = T;
      g.b_transposed = 1; g.alpha = 0.125f; g.nz1 = nb; g.nz2 = H;
      g.sa1 = (long long)T * 3 * D; g.sa2 = 64; g.sb1 = g.sa1; g.sb2 = 64; g.sc1 = (long long)H * T * T; g.sc2 = (long long)T * T;
      YOHO_TRY(launch_sgemm(g, s));
      YOHO_TRY(launch_softmax_rows(scores, (long long)nb * H * T, T, s));
      SgemmArgs g2;  // out = w v, heads merged (whisper.py:65)
      g2.a = scores; g2.b = qb + 2 * D; g2.c = att + size_t(b0) * T * D; g2.m = T; g2.n = 64; g2.k = T;
      g2.lda = T; g2.ldb = 3 * D; g2.ldc = D; g2.nz1 = nb; g2.nz2 = H;
      g2.sa1 = g.sc1; g2.sa2 = g.sc2; g2.sb1 = g.sa1; g2.sb2 = 64; g2.sc1 = (long long)T * D; g2.sc2 = 64;
      YOHO_TRY(launch_sgemm(g2, s));
    }
    YOHO_TRY(dense(att, D, D, F32(l.wo), D, F32(l.bo), x, D, 0, 1));
    YOHO_TRY(launch_layernorm_f32(x, F32(l.ln1_s), F32(l.ln1_b), xn, rows, D, s));
    YOHO_TRY(dense(xn, D, D, F32(l.w1), 4 * D, F32(l.b1), hbuf, 4 * D, 1, 0));
    YOHO_TRY(dense(hbuf, 4 * D, 4 * D, F32(l.w2), D, F32(l.b2), x, D, 0, 1));
  }
  return launch_layernorm_f32(x, F32(L.lnf_s), F32(L.lnf_b), out, rows, D, s);
}

}  // namespace yoho

using namespace yoho;

extern "C" int yoho_encoder_num_tensors(const yoho_enc_desc* desc) {
  if (!desc || desc->n_layers < 0) return fail(YOHO_ERR_INVALID, "yoho_encoder_num_tensors: bad desc");
  return kTensorsFixed + kTensorsPerLayer * desc->n_layers;
}

extern "C" long long yoho_encoder_tensor_elems(const yoho_enc_desc* desc, int idx) {
  if (check_desc(desc, "yoho_encoder_tensor_elems") != YOHO_OK) return YOHO_ERR_INVALID;
  if (idx < 0 || idx >= kTensorsFixed + kTensorsPerLayer * desc->n_layers) return fail(YOHO_ERR_INVALID, "yoho_encoder_tensor_elems: index %d", idx);
  return (long long)tensor_elems(*desc, idx);
}

extern "C" size_t yoho_encoder_weights_bytes(const yoho_enc_desc* desc) {
  if (check_desc(desc, "yoho_encoder_weights_bytes") != YOHO_OK) return 0;
  return plan_blob(*desc).total;
}

extern "C" int yoho_encoder_out_len(int frames) {
  if (frames <= 0) return fail(YOHO_ERR_INVALID, "yoho_encoder_out_len: frames=%d", frames);
  return conv_out_len(frames);
}

extern "C" int yoho_encoder_weights_pack(yoho_ctx* ctx, const yoho_enc_desc* desc, const float* const* tensors,
                                         int n_tensors, void* blob_, size_t blob_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !tensors || !blob_) return fail(YOHO_ERR_INVALID, "yoho_encoder_weights_pack: null argument");
  YOHO_TRY(check_desc(desc, "yoho_encoder_weights_pack"));
  const yoho_enc_desc& d = *desc;
  if (n_tensors != kTensorsFixed + kTensorsPerLayer * d.n_layers)
    return fail(YOHO_ERR_INVALID, "yoho_encoder_weights_pack: %d tensors given, %d expected", n_tensors, kTensorsFixed + kTensorsPerLayer * d.n_layers);
  for (int i = 0; i < n_tensors; ++i)
    if (!tensors[i]) return fail(YOHO_ERR_INVALID, "yoho_encoder_weights_pack: tensor %d is null", i);
  BlobLayout L = plan_blob(d);
  if (blob_bytes < L.total) return fail(YOHO_ERR_WORKSPACE, "yoho_encoder_weights_pack: blob %zu < %zu bytes", blob_bytes, L.total);
  char* blob = static_cast<char*>(blob_);
  const int D = d.dims;
  auto put = [&](size_t off, int idx) -> int {
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + off, tensors[idx], tensor_elems(d, idx) * sizeof(float), cudaMemcpyDefault, s));
    return YOHO_OK;
  };
  YOHO_TRY(put(L.conv0_w, 0)); YOHO_TRY(put(L.conv0_b, 1)); YOHO_TRY(put(L.conv1_w, 2)); YOHO_TRY(put(L.conv1_b, 3));
  YOHO_TRY(put(L.pos, 4)); YOHO_TRY(put(L.lnf_s, 5)); YOHO_TRY(put(L.lnf_b, 6));
  for (int i = 0; i < d.n_layers; ++i) {
    const LayerOff& l = L.layers[i];
    const size_t offs[kTensorsPerLayer] = {l.ln0_s, l.ln0_b, l.wq, l.bq, l.wk, l.wv, l.bv, l.wo, l.bo, l.ln1_s, l.ln1_b, l.w1, l.b1, l.w2, l.b2};
    for (int j = 0; j < kTensorsPerLayer; ++j) YOHO_TRY(put(offs[j], kTensorsFixed + i * kTensorsPerLayer + j));
  }
  using bf = __nv_bfloat16;
  auto F32 = [&](size_t off) { return reinterpret_cast<const float*>(blob + off); };
  auto H16 = [&](size_t off) { return reinterpret_cast<bf*>(blob + off); };
  // Flax kernels are (in, out) / (k, in, out); the tensor-core path wants (out, in) with `in` contiguous.
  for (int tap = 0; tap < 3; ++tap)   // (D, 3 * Kt): tap-major, each tap's n_mels columns followed by zeros up to Kt
    YOHO_TRY(launch_transpose_cast(F32(L.conv0_w) + size_t(tap) * d.n_mels * D, H16(L.conv0_w_h) + size_t(tap) * conv0_tap_k(d.n_mels),
                                   d.n_mels, D, conv0_k(d.n_mels), s, conv0_tap_k(d.n_mels)));
  YOHO_TRY(launch_transpose_cast(F32(L.conv1_w), H16(L.conv1_w_h), 3 * D, D, 3 * D, s));
  for (int i = 0; i < d.n_layers; ++i) {
    const LayerOff& l = L.layers[i];
    // fused bias [bq | 0 | bv]: key_proj has use_bias=False (whisper.py:33)
    YOHO_TRY(launch_fill_zero(blob + l.bqkv, size_t(3) * D * sizeof(float), s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.bqkv, blob + l.bq, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(blob + l.bqkv + size_t(2) * D * sizeof(float), blob + l.bv, D * sizeof(float), cudaMemcpyDeviceToDevice, s));
    if (ln_fuse()) {
      // LayerNorm_0 / LayerNorm_1 (whisper.py:77,82) are folded into the GEMMs that consume them: the weights carry the LN
      // scale and have their column means removed (the centring), the epilogue applies rstd, LN bias and Dense bias merge
      auto Fp = [&](size_t off) { return reinterpret_cast<float*>(blob + off); };
      YOHO_TRY(launch_ln_fold(F32(l.wq), F32(l.ln0_s), F32(l.ln0_b), F32(l.bqkv), H16(l.wqkv_h), Fp(l.bqkv_ln), D, D, s));
      YOHO_TRY(launch_ln_fold(F32(l.wk), F32(l.ln0_s), F32(l.ln0_b), F32(l.bqkv) + D, H16(l.wqkv_h) + size_t(D) * D, Fp(l.bqkv_ln) + D, D, D, s));
      YOHO_TRY(launch_ln_fold(F32(l.wv), F32(l.ln0_s), F32(l.ln0_b), F32(l.bqkv) + 2 * D, H16(l.wqkv_h) + size_t(2) * D * D, Fp(l.bqkv_ln) + 2 * D, D, D, s));
      YOHO_TRY(launch_ln_fold(F32(l.w1), F32(l.ln1_s), F32(l.ln1_b), F32(l.b1), H16(l.w1_h), Fp(l.b1_ln), D, 4 * D, s));
    } else {
      YOHO_TRY(launch_transpose_cast(F32(l.wq), H16(l.wqkv_h), D, D, D, s));
      YOHO_TRY(launch_transpose_cast(F32(l.wk), H16(l.wqkv_h) + size_t(D) * D, D, D, D, s));
      YOHO_TRY(launch_transpose_cast(F32(l.wv), H16(l.wqkv_h) + size_t(2) * D * D, D, D, D, s));
      YOHO_TRY(launch_transpose_cast(F32(l.w1), H16(l.w1_h), D, 4 * D, D, s));
    }
    YOHO_TRY(launch_transpose_cast(F32(l.wo), H16(l.wo_h), D, D, D, s));
    YOHO_TRY(launch_transpose_cast(F32(l.w2), H16(l.w2_h), 4 * D, D, 4 * D, s));
  }
  return YOHO_OK;
}

extern "C" size_t yoho_encoder_workspace_bytes(const yoho_enc_desc* desc, int batch, int frames, int compute) {
  if (check_desc(desc, "yoho_encoder_workspace_bytes") != YOHO_OK) return 0;
  if (batch <= 0 || frames <= 0) return 1024;
  return plan_ws(*desc, batch, frames, compute).total;
}

extern "C" int yoho_encoder_fwd(yoho_ctx* ctx, const yoho_enc_desc* desc, const void* blob, size_t blob_bytes, const float* mel,
                                int batch, int frames, void* out, int out_dtype, int compute, void* ws, size_t ws_bytes,
                                yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !blob || !mel || !out) return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: null argument");
  YOHO_TRY(check_desc(desc, "yoho_encoder_fwd"));
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "yoho_encoder_fwd: device is sm_%d%d, kernels are sm_100a only", ctx->cc_major, ctx->cc_minor);
  if (compute != YOHO_COMPUTE_FP32 && compute != YOHO_COMPUTE_BF16) return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: compute=%d", compute);
  if (out_dtype != YOHO_DTYPE_F32 && out_dtype != YOHO_DTYPE_BF16) return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: out_dtype=%d", out_dtype);
  if (out_dtype == YOHO_DTYPE_BF16 && compute != YOHO_COMPUTE_BF16)
    return fail(YOHO_ERR_UNSUPPORTED, "yoho_encoder_fwd: bf16 states are produced by the bf16 compute path only");
  if (batch < 0 || frames <= 0) return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: batch=%d frames=%d", batch, frames);
  const int T = conv_out_len(frames);
  if (T != desc->seq_len)
    return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: %d frames give %d positions but the sinusoid table has %d rows (whisper.py:23 broadcasts)", frames, T, desc->seq_len);
  if (batch == 0) return YOHO_OK;
  if ((long long)batch * 2 * T > 0x7fffffffLL / 8) return fail(YOHO_ERR_UNSUPPORTED, "yoho_encoder_fwd: batch too large for 32-bit row indices");
  WsPlan w = plan_ws(*desc, batch, frames, compute);
  if (!ws || ws_bytes < w.total) return fail(YOHO_ERR_WORKSPACE, "yoho_encoder_fwd: workspace %zu < %zu bytes", ws_bytes, w.total);
  if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0 || reinterpret_cast<uintptr_t>(blob) % 256 != 0)
    return fail(YOHO_ERR_INVALID, "yoho_encoder_fwd: workspace must be 1024-byte and blob 256-byte aligned");
  BlobLayout L = plan_blob(*desc);
  if (blob_bytes < L.total)
    return fail(YOHO_ERR_WORKSPACE, "yoho_encoder_fwd: blob of %zu bytes, this descriptor's layout needs %zu (packed for another shape?)", blob_bytes, L.total);
  if (compute == YOHO_COMPUTE_BF16)
    return forward_bf16(ctx, *desc, L, static_cast<const char*>(blob), mel, batch, frames, out, out_dtype, static_cast<char*>(ws), w, s);
  return forward_fp32(*desc, L, static_cast<const char*>(blob), mel, batch, frames, static_cast<float*>(out), static_cast<char*>(ws), w, s);
}

// ------------------------------------------------------------------------------ test hooks
extern "C" size_t yoho_test_gemm_workspace_bytes(int m, int n, int k) {
  if (m <= 0 || n <= 0 || k <= 0) return 1024;
  return align_up(size_t(m) * k * 2, 1024) + align_up(size_t(n) * k * 2, 1024) + 1024;
}

extern "C" int yoho_test_gemm_bf16(yoho_ctx* ctx, const float* a, const float* w, const float* bias, float* c, int m, int n,
                                   int k, int epilogue, void* ws, size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !a || !w || !c || !ws) return fail(YOHO_ERR_INVALID, "yoho_test_gemm_bf16: null argument");
  if (ws_bytes < yoho_test_gemm_workspace_bytes(m, n, k)) return fail(YOHO_ERR_WORKSPACE, "yoho_test_gemm_bf16: workspace too small");
  using bf = __nv_bfloat16;
  bf* a_h = static_cast<bf*>(ws);
  bf* w_h = reinterpret_cast<bf*>(static_cast<char*>(ws) + align_up(size_t(m) * k * 2, 1024));
  YOHO_TRY(launch_cast_bf16(a, a_h, size_t(m) * k, s));
  YOHO_TRY(launch_transpose_cast(w, w_h, k, n, k, s));
  GemmBf16Args g;
  g.a = a_h; g.a_ld = k; g.a_cols = k; g.batch = 1; g.rows_per_batch = m;
  g.w = w_h; g.n = n; g.k_per_tap = k;
  g.out_f32 = c; g.ldc = n; g.bias = bias; g.gelu = epilogue == 1;
  return launch_gemm_bf16(ctx, g, s);
}

extern "C" size_t yoho_test_attention_workspace_bytes(int batch, int t, int heads) {
  if (batch <= 0 || t <= 0 || heads <= 0) return 1024;
  size_t rows = size_t(batch) * t, d = size_t(heads) * 64;
  return align_up(rows * 3 * d * 2, 1024) + align_up(rows * d * 2, 1024) + 1024;
}

extern "C" int yoho_test_attention_bf16(yoho_ctx* ctx, const float* qkv, float* out, int batch, int t, int heads, void* ws,
                                        size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !qkv || !out || !ws) return fail(YOHO_ERR_INVALID, "yoho_test_attention_bf16: null argument");
  if (ws_bytes < yoho_test_attention_workspace_bytes(batch, t, heads)) return fail(YOHO_ERR_WORKSPACE, "yoho_test_attention_bf16: workspace too small");
  using bf = __nv_bfloat16;
  const size_t rows = size_t(batch) * t, d = size_t(heads) * 64;
  bf* qkv_h = static_cast<bf*>(ws);
  bf* out_h = reinterpret_cast<bf*>(static_cast<char*>(ws) + align_up(rows * 3 * d * 2, 1024));
  YOHO_TRY(launch_cast_bf16(qkv, qkv_h, rows * 3 * d, s));
  YOHO_TRY(launch_attention_bf16(ctx, qkv_h, out_h, batch, t, heads, s));
  return launch_cast_f32(out_h, out, rows * d, s);
}
