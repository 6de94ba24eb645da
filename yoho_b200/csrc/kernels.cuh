// Internal launch interface between the encoder driver (encoder.cu) and the kernel files.
#pragma once
#include "common.cuh"

namespace yoho {

// ---------------------------------------------------------------- elementwise.cu
// LayerNorm over the last axis (flax nn.LayerNorm: eps 1e-6, var = E[x^2]-E[x]^2 >= 0).
int launch_layernorm_f32(const float* x, const float* scale, const float* bias, float* y, int rows, int dims, cudaStream_t s);
int launch_layernorm_bf16(const float* x, const float* scale, const float* bias, __nv_bfloat16* y, int rows, int dims, cudaStream_t s);
// dst (n, k) bf16 = transpose of src (k, n) float32; dst row stride ld_dst (>= k), tail zero-filled.
int launch_transpose_cast(const float* src, __nv_bfloat16* dst, int k, int n, int ld_dst, cudaStream_t s, int k_pad = -1);   // k_pad >= 0: zero only columns k..k_pad-1
int launch_cast_bf16(const float* src, __nv_bfloat16* dst, size_t n, cudaStream_t s);
int launch_cast_f32(const __nv_bfloat16* src, float* dst, size_t n, cudaStream_t s);
int launch_fill_zero(void* p, size_t bytes, cudaStream_t s);

// ---------------------------------------------------------------- yoho_ops.cu (YOHO's own encoder: model.py / layers.py)
// y = [silu](x * rsqrt(mean(x^2) + 1e-6) * scale), one row at a time
int launch_rmsnorm_f32(const float* x, const float* scale, float* y, int rows, int dims, int silu, cudaStream_t s);
int launch_rmsnorm_bf16(const float* x, const float* scale, __nv_bfloat16* y, int rows, int dims, int silu, cudaStream_t s);
// (batch, in_len, dims) -> (batch, in_len / 2, dims): silu(rmsnorm(x)) then max over time pairs
int launch_rmsnorm_silu_pool_f32(const float* x, const float* scale, float* y, int batch, int in_len, int dims, cudaStream_t s);
int launch_rmsnorm_silu_pool_bf16(const float* x, const float* scale, __nv_bfloat16* y, int batch, int in_len, int dims, cudaStream_t s);
// RoPE in place on the first n_rot_heads 64-wide heads of each row; table (len, 2, 32) float32 {cos, sin}
int launch_rope_f32(float* qkv, const float* table, long long rows, int seq, int ld, int n_rot_heads, cudaStream_t s);
int launch_rope_bf16(__nv_bfloat16* qkv, const float* table, long long rows, int seq, int ld, int n_rot_heads, cudaStream_t s);
// g = silu(h[:, :hidden]) * h[:, hidden:]
int launch_swiglu_gate_f32(const float* h, float* g, long long rows, int hidden, cudaStream_t s);
int launch_swiglu_gate_bf16(const __nv_bfloat16* h, __nv_bfloat16* g, long long rows, int hidden, cudaStream_t s);
// dst (d_in, kv_heads*64; row stride ld_dst) = sum_g src (d_in, kv_heads, groups, 64)
int launch_group_sum(const float* src, float* dst, int d_in, int kv_heads, int groups, int ld_dst, cudaStream_t s);

// ---------------------------------------------------------------- encoder_fp32.cu (SIMT, 1e-4 path)
struct SgemmArgs {
  const float* a = nullptr; const float* b = nullptr; float* c = nullptr;
  int m = 0, n = 0, k = 0;
  int lda = 0, ldb = 0, ldc = 0;
  int b_transposed = 0;            // 0: b is (k, n) row-major; 1: b is (n, k) row-major
  // conv mode (a is (batch, conv_len, conv_ch)): row -> (clip, t), k -> (tap, ch), zero padding 1
  int conv = 0, conv_len = 0, conv_out = 0, conv_ch = 0, conv_stride = 1;
  // batching over blockIdx.z = z1 * nz2 + z2
  int nz1 = 1, nz2 = 1;
  long long sa1 = 0, sa2 = 0, sb1 = 0, sb2 = 0, sc1 = 0, sc2 = 0;
  // epilogue: v = alpha*acc + bias[n]; gelu; + pos[(row % pos_period), n]; c (+)= v
  float alpha = 1.0f;
  const float* bias = nullptr;
  int gelu = 0;
  const float* pos = nullptr; int pos_period = 1;
  int accumulate = 0;
};
int launch_sgemm(const SgemmArgs& a, cudaStream_t s);
// in-place softmax over the last axis; causal_period > 0: row r only sees columns <= r % causal_period (the
// triu(-inf, 1) mask of whisper.py:157), the rest become 0
int launch_softmax_rows(float* x, long long rows, int cols, cudaStream_t s, int causal_period = 0);

// ---------------------------------------------------------------- decoder_ops.cu (token-by-token text decoder)
int launch_embed(const int* tokens, int tok_ld, const int* pos_ptr, int batch, int s_len, const float* emb, const float* pos,
                 float* x, int dims, int vocab, cudaStream_t s);
struct SkinnyArgs {
  const float* x = nullptr; int ldx = 0;            // (b, k) activations
  const float* w = nullptr; int ldw = 0;            // (k, n) row-major weights (the Flax kernel layout)
  const float* bias = nullptr;                      // (n) or null
  float* out = nullptr; long long out_row_stride = 0;
  const int* row_off_ptr = nullptr; int row_off_mul = 0;   // + *row_off_ptr * row_off_mul elements (append at the device-side position)
  // optional second destination: columns >= n_split go to out2 (column c - n_split); the row offset then applies to out2
  float* out2 = nullptr; long long out2_row_stride = 0; int n_split = 0;
  // optional LayerNorm of the activation rows on load (flax: eps 1e-6, var = E[x^2] - E[x]^2 >= 0): scale / bias are (k)
  const float* ln_scale = nullptr; const float* ln_bias = nullptr;
  int b = 0, k = 0, n = 0;
  int gelu = 0, accumulate = 0;
};
int launch_skinny_gemm(const SkinnyArgs& a, cudaStream_t s);
struct DecodeAttArgs {
  const float* q = nullptr; int ldq = 0;            // (batch, heads * 64)
  const float* kv = nullptr; long long kv_clip_stride = 0; int kv_ld = 0; int k_off = 0, v_off = 0;   // cache rows [.. k .. v ..]
  const int* len_ptr = nullptr; int len_add = 0; int len_const = 0;
  float* out = nullptr; int ldo = 0;
};
int launch_decode_attention(const DecodeAttArgs& a, int batch, int heads, int max_len, cudaStream_t s);
// pos_ptr[0] is the position, pos_ptr[1] a ticket counter: the last CTA to finish advances the position (saturating)
int launch_greedy_pick(const float* logits, int ld, int vocab, int* tokens, int tok_ld, int max_len, int batch, int* pos_ptr,
                       const unsigned char* forced, int eot, cudaStream_t s);

// ---------------------------------------------------------------- gemm_sm100.cu (tcgen05)
struct TapSpec { int col_off; int row_off; };
struct GemmBf16Args {
  // A: bf16, viewed as (batch, rows_per_batch, a_cols) row-major, row stride a_ld elements
  const __nv_bfloat16* a = nullptr; int a_ld = 0; int a_cols = 0;
  int batch = 1; int rows_per_batch = 0;
  int flop_k = 0;                  // contraction length the profiler counts as algorithmic work (0: n_taps * k_per_tap); Conv_0 pads each tap's K
  int a_rows = 0;                  // rows of A that exist per batch entry (0: rows_per_batch); rows beyond are TMA zero fill
  // W: bf16 (n, k_total) row-major ("K-major"), k_total = n_taps * k_per_tap (zero padded to 64)
  const __nv_bfloat16* w = nullptr; int n = 0; int k_per_tap = 0; int n_taps = 1;
  TapSpec taps[3] = {{0, 0}, {0, 0}, {0, 0}};
  // output: row index = clip * rows_per_batch + r, row stride ldc elements
  __nv_bfloat16* out_bf16 = nullptr; float* out_f32 = nullptr; int ldc = 0;
  const float* bias = nullptr;     // (n) or null
  int gelu = 0;
  const float* pos = nullptr;      // (rows_per_batch.., n): + pos[r % pos_period][col]
  int pos_period = 1;
  int accumulate = 0;              // out_f32 += value (residual stream), a TMA reduce-add performed in L2
  int mask_period = 0, mask_valid = 0;  // if mask_period: rows with (r % mask_period) >= mask_valid store 0
  // ---- LayerNorm folded into the GEMMs on either side of it (whisper.py:77,82: x + f(LN(x))) -------------------------
  // producer (float32 output = the residual stream x): with add_resid the epilogue fetches the old rows of out_f32 by TMA
  // into its staging buffer, adds them and stores the new rows (instead of the reduce-add in L2, which never shows the
  // sum to the SM); a bf16 copy of the new rows goes to xb_out (row stride xb_ld) and per-row partial {sum, sum of
  // squares} over each epilogue warp's columns to ln_stats_out[row * gemm_ln_parts(n) + part]
  int add_resid = 0;
  __nv_bfloat16* xb_out = nullptr; int xb_ld = 0;
  float2* ln_stats_out = nullptr;
  // consumer: A holds the RAW rows x in bf16; W carries the LayerNorm scale and has its column means removed (the centring
  // x - mean(x) is the projection I - 11^T/D, folded into the weights: (P x)^T W' = x^T (P W')), bias = beta W + b; the
  // epilogue applies value = rstd * acc + bias[col] with rstd from the ln_parts_in partials of the row
  const float2* ln_stats_in = nullptr; int ln_parts_in = 0; int ln_dims = 0;
};
int launch_gemm_bf16(yoho_ctx* ctx, const GemmBf16Args& a, cudaStream_t s);
// partials per row that a producer GEMM with n output columns writes to ln_stats_out (two epilogue warps per column tile)
int gemm_ln_parts(int n);
// weight preparation for the consumer: with w'[k][n] = scale[k] * src[k][n] and m[n] = mean_k w'[k][n],
// dst (n, k) bf16 = transpose of (w' - m);  bias_out[n] = (bias_in ? bias_in[n] : 0) + sum_k beta[k] * src[k][n]
int launch_ln_fold(const float* src, const float* scale, const float* beta, const float* bias_in, __nv_bfloat16* dst,
                   float* bias_out, int k, int n, cudaStream_t s);

// ---------------------------------------------------------------- attention_sm100.cu (tcgen05)
// qkv (batch, t, 3*dims) bf16 with q | k | v column blocks, heads of 64; out (batch, t, dims) bf16.
int launch_attention_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s);
// attention3_sm100.cu: three query tiles (384 rows) per work item; launch_attention_bf16 dispatches to it when that
// pads the sequence no further than 256-row items do (YOHO_ATT_TILES=2|3 forces one of them)
int launch_attention3_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s);

}  // namespace yoho
