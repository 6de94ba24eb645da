// Cross-attention key / value pre-projection of the encoder states (the step after the path):
// every DecoderBlock of the reference applies key_proj / value_proj of its second MultiHeadAttention to the same
// encoder states (yoho/src/nn/whisper.py:104 -> :45-46), and the reference recomputes them for every generated token
// (:274-281, "TODO: add KV cache" :27).  They depend on the clip only, so they are one data-parallel GEMM per
// decoder layer over (B * T, d_in), computed once here:
//     kv[l] = states @ [key_proj.kernel | value_proj.kernel] + [0 | value_proj.bias]        (key_proj has no bias, :33)
// The d_head^-0.25 scaling of k (:54-57) belongs to qkv_attention and is left to the consumer.
#include "kernels.cuh"

namespace yoho {

struct CkvOff { size_t wk, wv, bv, wkv, bias, wkv_h; };

static size_t ckv_layer_bytes(const yoho_crosskv_desc& d, CkvOff* o) {
  size_t off = 0;
  auto take = [&](size_t bytes) { off = align_up(off, 256); size_t r = off; off += bytes; return r; };
  const size_t I = d.d_in, O = d.d_out, f = sizeof(float);
  CkvOff t;
  t.wk = take(I * O * f); t.wv = take(I * O * f); t.bv = take(O * f);
  t.wkv = take(I * 2 * O * f); t.bias = take(2 * O * f); t.wkv_h = take(2 * O * I * sizeof(__nv_bfloat16));
  if (o) *o = t;
  return align_up(off, 1024);
}

static int ckv_check(const yoho_crosskv_desc* d, const char* who) {
  if (!d) return fail(YOHO_ERR_INVALID, "%s: desc is null", who);
  if (d->d_in <= 0 || d->d_out <= 0 || d->n_layers < 0) return fail(YOHO_ERR_INVALID, "%s: desc has a non-positive field", who);
  if (d->d_in % 64 != 0 || d->d_out % 16 != 0) return fail(YOHO_ERR_UNSUPPORTED, "%s: d_in=%d must be a multiple of 64, d_out=%d of 16", who, d->d_in, d->d_out);
  return YOHO_OK;
}

__global__ void ckv_copy_cols_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows, int cols, int ld_dst, int c0) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += gridDim.x * blockDim.x) {
    const int r = i / cols, c = i - r * cols;
    dst[size_t(r) * ld_dst + c0 + c] = src[i];
  }
}

}  // namespace yoho

using namespace yoho;

#define YOHO_TRY(expr) do { int _r = (expr); if (_r != YOHO_OK) return _r; } while (0)

extern "C" int yoho_crosskv_num_tensors(const yoho_crosskv_desc* desc) {
  if (!desc || desc->n_layers < 0) return fail(YOHO_ERR_INVALID, "yoho_crosskv_num_tensors: bad desc");
  return 3 * desc->n_layers;
}

extern "C" long long yoho_crosskv_tensor_elems(const yoho_crosskv_desc* desc, int idx) {
  if (ckv_check(desc, "yoho_crosskv_tensor_elems") != YOHO_OK) return YOHO_ERR_INVALID;
  if (idx < 0 || idx >= 3 * desc->n_layers) return fail(YOHO_ERR_INVALID, "yoho_crosskv_tensor_elems: index %d", idx);
  return idx % 3 == 2 ? (long long)desc->d_out : (long long)desc->d_in * desc->d_out;
}

extern "C" size_t yoho_crosskv_weights_bytes(const yoho_crosskv_desc* desc) {
  if (ckv_check(desc, "yoho_crosskv_weights_bytes") != YOHO_OK) return 0;
  return ckv_layer_bytes(*desc, nullptr) * size_t(desc->n_layers) + 1024;
}

extern "C" int yoho_crosskv_weights_pack(yoho_ctx* ctx, const yoho_crosskv_desc* desc, const float* const* tensors, int n_tensors,
                                         void* blob_, size_t blob_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !tensors || !blob_) return fail(YOHO_ERR_INVALID, "yoho_crosskv_weights_pack: null argument");
  YOHO_TRY(ckv_check(desc, "yoho_crosskv_weights_pack"));
  const yoho_crosskv_desc& d = *desc;
  if (n_tensors != 3 * d.n_layers) return fail(YOHO_ERR_INVALID, "yoho_crosskv_weights_pack: %d tensors given, %d expected", n_tensors, 3 * d.n_layers);
  for (int i = 0; i < n_tensors; ++i)
    if (!tensors[i]) return fail(YOHO_ERR_INVALID, "yoho_crosskv_weights_pack: tensor %d is null", i);
  if (blob_bytes < yoho_crosskv_weights_bytes(desc)) return fail(YOHO_ERR_WORKSPACE, "yoho_crosskv_weights_pack: blob too small");
  char* blob = static_cast<char*>(blob_);
  CkvOff o;
  const size_t per_layer = ckv_layer_bytes(d, &o);
  const int I = d.d_in, O = d.d_out;
  for (int l = 0; l < d.n_layers; ++l) {
    char* b = blob + per_layer * l;
    YOHO_CUDA_CHECK(cudaMemcpyAsync(b + o.wk, tensors[3 * l], size_t(I) * O * sizeof(float), cudaMemcpyDefault, s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(b + o.wv, tensors[3 * l + 1], size_t(I) * O * sizeof(float), cudaMemcpyDefault, s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(b + o.bv, tensors[3 * l + 2], size_t(O) * sizeof(float), cudaMemcpyDefault, s));
    float* wkv = reinterpret_cast<float*>(b + o.wkv);
    const int blocks = ceil_div(I * O, 256);
    ckv_copy_cols_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<const float*>(b + o.wk), wkv, I, O, 2 * O, 0);
    YOHO_LAUNCH_CHECK("ckv_copy_cols_kernel");
    ckv_copy_cols_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<const float*>(b + o.wv), wkv, I, O, 2 * O, O);
    YOHO_LAUNCH_CHECK("ckv_copy_cols_kernel");
    YOHO_TRY(launch_fill_zero(b + o.bias, size_t(2) * O * sizeof(float), s));
    YOHO_CUDA_CHECK(cudaMemcpyAsync(b + o.bias + size_t(O) * sizeof(float), b + o.bv, size_t(O) * sizeof(float), cudaMemcpyDeviceToDevice, s));
    YOHO_TRY(launch_transpose_cast(wkv, reinterpret_cast<__nv_bfloat16*>(b + o.wkv_h), I, 2 * O, I, s));
  }
  return YOHO_OK;
}

extern "C" size_t yoho_crosskv_workspace_bytes(const yoho_crosskv_desc* desc, int batch, int t, int compute) {
  if (ckv_check(desc, "yoho_crosskv_workspace_bytes") != YOHO_OK) return 0;
  if (batch <= 0 || t <= 0 || compute != YOHO_COMPUTE_BF16) return 1024;
  return align_up(size_t(batch) * t * desc->d_in * sizeof(__nv_bfloat16), 1024);
}

extern "C" int yoho_crosskv_fwd(yoho_ctx* ctx, const yoho_crosskv_desc* desc, const void* blob_, const float* states, int batch, int t,
                                float* kv, int compute, void* ws, size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  if (!ctx || !blob_ || !states || !kv) return fail(YOHO_ERR_INVALID, "yoho_crosskv_fwd: null argument");
  YOHO_TRY(ckv_check(desc, "yoho_crosskv_fwd"));
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "yoho_crosskv_fwd: device is sm_%d%d, kernels are sm_100a only", ctx->cc_major, ctx->cc_minor);
  if (compute != YOHO_COMPUTE_FP32 && compute != YOHO_COMPUTE_BF16) return fail(YOHO_ERR_INVALID, "yoho_crosskv_fwd: compute=%d", compute);
  if (batch < 0 || t <= 0) return fail(YOHO_ERR_INVALID, "yoho_crosskv_fwd: batch=%d t=%d", batch, t);
  if (batch == 0 || desc->n_layers == 0) return YOHO_OK;
  const long long rows_ll = (long long)batch * t;
  if (rows_ll > 0x7fffffffLL / 8) return fail(YOHO_ERR_UNSUPPORTED, "yoho_crosskv_fwd: batch too large for 32-bit row indices");
  const int rows = int(rows_ll), I = desc->d_in, O = desc->d_out;
  if (reinterpret_cast<uintptr_t>(blob_) % 256 != 0) return fail(YOHO_ERR_INVALID, "yoho_crosskv_fwd: blob must be 256-byte aligned");
  const char* blob = static_cast<const char*>(blob_);
  CkvOff o;
  const size_t per_layer = ckv_layer_bytes(*desc, &o);
  const size_t out_per_layer = size_t(rows) * 2 * O;
  if (compute == YOHO_COMPUTE_BF16) {
    const size_t need = yoho_crosskv_workspace_bytes(desc, batch, t, compute);
    if (!ws || ws_bytes < need) return fail(YOHO_ERR_WORKSPACE, "yoho_crosskv_fwd: workspace %zu < %zu bytes", ws_bytes, need);
    if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0) return fail(YOHO_ERR_INVALID, "yoho_crosskv_fwd: workspace must be 1024-byte aligned");
    __nv_bfloat16* xh = static_cast<__nv_bfloat16*>(ws);
    YOHO_TRY(launch_cast_bf16(states, xh, size_t(rows) * I, s));
    for (int l = 0; l < desc->n_layers; ++l) {
      const char* b = blob + per_layer * l;
      GemmBf16Args g;
      g.a = xh; g.a_ld = I; g.a_cols = I; g.batch = 1; g.rows_per_batch = rows;
      g.w = reinterpret_cast<const __nv_bfloat16*>(b + o.wkv_h); g.n = 2 * O; g.k_per_tap = I;
      g.out_f32 = kv + out_per_layer * l; g.ldc = 2 * O; g.bias = reinterpret_cast<const float*>(b + o.bias);
      YOHO_TRY(launch_gemm_bf16(ctx, g, s));
    }
    return YOHO_OK;
  }
  if (rows > 65535 * 128) return fail(YOHO_ERR_UNSUPPORTED, "yoho_crosskv_fwd (fp32): %d rows exceed one launch", rows);
  for (int l = 0; l < desc->n_layers; ++l) {
    const char* b = blob + per_layer * l;
    SgemmArgs g;
    g.a = states; g.b = reinterpret_cast<const float*>(b + o.wkv); g.c = kv + out_per_layer * l;
    g.m = rows; g.n = 2 * O; g.k = I; g.lda = I; g.ldb = 2 * O; g.ldc = 2 * O; g.bias = reinterpret_cast<const float*>(b + o.bias);
    YOHO_TRY(launch_sgemm(g, s));
  }
  return YOHO_OK;
}
