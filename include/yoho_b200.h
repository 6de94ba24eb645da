/*
 * yoho_b200.h -- C ABI of the B200-native YOHO hot path (log-mel frontend +
 * Whisper audio-encoder forward).
 *
 * The reference (JirkaKlimes/YOHO) has no plugin / FFI interface: the boundary is four
 * Python call signatures traced under jax.jit / jax.pmap.  Each entry point below names
 * the reference call it stands in for (paths relative to the reference checkout); the
 * Python mirror in yoho_b200/ keeps the reference names and binds these symbols with
 * ctypes, and csrc/xla_ffi.cc wraps the same symbols as jax.ffi custom calls.
 *
 * Conventions
 *   - plain C, no C++ types, no exceptions cross the boundary;
 *   - every array pointer is a DEVICE pointer owned by the caller unless the name says
 *     `host_`; all arrays are dense, row-major, C-contiguous;
 *   - every compute call is asynchronous on the caller's stream (`yoho_stream_t` is a
 *     `cudaStream_t`), performs no allocation and no synchronisation; scratch memory is
 *     passed in (`ws`, size from the matching `*_workspace_bytes`);
 *   - return value: YOHO_OK (0) or a negative yoho_status; `yoho_last_error()` returns a
 *     thread-local message for the last failure on the calling thread;
 *   - there is no CPU fallback: on a device that is not sm_100 every compute call fails
 *     with YOHO_ERR_ARCH.
 */
#ifndef YOHO_B200_H_
#define YOHO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YOHO_ABI_VERSION 2

typedef struct yoho_ctx yoho_ctx;               /* per-device immutable state           */
typedef struct yoho_filterbank yoho_filterbank; /* banded mel filterbank on the device  */
typedef void* yoho_stream_t;                    /* cudaStream_t                         */

typedef enum yoho_status {
  YOHO_OK = 0,
  YOHO_ERR_INVALID = -1,     /* bad argument (null pointer, negative size, shape rule)  */
  YOHO_ERR_UNSUPPORTED = -2, /* valid in the reference, not implemented by the kernels  */
  YOHO_ERR_WORKSPACE = -3,   /* ws_bytes smaller than *_workspace_bytes()               */
  YOHO_ERR_CUDA = -4,        /* a CUDA runtime / driver call failed                     */
  YOHO_ERR_ARCH = -5         /* device is not compute capability 10.x                   */
} yoho_status;

typedef enum yoho_norm {
  YOHO_NORM_NONE = 0,      /* mel power, what mel_spectogram returns (audio.py:32-37)        */
  YOHO_NORM_REFERENCE = 1, /* normalize_spectogram: log10(max(x,1e-13)), (x-mean)/std (audio.py:53-58) */
  YOHO_NORM_WHISPER = 2    /* log10(max(x,1e-10)), max(x, clip_max-8), (x+4)/4  (north_star)  */
} yoho_norm;

typedef enum yoho_compute {
  YOHO_COMPUTE_FP32 = 0, /* fp32 operands and accumulation (SIMT), 1e-4 parity path      */
  YOHO_COMPUTE_BF16 = 1  /* bf16 operands, fp32 accumulation on tcgen05 tensor cores     */
} yoho_compute;

typedef enum yoho_dtype {
  YOHO_DTYPE_F32 = 0,  /* what the reference returns (float32 everywhere, x64 disabled)           */
  YOHO_DTYPE_BF16 = 1  /* encoder states rounded once by the final LayerNorm: half the D2H bytes */
} yoho_dtype;

/* ---- library / context ------------------------------------------------------------- */

int yoho_abi_version(void);
const char* yoho_last_error(void);
/* 1 when the library was built with the XLA FFI headers and exports the jax.ffi handler symbols
 * yoho_xla_logmel / yoho_xla_normalize / yoho_xla_encoder - see csrc/xla_ffi.cc and INTEGRATION.md; else 0. */
int yoho_xla_ffi_available(void);

/* Creates the per-device state (FFT twiddles, window, driver entry points).  May allocate
 * and synchronise.  One context per device; safe to share between host threads. */
int yoho_ctx_create(int device, yoho_ctx** out);
int yoho_ctx_destroy(yoho_ctx* ctx);

/* ---- frontend ---------------------------------------------------------------------- */

/* Frames produced by jax.scipy.signal.stft(x, nperseg=n_fft, noverlap=n_fft-hop,
 * boundary=None) with its default padded=True (audio.py:33): 480000 -> 2999.
 * Returns a negative yoho_status for invalid arguments. */
int yoho_stft_frames(int n_samples, int n_fft, int hop);

/* Replaces mel_filter_banks(sr, n_fft, n_mels) as a *device constant*
 * (mel_filterbanks.py:21-42, consumed at audio.py:35-36).  `host_filters` is the dense
 * row-major (n_mels, n_bins) float32 matrix on the HOST; it is packed into per-row
 * bands [lo, hi) covering every non-zero, sorted longest first into the round schedule
 * the log-mel kernel's ten warps walk, and uploaded.  Limits (YOHO_ERR_UNSUPPORTED
 * beyond them): n_mels <= 160, n_bins <= 204, n_mels x (longest band rounded up to 4)
 * <= 3072 weights.  Setup call: may allocate and synchronise. */
int yoho_filterbank_create(yoho_ctx* ctx, const float* host_filters, int n_mels, int n_bins,
                           yoho_filterbank** out);
int yoho_filterbank_destroy(yoho_filterbank* fb);

size_t yoho_logmel_workspace_bytes(int batch, int n_samples, int n_fft, int hop, int n_mels);

/* Replaces mel_spectogram(audio, n_fft, hop_len, sample_rate, n_mels) (audio.py:32-37;
 * norm = YOHO_NORM_NONE) and, fused, normalize_spectogram (audio.py:53-58;
 * YOHO_NORM_REFERENCE) or the Whisper normaliser (YOHO_NORM_WHISPER).
 *   pcm  (batch, n_samples) float32      out  (batch, frames, n_mels) float32
 * Kernels exist for n_fft = 400, hop = 160 (every configuration in the reference and in
 * BASELINE.json); other values return YOHO_ERR_UNSUPPORTED. */
int yoho_logmel_f32(yoho_ctx* ctx, const yoho_filterbank* fb, const float* pcm, int batch,
                    int n_samples, int n_fft, int hop, float* out, int norm, void* ws,
                    size_t ws_bytes, yoho_stream_t stream);

/* Same with interleaved-free mono int16 PCM as ffmpeg delivers it (audio.py:11-18); samples
 * are scaled by 1/32768 on load (whisper.py:249). */
int yoho_logmel_s16(yoho_ctx* ctx, const yoho_filterbank* fb, const int16_t* pcm, int batch,
                    int n_samples, int n_fft, int hop, float* out, int norm, void* ws,
                    size_t ws_bytes, yoho_stream_t stream);

/* No counterpart in the reference.  Same computation as yoho_logmel_f32 for
 * callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the per-tile
 * statistics scratch lives in a file-scope device array of the library (256 KB, held by every GPU that loads the
 * library until the process exits) instead of a caller-provided workspace, and the call performs no cudaMalloc / cudaFree
 * or any other allocation.  Limit: batch * ceil(frames / 32) <= yoho_logmel_static_max_tiles() = 32768 statistics slots
 * (e.g. 348 clips of 30 s); larger calls return YOHO_ERR_WORKSPACE.  All calls on a device share the array, which keeps
 * the last call's contents (nothing in it needs to start at zero): calls that overlap on the GPU (different streams) race;
 * serialise them on one stream or use yoho_logmel_f32. */
size_t yoho_logmel_static_max_tiles(void);
int yoho_logmel_f32_static(yoho_ctx* ctx, const yoho_filterbank* fb, const float* pcm, int batch,
                           int n_samples, int n_fft, int hop, float* out, int norm, yoho_stream_t stream);

/* Long-recording ingest (the step before the path: whisper.py:249-253, dataloaders.py:125-128): `recording`
 * holds total_samples mono samples on the device; clip b is recording[b * clip_samples, (b + 1) * clip_samples),
 * the last clip zero padded (np.pad, whisper.py:253).  Writes yoho_chunk_count(total_samples, clip_samples) =
 * ceil(total / clip) rows of (frames, n_mels); size the workspace with yoho_logmel_workspace_bytes(that count, ...).
 * No intermediate (clips, clip_samples) float32 copy is made: conversion, padding and framing happen in the load. */
int yoho_chunk_count(long long total_samples, int clip_samples);
int yoho_logmel_chunked_f32(yoho_ctx* ctx, const yoho_filterbank* fb, const float* recording,
                            long long total_samples, int clip_samples, int n_fft, int hop, float* out,
                            int norm, void* ws, size_t ws_bytes, yoho_stream_t stream);
int yoho_logmel_chunked_s16(yoho_ctx* ctx, const yoho_filterbank* fb, const int16_t* recording,
                            long long total_samples, int clip_samples, int n_fft, int hop, float* out,
                            int norm, void* ws, size_t ws_bytes, yoho_stream_t stream);

size_t yoho_normalize_workspace_bytes(int batch, int frames, int n_mels);

/* Stand-alone normalize_spectogram(spectogram) (audio.py:53-58) / Whisper normaliser on an
 * existing (batch, frames, n_mels) float32 mel power spectrogram.  `out` may alias `spec`.
 * norm must be YOHO_NORM_REFERENCE or YOHO_NORM_WHISPER. */
int yoho_normalize_f32(yoho_ctx* ctx, const float* spec, int batch, int frames, int n_mels,
                       float* out, int norm, void* ws, size_t ws_bytes, yoho_stream_t stream);

/* ---- encoder ----------------------------------------------------------------------- */

/* Shape of AudioEncoder(seq_len, dims, n_heads, n_layers) (whisper.py:117-122) plus the mel
 * width of its input.  dims / n_heads must be 64 (true for every Whisper size). */
typedef struct yoho_enc_desc {
  int32_t n_mels;   /* input channels of Conv_0                    */
  int32_t dims;     /* D                                            */
  int32_t n_heads;  /* H                                            */
  int32_t n_layers; /* L                                            */
  int32_t seq_len;  /* rows of the sinusoid table (1500)            */
} yoho_enc_desc;

/* Number and order of the float32 tensors yoho_encoder_weights_pack consumes:
 *   0 Conv_0.kernel (3, n_mels, D)   1 Conv_0.bias (D)
 *   2 Conv_1.kernel (3, D, D)        3 Conv_1.bias (D)
 *   4 sinusoid table (seq_len, D)    [SinPositionalEncoding has no params; built by the host
 *                                     in the reference's float32 op order, whisper.py:11-17]
 *   5 LayerNorm_0.scale (D)          6 LayerNorm_0.bias (D)            [final LN, whisper.py:134]
 *   then for each EncoderBlock_i, 15 tensors (whisper.py:75-88, 31-35):
 *     LayerNorm_0.scale, LayerNorm_0.bias,
 *     query_proj.kernel (D,D), query_proj.bias, key_proj.kernel (D,D) [no bias],
 *     value_proj.kernel (D,D), value_proj.bias, out_proj.kernel (D,D), out_proj.bias,
 *     LayerNorm_1.scale, LayerNorm_1.bias,
 *     Dense_0.kernel (D,4D), Dense_0.bias (4D), Dense_1.kernel (4D,D), Dense_1.bias (D)
 * Kernels keep the Flax layout (in, out) / (k, in, out). */
int yoho_encoder_num_tensors(const yoho_enc_desc* desc);
/* Elements of tensor `idx` in that order (negative yoho_status for a bad desc / index): lets a binding check the
 * leaves of a parameter pytree against the descriptor before yoho_encoder_weights_pack reads them. */
long long yoho_encoder_tensor_elems(const yoho_enc_desc* desc, int idx);

size_t yoho_encoder_weights_bytes(const yoho_enc_desc* desc);

/* Replaces "{'params': params}" of model.apply (whisper.py:258): copies the float32 tensors
 * (host or device pointers) into `blob` and derives the packed bf16 operand copies
 * ((out, in) K-major, q/k/v fused) on the device.  Asynchronous on `stream`; host source
 * buffers must stay valid until the stream reaches this point. */
int yoho_encoder_weights_pack(yoho_ctx* ctx, const yoho_enc_desc* desc, const float* const* tensors,
                              int n_tensors, void* blob, size_t blob_bytes, yoho_stream_t stream);

/* T = floor((frames + 2 - 3) / 2) + 1 (Conv_1: k=3, padding=1, strides=2; whisper.py:127). */
int yoho_encoder_out_len(int frames);

size_t yoho_encoder_workspace_bytes(const yoho_enc_desc* desc, int batch, int frames, int compute);

/* Replaces model.apply({"params": params}, mel, method=Whisper.encode_audio)
 * (whisper.py:194-195, 258) == AudioEncoder.__call__ (whisper.py:124-135).
 *   mel (batch, frames, n_mels) float32  ->  out (batch, T, D) float32 or bf16 (out_dtype: yoho_dtype)
 * yoho_encoder_out_len(frames) must equal desc->seq_len (whisper.py:23 broadcasts).  blob_bytes is the size of
 * the packed blob and must be >= yoho_encoder_weights_bytes(desc): a blob packed for another descriptor (another
 * seq_len moves every offset behind the sinusoid table) is refused instead of being read at wrong offsets.
 * YOHO_DTYPE_BF16 needs YOHO_COMPUTE_BF16.  The launch sequence only depends on (desc, batch, frames, pointers),
 * so the call can be captured in a CUDA graph (the Python mirror does so for single-clip latency). */
int yoho_encoder_fwd(yoho_ctx* ctx, const yoho_enc_desc* desc, const void* blob, size_t blob_bytes,
                     const float* mel, int batch, int frames, void* out, int out_dtype, int compute,
                     void* ws, size_t ws_bytes, yoho_stream_t stream);

/* Optional: all-gather of every rank's encoder states for a single consumer (the reference shards a batch over the
 * devices with common_utils.shard, train/stages/transcription_pretrain_no_voiceprints.py:231-233, and reads results back
 * per device).  The hot path itself has NO collective.  `nccl_comm` is the caller's ncclComm_t; NCCL is not linked, the
 * call resolves ncclAllGather from the libnccl already loaded in the process (YOHO_ERR_UNSUPPORTED if there is none).
 *   states (elems_per_rank) -> gathered (n_ranks * elems_per_rank), rank order; dtype: yoho_dtype. */
int yoho_allgather_states(yoho_ctx* ctx, void* nccl_comm, const void* states, void* gathered,
                          size_t elems_per_rank, int dtype, yoho_stream_t stream);

/* ---- cross-attention K/V pre-projection (the step after the path) ------------------------- */

/* Every DecoderBlock applies key_proj / value_proj of its cross MultiHeadAttention to the same encoder states
 * (whisper.py:104 -> :45-46) and the reference redoes that for every generated token (:274-281; "TODO: add KV
 * cache", :27).  They depend on the clip only: one data-parallel GEMM per decoder layer, done once here. */
typedef struct yoho_crosskv_desc {
  int32_t d_in;     /* encoder width (audio_dims)       */
  int32_t d_out;    /* decoder width (text_dims)        */
  int32_t n_layers; /* decoder layers                   */
} yoho_crosskv_desc;

/* 3 float32 tensors per decoder layer, in order: DecoderBlock_l/MultiHeadAttention_1/key_proj.kernel (d_in, d_out)
 * [no bias, whisper.py:33], value_proj.kernel (d_in, d_out), value_proj.bias (d_out). */
int yoho_crosskv_num_tensors(const yoho_crosskv_desc* desc);
long long yoho_crosskv_tensor_elems(const yoho_crosskv_desc* desc, int idx);
size_t yoho_crosskv_weights_bytes(const yoho_crosskv_desc* desc);
int yoho_crosskv_weights_pack(yoho_ctx* ctx, const yoho_crosskv_desc* desc, const float* const* tensors,
                              int n_tensors, void* blob, size_t blob_bytes, yoho_stream_t stream);
size_t yoho_crosskv_workspace_bytes(const yoho_crosskv_desc* desc, int batch, int t, int compute);
/* states (batch, t, d_in) float32 -> kv (n_layers, batch, t, 2 * d_out) float32 with [k | v] along the last axis:
 * k = states @ key_proj.kernel, v = states @ value_proj.kernel + value_proj.bias.  The d_head^-0.25 scaling of k
 * (whisper.py:54-57) is part of qkv_attention and left to the consumer. */
int yoho_crosskv_fwd(yoho_ctx* ctx, const yoho_crosskv_desc* desc, const void* blob, const float* states,
                     int batch, int t, float* kv, int compute, void* ws, size_t ws_bytes,
                     yoho_stream_t stream);

/* ---- Whisper text decoder with a KV cache -------------------------------- */

/* TextDecoder(vocab_size, seq_len, dims, n_heads, n_layers) (whisper.py:138-165) plus the shape of the encoder
 * states it attends to.  dims / n_heads must be 64.  float32 SIMT kernels throughout (1e-4 parity path): one
 * generated token is `batch` rows against every weight, i.e. weight streaming, not tensor-core work. */
typedef struct yoho_decoder_desc {
  int32_t vocab;      /* rows of Embed_0.embedding, tied with the logits (whisper.py:153, :163) */
  int32_t seq_len;    /* rows of positional_embedding = token buffer length (448)             */
  int32_t dims;       /* D (text_dims)                                                          */
  int32_t n_heads;
  int32_t n_layers;
  int32_t audio_dims; /* width of the encoder states                                            */
  int32_t audio_len;  /* encoder positions per clip (1500)                                      */
} yoho_decoder_desc;

/* 4 + 24 * n_layers float32 tensors, Flax layouts, in order:
 *   positional_embedding (seq_len, D), Embed_0.embedding (vocab, D), LayerNorm_0.scale, LayerNorm_0.bias (final, :162);
 *   per DecoderBlock_l: LayerNorm_0.{scale,bias};
 *     MultiHeadAttention_0 (masked self attention): query_proj.{kernel,bias}, key_proj.kernel, value_proj.{kernel,bias},
 *     out_proj.{kernel,bias};  LayerNorm_1.{scale,bias};
 *     MultiHeadAttention_1 (cross attention; key/value kernels are (audio_dims, D)): same seven;
 *     LayerNorm_2.{scale,bias};  Dense_0.{kernel (D,4D), bias};  Dense_1.{kernel (4D,D), bias}. */
int yoho_decoder_num_tensors(const yoho_decoder_desc* desc);
long long yoho_decoder_tensor_elems(const yoho_decoder_desc* desc, int idx);
size_t yoho_decoder_weights_bytes(const yoho_decoder_desc* desc);
int yoho_decoder_weights_pack(yoho_ctx* ctx, const yoho_decoder_desc* desc, const float* const* tensors,
                              int n_tensors, void* blob, size_t blob_bytes, yoho_stream_t stream);

/* Teacher-forced forward == Whisper.decode_text(encoded_audio, tokens) (whisper.py:197-198 -> :147-165):
 *   tokens (batch, s_len) int32, states (batch, audio_len, audio_dims) float32 -> logits (batch, s_len, vocab) float32.
 * s_len <= seq_len.  Token ids outside [0, vocab) are clamped. */
size_t yoho_decoder_workspace_bytes(const yoho_decoder_desc* desc, int batch, int s_len);
int yoho_decoder_fwd(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, const int32_t* tokens,
                     const float* states, int batch, int s_len, float* logits, void* ws, size_t ws_bytes,
                     yoho_stream_t stream);

/* Token-by-token generation, the loop of whisper.py:260-285 with the KV cache the reference lacks (:27).
 *   cache: yoho_decoder_cache_bytes(desc, batch) bytes, 1024-byte aligned: a device-side position counter, the self
 *          attention keys / values of every layer (n_layers, batch, seq_len, 2D) and the cross attention keys / values
 *          (n_layers, batch, audio_len, 2D).
 *   yoho_decoder_begin   computes the cross keys / values from `states` once and sets the position to 0.
 *   yoho_decoder_step    at position p (read on the device): consumes tokens[:, p], appends to the self cache, writes
 *                        the logits for position p + 1 to `logits` (batch, vocab) if not NULL, stores their first argmax
 *                        (jnp.argmax, :279) into tokens[:, p + 1] unless forced[p + 1] != 0 (forced: seq_len bytes on the
 *                        device or NULL; the reference pre-fills positions 0, 2, 3, :266-272), and advances p.  A row
 *                        whose token at p is `eot` keeps producing `eot` (pass eot < 0 to disable).
 *   yoho_decoder_generate  n_steps such steps back to back (no host synchronisation; CUDA-graph capturable).
 * tokens: (batch, seq_len) int32 on the device. */
size_t yoho_decoder_cache_bytes(const yoho_decoder_desc* desc, int batch);
size_t yoho_decoder_step_workspace_bytes(const yoho_decoder_desc* desc, int batch);
int yoho_decoder_begin(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, const float* states,
                       int batch, void* cache, size_t cache_bytes, yoho_stream_t stream);
int yoho_decoder_step(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, int32_t* tokens, int batch,
                      void* cache, size_t cache_bytes, float* logits, const unsigned char* forced, int eot,
                      void* ws, size_t ws_bytes, yoho_stream_t stream);
int yoho_decoder_generate(yoho_ctx* ctx, const yoho_decoder_desc* desc, const void* blob, int32_t* tokens,
                          int batch, void* cache, size_t cache_bytes, const unsigned char* forced, int eot,
                          int n_steps, void* ws, size_t ws_bytes, yoho_stream_t stream);

/* ---- YOHO's own audio encoder (next row after the Whisper path) ------------------------ */

/* Shape of AudioEncoder(q_heads, kv_heads, dims, seq_len, blocks) (yoho/src/nn/model.py:9-14) plus the mel
 * width of its input; Model.setup builds it with kv_heads = n_audio_heads // 2 (model.py:60-66).
 * dims / q_heads must be 64 and n_mels a multiple of 64 (the reference config: 384 / 6, 128 mels). */
typedef struct yoho_model_desc {
  int32_t n_mels;   /* input channels of Conv_0                              */
  int32_t dims;     /* D                                                      */
  int32_t q_heads;  /* query heads                                            */
  int32_t kv_heads; /* key / value heads (q_heads % kv_heads == 0)            */
  int32_t n_blocks; /* EncoderBlocks                                          */
  int32_t seq_len;  /* max_audio_len: the RoPE table has seq_len / 2 rows     */
} yoho_model_desc;

/* Number and order of the float32 tensors yoho_model_weights_pack consumes:
 *   0 Conv_0.kernel (3, n_mels, D)  1 Conv_0.bias (D)  2 RMSNorm_0.scale (D)
 *   3 Conv_1.kernel (3, D, D)       4 Conv_1.bias (D)  5 RMSNorm_1.scale (D)
 *   6 RoPE table (seq_len / 2, 2, 32): cos | sin of (s + 1) * 10^-i  [RoPE has no params; built by the
 *     host in the reference's float32 op order, layers.py:31-33]
 *   then for each EncoderBlock_i, 9 tensors (layers.py:101-109, 65-67, 88, 14-16):
 *     RMSNorm_0.scale, DenseGeneral_0.kernel (D, q_heads, 64), DenseGeneral_1.kernel (D, kv_heads, 64),
 *     DenseGeneral_2.kernel (D, kv_heads, 64), Dense_0.kernel (kv_heads*64, D), RMSNorm_1.scale,
 *     SwiGLU Dense_0.kernel (D, 3D), Dense_1.kernel (D, 3D), Dense_2.kernel (3D, D) */
int yoho_model_num_tensors(const yoho_model_desc* desc);
long long yoho_model_tensor_elems(const yoho_model_desc* desc, int idx);
size_t yoho_model_weights_bytes(const yoho_model_desc* desc);
int yoho_model_weights_pack(yoho_ctx* ctx, const yoho_model_desc* desc, const float* const* tensors,
                            int n_tensors, void* blob, size_t blob_bytes, yoho_stream_t stream);
/* frames / 2 (nn.max_pool(x, (1, 2), (1, 2)), model.py:20). */
int yoho_model_out_len(int frames);
size_t yoho_model_workspace_bytes(const yoho_model_desc* desc, int batch, int frames, int compute);

/* Replaces state.apply_fn({"params": params}, spectogram, method=Model.encode_audio)
 * (train/stages/transcription_pretrain_no_voiceprints.py:198-203; model.py:82-83, 16-29).
 *   mel (batch, frames, n_mels) float32  ->  out (batch, frames / 2, D) float32 */
int yoho_model_encoder_fwd(yoho_ctx* ctx, const yoho_model_desc* desc, const void* blob, const float* mel,
                           int batch, int frames, float* out, int compute, void* ws, size_t ws_bytes,
                           yoho_stream_t stream);

/* ---- measurement hooks (bench.py; not part of the reference surface) ------------------ */

/* Kernels launched by this library in this process so far (the `gpu_launches` evidence). */
long long yoho_launch_count(void);

/* Optional per-kernel-class CUDA-event profiler: while enabled every launch group is bracketed by
 * events on its own stream (no synchronisation); yoho_prof_collect synchronises on them and sums, per
 * class, the device time in ms, the algorithmic work (FLOP for the tcgen05 classes, bytes for the
 * memory-bound ones) and the number of groups.  yoho_prof_enable(0/1) clears previous records. */
int yoho_prof_enable(int on);
int yoho_prof_num_classes(void);
const char* yoho_prof_class_name(int cls);
int yoho_prof_collect(double* ms, double* work, long long* launches, int n);

/* ---- test hooks (exercised by tests/, not part of the reference surface) ------------- */

/* C (m, n) = A (m, k) @ W (k, n) + bias with the tcgen05 kernel: A, W float32 on the device,
 * rounded to bf16 inside; used to unit-test the tensor-core path in isolation. epilogue:
 * 0 bias, 1 bias+gelu(tanh). */
size_t yoho_test_gemm_workspace_bytes(int m, int n, int k);
int yoho_test_gemm_bf16(yoho_ctx* ctx, const float* a, const float* w, const float* bias, float* c,
                        int m, int n, int k, int epilogue, void* ws, size_t ws_bytes,
                        yoho_stream_t stream);

/* softmax(q k^T / 8) v per (batch, head) with the tcgen05 attention kernel; qkv (batch, t, 3, h, 64)
 * float32 (rounded to bf16 inside), out (batch, t, h*64) float32. */
size_t yoho_test_attention_workspace_bytes(int batch, int t, int heads);
int yoho_test_attention_bf16(yoho_ctx* ctx, const float* qkv, float* out, int batch, int t, int heads,
                             void* ws, size_t ws_bytes, yoho_stream_t stream);
/* which attention kernel serves the calls that follow: 0 = chosen by padding (default), 2 = two query tiles per work
 * item, 3 = three; returns the previous setting.  Process-wide, for tests and A/B timing only. */
int yoho_test_attention_tiles(int tiles);
/* out (b, n) (+)= act(x (b, k) @ w (k, n) + bias): the decoder's weight-streaming GEMM in isolation */
int yoho_test_skinny_gemm(yoho_ctx* ctx, const float* x, const float* w, const float* bias, float* out, int b,
                          int k, int n, int gelu, int accumulate, yoho_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* YOHO_B200_H_ */
