// jax.ffi (XLA FFI) custom-call handlers over the C ABI of include/yoho_b200.h.
//
// The reference calls the hot path inside jax.jit / jax.pmap traces (audio.py:32-58,
// whisper.py:194-195, transcription_pretrain_no_voiceprints.py:172-203), so the drop-in for the
// JAX pipeline is a set of XLA custom calls.  XLA owns every buffer (inputs, outputs and the scratch
// workspace, which is declared as an extra U8 result), supplies the stream, and calls the handlers
// from one host thread per device; errors come back as ffi::Error, never as exceptions.
//
// Opaque library state (yoho_ctx*, yoho_filterbank*) is created once per device from Python at
// registration time and passed as int64 attributes.  So is the packed weight blob: it is packed once, outside
// any trace, by the ctypes call yoho_encoder_weights_pack into a 256-byte-aligned device allocation the Python
// side keeps alive, and its address and size travel as the attributes "blob" / "blob_bytes"
// -- XLA never relocates it, so the alignment it was packed at is the alignment it is read at.
//
// STATUS: written against the XLA FFI API, never compiled or run (no jaxlib on either machine).
//
// This translation unit needs the XLA FFI headers that ship with jaxlib (jax.ffi.include_dir()).
// They exist neither in the build container nor on the GPU box (no jax / jaxlib, no network), so
// here the file compiles to an empty object; INTEGRATION.md has the build line and the Python side.
#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define YOHO_HAVE_XLA_FFI 1
#endif
#endif

#ifdef YOHO_HAVE_XLA_FFI
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

#include "../../include/yoho_b200.h"
#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

namespace {

// XLA aligns buffers to 64 B; the library wants its scratch at 1024 B: round up inside the (over-allocated by 1024 B
// on the Python side) buffer.
struct Scratch { void* ptr; size_t bytes; };
Scratch align_scratch(ffi::ResultBuffer<ffi::U8>& ws) {
  const uintptr_t raw = reinterpret_cast<uintptr_t>(ws->typed_data()), al = (raw + 1023) / 1024 * 1024;
  const size_t lost = size_t(al - raw), have = ws->size_bytes();
  return {reinterpret_cast<void*>(al), have > lost ? have - lost : 0};
}

ffi::Error status_to_error(int code, const char* what) {
  if (code >= 0) return ffi::Error::Success();
  return ffi::Error(code == YOHO_ERR_INVALID ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal,
                    std::string(what) + ": " + yoho_last_error());
}

// mel_spectogram / normalize_spectogram fused: pcm (B, N) f32 -> out (B, F, M) f32; ws: U8 scratch.
ffi::Error LogMelImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> pcm, ffi::ResultBuffer<ffi::F32> out,
                      ffi::ResultBuffer<ffi::U8> ws, int64_t ctx, int64_t filterbank, int32_t n_fft, int32_t hop,
                      int32_t norm) {
  auto dims = pcm.dimensions();
  if (dims.size() != 2) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "yoho_logmel: audio must be (batch, samples)");
  const Scratch sc = align_scratch(ws);
  int rc = yoho_logmel_f32(reinterpret_cast<yoho_ctx*>(ctx), reinterpret_cast<const yoho_filterbank*>(filterbank),
                           pcm.typed_data(), int(dims[0]), int(dims[1]), n_fft, hop, out->typed_data(), norm,
                           sc.ptr, sc.bytes, stream);
  return status_to_error(rc, "yoho_logmel_f32");
}

// normalize_spectogram alone: spec (B, F, M) f32 -> out (B, F, M) f32.
ffi::Error NormalizeImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> spec, ffi::ResultBuffer<ffi::F32> out,
                         ffi::ResultBuffer<ffi::U8> ws, int64_t ctx, int32_t norm) {
  auto dims = spec.dimensions();
  if (dims.size() < 2) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "yoho_normalize: need (..., frames, mels)");
  int64_t batch = 1;
  for (size_t i = 0; i + 2 < dims.size(); ++i) batch *= dims[i];
  const Scratch sc = align_scratch(ws);
  int rc = yoho_normalize_f32(reinterpret_cast<yoho_ctx*>(ctx), spec.typed_data(), int(batch), int(dims[dims.size() - 2]),
                              int(dims[dims.size() - 1]), out->typed_data(), norm, sc.ptr, sc.bytes, stream);
  return status_to_error(rc, "yoho_normalize_f32");
}

// Whisper.encode_audio: mel (B, F, M) f32 -> out (B, T, D) f32; the packed weight blob is a device allocation owned
// by the Python side (see the header comment), passed by address.
ffi::Error EncoderImpl(cudaStream_t stream, ffi::Buffer<ffi::F32> mel, ffi::ResultBuffer<ffi::F32> out,
                       ffi::ResultBuffer<ffi::U8> ws, int64_t ctx, int64_t blob, int64_t blob_bytes, int32_t dims,
                       int32_t n_heads, int32_t n_layers, int32_t seq_len, int32_t compute) {
  auto d = mel.dimensions();
  if (d.size() != 3) return ffi::Error(ffi::ErrorCode::kInvalidArgument, "yoho_encoder: mel must be (batch, frames, mels)");
  if (blob == 0 || (uintptr_t(blob) & 255u) != 0)
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "yoho_encoder: the packed weight blob must be a 256-byte-aligned device pointer");
  yoho_enc_desc desc{int32_t(d[2]), dims, n_heads, n_layers, seq_len};
  const Scratch sc = align_scratch(ws);
  int rc = yoho_encoder_fwd(reinterpret_cast<yoho_ctx*>(ctx), &desc, reinterpret_cast<const void*>(blob), size_t(blob_bytes),
                            mel.typed_data(), int(d[0]), int(d[1]), out->typed_data(), YOHO_DTYPE_F32, compute, sc.ptr, sc.bytes,
                            stream);
  return status_to_error(rc, "yoho_encoder_fwd");
}

}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(yoho_xla_logmel, LogMelImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("ctx")
                                  .Attr<int64_t>("filterbank")
                                  .Attr<int32_t>("n_fft")
                                  .Attr<int32_t>("hop")
                                  .Attr<int32_t>("norm"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(yoho_xla_normalize, NormalizeImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("ctx")
                                  .Attr<int32_t>("norm"));

XLA_FFI_DEFINE_HANDLER_SYMBOL(yoho_xla_encoder, EncoderImpl,
                              ffi::Ffi::Bind()
                                  .Ctx<ffi::PlatformStream<cudaStream_t>>()
                                  .Arg<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::F32>>()
                                  .Ret<ffi::Buffer<ffi::U8>>()
                                  .Attr<int64_t>("ctx")
                                  .Attr<int64_t>("blob")
                                  .Attr<int64_t>("blob_bytes")
                                  .Attr<int32_t>("dims")
                                  .Attr<int32_t>("n_heads")
                                  .Attr<int32_t>("n_layers")
                                  .Attr<int32_t>("seq_len")
                                  .Attr<int32_t>("compute"));
extern "C" int yoho_xla_ffi_available(void) { return 1; }
#else
// XLA FFI headers not available on this image: nothing else to compile (see the header comment).
extern "C" int yoho_xla_ffi_available(void) { return 0; }
#endif
