// Shared host/device helpers for libyoho_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/yoho_b200.h"

namespace yoho {

// ---- thread-local error message ---------------------------------------------------
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define YOHO_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return ::yoho::fail(YOHO_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                   \
                          cudaGetErrorString(_e), __FILE__, __LINE__);                     \
  } while (0)

#define YOHO_LAUNCH_CHECK(name)                                                            \
  do {                                                                                     \
    ::yoho::count_launch();                                                                \
    cudaError_t _e = cudaGetLastError();                                                   \
    if (_e != cudaSuccess)                                                                 \
      return ::yoho::fail(YOHO_ERR_CUDA, "launch of %s failed: %s", name,                  \
                          cudaGetErrorString(_e));                                         \
  } while (0)

// ---- launch accounting and the optional per-kernel-class event profiler (bench.py roofline) ----
enum ProfClass { kProfFrontend = 0, kProfNormalize, kProfGemm, kProfAttention, kProfLayerNorm, kProfOther, kProfDecGemm, kProfDecAttention, kProfNumClasses };
void count_launch();
// RAII: when profiling is enabled, brackets the launches issued in its lifetime with CUDA events on `s`.
struct ProfScope {
  ProfScope(cudaStream_t s, int cls, double work);
  ~ProfScope();
  cudaStream_t stream; int slot;
};

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// Bump allocator over a caller-provided workspace; every block is 1024-byte aligned.
struct Carver {
  char* base;
  size_t off = 0;
  explicit Carver(void* p) : base(static_cast<char*>(p)) {}
  template <class T> T* take(size_t n) {
    off = align_up(off, 1024);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
  size_t used() const { return align_up(off, 1024); }
};

}  // namespace yoho

// ---- per-device state -------------------------------------------------------------------
struct yoho_ctx {
  int device = -1;
  int num_sms = 0;
  int cc_major = 0, cc_minor = 0;
  // frontend constants for n_fft = 400 (built in double on the host, stored as float32)
  float* d_window = nullptr;    // periodic Hann (400)
  float2* d_twiddle = nullptr;  // W_400^(n2*k1) laid out [k1][n2] (20 x 20)
  double window_sum = 0.0;      // sum of the float32-rounded window
  void* encode_tiled = nullptr; // cuTensorMapEncodeTiled driver entry point
};

// Mel filterbank as the log-mel kernel consumes it: bands sorted longest first and dealt to the kernel's
// kMelWarps warps round by round (boustrophedon); every band of a round is walked 4 * round_n4 bins.
constexpr int kMelWarps = 10;          // warps of logmel_kernel
constexpr int kMelPRows = 204;         // rows of the kernel's power tile: 201 bins rounded up to the 4-bin step
constexpr int kMelMaxRounds = 16;      // n_mels <= 160
constexpr int kMelMaxWeights = 3072;   // floats of band weights held in shared memory
struct MelSchedule { int rounds; unsigned char n4[kMelMaxRounds]; };
struct yoho_filterbank {
  int device = -1;
  int n_mels = 0, n_bins = 0;
  int band_stride = 0;        // floats per row in d_weights (multiple of 4)
  MelSchedule sched = {};
  int2* d_bands = nullptr;    // sorted position -> {mel row, first bin walked}
  float* d_weights = nullptr; // (n_mels, band_stride) in sorted order, zero padded, already multiplied by the STFT power scale
  double power_scale = 0.0;   // 0.25 / window_sum^2 folded into d_weights
};
