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

// ---- programmatic dependent launch -----------------------------------------------------------------------------------
// Kernels of the encoder chain are launched with the programmatic-stream-serialization attribute: kernel B may be
// scheduled while its predecessor A is still running (on idle SMs, or SMs A's CTAs have left), runs its prologue (barrier
// init, TMEM allocation, tensor-map prefetch) and then blocks in pdl_wait() until A has completed and its writes are
// visible.  Every participating kernel calls pdl_launch_dependents() first thing and pdl_wait() before its first global
// memory access (reads AND writes: a buffer it overwrites may still be read by A).  YOHO_PDL=0 launches them plainly.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

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
  // frontend constants for n_fft = 400 (built in double on the host, stored as float32 in the kernel's constant bank:
  // frontend_upload_constants)
  double window_sum = 0.0;      // sum of the float32-rounded periodic Hann window
  void* encode_tiled = nullptr; // cuTensorMapEncodeTiled driver entry point
};

namespace yoho {
// Uploads the periodic Hann window and the W_400 twiddles into logmel_kernel's __constant__ tables on the current
// device (called by yoho_ctx_create); returns the sum of the float32-rounded window through *window_sum.
int frontend_upload_constants(double* window_sum);
}

// Mel filterbank as the log-mel kernel consumes it.  Every mel row is a band [first, first + trip) of FFT bins; bands are
// sorted longest first and taken four at a time (two per half warp, lane = frame pair), so the bands of a group are all
// walked `trip` bins (the longest one's length rounded up to even) with zero weights past their own end.
constexpr int kMelMaxBands = 128;      // n_mels <= 128 (band pairs * 2)
constexpr int kMelMaxWeights = 600;    // (w, w) entries held in shared memory: sum over bands of their pair's trip
constexpr int kMelBins = 201;          // n_fft / 2 + 1
constexpr int kMelGroup = 8;           // bands that share a trip count: one warp-round of the mel phase (4 lane slots x 2)
constexpr int kMelWarps = 6;           // warps of the frontend CTA: the host deals the band groups to them
constexpr int kMelMaxRounds = 5;       // rounds x kMelWarps schedule entries (32 bytes of shared memory)
struct yoho_filterbank {
  int device = -1;
  int n_mels = 0, n_bins = 0;
  int n_groups = 0;           // groups of kMelGroup bands, sorted by trip (descending)
  int n_rounds = 0;           // rounds of the mel phase: schedule[round][warp] = group or -1 (longest-processing-time first)
  int n_weights = 0;          // float2 entries in d_weights
  int4* d_bands = nullptr;    // [kMelGroup * n_groups] {mel row or -1, first bin walked, offset into d_weights, trip}, then 8 int4 = the schedule
  float2* d_weights = nullptr; // (w, w) per walked bin, already multiplied by the STFT power scale
  double power_scale = 0.0;   // 0.25 / window_sum^2 folded into d_weights
};
