// Log-mel frontend for sm_100a: framing -> periodic Hann -> 400-point real FFT -> power ->
// banded mel projection -> log10 -> per-clip statistics, one kernel; then a small in-place
// finalise kernel applies the per-clip normaliser while the clip's log-mel is still L2 resident.
//
// Replaces yoho/src/preprocessing/audio.py:32-37 (mel_spectogram) and :53-58
// (normalize_spectogram) of the reference, plus the Whisper clamp/(x+4)/4 normaliser that
// BASELINE.json's north_star names.
//
// Data layout
//   pcm   (B, N)        float32 or int16, row-major
//   out   (B, F, M)     float32, time-major / mel-last (what audio.py:36's einsum returns)
//   stats (B, tiles, 2) float32 per-tile partials {sum, sum of squares} or {max, -}
//
// One tile = 32 consecutive frames of one clip = 16 complex FFTs of length 400, two real frames
// packed as (re, im).  400 = 20 x 20: each FFT is run by 20 threads, each doing one radix-20
// butterfly per pass entirely in registers (radix-20 = 4 x 5 by the prime-factor map, so no
// internal twiddles), with one shared-memory transpose between the passes.  The kernel is
// persistent (two CTAs per SM walk the (clip, tile) list): the 5360 PCM samples of the NEXT tile
// are fetched with 16-byte cp.async into the second half of a double buffer while the current tile
// is transformed (each sample is used by 2.5 frames; int16 is converted when it is read).  The two
// real spectra are untangled from Z[k], Z[400-k] through registers and the power spectrum is
// written back over Z bin-major, P[k][frame].  The mel bands (<= 2 filters touch any FFT bin) are
// then applied with lane = frame: a warp takes whole bands (dealt longest-first, boustrophedon,
// so the ten warps get equal work), the band's weights are warp-uniform 128-bit loads, there is
// no divergence, and the (32 frames x n_mels) result is transposed through shared memory so the
// global stores are full 128-byte lines.
#include "common.cuh"

namespace yoho {

constexpr int kNfft = 400;
constexpr int kHop = 160;
constexpr int kBins = 201;
constexpr int kFramesPerCta = 32;
constexpr int kFftsPerCta = kFramesPerCta / 2;
constexpr int kThreads = kFftsPerCta * 20;                      // 320
constexpr int kTileSamples = (kFramesPerCta - 1) * kHop + kNfft;  // 5360
constexpr int kERow = 21;                                       // padded row (float2) of the transpose buffer
constexpr int kEFft = 20 * kERow;                               // 420 float2 per FFT
constexpr int kZFft = 404;                                      // float2 per FFT of Z: 808 words = 8 mod 32 keeps the 20-thread groups off each other's banks
constexpr int kPRow = kFramesPerCta + 1;                        // floats per bin row of P[k][frame]: odd, so bin-strided writes spread over the banks
constexpr int kPRows = kMelPRows;                               // 201 bins, rounded up: bands are walked four bins at a time
constexpr int kMaxMels = kMelWarps * kMelMaxRounds;             // the output transpose buffer holds 32 x (n_mels | 1) floats
constexpr int kWarps = kThreads / 32;
static_assert(kWarps == kMelWarps, "the filterbank's round schedule is built for this many warps");

template <typename T>
struct FrontendSmem {
  T raw[2][kTileSamples];          // double-buffered PCM tile, as it lies in global memory
  float2 ez[kFftsPerCta * kEFft];  // pass-1 -> pass-2 transpose; then Z in natural order; then the power spectrum
  float mel_w[kMelMaxWeights];     // band weights, sorted order (resident for the whole kernel)
  int2 mel_band[kMaxMels];
  float red[2 * 32];
};
static_assert(kFftsPerCta * kZFft <= kFftsPerCta * kEFft, "Z must fit in the transpose buffer");
static_assert(kPRows * kPRow + kFramesPerCta * (kMaxMels | 1) <= 2 * kFftsPerCta * kEFft, "P and the output tile overlay the transpose buffer");
static_assert((kTileSamples * sizeof(int16_t)) % 16 == 0, "second PCM buffer stays 16-byte aligned");

// ---- radix-20 butterfly (prime-factor 4 x 5), forward sign ------------------------------
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}
// multiply by -i
__device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }

__device__ __forceinline__ void dft4(float2 a, float2 b, float2 c, float2 d, float2& o0, float2& o1,
                                     float2& o2, float2& o3) {
  float2 s02 = cadd(a, c), d02 = csub(a, c), s13 = cadd(b, d), d13 = mul_mi(csub(b, d));
  o0 = cadd(s02, s13);
  o2 = csub(s02, s13);
  o1 = cadd(d02, d13);
  o3 = csub(d02, d13);
}

__device__ __forceinline__ void dft5(float2 x0, float2 x1, float2 x2, float2 x3, float2 x4, float2& o0,
                                     float2& o1, float2& o2, float2& o3, float2& o4) {
  constexpr float c1 = 0.30901699437494742f;   // cos(2 pi / 5)
  constexpr float c2 = -0.80901699437494742f;  // cos(4 pi / 5)
  constexpr float s1 = 0.95105651629515357f;   // sin(2 pi / 5)
  constexpr float s2 = 0.58778525229247313f;   // sin(4 pi / 5)
  float2 p1 = cadd(x1, x4), m1 = csub(x1, x4), p2 = cadd(x2, x3), m2 = csub(x2, x3);
  o0 = make_float2(x0.x + p1.x + p2.x, x0.y + p1.y + p2.y);
  float2 a1 = make_float2(fmaf(c2, p2.x, fmaf(c1, p1.x, x0.x)), fmaf(c2, p2.y, fmaf(c1, p1.y, x0.y)));
  float2 a2 = make_float2(fmaf(c1, p2.x, fmaf(c2, p1.x, x0.x)), fmaf(c1, p2.y, fmaf(c2, p1.y, x0.y)));
  float2 b1 = make_float2(fmaf(s2, m2.x, s1 * m1.x), fmaf(s2, m2.y, s1 * m1.y));
  float2 b2 = make_float2(fmaf(-s1, m2.x, s2 * m1.x), fmaf(-s1, m2.y, s2 * m1.y));
  // X1 = a1 - i b1, X4 = a1 + i b1, X2 = a2 - i b2, X3 = a2 + i b2
  o1 = make_float2(a1.x + b1.y, a1.y - b1.x);
  o4 = make_float2(a1.x - b1.y, a1.y + b1.x);
  o2 = make_float2(a2.x + b2.y, a2.y - b2.x);
  o3 = make_float2(a2.x - b2.y, a2.y + b2.x);
}

// v[n] -> v[k], 20-point forward DFT.  n = (5 n1 + 4 n2) mod 20, k = (5 k1 + 16 k2) mod 20.
__device__ __forceinline__ void dft20(float2 (&v)[20]) {
  float2 t[4][5];
#pragma unroll
  for (int n2 = 0; n2 < 5; ++n2)
    dft4(v[(4 * n2) % 20], v[(5 + 4 * n2) % 20], v[(10 + 4 * n2) % 20], v[(15 + 4 * n2) % 20], t[0][n2],
         t[1][n2], t[2][n2], t[3][n2]);
#pragma unroll
  for (int k1 = 0; k1 < 4; ++k1)
    dft5(t[k1][0], t[k1][1], t[k1][2], t[k1][3], t[k1][4], v[(5 * k1) % 20], v[(5 * k1 + 16) % 20],
         v[(5 * k1 + 32) % 20], v[(5 * k1 + 48) % 20], v[(5 * k1 + 64) % 20]);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T> __device__ __forceinline__ float pcm_to_float(T v);
template <> __device__ __forceinline__ float pcm_to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float pcm_to_float<int16_t>(int16_t v) { return float(v) * (1.0f / 32768.0f); }

__device__ __forceinline__ void cp_async_16_zfill(void* smem_dst, const void* gmem_src, int src_bytes) {
  uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Start copying samples [s0, s0 + kTileSamples) of one clip into shared memory, zero past `valid`
// (the stft tail padding, audio.py:33 `padded=True`, and the short last clip of a long recording).
// vec_ok: clip base and length are 16-byte multiples, so every chunk is one zero-filling cp.async.
template <typename T>
__device__ __forceinline__ void stage_pcm_async(T* dst, const T* __restrict__ pcm, size_t clip_off, int s0, int valid,
                                                bool vec_ok) {
  constexpr int kVec = 16 / sizeof(T);  // elements per 128-bit chunk
  const T* clip = pcm + clip_off;
  if (vec_ok && s0 + kTileSamples <= valid) {      // interior tile: every chunk is whole
    const T* src = clip + s0;
#pragma unroll
    for (int i = threadIdx.x * kVec, it = 0; it < (kTileSamples + kThreads * kVec - 1) / (kThreads * kVec); ++it, i += kThreads * kVec)
      if (i < kTileSamples) cp_async_16_zfill(dst + i, src + i, 16);
  } else if (vec_ok) {
    for (int i = threadIdx.x * kVec; i < kTileSamples; i += kThreads * kVec) {
      int g = s0 + i;
      int n = valid - g;
      n = n < 0 ? 0 : (n > kVec ? kVec : n);
      cp_async_16_zfill(dst + i, n > 0 ? clip + g : pcm, n * int(sizeof(T)));
    }
  } else {
    for (int i = threadIdx.x; i < kTileSamples; i += kThreads) {
      int g = s0 + i;
      dst[i] = g < valid ? __ldg(clip + g) : T(0);
    }
  }
}

// norm: 0 mel power, 1 log10(max(x,1e-13)) + {sum,sumsq} partials, 2 log10(max(x,1e-10)) + {max} partials
// Work item = clip0 * tiles_per_clip + i, i < n_items: clip = item / tiles_per_clip, tile = item % tiles_per_clip.
template <typename T>
__global__ void __launch_bounds__(kThreads, 2)
logmel_kernel(const T* __restrict__ pcm, int n_samples, long long total_samples, int n_frames, int clip0, int n_items,
              const float* __restrict__ window, const float2* __restrict__ twiddle,
              const int2* __restrict__ fb_bands, const float* __restrict__ fb_w, int band_stride, int n_mels,
              const __grid_constant__ MelSchedule sched, float* __restrict__ out,
              float2* __restrict__ stats, int tiles_per_clip, int norm, int vec_ok) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FrontendSmem<T>& sm = *reinterpret_cast<FrontendSmem<T>*>(smem_raw);

  const int tid = threadIdx.x;
  const int fft = tid / 20;   // which pair of frames
  const int lane20 = tid % 20;
  const int warp = tid >> 5, lane = tid & 31;
  const float log_floor = (norm == YOHO_NORM_WHISPER) ? 1e-10f : 1e-13f;
  const int o_stride = n_mels | 1;

  auto stage = [&](int item, int buf) {
    int clip = clip0 + item / tiles_per_clip, tile = item % tiles_per_clip;
    // samples of this clip that exist in `pcm` (a long recording's last clip is short: zero padded, whisper.py:253)
    const long long left = total_samples - (long long)clip * n_samples;
    const int valid = left >= n_samples ? n_samples : (left > 0 ? int(left) : 0);
    stage_pcm_async<T>(sm.raw[buf], pcm, size_t(clip) * n_samples, tile * kFramesPerCta * kHop, valid, vec_ok != 0);
  };

  int item = blockIdx.x;
  if (item >= n_items) return;
  float win[20];
#pragma unroll
  for (int n1 = 0; n1 < 20; ++n1) win[n1] = __ldg(window + 20 * n1 + lane20);
  stage(item, 0);
  cp_async_commit();
  for (int i = tid; i < n_mels * band_stride; i += kThreads) sm.mel_w[i] = __ldg(fb_w + i);
  for (int i = tid; i < n_mels; i += kThreads) sm.mel_band[i] = __ldg(fb_bands + i);

  for (int buf = 0; item < n_items; item += gridDim.x, buf ^= 1) {
    const int clip = clip0 + item / tiles_per_clip, tile = item % tiles_per_clip;
    cp_async_wait_all();
    __syncthreads();  // this tile's samples have landed; the previous tile's power rows and red[] are free
    if (item + int(gridDim.x) < n_items) stage(item + gridDim.x, buf ^ 1);
    cp_async_commit();

    float2 v[20];
    // ---- pass 1: this thread owns n2 = lane20; DFT over n1 of x[20 n1 + n2] -----------------
    {
      const T* fa = sm.raw[buf] + (2 * fft) * kHop + lane20;  // frame A -> real part
      const T* fb = fa + kHop;                                // frame B -> imaginary part
#pragma unroll
      for (int n1 = 0; n1 < 20; ++n1)
        v[n1] = make_float2(pcm_to_float<T>(fa[20 * n1]) * win[n1], pcm_to_float<T>(fb[20 * n1]) * win[n1]);
      dft20(v);
      float2* e = sm.ez + fft * kEFft + lane20;  // E[fft][k1][n2]
#pragma unroll
      for (int k1 = 0; k1 < 20; ++k1) {
        float2 tw = __ldg(twiddle + k1 * 20 + lane20);  // W_400^(n2 k1)
        e[k1 * kERow] = cmul(v[k1], tw);
      }
    }
    __syncthreads();
    // ---- pass 2: this thread owns k1 = lane20; DFT over n2 -> Z[k1 + 20 k2] ------------------
    {
      const float2* e = sm.ez + fft * kEFft + lane20 * kERow;
#pragma unroll
      for (int n2 = 0; n2 < 20; ++n2) v[n2] = e[n2];
      dft20(v);
    }
    __syncthreads();  // every thread has read its row before Z overwrites the buffer
    {
      float2* z = sm.ez + fft * kZFft + lane20;
#pragma unroll
      for (int k2 = 0; k2 < 20; ++k2) z[20 * k2] = v[k2];
    }
    __syncthreads();
    // ---- untangle the two real spectra and take |.|^2 (scale folded into the mel weights) ----
    // bins k = lane20 + 20 j of this thread's own FFT, through registers: the power rows of frames
    // 2 fft and 2 fft + 1 replace Z of the same FFT.
    {
      const float2* z = sm.ez + fft * kZFft;
      float pa[11], pb[11];
#pragma unroll
      for (int j = 0; j < 11; ++j) {
        int k = lane20 + 20 * j;
        if (j < 10 || lane20 == 0) {
          float2 a = z[k];
          float2 b = z[k == 0 ? 0 : 400 - k];
          float ar = a.x + b.x, ai = a.y - b.y;  // 2 * X_A[k]
          float br = a.y + b.y, bi = a.x - b.x;  // 2 * X_B[k] (up to a unit factor)
          pa[j] = fmaf(ar, ar, ai * ai);
          pb[j] = fmaf(br, br, bi * bi);
        }
      }
      __syncthreads();
      float* p = reinterpret_cast<float*>(sm.ez) + lane20 * kPRow + 2 * fft;   // P[k][frame]
#pragma unroll
      for (int j = 0; j < 11; ++j)
        if (j < 10 || lane20 == 0) { p[20 * j * kPRow] = pa[j]; p[20 * j * kPRow + 1] = pb[j]; }
    }
    __syncthreads();
    // ---- banded mel projection (lane = frame), log10, per-tile statistics ----------------------
    // rows 201..203 of P are whatever Z left there (finite); the weights beyond a band are zero.
    float acc_sum = 0.0f, acc_sq = 0.0f, acc_max = -INFINITY;
    const int frame0 = tile * kFramesPerCta;
    const float* power = reinterpret_cast<const float*>(sm.ez);
    float* otile = reinterpret_cast<float*>(sm.ez) + kPRows * kPRow;          // [frame][o_stride]
    const bool lane_valid = frame0 + lane < n_frames;
    for (int round = 0; round < sched.rounds; ++round) {
      const int r = round * kWarps + ((round & 1) ? kWarps - 1 - warp : warp);
      const int n4 = sched.n4[round];                                         // the same for every warp: a uniform loop
      if (r < n_mels) {
        const int2 band = sm.mel_band[r];                                     // {mel row, first bin}
        const float* pk = power + band.y * kPRow + lane;
        const float4* w4 = reinterpret_cast<const float4*>(sm.mel_w + r * band_stride);
        float a0 = 0.0f, a1 = 0.0f;
#pragma unroll 1
        for (int i = 0; i < n4; ++i) {
          const float4 w = w4[i];
          a0 = fmaf(w.x, pk[0], a0);
          a1 = fmaf(w.y, pk[kPRow], a1);
          a0 = fmaf(w.z, pk[2 * kPRow], a0);
          a1 = fmaf(w.w, pk[3 * kPRow], a1);
          pk += 4 * kPRow;
        }
        float val = a0 + a1;
        if (norm != YOHO_NORM_NONE) {
          val = lg2_approx(fmaxf(val, log_floor)) * 0.30102999566398120f;     // the floor keeps the argument normal
          if (lane_valid) {
            acc_sum += val;
            acc_sq = fmaf(val, val, acc_sq);
            acc_max = fmaxf(acc_max, val);
          }
        }
        otile[lane * o_stride + band.x] = val;
      }
    }
    if (norm != YOHO_NORM_NONE) {
      float a = (norm == YOHO_NORM_REFERENCE) ? warp_sum(acc_sum) : warp_max(acc_max);
      float b = (norm == YOHO_NORM_REFERENCE) ? warp_sum(acc_sq) : 0.0f;
      if (lane == 0) { sm.red[warp] = a; sm.red[32 + warp] = b; }
    }
    __syncthreads();
    // ---- the tile's (frames, n_mels) block is contiguous in `out`: full-line stores ----------------
    {
      float* out_tile = out + (size_t(clip) * n_frames + frame0) * n_mels;
      const int rows = min(kFramesPerCta, n_frames - frame0);
      for (int f = warp; f < rows; f += kWarps)
        for (int m = lane; m < n_mels; m += 32) out_tile[f * n_mels + m] = otile[f * o_stride + m];
    }
    if (norm != YOHO_NORM_NONE && warp == 0) {
      float a, b;
      if (norm == YOHO_NORM_REFERENCE) {
        a = warp_sum(lane < kWarps ? sm.red[lane] : 0.0f);
        b = warp_sum(lane < kWarps ? sm.red[32 + lane] : 0.0f);
      } else {
        a = warp_max(lane < kWarps ? sm.red[lane] : -INFINITY);
        b = 0.0f;
      }
      if (lane == 0) stats[size_t(clip) * tiles_per_clip + tile] = make_float2(a, b);
    }
  }
}

// ---- per-clip statistics of an existing mel power spectrogram (stand-alone normaliser) -------
constexpr int kStatThreads = 256;
constexpr int kStatElemsPerCta = kStatThreads * 32;

__global__ void __launch_bounds__(kStatThreads)
spec_stats_kernel(const float* __restrict__ spec, int elems_per_clip, float2* __restrict__ stats,
                  int tiles_per_clip, int norm) {
  __shared__ float red[64];
  const int clip = blockIdx.y, tile = blockIdx.x, tid = threadIdx.x;
  const float* src = spec + size_t(clip) * elems_per_clip;
  const float log_floor = (norm == YOHO_NORM_WHISPER) ? 1e-10f : 1e-13f;
  float s = 0.0f, q = 0.0f, mx = -INFINITY;
  int base = tile * kStatElemsPerCta;
  for (int i = base + tid; i < min(base + kStatElemsPerCta, elems_per_clip); i += kStatThreads) {
    float val = __log2f(fmaxf(__ldg(src + i), log_floor)) * 0.30102999566398120f;
    s += val;
    q = fmaf(val, val, q);
    mx = fmaxf(mx, val);
  }
  const int warp = tid >> 5, lane = tid & 31;
  float a = (norm == YOHO_NORM_REFERENCE) ? warp_sum(s) : warp_max(mx);
  float b = (norm == YOHO_NORM_REFERENCE) ? warp_sum(q) : 0.0f;
  if (lane == 0) { red[warp] = a; red[32 + warp] = b; }
  __syncthreads();
  if (warp == 0) {
    constexpr int kWarps = kStatThreads / 32;
    if (norm == YOHO_NORM_REFERENCE) {
      a = warp_sum(lane < kWarps ? red[lane] : 0.0f);
      b = warp_sum(lane < kWarps ? red[32 + lane] : 0.0f);
    } else {
      a = warp_max(lane < kWarps ? red[lane] : -INFINITY);
    }
    if (lane == 0) stats[size_t(clip) * tiles_per_clip + tile] = make_float2(a, b);
  }
}

// ---- finalise: reduce the tile partials of a clip (in double) and normalise in place ----------
// take_log = 1 when `src` still holds mel power (stand-alone normaliser), 0 when it already holds
// log10 values written by logmel_kernel.
constexpr int kFinThreads = 256;
constexpr int kFinElemsPerCta = kFinThreads * 4 * 8;

__global__ void __launch_bounds__(kFinThreads)
normalize_apply_kernel(const float* __restrict__ src, float* __restrict__ dst, int elems_per_clip,
                       const float2* __restrict__ stats, int tiles_per_clip, int clip0, int norm,
                       int take_log) {
  __shared__ double dred[2 * (kFinThreads / 32)];
  __shared__ float coef[2];
  const int clip = clip0 + blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float2* st = stats + size_t(clip) * tiles_per_clip;
  if (norm == YOHO_NORM_REFERENCE) {
    double s = 0.0, q = 0.0;
    for (int i = tid; i < tiles_per_clip; i += kFinThreads) { float2 p = st[i]; s += p.x; q += p.y; }
#pragma unroll
    for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); q += __shfl_xor_sync(0xffffffffu, q, o); }
    if (lane == 0) { dred[warp] = s; dred[kFinThreads / 32 + warp] = q; }
    __syncthreads();
    if (tid == 0) {
      s = 0.0; q = 0.0;
      for (int w = 0; w < kFinThreads / 32; ++w) { s += dred[w]; q += dred[kFinThreads / 32 + w]; }
      double mean = s / elems_per_clip;
      double var = q / elems_per_clip - mean * mean;
      double sd = var > 0.0 ? sqrt(var) : 0.0;
      coef[0] = float(mean);
      coef[1] = float(sd != 0.0 ? 1.0 / sd : 1.0);  // audio.py:57: where(std != 0, std, 1)
    }
  } else {
    float mx = -INFINITY;
    for (int i = tid; i < tiles_per_clip; i += kFinThreads) mx = fmaxf(mx, st[i].x);
    mx = warp_max(mx);
    if (lane == 0) dred[warp] = mx;
    __syncthreads();
    if (tid == 0) {
      for (int w = 0; w < kFinThreads / 32; ++w) mx = fmaxf(mx, float(dred[w]));
      coef[0] = mx - 8.0f;
      coef[1] = 0.0f;
    }
  }
  __syncthreads();
  const float c0 = coef[0], c1 = coef[1];
  const float log_floor = (norm == YOHO_NORM_WHISPER) ? 1e-10f : 1e-13f;
  const float* s_clip = src + size_t(clip) * elems_per_clip;
  float* d_clip = dst + size_t(clip) * elems_per_clip;
  const int base = blockIdx.x * kFinElemsPerCta;
  const int end = min(base + kFinElemsPerCta, elems_per_clip);
  auto xform = [&](float x) {
    if (take_log) x = __log2f(fmaxf(x, log_floor)) * 0.30102999566398120f;
    return norm == YOHO_NORM_REFERENCE ? (x - c0) * c1 : (fmaxf(x, c0) + 4.0f) * 0.25f;
  };
  const bool vec = (elems_per_clip % 4 == 0) && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) % 16 == 0);
  if (vec) {
    for (int i = base + tid * 4; i < end; i += kFinThreads * 4) {
      float4 x = *reinterpret_cast<const float4*>(s_clip + i);
      x.x = xform(x.x); x.y = xform(x.y); x.z = xform(x.z); x.w = xform(x.w);
      *reinterpret_cast<float4*>(d_clip + i) = x;
    }
  } else {
    for (int i = base + tid; i < end; i += kFinThreads) d_clip[i] = xform(s_clip[i]);
  }
}

// ---- host side ---------------------------------------------------------------------------------
static int frames_of(int n_samples, int n_fft, int hop) {
  if (n_samples < n_fft || n_fft <= 0 || hop <= 0 || hop > n_fft) return -1;
  int n_pad = ((-(n_samples - n_fft)) % hop + hop) % hop % n_fft;
  return (n_samples + n_pad - n_fft) / hop + 1;
}

static int logmel_tiles(int frames) { return ceil_div(frames, kFramesPerCta); }

// Clips per launch pair: keep one chunk's log-mel (written by logmel_kernel, rewritten in place by
// normalize_apply_kernel) well inside the 126 MB L2 so the second pass costs no DRAM traffic.
static int clips_per_chunk(int frames, int n_mels) {
  size_t per_clip = size_t(frames) * n_mels * sizeof(float);
  size_t c = (48u << 20) / (per_clip ? per_clip : 1);
  return int(c < 1 ? 1 : (c > 65535 ? 65535 : c));
}

template <typename T>
static int logmel_launch(yoho_ctx* ctx, const yoho_filterbank* fb, const T* pcm, int batch, int n_samples, long long total_samples,
                         int n_fft, int hop, float* out, int norm, void* ws, size_t ws_bytes,
                         cudaStream_t stream) {
  if (batch == 0 && ctx && fb) return YOHO_OK;      // an empty batch / recording has no buffers to point at
  if (!ctx || !fb || !pcm || !out) return fail(YOHO_ERR_INVALID, "yoho_logmel: null argument");
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "yoho_logmel: device %d is sm_%d%d, kernels are sm_100a only", ctx->device, ctx->cc_major, ctx->cc_minor);
  if (batch < 0 || n_samples <= 0) return fail(YOHO_ERR_INVALID, "yoho_logmel: batch=%d n_samples=%d", batch, n_samples);
  if (norm < YOHO_NORM_NONE || norm > YOHO_NORM_WHISPER) return fail(YOHO_ERR_INVALID, "yoho_logmel: norm=%d", norm);
  if (n_fft != kNfft || hop != kHop)
    return fail(YOHO_ERR_UNSUPPORTED, "yoho_logmel: kernels exist for n_fft=400, hop=160 (got %d, %d)", n_fft, hop);
  if (fb->n_mels > kMaxMels) return fail(YOHO_ERR_UNSUPPORTED, "yoho_logmel: n_mels=%d, the kernel's output tile holds up to %d", fb->n_mels, kMaxMels);
  if (fb->n_bins != kBins) return fail(YOHO_ERR_INVALID, "yoho_logmel: filterbank has %d bins, n_fft=400 needs 201", fb->n_bins);
  if (fb->device != ctx->device) return fail(YOHO_ERR_INVALID, "yoho_logmel: filterbank lives on device %d, context on %d", fb->device, ctx->device);
  int frames = frames_of(n_samples, n_fft, hop);
  if (frames <= 0) return fail(YOHO_ERR_INVALID, "yoho_logmel: n_samples=%d shorter than n_fft=%d", n_samples, n_fft);
  if (batch == 0) return YOHO_OK;
  size_t need = yoho_logmel_workspace_bytes(batch, n_samples, n_fft, hop, fb->n_mels);
  if (norm != YOHO_NORM_NONE && (ws == nullptr || ws_bytes < need))
    return fail(YOHO_ERR_WORKSPACE, "yoho_logmel: workspace %zu < %zu bytes", ws_bytes, need);

  static_assert(sizeof(FrontendSmem<T>) <= 113 * 1024, "two CTAs per SM");
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(logmel_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FrontendSmem<T>))));
  const int tiles = logmel_tiles(frames);
  float2* stats = static_cast<float2*>(ws);
  const int vec_ok = (n_samples % int(16 / sizeof(T)) == 0) && (reinterpret_cast<uintptr_t>(pcm) % 16 == 0);
  const int elems = frames * fb->n_mels;
  const int chunk = (norm == YOHO_NORM_NONE) ? 65535 : clips_per_chunk(frames, fb->n_mels);
  for (int c0 = 0; c0 < batch; c0 += chunk) {
    int nb = batch - c0 < chunk ? batch - c0 : chunk;
    {
      ProfScope prof(stream, kProfFrontend, double(nb) * (double(n_samples) * sizeof(T) + double(elems) * sizeof(float)));
      const long long items = (long long)nb * tiles;
      const int grid = int(items < 2LL * ctx->num_sms ? items : 2LL * ctx->num_sms);
      logmel_kernel<T><<<grid, kThreads, sizeof(FrontendSmem<T>), stream>>>(
          pcm, n_samples, total_samples, frames, c0, int(items), ctx->d_window, ctx->d_twiddle, fb->d_bands, fb->d_weights,
          fb->band_stride, fb->n_mels, fb->sched, out, stats, tiles, norm, vec_ok);
      YOHO_LAUNCH_CHECK("logmel_kernel");
    }
    if (norm != YOHO_NORM_NONE) {
      ProfScope prof(stream, kProfNormalize, double(nb) * double(elems) * 2.0 * sizeof(float));
      normalize_apply_kernel<<<dim3(ceil_div(elems, kFinElemsPerCta), nb), kFinThreads, 0, stream>>>(
          out, out, elems, stats, tiles, c0, norm, 0);
      YOHO_LAUNCH_CHECK("normalize_apply_kernel");
    }
  }
  return YOHO_OK;
}

}  // namespace yoho

using namespace yoho;

extern "C" int yoho_stft_frames(int n_samples, int n_fft, int hop) {
  int f = frames_of(n_samples, n_fft, hop);
  if (f <= 0) return fail(YOHO_ERR_INVALID, "yoho_stft_frames: n_samples=%d n_fft=%d hop=%d", n_samples, n_fft, hop);
  return f;
}

extern "C" size_t yoho_logmel_workspace_bytes(int batch, int n_samples, int n_fft, int hop, int n_mels) {
  (void)n_mels;
  int frames = frames_of(n_samples, n_fft, hop);
  if (frames <= 0 || batch <= 0) return 1024;
  return align_up(size_t(batch) * logmel_tiles(frames) * sizeof(float2), 1024);
}

extern "C" int yoho_logmel_f32(yoho_ctx* ctx, const yoho_filterbank* fb, const float* pcm, int batch,
                               int n_samples, int n_fft, int hop, float* out, int norm, void* ws,
                               size_t ws_bytes, yoho_stream_t stream) {
  return logmel_launch<float>(ctx, fb, pcm, batch, n_samples, (long long)batch * n_samples, n_fft, hop, out, norm, ws, ws_bytes,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int yoho_logmel_s16(yoho_ctx* ctx, const yoho_filterbank* fb, const int16_t* pcm, int batch,
                               int n_samples, int n_fft, int hop, float* out, int norm, void* ws,
                               size_t ws_bytes, yoho_stream_t stream) {
  return logmel_launch<int16_t>(ctx, fb, pcm, batch, n_samples, (long long)batch * n_samples, n_fft, hop, out, norm, ws, ws_bytes,
                                static_cast<cudaStream_t>(stream));
}

extern "C" int yoho_chunk_count(long long total_samples, int clip_samples) {
  if (total_samples < 0 || clip_samples <= 0) return fail(YOHO_ERR_INVALID, "yoho_chunk_count: total=%lld clip=%d", total_samples, clip_samples);
  long long n = (total_samples + clip_samples - 1) / clip_samples;
  if (n > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "yoho_chunk_count: %lld clips", n);
  return int(n);
}

template <typename T>
static int logmel_chunked(yoho_ctx* ctx, const yoho_filterbank* fb, const T* recording, long long total_samples, int clip_samples,
                          int n_fft, int hop, float* out, int norm, void* ws, size_t ws_bytes, yoho_stream_t stream) {
  if (total_samples < 0) return fail(YOHO_ERR_INVALID, "yoho_logmel_chunked: total_samples=%lld", total_samples);
  int clips = yoho_chunk_count(total_samples, clip_samples);
  if (clips < 0) return clips;
  return logmel_launch<T>(ctx, fb, recording, clips, clip_samples, total_samples, n_fft, hop, out, norm, ws, ws_bytes,
                          static_cast<cudaStream_t>(stream));
}

extern "C" int yoho_logmel_chunked_f32(yoho_ctx* ctx, const yoho_filterbank* fb, const float* recording, long long total_samples,
                                       int clip_samples, int n_fft, int hop, float* out, int norm, void* ws, size_t ws_bytes,
                                       yoho_stream_t stream) {
  return logmel_chunked<float>(ctx, fb, recording, total_samples, clip_samples, n_fft, hop, out, norm, ws, ws_bytes, stream);
}

extern "C" int yoho_logmel_chunked_s16(yoho_ctx* ctx, const yoho_filterbank* fb, const int16_t* recording, long long total_samples,
                                       int clip_samples, int n_fft, int hop, float* out, int norm, void* ws, size_t ws_bytes,
                                       yoho_stream_t stream) {
  return logmel_chunked<int16_t>(ctx, fb, recording, total_samples, clip_samples, n_fft, hop, out, norm, ws, ws_bytes, stream);
}

extern "C" size_t yoho_normalize_workspace_bytes(int batch, int frames, int n_mels) {
  if (batch <= 0 || frames <= 0 || n_mels <= 0) return 1024;
  long long elems = (long long)frames * n_mels;
  return align_up(size_t(batch) * size_t((elems + kStatElemsPerCta - 1) / kStatElemsPerCta) * sizeof(float2), 1024);
}

extern "C" int yoho_normalize_f32(yoho_ctx* ctx, const float* spec, int batch, int frames, int n_mels,
                                  float* out, int norm, void* ws, size_t ws_bytes, yoho_stream_t stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (!ctx || !spec || !out) return fail(YOHO_ERR_INVALID, "yoho_normalize_f32: null argument");
  if (ctx->cc_major != 10) return fail(YOHO_ERR_ARCH, "yoho_normalize_f32: device is sm_%d%d, kernels are sm_100a only", ctx->cc_major, ctx->cc_minor);
  if (norm != YOHO_NORM_REFERENCE && norm != YOHO_NORM_WHISPER) return fail(YOHO_ERR_INVALID, "yoho_normalize_f32: norm=%d", norm);
  if (batch < 0 || frames <= 0 || n_mels <= 0) return fail(YOHO_ERR_INVALID, "yoho_normalize_f32: bad shape (%d,%d,%d)", batch, frames, n_mels);
  if ((long long)frames * n_mels > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "yoho_normalize_f32: clip larger than 2^31 elements");
  if (batch == 0) return YOHO_OK;
  size_t need = yoho_normalize_workspace_bytes(batch, frames, n_mels);
  if (!ws || ws_bytes < need) return fail(YOHO_ERR_WORKSPACE, "yoho_normalize_f32: workspace %zu < %zu bytes", ws_bytes, need);
  const int elems = frames * n_mels;
  const int tiles = ceil_div(elems, kStatElemsPerCta);
  float2* stats = static_cast<float2*>(ws);
  for (int c0 = 0; c0 < batch; c0 += 65535) {
    int nb = batch - c0 < 65535 ? batch - c0 : 65535;
    ProfScope prof(stream, kProfNormalize, double(nb) * double(elems) * 3.0 * sizeof(float));
    spec_stats_kernel<<<dim3(tiles, nb), kStatThreads, 0, stream>>>(spec + size_t(c0) * elems, elems,
                                                                   stats + size_t(c0) * tiles, tiles, norm);
    YOHO_LAUNCH_CHECK("spec_stats_kernel");
    normalize_apply_kernel<<<dim3(ceil_div(elems, kFinElemsPerCta), nb), kFinThreads, 0, stream>>>(
        spec, out, elems, stats, tiles, c0, norm, 1);
    YOHO_LAUNCH_CHECK("normalize_apply_kernel");
  }
  return YOHO_OK;
}
