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
// The kernel is bound by shared-memory bandwidth and instruction issue, not by HBM (2.9 MB and ~30 MFLOP per clip), so
// its design minimises shared-memory wavefronts and issue slots per frame:
//
//   * One tile = 32 consecutive frames of one clip = 16 complex FFTs of length 400 (two real frames packed as re/im).
//     400 = 20 x 20; a thread does radix-20 butterflies (prime-factor 4 x 5, no internal twiddles) in registers.
//   * Every thread works on TWO columns at once in packed fp32 pairs (add/mul/fma.f32x2: one issue slot for two lanes
//     of work), structure-of-arrays: a register pair holds the real (or imaginary) parts of the two columns, so complex
//     arithmetic needs no swaps and the +-i rotations of the butterflies are free.
//   * lane <-> FFT (8 FFTs x 2 column pairs per half warp), warp <-> column pair: window and twiddle factors are the same for
//     all FFTs of a column pair and come from the constant bank, not from shared memory; every shared-memory access of the
//     two passes is bank-conflict free by construction (padding of 16 bytes per 320 samples, 21-row blocks, XOR swizzle of P).
//   * Pass 1 (columns 2t, 2t+1 of x[20 n1 + n2]): frame B = frame A shifted by 160 samples = 8 rows, so 28 row loads serve
//     both frames.  Its outputs are written for pass 2 as the pair (column j, column 20-j): E_j = D_j T_j and
//     conj(D_{20-j}) T_j with one twiddle T_j = W_400^(n2 j) for both.
//   * Pass 2 transforms the pair (E_j, conj-reversed E_{20-j}) in one packed radix-20: slot x then holds Z[k] and slot y
//     conj(Z[400-k]) at the SAME register index, so the two real spectra are untangled and squared in registers
//     (|2 X_A[k]|^2, |2 X_B[k]|^2 = U^2 + V^2 with U, V sums / differences of the halves): no Z round trip through shared
//     memory.  Columns 0 and 10 are their own partners and take a scalar variant in one half warp.
//   * Power goes to shared memory bin-major over the (dead) PCM tile; the mel bands (<= 2 filters touch any FFT bin) are
//     applied with lane = (four frames, band slot): a warp walks a group of eight bands per round (packed FMA on two frame
//     pairs, 128-bit loads), the groups dealt to the warps by the host; the (32 frames x n_mels) result is transposed
//     through shared memory so the global stores are full 128-byte lines.
//   * Six warps per CTA: five carry the ten column-pair tasks of the two passes, the sixth owns columns 0 and 10 of pass 2
//     (their own partners: a different untangle, which would otherwise make one warp run both variants in turn) and a share
//     of the mel groups; two CTAs per SM are twelve warps, three per scheduler.
//   * Persistent CTAs (two per SM) walk the (clip, tile) list; the PCM of the next tile arrives by 16-byte zero-filling
//     cp.async into the other slot while the current tile is transformed (int16 is converted when it is read).
#include "common.cuh"
#include "ptx_sm100.cuh"

#include <cmath>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace yoho {

using ptx::add_f32x2;
using ptx::fma_f32x2;
using ptx::mul_f32x2;
using ptx::pack_f32x2;
using ptx::unpack_f32x2;

constexpr int kNfft = 400;
constexpr int kHop = 160;
constexpr int kBins = kMelBins;                                   // 201
constexpr int kFramesPerCta = 32;
constexpr int kFftsPerCta = kFramesPerCta / 2;                    // 16
constexpr int kFftThreads = 160;                                  // 10 column-pair tasks x 16 FFTs
constexpr int kThreads = 192;                                     // + one warp: columns 0 / 10 of pass 2 on their own warp, and 2 x 6
constexpr int kWarps = kThreads / 32;                             //   warps per SM are three per scheduler (2 x 5 left two schedulers with three
                                                                  //   warps and two with two: the warps of the crowded ones were late at every barrier)
constexpr int kTileSamples = (kFramesPerCta - 1) * kHop + kNfft;  // 5360
constexpr int kEBlockRows = 21;                                   // rows per k1 block of E: 20 + 1 so that neighbouring blocks differ in row parity
constexpr int kERows = 20 * kEBlockRows;                          // 420 rows of 16 floats (one per FFT)
constexpr int kPRow = kFramesPerCta;                              // floats per bin row of P[k][frame]
constexpr int kPBytes = kBins * kPRow * 4;                        // 25 728

// window[t][n1] = (w[20 n1 + 2t], w[20 n1 + 2t + 1]); [0] plain, [1] times 1/32768 (int16 PCM, whisper.py:249)
__constant__ float2 c_win[2][10][20];
// twiddle[t][j] = (Re T lo, Re T hi, Im T lo, Im T hi), T = W_400^(n2 j) for n2 = 2t (lo) and 2t + 1 (hi), j = 1..10
__constant__ float4 c_tw[10][11];

template <typename T>
struct FrontendSmem {
  static constexpr int kPad = 16 / int(sizeof(T));                // elements of padding after every 320 samples
  static constexpr int kFftStride = 320 + kPad;                   // elements between the first samples of consecutive FFTs
  static constexpr int kRawElems = kTileSamples + 17 * kPad;
  static constexpr int kSlotBytes = kPBytes > int(kRawElems * sizeof(T)) ? kPBytes : int(kRawElems * sizeof(T));
  // slot[b]: PCM of the tile being staged / transformed; after pass 1 the samples are dead and the slot takes the tile's
  // power spectrum P[bin][frame]
  alignas(16) unsigned char slot[2][kSlotBytes];
  // pass-1 -> pass-2 transpose, structure of arrays: e_re / e_im [k1 block][position of n2][fft]; after pass 2 has
  // loaded it, the (32 frames x n_mels|1) output tile
  alignas(16) float e_re[kERows * kFftsPerCta];
  float e_im[kERows * kFftsPerCta];
  alignas(16) float4 tw[10][11];              // copies of c_tw / c_win[this T]: read with two addresses per warp (one per task),
  float2 win[10][20];                         // which shared memory broadcasts in one wavefront and the constant cache serialises
  alignas(16) float2 mel_w[kMelMaxWeights];   // (w, w) per walked bin (resident for the whole kernel)
  int4 mel_band[kMelMaxBands];
  signed char mel_sched[32];                  // [round][warp] -> group or -1
};
static_assert(kFramesPerCta * kMelMaxBands * 4 <= kERows * kFftsPerCta * 4, "the output tile overlays e_re");
static_assert(kWarps == kMelWarps, "the filterbank's [round][warp] layout is built for this CTA");

// ---- two complex numbers in packed fp32 pairs (structure of arrays) -------------------------------------------------
struct C2 { uint64_t re, im; };
__device__ __forceinline__ uint64_t sub_f32x2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ C2 cadd(C2 a, C2 b) { return {add_f32x2(a.re, b.re), add_f32x2(a.im, b.im)}; }
__device__ __forceinline__ C2 csub(C2 a, C2 b) { return {sub_f32x2(a.re, b.re), sub_f32x2(a.im, b.im)}; }
// a - i b and a + i b (the rotations cost nothing in this layout: they only exchange the roles of re and im)
__device__ __forceinline__ C2 sub_i(C2 a, C2 b) { return {add_f32x2(a.re, b.im), sub_f32x2(a.im, b.re)}; }
__device__ __forceinline__ C2 add_i(C2 a, C2 b) { return {sub_f32x2(a.re, b.im), add_f32x2(a.im, b.re)}; }
__device__ __forceinline__ uint64_t dup(float c) { return pack_f32x2(c, c); }

__device__ __forceinline__ void dft4(C2 a, C2 b, C2 c, C2 d, C2& o0, C2& o1, C2& o2, C2& o3) {
  const C2 s02 = cadd(a, c), d02 = csub(a, c), s13 = cadd(b, d), e = csub(b, d);
  o0 = cadd(s02, s13);
  o2 = csub(s02, s13);
  o1 = sub_i(d02, e);    // d02 - i (b - d)
  o3 = add_i(d02, e);
}

__device__ __forceinline__ void dft5(C2 x0, C2 x1, C2 x2, C2 x3, C2 x4, C2& o0, C2& o1, C2& o2, C2& o3, C2& o4) {
  const uint64_t c1 = dup(0.30901699437494742f);    // cos(2 pi / 5)
  const uint64_t c2 = dup(-0.80901699437494742f);   // cos(4 pi / 5)
  const uint64_t s1 = dup(0.95105651629515357f);    // sin(2 pi / 5)
  const uint64_t s2 = dup(0.58778525229247313f);    // sin(4 pi / 5)
  const uint64_t ns1 = dup(-0.95105651629515357f);
  const C2 p1 = cadd(x1, x4), m1 = csub(x1, x4), p2 = cadd(x2, x3), m2 = csub(x2, x3);
  o0 = cadd(x0, cadd(p1, p2));
  const C2 a1 = {fma_f32x2(c2, p2.re, fma_f32x2(c1, p1.re, x0.re)), fma_f32x2(c2, p2.im, fma_f32x2(c1, p1.im, x0.im))};
  const C2 a2 = {fma_f32x2(c1, p2.re, fma_f32x2(c2, p1.re, x0.re)), fma_f32x2(c1, p2.im, fma_f32x2(c2, p1.im, x0.im))};
  const C2 b1 = {fma_f32x2(s2, m2.re, mul_f32x2(s1, m1.re)), fma_f32x2(s2, m2.im, mul_f32x2(s1, m1.im))};
  const C2 b2 = {fma_f32x2(ns1, m2.re, mul_f32x2(s2, m1.re)), fma_f32x2(ns1, m2.im, mul_f32x2(s2, m1.im))};
  // X1 = a1 - i b1, X4 = a1 + i b1, X2 = a2 - i b2, X3 = a2 + i b2
  o1 = sub_i(a1, b1);
  o4 = add_i(a1, b1);
  o2 = sub_i(a2, b2);
  o3 = add_i(a2, b2);
}

// v[n] -> v[k], 20-point forward DFT of both packed columns.  n = (5 n1 + 4 n2) mod 20, k = (5 k1 + 16 k2) mod 20.
__device__ __forceinline__ void dft20(C2 (&v)[20]) {
  C2 t[4][5];
#pragma unroll
  for (int n2 = 0; n2 < 5; ++n2)
    dft4(v[(4 * n2) % 20], v[(5 + 4 * n2) % 20], v[(10 + 4 * n2) % 20], v[(15 + 4 * n2) % 20], t[0][n2], t[1][n2], t[2][n2],
         t[3][n2]);
#pragma unroll
  for (int k1 = 0; k1 < 4; ++k1)
    dft5(t[k1][0], t[k1][1], t[k1][2], t[k1][3], t[k1][4], v[(5 * k1) % 20], v[(5 * k1 + 16) % 20], v[(5 * k1 + 32) % 20],
         v[(5 * k1 + 48) % 20], v[(5 * k1 + 64) % 20]);
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

// two consecutive samples as a packed float pair
template <typename T> __device__ __forceinline__ uint64_t load_pair(const T* p);
template <> __device__ __forceinline__ uint64_t load_pair<float>(const float* p) { return *reinterpret_cast<const uint64_t*>(p); }
template <> __device__ __forceinline__ uint64_t load_pair<int16_t>(const int16_t* p) {
  const uint32_t w = *reinterpret_cast<const uint32_t*>(p);
  return pack_f32x2(float(int(short(w & 0xffffu))), float(int(w) >> 16));     // the 1/32768 lives in the window table
}

__device__ __forceinline__ uint64_t lds_u64(uint32_t addr) {
  uint64_t v;
  asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void lds_2u64(uint32_t addr, uint64_t& lo, uint64_t& hi) {
  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "r"(addr) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

__device__ __forceinline__ void cp_async_16_zfill(void* smem_dst, const void* gmem_src, int src_bytes) {
  uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gmem_src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Start copying samples [s0, s0 + kTileSamples) of one clip into a slot, zero past `valid` (the stft tail padding,
// audio.py:33 `padded=True`, and the short last clip of a long recording).  Sample n of the tile lands at element
// n + kPad * (n / 320): 16 bytes of padding after every 320 samples put consecutive FFTs 8 (float) / 4 (int16) banks apart.
// vec_ok: clip base and length are 16-byte multiples, so every chunk is one zero-filling cp.async.
template <typename T>
__device__ __forceinline__ void stage_pcm_async(T* dst, const T* __restrict__ pcm, size_t clip_off, int s0, int valid, bool vec_ok) {
  constexpr int kVec = 16 / sizeof(T);          // elements per 128-bit chunk
  constexpr int kChunks = kTileSamples / kVec;  // 1340 / 670
  constexpr int kChunksPerBlock = 320 / kVec;   // a chunk never straddles a 320-sample block
  static_assert(kTileSamples % kVec == 0 && 320 % kVec == 0, "chunking");
  const T* clip = pcm + clip_off;
  constexpr int kThreads = kFftThreads;         // the copy pattern is laid out for 160 threads; the sixth warp takes no part
  static_assert(kThreads % kChunksPerBlock == 0, "a thread's chunks advance by whole 320-sample blocks");
  if (threadIdx.x >= kThreads) return;
  if (vec_ok && s0 + kTileSamples <= valid) {
    // interior tile: chunk c = tid + it * kThreads lands at c * kVec + kVec * (c / kChunksPerBlock); both advance by constants
    T* d = dst + threadIdx.x * kVec + kVec * (threadIdx.x / kChunksPerBlock);
    const T* s = clip + s0 + threadIdx.x * kVec;
    constexpr int kFull = kChunks / kThreads, kRem = kChunks % kThreads;
#pragma unroll
    for (int it = 0; it < kFull; ++it)
      cp_async_16_zfill(d + it * (kThreads + kThreads / kChunksPerBlock) * kVec, s + it * kThreads * kVec, 16);
    if (kRem && int(threadIdx.x) < kRem)
      cp_async_16_zfill(d + kFull * (kThreads + kThreads / kChunksPerBlock) * kVec, s + kFull * kThreads * kVec, 16);
  } else if (vec_ok) {
#pragma unroll 1
    for (int c = threadIdx.x; c < kChunks; c += kThreads) {
      const int n = c * kVec, g = s0 + n;
      int left = valid - g;
      left = left < 0 ? 0 : (left > kVec ? kVec : left);
      cp_async_16_zfill(dst + n + kVec * (c / kChunksPerBlock), left > 0 ? clip + g : pcm, left * int(sizeof(T)));
    }
  } else {
    for (int n = threadIdx.x; n < kTileSamples; n += kThreads) {
      const int g = s0 + n;
      dst[n + kVec * (n / 320)] = g < valid ? __ldg(clip + g) : T(0);
    }
  }
}

// E[block][pos][fft]: the element of column n2 sits at row block * 21 + pos(n2), pos = n2 / 2 + 10 * (n2 & 1)
__device__ __forceinline__ int e_index(int block, int pos, int fft) { return (block * kEBlockRows + pos) * kFftsPerCta + fft; }

// Four frames (two packed pairs) against two bands (slots b and b + 4 of the warp's group of eight):
// sum_i P[first + i][frames] * w_i over kSteps x 2 bins each.  a0 / a1 address the even / odd steps' rows of a band (odd
// bins keep their 16-frame halves exchanged), wa its (w, w) table.  All loads first (128-bit, immediate offsets); eight
// independent accumulation chains.
struct MelBandAddr { uint32_t a0, a1, wa; };
template <int kSteps>
__device__ __forceinline__ void mel_oct(const MelBandAddr& x, const MelBandAddr& y, uint64_t (&sum_x)[2], uint64_t (&sum_y)[2]) {
  uint64_t xa[kSteps][2], xb[kSteps][2], ya[kSteps][2], yb[kSteps][2];     // [step][frame pair] of the even / odd bin
  float4 xw[kSteps], yw[kSteps];
#pragma unroll
  for (int i = 0; i < kSteps; ++i) {
    lds_2u64(x.a0 + uint32_t(i) * (2 * kPRow * 4), xa[i][0], xa[i][1]);
    lds_2u64(x.a1 + uint32_t(i) * (2 * kPRow * 4), xb[i][0], xb[i][1]);
    xw[i] = lds_f4(x.wa + uint32_t(i) * 16u);                 // (w_2i, w_2i, w_2i+1, w_2i+1)
    lds_2u64(y.a0 + uint32_t(i) * (2 * kPRow * 4), ya[i][0], ya[i][1]);
    lds_2u64(y.a1 + uint32_t(i) * (2 * kPRow * 4), yb[i][0], yb[i][1]);
    yw[i] = lds_f4(y.wa + uint32_t(i) * 16u);
  }
  uint64_t s[4], t[4];
  {
    const uint64_t xw0 = pack_f32x2(xw[0].x, xw[0].y), xw1 = pack_f32x2(xw[0].z, xw[0].w);
    const uint64_t yw0 = pack_f32x2(yw[0].x, yw[0].y), yw1 = pack_f32x2(yw[0].z, yw[0].w);
    s[0] = mul_f32x2(xa[0][0], xw0); s[1] = mul_f32x2(xa[0][1], xw0); s[2] = mul_f32x2(xb[0][0], xw1); s[3] = mul_f32x2(xb[0][1], xw1);
    t[0] = mul_f32x2(ya[0][0], yw0); t[1] = mul_f32x2(ya[0][1], yw0); t[2] = mul_f32x2(yb[0][0], yw1); t[3] = mul_f32x2(yb[0][1], yw1);
  }
#pragma unroll
  for (int i = 1; i < kSteps; ++i) {
    const uint64_t xw0 = pack_f32x2(xw[i].x, xw[i].y), xw1 = pack_f32x2(xw[i].z, xw[i].w);
    const uint64_t yw0 = pack_f32x2(yw[i].x, yw[i].y), yw1 = pack_f32x2(yw[i].z, yw[i].w);
    s[0] = fma_f32x2(xa[i][0], xw0, s[0]); s[1] = fma_f32x2(xa[i][1], xw0, s[1]);
    s[2] = fma_f32x2(xb[i][0], xw1, s[2]); s[3] = fma_f32x2(xb[i][1], xw1, s[3]);
    t[0] = fma_f32x2(ya[i][0], yw0, t[0]); t[1] = fma_f32x2(ya[i][1], yw0, t[1]);
    t[2] = fma_f32x2(yb[i][0], yw1, t[2]); t[3] = fma_f32x2(yb[i][1], yw1, t[3]);
  }
  sum_x[0] = add_f32x2(sum_x[0], add_f32x2(s[0], s[2])); sum_x[1] = add_f32x2(sum_x[1], add_f32x2(s[1], s[3]));
  sum_y[0] = add_f32x2(sum_y[0], add_f32x2(t[0], t[2])); sum_y[1] = add_f32x2(sum_y[1], add_f32x2(t[1], t[3]));
}

// |2 X_A[k]|^2 and |2 X_B[k]|^2 from a = Z[k], b = Z[400 - k] (frames A, B packed as re / im of one complex FFT)
__device__ __forceinline__ float2 untangle_power(float ax, float ay, float bx, float by) {
  const float ar = ax + bx, ai = ay - by, br = ay + by, bi = ax - bx;
  return make_float2(fmaf(ar, ar, ai * ai), fmaf(br, br, bi * bi));
}

// kNorm: 0 mel power, 1 log10(max(x,1e-13)) + {sum,sumsq} partials, 2 log10(max(x,1e-10)) + {max} partials
// Work item = clip0 * tiles_per_clip + i, i < n_items: clip = item / tiles_per_clip, tile = item % tiles_per_clip.
template <typename T, int kNorm>
__global__ void __launch_bounds__(kThreads, 2)
logmel_kernel(const T* __restrict__ pcm, int n_samples, long long total_samples, int n_frames, int clip0, int n_items,
              const int4* __restrict__ fb_bands, const float2* __restrict__ fb_w, int n_groups, int n_rounds, int n_weights, int n_mels,
              float* __restrict__ out, float2* __restrict__ stats, int tiles_per_clip, int vec_ok, int vec_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Smem = FrontendSmem<T>;
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  // FFT roles: half warp hw holds FFTs 8 hw .. 8 hw + 7, each twice: sub = 0 / 1 are the warp's two column-pair tasks
  const int fft = (lane & 7) + 8 * (lane >> 4);
  const int sub = (lane >> 3) & 1;
  const int task = 2 * warp + sub;                        // pass 1 (warps 0-4): columns 2 task, 2 task + 1
  // pass 2: columns task2, 20 - task2; columns 0 and 10 (task2 = 0, their own partners: a different untangle) belong to the
  // sixth warp, so that no warp runs both untangle variants one after the other; -1 = this half warp rests
  const int task2 = warp == kWarps - 1 ? (sub == 0 ? 0 : -1) : (task == 0 ? -1 : task);
  // mel roles: half warp h takes band h of the warp's current band pair, lane = frame pair
  const float log_floor = (kNorm == YOHO_NORM_WHISPER) ? 1e-10f : 1e-13f;
  constexpr int kWin = sizeof(T) == 2 ? 1 : 0;

  auto stage = [&](int item, int buf) {
    int clip = clip0 + item / tiles_per_clip, tile = item % tiles_per_clip;
    // samples of this clip that exist in `pcm` (a long recording's last clip is short: zero padded, whisper.py:253)
    const long long left = total_samples - (long long)clip * n_samples;
    const int valid = left >= n_samples ? n_samples : (left > 0 ? int(left) : 0);
    stage_pcm_async<T>(reinterpret_cast<T*>(sm.slot[buf]), pcm, size_t(clip) * n_samples, tile * kFramesPerCta * kHop, valid, vec_ok != 0);
  };

  pdl_launch_dependents();
  pdl_wait();                      // the mel / statistics buffers this kernel overwrites may still be read by the previous step's kernels
  int item = blockIdx.x;
  if (item >= n_items) return;
  stage(item, 0);
  cp_async_commit();
  for (int i = tid; i < 10 * 11; i += kThreads) (&sm.tw[0][0])[i] = (&c_tw[0][0])[i];
  for (int i = tid; i < 10 * 20; i += kThreads) (&sm.win[0][0])[i] = (&c_win[kWin][0][0])[i];
  for (int i = tid; i < n_weights; i += kThreads) sm.mel_w[i] = __ldg(fb_w + i);
  for (int i = tid; i < kMelGroup * n_groups; i += kThreads) sm.mel_band[i] = __ldg(fb_bands + i);
  if (tid < 32) sm.mel_sched[tid] = static_cast<signed char>(__ldg(reinterpret_cast<const int*>(fb_bands + kMelGroup * n_groups) + tid));

  // Statistics of the tile finished last.  Every thread leaves its partial {sum, sum of squares} (or {max}) in the tail of
  // the tile's slot (the power rows there are dead after the mel phase and the next staging only writes the PCM part);
  // after the next iteration's first barrier the sixth warp, which has no column pair in pass 1, adds the 192 partials in a
  // fixed order and writes the tile's entry.  No shuffle chain on the five FFT warps (it held 6 % of the stall samples).
  constexpr int kPartOffset = Smem::kSlotBytes - kThreads * int(sizeof(float2));
  static_assert(kPartOffset >= int(Smem::kRawElems * sizeof(T)) && kPartOffset % 8 == 0, "the partials sit behind the staged samples");
  int pending_stats = -1;
  auto collect_stats = [&](int slot_of_tile) {
    if (kNorm == YOHO_NORM_NONE || warp != kWarps - 1 || pending_stats < 0) return;
    const float2* part = reinterpret_cast<const float2*>(sm.slot[slot_of_tile] + kPartOffset);
    float a = (kNorm == YOHO_NORM_REFERENCE) ? 0.0f : -INFINITY, b = 0.0f;
#pragma unroll
    for (int i = 0; i < kThreads / 32; ++i) {
      const float2 v = part[lane + 32 * i];
      if (kNorm == YOHO_NORM_REFERENCE) { a += v.x; b += v.y; } else a = fmaxf(a, v.x);
    }
    a = (kNorm == YOHO_NORM_REFERENCE) ? warp_sum(a) : warp_max(a);
    if (kNorm == YOHO_NORM_REFERENCE) b = warp_sum(b);
    if (lane == 0) stats[pending_stats] = make_float2(a, b);
  };

  for (int buf = 0; item < n_items; item += gridDim.x, buf ^= 1) {
    const int clip = clip0 + item / tiles_per_clip, tile = item % tiles_per_clip;
    cp_async_wait_all();
    __syncthreads();  // this tile's samples have landed; the previous tile's power rows (other slot), output tile and partials are complete
    collect_stats(buf ^ 1);
    if (item + int(gridDim.x) < n_items) stage(item + gridDim.x, buf ^ 1);
    cp_async_commit();

    C2 v[20];
    // ---- pass 1: columns n2 = 2 task, 2 task + 1; DFT over n1 of x[20 n1 + n2], frames A / B as re / im ------------------
    if (warp < kFftThreads / 32) {
      const T* src = reinterpret_cast<const T*>(sm.slot[buf]) + fft * Smem::kFftStride + 2 * task;
      uint64_t r[28];          // rows 0..19 are frame A; frame B starts 160 samples = 8 rows later
#pragma unroll
      for (int i = 0; i < 28; ++i) r[i] = load_pair<T>(src + 20 * i + (i >= 16 ? Smem::kPad : 0));
#pragma unroll
      for (int n1 = 0; n1 < 20; ++n1) {
        const float2 wf = sm.win[task][n1];
        const uint64_t w = pack_f32x2(wf.x, wf.y);
        v[n1].re = mul_f32x2(r[n1], w);
        v[n1].im = mul_f32x2(r[n1 + 8], w);
      }
      dft20(v);              // v[k1] = D[k1][n2]
      auto store = [&](int block, C2 val) {
        float lo, hi;
        unpack_f32x2(val.re, lo, hi);
        sm.e_re[e_index(block, task, fft)] = lo;
        sm.e_re[e_index(block, task + 10, fft)] = hi;
        unpack_f32x2(val.im, lo, hi);
        sm.e_im[e_index(block, task, fft)] = lo;
        sm.e_im[e_index(block, task + 10, fft)] = hi;
      };
      store(0, v[0]);
#pragma unroll
      for (int j = 1; j <= 10; ++j) {
        const float4 tw = sm.tw[task][j];                     // T = W_400^(n2 j)
        const uint64_t tr = pack_f32x2(tw.x, tw.y), ti = pack_f32x2(tw.z, tw.w);
        const C2 d = v[j];
        // E_j = D_j T
        store(j, {sub_f32x2(mul_f32x2(d.re, tr), mul_f32x2(d.im, ti)), fma_f32x2(d.re, ti, mul_f32x2(d.im, tr))});
        if (j < 10) {
          // partner column: conj(D_{20-j}) T  (pass 2 then finds conj(Z[400 - k]) at the index where column j has Z[k])
          const C2 c = v[20 - j];
          store(20 - j, {fma_f32x2(c.re, tr, mul_f32x2(c.im, ti)), sub_f32x2(mul_f32x2(c.re, ti), mul_f32x2(c.im, tr))});
        }
      }
    }
    __syncthreads();
    // ---- pass 2: columns (task2, 20 - task2) [task2 0: columns 0 and 10] in the two slots; DFT over n2 -------------------
    if (task2 >= 0) {
      const int task = task2;                                 // shadows the pass-1 role
      const int bx = task, by = task == 0 ? 10 : 20 - task;
#pragma unroll
      for (int n2 = 0; n2 < 20; ++n2) {
        const int pos = (n2 >> 1) + 10 * (n2 & 1);
        v[n2].re = pack_f32x2(sm.e_re[e_index(bx, pos, fft)], sm.e_re[e_index(by, pos, fft)]);
        v[n2].im = pack_f32x2(sm.e_im[e_index(bx, pos, fft)], sm.e_im[e_index(by, pos, fft)]);
      }
      dft20(v);
      // power rows P[bin][frame]; odd bins keep their two 16-frame halves exchanged (the two tasks of a half warp write
      // neighbouring bins: they land in different banks)
      float* pw = reinterpret_cast<float*>(sm.slot[buf]);
      if (task != 0) {
        // slot x: Z[k], k = task + 20 k2; slot y: conj(Z[400 - k])
        float2* prow = reinterpret_cast<float2*>(pw + ((2 * fft) ^ ((task & 1) << 4)));
#pragma unroll
        for (int k2 = 0; k2 < 20; ++k2) {
          float xr, yr, xi, yi;
          unpack_f32x2(v[k2].re, xr, yr);
          unpack_f32x2(v[k2].im, xi, yi);
          const uint64_t u = pack_f32x2(xr + yr, xr - yr), w = pack_f32x2(xi + yi, xi - yi);
          const uint64_t p = fma_f32x2(w, w, mul_f32x2(u, u));      // (|2 X_A|^2, |2 X_B|^2)
          const int bin_off = k2 <= 9 ? 20 * k2 : 20 * (19 - k2) + 20;   // bin = +task or -task from here
          float2* dst = k2 <= 9 ? prow + (bin_off + task) * (kPRow / 2) : prow + (bin_off - task) * (kPRow / 2);
          *reinterpret_cast<uint64_t*>(dst) = p;
        }
      } else {
        // columns 0 (slot x) and 10 (slot y) are their own partners
        float zr0[20], zi0[20], zr1[20], zi1[20];
#pragma unroll
        for (int k2 = 0; k2 < 20; ++k2) { unpack_f32x2(v[k2].re, zr0[k2], zr1[k2]); unpack_f32x2(v[k2].im, zi0[k2], zi1[k2]); }
        float2* p_even = reinterpret_cast<float2*>(pw + 2 * fft);     // bins 20 k2 and 10 + 20 k2 are all even
#pragma unroll
        for (int k2 = 0; k2 <= 10; ++k2) {
          const int kb = (20 - k2) % 20;
          p_even[(20 * k2) * (kPRow / 2)] = untangle_power(zr0[k2], zi0[k2], zr0[kb], zi0[kb]);
        }
#pragma unroll
        for (int k2 = 0; k2 < 10; ++k2)
          p_even[(10 + 20 * k2) * (kPRow / 2)] = untangle_power(zr1[k2], zi1[k2], zr1[19 - k2], zi1[19 - k2]);
      }
    }
    __syncthreads();
    // ---- banded mel projection: lane = (four frames, band slot); a warp walks a group of eight bands per round ---------
    const int frame0 = tile * kFramesPerCta;
    // output tile [mel][32 frames] of mel POWER; E is dead: every thread loaded its rows before the barrier
    float* otile = sm.e_re;
    {
      const uint32_t pbase = ptx::smem_u32(sm.slot[buf]), wbase = ptx::smem_u32(sm.mel_w);
      const int fq = lane & 7, slot = lane >> 3;               // frames 4 fq .. 4 fq + 3; bands `slot` and `slot + 4` of the group
      const uint32_t col = uint32_t(fq) * 16u;
      auto address = [&](const int4& band) {                   // odd bins keep their 16-frame halves exchanged (64 bytes)
        const uint32_t odd = uint32_t(band.y & 1) << 6, row = pbase + uint32_t(band.y) * (kPRow * 4);
        return MelBandAddr{row + (col ^ odd), row + (kPRow * 4) + (col ^ odd ^ 64u), wbase + uint32_t(band.z) * 8u};
      };
      // the host has dealt the groups to the six warps (longest processing time first): schedule[round][warp]
      for (int round = 0; round < n_rounds; ++round) {
        const int group = sm.mel_sched[round * kWarps + warp];
        if (group < 0) continue;
        const int4 bx = sm.mel_band[kMelGroup * group + slot], by = sm.mel_band[kMelGroup * group + 4 + slot];
        int steps = bx.w >> 1;                                  // two bins a step; the same for the eight bands of the group
        MelBandAddr ax = address(bx), ay = address(by);
        uint64_t sum_x[2] = {0ull, 0ull}, sum_y[2] = {0ull, 0ull};   // a group of empty bands (no steps) still stores its zeros
        // one straight-line variant per step count up to four; longer bands take chunks of four steps first
        while (steps > 4) {
          mel_oct<4>(ax, ay, sum_x, sum_y);
          ax.a0 += 4 * 2 * kPRow * 4; ax.a1 += 4 * 2 * kPRow * 4; ax.wa += 4 * 16;
          ay.a0 += 4 * 2 * kPRow * 4; ay.a1 += 4 * 2 * kPRow * 4; ay.wa += 4 * 16;
          steps -= 4;
        }
        switch (steps) {
          case 0: break;
          case 1: mel_oct<1>(ax, ay, sum_x, sum_y); break;
          case 2: mel_oct<2>(ax, ay, sum_x, sum_y); break;
          case 3: mel_oct<3>(ax, ay, sum_x, sum_y); break;
          default: mel_oct<4>(ax, ay, sum_x, sum_y); break;
        }
        if (bx.x >= 0) *reinterpret_cast<ulonglong2*>(otile + bx.x * kFramesPerCta + 4 * fq) = make_ulonglong2(sum_x[0], sum_x[1]);
        if (by.x >= 0) *reinterpret_cast<ulonglong2*>(otile + by.x * kFramesPerCta + 4 * fq) = make_ulonglong2(sum_y[0], sum_y[1]);
      }
    }
    __syncthreads();
    // ---- log10, per-tile statistics and the stores: the tile's (frames, n_mels) block is contiguous in `out` -----------
    {
      float acc_sum = 0.0f, acc_sq = 0.0f, acc_max = -INFINITY;
      auto finish = [&](float v, bool valid) {
        if (kNorm == YOHO_NORM_NONE) return v;
        v = lg2_approx(fmaxf(v, log_floor)) * 0.30102999566398120f;      // the floor keeps the argument normal
        if (kNorm == YOHO_NORM_REFERENCE) { if (valid) { acc_sum += v; acc_sq = fmaf(v, v, acc_sq); } }
        else if (valid) acc_max = fmaxf(acc_max, v);
        return v;
      };
      float* out_tile = out + (size_t(clip) * n_frames + frame0) * n_mels;
      const int rows = min(kFramesPerCta, n_frames - frame0);
      if (vec_out) {
        // item = (4 mels) x (4 frames): four 128-bit reads along the frames, transposed in registers (by naming), four
        // 128-bit stores along the mels; consecutive lanes take consecutive frame groups (conflict-free reads) and
        // every 32-byte sector of `out` is written whole
        const int items = (n_mels >> 2) * 8;
        for (int idx = tid; idx < items; idx += kThreads) {
          const int f4 = idx & 7, m4 = idx >> 3;
          const float4* src = reinterpret_cast<const float4*>(otile + (4 * m4) * kFramesPerCta + 4 * f4);
          const float4 r0 = src[0], r1 = src[kFramesPerCta / 4], r2 = src[2 * (kFramesPerCta / 4)], r3 = src[3 * (kFramesPerCta / 4)];
          float* dst = out_tile + size_t(4 * f4) * n_mels + 4 * m4;
          const bool v0 = 4 * f4 + 0 < rows, v1 = 4 * f4 + 1 < rows, v2 = 4 * f4 + 2 < rows, v3 = 4 * f4 + 3 < rows;
          const float4 o0 = make_float4(finish(r0.x, v0), finish(r1.x, v0), finish(r2.x, v0), finish(r3.x, v0));
          const float4 o1 = make_float4(finish(r0.y, v1), finish(r1.y, v1), finish(r2.y, v1), finish(r3.y, v1));
          const float4 o2 = make_float4(finish(r0.z, v2), finish(r1.z, v2), finish(r2.z, v2), finish(r3.z, v2));
          const float4 o3 = make_float4(finish(r0.w, v3), finish(r1.w, v3), finish(r2.w, v3), finish(r3.w, v3));
          if (v0) *reinterpret_cast<float4*>(dst) = o0;
          if (v1) *reinterpret_cast<float4*>(dst + n_mels) = o1;
          if (v2) *reinterpret_cast<float4*>(dst + 2 * n_mels) = o2;
          if (v3) *reinterpret_cast<float4*>(dst + 3 * n_mels) = o3;
        }
      } else {
        for (int f = warp; f < rows; f += kWarps)
          for (int m = lane; m < n_mels; m += 32) out_tile[f * n_mels + m] = finish(otile[m * kFramesPerCta + f], true);
      }
      if (kNorm != YOHO_NORM_NONE) {
        reinterpret_cast<float2*>(sm.slot[buf] + kPartOffset)[tid] =
            make_float2(kNorm == YOHO_NORM_REFERENCE ? acc_sum : acc_max, kNorm == YOHO_NORM_REFERENCE ? acc_sq : 0.0f);
        pending_stats = clip * tiles_per_clip + tile;
      }
    }
    if (item + int(gridDim.x) >= n_items && kNorm != YOHO_NORM_NONE) {   // the CTA's last tile: nobody comes back to the top
      __syncthreads();
      collect_stats(buf);
    }
  }
}

// ---- constant tables ------------------------------------------------------------------------------------------------
int frontend_upload_constants(double* window_sum) {
  // Periodic Hann, scipy.signal.get_window('hann', 400): float64 then rounded to float32, exactly what
  // jax.scipy.signal.stft multiplies by (reference audio.py:33).
  const double kPi = 3.14159265358979323846;
  std::vector<float> win(400);
  double wsum = 0.0;
  for (int n = 0; n < 400; ++n) {
    win[n] = float(0.5 - 0.5 * std::cos(2.0 * kPi * n / 400.0));
    wsum += double(win[n]);
  }
  *window_sum = wsum;
  static float2 h_win[2][10][20];
  static float4 h_tw[10][11];
  for (int t = 0; t < 10; ++t) {
    for (int n1 = 0; n1 < 20; ++n1) {
      const float a = win[20 * n1 + 2 * t], b = win[20 * n1 + 2 * t + 1];
      h_win[0][t][n1] = make_float2(a, b);
      h_win[1][t][n1] = make_float2(a * (1.0f / 32768.0f), b * (1.0f / 32768.0f));   // exact (power of two)
    }
    for (int j = 0; j <= 10; ++j) {
      const double a0 = -2.0 * kPi * double(2 * t * j) / 400.0, a1 = -2.0 * kPi * double((2 * t + 1) * j) / 400.0;
      h_tw[t][j] = make_float4(float(std::cos(a0)), float(std::cos(a1)), float(std::sin(a0)), float(std::sin(a1)));
    }
  }
  YOHO_CUDA_CHECK(cudaMemcpyToSymbol(c_win, h_win, sizeof(h_win)));
  YOHO_CUDA_CHECK(cudaMemcpyToSymbol(c_tw, h_tw, sizeof(h_tw)));
  return YOHO_OK;
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
  pdl_launch_dependents();
  pdl_wait();
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

// Clips per launch pair.  Round 1 kept one chunk's log-mel (written by logmel_kernel, rewritten in place by
// normalize_apply_kernel) inside the L2 (48 MB chunks); with the round-2 kernel the launch tails of five chunks cost more
// than the DRAM traffic of an unchunked second pass (256 clips in the step: 0.68 + 0.12 ms chunked, 0.63 + 0.10 ms in
// one launch pair), so the default is one launch pair per call.  YOHO_LOGMEL_CHUNK_MB (read once) restores chunking for A/B.
static int clips_per_chunk(int frames, int n_mels) {
  static const size_t chunk_mb = [] { const char* e = getenv("YOHO_LOGMEL_CHUNK_MB"); long v = e ? atol(e) : 0; return size_t(v > 0 ? v : 0); }();
  if (chunk_mb == 0) return 65535;
  size_t per_clip = size_t(frames) * n_mels * sizeof(float);
  size_t c = (chunk_mb << 20) / (per_clip ? per_clip : 1);
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
  if (fb->n_mels > kMelMaxBands) return fail(YOHO_ERR_UNSUPPORTED, "yoho_logmel: n_mels=%d, the kernel's output tile holds up to %d", fb->n_mels, kMelMaxBands);
  if (fb->n_bins != kBins) return fail(YOHO_ERR_INVALID, "yoho_logmel: filterbank has %d bins, n_fft=400 needs 201", fb->n_bins);
  if (fb->device != ctx->device) return fail(YOHO_ERR_INVALID, "yoho_logmel: filterbank lives on device %d, context on %d", fb->device, ctx->device);
  int frames = frames_of(n_samples, n_fft, hop);
  if (frames <= 0) return fail(YOHO_ERR_INVALID, "yoho_logmel: n_samples=%d shorter than n_fft=%d", n_samples, n_fft);
  if (batch == 0) return YOHO_OK;
  size_t need = yoho_logmel_workspace_bytes(batch, n_samples, n_fft, hop, fb->n_mels);
  if (norm != YOHO_NORM_NONE && (ws == nullptr || ws_bytes < need))
    return fail(YOHO_ERR_WORKSPACE, "yoho_logmel: workspace %zu < %zu bytes", ws_bytes, need);

  static_assert(2 * (sizeof(FrontendSmem<T>) + 1024) <= 228 * 1024, "two CTAs per SM");
  auto kernel = norm == YOHO_NORM_NONE ? logmel_kernel<T, YOHO_NORM_NONE>
                : norm == YOHO_NORM_REFERENCE ? logmel_kernel<T, YOHO_NORM_REFERENCE> : logmel_kernel<T, YOHO_NORM_WHISPER>;
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FrontendSmem<T>))));
  // two CTAs of ~113 KB need the full 228 KB carve-out; left to itself the driver may size it for one
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  const int tiles = logmel_tiles(frames);
  float2* stats = static_cast<float2*>(ws);
  const int vec_ok = (n_samples % int(16 / sizeof(T)) == 0) && (reinterpret_cast<uintptr_t>(pcm) % 16 == 0);
  const int vec_out = (fb->n_mels % 4 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);
  const int elems = frames * fb->n_mels;
  const int chunk = (norm == YOHO_NORM_NONE) ? 65535 : clips_per_chunk(frames, fb->n_mels);
  for (int c0 = 0; c0 < batch; c0 += chunk) {
    int nb = batch - c0 < chunk ? batch - c0 : chunk;
    {
      ProfScope prof(stream, kProfFrontend, double(nb) * (double(n_samples) * sizeof(T) + double(elems) * sizeof(float)));
      const long long items = (long long)nb * tiles;
      const int grid = int(items < 2LL * ctx->num_sms ? items : 2LL * ctx->num_sms);
      YOHO_CUDA_CHECK(launch_pdl(kernel, dim3(grid), dim3(kThreads), sizeof(FrontendSmem<T>), stream,
          pcm, n_samples, total_samples, frames, c0, int(items), fb->d_bands, fb->d_weights, fb->n_groups, fb->n_rounds, fb->n_weights,
          fb->n_mels, out, stats, tiles, vec_ok, vec_out));
      YOHO_LAUNCH_CHECK("logmel_kernel");
    }
    if (norm != YOHO_NORM_NONE) {
      ProfScope prof(stream, kProfNormalize, double(nb) * double(elems) * 2.0 * sizeof(float));
      YOHO_CUDA_CHECK(launch_pdl(normalize_apply_kernel, dim3(ceil_div(elems, kFinElemsPerCta), nb), dim3(kFinThreads), 0, stream,
          out, out, elems, stats, tiles, c0, norm, 0));
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
    YOHO_CUDA_CHECK(launch_pdl(normalize_apply_kernel, dim3(ceil_div(elems, kFinElemsPerCta), nb), dim3(kFinThreads), 0, stream,
        spec, out, elems, stats, tiles, c0, norm, 1));
    YOHO_LAUNCH_CHECK("normalize_apply_kernel");
  }
  return YOHO_OK;
}

// ---- no-allocation variant of the fused log-mel call ------------------------------------------------------------------
// For callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): the
// per-tile statistics scratch that yoho_logmel_f32 takes as `ws` lives in a file-scope device array instead.  Every GPU
// that loads the library holds these 256 KB until the process exits.  All calls on a device share the array and it
// keeps the last call's contents; nothing in it has to start at zero (every (clip, tile) slot a call reads was written
// by the same call's logmel_kernel earlier on the same stream), but two calls that overlap on the GPU (different
// streams) race on it.
namespace {

constexpr int kStaticStatsSlots = 32768;                          // float2 per-tile partials: batch * ceil(frames / 32) <= 32768
static __device__ __align__(256) float2 g_logmel_static_stats[kStaticStatsSlots];

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;
float2* g_static_ptr[kStaticMaxDevices] = {};                     // device address of the array, per device (host code must not use the symbol's name)

// Device address of g_logmel_static_stats on `device`: looked up once per device, with that device current.
int static_stats_address(int device, float2** out) {
  if (device < 0 || device >= kStaticMaxDevices) return fail(YOHO_ERR_UNSUPPORTED, "yoho_logmel_f32_static: device %d (at most %d devices)", device, kStaticMaxDevices);
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_ptr[device]) {
    int prev = -1;
    YOHO_CUDA_CHECK(cudaGetDevice(&prev));
    if (prev != device) YOHO_CUDA_CHECK(cudaSetDevice(device));
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_logmel_static_stats);
    if (prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess || !p) return fail(YOHO_ERR_CUDA, "yoho_logmel_f32_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    g_static_ptr[device] = static_cast<float2*>(p);
  }
  *out = g_static_ptr[device];
  return YOHO_OK;
}

}  // namespace

extern "C" size_t yoho_logmel_static_max_tiles(void) { return size_t(kStaticStatsSlots); }

extern "C" int yoho_logmel_f32_static(yoho_ctx* ctx, const yoho_filterbank* fb, const float* pcm, int batch, int n_samples,
                                      int n_fft, int hop, float* out, int norm, yoho_stream_t stream) {
  if (!ctx) return fail(YOHO_ERR_INVALID, "yoho_logmel_f32_static: null context");
  const int frames = frames_of(n_samples, n_fft, hop);
  if (frames <= 0 || batch < 0) return fail(YOHO_ERR_INVALID, "yoho_logmel_f32_static: batch=%d n_samples=%d n_fft=%d hop=%d", batch, n_samples, n_fft, hop);
  const long long slots = (long long)batch * logmel_tiles(frames);
  if (slots > kStaticStatsSlots)
    return fail(YOHO_ERR_WORKSPACE, "yoho_logmel_f32_static: %d clips x %d tiles = %lld statistics slots, the built-in scratch holds %d "
                "(use yoho_logmel_f32 with a caller-provided workspace)", batch, logmel_tiles(frames), slots, kStaticStatsSlots);
  float2* scratch = nullptr;
  if (int rc = static_stats_address(ctx->device, &scratch)) return rc;
  return logmel_launch<float>(ctx, fb, pcm, batch, n_samples, (long long)batch * n_samples, n_fft, hop, out, norm, scratch,
                              size_t(kStaticStatsSlots) * sizeof(float2), static_cast<cudaStream_t>(stream));
}
