// bf16 x bf16 -> fp32 GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM),
// operands staged by TMA into 128-byte-swizzled shared memory, persistent, warp-specialised, CTA pairs.
//
// One kernel serves every dense contraction of the encoder (reference whisper.py):
//   Conv_0 (:125)  plain GEMM over the im2col'd mel rows (K = 3*n_mels padded to 64)
//   Conv_1 (:127)  implicit GEMM, 3 taps: the (B, 2T, D) activations are viewed as (B, T, 2D) row
//                  pairs, so stride 2 becomes three shifted TMA boxes (x[2t-1], x[2t], x[2t+1]);
//                  the zero padding is TMA out-of-bounds fill (row -1 of each clip)
//   q/k/v (:32-34) one GEMM against the fused (3D, D) weight
//   out_proj (:35), Dense_0/Dense_1 (:83,85)
// and fuses what follows each of them into the epilogue: bias, tanh-GELU, the sinusoid table add
// (:23), the residual add (:79,86) and the zeroing of padded rows.
//
// Cluster = 2 CTAs on one TPC computing a 256 x BN output tile with tcgen05 cta_group::2: each CTA stages
// its own 128 A rows and HALF of the B rows (BN/2) per 64-wide K step, the leader CTA issues
// M256 x N(BN) x K16 MMAs that read both CTAs' shared memory and write 128 accumulator rows into each
// CTA's TMEM.  This halves the B traffic out of L2 per FLOP compared with two independent 128-row CTAs.
//
// CTA = 384 threads: warp 0 TMA producer, warp 1 MMA issuer (leader only), warp 2 TMEM allocator,
// warps 4-11 epilogue (warp w owns TMEM lanes 32*(w%4)..+31; the two warps of a quadrant alternate column chunks).  The epilogue goes TMEM -> registers ->
// (bias / GELU / +pos / mask) -> 128B-swizzled shared staging -> TMA store; the residual add is a TMA
// reduce-add into the fp32 residual stream (performed in L2, the stream is never read by the SM).
// 2 TMEM accumulator stages overlap the epilogue of tile i with the MMAs of tile i+1.
#include "kernels.cuh"
#include "ptx_sm100.cuh"

namespace yoho {

using namespace ptx;

constexpr int kBlockM = 128;       // accumulator rows per CTA
constexpr int kPairM = 2 * kBlockM;  // rows per cluster tile
constexpr int kBlockK = 64;        // 64 bf16 = 128 B = one swizzle row
constexpr int kGemmThreads = 384;
constexpr int kEpiWarps = 8;        // two per TMEM lane quadrant: each takes every other column chunk of its 32 rows
constexpr int kStageBufBytes = 32 * 128;   // one staging buffer: 32 rows x 128 B

struct GemmParams {
  int batch, rows_per_batch, n;
  int tiles_m, tiles_n, total_tiles;
  int n_taps, kblocks_per_tap, k_per_tap;
  int tap_col[3], tap_row[3];
  int out_is_f32;
  const float* bias;
  int gelu;
  const float* pos;
  int pos_period;
  int accumulate;
  int mask_period, mask_valid;
};

template <int BN> struct GemmCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;          // 16 KB
  static constexpr int kBBytes = (BN / 2) * kBlockK * 2;         // this CTA's half of the B tile
  static constexpr int kStageBytes = kABytes + kBBytes;
  // Operand ring depth decides how much TMA latency the MMA stream tolerates: with 5 stages of 32 KB the issuer waited on
  // full_bar for ~2/3 of its time (ncu, tensor pipe 75 % active).  BN = 256 therefore gives the epilogue warps one
  // staging buffer each (they are far off the critical path) and spends the 32 KB on a sixth operand stage.
#ifndef YOHO_GEMM_STAGES_256
#define YOHO_GEMM_STAGES_256 6
#endif
  static constexpr int kStages = BN == 256 ? YOHO_GEMM_STAGES_256 : 6;
  static constexpr int kEpiBufs = (BN == 256 && kStages > 5) ? 1 : 2;     // staging buffers per epilogue warp
  static constexpr int kStagingBytes = kEpiWarps * kEpiBufs * kStageBufBytes;   // 32 or 64 KB
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(kSmemBytes <= 232448, "exceeds the 227 KB dynamic shared memory limit");
};

__device__ __forceinline__ float gelu_tanh_fast(float x) {
  // flax nn.gelu(approximate=True); tanh.approx.f32 (MUFU) is ample for the bf16 path
  float inner = 0.7978845608028654f * fmaf(0.044715f * x, x * x, x);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(inner));
  return 0.5f * x * (1.0f + t);
}

// the same on a pair of values with packed fp32 arithmetic (5 FFMA2 / FMUL2 + 2 MUFU per pair instead of 12 scalar
// operations): the GELU epilogue of fc1 is the longest of the encoder's epilogues and close to its main loop
__device__ __forceinline__ void gelu_tanh_fast2(float& a, float& b) {
  const uint64_t x = pack_f32x2(a, b);
  const uint64_t k0 = pack_f32x2(0.7978845608028654f, 0.7978845608028654f);
  const uint64_t k1 = pack_f32x2(0.7978845608028654f * 0.044715f, 0.7978845608028654f * 0.044715f);
  const uint64_t half = pack_f32x2(0.5f, 0.5f);
  float u0, u1;
  unpack_f32x2(mul_f32x2(x, fma_f32x2(mul_f32x2(x, x), k1, k0)), u0, u1);
  float t0, t1;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t0) : "f"(u0));
  asm("tanh.approx.f32 %0, %1;" : "=f"(t1) : "f"(u1));
  unpack_f32x2(mul_f32x2(x, fma_f32x2(pack_f32x2(t0, t1), half, half)), a, b);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

// Epilogue math on 32 consecutive output columns [col0, col0+32) of this thread's row.
__device__ __forceinline__ void epilogue_math(float (&v)[32], const GemmParams& p, int col0, const float* pos_row, bool masked) {
  if (p.bias) {
    const float4* bp = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 bv = __ldg(bp + j);
      v[4 * j] += bv.x; v[4 * j + 1] += bv.y; v[4 * j + 2] += bv.z; v[4 * j + 3] += bv.w;
    }
  }
  if (p.gelu) {
#pragma unroll
    for (int j = 0; j < 32; j += 2) gelu_tanh_fast2(v[j], v[j + 1]);
  }
  if (pos_row) {
    const float4* pp = reinterpret_cast<const float4*>(pos_row + col0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 pv = __ldg(pp + j);
      v[4 * j] += pv.x; v[4 * j + 1] += pv.y; v[4 * j + 2] += pv.z; v[4 * j + 3] += pv.w;
    }
  }
  if (masked) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = 0.0f;
  }
}

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  unsigned char* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  unsigned char* stage_base = smem;
  unsigned char* staging = smem + Cfg::kStages * Cfg::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
  uint64_t* empty_bar = full_bar + Cfg::kStages;
  uint64_t* tmem_full = empty_bar + Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  // warp index through a shuffle: ptxas then treats it (and everything in the role branches that derives from it) as
  // warp-uniform, which keeps the producer's and the MMA issuer's loops on the uniform datapath
  const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();          // 0 = leader (issues the MMAs)
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_w);
    prefetch_tmap(&map_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < Cfg::kStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 2 * kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2cta<Cfg::kTmemCols>(tmem_slot);
    tmem_relinquish_2cta();
  }
  tc_fence_before();
  cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / TMA signal
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  const int kblocks = p.n_taps * p.kblocks_per_tap;

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    int stage = 0; uint32_t phase = 0;
    for (int tile = cluster_id; tile < p.total_tiles; tile += n_clusters) {
      const int nb = tile % p.tiles_n, bm = tile / p.tiles_n;
      const int b = bm / p.tiles_m, mb = bm - b * p.tiles_m;
      const int row0 = mb * kPairM + int(rank) * kBlockM;
      const int ncol0 = nb * BN + int(rank) * (BN / 2);
      for (int tap = 0; tap < p.n_taps; ++tap) {
        for (int kb = 0; kb < p.kblocks_per_tap; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (lane == 0) {
            unsigned char* sa = stage_base + stage * Cfg::kStageBytes;
            unsigned char* sb = sa + Cfg::kABytes;
            // the leader's barrier collects the bytes of both CTAs' loads for this stage
            if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
            const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);
            tma_load_3d_2sm(sa, &map_a, bar, p.tap_col[tap] + kb * kBlockK, row0 + p.tap_row[tap], b);
            tma_load_3d_2sm(sb, &map_w, bar, tap * p.k_per_tap + kb * kBlockK, ncol0, 0);
          }
          __syncwarp();
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0) {
      // ===================== MMA issuer (leader CTA) =====================
      constexpr uint32_t idesc = make_idesc_bf16(kPairM, BN, 0, 0);
      const uint64_t adesc0 = make_desc_sw128(smem_u32(stage_base), 16, 1024);
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = cluster_id; tile < p.total_tiles; tile += n_clusters, ++it) {
        const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + uint32_t(as * BN);
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          // The warp stays converged and one elected lane issues (umma_*_w): a divergent `if (lane == 0)` block cost
          // ~120 dependent instructions per k-block, more than the 512 tensor-pipe cycles its four MMAs take, and made
          // this warp -- not the tensor pipe -- the limit (ncu: no wait on full_bar, pipe 70 % active).  The operand
          // descriptors are linear in the stage index.
          const uint64_t adesc = adesc0 + uint64_t(stage) * uint64_t(Cfg::kStageBytes >> 4);
          const uint64_t bdesc = adesc + uint64_t(Cfg::kABytes >> 4);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            umma_bf16_ss_2cta_w(d_tmem, adesc + uint64_t(k * 2), bdesc + uint64_t(k * 2), idesc, (kb | k) != 0);
          umma_commit_2cta_mcast_w(&empty_bar[stage], 0x3);          // both CTAs' smem slots reusable once these retire
          if (kb == kblocks - 1) umma_commit_2cta_mcast_w(&tmem_full[as], 0x3);
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> swizzled smem -> TMA store / reduce-add =====================
    const int q = warp & 3;          // the TMEM lane quadrant this warp may read
    const int half = (warp - 4) >> 2;  // which of the two warps of the quadrant: takes column chunks half, half + 2, ...
    unsigned char* my_stage = staging + (warp - 4) * Cfg::kEpiBufs * kStageBufBytes;
    const int sw = lane & 7;
    int buf = 0;
    int it = 0;
    for (int tile = cluster_id; tile < p.total_tiles; tile += n_clusters, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int nb = tile % p.tiles_n, bm = tile / p.tiles_n;
      const int b = bm / p.tiles_m, mb = bm - b * p.tiles_m;
      const int warp_row0 = mb * kPairM + int(rank) * kBlockM + q * 32;
      const int r_in_batch = warp_row0 + lane;
      const bool masked = p.mask_period > 0 && (r_in_batch % p.mask_period) >= p.mask_valid;
      const float* pos_row = p.pos ? p.pos + size_t(r_in_batch % p.pos_period) * p.n : nullptr;
      const bool warp_live = warp_row0 < p.rows_per_batch;     // warp-uniform: any row of this warp inside the batch
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      // The accumulator stage goes back to the MMA warp as soon as this warp's LAST chunk of it is in registers, not after
      // that chunk's math / staging / TMA store; the arrive carries no release fence (the payload is TMEM, ordered by the
      // tcgen05 fence).  With short K (512-768) the issuer used to wait ~45 % of its time for tmem_empty.
      const uint32_t empty_remote = mapa_shared(smem_u32(&tmem_empty[as]), 0);   // leader's barrier, both CTAs arrive
      bool released = false;
      auto release_acc = [&]() {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster_relaxed(empty_remote);
        released = true;
      };
      if (warp_live) {
        if (p.out_is_f32) {
          // 32 fp32 columns (= 128 B rows) per staging buffer
#pragma unroll 1
          for (int c = half; c < BN / 32; c += 2) {
            const int col0 = nb * BN + c * 32;
            if (col0 >= p.n) break;  // warp-uniform
            uint32_t acc[32];
            tmem_ld_32x32b_x32(t_addr + uint32_t(c * 32), acc);
            tmem_ld_wait();
            if (c + 2 >= BN / 32 || col0 + 64 >= p.n) release_acc();    // warp-uniform: no further chunk for this warp
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
            epilogue_math(v, p, col0, pos_row, masked);
            unsigned char* sbuf = my_stage + buf * kStageBufBytes;
            if (lane == 0) tma_store_wait_read<Cfg::kEpiBufs - 1>();     // the previous user of this buffer has been read out
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 8; ++j)
              *reinterpret_cast<float4*>(sbuf + lane * 128 + ((j ^ sw) << 4)) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (p.accumulate) tma_reduce_add_3d(&map_c, sbuf, col0, warp_row0, b);
              else tma_store_3d(&map_c, sbuf, col0, warp_row0, b);
              tma_store_commit();
            }
            buf = (buf + 1) % Cfg::kEpiBufs;
          }
        } else {
          // 64 bf16 columns (= 128 B rows) per staging buffer
#pragma unroll 1
          for (int c = half; c < BN / 64; c += 2) {
            const int col0 = nb * BN + c * 64;
            if (col0 >= p.n) break;  // warp-uniform
            uint32_t acc0[32], acc1[32];
            tmem_ld_32x32b_x32(t_addr + uint32_t(c * 64), acc0);
            tmem_ld_32x32b_x32(t_addr + uint32_t(c * 64 + 32), acc1);
            tmem_ld_wait();
            if (c + 2 >= BN / 64 || col0 + 128 >= p.n) release_acc();   // warp-uniform: no further chunk for this warp
            unsigned char* sbuf = my_stage + buf * kStageBufBytes;
            if (lane == 0) tma_store_wait_read<Cfg::kEpiBufs - 1>();
            __syncwarp();
            {
              float v[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc0[j]);
              epilogue_math(v, p, col0, pos_row, masked);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(sbuf + lane * 128 + ((j ^ sw) << 4)) =
                    make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                               pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
            }
            if (col0 + 32 < p.n) {
              float v[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc1[j]);
              epilogue_math(v, p, col0 + 32, pos_row, masked);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4*>(sbuf + lane * 128 + (((4 + j) ^ sw) << 4)) =
                    make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                               pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_3d(&map_c, sbuf, col0, warp_row0, b);   // columns >= n and rows >= rows_per_batch are clipped
              tma_store_commit();
            }
            buf = (buf + 1) % Cfg::kEpiBufs;
          }
        }
      }
      if (!released) release_acc();     // warps without a chunk in this tile
    }
    if (lane == 0) tma_store_wait0();
  }

  tc_fence_before();
  cluster_sync_all();   // the peer may not exit (or free TMEM) while the leader's MMAs / remote arrives are in flight
  if (warp == 2) tmem_dealloc_2cta<Cfg::kTmemCols>(tmem_base);
}

template <int BN>
static int launch_gemm_bn(yoho_ctx* ctx, const GemmBf16Args& a, cudaStream_t s) {
  using Cfg = GemmCfg<BN>;
  GemmParams p{};
  p.batch = a.batch; p.rows_per_batch = a.rows_per_batch; p.n = a.n;
  p.tiles_m = ceil_div(a.rows_per_batch, kPairM);
  p.tiles_n = ceil_div(a.n, BN);
  long long total = (long long)a.batch * p.tiles_m * p.tiles_n;
  if (total > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "gemm: %lld tiles", total);
  p.total_tiles = int(total);
  p.n_taps = a.n_taps; p.k_per_tap = a.k_per_tap; p.kblocks_per_tap = ceil_div(a.k_per_tap, kBlockK);
  for (int i = 0; i < 3; ++i) { p.tap_col[i] = a.taps[i].col_off; p.tap_row[i] = a.taps[i].row_off; }
  p.out_is_f32 = a.out_f32 != nullptr; p.bias = a.bias; p.gelu = a.gelu;
  p.pos = a.pos; p.pos_period = a.pos_period > 0 ? a.pos_period : 1; p.accumulate = a.accumulate;
  p.mask_period = a.mask_period; p.mask_valid = a.mask_valid;

  CUtensorMap map_a, map_w, map_c;
  const int a_rows = a.a_rows > 0 ? a.a_rows : a.rows_per_batch;      // rows past a_rows (and past either end of a clip, for the taps) read as zero
  int r = make_tmap_bf16_3d(ctx->encode_tiled, &map_a, a.a, uint64_t(a.a_cols), uint64_t(a_rows), uint64_t(a.batch),
                            uint64_t(a.a_ld), uint64_t(a.a_ld) * uint64_t(a_rows), kBlockK, kBlockM, 1);
  if (r) return fail(YOHO_ERR_CUDA, "gemm: cuTensorMapEncodeTiled(A) failed with CUresult %d (cols=%d rows=%d batch=%d ld=%d)", r, a.a_cols, a.rows_per_batch, a.batch, a.a_ld);
  const uint64_t k_total = uint64_t(a.n_taps) * uint64_t(a.k_per_tap);
  r = make_tmap_bf16_3d(ctx->encode_tiled, &map_w, a.w, k_total, uint64_t(a.n), 1, k_total, k_total * uint64_t(a.n), kBlockK, BN / 2, 1);
  if (r) return fail(YOHO_ERR_CUDA, "gemm: cuTensorMapEncodeTiled(W) failed with CUresult %d (k=%llu n=%d)", r, (unsigned long long)k_total, a.n);
  if (a.out_f32)
    r = make_tmap_3d(ctx->encode_tiled, &map_c, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, a.out_f32, uint64_t(a.n), uint64_t(a.rows_per_batch), uint64_t(a.batch),
                     uint64_t(a.ldc) * 4, uint64_t(a.ldc) * 4 * uint64_t(a.rows_per_batch), 32, 32, 1);
  else
    r = make_tmap_3d(ctx->encode_tiled, &map_c, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, a.out_bf16, uint64_t(a.n), uint64_t(a.rows_per_batch), uint64_t(a.batch),
                     uint64_t(a.ldc) * 2, uint64_t(a.ldc) * 2 * uint64_t(a.rows_per_batch), 64, 32, 1);
  if (r) return fail(YOHO_ERR_CUDA, "gemm: cuTensorMapEncodeTiled(C) failed with CUresult %d (n=%d rows=%d batch=%d ldc=%d)", r, a.n, a.rows_per_batch, a.batch, a.ldc);

  YOHO_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  const int max_clusters = ctx->num_sms / 2;
  const int clusters = p.total_tiles < max_clusters ? p.total_tiles : max_clusters;
  ProfScope prof(s, kProfGemm, 2.0 * double(a.batch) * double(a.rows_per_batch) * double(a.n) * double(a.flop_k > 0 ? a.flop_k : a.n_taps * a.k_per_tap));
  gemm_bf16_kernel<BN><<<2 * clusters, kGemmThreads, Cfg::kSmemBytes, s>>>(map_a, map_w, map_c, p);
  YOHO_LAUNCH_CHECK("gemm_bf16_kernel");
  return YOHO_OK;
}

int launch_gemm_bf16(yoho_ctx* ctx, const GemmBf16Args& a, cudaStream_t s) {
  if (!a.a || !a.w || (!a.out_bf16 && !a.out_f32)) return fail(YOHO_ERR_INVALID, "gemm: null operand");
  if (a.batch <= 0 || a.rows_per_batch <= 0) return YOHO_OK;
  if (a.n <= 0 || a.n % 32 != 0) return fail(YOHO_ERR_UNSUPPORTED, "gemm: n=%d must be a positive multiple of 32", a.n);
  if (a.n_taps < 1 || a.n_taps > 3) return fail(YOHO_ERR_INVALID, "gemm: n_taps=%d", a.n_taps);
  if (a.a_rows < 0 || a.a_rows > a.rows_per_batch) return fail(YOHO_ERR_INVALID, "gemm: a_rows=%d of %d", a.a_rows, a.rows_per_batch);
  if (a.k_per_tap <= 0 || a.k_per_tap % 8 != 0 || (a.n_taps > 1 && a.k_per_tap % kBlockK != 0))
    return fail(YOHO_ERR_UNSUPPORTED, "gemm: k_per_tap=%d (multiple of 8; multiple of 64 with taps)", a.k_per_tap);
  if (a.a_ld % 8 != 0 || a.ldc % 8 != 0) return fail(YOHO_ERR_UNSUPPORTED, "gemm: leading dimensions must be multiples of 8 (a_ld=%d ldc=%d)", a.a_ld, a.ldc);
  if ((reinterpret_cast<uintptr_t>(a.a) | reinterpret_cast<uintptr_t>(a.w)) % 16 != 0) return fail(YOHO_ERR_INVALID, "gemm: operands must be 16-byte aligned");
  const void* out = a.out_f32 ? static_cast<const void*>(a.out_f32) : static_cast<const void*>(a.out_bf16);
  if (reinterpret_cast<uintptr_t>(out) % 16 != 0) return fail(YOHO_ERR_INVALID, "gemm: output must be 16-byte aligned");
  if (a.accumulate && !a.out_f32) return fail(YOHO_ERR_INVALID, "gemm: accumulate needs a float32 output");
  return (a.n % 256 == 0 || a.n > 1024) ? launch_gemm_bn<256>(ctx, a, s) : launch_gemm_bn<128>(ctx, a, s);
}

}  // namespace yoho
