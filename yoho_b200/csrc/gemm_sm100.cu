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
  // LayerNorm folding (kLn kernels only; see GemmBf16Args)
  int add_resid;
  __nv_bfloat16* xb_out; int xb_ld;
  float2* ln_stats_out; int ln_parts_out;
  const float2* ln_stats_in; int ln_parts_in; float ln_inv_dims;
};

// kLn: 0 plain; 1 LayerNorm consumer (rows scaled by rstd from the partials); 2 LayerNorm producer (residual rows fetched by
// TMA, bf16 copy, partials)
// kDeep (producers with a short K, i.e. memory-bound ones): three staging buffers per epilogue warp keep two residual
// loads in flight per warp (one is not enough: 8 warps x 4 KB per SM over ~1.2 us of DRAM latency caps at ~4 TB/s), paid
// for with operand stages the short main loop does not need
template <int BN, int kLn = 0, bool kDeep = false> struct GemmCfg {
  static constexpr int kABytes = kBlockM * kBlockK * 2;          // 16 KB
  static constexpr int kBBytes = (BN / 2) * kBlockK * 2;         // this CTA's half of the B tile
  static constexpr int kStageBytes = kABytes + kBBytes;
  // Operand ring depth decides how much TMA latency the MMA stream tolerates: with 5 stages of 32 KB the issuer waited on
  // full_bar for ~2/3 of its time (ncu, tensor pipe 75 % active).  BN = 256 therefore gives the epilogue warps one
  // staging buffer each (they are far off the critical path) and spends the 32 KB on a sixth operand stage.
#ifndef YOHO_GEMM_STAGES_256
#define YOHO_GEMM_STAGES_256 6
#endif
  // the producer's epilogue needs two staging buffers per warp (the next chunk's residual rows arrive while this chunk is
  // finished): it runs on five operand stages
  static constexpr int kStages = kDeep ? (BN == 256 ? 4 : 5) : (BN == 256 ? (kLn == 2 ? 5 : YOHO_GEMM_STAGES_256) : 6);
  static constexpr int kEpiBufs = kDeep ? 3 : ((BN == 256 && kStages > 5) ? 1 : 2);     // staging buffers per epilogue warp
  static constexpr int kStagingBytes = kEpiWarps * kEpiBufs * kStageBufBytes;   // 32 or 64 KB
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kBarrierBytes = 512;   // 2 * kStages + 4 pipeline barriers, the TMEM slot, 2 * kEpiWarps residual barriers
  static constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + kBarrierBytes;
  static_assert((2 * kStages + 4 + 1 + kEpiBufs * kEpiWarps) * 8 <= kBarrierBytes, "barrier area");
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
__device__ __forceinline__ void epilogue_math(float (&v)[32], const GemmParams& p, int col0, const float* pos_row, bool masked,
                                              bool ln_scale = false, float ln_rstd = 1.0f) {
  if (ln_scale) {
    // LN(x) W + b = rstd * (x W'') + b': the GEMM ran on the raw rows and on weights that carry the LN scale and have their
    // column means removed; b' = beta W + b.  One packed FMA per pair replaces the bias add.
    const float4* bp = reinterpret_cast<const float4*>(p.bias + col0);
    const uint64_t r2 = pack_f32x2(ln_rstd, ln_rstd);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 bv = __ldg(bp + j);
      unpack_f32x2(fma_f32x2(pack_f32x2(v[4 * j], v[4 * j + 1]), r2, pack_f32x2(bv.x, bv.y)), v[4 * j], v[4 * j + 1]);
      unpack_f32x2(fma_f32x2(pack_f32x2(v[4 * j + 2], v[4 * j + 3]), r2, pack_f32x2(bv.z, bv.w)), v[4 * j + 2], v[4 * j + 3]);
    }
  } else if (p.bias) {
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

template <int BN, int kLn, bool kDeep>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
  using Cfg = GemmCfg<BN, kLn, kDeep>;
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
  uint64_t* resid_bar = tmem_empty + 3;            // [epilogue warp][staging buffer]: the TMA of the old residual rows has landed

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
    if (kLn == 2) for (int i = 0; i < Cfg::kEpiBufs * kEpiWarps; ++i) mbar_init(&resid_bar[i], 1);
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
  // everything above (barriers, TMEM, descriptor prefetch) may run while the previous kernel of the stream is finishing
  pdl_launch_dependents();
  pdl_wait();

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
    uint64_t* my_rbar = resid_bar + (warp - 4) * Cfg::kEpiBufs;     // producer: one barrier per staging buffer
    uint32_t rphase = 0;                               // bit b: parity the next wait on my_rbar[b] uses
    // consumer: up to ten {sum, sum of squares} partials of this lane's row (dims <= 1280), two per load, fetched one tile
    // ahead: the epilogue is the critical path of the GELU GEMM and must not wait for an L2 round trip per tile
    float4 stat[5];
    auto load_stats = [&](int t) {
      const int bm2 = t / p.tiles_n, b2 = bm2 / p.tiles_m, mb2 = bm2 - b2 * p.tiles_m;
      const int r = mb2 * kPairM + int(rank) * kBlockM + q * 32 + lane;
      const bool ok = r < p.rows_per_batch;
      const float4* st = reinterpret_cast<const float4*>(p.ln_stats_in + (size_t(b2) * size_t(p.rows_per_batch) + size_t(ok ? r : 0)) * size_t(p.ln_parts_in));
#pragma unroll
      for (int i = 0; i < 5; ++i) stat[i] = (ok && 2 * i < p.ln_parts_in) ? __ldg(st + i) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    };
    if (kLn == 1 && cluster_id < p.total_tiles) load_stats(cluster_id);
    for (int tile = cluster_id; tile < p.total_tiles; tile += n_clusters, ++it) {
      const int as = it & 1; const uint32_t aphase = (it >> 1) & 1;
      const int nb = tile % p.tiles_n, bm = tile / p.tiles_n;
      const int b = bm / p.tiles_m, mb = bm - b * p.tiles_m;
      const int warp_row0 = mb * kPairM + int(rank) * kBlockM + q * 32;
      const int r_in_batch = warp_row0 + lane;
      const bool masked = p.mask_period > 0 && (r_in_batch % p.mask_period) >= p.mask_valid;
      const float* pos_row = p.pos ? p.pos + size_t(r_in_batch % p.pos_period) * p.n : nullptr;
      const bool warp_live = warp_row0 < p.rows_per_batch;     // warp-uniform: any row of this warp inside the batch
      // ---- LayerNorm folding: whatever does not need the accumulator is fetched before the wait for it ----
      const bool row_ok = r_in_batch < p.rows_per_batch;
      const size_t grow = size_t(b) * size_t(p.rows_per_batch) + size_t(r_in_batch);     // row of the (rows, n) output
      float ln_rstd = 1.0f;
      const bool fetch = kLn == 2 && p.add_resid != 0;
      // old residual rows of a 32-column chunk -> staging buffer sb (TMA, zero fill past the last row); the buffer's previous
      // store must have left shared memory
      auto issue_resid = [&](int sb, int col) {
        if (lane == 0) {
          tma_store_wait_read<Cfg::kEpiBufs - 2>();      // two buffers: the previous chunk's store; three: the one before it
          mbar_arrive_expect_tx(&my_rbar[sb], kStageBufBytes);
          tma_load_3d(my_stage + sb * kStageBufBytes, &map_c, &my_rbar[sb], col, warp_row0, b);
        }
      };
      // chunks of this warp in this tile: columns nb * BN + (half + 2 j) * 32, j < nchunks; chunk j lives in staging buffer
      // (buf + j) % kEpiBufs and its old rows are requested one chunk ahead.  With three buffers the buffer that is refilled
      // belongs to the chunk before the previous one, so the request never waits for a store to drain.
      constexpr int kAhead = 1;
      int nchunks = 0;
      if (kLn == 2 && fetch && warp_live) {
        const int cols_left = p.n - nb * BN - half * 32;          // > 0: this warp's first chunk exists
        nchunks = cols_left > 0 ? min((BN / 32 - half + 1) / 2, (cols_left + 63) / 64) : 0;
#pragma unroll
        for (int j = 0; j < kAhead; ++j)
          if (j < nchunks) issue_resid((buf + j) % Cfg::kEpiBufs, nb * BN + (half + 2 * j) * 32);
      }
      int chunk_j = 0;
      float st_sum = 0.0f, st_sq = 0.0f;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      if (kLn == 1) {
        const float ssum = ((stat[0].x + stat[0].z) + (stat[1].x + stat[1].z)) + ((stat[2].x + stat[2].z) + (stat[3].x + stat[3].z)) + (stat[4].x + stat[4].z);
        const float ssq = ((stat[0].y + stat[0].w) + (stat[1].y + stat[1].w)) + ((stat[2].y + stat[2].w) + (stat[3].y + stat[3].w)) + (stat[4].y + stat[4].w);
        const float mean = ssum * p.ln_inv_dims;
        ln_rstd = rsqrtf(fmaxf(0.0f, ssq * p.ln_inv_dims - mean * mean) + 1e-6f);   // flax: var = E[x^2] - E[x]^2 >= 0, eps 1e-6
        if (tile + n_clusters < p.total_tiles) load_stats(tile + n_clusters);
      }
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
            const bool last_chunk = c + 2 >= BN / 32 || col0 + 64 >= p.n;   // warp-uniform: no further chunk for this warp
            uint32_t acc[32];
            tmem_ld_32x32b_x32(t_addr + uint32_t(c * 32), acc);
            unsigned char* sbuf = my_stage + buf * kStageBufBytes;
            float4 bias4[8];                                                // fetched ahead of the waits below (producer: the chain per chunk is the critical path)
            if (kLn == 2 && p.bias) {
#pragma unroll
              for (int j = 0; j < 8; ++j) bias4[j] = __ldg(reinterpret_cast<const float4*>(p.bias + col0) + j);
            }
            if (kLn == 2 && fetch) {
              if (Cfg::kEpiBufs >= 3 && chunk_j + kAhead < nchunks) issue_resid((buf + kAhead) % Cfg::kEpiBufs, col0 + 64 * kAhead);
              mbar_wait(&my_rbar[buf], (rphase >> buf) & 1u);             // this chunk's old rows (requested a chunk ago)
              rphase ^= 1u << buf;
            }
            tmem_ld_wait();
            if (last_chunk) release_acc();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(acc[j]);
            if (kLn == 2 && p.bias && !p.gelu && !pos_row && !masked) {
#pragma unroll
              for (int j = 0; j < 8; ++j) { v[4 * j] += bias4[j].x; v[4 * j + 1] += bias4[j].y; v[4 * j + 2] += bias4[j].z; v[4 * j + 3] += bias4[j].w; }
            } else {
              epilogue_math(v, p, col0, pos_row, masked);
            }
            if (kLn == 2) {
              if (fetch) {                                                  // x_new = x_old + value
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float4 xo = *reinterpret_cast<const float4*>(sbuf + lane * 128 + ((j ^ sw) << 4));
                  v[4 * j] += xo.x; v[4 * j + 1] += xo.y; v[4 * j + 2] += xo.z; v[4 * j + 3] += xo.w;
                }
              }
              if (p.ln_stats_out) {
                float s0 = 0.0f, s1 = 0.0f, q0 = 0.0f, q1 = 0.0f;
#pragma unroll
                for (int j = 0; j < 32; j += 2) { s0 += v[j]; s1 += v[j + 1]; q0 = fmaf(v[j], v[j], q0); q1 = fmaf(v[j + 1], v[j + 1], q1); }
                st_sum += s0 + s1;
                st_sq += q0 + q1;
              }
              if (p.xb_out && row_ok) {                                     // bf16 copy of the new rows: 64 contiguous bytes per lane
                uint4* xb = reinterpret_cast<uint4*>(p.xb_out + grow * size_t(p.xb_ld) + col0);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  xb[j] = make_uint4(pack_bf16(v[8 * j], v[8 * j + 1]), pack_bf16(v[8 * j + 2], v[8 * j + 3]),
                                     pack_bf16(v[8 * j + 4], v[8 * j + 5]), pack_bf16(v[8 * j + 6], v[8 * j + 7]));
              }
            }
            if (kLn == 2 && fetch) {
              // request the rows of chunk j + kAhead: they go to the buffer of chunk j - 1, whose store was committed a whole
              // chunk of arithmetic ago
              if (Cfg::kEpiBufs < 3 && chunk_j + kAhead < nchunks) issue_resid((buf + kAhead) % Cfg::kEpiBufs, col0 + 64 * kAhead);
              ++chunk_j;
            }
            if (!(kLn == 2 && fetch)) {                                     // (with fetch the landed load already proves the buffer free)
              if (lane == 0) tma_store_wait_read<Cfg::kEpiBufs - 1>();     // the previous user of this buffer has been read out
              __syncwarp();
            }
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
          if (kLn == 2 && p.ln_stats_out && row_ok)                         // a warp without a chunk in this tile contributes zeros
            p.ln_stats_out[grow * size_t(p.ln_parts_out) + size_t(nb * 2 + half)] = make_float2(st_sum, st_sq);
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
              epilogue_math(v, p, col0, pos_row, masked, kLn == 1, ln_rstd);
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
              epilogue_math(v, p, col0 + 32, pos_row, masked, kLn == 1, ln_rstd);
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

static bool gemm_picks_bn256(int n) { return n % 256 == 0 || n > 1024; }
int gemm_ln_parts(int n) { return 2 * ceil_div(n, gemm_picks_bn256(n) ? 256 : 128); }

template <int BN, int kLn, bool kDeep = false>
static int launch_gemm_bn(yoho_ctx* ctx, const GemmBf16Args& a, cudaStream_t s) {
  using Cfg = GemmCfg<BN, kLn, kDeep>;
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
  p.add_resid = a.add_resid; p.xb_out = a.xb_out; p.xb_ld = a.xb_ld;
  p.ln_stats_out = a.ln_stats_out; p.ln_parts_out = 2 * p.tiles_n;
  p.ln_stats_in = a.ln_stats_in; p.ln_parts_in = a.ln_parts_in; p.ln_inv_dims = a.ln_dims > 0 ? 1.0f / float(a.ln_dims) : 0.0f;

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

  YOHO_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<BN, kLn, kDeep>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
  const int max_clusters = ctx->num_sms / 2;
  const int clusters = p.total_tiles < max_clusters ? p.total_tiles : max_clusters;
  ProfScope prof(s, kProfGemm, 2.0 * double(a.batch) * double(a.rows_per_batch) * double(a.n) * double(a.flop_k > 0 ? a.flop_k : a.n_taps * a.k_per_tap));
  YOHO_CUDA_CHECK(launch_pdl(gemm_bf16_kernel<BN, kLn, kDeep>, dim3(2 * clusters), dim3(kGemmThreads), size_t(Cfg::kSmemBytes), s, map_a, map_w, map_c, p));
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
  const bool producer = a.add_resid || a.xb_out || a.ln_stats_out, consumer = a.ln_stats_in != nullptr;
  if (producer) {
    if (consumer) return fail(YOHO_ERR_INVALID, "gemm: one GEMM is a LayerNorm producer or a consumer, not both");
    if (!a.out_f32 || a.accumulate) return fail(YOHO_ERR_INVALID, "gemm: the LayerNorm producer outputs need a float32 output without TMA accumulation");
    if (a.xb_out && (a.xb_ld % 8 != 0 || reinterpret_cast<uintptr_t>(a.xb_out) % 16 != 0)) return fail(YOHO_ERR_INVALID, "gemm: xb_out must be 16-byte aligned with a row stride that is a multiple of 8");
    if (a.add_resid && a.n_taps * a.k_per_tap <= 1024)     // memory-bound: deep residual prefetch, fewer operand stages
      return gemm_picks_bn256(a.n) ? launch_gemm_bn<256, 2, true>(ctx, a, s) : launch_gemm_bn<128, 2, true>(ctx, a, s);
    return gemm_picks_bn256(a.n) ? launch_gemm_bn<256, 2>(ctx, a, s) : launch_gemm_bn<128, 2>(ctx, a, s);
  }
  if (consumer) {
    if (!a.out_bf16 || !a.bias || a.ln_dims <= 0 || a.ln_parts_in <= 0)
      return fail(YOHO_ERR_INVALID, "gemm: the LayerNorm consumer needs a bf16 output, the folded bias, ln_dims and ln_parts_in");
    return gemm_picks_bn256(a.n) ? launch_gemm_bn<256, 1>(ctx, a, s) : launch_gemm_bn<128, 1>(ctx, a, s);
  }
  return gemm_picks_bn256(a.n) ? launch_gemm_bn<256, 0>(ctx, a, s) : launch_gemm_bn<128, 0>(ctx, a, s);
}

}  // namespace yoho
