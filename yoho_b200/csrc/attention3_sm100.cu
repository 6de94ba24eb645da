// Fused self-attention forward, three query tiles per CTA (sm_100a).  Same contract as attention_sm100.cu:
//   out = softmax(q k^T / sqrt(64)) v   per (clip, head), no mask      (reference whisper.py:52-67)
//
// Why a second kernel: at d_head = 64 the bound is the SFU (16 ex2/clk/SM = 1024 clk per 128 x 128 score tile
// against 512 clk of MMA).  The two-tile kernel keeps two softmax warps per sub-partition; they run in phase
// (both wait for the same K/V tile), so a step pair costs 2 E + N clocks (E = 512 clk of exp2, N = ~450 clk of
// mbarrier wait -> tcgen05.ld -> row max -> vote -> ... -> tcgen05.st -> arrive latency during which the SFU idles).
// Here a work item is (clip, head, 384 query rows) = THREE 128-row query tiles sharing every K/V tile, i.e. three
// softmax warps per sub-partition whose N phases hide behind the other warps' exp2 blocks, and the tensor-core
// round trip is taken off the softmax warps' path altogether (below).
//
// CTA = 16 warps (512 threads, up to 128 registers each):
//   warps 0-11   softmax of tile t = warp / 4, TMEM lanes 32 (warp % 4) .. +31  (thread = one query row)
//   warp 12      TMA producer: Q tiles (double buffered across items), K_j / V_j rings of 64 keys (= 1 step)
//   warps 13-15  MMA issuer of tile t = warp - 13 (warp 13 also allocates TMEM)
// Per step g (64 keys) and tile t:   S_t = Q_t K_g^T   (M128 N64 K64, operands in shared memory)
//                                    O_t += P_t V_g    (M128 N64 K64, P read from TMEM, V MN-major)
// TMEM: S_t = 64 fp32 columns at 64 t, O_t = 64 columns at 192 + 64 t, P_t = 32 columns of packed bf16 at
// 384 + 32 t.  P has columns of its own, so S_t is handed back (s_free) as soon as the softmax warps hold the 64
// scores of step g in registers: Q K^T of step g+1 runs while they exponentiate step g and is finished when they
// get there -- one S buffer per tile is enough and the MMA issue / pipe latency never shows.  O stays in TMEM and
// is rescaled only when the running maximum grows by more than 2^8 (lazy rescaling).
#include "kernels.cuh"
#include "ptx_sm100.cuh"

namespace yoho {

using namespace ptx;

namespace att3 {

constexpr int kTiles = 3;
constexpr int kSoftmaxWarps = 4 * kTiles;
constexpr int kThreads = (kSoftmaxWarps + 1 + kTiles) * 32;   // 512
constexpr int kTileQ = 128, kStepKV = 64, kHeadDim = 64;
constexpr int kItemQ = kTiles * kTileQ;          // 384 query rows per work item
constexpr int kQBytes = kTileQ * kHeadDim * 2;   // 16 KB
constexpr int kKvBytes = kStepKV * kHeadDim * 2; // 8 KB
#ifndef YOHO_ATT3_KV_STAGES
#define YOHO_ATT3_KV_STAGES 6
#endif
constexpr int kKvStages = YOHO_ATT3_KV_STAGES;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemS = 0, kTmemO = 64 * kTiles, kTmemP = 128 * kTiles;   // S_t at 64 t, O_t at 192 + 64 t, P_t at 384 + 32 t
constexpr float kRescaleLog2 = 8.0f;

struct Smem {
  unsigned char q[2][kTiles][kQBytes];   // [item slot][query tile]       96 KB
  unsigned char k[kKvStages][kKvBytes];  //                               48 KB
  unsigned char v[kKvStages][kKvBytes];  //                               48 KB
  uint64_t q_full[2][kTiles], q_empty[2][kTiles];
  uint64_t k_full[kKvStages], k_empty[kKvStages], v_full[kKvStages], v_empty[kKvStages];
  uint64_t s_full[kTiles], p_full[kTiles];
  uint64_t s_free[kTiles], p_free[kTiles];   // S_t is in the softmax warps' registers / P_t has been consumed by its P V
  uint64_t o_last[kTiles], o_free[kTiles];   // last P V of an item has retired / O has been read out
  uint32_t tmem_slot;
};
static_assert(sizeof(Smem) + 1024 <= 232448, "attention x3: shared memory");

struct Params {
  int t_len, dims, heads;
  int q_blocks;      // ceil(t_len / 384)
  int n_items;       // batch * heads * q_blocks
  int n_steps;       // ceil(t_len / 64)
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float y;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(a), "f"(b), "f"(c));
  return y;
}

// exp2 of a pair on the FMA pipe (no MUFU): x = n + f with n = round(x), f in [-0.5, 0.5]; 2^f by a degree-3
// minimax polynomial (7.5e-5 relative, well under the bf16 rounding of P), 2^n by adding n to the exponent field.
// t = x + 1.5 * 2^23 holds n in its low mantissa bits; x is clamped to -126 first (masked keys are -inf).
__device__ __forceinline__ void exp2_poly_pair(float& lo, float& hi) {
  const uint64_t magic = pack_f32x2(12582912.0f, 12582912.0f), neg_magic = pack_f32x2(-12582912.0f, -12582912.0f);
  const uint64_t neg_one = pack_f32x2(-1.0f, -1.0f);
  const uint64_t k0 = pack_f32x2(0.9999280571937561f, 0.9999280571937561f), k1 = pack_f32x2(0.6932609677314758f, 0.6932609677314758f);
  const uint64_t k2 = pack_f32x2(0.2426111251115799f, 0.2426111251115799f), k3 = pack_f32x2(0.05517164617776871f, 0.05517164617776871f);
  const uint64_t x = pack_f32x2(fmaxf(lo, -126.0f), fmaxf(hi, -126.0f));
  const uint64_t t = add_f32x2(x, magic);
  const uint64_t f = fma_f32x2(add_f32x2(t, neg_magic), neg_one, x);
  const uint64_t q = fma_f32x2(fma_f32x2(fma_f32x2(k3, f, k2), f, k1), f, k0);
  float t_lo, t_hi, q_lo, q_hi;
  unpack_f32x2(t, t_lo, t_hi);
  unpack_f32x2(q, q_lo, q_hi);
  lo = __int_as_float(__float_as_int(q_lo) + (__float_as_int(t_lo) << 23));
  hi = __int_as_float(__float_as_int(q_hi) + (__float_as_int(t_hi) << 23));
}

#ifndef YOHO_ATT_POLY_EVERY
#define YOHO_ATT_POLY_EVERY 4
#endif
// every kPolyEvery-th pair of a row takes the polynomial instead of the SFU (0 = none): the SFU is the bound of this
// kernel (ncu: MUFU queue full for most of the softmax warps' stall samples) while the FMA pipe is ~13 % busy
constexpr int kPolyEvery = YOHO_ATT_POLY_EVERY;
#ifndef YOHO_ATT_POLY_PHASE
#define YOHO_ATT_POLY_PHASE (YOHO_ATT_POLY_EVERY - 1)
#endif
constexpr int kPolyPhase = YOHO_ATT_POLY_PHASE;   // which pair of each group of kPolyEvery
__device__ __forceinline__ constexpr bool poly_pair(int i) { return kPolyEvery > 0 && ((i >> 1) % (kPolyEvery > 0 ? kPolyEvery : 1)) == kPolyPhase; }

// exp2(s c - mc) of 32 scores -> 16 packed bf16x2 words; returns the sum of the 32 (unrounded) exponentials.
// Scale / shift and the sum run on packed fp32 pairs (FFMA2 / FADD2); the exponentials are issued one block of 8
// ahead of their consumers so that the in-order warp does not wait on a MUFU result with the SFU idle.
__device__ __forceinline__ float exp_pack32(const uint32_t (&s)[32], uint64_t c2, uint64_t nmc2, uint32_t* packed) {
  float x[32];
#pragma unroll
  for (int i = 0; i < 32; i += 2)
    unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(s[i]), __uint_as_float(s[i + 1])), c2, nmc2), x[i], x[i + 1]);
  constexpr int kBlk = 8;
#pragma unroll
  for (int i = 0; i < kBlk; ++i) if (!poly_pair(i)) x[i] = ex2_approx(x[i]);
  uint64_t l01 = 0, l23 = 0;
#pragma unroll
  for (int blk = 0; blk < 32 / kBlk; ++blk) {
    if (blk + 1 < 32 / kBlk) {
#pragma unroll
      for (int i = 0; i < kBlk; ++i) if (!poly_pair((blk + 1) * kBlk + i)) x[(blk + 1) * kBlk + i] = ex2_approx(x[(blk + 1) * kBlk + i]);
    }
#pragma unroll
    for (int i = blk * kBlk; i < (blk + 1) * kBlk; i += 2) {
      if (poly_pair(i)) exp2_poly_pair(x[i], x[i + 1]);
      if (blk & 1) l23 = add_f32x2(l23, pack_f32x2(x[i], x[i + 1]));
      else l01 = add_f32x2(l01, pack_f32x2(x[i], x[i + 1]));
      __nv_bfloat162 u = __floats2bfloat162_rn(x[i], x[i + 1]);
      packed[i >> 1] = *reinterpret_cast<uint32_t*>(&u);
    }
  }
  float l0, l1;
  unpack_f32x2(add_f32x2(l01, l23), l0, l1);
  return l0 + l1;
}

__device__ __forceinline__ float max32(const uint32_t (&s)[32]) {
  float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
  for (int i = 0; i < 32; i += 8) {
    m0 = fmax3(m0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
    m1 = fmax3(m1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
    m2 = fmax3(m2, __uint_as_float(s[i + 4]), __uint_as_float(s[i + 5]));
    m3 = fmax3(m3, __uint_as_float(s[i + 6]), __uint_as_float(s[i + 7]));
  }
  return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}

__global__ void __launch_bounds__(kThreads, 1)
attention3_bf16_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, const Params p) {
  extern __shared__ __align__(1024) unsigned char att3_smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(att3_smem_raw);
  // warp index through a shuffle: warp-uniform for ptxas (role branches and the MMA warps' loops stay on the uniform datapath)
  const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    if (smem_u32(att3_smem_raw) & 1023u) __trap();  // operand tiles need 1024-byte alignment (128B swizzle)
    for (int s = 0; s < 2; ++s)
      for (int t = 0; t < kTiles; ++t) { mbar_init(&sm.q_full[s][t], 1); mbar_init(&sm.q_empty[s][t], 1); }
    for (int i = 0; i < kKvStages; ++i) {
      mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], kTiles);   // released by the four MMA warps
      mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], kTiles);
    }
    for (int t = 0; t < kTiles; ++t) {
      mbar_init(&sm.s_full[t], 1); mbar_init(&sm.p_full[t], 4);
      mbar_init(&sm.s_free[t], 4); mbar_init(&sm.p_free[t], 1);
      mbar_init(&sm.o_last[t], 1); mbar_init(&sm.o_free[t], 4);
    }
    fence_mbar_init();
  }
  if (warp == kSoftmaxWarps && lane == 0) prefetch_tmap(&map_qkv);
  if (warp == kSoftmaxWarps + 1) {
    tmem_alloc<kTmemCols>(&sm.tmem_slot);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, sm.tmem_slot, 0);
  pdl_launch_dependents();
  pdl_wait();                      // the prologue above overlaps the tail of the q | k | v GEMM
  const int n_steps = p.n_steps;
  const int stride = int(gridDim.x);
  const int n_my = int(blockIdx.x) < p.n_items ? (p.n_items - int(blockIdx.x) + stride - 1) / stride : 0;   // items of this CTA

  if (warp == kSoftmaxWarps) {
    // ===================== TMA producer (one lane) =====================
    if (lane == 0) {
      int st = 0; uint32_t ph = 0;   // K/V ring position
      for (int it = 0; it < n_my; ++it) {
        const int item = int(blockIdx.x) + it * stride;
        const int qb = item % p.q_blocks, ch = item / p.q_blocks;
        const int head = ch % p.heads, clip = ch / p.heads;
        const int slot = it & 1; const uint32_t qph = (it >> 1) & 1;
        for (int t = 0; t < kTiles; ++t) {
          mbar_wait_idle(&sm.q_empty[slot][t], qph ^ 1);
          mbar_arrive_expect_tx(&sm.q_full[slot][t], kQBytes);
          // the tensor map's box is 64 rows (one K/V step): a query tile is two boxes
          tma_load_3d(sm.q[slot][t], &map_qkv, &sm.q_full[slot][t], head * kHeadDim, qb * kItemQ + t * kTileQ, clip);
          tma_load_3d(sm.q[slot][t] + kKvBytes, &map_qkv, &sm.q_full[slot][t], head * kHeadDim, qb * kItemQ + t * kTileQ + 64, clip);
        }
        for (int j = 0; j < n_steps; ++j) {
          mbar_wait_idle(&sm.k_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.k_full[st], kKvBytes);
          tma_load_3d(sm.k[st], &map_qkv, &sm.k_full[st], p.dims + head * kHeadDim, j * kStepKV, clip);
          mbar_wait_idle(&sm.v_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.v_full[st], kKvBytes);
          tma_load_3d(sm.v[st], &map_qkv, &sm.v_full[st], 2 * p.dims + head * kHeadDim, j * kStepKV, clip);
          if (++st == kKvStages) { st = 0; ph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp > kSoftmaxWarps) {
    // ===================== MMA issuers: warp 13 + t drives query tile t =====================
    // The warp stays converged and one elected lane issues (umma_*_w); descriptors are linear in (slot, stage, k step).
    const int t = warp - (kSoftmaxWarps + 1);
    constexpr uint32_t idesc_qk = make_idesc_bf16(kTileQ, kStepKV, 0, 0);   // A = Q (K-major), B = K (K-major)
    constexpr uint32_t idesc_pv = make_idesc_bf16(kTileQ, kHeadDim, 0, 1);  // A = P (TMEM), B = V (MN-major)
    constexpr uint64_t kQSlotDesc = (kTiles * kQBytes) >> 4, kKvStageDesc = kKvBytes >> 4;
    const uint64_t qdesc0 = make_desc_sw128(smem_u32(sm.q[0][t]), 16, 1024);       // slot s: + s kQSlotDesc
    const uint64_t kdesc0 = make_desc_sw128(smem_u32(sm.k[0]), 16, 1024);          // stage st: + st kKvStageDesc
    const uint64_t vdesc0 = make_desc_sw128(smem_u32(sm.v[0]), 1024, 1024);
    const uint32_t s_tmem = tmem_base + kTmemS + uint32_t(t * 64), o_tmem = tmem_base + kTmemO + uint32_t(t * 64);
    const uint32_t p_tmem = tmem_base + kTmemP + uint32_t(t * 32);
    const int g_total = n_my * n_steps;
    // Q K^T cursor (one step ahead of P V)
    int q_it = 0, q_j = 0, q_st = 0; uint32_t q_ph = 0;
    auto qk_operands = [&]() {           // wait for the operands of the next Q K^T (off the critical path)
      if (q_j == 0) mbar_wait(&sm.q_full[q_it & 1][t], (q_it >> 1) & 1);
      mbar_wait(&sm.k_full[q_st], q_ph);
    };
    auto qk_mma = [&]() {
      const uint64_t adesc = qdesc0 + uint64_t(q_it & 1) * kQSlotDesc;
      const uint64_t bdesc = kdesc0 + uint64_t(q_st) * kKvStageDesc;
#pragma unroll
      for (int k = 0; k < kHeadDim / 16; ++k) umma_bf16_ss_w(s_tmem, adesc + uint64_t(2 * k), bdesc + uint64_t(2 * k), idesc_qk, k != 0);
      umma_commit_w(&sm.s_full[t]);
    };
    auto qk_release = [&]() {
      const bool last = q_j == n_steps - 1;
      if (last) umma_commit_w(&sm.q_empty[q_it & 1][t]);
      umma_commit_w(&sm.k_empty[q_st]);
      if (++q_st == kKvStages) { q_st = 0; q_ph ^= 1; }
      if (last) { q_j = 0; ++q_it; } else { ++q_j; }
    };
    if (g_total > 0) { qk_operands(); tc_fence_after(); qk_mma(); qk_release(); }
    int p_it = 0, p_j = 0, p_st = 0; uint32_t p_ph = 0;
    for (int g = 0; g < g_total; ++g) {
      const bool last = p_j == n_steps - 1;
      // Q K^T one step ahead: S_t is free as soon as the softmax warps have read the scores of step g
      if (g + 1 < g_total) {
        qk_operands();
        mbar_wait(&sm.s_free[t], g & 1);
        tc_fence_after();
        qk_mma();
        qk_release();
      }
      mbar_wait(&sm.v_full[p_st], p_ph);
      if (p_j == 0 && p_it > 0) mbar_wait(&sm.o_free[t], (p_it - 1) & 1);    // previous item's O has been read out
      const uint64_t bdesc = vdesc0 + uint64_t(p_st) * kKvStageDesc;
      mbar_wait_idle(&sm.p_full[t], g & 1);                                       // P in TMEM, O rescaled if needed
      tc_fence_after();
      // A: 16 keys = 8 packed columns.  B: V rows are keys (128 B each): 16 keys = 2 groups of 8 rows, 1024 B apart
#pragma unroll
      for (int ks = 0; ks < kStepKV / 16; ++ks)
        umma_bf16_ts_w(o_tmem, p_tmem + uint32_t(ks * 8), bdesc + uint64_t(ks * (2048 >> 4)), idesc_pv, (p_j | ks) != 0);
      umma_commit_w(&sm.p_free[t]);
      if (last) umma_commit_w(&sm.o_last[t]);
      umma_commit_w(&sm.v_empty[p_st]);
      if (++p_st == kKvStages) { p_st = 0; p_ph ^= 1; }
      if (last) { p_j = 0; ++p_it; } else { ++p_j; }
    }
    __syncwarp();
  } else {
    // ===================== softmax + output: warp w -> tile w / 4, TMEM lanes 32 (w % 4) .. =====================
    const int t = warp >> 2;
    const int row = (warp & 3) * 32 + lane;                   // query row in the tile == TMEM lane
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    const uint32_t s_addr = lane_base + kTmemS + uint32_t(t * 64);
    const uint32_t o_addr = lane_base + kTmemO + uint32_t(t * 64);
    const uint32_t p_addr = lane_base + kTmemP + uint32_t(t * 32);
    const float c = 0.125f * 1.4426950408889634f;            // d_head^-0.5 * log2(e)   (whisper.py:54-57)
    const uint64_t c2 = pack_f32x2(c, c);
    const int kv_tail = p.t_len - (n_steps - 1) * kStepKV;    // valid keys of the last step
    int g = 0;
    for (int it = 0; it < n_my; ++it) {
      const int item = int(blockIdx.x) + it * stride;
      const int qb = item % p.q_blocks, ch = item / p.q_blocks;
      const int head = ch % p.heads, clip = ch / p.heads;
      float m_ref = -INFINITY, l_run = 0.0f;
      // Warps whose 32 query rows all lie past the end of the clip (1500 = 3 * 384 + 348: the last warp of the last
      // item's third tile) only keep the barrier protocol going: their P / O rows are never stored, and tcgen05.mma
      // rows are independent, so whatever stale bits sit in their TMEM lanes cannot reach a valid row.
      const bool warp_dead = qb * kItemQ + t * kTileQ + (warp & 3) * 32 >= p.t_len;     // warp-uniform
      if (warp_dead) {
        for (int j = 0; j < n_steps; ++j, ++g) {
          mbar_wait_idle(&sm.s_full[t], g & 1);
          if (lane == 0) mbar_arrive(&sm.s_free[t]);
          if (g > 0) mbar_wait_idle(&sm.p_free[t], (g - 1) & 1);
          if (lane == 0) mbar_arrive(&sm.p_full[t]);
          __syncwarp();
        }
        mbar_wait_idle(&sm.o_last[t], it & 1);
        if (lane == 0) mbar_arrive(&sm.o_free[t]);
        __syncwarp();
        continue;
      }
      for (int j = 0; j < n_steps; ++j, ++g) {
        mbar_wait_idle(&sm.s_full[t], g & 1);
        tc_fence_after();
        uint32_t a[32], b[32];
        tmem_ld_32x32b_x32(s_addr, a);
        tmem_ld_32x32b_x32(s_addr + 32, b);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.s_free[t]);            // the scores are in registers: S_t may take Q K^T of step g + 1
        if (j == n_steps - 1 && kv_tail < kStepKV) {          // CTA-uniform
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (i >= kv_tail) a[i] = 0xff800000u;               // -inf: exp2 gives exactly 0
            if (32 + i >= kv_tail) b[i] = 0xff800000u;
          }
        }
        const float m_tile = fmaxf(max32(a), max32(b));
        if (j == 0) {
          m_ref = m_tile;
        } else if (__any_sync(0xffffffffu, (m_tile - m_ref) * c > kRescaleLog2)) {
          // rare: the running maximum grew by more than 2^8 somewhere in this warp -> rescale O and l
          const float m_new = fmaxf(m_ref, m_tile);
          const float alpha = ex2_approx((m_ref - m_new) * c);
          m_ref = m_new;
          l_run *= alpha;
          mbar_wait(&sm.p_free[t], (g - 1) & 1);              // P V of the previous step has retired: O is quiescent
          tc_fence_after();
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            uint32_t o[16];
            tmem_ld_32x32b_x16(o_addr + h * 16, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x16(o_addr + h * 16, o);
          }
          tmem_st_wait();
        }
        const uint64_t nmc2 = pack_f32x2(-m_ref * c, -m_ref * c);
        uint32_t packed[32];
        l_run += exp_pack32(a, c2, nmc2, packed);
        if (j == n_steps - 1 && kv_tail <= 32) {              // CTA-uniform: keys 32-63 of a ragged last step are all masked
#pragma unroll
          for (int i = 16; i < 32; ++i) packed[i] = 0u;
        } else {
          l_run += exp_pack32(b, c2, nmc2, packed + 16);
        }
        if (g > 0) {                                          // P_t feeds P V of step g - 1 until that MMA retires (long done)
          mbar_wait(&sm.p_free[t], (g - 1) & 1);
          tc_fence_after();
        }
        tmem_st_32x32b_x32(p_addr, packed);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.p_full[t]);
      }
      // last P V of the item: pull O into registers (two halves), hand O_t back to the MMA warp, normalise and store
      mbar_wait_idle(&sm.o_last[t], it & 1);
      tc_fence_after();
      const float inv_l = 1.0f / l_run;
      const int t_row = qb * kItemQ + t * kTileQ + row;
      __nv_bfloat16* dst = out + (size_t(clip) * p.t_len + (t_row < p.t_len ? t_row : 0)) * p.dims + head * kHeadDim;
      uint32_t w[32];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(o_addr + h * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          __nv_bfloat162 hv = __floats2bfloat162_rn(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l);
          w[h * 16 + (i >> 1)] = *reinterpret_cast<uint32_t*>(&hv);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.o_free[t]);
      if (t_row < p.t_len) {
#pragma unroll
        for (int q4 = 0; q4 < 8; ++q4)
          *reinterpret_cast<uint4*>(dst + q4 * 8) = make_uint4(w[4 * q4], w[4 * q4 + 1], w[4 * q4 + 2], w[4 * q4 + 3]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kSoftmaxWarps + 1) tmem_dealloc<kTmemCols>(tmem_base);
}

}  // namespace att3

int launch_attention3_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s) {
  using namespace att3;
  if (!qkv || !out) return fail(YOHO_ERR_INVALID, "attention: null operand");
  if (batch <= 0 || t <= 0 || heads <= 0) return YOHO_OK;
  const int dims = heads * kHeadDim;
  Params p{};
  p.t_len = t; p.dims = dims; p.heads = heads;
  p.q_blocks = ceil_div(t, kItemQ);
  p.n_steps = ceil_div(t, kStepKV);
  const long long items = (long long)batch * heads * p.q_blocks;
  if (items > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "attention: %lld work items", items);
  p.n_items = int(items);
  CUtensorMap map;
  int r = make_tmap_bf16_3d(ctx->encode_tiled, &map, qkv, uint64_t(3 * dims), uint64_t(t), uint64_t(batch), uint64_t(3 * dims),
                            uint64_t(3 * dims) * uint64_t(t), kHeadDim, kStepKV, 1);
  if (r) return fail(YOHO_ERR_CUDA, "attention: cuTensorMapEncodeTiled failed with CUresult %d", r);
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(attention3_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(Smem))));
  const int grid = p.n_items < ctx->num_sms ? p.n_items : ctx->num_sms;
  ProfScope prof(s, kProfAttention, 4.0 * double(batch) * double(t) * double(t) * double(dims));
  YOHO_CUDA_CHECK(launch_pdl(attention3_bf16_kernel, dim3(grid), dim3(kThreads), sizeof(Smem), s, map, out, p));
  YOHO_LAUNCH_CHECK("attention3_bf16_kernel");
  return YOHO_OK;
}

}  // namespace yoho
