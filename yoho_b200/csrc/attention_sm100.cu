// Fused self-attention forward for the encoder on tcgen05 tensor cores (sm_100a):
//   out = softmax(q k^T / sqrt(64)) v   per (clip, head), no mask      (reference whisper.py:52-67)
// The reference materialises the (B, H, 1500, 1500) score tensor; here scores never leave the SM.
// This is the two-query-tile kernel.  launch_attention_bf16 (bottom of the file) sends every length whose 384-row
// padding is acceptable -- 1500, 513, ... -- to the faster three-tile kernel in attention3_sm100.cu and keeps this one
// for the rest (short sequences, lengths just above a multiple of 384).
//
// Persistent kernel, one CTA per SM, 11 warps.  A work item is (clip, head, 256 query rows) = two
// 128-row query tiles A and B that share every K/V tile; a step is 64 keys of an item:
//   warps 0-3  softmax of tile A   (thread i owns query row i = TMEM lane i: row max / sum need no shuffles)
//   warps 4-7  softmax of tile B
//   warp 8     TMA producer: Q tiles (double buffered across items), K_j / V_j rings of 128 keys (= 2 steps)
//   warp 9/10  MMA issuer of tile A / tile B (warp 9 also allocates TMEM)
// Per step g and query tile t:  S_t[g&1] = Q_t K_g^T (M128 N64 K64, operands in shared memory) -- issued 2 steps ahead
//                               O_t += P_t V_g        (M128 N64 K64, P read from TMEM, V MN-major from its TMA tile)
// TMEM: S_t[b] = 64 fp32 columns at 128 t + 64 b, O_t = 64 columns at 256 + 64 t.  S is double buffered per
// query tile so that Q K^T of step g+1 is already in TMEM when the softmax of step g retires: the softmax warps
// never wait for the tensor core, only for the SFU (16 ex2/clk/SM is the bound at d_head = 64).  A softmax
// thread reads its 64 scores once, takes the row maximum, writes P as packed bf16 over the first 32 columns of
// the same S buffer (tcgen05.st) and keeps the row sum in registers.  O stays in TMEM across steps: it is
// only rescaled when the running maximum grows by more than 2^8 (lazy rescaling; P is then bounded by 256), so
// the common case has no correction.  Keys past the clip end are TMA zero-filled and masked to -inf.
#include <cstdlib>
#include "kernels.cuh"
#include "ptx_sm100.cuh"

namespace yoho {

using namespace ptx;

constexpr int kAttThreads = 352;
constexpr int kTileQ = 128, kTileKV = 128, kStepKV = 64, kHeadDim = 64;
constexpr int kItemQ = 2 * kTileQ;        // query rows per work item
constexpr int kTileBytes = 128 * 64 * 2;  // one (128 rows x 64 bf16) swizzled tile = 16 KB
constexpr int kKvStages = 4;
constexpr uint32_t kAttTmemCols = 512;
constexpr uint32_t kTmemS = 0, kTmemO = 256;   // S_t[b] at kTmemS + 128 t + 64 b, O_t at kTmemO + 64 t
constexpr float kRescaleLog2 = 8.0f;

struct AttSmem {
  unsigned char q[2][2][kTileBytes];   // [item slot][query tile]
  unsigned char k[kKvStages][kTileBytes];
  unsigned char v[kKvStages][kTileBytes];
  uint64_t q_full[2][2], q_empty[2][2];
  uint64_t k_full[kKvStages], k_empty[kKvStages], v_full[kKvStages], v_empty[kKvStages];
  uint64_t s_full[2][2], p_full[2][2];   // [query tile][S buffer]
  uint64_t o_full[2], o_last[2], o_free[2];   // [query tile]: every P V / last P V of an item / O read out
  uint32_t tmem_slot;
};

struct AttParams {
  int t_len, dims, heads;
  int q_blocks;      // ceil(t_len / 256)
  int n_items;       // batch * heads * q_blocks
  int n_kv;          // ceil(t_len / 128): K/V tiles per item
  int n_steps;       // ceil(t_len / 64): softmax steps per item
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

#ifdef YOHO_ATT_TIMING
__device__ long long g_att_phase[16];   // [0, 8) softmax warp 0, [8, 16) MMA warp of tile A; CTA 0
#define ATT_T(i) do { long long now_ = clock64(); tacc[i] += now_ - tprev; tprev = now_; } while (0)
#else
#define ATT_T(i) do {} while (0)
#endif

__global__ void __launch_bounds__(kAttThreads, 1)
attention_bf16_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, const AttParams p) {
  extern __shared__ __align__(1024) unsigned char att_smem_raw[];
  AttSmem& sm = *reinterpret_cast<AttSmem*>(att_smem_raw);
  const int warp = __shfl_sync(0xffffffffu, int(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    if (smem_u32(att_smem_raw) & 1023u) __trap();  // operand tiles need 1024-byte alignment (128B swizzle)
    for (int s = 0; s < 2; ++s)
      for (int t = 0; t < 2; ++t) { mbar_init(&sm.q_full[s][t], 1); mbar_init(&sm.q_empty[s][t], 1); }
    for (int i = 0; i < kKvStages; ++i) {
      mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 2);   // released by both MMA warps
      mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], 2);
    }
    for (int t = 0; t < 2; ++t) {
      for (int b = 0; b < 2; ++b) { mbar_init(&sm.s_full[t][b], 1); mbar_init(&sm.p_full[t][b], 4); }
      mbar_init(&sm.o_full[t], 1); mbar_init(&sm.o_last[t], 1); mbar_init(&sm.o_free[t], 4);
    }
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) prefetch_tmap(&map_qkv);
  if (warp == 9) {
    tmem_alloc<kAttTmemCols>(&sm.tmem_slot);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, sm.tmem_slot, 0);
  pdl_launch_dependents();
  pdl_wait();
  const int n_kv = p.n_kv, n_steps = p.n_steps;
  const int stride = int(gridDim.x);
  const int n_my = int(blockIdx.x) < p.n_items ? (p.n_items - int(blockIdx.x) + stride - 1) / stride : 0;   // items of this CTA

  if (warp == 8) {
    // ===================== TMA producer (one lane) =====================
    if (lane == 0) {
      int kc = 0;   // K/V tiles issued so far (ring position)
      for (int it = 0; it < n_my; ++it) {
        const int item = int(blockIdx.x) + it * stride;
        const int qb = item % p.q_blocks, ch = item / p.q_blocks;
        const int head = ch % p.heads, clip = ch / p.heads;
        const int slot = it & 1; const uint32_t qph = (it >> 1) & 1;
        for (int t = 0; t < 2; ++t) {
          mbar_wait(&sm.q_empty[slot][t], qph ^ 1);
          mbar_arrive_expect_tx(&sm.q_full[slot][t], kTileBytes);
          tma_load_3d(sm.q[slot][t], &map_qkv, &sm.q_full[slot][t], head * kHeadDim, qb * kItemQ + t * kTileQ, clip);
        }
        for (int j = 0; j < n_kv; ++j, ++kc) {
          const int st = kc % kKvStages; const uint32_t ph = (kc / kKvStages) & 1;
          mbar_wait(&sm.k_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.k_full[st], kTileBytes);
          tma_load_3d(sm.k[st], &map_qkv, &sm.k_full[st], p.dims + head * kHeadDim, j * kTileKV, clip);
          mbar_wait(&sm.v_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&sm.v_full[st], kTileBytes);
          tma_load_3d(sm.v[st], &map_qkv, &sm.v_full[st], 2 * p.dims + head * kHeadDim, j * kTileKV, clip);
        }
      }
    }
    __syncwarp();
  } else if (warp >= 9) {
    // ===================== MMA issuers: warp 9 = query tile A, warp 10 = query tile B =====================
    // The whole warp stays converged and one elected lane issues (umma_*_w).  Every MMA here is small (M128 N64 K16
    // = 32 tensor-pipe cycles), so the issuing instruction stream is kept short: shared-memory descriptors are
    // linear in (ring stage, half, k step) and advanced by constant adds, ring positions are carried, not divided.
    const int t = warp - 9;
    constexpr uint32_t idesc_qk = make_idesc_bf16(kTileQ, kStepKV, 0, 0);   // A = Q (K-major), B = K (K-major)
    constexpr uint32_t idesc_pv = make_idesc_bf16(kTileQ, kHeadDim, 0, 1);  // A = P (TMEM), B = V (MN-major)
    constexpr uint64_t kStageDesc = kTileBytes >> 4, kHalfDesc = (kStepKV * 128) >> 4;
    const uint64_t qdesc0 = make_desc_sw128(smem_u32(sm.q[0][t]), 16, 1024);       // slot s: + 2 s kStageDesc
    const uint64_t kdesc0 = make_desc_sw128(smem_u32(sm.k[0]), 16, 1024);          // stage st: + st kStageDesc
    const uint64_t vdesc0 = make_desc_sw128(smem_u32(sm.v[0]), 1024, 1024);
    const uint32_t s_tmem = tmem_base + kTmemS + uint32_t(t * 128), o_tmem = tmem_base + kTmemO + uint32_t(t * 64);
    const int g_total = n_my * n_steps;
    // Q K^T cursor (runs two steps ahead of P V)
    int q_it = 0, q_j = 0, q_st = 0; uint32_t q_ph = 0;
    auto issue_qk = [&](int g) {
      const int sb = g & 1, half = q_j & 1;
      const bool last = q_j == n_steps - 1;
      if (q_j == 0) mbar_wait(&sm.q_full[q_it & 1][t], (q_it >> 1) & 1);
      mbar_wait(&sm.k_full[q_st], q_ph);
      tc_fence_after();
      const uint64_t adesc = qdesc0 + uint64_t(q_it & 1) * (2 * kStageDesc);
      const uint64_t bdesc = kdesc0 + uint64_t(q_st) * kStageDesc + uint64_t(half) * kHalfDesc;
      const uint32_t d = s_tmem + uint32_t(sb * 64);
#pragma unroll
      for (int k = 0; k < kHeadDim / 16; ++k) umma_bf16_ss_w(d, adesc + uint64_t(2 * k), bdesc + uint64_t(2 * k), idesc_qk, k != 0);
      umma_commit_w(&sm.s_full[t][sb]);
      if (last) umma_commit_w(&sm.q_empty[q_it & 1][t]);
      if (half == 1 || last) {
        umma_commit_w(&sm.k_empty[q_st]);
        if (++q_st == kKvStages) { q_st = 0; q_ph ^= 1; }
      }
      if (last) { q_j = 0; ++q_it; } else { ++q_j; }
    };
    for (int g = 0; g < 2 && g < g_total; ++g) issue_qk(g);
    int p_it = 0, p_j = 0, p_st = 0; uint32_t p_ph = 0;
#ifdef YOHO_ATT_TIMING
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#endif
    for (int g = 0; g < g_total; ++g) {
      const int sb = g & 1, half = p_j & 1;
      const bool last = p_j == n_steps - 1;
      ATT_T(0);
      mbar_wait(&sm.p_full[t][sb], (g >> 1) & 1);                        // P in TMEM, S consumed, O rescaled if needed
      ATT_T(1);
      if (p_j == 0 && p_it > 0) mbar_wait(&sm.o_free[t], (p_it - 1) & 1);  // previous item's O has been read out
      mbar_wait(&sm.v_full[p_st], p_ph);
      tc_fence_after();
      ATT_T(2);
      // A: 16 keys = 8 packed columns.  B: V rows are keys (128 B each): 16 keys = 2 groups of 8 rows, 1024 B apart
      const uint64_t bdesc = vdesc0 + uint64_t(p_st) * kStageDesc + uint64_t(half) * kHalfDesc;
      const uint32_t p_addr = s_tmem + uint32_t(sb * 64);
#pragma unroll
      for (int ks = 0; ks < kStepKV / 16; ++ks)
        umma_bf16_ts_w(o_tmem, p_addr + uint32_t(ks * 8), bdesc + uint64_t(ks * (2048 >> 4)), idesc_pv, (p_j | ks) != 0);
      umma_commit_w(&sm.o_full[t]);
      if (last) umma_commit_w(&sm.o_last[t]);
      if (half == 1 || last) {
        umma_commit_w(&sm.v_empty[p_st]);
        if (++p_st == kKvStages) { p_st = 0; p_ph ^= 1; }
      }
      if (last) { p_j = 0; ++p_it; } else { ++p_j; }
      ATT_T(3);
#if defined(YOHO_ATT_TIMING) && YOHO_ATT_TIMING >= 2
      mbar_wait(&sm.o_full[t], g & 1);      // measurement only: P V chain latency (serialises the MMA warp)
      ATT_T(4);
#endif
      if (g + 2 < g_total) issue_qk(g + 2);
      ATT_T(5);
#if defined(YOHO_ATT_TIMING) && YOHO_ATT_TIMING >= 2
      if (g + 2 < g_total) mbar_wait(&sm.s_full[t][sb], ((g + 2) >> 1) & 1);      // measurement only: Q K^T chain latency
      ATT_T(6);
#endif
    }
#ifdef YOHO_ATT_TIMING
    if (blockIdx.x == 0 && t == 0 && lane == 0)
      for (int i = 0; i < 8; ++i) g_att_phase[8 + i] = tacc[i];
#endif
    __syncwarp();
  } else {
    // ===================== softmax + output: warps 0-3 tile A, warps 4-7 tile B =====================
    const int t = warp >> 2;
    const int row = (warp & 3) * 32 + lane;                   // query row in the tile == TMEM lane
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    const uint32_t s_base = lane_base + kTmemS + uint32_t(t * 128);
    const uint32_t o_addr = lane_base + kTmemO + uint32_t(t * 64);
    const float c = 0.125f * 1.4426950408889634f;            // d_head^-0.5 * log2(e)   (whisper.py:54-57)
    const int kv_tail = p.t_len - (n_steps - 1) * kStepKV;    // valid keys of the last step
    int g = 0;
#ifdef YOHO_ATT_TIMING
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#endif
    for (int it = 0; it < n_my; ++it) {
      const int item = int(blockIdx.x) + it * stride;
      const int qb = item % p.q_blocks, ch = item / p.q_blocks;
      const int head = ch % p.heads, clip = ch / p.heads;
      float m_ref = -INFINITY, l_run = 0.0f;
      for (int j = 0; j < n_steps; ++j, ++g) {
        const int sb = g & 1;
        const uint32_t s_addr = s_base + uint32_t(sb * 64);
        ATT_T(0);
        mbar_wait(&sm.s_full[t][sb], (g >> 1) & 1);
        tc_fence_after();
        ATT_T(1);
        uint32_t a[32], b[32];
        tmem_ld_32x32b_x32(s_addr, a);
        tmem_ld_32x32b_x32(s_addr + 32, b);
        tmem_ld_wait();
        ATT_T(2);
        if (j == n_steps - 1 && kv_tail < kStepKV) {          // CTA-uniform
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (i >= kv_tail) a[i] = 0xff800000u;               // -inf: exp2 gives exactly 0
            if (32 + i >= kv_tail) b[i] = 0xff800000u;
          }
        }
        float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          m0 = fmax3(m0, __uint_as_float(a[i]), __uint_as_float(a[i + 1]));
          m1 = fmax3(m1, __uint_as_float(a[i + 2]), __uint_as_float(a[i + 3]));
          m2 = fmax3(m2, __uint_as_float(b[i]), __uint_as_float(b[i + 1]));
          m3 = fmax3(m3, __uint_as_float(b[i + 2]), __uint_as_float(b[i + 3]));
        }
        const float m_tile = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
        if (j == 0) {
          m_ref = m_tile;
        } else if (__any_sync(0xffffffffu, (m_tile - m_ref) * c > kRescaleLog2)) {
          // rare: the running maximum grew by more than 2^8 somewhere in this warp -> rescale O and l
          const float m_new = fmaxf(m_ref, m_tile);
          const float alpha = ex2_approx((m_ref - m_new) * c);
          m_ref = m_new;
          l_run *= alpha;
          mbar_wait(&sm.o_full[t], (g - 1) & 1);                  // P V of the previous step has landed in O
          tc_fence_after();
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t o[32];
            tmem_ld_32x32b_x32(o_addr + h * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x32(o_addr + h * 32, o);
          }
          tmem_st_wait();
        }
        ATT_T(3);
        // scale / shift and the row sum run on packed fp32 pairs (FFMA2 / FADD2: half the issue slots)
        const uint64_t c2 = pack_f32x2(c, c), nmc2 = pack_f32x2(-m_ref * c, -m_ref * c);
        uint64_t l01 = 0, l23 = 0;
        uint32_t packed[32];
        // x[i] = key i (a) for i < 32, key i (b) beyond; exponentials are issued one block of kExpBlock ahead of their
        // consumers so that the in-order warp never waits on a MUFU result with the SFU idle
        float x[64];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(a[i]), __uint_as_float(a[i + 1])), c2, nmc2), x[i], x[i + 1]);
          unpack_f32x2(fma_f32x2(pack_f32x2(__uint_as_float(b[i]), __uint_as_float(b[i + 1])), c2, nmc2), x[32 + i], x[33 + i]);
        }
        constexpr int kExpBlock = 8;
#pragma unroll
        for (int i = 0; i < kExpBlock; ++i) x[i] = ex2_approx(x[i]);
#pragma unroll
        for (int blk = 0; blk < 64 / kExpBlock; ++blk) {
          if (blk + 1 < 64 / kExpBlock) {
#pragma unroll
            for (int i = 0; i < kExpBlock; ++i) x[(blk + 1) * kExpBlock + i] = ex2_approx(x[(blk + 1) * kExpBlock + i]);
          }
#pragma unroll
          for (int i = blk * kExpBlock; i < (blk + 1) * kExpBlock; i += 2) {
            if (blk & 1) l23 = add_f32x2(l23, pack_f32x2(x[i], x[i + 1]));
            else l01 = add_f32x2(l01, pack_f32x2(x[i], x[i + 1]));
            __nv_bfloat162 u = __floats2bfloat162_rn(x[i], x[i + 1]);
            packed[i >> 1] = *reinterpret_cast<uint32_t*>(&u);          // keys i, i+1
          }
        }
        {
          float l0, l1;
          unpack_f32x2(add_f32x2(l01, l23), l0, l1);
          l_run += l0 + l1;
        }
        ATT_T(4);
        tmem_st_32x32b_x32(s_addr, packed);      // P over the first 32 (already consumed) columns of this S buffer
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.p_full[t][sb]);
        ATT_T(5);
      }
      // last P V of the item: pull O into registers, hand O_t back to the MMA warp, then normalise and store.
      // (o_full cannot be used here: the softmax may be two P V commits ahead, and a parity wait only tells the
      // current phase from the previous one.  In the rescale path above s_full(g) implies P V(g-2) has retired.)
      mbar_wait(&sm.o_last[t], it & 1);
      tc_fence_after();
      ATT_T(6);
      uint32_t o0[32], o1[32];
      tmem_ld_32x32b_x32(o_addr, o0);
      tmem_ld_32x32b_x32(o_addr + 32, o1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.o_free[t]);
      const float inv_l = 1.0f / l_run;
      const int t_row = qb * kItemQ + t * kTileQ + row;
      if (t_row < p.t_len) {
        __nv_bfloat16* dst = out + (size_t(clip) * p.t_len + t_row) * p.dims + head * kHeadDim;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint32_t w0[4], w1[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int i = q4 * 8 + e * 2;
            __nv_bfloat162 h0 = __floats2bfloat162_rn(__uint_as_float(o0[i]) * inv_l, __uint_as_float(o0[i + 1]) * inv_l);
            __nv_bfloat162 h1 = __floats2bfloat162_rn(__uint_as_float(o1[i]) * inv_l, __uint_as_float(o1[i + 1]) * inv_l);
            w0[e] = *reinterpret_cast<uint32_t*>(&h0);
            w1[e] = *reinterpret_cast<uint32_t*>(&h1);
          }
          *reinterpret_cast<uint4*>(dst + q4 * 8) = make_uint4(w0[0], w0[1], w0[2], w0[3]);
          *reinterpret_cast<uint4*>(dst + 32 + q4 * 8) = make_uint4(w1[0], w1[1], w1[2], w1[3]);
        }
      }
      ATT_T(7);
    }
#ifdef YOHO_ATT_TIMING
    if (blockIdx.x == 0 && threadIdx.x == 0)
      for (int i = 0; i < 8; ++i) g_att_phase[i] = tacc[i];
#endif
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<kAttTmemCols>(tmem_base);
}

int g_att_tiles_override = 0;   // tests: 0 = by padding, 2 / 3 = that kernel (yoho_test_attention_tiles)

int launch_attention_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s) {
  if (!qkv || !out) return fail(YOHO_ERR_INVALID, "attention: null operand");
  if (batch <= 0 || t <= 0 || heads <= 0) return YOHO_OK;
  {
    // Three-tile kernel (attention3_sm100.cu) unless its 384-row work items pad the queries so much further than 256-row
    // items that the faster kernel loses: cost ~ padded query rows / stand-alone throughput (816 vs 700 TFLOP/s).
    static const int env_forced = [] { const char* e = getenv("YOHO_ATT_TILES"); return e ? atoi(e) : 0; }();
    const int forced = g_att_tiles_override ? g_att_tiles_override : env_forced;
    const bool pad_ok = (long long)ceil_div(t, 384) * 384 * 700 <= (long long)ceil_div(t, kItemQ) * kItemQ * 816;
    if (forced == 3 || (forced != 2 && pad_ok)) return launch_attention3_bf16(ctx, qkv, out, batch, t, heads, s);
  }
  const int dims = heads * kHeadDim;
  AttParams p{};
  p.t_len = t; p.dims = dims; p.heads = heads;
  p.q_blocks = ceil_div(t, kItemQ);
  p.n_kv = ceil_div(t, kTileKV);
  p.n_steps = ceil_div(t, kStepKV);
  const long long items = (long long)batch * heads * p.q_blocks;
  if (items > 0x7fffffffLL) return fail(YOHO_ERR_UNSUPPORTED, "attention: %lld work items", items);
  p.n_items = int(items);
  CUtensorMap map;
  int r = make_tmap_bf16_3d(ctx->encode_tiled, &map, qkv, uint64_t(3 * dims), uint64_t(t), uint64_t(batch), uint64_t(3 * dims),
                            uint64_t(3 * dims) * uint64_t(t), kHeadDim, 128, 1);
  if (r) return fail(YOHO_ERR_CUDA, "attention: cuTensorMapEncodeTiled failed with CUresult %d", r);
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(attention_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(AttSmem))));
  const int grid = p.n_items < ctx->num_sms ? p.n_items : ctx->num_sms;
  ProfScope prof(s, kProfAttention, 4.0 * double(batch) * double(t) * double(t) * double(dims));
  YOHO_CUDA_CHECK(launch_pdl(attention_bf16_kernel, dim3(grid), dim3(kAttThreads), sizeof(AttSmem), s, map, out, p));
  YOHO_LAUNCH_CHECK("attention_bf16_kernel");
  return YOHO_OK;
}

}  // namespace yoho

#ifdef YOHO_ATT_TIMING
// debug builds only (-DYOHO_ATT_TIMING): clock64 totals of the softmax phases of CTA 0 / warp 0 of the last launch
extern "C" int yoho_debug_att_phases(long long* out16) {
  return int(cudaMemcpyFromSymbol(out16, yoho::g_att_phase, 16 * sizeof(long long)));
}
#endif

// test hook: pick the attention kernel (0 = automatic, 2 = two-tile, 3 = three-tile); returns the previous setting
extern "C" int yoho_test_attention_tiles(int tiles) {
  const int prev = yoho::g_att_tiles_override;
  yoho::g_att_tiles_override = (tiles == 2 || tiles == 3) ? tiles : 0;
  return prev;
}
