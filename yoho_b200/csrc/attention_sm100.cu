// Fused self-attention forward for the encoder on tcgen05 tensor cores (sm_100a):
//   out = softmax(q k^T / sqrt(64)) v   per (clip, head), no mask      (reference whisper.py:52-67)
// The reference materialises the (B, H, 1500, 1500) score tensor; here scores never leave the SM.
//
// Persistent kernel, one CTA per SM, 10 warps.  A work item is (clip, head, 256 query rows) = two
// 128-row query tiles A and B that share every K/V tile:
//   warps 0-3  softmax of tile A   (thread i owns query row i = TMEM lane i: row max / sum need no shuffles)
//   warps 4-7  softmax of tile B
//   warp 8     TMA producer: Q tiles (double buffered across items), K_j / V_j rings of 128 keys
//   warp 9     MMA issuer + TMEM allocator
// Per KV tile j and query tile t:  S_t = Q_t K_j^T (M128 N128 K64, operands in shared memory)
//                                  O_t += P_t V_j  (M128 N64 K128, P read from TMEM, V MN-major from its TMA tile)
// TMEM: S_A | S_B (128 fp32 columns each), O_A | O_B (64 columns each).  The softmax threads read S twice
// (row max, then exp2), write P as packed bf16 over the first 64 columns of S (tcgen05.st) and keep the row
// sum in registers.  O stays in TMEM across KV tiles: it is only rescaled when the running maximum
// grows by more than 2^8 (lazy rescaling; P is then bounded by 256), so the common case has no correction.
// While one tile's softmax runs on the SFU (16 ex2/clk/SM is the bound at d_head = 64), the tensor core
// runs the other tile's MMAs.  Keys past the clip end are TMA zero-filled and masked to -inf.
#include "kernels.cuh"
#include "ptx_sm100.cuh"

namespace yoho {

using namespace ptx;

constexpr int kAttThreads = 320;
constexpr int kTileQ = 128, kTileKV = 128, kHeadDim = 64;
constexpr int kItemQ = 2 * kTileQ;        // query rows per work item
constexpr int kTileBytes = 128 * 64 * 2;  // one (128 rows x 64 bf16) swizzled tile = 16 KB
constexpr int kKvStages = 4;
constexpr uint32_t kAttTmemCols = 512;
constexpr uint32_t kTmemS = 0, kTmemO = 256;   // S_t at kTmemS + 128 t, O_t at kTmemO + 64 t
constexpr float kRescaleLog2 = 8.0f;

struct AttSmem {
  unsigned char q[2][2][kTileBytes];   // [item slot][query tile]
  unsigned char k[kKvStages][kTileBytes];
  unsigned char v[kKvStages][kTileBytes];
  uint64_t q_full[2][2], q_empty[2][2];
  uint64_t k_full[kKvStages], k_empty[kKvStages], v_full[kKvStages], v_empty[kKvStages];
  uint64_t s_full[2], p_full[2], o_full[2];
  uint32_t tmem_slot;
};

struct AttParams {
  int t_len, dims, heads;
  int q_blocks;      // ceil(t_len / 256)
  int n_items;       // batch * heads * q_blocks
  int n_kv;          // ceil(t_len / 128)
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float y;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(y) : "f"(a), "f"(b), "f"(c));
  return y;
}

// Row maximum of one 128-column S tile (this thread's TMEM lane); columns >= kv_valid are ignored when kMask.
template <bool kMask>
__device__ __forceinline__ float tile_row_max(uint32_t s_addr, int kv_valid) {
  float m = -INFINITY;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t a[32], b[32];
    tmem_ld_32x32b_x32(s_addr + h * 64, a);
    tmem_ld_32x32b_x32(s_addr + h * 64 + 32, b);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float x0 = __uint_as_float(a[i]), x1 = __uint_as_float(a[i + 1]);
      float y0 = __uint_as_float(b[i]), y1 = __uint_as_float(b[i + 1]);
      if (kMask) {
        if (h * 64 + i >= kv_valid) x0 = -INFINITY;
        if (h * 64 + i + 1 >= kv_valid) x1 = -INFINITY;
        if (h * 64 + 32 + i >= kv_valid) y0 = -INFINITY;
        if (h * 64 + 32 + i + 1 >= kv_valid) y1 = -INFINITY;
      }
      m = fmax3(m, x0, x1);
      m = fmax3(m, y0, y1);
    }
  }
  return m;
}

// p = exp2(s c - mc) for the 128 columns of this thread's row; P is written as packed bf16 over the first 64
// columns of the S tile (columns already consumed); returns the row sum of the unrounded p.
template <bool kMask>
__device__ __forceinline__ float tile_exp_store(uint32_t s_addr, float c, float mc, int kv_valid) {
  float l0 = 0.0f, l1 = 0.0f;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t a[32], b[32];
    tmem_ld_32x32b_x32(s_addr + h * 64, a);
    tmem_ld_32x32b_x32(s_addr + h * 64 + 32, b);
    tmem_ld_wait();
    uint32_t packed[32];
#pragma unroll
    for (int i = 0; i < 32; i += 2) {
      float p0 = ex2_approx(fmaf(__uint_as_float(a[i]), c, -mc));
      float p1 = ex2_approx(fmaf(__uint_as_float(a[i + 1]), c, -mc));
      float r0 = ex2_approx(fmaf(__uint_as_float(b[i]), c, -mc));
      float r1 = ex2_approx(fmaf(__uint_as_float(b[i + 1]), c, -mc));
      if (kMask) {
        if (h * 64 + i >= kv_valid) p0 = 0.0f;
        if (h * 64 + i + 1 >= kv_valid) p1 = 0.0f;
        if (h * 64 + 32 + i >= kv_valid) r0 = 0.0f;
        if (h * 64 + 32 + i + 1 >= kv_valid) r1 = 0.0f;
      }
      l0 += p0 + p1;
      l1 += r0 + r1;
      __nv_bfloat162 u = __floats2bfloat162_rn(p0, p1), w = __floats2bfloat162_rn(r0, r1);
      packed[i >> 1] = *reinterpret_cast<uint32_t*>(&u);          // keys h*64 + i, i+1
      packed[16 + (i >> 1)] = *reinterpret_cast<uint32_t*>(&w);   // keys h*64 + 32 + i, i+1
    }
    tmem_st_32x32b_x32(s_addr + h * 32, packed);                  // P columns h*32 .. +31 = keys h*64 .. +63
  }
  tmem_st_wait();
  return l0 + l1;
}

__global__ void __launch_bounds__(kAttThreads, 1)
attention_bf16_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, const AttParams p) {
  extern __shared__ __align__(1024) unsigned char att_smem_raw[];
  AttSmem& sm = *reinterpret_cast<AttSmem*>(att_smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    if (smem_u32(att_smem_raw) & 1023u) __trap();  // operand tiles need 1024-byte alignment (128B swizzle)
    for (int s = 0; s < 2; ++s)
      for (int t = 0; t < 2; ++t) { mbar_init(&sm.q_full[s][t], 1); mbar_init(&sm.q_empty[s][t], 1); }
    for (int i = 0; i < kKvStages; ++i) {
      mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 1);
      mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], 1);
    }
    for (int t = 0; t < 2; ++t) { mbar_init(&sm.s_full[t], 1); mbar_init(&sm.p_full[t], 4); mbar_init(&sm.o_full[t], 1); }
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
  const uint32_t tmem_base = sm.tmem_slot;
  const int n_kv = p.n_kv;

  if (warp == 8) {
    // ===================== TMA producer (one lane) =====================
    if (lane == 0) {
      int kc = 0;   // K/V tiles issued so far (ring position)
      int it = 0;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++it) {
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
  } else if (warp == 9) {
    // ===================== MMA issuer (one lane) =====================
    if (lane == 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(kTileQ, kTileKV, 0, 0);   // A = Q (K-major), B = K (K-major)
      constexpr uint32_t idesc_pv = make_idesc_bf16(kTileQ, kHeadDim, 0, 1);  // A = P (TMEM), B = V (MN-major)
      int kq = 0;   // K tiles consumed by QK so far (ring position of the NEXT QK)
      int kv = 0;   // V tiles consumed by PV so far
      int g = 0;    // KV steps completed (phase of s_full / p_full / o_full)
      // S_t = Q_t K^T for the K tile at ring position kq; releases the K slot after tile B.
      auto issue_qk = [&](int t, int slot, bool last_of_item) {
        const int st = kq % kKvStages;
        if (t == 0) { mbar_wait(&sm.k_full[st], (kq / kKvStages) & 1); }
        tc_fence_after();
        const uint64_t adesc = make_desc_sw128(smem_u32(sm.q[slot][t]), 16, 1024);
        const uint64_t bdesc = make_desc_sw128(smem_u32(sm.k[st]), 16, 1024);
#pragma unroll
        for (int k = 0; k < kHeadDim / 16; ++k)
          umma_bf16_ss(tmem_base + kTmemS + uint32_t(t * 128), adesc + uint64_t(2 * k), bdesc + uint64_t(2 * k), idesc_qk, k != 0);
        umma_commit(&sm.s_full[t]);
        if (last_of_item) umma_commit(&sm.q_empty[slot][t]);
        if (t == 1) { umma_commit(&sm.k_empty[st]); ++kq; }
      };
      auto issue_pv = [&](int t, int j) {
        const int st = kv % kKvStages;
        mbar_wait(&sm.p_full[t], g & 1);      // P_t in TMEM, S_t consumed, O_t rescaled if needed
        if (t == 0) mbar_wait(&sm.v_full[st], (kv / kKvStages) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sm.v[st]);
#pragma unroll
        for (int ks = 0; ks < kTileKV / 16; ++ks) {
          // B: V rows are keys (128 B each): 16 keys = 2 groups of 8 rows, 1024 B apart
          const uint64_t bdesc = make_desc_sw128(v_addr + ks * 2048, 1024, 1024);
          umma_bf16_ts(tmem_base + kTmemO + uint32_t(t * 64), tmem_base + kTmemS + uint32_t(t * 128 + ks * 8), bdesc, idesc_pv,
                       (j | ks) != 0);
        }
        umma_commit(&sm.o_full[t]);
        if (t == 1) { umma_commit(&sm.v_empty[st]); ++kv; }
      };
      int it = 0;
      const int first = blockIdx.x;
      if (first < p.n_items) {
        mbar_wait(&sm.q_full[0][0], 0);
        issue_qk(0, 0, n_kv == 1);
        mbar_wait(&sm.q_full[0][1], 0);
        issue_qk(1, 0, n_kv == 1);
      }
      for (int item = first; item < p.n_items; item += gridDim.x, ++it) {
        const int slot = it & 1;
        const bool has_next_item = item + int(gridDim.x) < p.n_items;
        const int nslot = (it + 1) & 1; const uint32_t nqph = ((it + 1) >> 1) & 1;
        for (int j = 0; j < n_kv; ++j, ++g) {
          const bool in_item = j + 1 < n_kv;
          const bool next = in_item || has_next_item;
          const int qslot = in_item ? slot : nslot;
          const bool next_last = in_item ? (j + 2 == n_kv) : (n_kv == 1);
          issue_pv(0, j);
          if (next) {
            if (!in_item) mbar_wait(&sm.q_full[nslot][0], nqph);
            issue_qk(0, qslot, next_last);
          }
          issue_pv(1, j);
          if (next) {
            if (!in_item) mbar_wait(&sm.q_full[nslot][1], nqph);
            issue_qk(1, qslot, next_last);
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== softmax + output: warps 0-3 tile A, warps 4-7 tile B =====================
    const int t = warp >> 2;
    const int row = (warp & 3) * 32 + lane;                   // query row in the tile == TMEM lane
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    const uint32_t s_addr = lane_base + kTmemS + uint32_t(t * 128);
    const uint32_t o_addr = lane_base + kTmemO + uint32_t(t * 64);
    const float c = 0.125f * 1.4426950408889634f;            // d_head^-0.5 * log2(e)   (whisper.py:54-57)
    const int kv_tail = p.t_len - (n_kv - 1) * kTileKV;       // valid keys of the last KV tile
    int g = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      const int qb = item % p.q_blocks, ch = item / p.q_blocks;
      const int head = ch % p.heads, clip = ch / p.heads;
      float m_ref = -INFINITY, l_run = 0.0f;
      for (int j = 0; j < n_kv; ++j, ++g) {
        const bool tail = (j == n_kv - 1) && kv_tail < kTileKV;   // CTA-uniform
        mbar_wait(&sm.s_full[t], g & 1);
        tc_fence_after();
        const float m_tile = tail ? tile_row_max<true>(s_addr, kv_tail) : tile_row_max<false>(s_addr, kv_tail);
        if (j == 0) {
          m_ref = m_tile;
        } else if (__any_sync(0xffffffffu, (m_tile - m_ref) * c > kRescaleLog2)) {
          // rare: the running maximum grew by more than 2^8 somewhere in this warp -> rescale O and l
          const float m_new = fmaxf(m_ref, m_tile);
          const float alpha = ex2_approx((m_ref - m_new) * c);
          m_ref = m_new;
          l_run *= alpha;
          mbar_wait(&sm.o_full[t], (g - 1) & 1);                  // P V of the previous KV tile has landed in O
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
        const float mc = m_ref * c;
        l_run += tail ? tile_exp_store<true>(s_addr, c, mc, kv_tail) : tile_exp_store<false>(s_addr, c, mc, kv_tail);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.p_full[t]);
      }
      // last P V of the item, normalise, store
      mbar_wait(&sm.o_full[t], (g - 1) & 1);
      tc_fence_after();
      const float inv_l = 1.0f / l_run;
      const int t_row = qb * kItemQ + t * kTileQ + row;
      __nv_bfloat16* dst = out + (size_t(clip) * p.t_len + t_row) * p.dims + head * kHeadDim;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(o_addr + h * 32, o);
        tmem_ld_wait();
        if (t_row < p.t_len) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int i = q4 * 8 + e * 2;
              __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(o[i]) * inv_l, __uint_as_float(o[i + 1]) * inv_l);
              w[e] = *reinterpret_cast<uint32_t*>(&hh);
            }
            *reinterpret_cast<uint4*>(dst + h * 32 + q4 * 8) = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
      }
      // O_t / S_t are overwritten by the next item's MMAs only after this warp arrives on p_full again
      tc_fence_before();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<kAttTmemCols>(tmem_base);
}

int launch_attention_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s) {
  if (!qkv || !out) return fail(YOHO_ERR_INVALID, "attention: null operand");
  if (batch <= 0 || t <= 0 || heads <= 0) return YOHO_OK;
  const int dims = heads * kHeadDim;
  AttParams p{};
  p.t_len = t; p.dims = dims; p.heads = heads;
  p.q_blocks = ceil_div(t, kItemQ);
  p.n_kv = ceil_div(t, kTileKV);
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
  attention_bf16_kernel<<<grid, kAttThreads, sizeof(AttSmem), s>>>(map, out, p);
  YOHO_LAUNCH_CHECK("attention_bf16_kernel");
  return YOHO_OK;
}

}  // namespace yoho
