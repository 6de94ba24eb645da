// Fused self-attention forward for the encoder on tcgen05 tensor cores (sm_100a):
//   out = softmax(q k^T / sqrt(64)) v   per (clip, head), no mask      (reference whisper.py:52-67)
// The reference materialises the (B, H, 1500, 1500) score tensor; here scores never leave the SM.
//
// CTA = (clip, head, 128 query rows), 6 warps:
//   warps 0-3  softmax: thread i owns query row i (= TMEM lane i), so row max / row sum need no shuffles
//   warp 4     TMA producer: Q once, then K_j / V_j tiles of 128 keys, double buffered
//   warp 5     MMA issuer + TMEM allocator: S_j = Q K_j^T (M128 N128 K64), O_j = P_j V_j (M128 N64 K128)
// S (128 columns) and the per-tile O_j (64 columns) live in TMEM (256 columns allocated, two CTAs
// share an SM and overlap each other's softmax and MMA phases).  P_j is written to shared memory as
// bf16 in the K-major 128-byte-swizzled operand layout; V_j is consumed straight from its TMA tile as
// an MN-major operand.  The running output is kept in fp32 registers and rescaled online.
// Keys past the clip end (1500 is not a multiple of 128) are TMA zero-filled and masked to -inf.
#include "kernels.cuh"
#include "ptx_sm100.cuh"

namespace yoho {

using namespace ptx;

constexpr int kAttThreads = 192;
constexpr int kTileQ = 128, kTileKV = 128, kHeadDim = 64;
constexpr int kTileBytes = 128 * 64 * 2;  // one (128 rows x 64 bf16) swizzled tile = 16 KB
constexpr uint32_t kAttTmemCols = 256;
constexpr uint32_t kTmemS = 0, kTmemO = 128;

struct AttSmem {
  unsigned char q[kTileBytes];
  unsigned char k[2][kTileBytes];
  unsigned char v[2][kTileBytes];
  unsigned char p[2][kTileBytes];  // P_j: keys 0-63 | keys 64-127
  uint64_t q_full, k_full[2], k_empty[2], v_full[2], v_empty[2], s_full, p_full, o_full;
  uint32_t tmem_slot;
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kAttThreads, 2)
attention_bf16_kernel(const __grid_constant__ CUtensorMap map_qkv, __nv_bfloat16* __restrict__ out, int t_len, int dims) {
  extern __shared__ __align__(1024) unsigned char att_smem_raw[];
  AttSmem& sm = *reinterpret_cast<AttSmem*>(att_smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * kTileQ, head = blockIdx.y, clip = blockIdx.z;
  const int n_kv = (t_len + kTileKV - 1) / kTileKV;

  if (threadIdx.x == 0) {
    if (smem_u32(att_smem_raw) & 1023u) __trap();  // operand tiles need 1024-byte alignment (128B swizzle)
    mbar_init(&sm.q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sm.k_full[i], 1); mbar_init(&sm.k_empty[i], 1);
      mbar_init(&sm.v_full[i], 1); mbar_init(&sm.v_empty[i], 1);
    }
    mbar_init(&sm.s_full, 1);
    mbar_init(&sm.p_full, 4);
    mbar_init(&sm.o_full, 1);
    fence_mbar_init();
  }
  if (warp == 4 && lane == 0) prefetch_tmap(&map_qkv);
  if (warp == 5) {
    tmem_alloc<kAttTmemCols>(&sm.tmem_slot);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_slot;

  if (warp == 4) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_arrive_expect_tx(&sm.q_full, kTileBytes);
      tma_load_3d(sm.q, &map_qkv, &sm.q_full, head * kHeadDim, q0, clip);
    }
    for (int j = 0; j < n_kv; ++j) {
      const int slot = j & 1; const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&sm.k_empty[slot], ph ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(&sm.k_full[slot], kTileBytes);
        tma_load_3d(sm.k[slot], &map_qkv, &sm.k_full[slot], dims + head * kHeadDim, j * kTileKV, clip);
      }
      mbar_wait(&sm.v_empty[slot], ph ^ 1);
      if (lane == 0) {
        mbar_arrive_expect_tx(&sm.v_full[slot], kTileBytes);
        tma_load_3d(sm.v[slot], &map_qkv, &sm.v_full[slot], 2 * dims + head * kHeadDim, j * kTileKV, clip);
      }
      __syncwarp();
    }
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc_qk = make_idesc_bf16(kTileQ, kTileKV, 0, 0);   // A = Q (K-major), B = K (K-major)
    constexpr uint32_t idesc_pv = make_idesc_bf16(kTileQ, kHeadDim, 0, 1);  // A = P (K-major), B = V (MN-major)
    const uint32_t q_addr = smem_u32(sm.q);
    auto issue_qk = [&](int j) {
      const int slot = j & 1; const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&sm.k_full[slot], ph);
      tc_fence_after();
      if (lane == 0) {
        const uint64_t adesc = make_desc_sw128(q_addr, 16, 1024);
        const uint64_t bdesc = make_desc_sw128(smem_u32(sm.k[slot]), 16, 1024);
#pragma unroll
        for (int k = 0; k < kHeadDim / 16; ++k)
          umma_bf16_ss(tmem_base + kTmemS, adesc + uint64_t(2 * k), bdesc + uint64_t(2 * k), idesc_qk, k != 0);
        umma_commit(&sm.s_full);
        umma_commit(&sm.k_empty[slot]);
      }
      __syncwarp();
    };
    mbar_wait(&sm.q_full, 0);
    issue_qk(0);
    for (int j = 0; j < n_kv; ++j) {
      const int slot = j & 1; const uint32_t ph = (j >> 1) & 1;
      mbar_wait(&sm.p_full, j & 1);      // P_j in smem; S_j and O_{j-1} have been read out of TMEM
      mbar_wait(&sm.v_full[slot], ph);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t p_addr = smem_u32(sm.p[0]);
        const uint32_t v_addr = smem_u32(sm.v[slot]);
#pragma unroll
        for (int ks = 0; ks < kTileKV / 16; ++ks) {
          // A: P tile (ks / 4), 32 B per 16-key step inside the swizzle row
          const uint64_t adesc = make_desc_sw128(p_addr + (ks >> 2) * kTileBytes + (ks & 3) * 32, 16, 1024);
          // B: V rows are keys (128 B each): 16 keys = 2 groups of 8 rows, 1024 B apart
          const uint64_t bdesc = make_desc_sw128(v_addr + ks * 2048, 1024, 1024);
          umma_bf16_ss(tmem_base + kTmemO, adesc, bdesc, idesc_pv, ks != 0);
        }
        umma_commit(&sm.o_full);
        umma_commit(&sm.v_empty[slot]);
      }
      __syncwarp();
      if (j + 1 < n_kv) issue_qk(j + 1);
    }
  } else {
    // ===================== softmax + output (warps 0-3) =====================
    const int row = warp * 32 + lane;                       // query row in the tile == TMEM lane
    const uint32_t lane_addr = tmem_base + (uint32_t(warp * 32) << 16);
    const float c = 0.125f * 1.4426950408889634f;            // d_head^-0.5 * log2(e)   (whisper.py:54-57)
    float m_run = -INFINITY, l_run = 0.0f;
    float o[kHeadDim];
#pragma unroll
    for (int i = 0; i < kHeadDim; ++i) o[i] = 0.0f;
    unsigned char* p_row = sm.p[0] + row * 128;
    const int sw = row & 7;

    for (int j = 0; j < n_kv; ++j) {
      const int kv_valid = t_len - j * kTileKV;              // keys of this tile inside the clip
      mbar_wait(&sm.s_full, j & 1);
      tc_fence_after();
      // pass A: row maximum
      float m_tile = -INFINITY;
#pragma unroll 1
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t s[32];
        tmem_ld_32x32b_x32(lane_addr + kTmemS + ch * 32, s);
        tmem_ld_wait();
        if (kv_valid >= (ch + 1) * 32) {
#pragma unroll
          for (int i = 0; i < 32; ++i) m_tile = fmaxf(m_tile, __uint_as_float(s[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (ch * 32 + i < kv_valid) m_tile = fmaxf(m_tile, __uint_as_float(s[i]));
        }
      }
      const float m_new = fmaxf(m_run, m_tile);
      const float alpha = ex2_approx((m_run - m_new) * c);
      const float mc = m_new * c;
      m_run = m_new;
      // fold the previous tile's P V product into the running output, then rescale
      if (j > 0) {
        mbar_wait(&sm.o_full, (j - 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {
          uint32_t t[32];
          tmem_ld_32x32b_x32(lane_addr + kTmemO + ch * 32, t);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[ch * 32 + i] = (o[ch * 32 + i] + __uint_as_float(t[i])) * alpha;
        }
      }
      // pass B: p = exp2(s c - m c), row sum, bf16 P tile into swizzled smem
      float l_tile = 0.0f;
#pragma unroll 1
      for (int ch = 0; ch < 4; ++ch) {
        uint32_t s[32];
        tmem_ld_32x32b_x32(lane_addr + kTmemS + ch * 32, s);
        tmem_ld_wait();
        uint32_t packed[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float p0 = ex2_approx(fmaf(__uint_as_float(s[2 * i]), c, -mc));
          float p1 = ex2_approx(fmaf(__uint_as_float(s[2 * i + 1]), c, -mc));
          if (ch * 32 + 2 * i >= kv_valid) p0 = 0.0f;
          if (ch * 32 + 2 * i + 1 >= kv_valid) p1 = 0.0f;
          l_tile += p0 + p1;
          __nv_bfloat162 h = __floats2bfloat162_rn(p0, p1);
          packed[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        // keys ch*32 .. +31 -> P tile (ch / 2), 16-byte chunks (ch % 2) * 4 + {0..3}, XOR-swizzled by row
        unsigned char* tile = p_row + (ch >> 1) * kTileBytes;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = ((ch & 1) * 4 + g) ^ sw;
          *reinterpret_cast<uint4*>(tile + chunk * 16) = make_uint4(packed[4 * g], packed[4 * g + 1], packed[4 * g + 2], packed[4 * g + 3]);
        }
      }
      l_run = l_run * alpha + l_tile;
      fence_proxy_async_smem();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.p_full);
    }
    // last tile's product, normalise, store
    mbar_wait(&sm.o_full, (n_kv - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.0f / l_run;
    const int t_row = q0 + row;
    __nv_bfloat16* dst = out + (size_t(clip) * t_len + t_row) * dims + head * kHeadDim;
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
      uint32_t t[32];
      tmem_ld_32x32b_x32(lane_addr + kTmemO + ch * 32, t);
      tmem_ld_wait();
      if (t_row < t_len) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            int i = g * 8 + e * 2;
            float a = (o[ch * 32 + i] + __uint_as_float(t[i])) * inv_l;
            float b = (o[ch * 32 + i + 1] + __uint_as_float(t[i + 1])) * inv_l;
            __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
            w[e] = *reinterpret_cast<uint32_t*>(&h);
          }
          *reinterpret_cast<uint4*>(dst + ch * 32 + g * 8) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc<kAttTmemCols>(tmem_base);
}

int launch_attention_bf16(yoho_ctx* ctx, const __nv_bfloat16* qkv, __nv_bfloat16* out, int batch, int t, int heads, cudaStream_t s) {
  if (!qkv || !out) return fail(YOHO_ERR_INVALID, "attention: null operand");
  if (batch <= 0 || t <= 0 || heads <= 0) return YOHO_OK;
  if (heads > 65535 || batch > 65535) return fail(YOHO_ERR_UNSUPPORTED, "attention: heads=%d batch=%d exceed one launch", heads, batch);
  const int dims = heads * kHeadDim;
  CUtensorMap map;
  int r = make_tmap_bf16_3d(ctx->encode_tiled, &map, qkv, uint64_t(3 * dims), uint64_t(t), uint64_t(batch), uint64_t(3 * dims),
                            uint64_t(3 * dims) * uint64_t(t), kHeadDim, 128, 1);
  if (r) return fail(YOHO_ERR_CUDA, "attention: cuTensorMapEncodeTiled failed with CUresult %d", r);
  YOHO_CUDA_CHECK(cudaFuncSetAttribute(attention_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(AttSmem))));
  dim3 grid(ceil_div(t, kTileQ), heads, batch);
  ProfScope prof(s, kProfAttention, 4.0 * double(batch) * double(t) * double(t) * double(dims));
  attention_bf16_kernel<<<grid, kAttThreads, sizeof(AttSmem), s>>>(map, out, t, dims);
  YOHO_LAUNCH_CHECK("attention_bf16_kernel");
  return YOHO_OK;
}

}  // namespace yoho
