// K6: causal flash attention, forward, for head_dim 128 with GQA, on tcgen05.
//
// Operates directly on the fused QKV activation [T, (Hq+2Hkv)*128] (RoPE already applied) through
// ONE TMA tensor map: Q, K and V tiles are just different column offsets of the same matrix.
//
// One CTA = one 128-row query tile of one (batch, head). Roles:
//   warp 0      TMA producer : Q once; K_j / V_j tiles into 2-stage rings (128B swizzle)
//   warp 1      MMA issuer   : S_j = Q K_j^T  (UMMA 128x128x16, fp32 in TMEM, double-buffered)
//                              O += P_j V_j   (A = P_j from smem K-major, B = V_j MN-major)
//   warps 2..5  softmax      : one thread per query row; tcgen05.ld S -> online softmax in the
//                              exp2 domain -> P (bf16) into swizzled smem for the second MMA.
// O stays in TMEM for the whole KV loop. The running max is updated lazily: O and l are rescaled
// only when the max grows by more than 2^8 (exact after the final normalisation; keeps the
// TMEM round trip off the common path). Heavy (late) query tiles are scheduled first.
//
// The reference has no attention code (it ships no model: SURVEY.md §2.3 F2).
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int FA_HD = 128;
constexpr int FA_BM = 128;  // query rows per CTA
constexpr int FA_BN = 128;  // kv rows per block
constexpr int FA_THREADS = 192;
constexpr int FA_TILE_BYTES = 128 * 128 * 2;  // 32 KB: two 64-column halves of 16 KB
constexpr int FA_HALF_BYTES = 128 * 64 * 2;
constexpr int FA_SMEM = 6 * FA_TILE_BYTES /*Q,K0,K1,V0,V1,P*/ + 256 + 1024;
constexpr float FA_RESCALE_THRESHOLD = 8.f;

struct FaBars {
  // byte offsets (from bar_base) of every mbarrier
  static constexpr int q_full = 0, k_full = 8, k_empty = 24, v_full = 40, v_empty = 56, s_full = 72, s_empty = 88,
                       p_full = 104, pv_done = 112, tmem_slot = 120;
};

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(FA_THREADS, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
                 int B, int S, int Hq, int Hk, int ld_out, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t q_s = base;
  const uint32_t k_s = base + FA_TILE_BYTES;       // 2 stages
  const uint32_t v_s = base + 3 * FA_TILE_BYTES;   // 2 stages
  const uint32_t p_s = base + 5 * FA_TILE_BYTES;
  const uint32_t bar = base + 6 * FA_TILE_BYTES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = S / FA_BM;
  const int qi = nq - 1 - (int)(blockIdx.x % nq);  // heavy tiles first
  const int h = (int)(blockIdx.x / nq) % Hq;
  const int b = (int)(blockIdx.x / nq) / Hq;
  const int g = h / (Hq / Hk);
  const int row0 = b * S + qi * FA_BM;
  const int nblk = qi + 1;  // causal: kv blocks 0..qi
  const int qcol = h * FA_HD, kcol = (Hq + g) * FA_HD, vcol = (Hq + Hk + g) * FA_HD;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm);
    mbar_init(bar + FaBars::q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar + FaBars::k_full + 8 * i, 1);
      mbar_init(bar + FaBars::k_empty + 8 * i, 1);
      mbar_init(bar + FaBars::v_full + 8 * i, 1);
      mbar_init(bar + FaBars::v_empty + 8 * i, 1);
      mbar_init(bar + FaBars::s_full + 8 * i, 1);
      mbar_init(bar + FaBars::s_empty + 8 * i, 128);
    }
    mbar_init(bar + FaBars::p_full, 128);
    mbar_init(bar + FaBars::pv_done, 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<1>(bar + FaBars::tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(bar + FaBars::tmem_slot));
  const uint32_t tm_s[2] = {tmem, tmem + 128};
  const uint32_t tm_o = tmem + 256;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar + FaBars::q_full, FA_TILE_BYTES);
      tma_load_2d(&tm, bar + FaBars::q_full, q_s, qcol, row0);
      tma_load_2d(&tm, bar + FaBars::q_full, q_s + FA_HALF_BYTES, qcol + 64, row0);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const int kv0 = b * S + j * FA_BN;
        mbar_wait(bar + FaBars::k_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + FaBars::k_full + 8 * st, FA_TILE_BYTES);
        tma_load_2d(&tm, bar + FaBars::k_full + 8 * st, k_s + st * FA_TILE_BYTES, kcol, kv0);
        tma_load_2d(&tm, bar + FaBars::k_full + 8 * st, k_s + st * FA_TILE_BYTES + FA_HALF_BYTES, kcol + 64, kv0);
        mbar_wait(bar + FaBars::v_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + FaBars::v_full + 8 * st, FA_TILE_BYTES);
        tma_load_2d(&tm, bar + FaBars::v_full + 8 * st, v_s + st * FA_TILE_BYTES, vcol, kv0);
        tma_load_2d(&tm, bar + FaBars::v_full + 8 * st, v_s + st * FA_TILE_BYTES + FA_HALF_BYTES, vcol + 64, kv0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 128, false, true);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(bar + FaBars::k_full + 8 * st, ph);
        mbar_wait(bar + FaBars::s_empty + 8 * st, ph ^ 1u);  // softmax finished reading S[j-2]
        tc_fence_after();
        const uint32_t kb = k_s + st * FA_TILE_BYTES;
#pragma unroll
        for (int k = 0; k < FA_HD / 16; ++k) {
          const uint32_t off = (k >> 2) * FA_HALF_BYTES + (k & 3) * 32;
          umma_f16<1>(tm_s[st], umma_smem_desc_sw128(q_s + off, 16, 1024), umma_smem_desc_sw128(kb + off, 16, 1024),
                      idesc_qk, k != 0 ? 1u : 0u);
        }
        umma_commit<1>(bar + FaBars::k_empty + 8 * st);
        umma_commit<1>(bar + FaBars::s_full + 8 * st);
      };
      mbar_wait(bar + FaBars::q_full, 0);
      issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_qk(j + 1);  // overlaps softmax of block j
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(bar + FaBars::v_full + 8 * st, ph);
        mbar_wait(bar + FaBars::p_full, j & 1);
        tc_fence_after();
        const uint32_t vb = v_s + st * FA_TILE_BYTES;
#pragma unroll
        for (int k = 0; k < FA_BN / 16; ++k) {
          const uint32_t aoff = (k >> 2) * FA_HALF_BYTES + (k & 3) * 32;
          umma_f16<1>(tm_o, umma_smem_desc_sw128(p_s + aoff, 16, 1024),
                      umma_smem_desc_sw128(vb + k * 2048, FA_HALF_BYTES, 1024), idesc_pv, (j | k) != 0 ? 1u : 0u);
        }
        umma_commit<1>(bar + FaBars::v_empty + 8 * st);
        umma_commit<1>(bar + FaBars::pv_done);
      }
    }
  } else {
    // ------------------------------ softmax: one thread per query row ------------------------------
    const int q = warp & 3;
    const int r = q * 32 + lane;            // row inside the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float m_used = -INFINITY;               // exponent reference currently baked into O and l (log2 domain)
    float l = 0.f;
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const bool diag = (j == nblk - 1);
      mbar_wait(bar + FaBars::s_full + 8 * st, ph);
      tc_fence_after();
      // pass 1: row max
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tm_s[st] + lane_off + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float s = __uint_as_float(v[e]);
          if (!diag || (c * 32 + e) <= r) mx = fmaxf(mx, s);
        }
      }
      const float m_new = fmaxf(m_used, mx * scale_log2);
      // lazy rescale: only when the reference exponent is too stale (or on the very first block)
      const bool need = (m_new - m_used) > FA_RESCALE_THRESHOLD;   // true when m_used == -inf
      if (j > 0) mbar_wait(bar + FaBars::pv_done, (j - 1) & 1);    // O and the P buffer are free again
      if (__any_sync(0xffffffffu, need) && j > 0) {
        tc_fence_after();
        const float alpha = need ? exp2f(m_used - m_new) : 1.f;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tm_o + lane_off + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) * alpha);
          tmem_st_32x32(tm_o + lane_off + c * 32, v);
        }
        tmem_st_wait();
        l *= alpha;
      }
      if (need) m_used = m_new;
      // pass 2: P = exp2(s*scale - m_used) -> bf16 -> swizzled smem (K-major, 2 halves of 64 kv columns)
      float rs = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(tm_s[st] + lane_off + c * 32, v);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float p0 = exp2f(__uint_as_float(v[e]) * scale_log2 - m_used);
          float p1 = exp2f(__uint_as_float(v[e + 1]) * scale_log2 - m_used);
          if (diag) {
            if (c * 32 + e > r) p0 = 0.f;
            if (c * 32 + e + 1 > r) p1 = 0.f;
          }
          rs += p0 + p1;
          pk[e >> 1] = pack_bf16x2(p0, p1);
        }
        // columns [c*32, c*32+32) -> half (c>>1), 16-byte chunks ((c&1)*4 + i), XOR-swizzled by row
        const uint32_t rowb = p_s + (c >> 1) * FA_HALF_BYTES + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int chunk = ((c & 1) * 4 + i) ^ (r & 7);
          st_shared_v4(rowb + chunk * 16, pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        }
      }
      l += rs;
      tc_fence_before();
      mbar_arrive(bar + FaBars::s_empty + 8 * st);   // S[st] may be overwritten by block j+2
      fence_proxy_async_smem();                      // make P visible to the tensor core (async proxy)
      mbar_arrive(bar + FaBars::p_full);
    }
    // epilogue: O / l -> bf16, LSE
    mbar_wait(bar + FaBars::pv_done, (nblk - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l;
    __nv_bfloat16* orow = out + (size_t)(row0 + r) * ld_out + h * FA_HD;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tm_o + lane_off + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        uint4 o;
        o.x = pack_bf16x2(__uint_as_float(v[8 * i]) * inv_l, __uint_as_float(v[8 * i + 1]) * inv_l);
        o.y = pack_bf16x2(__uint_as_float(v[8 * i + 2]) * inv_l, __uint_as_float(v[8 * i + 3]) * inv_l);
        o.z = pack_bf16x2(__uint_as_float(v[8 * i + 4]) * inv_l, __uint_as_float(v[8 * i + 5]) * inv_l);
        o.w = pack_bf16x2(__uint_as_float(v[8 * i + 6]) * inv_l, __uint_as_float(v[8 * i + 7]) * inv_l);
        *reinterpret_cast<uint4*>(orow + c * 32 + i * 8) = o;
      }
    }
    if (lse != nullptr) lse[(size_t)h * ((size_t)B * S) + row0 + r] = (m_used + log2f(l)) * 0.6931471805599453f;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem, 512);
  }
}

}  // namespace bz

using namespace bz;

extern "C" int bz_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int Hq, int Hk, int head_dim,
                                 float softmax_scale, void* stream) {
  BZ_REQUIRE(head_dim == FA_HD, "flash_attn_fwd: head_dim %d unsupported (128 only)", head_dim);
  BZ_REQUIRE(S % FA_BM == 0 && S > 0, "flash_attn_fwd: seq_len %d must be a multiple of %d", S, FA_BM);
  BZ_REQUIRE(Hq % Hk == 0, "flash_attn_fwd: Hq=%d not a multiple of Hkv=%d", Hq, Hk);
  const int ld = (Hq + 2 * Hk) * FA_HD;
  CUtensorMap tm;
  if (int rc = make_tmap_2d_bf16(&tm, qkv, (uint64_t)B * S, ld, ld, 128, 64)) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM));
    attr_done = true;
  }
  const int grid = (S / FA_BM) * Hq * B;
  const float scale_log2 = softmax_scale * 1.4426950408889634f;
  flash_fwd_kernel<<<grid, FA_THREADS, FA_SMEM, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, lse, B, S, Hq, Hk,
                                                                       Hq * FA_HD, scale_log2);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
