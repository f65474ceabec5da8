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
constexpr int FA_THREADS = 320;  // TMA warp, MMA warp, 8 softmax warps (two threads per query row)
constexpr int FA_TILE_BYTES = 128 * 128 * 2;  // 32 KB: two 64-column halves of 16 KB
constexpr int FA_HALF_BYTES = 128 * 64 * 2;
constexpr int FA_SMEM = 6 * FA_TILE_BYTES /*Q,K0,K1,V0,V1,P*/ + 256 + 3 * 2 * 128 * 4 /*pair exchange*/ + 1024;
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
      mbar_init(bar + FaBars::s_empty + 8 * i, 256);
    }
    mbar_init(bar + FaBars::p_full, 256);
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
    // ---------------- softmax: two threads per query row, each owning 64 of the 128 kv columns ----------------
    const int q = warp & 3;                       // TMEM lane quadrant of this warp
    const int hf = (warp - 2) >> 2;               // which half of the columns (warps 2-5: 0, warps 6-9: 1)
    const int r = q * 32 + lane;                  // query row inside the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float* xch = reinterpret_cast<float*>(smem_raw + (bar + 256 - smem_u32(smem_raw)));  // [3][2][128]
    float m_used = -INFINITY;                     // exponent reference baked into O and l (log2 domain)
    float l = 0.f;                                // this thread's share of the row sum
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const bool diag = (j == nblk - 1);
      mbar_wait(bar + FaBars::s_full + 8 * st, ph);
      tc_fence_after();
      uint32_t sv[64];
      tmem_ld_32x64(tm_s[st] + lane_off + hf * 64, sv);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar + FaBars::s_empty + 8 * st);          // S[st] is in registers: block j+2 may overwrite it
      float mx = -INFINITY;
#pragma unroll
      for (int e = 0; e < 64; ++e) {
        float s = __uint_as_float(sv[e]) * scale_log2;
        if (diag && (hf * 64 + e) > r) s = -INFINITY;
        sv[e] = __float_as_uint(s);
        mx = fmaxf(mx, s);
      }
      // combine the two halves of the row
      float* xm = xch + (j & 1) * 256;
      xm[hf * 128 + r] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
      mx = fmaxf(mx, xm[(hf ^ 1) * 128 + r]);
      const float m_new = fmaxf(m_used, mx);
      const bool need = (m_new - m_used) > FA_RESCALE_THRESHOLD;     // lazy rescale (always true on block 0)
      if (j > 0) mbar_wait(bar + FaBars::pv_done, (j - 1) & 1);     // O and the P buffer are free again
      if (__any_sync(0xffffffffu, need) && j > 0) {
        tc_fence_after();
        const float alpha = need ? fast_exp2(m_used - m_new) : 1.f;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(tm_o + lane_off + hf * 64 + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) * alpha);
          tmem_st_32x32(tm_o + lane_off + hf * 64 + c * 32, v);
        }
        tmem_st_wait();
        tc_fence_before();
        l *= alpha;
      }
      if (need) m_used = m_new;
      float rs = 0.f;
      uint32_t pk[32];
#pragma unroll
      for (int e = 0; e < 64; e += 2) {
        const float p0 = fast_exp2(__uint_as_float(sv[e]) - m_used);
        const float p1 = fast_exp2(__uint_as_float(sv[e + 1]) - m_used);
        rs += p0 + p1;
        pk[e >> 1] = pack_bf16x2(p0, p1);
      }
      l += rs;
      const uint32_t rowb = p_s + hf * FA_HALF_BYTES + r * 128;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        st_shared_v4(rowb + ((i ^ (r & 7)) * 16), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      fence_proxy_async_smem();                      // make P visible to the tensor core (async proxy)
      mbar_arrive(bar + FaBars::p_full);
    }
    // epilogue: combine l, O / l -> bf16, LSE
    float* xl = xch + 2 * 256;
    xl[hf * 128 + r] = l;
    asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
    l += xl[(hf ^ 1) * 128 + r];
    mbar_wait(bar + FaBars::pv_done, (nblk - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l;
    __nv_bfloat16* orow = out + (size_t)(row0 + r) * ld_out + h * FA_HD + hf * 64;
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(tm_o + lane_off + hf * 64 + c * 32, v);
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
    if (lse != nullptr && hf == 0)
      lse[(size_t)h * ((size_t)B * S) + row0 + r] = (m_used + log2f(l)) * 0.6931471805599453f;
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

// =====================================================================================
// Backward. One CTA owns a 128-row KV block of one (batch, kv head) and walks every query tile
// (64 rows) of every query head in the GQA group that can see it. All five GEMMs of the flash
// backward run on tcgen05 with the *transposed* score tile, so that every operand already has a
// layout the tensor core accepts (no register or smem transposes):
//     S^T  = K Q^T                    (M=kv 128, N=q 64)     A=K   (K-major)  B=Q   (K-major)
//     dP^T = V dO^T                                          A=V   (K-major)  B=dO  (K-major)
//     dV  += P^T dO                   (M=kv, N=hd 128, K=q)  A=P^T (smem)     B=dO  (MN-major)
//     dK  += dS^T Q                                          A=dS^T(smem)     B=Q   (MN-major)
//     dQ^T = K^T dS^T                 (M=hd, N=q, K=kv)      A=K   (MN-major) B=dS^T(MN-major)
// dV and dK accumulate in TMEM across the whole loop (no atomics, GQA heads included). Each
// dQ^T tile goes TMEM -> fp32 smem staging -> one cp.reduce.async.bulk (add.f32) into a
// tile-contiguous fp32 workspace; a small kernel converts the workspace to bf16 afterwards.
// P is recomputed from the saved log-sum-exp; Delta = rowsum(dO * O) comes from a pre-pass.
// =====================================================================================
namespace bz {

constexpr int FB_QT = 64;                       // query rows per inner tile
constexpr int FB_THREADS = 320;                // TMA warp, MMA warp, 8 compute warps (two threads per TMEM lane)
constexpr int FB_KV_BYTES = 128 * 128 * 2;      // 32 KB, two halves of 16 KB
constexpr int FB_KV_HALF = 128 * 64 * 2;
constexpr int FB_Q_BYTES = FB_QT * 128 * 2;     // 16 KB, two halves of 8 KB
constexpr int FB_Q_HALF = FB_QT * 64 * 2;
constexpr int FB_PT_BYTES = 128 * FB_QT * 2;    // 16 KB
constexpr int FB_STAGE_F32 = FB_QT * 128 * 4;   // 32 KB dQ staging
constexpr int FB_SMEM = 2 * FB_KV_BYTES + 4 * FB_Q_BYTES + 2 * FB_PT_BYTES + 2 * FB_STAGE_F32 + 2 * 2 * FB_QT * 4 + 256 + 1024;

struct FbBars {
  static constexpr int kv_full = 0, qdo_full = 8, qdo_empty = 24, st_full = 40, dp_full = 48, pt_ready = 56,
                       dst_ready = 64, dq_full = 72, sdp_empty = 80, dqt_empty = 88, acc_full = 96, tmem_slot = 104;
};

__global__ void __launch_bounds__(FB_THREADS, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmKV, const __grid_constant__ CUtensorMap tmQ,
                 const __grid_constant__ CUtensorMap tmdO, const float* __restrict__ lse,
                 const float* __restrict__ delta, __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dq_acc, int B,
                 int S, int Hq, int Hk, int ld, float scale, float scale_log2) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t k_s = base;
  const uint32_t v_s = base + FB_KV_BYTES;
  const uint32_t q_s = base + 2 * FB_KV_BYTES;                    // 2 stages
  const uint32_t do_s = q_s + 2 * FB_Q_BYTES;                     // 2 stages
  const uint32_t pt_s = do_s + 2 * FB_Q_BYTES;
  const uint32_t dst_s = pt_s + FB_PT_BYTES;
  const uint32_t stg_s = dst_s + FB_PT_BYTES;
  const uint32_t stat_s = stg_s + 2 * FB_STAGE_F32;               // (staging is double-buffered) [2][lse 64 | delta 64] floats
  const uint32_t bar = stat_s + 2 * 2 * FB_QT * 4;
  float* stat = reinterpret_cast<float*>(smem_raw + (stat_s - smem_u32(smem_raw)));
  float* stg = reinterpret_cast<float*>(smem_raw + (stg_s - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkv = S / 128;
  const int j = (int)(blockIdx.x % nkv);                           // small j = most query tiles = scheduled first
  const int g = (int)(blockIdx.x / nkv) % Hk;
  const int b = (int)(blockIdx.x / nkv) / Hk;
  const int rep = Hq / Hk;
  const int nqt = S / FB_QT;
  const int i0 = (j * 128) / FB_QT;                                // first query tile that sees this KV block
  const int per_head = nqt - i0;
  const int iters = rep * per_head;
  const int kvrow0 = b * S + j * 128;
  const int kcol = (Hq + g) * FA_HD, vcol = (Hq + Hk + g) * FA_HD;
  const long long T = (long long)B * S;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmdO);
    mbar_init(bar + FbBars::kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar + FbBars::qdo_full + 8 * i, 1);
      mbar_init(bar + FbBars::qdo_empty + 8 * i, 1);
    }
    mbar_init(bar + FbBars::st_full, 1);
    mbar_init(bar + FbBars::dp_full, 1);
    mbar_init(bar + FbBars::pt_ready, 256);
    mbar_init(bar + FbBars::dst_ready, 256);
    mbar_init(bar + FbBars::dq_full, 1);
    mbar_init(bar + FbBars::sdp_empty, 256);
    mbar_init(bar + FbBars::dqt_empty, 256);
    mbar_init(bar + FbBars::acc_full, 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<1>(bar + FbBars::tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(bar + FbBars::tmem_slot));
  const uint32_t tm_st = tmem, tm_dp = tmem + 64, tm_dv = tmem + 128, tm_dk = tmem + 256, tm_dq = tmem + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar + FbBars::kv_full, 2 * FB_KV_BYTES);
      tma_load_2d(&tmKV, bar + FbBars::kv_full, k_s, kcol, kvrow0);
      tma_load_2d(&tmKV, bar + FbBars::kv_full, k_s + FB_KV_HALF, kcol + 64, kvrow0);
      tma_load_2d(&tmKV, bar + FbBars::kv_full, v_s, vcol, kvrow0);
      tma_load_2d(&tmKV, bar + FbBars::kv_full, v_s + FB_KV_HALF, vcol + 64, kvrow0);
      for (int it = 0; it < iters; ++it) {
        const int st = it & 1;
        const uint32_t ph = (it >> 1) & 1;
        const int h = g * rep + it / per_head;
        const int qrow0 = b * S + (i0 + it % per_head) * FB_QT;
        mbar_wait(bar + FbBars::qdo_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + FbBars::qdo_full + 8 * st, 2 * FB_Q_BYTES);
        tma_load_2d(&tmQ, bar + FbBars::qdo_full + 8 * st, q_s + st * FB_Q_BYTES, h * FA_HD, qrow0);
        tma_load_2d(&tmQ, bar + FbBars::qdo_full + 8 * st, q_s + st * FB_Q_BYTES + FB_Q_HALF, h * FA_HD + 64, qrow0);
        tma_load_2d(&tmdO, bar + FbBars::qdo_full + 8 * st, do_s + st * FB_Q_BYTES, h * FA_HD, qrow0);
        tma_load_2d(&tmdO, bar + FbBars::qdo_full + 8 * st, do_s + st * FB_Q_BYTES + FB_Q_HALF, h * FA_HD + 64, qrow0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t id_s = umma_idesc_bf16(128, FB_QT, false, false);   // S^T, dP^T
      constexpr uint32_t id_acc = umma_idesc_bf16(128, 128, false, true);    // dV, dK
      constexpr uint32_t id_dq = umma_idesc_bf16(128, FB_QT, true, true);    // dQ^T
      mbar_wait(bar + FbBars::kv_full, 0);
      for (int it = 0; it < iters; ++it) {
        const int st = it & 1;
        const uint32_t ph = (it >> 1) & 1, par = it & 1;
        const uint32_t qb = q_s + st * FB_Q_BYTES, dob = do_s + st * FB_Q_BYTES;
        mbar_wait(bar + FbBars::qdo_full + 8 * st, ph);
        mbar_wait(bar + FbBars::sdp_empty, par ^ 1u);       // S^T / dP^T columns drained by the previous tile
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_f16<1>(tm_st, umma_smem_desc_sw128(k_s + (k >> 2) * FB_KV_HALF + (k & 3) * 32, 16, 1024),
                      umma_smem_desc_sw128(qb + (k >> 2) * FB_Q_HALF + (k & 3) * 32, 16, 1024), id_s, k != 0);
        umma_commit<1>(bar + FbBars::st_full);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_f16<1>(tm_dp, umma_smem_desc_sw128(v_s + (k >> 2) * FB_KV_HALF + (k & 3) * 32, 16, 1024),
                      umma_smem_desc_sw128(dob + (k >> 2) * FB_Q_HALF + (k & 3) * 32, 16, 1024), id_s, k != 0);
        umma_commit<1>(bar + FbBars::dp_full);
        mbar_wait(bar + FbBars::pt_ready, par);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)                          // dV += P^T dO
          umma_f16<1>(tm_dv, umma_smem_desc_sw128(pt_s + k * 32, 16, 1024),
                      umma_smem_desc_sw128(dob + k * 2048, FB_Q_HALF, 1024), id_acc, (it | k) != 0);
        mbar_wait(bar + FbBars::dst_ready, par);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)                          // dK += dS^T Q
          umma_f16<1>(tm_dk, umma_smem_desc_sw128(dst_s + k * 32, 16, 1024),
                      umma_smem_desc_sw128(qb + k * 2048, FB_Q_HALF, 1024), id_acc, (it | k) != 0);
        mbar_wait(bar + FbBars::dqt_empty, par ^ 1u);        // previous dQ^T tile read out of TMEM
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)                          // dQ^T = K^T dS^T
          umma_f16<1>(tm_dq, umma_smem_desc_sw128(k_s + k * 2048, FB_KV_HALF, 1024),
                      umma_smem_desc_sw128(dst_s + k * 2048, FB_KV_HALF, 1024), id_dq, k != 0);
        umma_commit<1>(bar + FbBars::qdo_empty + 8 * st);
        umma_commit<1>(bar + FbBars::dq_full);
      }
      umma_commit<1>(bar + FbBars::acc_full);
    }
  } else {
    const int q4 = warp & 3;
    const int hf = (warp - 2) >> 2;                          // which 32 of the 64 query columns this thread owns
    const int r = q4 * 32 + lane;                            // kv row (S^T, dP^T, dV, dK) / hd index (dQ^T)
    const int ct = (warp - 2) * 32 + lane;                   // 0..255 among the compute warps
    const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
    const int kv_pos = j * 128 + r;
    for (int it = 0; it < iters; ++it) {
      const uint32_t par = it & 1;
      const int h = g * rep + it / per_head;
      const int qi = i0 + it % per_head;
      const int q_pos0 = qi * FB_QT;
      // stage lse / delta of this query tile
      float* sl = stat + (it & 1) * 2 * FB_QT;
      if (ct < FB_QT) sl[ct] = lse[(size_t)h * T + (size_t)b * S + q_pos0 + ct] * 1.4426950408889634f;
      else if (ct < 2 * FB_QT) sl[ct] = delta[(size_t)h * T + (size_t)b * S + q_pos0 + (ct - FB_QT)];
      asm volatile("bar.sync 2, 256;" ::: "memory");
      const bool diag = q_pos0 < j * 128 + 128;
      const int c0 = hf * 32;                                // first query column of this thread
      // ---- P^T ----
      mbar_wait(bar + FbBars::st_full, par);
      tc_fence_after();
      uint32_t pk[16];                                       // 32 probabilities as packed bf16
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_st + lane_off + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float p0 = fast_exp2(__uint_as_float(v[e]) * scale_log2 - sl[c0 + e]);
          float p1 = fast_exp2(__uint_as_float(v[e + 1]) * scale_log2 - sl[c0 + e + 1]);
          if (diag) {
            if (kv_pos > q_pos0 + c0 + e) p0 = 0.f;
            if (kv_pos > q_pos0 + c0 + e + 1) p1 = 0.f;
          }
          pk[e >> 1] = pack_bf16x2(p0, p1);
        }
      }
      {
        const uint32_t rowb = pt_s + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          st_shared_v4(rowb + (((hf * 4 + i) ^ (r & 7)) * 16), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      }
      fence_proxy_async_smem();
      mbar_arrive(bar + FbBars::pt_ready);
      // ---- dS^T = P^T * (dP^T - Delta) * scale ----
      mbar_wait(bar + FbBars::dp_full, par);
      tc_fence_after();
      uint32_t dk_[16];
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_dp + lane_off + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float2 p = unpack_bf16x2(pk[e >> 1]);
          const float d0 = p.x * (__uint_as_float(v[e]) - sl[FB_QT + c0 + e]) * scale;
          const float d1 = p.y * (__uint_as_float(v[e + 1]) - sl[FB_QT + c0 + e + 1]) * scale;
          dk_[e >> 1] = pack_bf16x2(d0, d1);
        }
      }
      tc_fence_before();
      mbar_arrive(bar + FbBars::sdp_empty);
      {
        const uint32_t rowb = dst_s + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          st_shared_v4(rowb + (((hf * 4 + i) ^ (r & 7)) * 16), dk_[4 * i], dk_[4 * i + 1], dk_[4 * i + 2], dk_[4 * i + 3]);
      }
      fence_proxy_async_smem();
      mbar_arrive(bar + FbBars::dst_ready);
      // ---- dQ^T tile: TMEM -> fp32 staging [q][hd] -> bulk reduce-add into the workspace ----
      mbar_wait(bar + FbBars::dq_full, par);
      tc_fence_after();
      if (ct == 0) tma_store_wait_read<1>();                 // the reduce issued two tiles ago has read its staging buffer
      asm volatile("bar.sync 2, 256;" ::: "memory");
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_dq + lane_off + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) stg[(it & 1) * (FB_QT * 128) + (c0 + e) * 128 + r] = __uint_as_float(v[e]);
      }
      tc_fence_before();
      mbar_arrive(bar + FbBars::dqt_empty);
      fence_proxy_async_smem();
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (ct == 0) {
        float* gdst = dq_acc + ((size_t)(b * Hq + h) * nqt + qi) * (FB_QT * 128);
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gdst),
                     "r"(stg_s + (it & 1) * FB_STAGE_F32), "n"(FB_STAGE_F32)
                     : "memory");
        tma_store_commit();
      }
    }
    // ---- dV, dK epilogue: each thread converts 64 of the 128 columns of its row ----
    mbar_wait(bar + FbBars::acc_full, 0);
    tc_fence_after();
    __nv_bfloat16* vrow = dqkv + (size_t)(kvrow0 + r) * ld + vcol + hf * 64;
    __nv_bfloat16* krow = dqkv + (size_t)(kvrow0 + r) * ld + kcol + hf * 64;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      const uint32_t src = (pass == 0 ? tm_dv : tm_dk) + hf * 64;
      __nv_bfloat16* drow = pass == 0 ? vrow : krow;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(src + lane_off + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 o;
          o.x = pack_bf16x2(__uint_as_float(v[8 * i]), __uint_as_float(v[8 * i + 1]));
          o.y = pack_bf16x2(__uint_as_float(v[8 * i + 2]), __uint_as_float(v[8 * i + 3]));
          o.z = pack_bf16x2(__uint_as_float(v[8 * i + 4]), __uint_as_float(v[8 * i + 5]));
          o.w = pack_bf16x2(__uint_as_float(v[8 * i + 6]), __uint_as_float(v[8 * i + 7]));
          *reinterpret_cast<uint4*>(drow + c * 32 + i * 8) = o;
        }
      }
    }
    if (ct == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // all dQ reductions globally done
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem, 512);
  }
}

// Delta[h, t] = sum_d dO[t, h, d] * O[t, h, d]
__global__ void __launch_bounds__(256)
flash_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                   float* __restrict__ delta, long long T, int Hq) {
  // one warp per (t, h): 128 elements = 32 lanes x 4
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= T * Hq) return;
  const long long t = w / Hq;
  const int h = (int)(w - t * Hq);
  const size_t off = (size_t)t * Hq * FA_HD + (size_t)h * FA_HD + lane * 4;
  const uint2 a = *reinterpret_cast<const uint2*>(o + off);
  const uint2 d = *reinterpret_cast<const uint2*>(dout + off);
  const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), d0 = unpack_bf16x2(d.x), d1 = unpack_bf16x2(d.y);
  float s = a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y;
  s = warp_sum(s);
  if (lane == 0) delta[(size_t)h * T + t] = s;
}

// dq workspace (tile-contiguous fp32) -> bf16 q columns of dqkv
__global__ void __launch_bounds__(256)
flash_dq_convert_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int B, int S, int Hq,
                        int ld) {
  const long long total = (long long)B * S * Hq * (FA_HD / 8);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(idx % (FA_HD / 8));
    long long rest = idx / (FA_HD / 8);
    const int h = (int)(rest % Hq);
    const long long t = rest / Hq;
    const int b = (int)(t / S), s = (int)(t % S);
    const float* src = dq_acc + ((size_t)(b * Hq + h) * (S / FB_QT) + s / FB_QT) * (FB_QT * 128) + (s % FB_QT) * 128 + v * 8;
    const float4 x = *reinterpret_cast<const float4*>(src), y = *reinterpret_cast<const float4*>(src + 4);
    uint4 o;
    o.x = pack_bf16x2(x.x, x.y); o.y = pack_bf16x2(x.z, x.w);
    o.z = pack_bf16x2(y.x, y.y); o.w = pack_bf16x2(y.z, y.w);
    *reinterpret_cast<uint4*>(dqkv + (size_t)t * ld + h * FA_HD + v * 8) = o;
  }
}

}  // namespace bz

// workspace_f32: >= B*S*Hq*128 (dQ accumulator) + 2... caller passes delta separately.
extern "C" long long bz_flash_attn_bwd_workspace_floats(int B, int S, int Hq) {
  return (long long)B * S * Hq * FA_HD + (long long)B * S * Hq;   // dq accumulator + delta
}

extern "C" int bz_flash_attn_bwd(const void* qkv, const void* out, const float* lse, const void* dout, void* dqkv,
                                 float* workspace, int B, int S, int Hq, int Hk, int head_dim, float softmax_scale,
                                 void* stream) {
  BZ_REQUIRE(head_dim == FA_HD, "flash_attn_bwd: head_dim %d unsupported (128 only)", head_dim);
  BZ_REQUIRE(S % 128 == 0 && S > 0, "flash_attn_bwd: seq_len %d must be a multiple of 128", S);
  BZ_REQUIRE(Hq % Hk == 0, "flash_attn_bwd: Hq=%d not a multiple of Hkv=%d", Hq, Hk);
  const int ld = (Hq + 2 * Hk) * FA_HD;
  const long long T = (long long)B * S;
  cudaStream_t st = (cudaStream_t)stream;
  float* dq_acc = workspace;
  float* delta = workspace + T * Hq * FA_HD;
  CUtensorMap tmKV, tmQ, tmdO;
  if (int rc = make_tmap_2d_bf16(&tmKV, qkv, T, ld, ld, 128, 64)) return rc;
  if (int rc = make_tmap_2d_bf16(&tmQ, qkv, T, ld, ld, FB_QT, 64)) return rc;
  if (int rc = make_tmap_2d_bf16(&tmdO, dout, T, (uint64_t)Hq * FA_HD, (uint64_t)Hq * FA_HD, FB_QT, 64)) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FB_SMEM));
    attr_done = true;
  }
  BZ_CUDA_CHECK(cudaMemsetAsync(dq_acc, 0, sizeof(float) * T * Hq * FA_HD, st));
  {
    const long long warps = T * Hq;
    const int blocks = (int)((warps * 32 + 255) / 256);
    flash_delta_kernel<<<blocks, 256, 0, st>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, T, Hq);
    BZ_CUDA_CHECK(cudaGetLastError());
  }
  const int grid = (S / 128) * Hk * B;
  flash_bwd_kernel<<<grid, FB_THREADS, FB_SMEM, st>>>(tmKV, tmQ, tmdO, lse, delta, (__nv_bfloat16*)dqkv, dq_acc, B, S, Hq,
                                                     Hk, ld, softmax_scale, softmax_scale * 1.4426950408889634f);
  BZ_CUDA_CHECK(cudaGetLastError());
  flash_dq_convert_kernel<<<device_sm_count() * 8, 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dqkv, B, S, Hq, ld);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
