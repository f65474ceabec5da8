// K6: causal flash attention, forward, head_dim 128 with GQA, on tcgen05.
//
// Operates directly on the fused QKV activation [T, (Hq+2Hkv)*128] (RoPE already applied) through ONE TMA tensor
// map: Q, K and V tiles are different column offsets of the same matrix. One CTA = one 128-row query tile of one
// (batch, head). Roles: warp 0 TMA producer (Q once; K_j / V_j into 2-stage rings, 128B swizzle); warp 1 MMA issuer
// (S_j = Q K_j^T into double-buffered TMEM columns, O += P_j V_j with O resident in TMEM for the whole KV loop);
// warps 2..9 softmax (two threads per query row: tcgen05.ld S -> online softmax in the exp2 domain -> P as bf16 into
// swizzled smem). The running max is updated lazily: O and l are rescaled only when the max grows by more than 2^8.
//
// The dependency chain between the softmax warps and the tensor core is what this generation is about:
//   before:  exp/pack of block j+1 waited for  O += P_j V_j  (single P buffer; O must be quiet for a rescale)
//            => per KV block  ~ PV (512 tensor cycles)  +  exp phase (>= 1024 MUFU cycles)  in series
//   now:     P is double-buffered and pv_done is split per buffer, so the exp phase of block j+1 only waits
//            for P_{j-1}'s consumer (long finished) and runs WHILE the tensor core does P_j V_j; the wait for
//            the latest PV happens only on the (rare, lazy) rescale path.
//            => per KV block  ~ max(QK + PV = 1024 tensor cycles, softmax)
// Also: the score scale is folded into the exponent FFMA (max is taken on raw scores), which removes one
// multiply pass over the row. Shared memory: Q, 2xK, 2xV, 2xP tiles of 32 KB = 224 KB + 2.25 KB of
// barriers / pair-exchange — fits the 227 KB limit only because the dynamic segment is 1024-aligned.
// CTAs are ordered longest-first over the whole grid. Measured on B200 (B=2, S=4096, 32q/8kv): 844 TFLOP/s
// (first generation: 686); tensor pipe 40 % active inside a training step (profiles/r2/ncu/).
//
// The reference has no attention code (it ships no model: SURVEY.md §2.3 F2).
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int F3_HD = 128;
constexpr int F3_BM = 128;
constexpr int F3_BN = 128;
constexpr int F3_THREADS = 320;                 // TMA warp, MMA warp, 8 softmax warps (two threads per query row)
constexpr int F3_TILE = 128 * 128 * 2;          // 32 KB: two 64-column halves of 16 KB
constexpr int F3_HALF = 128 * 64 * 2;
constexpr int F3_XCH_BYTES = 2 * 2 * 128 * 4;   // pair exchange, double-buffered: [2][half][row] floats
constexpr int F3_SMEM_USED = 7 * F3_TILE + 256 + F3_XCH_BYTES;
constexpr int F3_SMEM = 232448;                 // everything the SM offers; slack covers <= 768 B of misalignment
constexpr float F3_RESCALE_THRESHOLD = 8.f;
static_assert(F3_SMEM_USED + 768 <= F3_SMEM, "v3 shared-memory budget");

struct F3Bars {
  static constexpr int q_full = 0, k_full = 8, k_empty = 24, v_full = 40, v_empty = 56, s_full = 72, s_empty = 88,
                       p_full = 104, pv_done = 120, tmem_slot = 136;   // p_full[2], pv_done[2]
};

__device__ __forceinline__ void f3_st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// 2^x on the FMA pipe (no MUFU): round-to-nearest split x = n + f, f in [-0.5, 0.5], degree-3 minimax polynomial
// for 2^f (max relative error 7.5e-5, far below bf16's 3.9e-3), n added straight into the exponent field.
// The B200 SFU does 16 ex2/clk/SM against 128 FMA lanes/clk/SM, and a 128x128 score tile needs 16 384
// exponentials = 1 024 SFU cycles — as long as the tensor core needs for the tile's two GEMMs. Sending every
// POLY-th element through this routine moves the exponentials off the critical pipe.
__device__ __forceinline__ float f3_exp2_poly(float x) {
  x = fmaxf(x, -126.f);
  const float xr = x + 12582912.f;            // 1.5 * 2^23: round(x) lands in the low mantissa bits
  const float f = x - (xr - 12582912.f);
  float p = fmaf(f, 0.05517165f, 0.24261113f);
  p = fmaf(f, p, 0.69326097f);
  p = fmaf(f, p, 0.99992806f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xr) << 23));
}

template <int POLY>   // 0: every exponential on the SFU; n > 0: elements with (index % n) == n - 1 use f3_exp2_poly
__global__ void __launch_bounds__(F3_THREADS, 1)
flash_fwd_v3_kernel(const __grid_constant__ CUtensorMap tm, __nv_bfloat16* __restrict__ out, float* __restrict__ lse,
                    int* __restrict__ error_flag, int B, int S, int Hq, int Hk, int ld_out, float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  if (base - raw > 768u) {                       // cannot happen with a 1024-aligned dynamic segment; never overrun
    if (threadIdx.x == 0 && error_flag != nullptr) atomicExch(error_flag, 1);
    return;
  }
  const uint32_t q_s = base;
  const uint32_t k_s = base + F3_TILE;           // 2 stages
  const uint32_t v_s = base + 3 * F3_TILE;       // 2 stages
  const uint32_t p_s = base + 5 * F3_TILE;       // 2 buffers
  const uint32_t bar = base + 7 * F3_TILE;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = S / F3_BM;
  // heavy tiles first over the whole grid (query tile qi visits qi + 1 KV blocks): every (head, batch) pair's last
  // query tile is scheduled before anybody's second-to-last one; the 4 query heads of a KV group stay adjacent
  const int groups = Hq * B;
  const int qi = nq - 1 - (int)(blockIdx.x / groups);
  const int h = (int)(blockIdx.x % groups) % Hq;
  const int b = (int)(blockIdx.x % groups) / Hq;
  const int g = h / (Hq / Hk);
  const int row0 = b * S + qi * F3_BM;
  const int nblk = qi + 1;                         // causal: kv blocks 0..qi
  const int qcol = h * F3_HD, kcol = (Hq + g) * F3_HD, vcol = (Hq + Hk + g) * F3_HD;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm);
    mbar_init(bar + F3Bars::q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar + F3Bars::k_full + 8 * i, 1);
      mbar_init(bar + F3Bars::k_empty + 8 * i, 1);
      mbar_init(bar + F3Bars::v_full + 8 * i, 1);
      mbar_init(bar + F3Bars::v_empty + 8 * i, 1);
      mbar_init(bar + F3Bars::s_full + 8 * i, 1);
      mbar_init(bar + F3Bars::s_empty + 8 * i, 256);
      mbar_init(bar + F3Bars::p_full + 8 * i, 256);
      mbar_init(bar + F3Bars::pv_done + 8 * i, 1);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<1>(bar + F3Bars::tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(bar + F3Bars::tmem_slot));
  const uint32_t tm_s[2] = {tmem, tmem + 128};
  const uint32_t tm_o = tmem + 256;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar + F3Bars::q_full, F3_TILE);
      tma_load_2d(&tm, bar + F3Bars::q_full, q_s, qcol, row0);
      tma_load_2d(&tm, bar + F3Bars::q_full, q_s + F3_HALF, qcol + 64, row0);
      for (int j = 0; j < nblk; ++j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        const int kv0 = b * S + j * F3_BN;
        mbar_wait(bar + F3Bars::k_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + F3Bars::k_full + 8 * st, F3_TILE);
        tma_load_2d(&tm, bar + F3Bars::k_full + 8 * st, k_s + st * F3_TILE, kcol, kv0);
        tma_load_2d(&tm, bar + F3Bars::k_full + 8 * st, k_s + st * F3_TILE + F3_HALF, kcol + 64, kv0);
        mbar_wait(bar + F3Bars::v_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + F3Bars::v_full + 8 * st, F3_TILE);
        tma_load_2d(&tm, bar + F3Bars::v_full + 8 * st, v_s + st * F3_TILE, vcol, kv0);
        tma_load_2d(&tm, bar + F3Bars::v_full + 8 * st, v_s + st * F3_TILE + F3_HALF, vcol + 64, kv0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 128, false, true);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(bar + F3Bars::k_full + 8 * st, ph);
        mbar_wait(bar + F3Bars::s_empty + 8 * st, ph ^ 1u);   // softmax finished reading S[j-2]
        tc_fence_after();
        const uint32_t kb = k_s + st * F3_TILE;
#pragma unroll
        for (int k = 0; k < F3_HD / 16; ++k) {
          const uint32_t off = (k >> 2) * F3_HALF + (k & 3) * 32;
          umma_f16<1>(tm_s[st], umma_smem_desc_sw128(q_s + off, 16, 1024), umma_smem_desc_sw128(kb + off, 16, 1024),
                      idesc_qk, k != 0 ? 1u : 0u);
        }
        umma_commit<1>(bar + F3Bars::k_empty + 8 * st);
        umma_commit<1>(bar + F3Bars::s_full + 8 * st);
      };
      mbar_wait(bar + F3Bars::q_full, 0);
      issue_qk(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_qk(j + 1);                     // overlaps softmax of block j
        const int st = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(bar + F3Bars::v_full + 8 * st, ph);
        mbar_wait(bar + F3Bars::p_full + 8 * st, ph);          // P_j sits in buffer j & 1
        tc_fence_after();
        const uint32_t vb = v_s + st * F3_TILE;
        const uint32_t pb = p_s + st * F3_TILE;
#pragma unroll
        for (int k = 0; k < F3_BN / 16; ++k) {
          const uint32_t aoff = (k >> 2) * F3_HALF + (k & 3) * 32;
          umma_f16<1>(tm_o, umma_smem_desc_sw128(pb + aoff, 16, 1024),
                      umma_smem_desc_sw128(vb + k * 2048, F3_HALF, 1024), idesc_pv, (j | k) != 0 ? 1u : 0u);
        }
        umma_commit<1>(bar + F3Bars::v_empty + 8 * st);
        umma_commit<1>(bar + F3Bars::pv_done + 8 * st);        // buffer j & 1 free again; O includes block j
      }
    }
  } else {
    // ---------------- softmax: two threads per query row, each owning 64 of the 128 kv columns ----------------
    const int q = warp & 3;                       // TMEM lane quadrant of this warp
    const int hf = (warp - 2) >> 2;               // which half of the columns (warps 2-5: 0, warps 6-9: 1)
    const int r = q * 32 + lane;                  // query row inside the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    float* xch = reinterpret_cast<float*>(smem_raw + (bar + 256 - raw));  // [2][2][128]
    float m_used = -INFINITY;                     // exponent reference baked into O and l (log2 domain)
    float l = 0.f;                                // this thread's share of the row sum
    for (int j = 0; j < nblk; ++j) {
      const int st = j & 1;
      const uint32_t ph = (j >> 1) & 1;
      const bool diag = (j == nblk - 1);
      mbar_wait(bar + F3Bars::s_full + 8 * st, ph);
      tc_fence_after();
      uint32_t sv[64];
      tmem_ld_32x64(tm_s[st] + lane_off + hf * 64, sv);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(bar + F3Bars::s_empty + 8 * st);          // S[st] is in registers: block j+2 may overwrite it
      float mx = -INFINITY;                                  // max of the RAW scores (scale > 0 commutes with max)
      if (diag) {
#pragma unroll
        for (int e = 0; e < 64; ++e) {
          if ((hf * 64 + e) > r) sv[e] = 0xff800000u;        // -inf
          mx = fmaxf(mx, __uint_as_float(sv[e]));
        }
      } else {
#pragma unroll
        for (int e = 0; e < 64; ++e) mx = fmaxf(mx, __uint_as_float(sv[e]));
      }
      // combine the two halves of the row
      float* xm = xch + st * 256;
      xm[hf * 128 + r] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
      mx = fmaxf(mx, xm[(hf ^ 1) * 128 + r]);
      const float m_new = fmaxf(m_used, mx * scale_log2);
      const bool need = (m_new - m_used) > F3_RESCALE_THRESHOLD;     // lazy rescale (always true on block 0)
      // P buffer `st` was last read by PV(j-2): long finished in steady state
      if (j >= 2) mbar_wait(bar + F3Bars::pv_done + 8 * st, ph ^ 1u);
      if (__any_sync(0xffffffffu, need) && j > 0) {
        // rare path: O must be quiet, i.e. every PV up to block j-1 has completed
        mbar_wait(bar + F3Bars::pv_done + 8 * ((j - 1) & 1), (uint32_t)(((j - 1) >> 1) & 1));
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
      const float neg_m = -m_used;
      float rs = 0.f;
      uint32_t pk[32];
#pragma unroll
      for (int e = 0; e < 64; e += 2) {
        const float x0 = fmaf(__uint_as_float(sv[e]), scale_log2, neg_m);
        const float x1 = fmaf(__uint_as_float(sv[e + 1]), scale_log2, neg_m);
        const float p0 = (POLY > 0 && (e % POLY) == POLY - 1) ? f3_exp2_poly(x0) : fast_exp2(x0);
        const float p1 = (POLY > 0 && ((e + 1) % POLY) == POLY - 1) ? f3_exp2_poly(x1) : fast_exp2(x1);
        rs += p0 + p1;
        pk[e >> 1] = pack_bf16x2(p0, p1);
      }
      l += rs;
      const uint32_t rowb = p_s + st * F3_TILE + hf * F3_HALF + r * 128;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        f3_st_shared_v4(rowb + ((i ^ (r & 7)) * 16), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      fence_proxy_async_smem();                      // make P visible to the tensor core (async proxy)
      mbar_arrive(bar + F3Bars::p_full + 8 * st);
    }
    // epilogue: combine l, O / l -> bf16, LSE. The exchange buffer not used by the last iteration is free.
    float* xl = xch + (nblk & 1) * 256;
    xl[hf * 128 + r] = l;
    asm volatile("bar.sync %0, 64;" ::"r"(2 + q) : "memory");
    l += xl[(hf ^ 1) * 128 + r];
    mbar_wait(bar + F3Bars::pv_done + 8 * ((nblk - 1) & 1), (uint32_t)(((nblk - 1) >> 1) & 1));
    tc_fence_after();
    const float inv_l = 1.f / l;
    __nv_bfloat16* orow = out + (size_t)(row0 + r) * ld_out + h * F3_HD + hf * 64;
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

// Same contract as bz_flash_attn_fwd. error_flag (device int, may be NULL) is set to 1 if the kernel refused
// to run because its shared-memory segment was not aligned as required. poly_stride: 0 = all exponentials on
// the SFU; 2..8 = every poly_stride-th one on the FMA pipe (4 balances the two pipes on B200).
extern "C" int bz_flash_attn_fwd_v3(const void* qkv, void* out, float* lse, int* error_flag, int B, int S, int Hq,
                                    int Hk, int head_dim, float softmax_scale, int poly_stride, void* stream) {
  BZ_REQUIRE(head_dim == F3_HD, "flash_attn_fwd_v3: head_dim %d unsupported (128 only)", head_dim);
  BZ_REQUIRE(S % F3_BM == 0 && S > 0, "flash_attn_fwd_v3: seq_len %d must be a multiple of %d", S, F3_BM);
  BZ_REQUIRE(Hq % Hk == 0, "flash_attn_fwd_v3: Hq=%d not a multiple of Hkv=%d", Hq, Hk);
  BZ_REQUIRE(poly_stride == 0 || poly_stride == 2 || poly_stride == 4 || poly_stride == 8,
             "flash_attn_fwd_v3: poly_stride %d not in {0, 2, 4, 8}", poly_stride);
  const int ld = (Hq + 2 * Hk) * F3_HD;
  CUtensorMap tm;
  if (int rc = make_tmap_2d_bf16(&tm, qkv, (uint64_t)B * S, ld, ld, 128, 64)) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_fwd_v3_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_fwd_v3_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_fwd_v3_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_fwd_v3_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SMEM));
    attr_done = true;
  }
  const int grid = (S / F3_BM) * Hq * B;
  const float scale_log2 = softmax_scale * 1.4426950408889634f;
  cudaStream_t st = (cudaStream_t)stream;
  __nv_bfloat16* o = (__nv_bfloat16*)out;
  switch (poly_stride) {
    case 0: flash_fwd_v3_kernel<0><<<grid, F3_THREADS, F3_SMEM, st>>>(tm, o, lse, error_flag, B, S, Hq, Hk, Hq * F3_HD, scale_log2); break;
    case 2: flash_fwd_v3_kernel<2><<<grid, F3_THREADS, F3_SMEM, st>>>(tm, o, lse, error_flag, B, S, Hq, Hk, Hq * F3_HD, scale_log2); break;
    case 4: flash_fwd_v3_kernel<4><<<grid, F3_THREADS, F3_SMEM, st>>>(tm, o, lse, error_flag, B, S, Hq, Hk, Hq * F3_HD, scale_log2); break;
    default: flash_fwd_v3_kernel<8><<<grid, F3_THREADS, F3_SMEM, st>>>(tm, o, lse, error_flag, B, S, Hq, Hk, Hq * F3_HD, scale_log2); break;
  }
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
