// K6: causal flash attention, backward, head_dim 128 with GQA, on tcgen05. One CTA = one 128-row KV
// block of one (batch, kv head), looping over the query tiles (64 rows) of its query heads: five tcgen05 GEMMs on the
// transposed score tile, dK/dV resident in TMEM, dQ^T through fp32 staging + cp.reduce.async.bulk. This generation's
// subject is the schedule (the first generation, removed, ran the steps one after another: 508 TFLOP/s; this: 712).
//
//   v2, per (KV block, 64-row query tile):   S^T, dP^T | P | dV | dS | dK, dQ^T | stage dQ      — one after another:
//        the tensor core idles while the compute warps work and vice versa (measured: tensor pipe 31 % active,
//        about 4 600 cycles per tile for 1 280 cycles of tensor work).
//   v3:  tensor pipe   ... dV(i)  S^T(i+1) | dP^T(i+1)  dK(i)  dQ^T(i) | dV(i+1) ...
//        compute warps ... P(i) | dS(i) | stage dQ(i-1) | P(i+1) ...
//        S^T of the next tile is issued right after dV of the current one (double-buffered S^T columns), dP^T of the
//        next tile right after the compute warps released the dP^T columns, and the compute warps read dQ^T of tile
//        i-1 out of TMEM only after they have produced dS(i) — so each side always has work that does not depend on
//        what the other side is doing at that moment.
//
// Resources: TMEM  S^T[2] 2x64 | dP^T 64 | dQ^T 64 | dV 128 | dK 128 = 512 columns.
//            smem  K, V 64 KB | Q, dO 3 stages 96 KB (S^T(i+1) is wanted a tile earlier than in v2) | P^T, dS^T 32 KB |
//                  one dQ staging buffer 32 KB = 224 KB, plus 1.25 KB of statistics and barriers.
// Single-buffered resources and what guards them: P^T smem — the in-order tensor pipe (S^T(i+1) is issued after
// dV(i), so st_full(i+1) implies dV(i) has read it); dS^T smem — dq_full(i) before dS(i+1) is stored; the dP^T
// columns — dP^T(i+1) is issued after dst_ready(i); the dQ^T columns — dqt_empty(i-1) before dQ^T(i) is issued.
// tests/test_flash_bwd_v3_protocol_model.py executes exactly this protocol with hardware mbarrier semantics.
//
// The reference has no attention code (it ships no model: SURVEY.md §2.3 F2).
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int B3_HD = 128;
constexpr int B3_QT = 64;                        // query rows per inner tile
constexpr int B3_THREADS = 320;                  // TMA warp, MMA warp, 8 compute warps (two threads per TMEM lane)
constexpr int B3_KV_BYTES = 128 * 128 * 2;       // 32 KB, two halves of 16 KB
constexpr int B3_KV_HALF = 128 * 64 * 2;
constexpr int B3_Q_BYTES = B3_QT * 128 * 2;      // 16 KB, two halves of 8 KB
constexpr int B3_Q_HALF = B3_QT * 64 * 2;
constexpr int B3_PT_BYTES = 128 * B3_QT * 2;     // 16 KB
constexpr int B3_STAGE_F32 = B3_QT * 128 * 4;    // 32 KB dQ staging
constexpr int B3_NSTAGE = 3;
constexpr int B3_SMEM_USED = 2 * B3_KV_BYTES + 2 * B3_NSTAGE * B3_Q_BYTES + 2 * B3_PT_BYTES + B3_STAGE_F32 +
                             2 * 2 * B3_QT * 4 + 256;
constexpr int B3_SMEM = 232448;
static_assert(B3_SMEM_USED + 768 <= B3_SMEM, "bwd v3 shared-memory budget");

struct B3Bars {
  static constexpr int kv_full = 0, qdo_full = 8, qdo_empty = 32, st_full = 56, dp_full = 72, pt_ready = 80,
                       dst_ready = 88, dq_full = 96, dqt_empty = 104, acc_full = 112, tmem_slot = 120;
};

__device__ __forceinline__ void b3_st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__global__ void __launch_bounds__(B3_THREADS, 1)
flash_bwd_v3_kernel(const __grid_constant__ CUtensorMap tmKV, const __grid_constant__ CUtensorMap tmQ,
                    const __grid_constant__ CUtensorMap tmdO, const float* __restrict__ lse,
                    const float* __restrict__ delta, __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dq_acc,
                    int* __restrict__ error_flag, int B, int S, int Hq, int Hk, int ld, float scale,
                    float scale_log2) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  if (base - raw > 768u) {                        // cannot happen with a 1024-aligned dynamic segment; never overrun
    if (threadIdx.x == 0 && error_flag != nullptr) atomicExch(error_flag, 1);
    return;
  }
  const uint32_t k_s = base;
  const uint32_t v_s = base + B3_KV_BYTES;
  const uint32_t q_s = base + 2 * B3_KV_BYTES;                    // 3 stages
  const uint32_t do_s = q_s + B3_NSTAGE * B3_Q_BYTES;             // 3 stages
  const uint32_t pt_s = do_s + B3_NSTAGE * B3_Q_BYTES;
  const uint32_t dst_s = pt_s + B3_PT_BYTES;
  const uint32_t stg_s = dst_s + B3_PT_BYTES;
  const uint32_t stat_s = stg_s + B3_STAGE_F32;                   // [2][lse 64 | delta 64] floats
  const uint32_t bar = stat_s + 2 * 2 * B3_QT * 4;
  float* stat = reinterpret_cast<float*>(smem_raw + (stat_s - raw));
  float* stg = reinterpret_cast<float*>(smem_raw + (stg_s - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkv = S / 128;
  // Longest work first over the WHOLE grid: a KV block j is visited by (S/64 - 2j)·rep query tiles, so all CTAs of
  // j = 0 come first, then j = 1, ... Block indices are handed to free SMs in order; with j fastest-varying the last
  // (head, batch) group's j = 0 CTA started late and ran alone at the end (list-scheduling model at S = 4096,
  // 8 kv heads, batch 2: 626 tile-times against 498 in this order, 457 ideal).
  const int groups = Hk * B;
  const int j = (int)(blockIdx.x / groups);
  const int g = (int)(blockIdx.x % groups) % Hk;
  const int b = (int)(blockIdx.x % groups) / Hk;
  const int rep = Hq / Hk;
  const int nqt = S / B3_QT;
  const int i0 = (j * 128) / B3_QT;                                // first query tile that sees this KV block
  const int per_head = nqt - i0;
  const int iters = rep * per_head;
  const int kvrow0 = b * S + j * 128;
  const int kcol = (Hq + g) * B3_HD, vcol = (Hq + Hk + g) * B3_HD;
  const long long T = (long long)B * S;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmKV);
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmdO);
    mbar_init(bar + B3Bars::kv_full, 1);
    for (int i = 0; i < B3_NSTAGE; ++i) {
      mbar_init(bar + B3Bars::qdo_full + 8 * i, 1);
      mbar_init(bar + B3Bars::qdo_empty + 8 * i, 1);
    }
    mbar_init(bar + B3Bars::st_full, 1);
    mbar_init(bar + B3Bars::st_full + 8, 1);
    mbar_init(bar + B3Bars::dp_full, 1);
    mbar_init(bar + B3Bars::pt_ready, 256);
    mbar_init(bar + B3Bars::dst_ready, 256);
    mbar_init(bar + B3Bars::dq_full, 1);
    mbar_init(bar + B3Bars::dqt_empty, 256);
    mbar_init(bar + B3Bars::acc_full, 1);
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<1>(bar + B3Bars::tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(bar + B3Bars::tmem_slot));
  const uint32_t tm_st[2] = {tmem, tmem + 64};
  const uint32_t tm_dp = tmem + 128, tm_dq = tmem + 192, tm_dv = tmem + 256, tm_dk = tmem + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(bar + B3Bars::kv_full, 2 * B3_KV_BYTES);
      tma_load_2d(&tmKV, bar + B3Bars::kv_full, k_s, kcol, kvrow0);
      tma_load_2d(&tmKV, bar + B3Bars::kv_full, k_s + B3_KV_HALF, kcol + 64, kvrow0);
      tma_load_2d(&tmKV, bar + B3Bars::kv_full, v_s, vcol, kvrow0);
      tma_load_2d(&tmKV, bar + B3Bars::kv_full, v_s + B3_KV_HALF, vcol + 64, kvrow0);
      for (int it = 0; it < iters; ++it) {
        const int st = it % B3_NSTAGE;
        const uint32_t ph = (uint32_t)(it / B3_NSTAGE) & 1u;
        const int h = g * rep + it / per_head;
        const int qrow0 = b * S + (i0 + it % per_head) * B3_QT;
        mbar_wait(bar + B3Bars::qdo_empty + 8 * st, ph ^ 1u);
        mbar_expect_tx(bar + B3Bars::qdo_full + 8 * st, 2 * B3_Q_BYTES);
        tma_load_2d(&tmQ, bar + B3Bars::qdo_full + 8 * st, q_s + st * B3_Q_BYTES, h * B3_HD, qrow0);
        tma_load_2d(&tmQ, bar + B3Bars::qdo_full + 8 * st, q_s + st * B3_Q_BYTES + B3_Q_HALF, h * B3_HD + 64, qrow0);
        tma_load_2d(&tmdO, bar + B3Bars::qdo_full + 8 * st, do_s + st * B3_Q_BYTES, h * B3_HD, qrow0);
        tma_load_2d(&tmdO, bar + B3Bars::qdo_full + 8 * st, do_s + st * B3_Q_BYTES + B3_Q_HALF, h * B3_HD + 64, qrow0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t id_s = umma_idesc_bf16(128, B3_QT, false, false);   // S^T, dP^T
      constexpr uint32_t id_acc = umma_idesc_bf16(128, 128, false, true);    // dV, dK
      constexpr uint32_t id_dq = umma_idesc_bf16(128, B3_QT, true, true);    // dQ^T
      auto issue_st = [&](int it) {                                          // S^T(it) = K Q^T -> ST[it & 1]
        const uint32_t qb = q_s + (it % B3_NSTAGE) * B3_Q_BYTES;
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_f16<1>(tm_st[it & 1], umma_smem_desc_sw128(k_s + (k >> 2) * B3_KV_HALF + (k & 3) * 32, 16, 1024),
                      umma_smem_desc_sw128(qb + (k >> 2) * B3_Q_HALF + (k & 3) * 32, 16, 1024), id_s, k != 0);
        umma_commit<1>(bar + B3Bars::st_full + 8 * (it & 1));
      };
      auto issue_dp = [&](int it) {                                          // dP^T(it) = V dO^T -> DP
        const uint32_t dob = do_s + (it % B3_NSTAGE) * B3_Q_BYTES;
#pragma unroll
        for (int k = 0; k < 8; ++k)
          umma_f16<1>(tm_dp, umma_smem_desc_sw128(v_s + (k >> 2) * B3_KV_HALF + (k & 3) * 32, 16, 1024),
                      umma_smem_desc_sw128(dob + (k >> 2) * B3_Q_HALF + (k & 3) * 32, 16, 1024), id_s, k != 0);
        umma_commit<1>(bar + B3Bars::dp_full);
      };
      mbar_wait(bar + B3Bars::kv_full, 0);
      mbar_wait(bar + B3Bars::qdo_full, 0);
      tc_fence_after();
      issue_st(0);
      issue_dp(0);
      for (int it = 0; it < iters; ++it) {
        const int st = it % B3_NSTAGE;
        const uint32_t par = it & 1;
        const uint32_t qb = q_s + st * B3_Q_BYTES, dob = do_s + st * B3_Q_BYTES;
        mbar_wait(bar + B3Bars::pt_ready, par);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)                          // dV += P^T dO
          umma_f16<1>(tm_dv, umma_smem_desc_sw128(pt_s + k * 32, 16, 1024),
                      umma_smem_desc_sw128(dob + k * 2048, B3_Q_HALF, 1024), id_acc, (it | k) != 0);
        if (it + 1 < iters) {                                // S^T of the next tile: its columns were drained before
          const int st1 = (it + 1) % B3_NSTAGE;              // pt_ready(it-1), its P^T consumer is the dV above
          mbar_wait(bar + B3Bars::qdo_full + 8 * st1, (uint32_t)((it + 1) / B3_NSTAGE) & 1u);
          tc_fence_after();
          issue_st(it + 1);
        }
        mbar_wait(bar + B3Bars::dst_ready, par);             // dS^T(it) stored; the dP^T columns have been read
        tc_fence_after();
        if (it + 1 < iters) issue_dp(it + 1);
#pragma unroll
        for (int k = 0; k < 4; ++k)                          // dK += dS^T Q
          umma_f16<1>(tm_dk, umma_smem_desc_sw128(dst_s + k * 32, 16, 1024),
                      umma_smem_desc_sw128(qb + k * 2048, B3_Q_HALF, 1024), id_acc, (it | k) != 0);
        mbar_wait(bar + B3Bars::dqt_empty, par ^ 1u);        // dQ^T(it-1) has been read out of its columns
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < 8; ++k)                          // dQ^T = K^T dS^T
          umma_f16<1>(tm_dq, umma_smem_desc_sw128(k_s + k * 2048, B3_KV_HALF, 1024),
                      umma_smem_desc_sw128(dst_s + k * 2048, B3_KV_HALF, 1024), id_dq, k != 0);
        umma_commit<1>(bar + B3Bars::qdo_empty + 8 * st);
        umma_commit<1>(bar + B3Bars::dq_full);
      }
      umma_commit<1>(bar + B3Bars::acc_full);
    }
  } else {
    const int q4 = warp & 3;
    const int hf = (warp - 2) >> 2;                          // which 32 of the 64 query columns this thread owns
    const int r = q4 * 32 + lane;                            // kv row (S^T, dP^T, dV, dK) / hd index (dQ^T)
    const int ct = (warp - 2) * 32 + lane;                   // 0..255 among the compute warps
    const uint32_t lane_off = (uint32_t)(q4 * 32) << 16;
    const int kv_pos = j * 128 + r;
    const int c0 = hf * 32;                                  // first query column of this thread

    // dQ^T of tile k: TMEM -> fp32 staging [q][hd] -> one bulk reduce-add into the workspace. The caller has
    // waited dq_full for tile k (and issued the tcgen05 fence after that wait).
    auto stage_dq = [&](int k) {
      const int h = g * rep + k / per_head;
      const int qi = i0 + k % per_head;
      if (ct == 0) tma_store_wait_read<0>();                 // the previous reduce has read the staging buffer
      asm volatile("bar.sync 2, 256;" ::: "memory");
      tc_fence_after();
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_dq + lane_off + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) stg[(c0 + e) * 128 + r] = __uint_as_float(v[e]);
      }
      tc_fence_before();
      mbar_arrive(bar + B3Bars::dqt_empty);
      fence_proxy_async_smem();
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (ct == 0) {
        float* gdst = dq_acc + ((size_t)(b * Hq + h) * nqt + qi) * (B3_QT * 128);
        asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(gdst),
                     "r"(stg_s), "n"(B3_STAGE_F32)
                     : "memory");
        tma_store_commit();
      }
    };

    // lse (threads 0..63) / delta (threads 64..127) of a query tile: read from global one tile ahead, so the load's
    // latency is hidden behind the previous tile instead of sitting in front of the barrier below (ncu: the load and
    // that barrier were 15 % of all stall samples)
    auto load_stat = [&](int k) -> float {
      const int hh = g * rep + k / per_head;
      const size_t base_t = (size_t)hh * T + (size_t)b * S + (size_t)(i0 + k % per_head) * B3_QT;
      if (ct < B3_QT) return lse[base_t + ct] * 1.4426950408889634f;
      if (ct < 2 * B3_QT) return delta[base_t + (ct - B3_QT)];
      return 0.f;
    };
    float stat_next = load_stat(0);
    for (int it = 0; it < iters; ++it) {
      const int s = it & 1;
      const uint32_t ph = (it >> 1) & 1, par = it & 1;
      const int qi = i0 + it % per_head;
      const int q_pos0 = qi * B3_QT;
      // stage lse / delta of this query tile (the other half of `stat` may still be read by slower threads of tile it-1)
      float* sl = stat + s * 2 * B3_QT;
      if (ct < 2 * B3_QT) sl[ct] = stat_next;
      asm volatile("bar.sync 2, 256;" ::: "memory");
      if (it + 1 < iters) stat_next = load_stat(it + 1);
      const bool diag = q_pos0 < j * 128 + 128;
      // ---- P^T ----
      mbar_wait(bar + B3Bars::st_full + 8 * s, ph);
      tc_fence_after();
      uint32_t pk[16];                                       // 32 probabilities as packed bf16
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_st[s] + lane_off + c0, v);
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
        // P^T smem is free: S^T(it) was issued behind dV(it-1) on the in-order tensor pipe
        const uint32_t rowb = pt_s + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          b3_st_shared_v4(rowb + (((hf * 4 + i) ^ (r & 7)) * 16), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar + B3Bars::pt_ready);
      // ---- dS^T = P^T * (dP^T - Delta) * scale ----
      mbar_wait(bar + B3Bars::dp_full, par);
      tc_fence_after();
      uint32_t dk_[16];
      {
        uint32_t v[32];
        tmem_ld_32x32(tm_dp + lane_off + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const float2 p = unpack_bf16x2(pk[e >> 1]);
          const float d0 = p.x * (__uint_as_float(v[e]) - sl[B3_QT + c0 + e]) * scale;
          const float d1 = p.y * (__uint_as_float(v[e + 1]) - sl[B3_QT + c0 + e + 1]) * scale;
          dk_[e >> 1] = pack_bf16x2(d0, d1);
        }
      }
      if (it >= 1) {                                         // dS^T smem free: dK(it-1) and dQ^T(it-1) have read it
        mbar_wait(bar + B3Bars::dq_full, par ^ 1u);
        tc_fence_after();
      }
      {
        const uint32_t rowb = dst_s + r * 128;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          b3_st_shared_v4(rowb + (((hf * 4 + i) ^ (r & 7)) * 16), dk_[4 * i], dk_[4 * i + 1], dk_[4 * i + 2], dk_[4 * i + 3]);
      }
      tc_fence_before();
      fence_proxy_async_smem();
      mbar_arrive(bar + B3Bars::dst_ready);
      // ---- dQ^T of the PREVIOUS tile (its GEMM finished while this tile's P / dS were computed) ----
      if (it >= 1) stage_dq(it - 1);
    }
    mbar_wait(bar + B3Bars::dq_full, (uint32_t)((iters - 1) & 1));
    tc_fence_after();
    stage_dq(iters - 1);
    // ---- dV, dK epilogue: each thread converts 64 of the 128 columns of its row ----
    mbar_wait(bar + B3Bars::acc_full, 0);
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

// Delta[h, t] = sum_d dO[t, h, d] * O[t, h, d]   (one warp per (t, h))
__global__ void __launch_bounds__(256)
b3_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout, float* __restrict__ delta,
                long long T, int Hq) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= T * Hq) return;
  const long long t = w / Hq;
  const int h = (int)(w - t * Hq);
  const size_t off = (size_t)t * Hq * B3_HD + (size_t)h * B3_HD + lane * 4;
  const uint2 a = *reinterpret_cast<const uint2*>(o + off);
  const uint2 d = *reinterpret_cast<const uint2*>(dout + off);
  const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), d0 = unpack_bf16x2(d.x), d1 = unpack_bf16x2(d.y);
  float s = a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y;
  s = warp_sum(s);
  if (lane == 0) delta[(size_t)h * T + t] = s;
}

// dq workspace (tile-contiguous fp32) -> bf16 q columns of dqkv
__global__ void __launch_bounds__(256)
b3_dq_convert_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, int B, int S, int Hq, int ld) {
  const long long total = (long long)B * S * Hq * (B3_HD / 8);
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(idx % (B3_HD / 8));
    long long rest = idx / (B3_HD / 8);
    const int h = (int)(rest % Hq);
    const long long t = rest / Hq;
    const int b = (int)(t / S), s = (int)(t % S);
    const float* src = dq_acc + ((size_t)(b * Hq + h) * (S / B3_QT) + s / B3_QT) * (B3_QT * 128) + (s % B3_QT) * 128 + v * 8;
    const float4 x = *reinterpret_cast<const float4*>(src), y = *reinterpret_cast<const float4*>(src + 4);
    uint4 o;
    o.x = pack_bf16x2(x.x, x.y); o.y = pack_bf16x2(x.z, x.w);
    o.z = pack_bf16x2(y.x, y.y); o.w = pack_bf16x2(y.z, y.w);
    *reinterpret_cast<uint4*>(dqkv + (size_t)t * ld + h * B3_HD + v * 8) = o;
  }
}

}  // namespace bz

using namespace bz;

// Same contract and workspace (bz_flash_attn_bwd_workspace_floats) as bz_flash_attn_bwd. error_flag (device int,
// may be NULL) becomes 1 if the kernel declined to run because its shared-memory segment was not 1024-aligned.
extern "C" int bz_flash_attn_bwd_v3(const void* qkv, const void* out, const float* lse, const void* dout, void* dqkv,
                                    float* workspace, int* error_flag, int B, int S, int Hq, int Hk, int head_dim,
                                    float softmax_scale, void* stream) {
  BZ_REQUIRE(head_dim == B3_HD, "flash_attn_bwd_v3: head_dim %d unsupported (128 only)", head_dim);
  BZ_REQUIRE(S % 128 == 0 && S > 0, "flash_attn_bwd_v3: seq_len %d must be a multiple of 128", S);
  BZ_REQUIRE(Hq % Hk == 0, "flash_attn_bwd_v3: Hq=%d not a multiple of Hkv=%d", Hq, Hk);
  const int ld = (Hq + 2 * Hk) * B3_HD;
  const long long T = (long long)B * S;
  cudaStream_t st = (cudaStream_t)stream;
  float* dq_acc = workspace;
  float* delta = workspace + T * Hq * B3_HD;
  CUtensorMap tmKV, tmQ, tmdO;
  if (int rc = make_tmap_2d_bf16(&tmKV, qkv, T, ld, ld, 128, 64)) return rc;
  if (int rc = make_tmap_2d_bf16(&tmQ, qkv, T, ld, ld, B3_QT, 64)) return rc;
  if (int rc = make_tmap_2d_bf16(&tmdO, dout, T, (uint64_t)Hq * B3_HD, (uint64_t)Hq * B3_HD, B3_QT, 64)) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(flash_bwd_v3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B3_SMEM));
    attr_done = true;
  }
  BZ_CUDA_CHECK(cudaMemsetAsync(dq_acc, 0, sizeof(float) * T * Hq * B3_HD, st));
  {
    const long long warps = T * Hq;
    const int blocks = (int)((warps * 32 + 255) / 256);
    b3_delta_kernel<<<blocks, 256, 0, st>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, T, Hq);
    BZ_CUDA_CHECK(cudaGetLastError());
  }
  const int grid = (S / 128) * Hk * B;
  flash_bwd_v3_kernel<<<grid, B3_THREADS, B3_SMEM, st>>>(tmKV, tmQ, tmdO, lse, delta, (__nv_bfloat16*)dqkv, dq_acc,
                                                        error_flag, B, S, Hq, Hk, ld, softmax_scale,
                                                        softmax_scale * 1.4426950408889634f);
  BZ_CUDA_CHECK(cudaGetLastError());
  b3_dq_convert_kernel<<<device_sm_count() * 8, 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dqkv, B, S, Hq, ld);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
