// K2, dgrad flavour (`ag_gemm_dgrad`): all-gather -> GEMM in ONE kernel for the backward
// data gradient of a ZeRO-3 row-partitioned linear layer:
//
//   dX[T,K] = dY[T,N] · W[N,K]        rank p holds rows [p·N/ws, (p+1)·N/ws) of W in its peer-mapped shard.
//
// Here the partitioned dimension N is the REDUCTION dimension of the GEMM, so a gathered chunk (256 rows of W)
// is a k-slice that every output tile needs. The communication warps of all CTAs pull the chunks exactly as in
// ag_gemm.cu (own rows first, then rank+1, ... — write-through to the local gathered buffer, release counter per
// chunk); the TMA producer walks the k-loop in the same rotated chunk order and acquires a chunk's counter before
// loading its four k-blocks, so the first wave of tiles consumes W as it lands and later tiles find it in L2.
// B (= W) is consumed MN-major straight from the row-major weight: no transposed copy of W exists anywhere.
//
// Each weight is used exactly once in backward (by its dgrad) when nothing is recomputed, so with this kernel the
// backward pass needs no bulk gather and no prefetch stream. Validated on B200 (emulated ranks and inside the
// engine on 2 GPUs); end to end it measured 1–3 % slower than prefetched pull gathers + plain GEMMs (2 and 8 GPUs), so
// the engine uses it only with B200ZERO_AG_DGRAD=1.
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int AG_BM = 128, AG_BN = 256, AG_BK = 64, AG_UK = 16;
constexpr int AG_GEMM_WARPS = 6;
constexpr int AG_COMM_WARPS = 4;
constexpr int AG_THREADS = (AG_GEMM_WARPS + AG_COMM_WARPS) * 32;
constexpr int AG_MAX_WS = 8;

template <int CG>
struct AdCfg {
  static constexpr int LOAD_N = AG_BN / CG;
  static constexpr int A_BYTES = AG_BM * AG_BK * 2;
  static constexpr int B_BYTES = LOAD_N * AG_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (CG == 1) ? 4 : 6;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 8 * (2 * STAGES + 4) + 16 + 1024;
};

struct AdPeers {
  const __nv_bfloat16* shard[AG_MAX_WS];  // rank p's rows of W (peer-mapped), index = rank
};

__device__ __forceinline__ uint4 ad_ld_peer(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ad_ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// chunk visiting order: own rows first, then rank+1, rank+2, ... (position -> chunk index)
__device__ __forceinline__ int ad_chunk_of_position(int pos, int n_tiles, int first_tile) {
  int t = first_tile + pos;
  return t >= n_tiles ? t - n_tiles : t;
}

template <int CG>
__global__ void __launch_bounds__(AG_THREADS, 1)
ag_gemm_dgrad_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ AdPeers peers, __nv_bfloat16* __restrict__ Wfull,
               __nv_bfloat16* __restrict__ Cp, uint32_t* __restrict__ chunk_ready, int M, int NW, int KW, int ldc,
               int rank, int ws) {
  // GEMM view: output columns N = KW (columns of W), reduction K = NW (rows of W, the partitioned dimension)
  const int N = KW, K = NW;
  using Cfg = AdCfg<CG>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_base = smem_base;
  const uint32_t b_base = smem_base + STAGES * Cfg::A_BYTES;
  const uint32_t bar_base = smem_base + STAGES * Cfg::STAGE_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int i) { return bar_base + 8u * (2 * STAGES + i); };
  auto tempty_bar = [&](int i) { return bar_base + 8u * (2 * STAGES + 2 + i); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 1) ? 0u : cluster_ctarank();
  const bool is_leader = (cta_rank == 0);
  const int m_tiles = (M + AG_BM * CG - 1) / (AG_BM * CG);
  const int n_tiles = (N + AG_BN - 1) / AG_BN;
  const int total_tiles = m_tiles * n_tiles;
  const int num_kb = (K + AG_BK - 1) / AG_BK;
  const int cluster_id = blockIdx.x / CG;
  const int num_clusters = gridDim.x / CG;
  // chunks of AG_BN rows of W; the first chunk whose rows are (mostly) ours
  constexpr int KB_PER_CHUNK = AG_BN / AG_BK;
  const int n_chunks = (NW + AG_BN - 1) / AG_BN;
  const long long rows_per_rank = NW / ws;
  const int first_chunk = (int)(((long long)rank * rows_per_rank) / AG_BN);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull_bar(i), 1);
      mbar_init(tempty_bar(i), 4 * CG);
    }
    fence_barrier_init();
  } else if (warp == 1) {
    tmem_alloc<CG>(tmem_slot, 512);
  }
  tc_fence_before();
  __syncwarp();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp >= AG_GEMM_WARPS) {
    // =========================== communication warps ===========================
    // All CTAs split every chunk evenly; a chunk is ready when gridDim.x slices have arrived.
    const int ct = (warp - AG_GEMM_WARPS) * 32 + lane;         // 0..127 inside the CTA
    constexpr int CT = AG_COMM_WARPS * 32;
    const long long n_shard = rows_per_rank * KW;              // elements per rank
    const long long total_elems = (long long)NW * KW;
    for (int pos = 0; pos < n_chunks; ++pos) {
      const int tile = ad_chunk_of_position(pos, n_chunks, first_chunk);
      const long long e0 = (long long)tile * AG_BN * KW;
      long long e1 = e0 + (long long)AG_BN * KW;
      if (e1 > total_elems) e1 = total_elems;
      const long long nvec = (e1 - e0) >> 3;
      const long long per_cta = (nvec + gridDim.x - 1) / gridDim.x;
      const long long v0 = (long long)blockIdx.x * per_cta;
      long long v1 = v0 + per_cta;
      if (v1 > nvec) v1 = nvec;
      constexpr int UNR = 8;
      for (long long vb = v0 + ct; vb < v1; vb += (long long)CT * UNR) {
        uint4 val[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const long long v = vb + (long long)u * CT;
          if (v < v1) {
            const long long e = e0 + v * 8;
            const int p = (int)(e / n_shard);
            val[u] = ad_ld_peer(peers.shard[p] + (e - (long long)p * n_shard));
          }
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          const long long v = vb + (long long)u * CT;
          if (v < v1) *reinterpret_cast<uint4*>(Wfull + e0 + v * 8) = val[u];
        }
      }
      __threadfence();                                         // slices visible GPU-wide before the count
      asm volatile("bar.sync 1, %0;" ::"n"(CT) : "memory");    // only the comm warps
      if (ct == 0) atomicAdd(chunk_ready + tile, 1u);
    }
  } else if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      bool all_seen = false;                                   // after one full k-loop every chunk is known to be there
      for (int t = cluster_id; t < total_tiles; t += num_clusters) {
        const int nt = t / m_tiles, mt = t - nt * m_tiles;
        const int m0 = (mt * CG + (int)cta_rank) * AG_BM;
        const int nb0 = nt * AG_BN + (int)cta_rank * Cfg::LOAD_N;
        for (int pos = 0; pos < n_chunks; ++pos) {
          const int chunk = ad_chunk_of_position(pos, n_chunks, first_chunk);
          if (!all_seen) {
            // acquire the chunk: all gridDim.x slices have been written through to Wfull
            unsigned ns = 64;
            while (ad_ld_acquire_gpu(chunk_ready + chunk) < gridDim.x) {
              __nanosleep(ns);
              if (ns < 2048) ns <<= 1;
            }
            asm volatile("fence.proxy.async.global;" ::: "memory");  // generic-proxy writes -> async-proxy (TMA) reads
          }
          for (int w = 0; w < KB_PER_CHUNK; ++w) {
            const int kb = chunk * KB_PER_CHUNK + w;
            if (kb >= num_kb) break;
            mbar_wait(empty_bar(s), ph ^ 1u);
            const uint32_t a_dst = a_base + s * Cfg::A_BYTES, b_dst = b_base + s * Cfg::B_BYTES;
            const int k0 = kb * AG_BK;
            if constexpr (CG == 1) {
              mbar_expect_tx(full_bar(s), Cfg::STAGE_BYTES);
              tma_load_2d(&tmA, full_bar(s), a_dst, k0, m0);
#pragma unroll
              for (int c = 0; c < Cfg::LOAD_N / 64; ++c)
                tma_load_2d(&tmB, full_bar(s), b_dst + c * (AG_BK * 128), nb0 + 64 * c, k0);
            } else {
              const uint32_t bar = mapa_u32(full_bar(s), 0);
              if (is_leader) mbar_expect_tx(full_bar(s), Cfg::STAGE_BYTES * CG);
              tma_load_2d_cg2(&tmA, bar, a_dst, k0, m0);
#pragma unroll
              for (int c = 0; c < Cfg::LOAD_N / 64; ++c)
                tma_load_2d_cg2(&tmB, bar, b_dst + c * (AG_BK * 128), nb0 + 64 * c, k0);
            }
            if (++s == STAGES) { s = 0; ph ^= 1u; }
          }
        }
        all_seen = true;
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0 && is_leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(AG_BM * CG, AG_BN, false, true);   // B = W consumed MN-major
      int s = 0, it = 0;
      uint32_t ph = 0;
      for (int t = cluster_id; t < total_tiles; t += num_clusters, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        if constexpr (CG == 1) mbar_wait(tempty_bar(as), aph ^ 1u); else mbar_wait_cluster(tempty_bar(as), aph ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * AG_BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(s), ph);
          tc_fence_after();
          const uint32_t a_src = a_base + s * Cfg::A_BYTES, b_src = b_base + s * Cfg::B_BYTES;
#pragma unroll
          for (int k = 0; k < AG_BK / AG_UK; ++k)
            umma_f16<CG>(tmem_d, umma_smem_desc_sw128(a_src + k * (AG_UK * 2), 16, 1024),
                         umma_smem_desc_sw128(b_src + k * (AG_UK * 128), AG_BK * 128, 1024), idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit<CG>(empty_bar(s));
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit<CG>(tfull_bar(as));
      }
    }
  } else {
    // =========================== epilogue (warps 2..5) ===========================
    const int q = warp & 3;
    const bool vec_ok = (N % 8 == 0) && (ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(Cp) & 15) == 0);
    int it = 0;
    for (int t = cluster_id; t < total_tiles; t += num_clusters, ++it) {
      const int nt = t / m_tiles, mt = t - nt * m_tiles;
      const int m0 = (mt * CG + (int)cta_rank) * AG_BM, n0 = nt * AG_BN;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      mbar_wait(tfull_bar(as), aph);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * AG_BN;
#pragma unroll 1
      for (int c = 0; c < AG_BN / 32; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32(taddr + c * 32, acc);
        tmem_ld_wait();
        const int col0 = n0 + c * 32;
        if (row < M && col0 < N) {
          __nv_bfloat16* crow = Cp + (size_t)row * ldc + col0;
          if (vec_ok) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (col0 + 8 * j < N) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(acc[8 * j]), __uint_as_float(acc[8 * j + 1]));
                o.y = pack_bf16x2(__uint_as_float(acc[8 * j + 2]), __uint_as_float(acc[8 * j + 3]));
                o.z = pack_bf16x2(__uint_as_float(acc[8 * j + 4]), __uint_as_float(acc[8 * j + 5]));
                o.w = pack_bf16x2(__uint_as_float(acc[8 * j + 6]), __uint_as_float(acc[8 * j + 7]));
                *reinterpret_cast<uint4*>(crow + 8 * j) = o;
              }
            }
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (col0 + e < N) crow[e] = __float2bfloat16_rn(__uint_as_float(acc[e]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 1) mbar_arrive(tempty_bar(as)); else mbar_arrive_cluster(mapa_u32(tempty_bar(as), 0));
      }
    }
  }

  tc_fence_before();
  __syncwarp();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, 512);
  }
}

template <int CG>
static int launch_ag_gemm_dgrad(const void* dY, void* const* shard_ptrs, void* Wfull, void* dX, uint32_t* chunk_ready,
                                int M, int NW, int KW, int ldy, int ldc, int rank, int ws, cudaStream_t st) {
  using Cfg = AdCfg<CG>;
  CUtensorMap tmA, tmB;
  if (int rc = make_tmap_2d_bf16(&tmA, dY, M, NW, ldy, AG_BM, AG_BK)) return rc;           // A = dY, K-major
  if (int rc = make_tmap_2d_bf16(&tmB, Wfull, NW, KW, KW, AG_BK, 64)) return rc;            // B = W, MN-major boxes
  AdPeers peers;
  for (int i = 0; i < AG_MAX_WS; ++i) peers.shard[i] = i < ws ? (const __nv_bfloat16*)shard_ptrs[i] : nullptr;
  auto kern = ag_gemm_dgrad_kernel<CG>;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_done = true;
  }
  const int n_chunks = (NW + AG_BN - 1) / AG_BN;
  BZ_CUDA_CHECK(cudaMemsetAsync(chunk_ready, 0, sizeof(uint32_t) * n_chunks, st));
  // Every CTA carries communication warps, so the grid is always the full machine: chunk completion counts
  // gridDim.x slices.
  const int sms = device_sm_count();
  const int clusters = sms / CG;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CG);
  cfg.blockDim = dim3(AG_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BZ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, peers, (__nv_bfloat16*)Wfull, (__nv_bfloat16*)dX, chunk_ready,
                                   M, NW, KW, ldc, rank, ws));
  return 0;
}

}  // namespace bz

// dY [M,N] bf16 (ldy). W [N,K] row-partitioned over `ws` ranks: shard_ptrs[p] = rank p's rows ([N/ws, K] contiguous,
// peer-mapped; shard_ptrs[rank] is local). Wfull [N,K]: local buffer that receives the gathered weight (valid after
// the kernel). dX [M,K] bf16 (ldc). chunk_ready: >= ceil(N/256) uint32 of device scratch (zeroed by this call).
extern "C" int bz_ag_gemm_dgrad_bf16(const void* dY, void* const* shard_ptrs, void* Wfull, void* dX, void* chunk_ready,
                                     int M, int N, int K, int ldy, int ldc, int rank, int ws, int cta_group,
                                     void* stream) {
  using namespace bz;
  BZ_REQUIRE(M > 0 && N > 0 && K > 0, "ag_gemm_dgrad: empty problem");
  BZ_REQUIRE(ws >= 1 && ws <= AG_MAX_WS && rank >= 0 && rank < ws, "ag_gemm_dgrad: bad rank/world %d/%d", rank, ws);
  BZ_REQUIRE(N % ws == 0 && ((long long)(N / ws) * K) % 8 == 0, "ag_gemm_dgrad: N=%d not partitionable over %d ranks", N, ws);
  BZ_REQUIRE(K % 8 == 0 && N % 8 == 0, "ag_gemm_dgrad: N=%d and K=%d must be multiples of 8", N, K);
  BZ_REQUIRE((N + AG_BN - 1) / AG_BN <= 4096, "ag_gemm_dgrad: too many chunks");
  cudaStream_t st = (cudaStream_t)stream;
  if (cta_group == 0) cta_group = (M > AG_BM) ? 2 : 1;
  if (cta_group == 2)
    return launch_ag_gemm_dgrad<2>(dY, shard_ptrs, Wfull, dX, (uint32_t*)chunk_ready, M, N, K, ldy, ldc, rank, ws, st);
  return launch_ag_gemm_dgrad<1>(dY, shard_ptrs, Wfull, dX, (uint32_t*)chunk_ready, M, N, K, ldy, ldc, rank, ws, st);
}
