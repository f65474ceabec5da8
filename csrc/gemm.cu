// K4: persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   C[M,N] (+)= A * B      fp32 accumulation in TMEM, bf16 or fp32 output
//
// Operand storage ("major-ness") is selectable per operand so that forward,
// dgrad and wgrad all run without materialising a transpose:
//   A K-major : A stored [M,K] row-major          A MN-major: A stored [K,M] row-major
//   B K-major : B stored [N,K] row-major (= W)    B MN-major: B stored [K,N] row-major
//     fwd   Y[T,N]  = X[T,K]  W[N,K]^T   -> A K,  B K
//     dgrad dX[T,K] = dY[T,N] W[N,K]     -> A K,  B MN   (reduction dim = N)
//     wgrad dW[N,K] = dY[T,N]^T X[T,K]   -> A MN, B MN   (reduction dim = T)
//
// Structure (one CTA, or a cta_group::2 CTA pair, per SM; persistent over tiles):
//   warp 0      TMA producer  : cp.async.bulk.tensor tiles -> 128B-swizzled smem ring
//   warp 1      MMA issuer    : one lane issues tcgen05.mma (UMMA 128x256x16, or 256x256x16
//                               across the CTA pair), accumulators double-buffered in TMEM
//   warps 2..5  epilogue      : tcgen05.ld TMEM -> registers -> (accumulate) -> global
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty (MMA <-> epilogue);
// the epilogue of tile i overlaps the main loop of tile i+1.
//
// The reference delegates these GEMMs to cuBLAS through a user script it does not
// ship; there is no reference code for this file.
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int BM = 128;  // rows of C per CTA
constexpr int BN = 256;  // UMMA N (columns of C per tile)
constexpr int BK = 64;   // 64 bf16 = one 128-byte swizzle span
constexpr int UK = 16;   // UMMA K for 16-bit inputs
constexpr int GEMM_THREADS = 192;
constexpr int RASTER_GROUP = 8;

template <int CG>
struct GemmCfg {
  static constexpr int LOAD_N = BN / CG;  // rows of B each CTA stages
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = LOAD_N * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (CG == 1) ? 4 : 6;
  static constexpr int BAR_BYTES = 8 * (2 * STAGES + 4) + 16;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + BAR_BYTES + 1024;  // +1024 alignment slack
};

enum OutMode { OUT_BF16 = 0, OUT_BF16_ACC = 1, OUT_F32 = 2, OUT_F32_ACC = 3 };

struct TileCoord {
  int mt, nt;
};
__device__ __forceinline__ TileCoord raster(int tile, int m_tiles, int n_tiles) {
  const int group_size = RASTER_GROUP * n_tiles;
  const int g = tile / group_size;
  const int first_m = g * RASTER_GROUP;
  const int gm = min(RASTER_GROUP, m_tiles - first_m);
  const int r = tile - g * group_size;
  TileCoord t;
  t.mt = first_m + r % gm;
  t.nt = r / gm;
  return t;
}

template <int OUT>
__device__ __forceinline__ void epilogue_store_32(void* __restrict__ Cp, size_t ldc, int row, int col0, int M,
                                                  int N, const uint32_t (&acc)[32], bool vec_ok) {
  if (row >= M || col0 >= N) return;
  if constexpr (OUT == OUT_BF16 || OUT == OUT_BF16_ACC) {
    __nv_bfloat16* crow = reinterpret_cast<__nv_bfloat16*>(Cp) + (size_t)row * ldc + col0;
    if (vec_ok) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (col0 + 8 * j < N) {
          uint4 o;
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(acc[8 * j + e]);
          if constexpr (OUT == OUT_BF16_ACC) {
            const uint4 old = *reinterpret_cast<const uint4*>(crow + 8 * j);
            float2 t;
            t = unpack_bf16x2(old.x); f[0] += t.x; f[1] += t.y;
            t = unpack_bf16x2(old.y); f[2] += t.x; f[3] += t.y;
            t = unpack_bf16x2(old.z); f[4] += t.x; f[5] += t.y;
            t = unpack_bf16x2(old.w); f[6] += t.x; f[7] += t.y;
          }
          o.x = pack_bf16x2(f[0], f[1]);
          o.y = pack_bf16x2(f[2], f[3]);
          o.z = pack_bf16x2(f[4], f[5]);
          o.w = pack_bf16x2(f[6], f[7]);
          *reinterpret_cast<uint4*>(crow + 8 * j) = o;
        }
      }
    } else {
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        if (col0 + e < N) {
          float v = __uint_as_float(acc[e]);
          if constexpr (OUT == OUT_BF16_ACC) v += __bfloat162float(crow[e]);
          crow[e] = __float2bfloat16_rn(v);
        }
      }
    }
  } else {
    float* crow = reinterpret_cast<float*>(Cp) + (size_t)row * ldc + col0;
    if (vec_ok) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (col0 + 4 * j < N) {
          float4 o = make_float4(__uint_as_float(acc[4 * j]), __uint_as_float(acc[4 * j + 1]),
                                 __uint_as_float(acc[4 * j + 2]), __uint_as_float(acc[4 * j + 3]));
          if constexpr (OUT == OUT_F32_ACC) {
            const float4 old = *reinterpret_cast<const float4*>(crow + 4 * j);
            o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
          }
          *reinterpret_cast<float4*>(crow + 4 * j) = o;
        }
      }
    } else {
#pragma unroll
      for (int e = 0; e < 32; ++e) {
        if (col0 + e < N) {
          float v = __uint_as_float(acc[e]);
          if constexpr (OUT == OUT_F32_ACC) v += crow[e];
          crow[e] = v;
        }
      }
    }
  }
}

template <int CG, bool A_MN, bool B_MN, int OUT>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  void* __restrict__ Cp, int M, int N, int K, int ldc) {
  using Cfg = GemmCfg<CG>;
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

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = (CG == 1) ? 0u : cluster_ctarank();
  const bool is_leader = (cta_rank == 0);

  const int m_tiles = (M + BM * CG - 1) / (BM * CG);
  const int n_tiles = (N + BN - 1) / BN;
  const int total_tiles = m_tiles * n_tiles;
  const int num_kb = (K + BK - 1) / BK;
  const int cluster_id = blockIdx.x / CG;
  const int num_clusters = gridDim.x / CG;

  // ---------------- one-time setup ----------------
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(tfull_bar(i), 1);
      mbar_init(tempty_bar(i), 4 * CG);  // one arrive per epilogue warp of every CTA in the group
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

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters) {
        const TileCoord tc = raster(tile, m_tiles, n_tiles);
        const int m0 = (tc.mt * CG + (int)cta_rank) * BM;
        const int nb0 = tc.nt * BN + (int)cta_rank * Cfg::LOAD_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t a_dst = a_base + s * Cfg::A_BYTES;
          const uint32_t b_dst = b_base + s * Cfg::B_BYTES;
          const int k0 = kb * BK;
          if constexpr (CG == 1) {
            const uint32_t bar = full_bar(s);
            mbar_expect_tx(bar, Cfg::STAGE_BYTES);
            if constexpr (!A_MN) {
              tma_load_2d(&tmA, bar, a_dst, k0, m0);
            } else {
#pragma unroll
              for (int c = 0; c < BM / 64; ++c) tma_load_2d(&tmA, bar, a_dst + c * (BK * 128), m0 + 64 * c, k0);
            }
            if constexpr (!B_MN) {
              tma_load_2d(&tmB, bar, b_dst, k0, nb0);
            } else {
#pragma unroll
              for (int c = 0; c < Cfg::LOAD_N / 64; ++c)
                tma_load_2d(&tmB, bar, b_dst + c * (BK * 128), nb0 + 64 * c, k0);
            }
          } else {
            // Both CTAs stage their halves; all bytes are accounted on the leader's barrier.
            const uint32_t bar = mapa_u32(full_bar(s), 0);
            if (is_leader) mbar_expect_tx(full_bar(s), Cfg::STAGE_BYTES * CG);
            if constexpr (!A_MN) {
              tma_load_2d_cg2(&tmA, bar, a_dst, k0, m0);
            } else {
#pragma unroll
              for (int c = 0; c < BM / 64; ++c)
                tma_load_2d_cg2(&tmA, bar, a_dst + c * (BK * 128), m0 + 64 * c, k0);
            }
            if constexpr (!B_MN) {
              tma_load_2d_cg2(&tmB, bar, b_dst, k0, nb0);
            } else {
#pragma unroll
              for (int c = 0; c < Cfg::LOAD_N / 64; ++c)
                tma_load_2d_cg2(&tmB, bar, b_dst + c * (BK * 128), nb0 + 64 * c, k0);
            }
          }
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (lane == 0 && is_leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      int s = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++it) {
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        if constexpr (CG == 1) mbar_wait(tempty_bar(as), aph ^ 1u); else mbar_wait_cluster(tempty_bar(as), aph ^ 1u);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(s), ph);
          tc_fence_after();
          const uint32_t a_src = a_base + s * Cfg::A_BYTES;
          const uint32_t b_src = b_base + s * Cfg::B_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            const uint64_t adesc = A_MN ? umma_smem_desc_sw128(a_src + k * (UK * 128), BK * 128, 1024)
                                        : umma_smem_desc_sw128(a_src + k * (UK * 2), 16, 1024);
            const uint64_t bdesc = B_MN ? umma_smem_desc_sw128(b_src + k * (UK * 128), BK * 128, 1024)
                                        : umma_smem_desc_sw128(b_src + k * (UK * 2), 16, 1024);
            umma_f16<CG>(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit<CG>(empty_bar(s));  // smem slot reusable once these MMAs retire
          if (++s == STAGES) { s = 0; ph ^= 1u; }
        }
        umma_commit<CG>(tfull_bar(as));  // accumulator ready for the epilogue
      }
    }
  } else {
    // =========================== epilogue ===========================
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    const bool vec_ok = (OUT == OUT_BF16 || OUT == OUT_BF16_ACC)
                            ? ((N % 8 == 0) && (ldc % 8 == 0) && ((reinterpret_cast<uintptr_t>(Cp) & 15) == 0))
                            : ((N % 4 == 0) && (ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(Cp) & 15) == 0));
    int it = 0;
    for (int tile = cluster_id; tile < total_tiles; tile += num_clusters, ++it) {
      const TileCoord tc = raster(tile, m_tiles, n_tiles);
      const int m0 = (tc.mt * CG + (int)cta_rank) * BM;
      const int n0 = tc.nt * BN;
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      mbar_wait(tfull_bar(as), aph);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t acc[32];
        tmem_ld_32x32(taddr + c * 32, acc);
        tmem_ld_wait();
        epilogue_store_32<OUT>(Cp, (size_t)ldc, row, n0 + c * 32, M, N, acc, vec_ok);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CG == 1) mbar_arrive(tempty_bar(as)); else mbar_arrive_cluster(mapa_u32(tempty_bar(as), 0));
      }
    }
  }

  // ---------------- teardown ----------------
  tc_fence_before();
  __syncwarp();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<CG>(tmem_base, 512);
  }
}

template <int CG, bool A_MN, bool B_MN, int OUT>
static int launch_gemm(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                       cudaStream_t stream, int max_ctas) {
  using Cfg = GemmCfg<CG>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!A_MN) rc = make_tmap_2d_bf16(&tmA, A, M, K, lda, BM, BK);
  else       rc = make_tmap_2d_bf16(&tmA, A, K, M, lda, BK, 64);
  if (rc) return rc;
  if (!B_MN) rc = make_tmap_2d_bf16(&tmB, B, N, K, ldb, Cfg::LOAD_N, BK);
  else       rc = make_tmap_2d_bf16(&tmB, B, K, N, ldb, BK, 64);
  if (rc) return rc;

  auto kern = gemm_bf16_tcgen05<CG, A_MN, B_MN, OUT>;
  static bool attr_done = false;  // per template instantiation
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_done = true;
  }
  const int m_tiles = (M + BM * CG - 1) / (BM * CG);
  const int n_tiles = (N + BN - 1) / BN;
  const int total = m_tiles * n_tiles;
  int sms = device_sm_count();
  if (max_ctas > 0 && max_ctas < sms) sms = max_ctas;
  int clusters = sms / CG;
  if (clusters > total) clusters = total;
  if (clusters < 1) clusters = 1;

  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(clusters * CG);
  cfg.blockDim = dim3(GEMM_THREADS);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BZ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, C, M, N, K, ldc));
  return 0;
}

template <int CG, bool A_MN, bool B_MN>
static int dispatch_out(int out_mode, const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                        int ldc, cudaStream_t st, int max_ctas) {
  switch (out_mode) {
    case OUT_BF16:     return launch_gemm<CG, A_MN, B_MN, OUT_BF16>(A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
    case OUT_BF16_ACC: return launch_gemm<CG, A_MN, B_MN, OUT_BF16_ACC>(A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
    case OUT_F32:      return launch_gemm<CG, A_MN, B_MN, OUT_F32>(A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
    case OUT_F32_ACC:  return launch_gemm<CG, A_MN, B_MN, OUT_F32_ACC>(A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
  }
  set_last_error("bz_gemm_bf16: bad out_mode %d", out_mode);
  return -1;
}

template <int CG>
static int dispatch_major(int a_mn, int b_mn, int out_mode, const void* A, const void* B, void* C, int M, int N,
                          int K, int lda, int ldb, int ldc, cudaStream_t st, int max_ctas) {
  if (!a_mn && !b_mn) return dispatch_out<CG, false, false>(out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
  if (!a_mn && b_mn)  return dispatch_out<CG, false, true>(out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
  if (a_mn && !b_mn)  return dispatch_out<CG, true, false>(out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
  return dispatch_out<CG, true, true>(out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
}

}  // namespace bz

extern "C" int bz_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                            int a_mn_major, int b_mn_major, int out_mode, int cta_group, int max_ctas,
                            void* stream) {
  using namespace bz;
  BZ_REQUIRE(M > 0 && N > 0 && K > 0, "bz_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  BZ_REQUIRE(A && B && C, "bz_gemm_bf16: null operand");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (cta_group == 0) cta_group = (M > BM) ? 2 : 1;
  if (cta_group == 2)
    return dispatch_major<2>(a_mn_major, b_mn_major, out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
  return dispatch_major<1>(a_mn_major, b_mn_major, out_mode, A, B, C, M, N, K, lda, ldb, ldc, st, max_ctas);
}
