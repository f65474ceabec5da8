// Pull all-gather, second generation (opt-in: B200ZERO_PULL_AG=2). Same contract as bz_pull_allgather in
// fused_comm.cu: every rank copies, for each partitioned tensor of a unit, the rows held by every rank's
// (peer-mapped) shard buffer into its local gathered buffer; replicated tensors are copied from the local shard.
//
// v1 interleaves the peers through a flat index j and recovers (peer, vector) with a 64-bit `j % ws` and `j / ws`
// per 16-byte vector — on a runtime ws that is two integer-division sequences per load, and the kernel was measured
// at 0.56-0.58 of the NVLink roofline (443 GB/s) where the reduce-scatter, whose inner loop has no division, reaches
// 0.76. v2 walks the vector index and unrolls over the peers instead: one thread issues its vector from ALL peers
// back to back (ws independent 16-byte loads in flight, every link busy, starting at rank+1 so that the ranks do
// not all hit the same peer at once), with no division anywhere. The bytes moved and the result are identical.
//
// Validated on B200 (byte-identical output on emulated ranks; in the engine on 2 GPUs). Within noise of the first
// version end to end at 8 GPUs (396 vs 401 ms per step), so it is selected with B200ZERO_PULL_AG=2 only.
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int C2_MAX_WS = 8;
constexpr int C2_THREADS = 256;

struct C2Peers {
  const __nv_bfloat16* data[C2_MAX_WS];   // peer p's shard parameter buffer (bf16), peer-mapped
};

struct C2Segment {      // must match Segment in fused_comm.cu (32 bytes)
  long long full_off;
  long long numel;
  long long shard_off;
  int partitioned;
  int push;
};

__device__ __forceinline__ uint4 c2_ld_peer(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

__global__ void __launch_bounds__(C2_THREADS)
pull_allgather_v2_kernel(const __grid_constant__ C2Peers pt, const C2Segment* __restrict__ segs, int nseg,
                         __nv_bfloat16* __restrict__ dst_unit, int rank, int ws) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (int s = 0; s < nseg; ++s) {
    const C2Segment sg = segs[s];
    if (!sg.partitioned) {
      const long long nvec = sg.numel >> 3;
      for (long long i = tid; i < nvec; i += stride)
        *reinterpret_cast<uint4*>(dst_unit + sg.full_off + i * 8) =
            *reinterpret_cast<const uint4*>(pt.data[rank] + sg.shard_off + i * 8);
      continue;
    }
    const long long n = sg.numel / ws;          // one division per tensor, not per vector
    const long long nvec = n >> 3;
    for (long long i = tid; i < nvec; i += stride) {
      uint4 val[C2_MAX_WS];
#pragma unroll
      for (int q = 0; q < C2_MAX_WS; ++q) {
        if (q < ws) {
          int p = rank + 1 + q;                 // rank+1, rank+2, ..., rank (own rows last: they are local)
          if (p >= ws) p -= ws;
          val[q] = c2_ld_peer(pt.data[p] + sg.shard_off + i * 8);
        }
      }
#pragma unroll
      for (int q = 0; q < C2_MAX_WS; ++q) {
        if (q < ws) {
          int p = rank + 1 + q;
          if (p >= ws) p -= ws;
          *reinterpret_cast<uint4*>(dst_unit + sg.full_off + (long long)p * n + i * 8) = val[q];
        }
      }
    }
  }
}

}  // namespace bz

using namespace bz;

extern "C" int bz_pull_allgather_v2(void* const* shard_param_ptrs, const void* segs, int nseg, void* dst_unit, int rank,
                                    int ws, int ctas, void* stream) {
  BZ_REQUIRE(ws >= 1 && ws <= C2_MAX_WS && rank >= 0 && rank < ws, "pull_allgather_v2: bad rank/world %d/%d", rank, ws);
  static_assert(sizeof(C2Segment) == 32, "segment layout");
  C2Peers pt;
  for (int i = 0; i < C2_MAX_WS; ++i) pt.data[i] = i < ws ? (const __nv_bfloat16*)shard_param_ptrs[i] : nullptr;
  static bool attr_done = false;
  if (!attr_done) {   // shares SMs with persistent GEMM CTAs: do not pin the L1/shared split to the L1-heavy default
    BZ_CUDA_CHECK(cudaFuncSetAttribute(pull_allgather_v2_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       cudaSharedmemCarveoutMaxShared));
    attr_done = true;
  }
  if (ctas <= 0) ctas = 64;
  if (ctas > 1024) ctas = 1024;
  pull_allgather_v2_kernel<<<ctas, C2_THREADS, 0, (cudaStream_t)stream>>>(pt, (const C2Segment*)segs, nseg,
                                                                        (__nv_bfloat16*)dst_unit, rank, ws);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// -------------------------------------------------------------------------------------------------
// Per-step control exchange without NCCL (opt-in: B200ZERO_STATS_PEER=1).
// After the last reduce-scatter every rank holds four floats — its partial sum of g^2 over the partitioned
// region, its non-finite flag, its loss share, one spare — and every rank needs their sum. v1 sends them through
// a 16-byte ncclAllReduce; this kernel stores them into every peer's signal pad (st.relaxed.sys of one 16-byte
// vector, then st.release.sys of an epoch flag), waits for every peer's flag (ld.acquire.sys) and adds the eight
// vectors in rank order, so every rank computes the bit-identical sum and takes the identical clip / skip decision.
// The payload area is the 64 spare words behind the flag table of the signal pad, double-buffered by epoch parity:
// a rank cannot run two exchanges ahead of a peer because each exchange waits for all peers.
// -------------------------------------------------------------------------------------------------
namespace bz {

constexpr int C2_SIG_CHANNELS = 8, C2_SIG_PHASES = 3;          // must match fused_comm.cu
constexpr int C2_FLAG_WORDS = C2_SIG_CHANNELS * C2_SIG_PHASES * C2_MAX_WS;

struct C2Pads {
  uint32_t* pad[C2_MAX_WS];
};

__global__ void __launch_bounds__(32)
stats_exchange_kernel(const __grid_constant__ C2Pads pads, float* __restrict__ vals4, uint32_t epoch, int channel,
                      int rank, int ws) {
  const int lane = threadIdx.x;
  const int buf = (int)(epoch & 1u);
  const float4 mine = *reinterpret_cast<const float4*>(vals4);
  if (lane < ws) {
    // my four floats -> slot [buf][rank] of peer `lane`, then the flag that publishes them
    float* dst = reinterpret_cast<float*>(pads.pad[lane] + C2_FLAG_WORDS) + (buf * C2_MAX_WS + rank) * 4;
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(mine.x), "f"(mine.y),
                 "f"(mine.z), "f"(mine.w)
                 : "memory");
    uint32_t* flag = pads.pad[lane] + (channel * C2_SIG_PHASES + 0) * C2_MAX_WS + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
    // wait for peer `lane`'s vector in my own pad
    const uint32_t* f = pads.pad[rank] + (channel * C2_SIG_PHASES + 0) * C2_MAX_WS + lane;
    unsigned ns = 32;
    for (;;) {
      uint32_t v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if ((int32_t)(v - epoch) >= 0) break;
      __nanosleep(ns);
      if (ns < 1024) ns <<= 1;
    }
  }
  __syncwarp();
  if (lane == 0) {
    const float* src = reinterpret_cast<const float*>(pads.pad[rank] + C2_FLAG_WORDS) + buf * C2_MAX_WS * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int p = 0; p < ws; ++p) {                      // rank order: identical sum on every rank
      float4 v;
      asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                   : "l"(src + p * 4)
                   : "memory");
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    *reinterpret_cast<float4*>(vals4) = acc;
  }
}

}  // namespace bz

// vals4: four device floats (16-byte aligned), replaced by their sum over all ranks. signal_ptrs: every rank's
// signal pad (bz_comm_signal_pad_bytes, peer-mapped). One call per step on every rank with the same epoch (>= 1,
// growing) and a channel no other kernel uses.
extern "C" int bz_peer_allreduce4(void* const* signal_ptrs, float* vals4, unsigned int epoch, int channel, int rank,
                                  int ws, void* stream) {
  BZ_REQUIRE(ws >= 1 && ws <= C2_MAX_WS && rank >= 0 && rank < ws, "peer_allreduce4: bad rank/world %d/%d", rank, ws);
  BZ_REQUIRE(channel >= 0 && channel < C2_SIG_CHANNELS, "peer_allreduce4: channel %d out of range", channel);
  BZ_REQUIRE(((uintptr_t)vals4 & 15) == 0, "peer_allreduce4: vals4 must be 16-byte aligned");
  C2Pads pads;
  for (int i = 0; i < C2_MAX_WS; ++i) pads.pad[i] = i < ws ? reinterpret_cast<uint32_t*>(signal_ptrs[i]) : nullptr;
  static bool attr_done = false;
  if (!attr_done) {
    BZ_CUDA_CHECK(cudaFuncSetAttribute(stats_exchange_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                       cudaSharedmemCarveoutMaxShared));
    attr_done = true;
  }
  stats_exchange_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, vals4, epoch, channel, rank, ws);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
