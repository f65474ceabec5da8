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
// Status: written after this round's GPU time was spent; checked by a child-process test on emulated ranks.
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
  int pad;
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
