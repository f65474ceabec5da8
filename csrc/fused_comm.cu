// K1 / K3 + the peer-memory runtime they need (H2).
//
//   bz_fused_reduce_scatter : every rank pulls ITS rows of a unit's gradients straight out of
//       all peers' buffers (128-bit loads on peer-mapped addresses over NVLink 5 / NVSwitch),
//       sums in fp32 in rank order (deterministic), scales, accumulates into the rank's gradient
//       shard, and produces the by-products the step needs with no extra pass: a ‖g‖² partial per
//       CTA and a non-finite flag (feeds LossSpikeMonitor / the predicated Adam update).
//       Replaces ncclReduceScatter + cast + _foreach_norm + isfinite (X1, X4, X5, X6, F9, F10).
//   bz_fused_adam_push      : AdamW on the rank's fp32 shard (predicated on the non-finite flag,
//       gradient scaled by an on-device clip coefficient) whose epilogue stores the updated bf16
//       rows directly into EVERY peer's parameter buffer (push all-gather, X3) — no separate
//       all-gather pass, no copy engine.
//   Cross-GPU ordering uses monotonically increasing epoch flags in a peer-mapped signal pad:
//   st.release.sys by the producer, ld.acquire.sys polling (with __nanosleep back-off) by the
//   consumer; one flag per (channel, phase, source rank). Kernels of different ranks run on
//   different GPUs, so the spins cannot starve one another (never launch two ranks' kernels on
//   one device — B200_PROFILING.md).
//
// The reference reaches none of this: its launcher only writes `reduce_scatter: true` into a
// JSON file for DeepSpeed (ref ai_engine/deepspeed_launcher.py:133-142).
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int MAX_WS = 8;
constexpr int COMM_THREADS = 256;  // small footprint: must co-reside with a persistent GEMM CTA on the same SM
constexpr int SIG_CHANNELS = 8;
constexpr int SIG_PHASES = 3;  // 0: data ready, 1: done reading, 2: pushed

struct PeerTable {
  void* data[MAX_WS];         // peer-mapped base of the buffer being communicated
  uint32_t* signal[MAX_WS];   // peer-mapped signal pads: [SIG_CHANNELS][SIG_PHASES][MAX_WS] uint32
};

__device__ __forceinline__ int sig_index(int channel, int phase, int src) {
  return (channel * SIG_PHASES + phase) * MAX_WS + src;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_v4(void* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.L1::no_allocate.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// Announce `epoch` on (channel, phase) to every rank, then wait until every rank announced it.
// Called by all threads of a CTA; only CTA 0 announces. Safe against reordering of this CTA's
// earlier global writes (fence) and later reads (acquire + barrier).
__device__ __forceinline__ void signal_all_and_wait(const PeerTable& pt, int channel, int phase, int rank, int ws,
                                                    uint32_t epoch, bool announce) {
  __syncthreads();
  if (announce && threadIdx.x < ws) {
    __threadfence_system();
    st_release_sys(pt.signal[threadIdx.x] + sig_index(channel, phase, rank), epoch);
  }
  if (threadIdx.x < ws) {
    const uint32_t* f = pt.signal[rank] + sig_index(channel, phase, threadIdx.x);
    unsigned ns = 32;
    // epochs only grow; signed distance handles wrap-around
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      __nanosleep(ns);
      if (ns < 1024) ns <<= 1;
    }
  }
  __syncthreads();
}

struct Segment {
  long long full_off;   // element offset of the tensor inside the unit buffer
  long long numel;      // full tensor size
  long long shard_off;  // element offset inside the rank's shard buffers
  int partitioned;      // 1: this rank owns rows [rank*numel/ws, (rank+1)*numel/ws); 0: replicated
  int push;             // K3 only: publish this tensor's updated rows even when the launch as a whole does not push
};

// -------------------------------------------------------------------------------------------------
// K1: reduce-scatter (+ all-reduce of replicated tensors) fused with scale / accumulate / ‖g‖² / isfinite
// -------------------------------------------------------------------------------------------------
template <bool ACC_F32>
__global__ void __launch_bounds__(COMM_THREADS)
fused_reduce_scatter_kernel(const __grid_constant__ PeerTable pt, const Segment* __restrict__ segs, int nseg,
                            void* __restrict__ shard_grads, float scale, float* __restrict__ sumsq_part,
                            float* __restrict__ sumsq_repl, int* __restrict__ nonfinite, unsigned int* block_counter,
                            uint32_t epoch, int channel, int rank, int ws, int use_barrier) {
  __shared__ float red[32];
  __shared__ int is_last;
  // use_barrier: 1 = phase 0 in this kernel (every CTA waits until all ranks' buffers are ready) + phase 1;
  //              2 = phase 0 was done by comm_ready_kernel in front of this launch, only phase 1 here
  if (use_barrier == 1) signal_all_and_wait(pt, channel, 0, rank, ws, epoch, blockIdx.x == 0);

  float ss_part = 0.f, ss_repl = 0.f;
  for (int s = 0; s < nseg; ++s) {
    const Segment sg = segs[s];
    const long long n = sg.partitioned ? sg.numel / ws : sg.numel;
    const long long src0 = sg.full_off + (sg.partitioned ? (long long)rank * n : 0);
    const long long nvec = n >> 3;  // 8 bf16 per 128-bit load (layout guarantees 16-element alignment)
    float ss = 0.f;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
      float acc[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = 0.f;
      uint4 in[MAX_WS];
#pragma unroll
      for (int p = 0; p < MAX_WS; ++p) {  // issue every peer load before consuming any (NVLink latency ~2 us)
        if (p < ws) in[p] = ld_peer_v4(reinterpret_cast<const __nv_bfloat16*>(pt.data[p]) + src0 + v * 8);
      }
#pragma unroll
      for (int p = 0; p < MAX_WS; ++p) {
        if (p < ws) {
          float2 t;
          t = unpack_bf16x2(in[p].x); acc[0] += t.x; acc[1] += t.y;
          t = unpack_bf16x2(in[p].y); acc[2] += t.x; acc[3] += t.y;
          t = unpack_bf16x2(in[p].z); acc[4] += t.x; acc[5] += t.y;
          t = unpack_bf16x2(in[p].w); acc[6] += t.x; acc[7] += t.y;
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] *= scale;
      const long long d = sg.shard_off + v * 8;
      if constexpr (ACC_F32) {
        float* o = reinterpret_cast<float*>(shard_grads) + d;
        float4 a = *reinterpret_cast<float4*>(o), b = *reinterpret_cast<float4*>(o + 4);
        a.x += acc[0]; a.y += acc[1]; a.z += acc[2]; a.w += acc[3];
        b.x += acc[4]; b.y += acc[5]; b.z += acc[6]; b.w += acc[7];
        *reinterpret_cast<float4*>(o) = a;
        *reinterpret_cast<float4*>(o + 4) = b;
        ss += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
      } else {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(shard_grads) + d;
        const uint4 old = *reinterpret_cast<const uint4*>(o);
        float2 t;
        t = unpack_bf16x2(old.x); acc[0] += t.x; acc[1] += t.y;
        t = unpack_bf16x2(old.y); acc[2] += t.x; acc[3] += t.y;
        t = unpack_bf16x2(old.z); acc[4] += t.x; acc[5] += t.y;
        t = unpack_bf16x2(old.w); acc[6] += t.x; acc[7] += t.y;
        uint4 w;
        w.x = pack_bf16x2(acc[0], acc[1]); w.y = pack_bf16x2(acc[2], acc[3]);
        w.z = pack_bf16x2(acc[4], acc[5]); w.w = pack_bf16x2(acc[6], acc[7]);
        *reinterpret_cast<uint4*>(o) = w;
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += acc[e] * acc[e];
      }
    }
    if (sg.partitioned) ss_part += ss; else ss_repl += ss;
  }
  // per-CTA partials: [0, grid) partitioned region, [grid, 2*grid) replicated region
  const float tp = block_sum(ss_part, red);
  const float tr = block_sum(ss_repl, red);
  if (threadIdx.x == 0) {
    sumsq_part[blockIdx.x] = tp;
    sumsq_repl[blockIdx.x] = tr;
    if (!(isfinite(tp) && isfinite(tr)) && nonfinite != nullptr) atomicOr(nonfinite, 1);
  }
  if (use_barrier) {
    // phase 1: tell every peer we are done reading its buffer; the last CTA waits for all peers so
    // that kernel completion on this rank means "my gradient buffer may be overwritten".
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      is_last = (atomicAdd(block_counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
      if (threadIdx.x == 0) *block_counter = 0;
      signal_all_and_wait(pt, channel, 1, rank, ws, epoch, true);
    }
  }
}

// Phase 0 of K1 as a launch of its own: one warp announces "my buffer for this epoch is complete" to every rank and
// waits until every rank has said the same; the reduce kernel behind it on the stream then starts with all inputs
// ready and none of its CTAs ever spins for a peer. Done inside the reduce kernel (use_barrier == 1), all of its
// CTAs sit on their SMs polling until the slowest rank arrives — registers and thread slots the compute kernels of
// the backward pass could be using, and at 8 ranks long runs stalled that way.
__global__ void __launch_bounds__(32)
comm_ready_kernel(const __grid_constant__ PeerTable pt, int channel, int rank, int ws, uint32_t epoch) {
  signal_all_and_wait(pt, channel, 0, rank, ws, epoch, true);
}

// -------------------------------------------------------------------------------------------------
// K1, NVLS flavour (opt-in): the gradient buffer is additionally bound to an NVSwitch multicast object
// (`mc` = multicast VA of pt.data[*]). Partitioned tensors are summed *inside the switch* by
// multimem.ld_reduce — one 128-bit instruction returns the bf16 sum (fp32-accumulated) of the ws copies,
// so this GPU receives 1/(ws-1) of the NVLink bytes and needs far fewer CTAs to keep the links busy.
// Replicated tensors stay on the rank-ordered point-to-point path: every rank must produce the
// bit-identical sum. A separate kernel so the default point-to-point kernel is untouched.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc_addr) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc_addr)
               : "memory");
  return v;
}

// shard[d .. d+8) += acc; returns the sum of squares of the new values
template <bool ACC_F32>
__device__ __forceinline__ float accumulate_shard(void* __restrict__ shard_grads, long long d, float (&acc)[8]) {
  float ss = 0.f;
  if constexpr (ACC_F32) {
    float* o = reinterpret_cast<float*>(shard_grads) + d;
    float4 a = *reinterpret_cast<float4*>(o), b = *reinterpret_cast<float4*>(o + 4);
    a.x += acc[0]; a.y += acc[1]; a.z += acc[2]; a.w += acc[3];
    b.x += acc[4]; b.y += acc[5]; b.z += acc[6]; b.w += acc[7];
    *reinterpret_cast<float4*>(o) = a;
    *reinterpret_cast<float4*>(o + 4) = b;
    ss = a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(shard_grads) + d;
    const uint4 old = *reinterpret_cast<const uint4*>(o);
    float2 t;
    t = unpack_bf16x2(old.x); acc[0] += t.x; acc[1] += t.y;
    t = unpack_bf16x2(old.y); acc[2] += t.x; acc[3] += t.y;
    t = unpack_bf16x2(old.z); acc[4] += t.x; acc[5] += t.y;
    t = unpack_bf16x2(old.w); acc[6] += t.x; acc[7] += t.y;
    uint4 w;
    w.x = pack_bf16x2(acc[0], acc[1]); w.y = pack_bf16x2(acc[2], acc[3]);
    w.z = pack_bf16x2(acc[4], acc[5]); w.w = pack_bf16x2(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(o) = w;
#pragma unroll
    for (int e = 0; e < 8; ++e) ss += acc[e] * acc[e];
  }
  return ss;
}

template <bool ACC_F32>
__global__ void __launch_bounds__(COMM_THREADS)
fused_reduce_scatter_nvls_kernel(const __grid_constant__ PeerTable pt, const void* __restrict__ mc_base,
                                 const Segment* __restrict__ segs, int nseg, void* __restrict__ shard_grads, float scale,
                                 float* __restrict__ sumsq_part, float* __restrict__ sumsq_repl,
                                 int* __restrict__ nonfinite, unsigned int* block_counter, uint32_t epoch, int channel,
                                 int rank, int ws, int use_barrier) {
  __shared__ float red[32];
  __shared__ int is_last;
  if (use_barrier) signal_all_and_wait(pt, channel, 0, rank, ws, epoch, blockIdx.x == 0);

  float ss_part = 0.f, ss_repl = 0.f;
  const long long vstride = (long long)gridDim.x * blockDim.x;
  for (int s = 0; s < nseg; ++s) {
    const Segment sg = segs[s];
    const long long n = sg.partitioned ? sg.numel / ws : sg.numel;
    const long long src0 = sg.full_off + (sg.partitioned ? (long long)rank * n : 0);
    const long long nvec = n >> 3;
    float ss = 0.f;
    if (sg.partitioned) {
      // the round trip through the switch is longer than a peer load: keep NVLS_UNR requests in flight
      constexpr int NVLS_UNR = 8;
      const __nv_bfloat16* mc = reinterpret_cast<const __nv_bfloat16*>(mc_base) + src0;
      for (long long v0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; v0 < nvec; v0 += NVLS_UNR * vstride) {
        uint4 in[NVLS_UNR];
#pragma unroll
        for (int u = 0; u < NVLS_UNR; ++u) {
          const long long v = v0 + u * vstride;
          if (v < nvec) in[u] = multimem_ld_reduce_bf16x8(mc + v * 8);
        }
#pragma unroll
        for (int u = 0; u < NVLS_UNR; ++u) {
          const long long v = v0 + u * vstride;
          if (v >= nvec) continue;
          float acc[8];
          float2 t;
          t = unpack_bf16x2(in[u].x); acc[0] = t.x * scale; acc[1] = t.y * scale;
          t = unpack_bf16x2(in[u].y); acc[2] = t.x * scale; acc[3] = t.y * scale;
          t = unpack_bf16x2(in[u].z); acc[4] = t.x * scale; acc[5] = t.y * scale;
          t = unpack_bf16x2(in[u].w); acc[6] = t.x * scale; acc[7] = t.y * scale;
          ss += accumulate_shard<ACC_F32>(shard_grads, sg.shard_off + v * 8, acc);
        }
      }
      ss_part += ss;
    } else {
      for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += vstride) {
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        uint4 in[MAX_WS];
#pragma unroll
        for (int p = 0; p < MAX_WS; ++p)
          if (p < ws) in[p] = ld_peer_v4(reinterpret_cast<const __nv_bfloat16*>(pt.data[p]) + src0 + v * 8);
#pragma unroll
        for (int p = 0; p < MAX_WS; ++p) {   // fixed rank order: deterministic, identical on every rank
          if (p < ws) {
            float2 t;
            t = unpack_bf16x2(in[p].x); acc[0] += t.x; acc[1] += t.y;
            t = unpack_bf16x2(in[p].y); acc[2] += t.x; acc[3] += t.y;
            t = unpack_bf16x2(in[p].z); acc[4] += t.x; acc[5] += t.y;
            t = unpack_bf16x2(in[p].w); acc[6] += t.x; acc[7] += t.y;
          }
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] *= scale;
        ss += accumulate_shard<ACC_F32>(shard_grads, sg.shard_off + v * 8, acc);
      }
      ss_repl += ss;
    }
  }
  const float tp = block_sum(ss_part, red);
  const float tr = block_sum(ss_repl, red);
  if (threadIdx.x == 0) {
    sumsq_part[blockIdx.x] = tp;
    sumsq_repl[blockIdx.x] = tr;
    if (!(isfinite(tp) && isfinite(tr)) && nonfinite != nullptr) atomicOr(nonfinite, 1);
  }
  if (use_barrier) {   // phase 1, as in the point-to-point kernel
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      is_last = (atomicAdd(block_counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
      if (threadIdx.x == 0) *block_counter = 0;
      signal_all_and_wait(pt, channel, 1, rank, ws, epoch, true);
    }
  }
}


__global__ void fold_partials_kernel(const float* __restrict__ part, int P, float* __restrict__ out_part,
                                     float* __restrict__ out_repl, int accumulate) {
  __shared__ float red[32];
  float a = 0.f, b = 0.f;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    a += part[i];
    b += part[P + i];
  }
  a = block_sum(a, red);
  b = block_sum(b, red);
  if (threadIdx.x == 0) {
    out_part[0] = accumulate ? out_part[0] + a : a;
    out_repl[0] = accumulate ? out_repl[0] + b : b;
  }
}

// -------------------------------------------------------------------------------------------------
// K3: AdamW on the shard + push of the updated bf16 rows into every peer's parameter buffer
// -------------------------------------------------------------------------------------------------
template <bool GRAD_F32>
__global__ void __launch_bounds__(COMM_THREADS)
fused_adam_push_kernel(const __grid_constant__ PeerTable pt, const Segment* __restrict__ segs, int nseg,
                       float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                       void* __restrict__ shard_grads, __nv_bfloat16* __restrict__ shard_params, float lr,
                       float beta1, float beta2, float eps, float weight_decay, float bc1, float bc2,
                       const float* __restrict__ grad_scale_dev, const int* __restrict__ skip_flag,
                       unsigned int* block_counter, uint32_t epoch, int channel, int rank, int ws, int push,
                       int use_barrier, int zero_grads) {
  __shared__ int is_last;
  const bool skip = (skip_flag != nullptr && *skip_flag != 0);
  const float gs = grad_scale_dev != nullptr ? *grad_scale_dev : 1.f;
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const float decay = 1.f - lr * weight_decay;
  for (int s = 0; s < nseg; ++s) {
    const Segment sg = segs[s];
    const long long n = sg.partitioned ? sg.numel / ws : sg.numel;
    const long long dst0 = sg.full_off + (sg.partitioned ? (long long)rank * n : 0);
    const long long nvec = n >> 3;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
      const long long o = sg.shard_off + i * 8;
      float p[8], mm[8], vv[8], g[8];
      *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(master + o);
      *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(master + o + 4);
      if (!skip) {
        *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(m + o);
        *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(m + o + 4);
        *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(v + o);
        *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(v + o + 4);
        if constexpr (GRAD_F32) {
          const float* gp = reinterpret_cast<const float*>(shard_grads) + o;
          *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(gp);
          *reinterpret_cast<float4*>(g + 4) = *reinterpret_cast<const float4*>(gp + 4);
        } else {
          const uint4 u = *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(shard_grads) + o);
          float2 t;
          t = unpack_bf16x2(u.x); g[0] = t.x; g[1] = t.y;
          t = unpack_bf16x2(u.y); g[2] = t.x; g[3] = t.y;
          t = unpack_bf16x2(u.z); g[4] = t.x; g[5] = t.y;
          t = unpack_bf16x2(u.w); g[6] = t.x; g[7] = t.y;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float ge = g[e] * gs;
          mm[e] = beta1 * mm[e] + (1.f - beta1) * ge;
          vv[e] = beta2 * vv[e] + (1.f - beta2) * ge * ge;
          p[e] = p[e] * decay - step_size * (mm[e] / (sqrtf(vv[e]) * inv_sqrt_bc2 + eps));
        }
        *reinterpret_cast<float4*>(master + o) = *reinterpret_cast<float4*>(p);
        *reinterpret_cast<float4*>(master + o + 4) = *reinterpret_cast<float4*>(p + 4);
        *reinterpret_cast<float4*>(m + o) = *reinterpret_cast<float4*>(mm);
        *reinterpret_cast<float4*>(m + o + 4) = *reinterpret_cast<float4*>(mm + 4);
        *reinterpret_cast<float4*>(v + o) = *reinterpret_cast<float4*>(vv);
        *reinterpret_cast<float4*>(v + o + 4) = *reinterpret_cast<float4*>(vv + 4);
      }
      if (zero_grads) {
        if constexpr (GRAD_F32) {
          float* gp = reinterpret_cast<float*>(shard_grads) + o;
          *reinterpret_cast<float4*>(gp) = make_float4(0.f, 0.f, 0.f, 0.f);
          *reinterpret_cast<float4*>(gp + 4) = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(shard_grads) + o) = make_uint4(0, 0, 0, 0);
        }
      }
      uint4 w;
      w.x = pack_bf16x2(p[0], p[1]); w.y = pack_bf16x2(p[2], p[3]);
      w.z = pack_bf16x2(p[4], p[5]); w.w = pack_bf16x2(p[6], p[7]);
      if (shard_params != nullptr) *reinterpret_cast<uint4*>(shard_params + o) = w;
      if (push || sg.push) {
        if (sg.partitioned) {
#pragma unroll
          for (int q = 0; q < MAX_WS; ++q)   // own rows -> every rank's gathered parameter buffer
            if (q < ws) st_peer_v4(reinterpret_cast<__nv_bfloat16*>(pt.data[q]) + dst0 + i * 8, w);
        } else {
          *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(pt.data[rank]) + dst0 + i * 8) = w;
        }
      }
    }
  }
  if (use_barrier) {
    // phase 2: every rank's rows are final (and, when pushing, have landed everywhere) before any
    // rank's next gather / forward may read them
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();
      is_last = (atomicAdd(block_counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
      if (threadIdx.x == 0) *block_counter = 0;
      signal_all_and_wait(pt, channel, 2, rank, ws, epoch, true);
    }
  }
}

// -------------------------------------------------------------------------------------------------
// pull all-gather (ZeRO-3 baseline for K2 and the stage-3 gather of non-GEMM tensors):
// every rank copies each peer's rows of each partitioned tensor into its own gathered buffer.
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(COMM_THREADS)
pull_allgather_kernel(const __grid_constant__ PeerTable pt, const Segment* __restrict__ segs, int nseg,
                      __nv_bfloat16* __restrict__ dst_unit, int rank, int ws) {
  // pt.data[p] = peer p's SHARD parameter buffer (bf16); segments give shard/full offsets
  for (int s = 0; s < nseg; ++s) {
    const Segment sg = segs[s];
    if (!sg.partitioned) {
      const long long nvec = sg.numel >> 3;
      for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x)
        *reinterpret_cast<uint4*>(dst_unit + sg.full_off + i * 8) =
            *reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(pt.data[rank]) + sg.shard_off + i * 8);
      continue;
    }
    const long long n = sg.numel / ws;
    const long long nvec = n >> 3;
    const long long total = nvec * ws;
    const long long stride = (long long)gridDim.x * blockDim.x;
    constexpr int UNR = 4;  // four 128-bit peer loads in flight per thread
    for (long long j0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; j0 < total; j0 += stride * UNR) {
      uint4 val[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long j = j0 + u * stride;
        if (j < total) {
          // interleave peers so that all NVLink ports are busy at once; start at our own neighbour
          const int p = (int)((j + rank) % ws);
          const long long i = j / ws;
          val[u] = ld_peer_v4(reinterpret_cast<const __nv_bfloat16*>(pt.data[p]) + sg.shard_off + i * 8);
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long j = j0 + u * stride;
        if (j < total) {
          const int p = (int)((j + rank) % ws);
          const long long i = j / ws;
          *reinterpret_cast<uint4*>(dst_unit + sg.full_off + (long long)p * n + i * 8) = val[u];
        }
      }
    }
  }
}

// The communication kernels are meant to share SMs with a persistent GEMM CTA that needs ~200 KB of
// shared memory. An SM's L1/shared split can only change while the SM is idle, so a comm CTA that
// got there first with the default (L1-heavy) carveout would keep the GEMM CTA off that SM until it
// exits — and with a statically partitioned persistent GEMM that doubles the GEMM's run time
// (measured: 723 instead of 1400 TFLOP/s on the first GEMM of every backward block). Asking for
// the max-shared carveout makes the two kernels co-resident whichever arrives first.
template <typename K>
static void prefer_max_shared_carveout(K kernel) {
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}
static void comm_kernels_setup_once() {
  static bool done = false;
  if (done) return;
  prefer_max_shared_carveout(fused_reduce_scatter_kernel<true>);
  prefer_max_shared_carveout(fused_reduce_scatter_kernel<false>);
  prefer_max_shared_carveout(fused_reduce_scatter_nvls_kernel<true>);
  prefer_max_shared_carveout(fused_reduce_scatter_nvls_kernel<false>);
  prefer_max_shared_carveout(fused_adam_push_kernel<true>);
  prefer_max_shared_carveout(fused_adam_push_kernel<false>);
  prefer_max_shared_carveout(pull_allgather_kernel);
  prefer_max_shared_carveout(fold_partials_kernel);
  prefer_max_shared_carveout(comm_ready_kernel);
  done = true;
}

static int fill_table(PeerTable& pt, void* const* data_ptrs, void* const* signal_ptrs, int ws) {
  comm_kernels_setup_once();
  BZ_REQUIRE(ws >= 1 && ws <= MAX_WS, "fused comm: world size %d not in [1, %d]", ws, MAX_WS);
  for (int i = 0; i < MAX_WS; ++i) {
    pt.data[i] = i < ws ? data_ptrs[i] : nullptr;
    pt.signal[i] = (i < ws && signal_ptrs) ? reinterpret_cast<uint32_t*>(signal_ptrs[i]) : nullptr;
  }
  return 0;
}

}  // namespace bz

using namespace bz;

extern "C" {

int bz_comm_max_world(void) { return MAX_WS; }
int bz_comm_signal_pad_bytes(void) { return (SIG_CHANNELS * SIG_PHASES * MAX_WS + 64) * (int)sizeof(uint32_t); }
int bz_comm_segment_bytes(void) { return (int)sizeof(Segment); }
int bz_comm_max_ctas(void) { return 1024; }

// ---- peer-memory runtime (H2) ------------------------------------------------------------
int bz_symm_alloc(unsigned long long bytes, void** out) {
  BZ_CUDA_CHECK(cudaMalloc(out, bytes));
  BZ_CUDA_CHECK(cudaMemset(*out, 0, bytes));
  return 0;
}
int bz_symm_free(void* p) {
  BZ_CUDA_CHECK(cudaFree(p));
  return 0;
}
int bz_ipc_export(void* p, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  BZ_CUDA_CHECK(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}
int bz_ipc_import(const unsigned char* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  BZ_CUDA_CHECK(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int bz_ipc_close(void* p) {
  BZ_CUDA_CHECK(cudaIpcCloseMemHandle(p));
  return 0;
}
int bz_enable_peer_access(int peer_device) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  BZ_CUDA_CHECK(e);
  return 0;
}

// ---- K1 ----------------------------------------------------------------------------------
// grad_ptrs[ws]: every rank's unit gradient buffer (bf16), peer-mapped, index = rank.
// signal_ptrs[ws]: every rank's signal pad. segs: device array of `nseg` Segment records.
// workspace: >= 2*ctas floats + 1 uint32 (CTA partials + CTA counter), zero-initialised once.
// sumsq_part/sumsq_repl: device scalars accumulated (+=) with this unit's contribution.
int bz_fused_reduce_scatter(void* const* grad_ptrs, void* const* signal_ptrs, const void* segs, int nseg,
                            void* shard_grads, int shard_is_f32, float scale, float* workspace, float* sumsq_part,
                            float* sumsq_repl, int* nonfinite_flag, unsigned int epoch, int channel, int rank, int ws,
                            int use_barrier, int ctas, void* stream) {
  PeerTable pt;
  if (int rc = fill_table(pt, grad_ptrs, signal_ptrs, ws)) return rc;
  BZ_REQUIRE(channel >= 0 && channel < SIG_CHANNELS, "fused_reduce_scatter: channel %d out of range", channel);
  BZ_REQUIRE(!use_barrier || signal_ptrs != nullptr, "fused_reduce_scatter: barrier requested without signal pads");
  if (ctas <= 0) ctas = 64;
  if (ctas > 1024) ctas = 1024;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counter = reinterpret_cast<unsigned int*>(workspace + 2 * 1024);
  if (use_barrier == 3) {          // phase 0 in front of the reduce kernel instead of inside every one of its CTAs
    comm_ready_kernel<<<1, 32, 0, st>>>(pt, channel, rank, ws, epoch);
    BZ_CUDA_CHECK(cudaGetLastError());
    use_barrier = 2;
  }
  if (shard_is_f32)
    fused_reduce_scatter_kernel<true><<<ctas, COMM_THREADS, 0, st>>>(pt, (const Segment*)segs, nseg, shard_grads, scale,
                                                                     workspace, workspace + ctas, nonfinite_flag, counter,
                                                                     epoch, channel, rank, ws, use_barrier);
  else
    fused_reduce_scatter_kernel<false><<<ctas, COMM_THREADS, 0, st>>>(pt, (const Segment*)segs, nseg, shard_grads, scale,
                                                                      workspace, workspace + ctas, nonfinite_flag, counter,
                                                                      epoch, channel, rank, ws, use_barrier);
  BZ_CUDA_CHECK(cudaGetLastError());
  fold_partials_kernel<<<1, 256, 0, st>>>(workspace, ctas, sumsq_part, sumsq_repl, 1);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// NVLS variant (opt-in): grad_mc is the multicast mapping of the same gradient buffer on all ranks.
int bz_fused_reduce_scatter_mc(void* const* grad_ptrs, void* const* signal_ptrs, void* grad_mc, const void* segs,
                               int nseg, void* shard_grads, int shard_is_f32, float scale, float* workspace,
                               float* sumsq_part, float* sumsq_repl, int* nonfinite_flag, unsigned int epoch,
                               int channel, int rank, int ws, int use_barrier, int ctas, void* stream) {
  PeerTable pt;
  if (int rc = fill_table(pt, grad_ptrs, signal_ptrs, ws)) return rc;
  BZ_REQUIRE(grad_mc != nullptr, "fused_reduce_scatter_mc: null multicast pointer");
  BZ_REQUIRE(channel >= 0 && channel < SIG_CHANNELS, "fused_reduce_scatter_mc: channel %d out of range", channel);
  BZ_REQUIRE(!use_barrier || signal_ptrs != nullptr, "fused_reduce_scatter_mc: barrier requested without signal pads");
  if (ctas <= 0) ctas = 32;
  if (ctas > 1024) ctas = 1024;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned int* counter = reinterpret_cast<unsigned int*>(workspace + 2 * 1024);
  if (shard_is_f32)
    fused_reduce_scatter_nvls_kernel<true><<<ctas, COMM_THREADS, 0, st>>>(
        pt, grad_mc, (const Segment*)segs, nseg, shard_grads, scale, workspace, workspace + ctas, nonfinite_flag,
        counter, epoch, channel, rank, ws, use_barrier);
  else
    fused_reduce_scatter_nvls_kernel<false><<<ctas, COMM_THREADS, 0, st>>>(
        pt, grad_mc, (const Segment*)segs, nseg, shard_grads, scale, workspace, workspace + ctas, nonfinite_flag,
        counter, epoch, channel, rank, ws, use_barrier);
  BZ_CUDA_CHECK(cudaGetLastError());
  fold_partials_kernel<<<1, 256, 0, st>>>(workspace, ctas, sumsq_part, sumsq_repl, 1);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int bz_fused_comm_workspace_bytes(void) { return (2 * 1024 + 16) * (int)sizeof(float); }

// ---- K3 ----------------------------------------------------------------------------------
// param_ptrs[ws]: every rank's gathered parameter buffer for this unit (bf16), peer-mapped; used
// when push != 0. shard_params (may be NULL): local bf16 copy of the shard (ZeRO-3).
int bz_fused_adam_push(void* const* param_ptrs, void* const* signal_ptrs, const void* segs, int nseg, float* master,
                       float* m, float* v, void* shard_grads, int grads_are_f32, void* shard_params, float lr,
                       float beta1, float beta2, float eps, float weight_decay, int step,
                       const float* grad_scale_dev, const int* skip_flag, float* workspace, unsigned int epoch,
                       int channel, int rank, int ws, int push, int use_barrier, int zero_grads, int ctas,
                       void* stream) {
  PeerTable pt;
  void* nulls[MAX_WS] = {nullptr};
  // the table is also needed when only some segments publish their rows (Segment::push) in a launch with push == 0
  BZ_REQUIRE(!push || param_ptrs != nullptr, "fused_adam_push: push requested without parameter buffers");
  if (int rc = fill_table(pt, param_ptrs != nullptr ? param_ptrs : nulls, signal_ptrs, ws)) return rc;
  BZ_REQUIRE(step >= 1, "fused_adam_push: step must be >= 1");
  BZ_REQUIRE(channel >= 0 && channel < SIG_CHANNELS, "fused_adam_push: channel %d out of range", channel);
  if (ctas <= 0) ctas = 128;
  if (ctas > 1024) ctas = 1024;
  const float bc1 = 1.f - powf(beta1, (float)step), bc2 = 1.f - powf(beta2, (float)step);
  unsigned int* counter = reinterpret_cast<unsigned int*>(workspace + 2 * 1024) + 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (grads_are_f32)
    fused_adam_push_kernel<true><<<ctas, COMM_THREADS, 0, st>>>(
        pt, (const Segment*)segs, nseg, master, m, v, shard_grads, (__nv_bfloat16*)shard_params, lr, beta1, beta2, eps,
        weight_decay, bc1, bc2, grad_scale_dev, skip_flag, counter, epoch, channel, rank, ws, push, use_barrier, zero_grads);
  else
    fused_adam_push_kernel<false><<<ctas, COMM_THREADS, 0, st>>>(
        pt, (const Segment*)segs, nseg, master, m, v, shard_grads, (__nv_bfloat16*)shard_params, lr, beta1, beta2, eps,
        weight_decay, bc1, bc2, grad_scale_dev, skip_flag, counter, epoch, channel, rank, ws, push, use_barrier, zero_grads);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// ---- pull all-gather ---------------------------------------------------------------------
int bz_pull_allgather(void* const* shard_param_ptrs, const void* segs, int nseg, void* dst_unit, int rank, int ws,
                      int ctas, void* stream) {
  PeerTable pt;
  if (int rc = fill_table(pt, shard_param_ptrs, nullptr, ws)) return rc;
  if (ctas <= 0) ctas = 64;
  pull_allgather_kernel<<<ctas, COMM_THREADS, 0, (cudaStream_t)stream>>>(pt, (const Segment*)segs, nseg,
                                                                       (__nv_bfloat16*)dst_unit, rank, ws);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
