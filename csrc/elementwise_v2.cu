// K5, second generation (opt-in: B200ZERO_RMSNORM=2): the RMSNorm kernels of elementwise.cu with the loads of
// the NEXT row issued before the reductions of the current one.
//
// Why: inside a full training step rmsnorm_bwd ran at 36 % of the HBM roofline (114 us for 268 MB) and rmsnorm_fwd at
// 66 %. The backward keeps its per-column dw partial sums in registers, which limits it to two 256-thread CTAs per
// SM; each thread has only 4 x 16 B in flight per row and nothing at all during the two block reductions, i.e.
// ~32 KB per SM against the ~35 KB per SM that 6.5 TB/s x ~800 ns of latency ask for. Holding the next row's
// operands as packed bf16 (4 registers per 16 B) doubles the bytes in flight and keeps them in flight across the
// reductions for 24 more registers per thread. Arithmetic and its order are unchanged: results are bit-identical
// to the v1 kernels (the GPU test asserts that).
//
// Validated on B200 (bit-identical to the first generation where promised) and the default since round 2.
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int E2_THREADS = 256;

__device__ __forceinline__ void e2_unpack8(const uint4 u, float (&f)[8]) {
  float2 t;
  t = unpack_bf16x2(u.x); f[0] = t.x; f[1] = t.y;
  t = unpack_bf16x2(u.y); f[2] = t.x; f[3] = t.y;
  t = unpack_bf16x2(u.z); f[4] = t.x; f[5] = t.y;
  t = unpack_bf16x2(u.w); f[6] = t.x; f[7] = t.y;
}
__device__ __forceinline__ uint4 e2_pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]);
  u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]);
  u.w = pack_bf16x2(f[6], f[7]);
  return u;
}
__device__ __forceinline__ uint4 e2_ldg(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }

// ---------------------------------------------------------------------------------
// forward: xo = x + delta (optional) ; y = xo * rstd * w
// ---------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(E2_THREADS)
rmsnorm_fwd_v2_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ delta,
                      const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ xo,
                      __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int T, int D, float eps) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  float wv[VPT][8];
  uint4 nx[VPT], nd[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * E2_THREADS;
    if (vi < nvec) e2_unpack8(e2_ldg(w + vi * 8), wv[i]);
  }
  auto issue = [&](int row) {
    const size_t base = (size_t)row * D;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        nx[i] = e2_ldg(x + base + vi * 8);
        if (delta != nullptr) nd[i] = e2_ldg(delta + base + vi * 8);
      }
    }
  };
  int row = blockIdx.x;
  if (row < T) issue(row);
  for (; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    float vals[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        e2_unpack8(nx[i], vals[i]);
        if (delta != nullptr) {
          float d[8];
          e2_unpack8(nd[i], d);
#pragma unroll
          for (int e = 0; e < 8; ++e) vals[i][e] += d[e];
          // round through bf16 so the statistics match what backward will re-read
#pragma unroll
          for (int e = 0; e < 8; ++e) vals[i][e] = __bfloat162float(__float2bfloat16_rn(vals[i][e]));
          *reinterpret_cast<uint4*>(xo + base + vi * 8) = e2_pack8(vals[i]);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += vals[i][e] * vals[i][e];
      }
    }
    if (row + (int)gridDim.x < T) issue(row + gridDim.x);     // next row's loads fly during the reduction below
    const float tot = block_sum(ss, red);
    const float rstd = rsqrtf(tot / (float)D + eps);
    if (threadIdx.x == 0 && rstd_out != nullptr) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = vals[i][e] * rstd * wv[i][e];
        *reinterpret_cast<uint4*>(y + base + vi * 8) = e2_pack8(o);
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// backward: xhat = x*rstd, g = dy*w:  dx = dres + rstd * (g - xhat * mean(g*xhat)) ;  dw[c] = sum_rows dy*xhat
// ---------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(E2_THREADS)
rmsnorm_bwd_v2_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                      const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dy,
                      const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                      float* __restrict__ dw_partial, int T, int D, float eps) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  float dwacc[VPT][8];
  float wv[VPT][8];
  uint4 nx[VPT], ng[VPT], nr[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * E2_THREADS;
#pragma unroll
    for (int e = 0; e < 8; ++e) dwacc[i][e] = 0.f;
    if (vi < nvec) e2_unpack8(e2_ldg(w + vi * 8), wv[i]);
  }
  auto issue = [&](int row) {
    const size_t base = (size_t)row * D;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        nx[i] = e2_ldg(x + base + vi * 8);
        ng[i] = e2_ldg(dy + base + vi * 8);
        if (dres != nullptr) nr[i] = e2_ldg(dres + base + vi * 8);
      }
    }
  };
  int row = blockIdx.x;
  if (row < T) issue(row);
  for (; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    float xv[VPT][8], gv[VPT][8];
    uint4 cr[VPT];
    float dot = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        e2_unpack8(nx[i], xv[i]);
        e2_unpack8(ng[i], gv[i]);
        cr[i] = nr[i];
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += xv[i][e] * xv[i][e];
      }
    }
    if (row + (int)gridDim.x < T) issue(row + gridDim.x);     // next row's loads fly during the reductions below
    float rstd;
    if (rstd_in != nullptr) {
      rstd = rstd_in[row];
    } else {
      rstd = rsqrtf(block_sum(ss, red) / (float)D + eps);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xhat = xv[i][e] * rstd;
          dwacc[i][e] += gv[i][e] * xhat;  // dy * xhat
          gv[i][e] *= wv[i][e];            // g = dy * w
          dot += gv[i][e] * xhat;
          xv[i][e] = xhat;
        }
      }
    }
    const float mean_gx = block_sum(dot, red) / (float)D;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        float o[8];
        if (dres != nullptr) e2_unpack8(cr[i], o);
        else {
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += rstd * (gv[i][e] - xv[i][e] * mean_gx);
        *reinterpret_cast<uint4*>(dx + base + vi * 8) = e2_pack8(o);
      }
    }
  }
  if (dw_partial != nullptr) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * E2_THREADS;
      if (vi < nvec) {
        float* dst = dw_partial + (size_t)blockIdx.x * D + vi * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
      }
    }
  }
}

__global__ void e2_dw_reduce_kernel(const float* __restrict__ partial, int P, int D, void* __restrict__ dw, int dw_is_f32,
                                    int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  float s = 0.f;
  for (int p = 0; p < P; ++p) s += partial[(size_t)p * D + c];
  if (dw_is_f32) {
    float* o = reinterpret_cast<float*>(dw);
    o[c] = accumulate ? o[c] + s : s;
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(dw);
    o[c] = __float2bfloat16_rn(accumulate ? __bfloat162float(o[c]) + s : s);
  }
}

// ---------------------------------------------------------------------------------
// SwiGLU v2 on the fused [T, 2F] gate|up buffer. v1 walks a flat item index and pays two 64-bit integer
// divisions per 8-element vector (~100 instructions) plus an IEEE division per element; at 117 M elements per
// call that is more ALU time than the 108 us the memory traffic needs, which is why it was measured at 52 % of the
// HBM roofline. v2 indexes by (row, vector) with no division and uses the correctly rounded reciprocal
// (__frcp_rn(x) == 1.f / x bit for bit), so results are identical to v1.
// ---------------------------------------------------------------------------------
template <bool FAST>   // FAST: rcp.approx (2 ulp) instead of the correctly rounded reciprocal — not bit-identical to v1
__global__ void __launch_bounds__(E2_THREADS)
swiglu_fwd_v2_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act, int T, int F) {
  const int vecs = F >> 3;
  constexpr int UNR = 4;
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const __nv_bfloat16* grow = gu + (size_t)t * 2 * F;
    __nv_bfloat16* arow = act + (size_t)t * F;
    for (int v0 = threadIdx.x; v0 < vecs; v0 += E2_THREADS * UNR) {
      uint4 g[UNR], u[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int v = v0 + k * E2_THREADS;
        if (v < vecs) {
          g[k] = e2_ldg(grow + v * 8);
          u[k] = e2_ldg(grow + F + v * 8);
        }
      }
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int v = v0 + k * E2_THREADS;
        if (v < vecs) {
          float gf[8], uf[8], o[8];
          e2_unpack8(g[k], gf);
          e2_unpack8(u[k], uf);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float den = 1.f + __expf(-gf[e]);
            const float sig = FAST ? __fdividef(1.f, den) : __frcp_rn(den);
            o[e] = gf[e] * sig * uf[e];
          }
          *reinterpret_cast<uint4*>(arow + v * 8) = e2_pack8(o);
        }
      }
    }
  }
}

template <bool FAST>
__global__ void __launch_bounds__(E2_THREADS)
swiglu_bwd_v2_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dact,
                     __nv_bfloat16* __restrict__ dgu, int T, int F) {
  const int vecs = F >> 3;
  constexpr int UNR = 2;
  for (int t = blockIdx.x; t < T; t += gridDim.x) {
    const __nv_bfloat16* grow = gu + (size_t)t * 2 * F;
    const __nv_bfloat16* drow = dact + (size_t)t * F;
    __nv_bfloat16* orow = dgu + (size_t)t * 2 * F;
    for (int v0 = threadIdx.x; v0 < vecs; v0 += E2_THREADS * UNR) {
      uint4 g[UNR], u[UNR], d[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int v = v0 + k * E2_THREADS;
        if (v < vecs) {
          g[k] = e2_ldg(grow + v * 8);
          u[k] = e2_ldg(grow + F + v * 8);
          d[k] = e2_ldg(drow + v * 8);
        }
      }
#pragma unroll
      for (int k = 0; k < UNR; ++k) {
        const int v = v0 + k * E2_THREADS;
        if (v < vecs) {
          float gf[8], uf[8], df[8], dg[8], du[8];
          e2_unpack8(g[k], gf);
          e2_unpack8(u[k], uf);
          e2_unpack8(d[k], df);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float den = 1.f + __expf(-gf[e]);
            const float sig = FAST ? __fdividef(1.f, den) : __frcp_rn(den);
            dg[e] = df[e] * uf[e] * sig * (1.f + gf[e] * (1.f - sig));
            du[e] = df[e] * gf[e] * sig;
          }
          *reinterpret_cast<uint4*>(orow + v * 8) = e2_pack8(dg);
          *reinterpret_cast<uint4*>(orow + F + v * 8) = e2_pack8(du);
        }
      }
    }
  }
}

static inline int e2_vpt_for(int D) {
  const int nvec = D >> 3;
  if (nvec <= E2_THREADS) return 1;
  if (nvec <= 2 * E2_THREADS) return 2;
  if (nvec <= 4 * E2_THREADS) return 4;
  return 0;
}

}  // namespace bz

using namespace bz;

// Same contracts as bz_rmsnorm_fwd / bz_rmsnorm_bwd (including the workspace size of bz_rmsnorm_bwd_workspace_floats).
extern "C" int bz_rmsnorm_fwd_v2(const void* x, const void* delta, const void* w, void* x_out, void* y, float* rstd,
                                 int T, int D, float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "rmsnorm_fwd_v2: D=%d must be a multiple of 8", D);
  BZ_REQUIRE(delta == nullptr || x_out != nullptr, "rmsnorm_fwd_v2: delta given without x_out");
  const int vpt = e2_vpt_for(D);
  BZ_REQUIRE(vpt != 0, "rmsnorm_fwd_v2: D=%d too large (max %d)", D, 8 * 4 * E2_THREADS);
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = min(T, device_sm_count() * 8);
  auto X = (const __nv_bfloat16*)x; auto DL = (const __nv_bfloat16*)delta; auto W = (const __nv_bfloat16*)w;
  auto XO = (__nv_bfloat16*)x_out; auto Y = (__nv_bfloat16*)y;
  if (vpt == 1) rmsnorm_fwd_v2_kernel<1><<<grid, E2_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  else if (vpt == 2) rmsnorm_fwd_v2_kernel<2><<<grid, E2_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  else rmsnorm_fwd_v2_kernel<4><<<grid, E2_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_rmsnorm_bwd_v2(const void* x, const void* w, const float* rstd, const void* dy, const void* dres,
                                 void* dx, void* dw, int dw_is_f32, int dw_accumulate, float* workspace, int T, int D,
                                 float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "rmsnorm_bwd_v2: D=%d must be a multiple of 8", D);
  const int vpt = e2_vpt_for(D);
  BZ_REQUIRE(vpt != 0, "rmsnorm_bwd_v2: D=%d too large", D);
  BZ_REQUIRE(dw == nullptr || workspace != nullptr, "rmsnorm_bwd_v2: dw requested without workspace");
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = min(T, device_sm_count() * 2);          // same grid as v1: the partial-sum layout is shared
  float* part = dw ? workspace : nullptr;
  auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w; auto DY = (const __nv_bfloat16*)dy;
  auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
  if (vpt == 1) rmsnorm_bwd_v2_kernel<1><<<grid, E2_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else if (vpt == 2) rmsnorm_bwd_v2_kernel<2><<<grid, E2_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else rmsnorm_bwd_v2_kernel<4><<<grid, E2_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  BZ_CUDA_CHECK(cudaGetLastError());
  if (dw) {
    e2_dw_reduce_kernel<<<(D + 255) / 256, 256, 0, st>>>(part, grid, D, dw, dw_is_f32, dw_accumulate);
    BZ_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

// fast != 0 selects rcp.approx for the sigmoid (the forward is instruction-issue-bound with the exact reciprocal:
// ~244 instructions per 8-element vector against the ~209 that fit under its memory time).
extern "C" int bz_swiglu_fwd_v2(const void* gate_up, void* act, long long T, int F, int fast, void* stream) {
  BZ_REQUIRE(F % 8 == 0, "swiglu_fwd_v2: F=%d must be a multiple of 8", F);
  BZ_REQUIRE(T < (1LL << 31), "swiglu_fwd_v2: T too large");
  if (T <= 0) return 0;
  const int grid = (int)min((long long)device_sm_count() * 8, T);
  cudaStream_t st = (cudaStream_t)stream;
  if (fast) swiglu_fwd_v2_kernel<true><<<grid, E2_THREADS, 0, st>>>((const __nv_bfloat16*)gate_up, (__nv_bfloat16*)act, (int)T, F);
  else swiglu_fwd_v2_kernel<false><<<grid, E2_THREADS, 0, st>>>((const __nv_bfloat16*)gate_up, (__nv_bfloat16*)act, (int)T, F);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_swiglu_bwd_v2(const void* gate_up, const void* dact, void* dgate_up, long long T, int F, int fast,
                                void* stream) {
  BZ_REQUIRE(F % 8 == 0, "swiglu_bwd_v2: F=%d must be a multiple of 8", F);
  BZ_REQUIRE(T < (1LL << 31), "swiglu_bwd_v2: T too large");
  if (T <= 0) return 0;
  const int grid = (int)min((long long)device_sm_count() * 8, T);
  cudaStream_t st = (cudaStream_t)stream;
  auto G = (const __nv_bfloat16*)gate_up; auto D = (const __nv_bfloat16*)dact; auto O = (__nv_bfloat16*)dgate_up;
  if (fast) swiglu_bwd_v2_kernel<true><<<grid, E2_THREADS, 0, st>>>(G, D, O, (int)T, F);
  else swiglu_bwd_v2_kernel<false><<<grid, E2_THREADS, 0, st>>>(G, D, O, (int)T, F);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
