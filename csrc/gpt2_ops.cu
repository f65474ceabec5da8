// F3 / F5, GPT-2 flavour: LayerNorm forward/backward and bias + GELU(tanh)
// forward/backward for the GPT-2 family (BASELINE.json configs 1-2). Same recipe as the Llama
// element-wise kernels (csrc/elementwise.cu): one CTA walks rows with a grid stride, 128-bit
// vector accesses, fp32 statistics, per-column gradient partials kept in registers across the
// CTA's rows and folded by a second tiny kernel — no atomics, deterministic.
// The reference has no kernels; these replace torch-eager LayerNorm / GELU in a DeepSpeed user script.
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int G2_THREADS = 256;

__device__ __forceinline__ void g2_load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 t;
  t = unpack_bf16x2(u.x); f[0] = t.x; f[1] = t.y;
  t = unpack_bf16x2(u.y); f[2] = t.x; f[3] = t.y;
  t = unpack_bf16x2(u.z); f[4] = t.x; f[5] = t.y;
  t = unpack_bf16x2(u.w); f[6] = t.x; f[7] = t.y;
}
__device__ __forceinline__ void g2_store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]);
  u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]);
  u.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ---------------------------------------------------------------------------------
// LayerNorm forward: y = (x - mean) * rstd * w + b ; mean / rstd saved (fp32) for backward.
// Two-pass statistics on register-cached values (mean first, then centred sum of squares).
// ---------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(G2_THREADS)
layernorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                     const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ y, float* __restrict__ mean_out,
                     float* __restrict__ rstd_out, int T, int D, float eps) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    float vals[VPT][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
        g2_load8(x + base + vi * 8, vals[i]);
#pragma unroll
        for (int e = 0; e < 8; ++e) s += vals[i][e];
      }
    }
    const float mean = block_sum(s, red) / (float)D;
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          vals[i][e] -= mean;
          ss += vals[i][e] * vals[i][e];
        }
      }
    }
    const float rstd = rsqrtf(block_sum(ss, red) / (float)D + eps);
    if (threadIdx.x == 0) {
      if (mean_out != nullptr) mean_out[row] = mean;
      if (rstd_out != nullptr) rstd_out[row] = rstd;
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
        float wv[8], bv[8], o[8];
        g2_load8(w + vi * 8, wv);
        g2_load8(b + vi * 8, bv);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = vals[i][e] * rstd * wv[e] + bv[e];
        g2_store8(y + base + vi * 8, o);
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// LayerNorm backward. xhat = (x - mean) * rstd, g = dy * w:
//   dx = rstd * (g - mean(g) - xhat * mean(g * xhat))     dw[c] = sum_rows dy * xhat     db[c] = sum_rows dy
// partial layout: [gridDim.x][2][D] fp32 (dw then db).
// ---------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(G2_THREADS)
layernorm_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                     const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                     const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                     int T, int D) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  float dwacc[VPT][8], dbacc[VPT][8], wv[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * G2_THREADS;
#pragma unroll
    for (int e = 0; e < 8; ++e) dwacc[i][e] = dbacc[i][e] = 0.f;
    if (vi < nvec) g2_load8(w + vi * 8, wv[i]);
  }
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    const float mean = mean_in[row], rstd = rstd_in[row];
    float xh[VPT][8], gv[VPT][8];
    float sg = 0.f, sgx = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
        g2_load8(x + base + vi * 8, xh[i]);
        g2_load8(dy + base + vi * 8, gv[i]);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float xhat = (xh[i][e] - mean) * rstd;
          dwacc[i][e] += gv[i][e] * xhat;
          dbacc[i][e] += gv[i][e];
          gv[i][e] *= wv[i][e];
          sg += gv[i][e];
          sgx += gv[i][e] * xhat;
          xh[i][e] = xhat;
        }
      }
    }
    const float mg = block_sum(sg, red) / (float)D;
    const float mgx = block_sum(sgx, red) / (float)D;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = rstd * (gv[i][e] - mg - xh[i][e] * mgx);
        g2_store8(dx + base + vi * 8, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * G2_THREADS;
    if (vi < nvec) {
      float* dst = partial + (size_t)blockIdx.x * 2 * D + vi * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
      dst += D;
      *reinterpret_cast<float4*>(dst) = make_float4(dbacc[i][0], dbacc[i][1], dbacc[i][2], dbacc[i][3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(dbacc[i][4], dbacc[i][5], dbacc[i][6], dbacc[i][7]);
    }
  }
}

// out[k][c] (+)= sum_p partial[p][k][c] for k < K; outputs bf16 or fp32.
__global__ void g2_fold_partials_kernel(const float* __restrict__ partial, int P, int K, int D, void* const out0,
                                        void* const out1, int out_is_f32, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  for (int k = 0; k < K; ++k) {
    void* out = k == 0 ? out0 : out1;
    if (out == nullptr) continue;
    float s = 0.f;
    for (int p = 0; p < P; ++p) s += partial[((size_t)p * K + k) * D + c];
    if (out_is_f32) {
      float* o = reinterpret_cast<float*>(out);
      o[c] = accumulate ? o[c] + s : s;
    } else {
      __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(out);
      o[c] = __float2bfloat16_rn(accumulate ? __bfloat162float(o[c]) + s : s);
    }
  }
}

// ---------------------------------------------------------------------------------
// bias + GELU (tanh approximation, GPT-2's "gelu_new"):  y = gelu(x + b)
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float gelu_tanh(float v) {
  const float k = 0.7978845608028654f, c = 0.044715f;
  return 0.5f * v * (1.f + tanhf(k * (v + c * v * v * v)));
}
__device__ __forceinline__ float gelu_tanh_grad(float v) {
  const float k = 0.7978845608028654f, c = 0.044715f;
  const float t = tanhf(k * (v + c * v * v * v));
  return 0.5f * (1.f + t) + 0.5f * v * (1.f - t * t) * k * (1.f + 3.f * c * v * v);
}

__global__ void __launch_bounds__(G2_THREADS)
bias_gelu_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ b,
                     __nv_bfloat16* __restrict__ y, int T, int F) {
  const int nvec = F >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {     // (row, vector) indexing: no integer division per vector
    const size_t base = (size_t)row * F;
    for (int vc = threadIdx.x; vc < nvec; vc += G2_THREADS) {
      float xv[8], bv[8], o[8];
      g2_load8(x + base + vc * 8, xv);
      g2_load8(b + vc * 8, bv);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = gelu_tanh(xv[e] + bv[e]);
      g2_store8(y + base + vc * 8, o);
    }
  }
}

// dx = dy * gelu'(x + b); db[c] = sum_rows dx (partials [gridDim.x][1][F]).
template <int VPT>
__global__ void __launch_bounds__(G2_THREADS)
bias_gelu_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ b,
                     const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, float* __restrict__ partial,
                     int T, int F) {
  const int nvec = F >> 3;
  float dbacc[VPT][8], bv[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * G2_THREADS;
#pragma unroll
    for (int e = 0; e < 8; ++e) dbacc[i][e] = 0.f;
    if (vi < nvec) g2_load8(b + vi * 8, bv[i]);
  }
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * F;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * G2_THREADS;
      if (vi < nvec) {
        float xv[8], gv[8], o[8];
        g2_load8(x + base + vi * 8, xv);
        g2_load8(dy + base + vi * 8, gv);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          o[e] = gv[e] * gelu_tanh_grad(xv[e] + bv[i][e]);
          // the bias gradient sums what the next GEMM will actually read (bf16-rounded dx)
          dbacc[i][e] += __bfloat162float(__float2bfloat16_rn(o[e]));
        }
        g2_store8(dx + base + vi * 8, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * G2_THREADS;
    if (vi < nvec) {
      float* dst = partial + (size_t)blockIdx.x * F + vi * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(dbacc[i][0], dbacc[i][1], dbacc[i][2], dbacc[i][3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(dbacc[i][4], dbacc[i][5], dbacc[i][6], dbacc[i][7]);
    }
  }
}

static inline int g2_vpt_for(int D) {
  const int nvec = D >> 3;
  const int v = (nvec + G2_THREADS - 1) / G2_THREADS;
  return v <= 1 ? 1 : v <= 2 ? 2 : v <= 4 ? 4 : v <= 8 ? 8 : 0;
}
static inline int g2_row_grid(int T) {
  const int sms = device_sm_count();
  return T < 2 * sms ? (T < 1 ? 1 : T) : 2 * sms;
}

}  // namespace bz

using namespace bz;

#define G2_DISPATCH_VPT(vpt, CALL)                    \
  switch (vpt) {                                      \
    case 1: { constexpr int V = 1; CALL; } break;     \
    case 2: { constexpr int V = 2; CALL; } break;     \
    case 4: { constexpr int V = 4; CALL; } break;     \
    default: { constexpr int V = 8; CALL; } break;    \
  }

extern "C" int bz_layernorm_fwd(const void* x, const void* w, const void* b, void* y, float* mean, float* rstd, int T,
                                int D, float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0 && D > 0, "layernorm_fwd: D=%d must be a positive multiple of 8", D);
  const int vpt = g2_vpt_for(D);
  BZ_REQUIRE(vpt != 0, "layernorm_fwd: D=%d too large (max %d)", D, 8 * 8 * G2_THREADS);
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = g2_row_grid(T);
  G2_DISPATCH_VPT(vpt, (layernorm_fwd_kernel<V><<<grid, G2_THREADS, 0, st>>>(
                           (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (const __nv_bfloat16*)b, (__nv_bfloat16*)y,
                           mean, rstd, T, D, eps)));
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// workspace floats needed by bz_layernorm_bwd / bz_bias_gelu_bwd for T rows of width D
extern "C" long long bz_gpt2_bwd_workspace_floats(int T, int D) { return (long long)g2_row_grid(T) * 2 * D; }

// dw / db: bf16 or fp32 ([D]); accumulate != 0 adds into them (gradient accumulation).
extern "C" int bz_layernorm_bwd(const void* x, const void* w, const float* mean, const float* rstd, const void* dy,
                                void* dx, void* dw, void* db, int dwb_is_f32, int accumulate, float* workspace, int T,
                                int D, void* stream) {
  BZ_REQUIRE(D % 8 == 0 && D > 0, "layernorm_bwd: D=%d must be a positive multiple of 8", D);
  const int vpt = g2_vpt_for(D);
  BZ_REQUIRE(vpt != 0, "layernorm_bwd: D=%d too large", D);
  BZ_REQUIRE(workspace != nullptr && mean != nullptr && rstd != nullptr, "layernorm_bwd: null workspace / statistics");
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = g2_row_grid(T);
  G2_DISPATCH_VPT(vpt, (layernorm_bwd_kernel<V><<<grid, G2_THREADS, 0, st>>>(
                           (const __nv_bfloat16*)x, (const __nv_bfloat16*)w, mean, rstd, (const __nv_bfloat16*)dy,
                           (__nv_bfloat16*)dx, workspace, T, D)));
  BZ_CUDA_CHECK(cudaGetLastError());
  g2_fold_partials_kernel<<<(D + 255) / 256, 256, 0, st>>>(workspace, grid, 2, D, dw, db, dwb_is_f32, accumulate);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_bias_gelu_fwd(const void* x, const void* bias, void* y, long long T, int F, void* stream) {
  BZ_REQUIRE(F % 8 == 0 && F > 0, "bias_gelu_fwd: F=%d must be a positive multiple of 8", F);
  BZ_REQUIRE(T < (1LL << 31), "bias_gelu_fwd: T too large");
  if (T <= 0) return 0;
  const int grid = (int)min((long long)device_sm_count() * 8, T);
  bias_gelu_fwd_kernel<<<grid, G2_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, (int)T, F);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_bias_gelu_bwd(const void* x, const void* bias, const void* dy, void* dx, void* dbias, int db_is_f32,
                                int accumulate, float* workspace, int T, int F, void* stream) {
  BZ_REQUIRE(F % 8 == 0 && F > 0, "bias_gelu_bwd: F=%d must be a positive multiple of 8", F);
  const int vpt = g2_vpt_for(F);
  BZ_REQUIRE(vpt != 0, "bias_gelu_bwd: F=%d too large (max %d)", F, 8 * 8 * G2_THREADS);
  BZ_REQUIRE(workspace != nullptr, "bias_gelu_bwd: null workspace");
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = g2_row_grid(T);
  G2_DISPATCH_VPT(vpt, (bias_gelu_bwd_kernel<V><<<grid, G2_THREADS, 0, st>>>(
                           (const __nv_bfloat16*)x, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)dy,
                           (__nv_bfloat16*)dx, workspace, T, F)));
  BZ_CUDA_CHECK(cudaGetLastError());
  g2_fold_partials_kernel<<<(F + 255) / 256, 256, 0, st>>>(workspace, grid, 1, F, dbias, nullptr, db_is_f32, accumulate);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
