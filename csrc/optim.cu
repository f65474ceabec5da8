// K11 + F9/F10: fused AdamW on flat fp32 partitions with bf16 write-back,
// squared-norm + isfinite reduction, clip-coefficient computation on device, casts.
// The update is predicated on a device-side skip flag so a non-finite gradient detected
// in the same step never touches the master weights and no host round-trip is needed
// (spike-to-halt path). Replaces DeepSpeed's multi_tensor_adam.cu that the
// reference reaches only through its `optimizer` config block (ref deepspeed_launcher.py:156-164).
#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int OPT_THREADS = 256;
constexpr int MAX_PARTIALS = 1024;

template <bool GRAD_F32>
__global__ void __launch_bounds__(OPT_THREADS)
adamw_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v, const void* __restrict__ grad,
             __nv_bfloat16* __restrict__ param_out, long long n, float lr, float beta1, float beta2, float eps,
             float weight_decay, float bc1, float bc2, float grad_scale, const float* __restrict__ grad_scale_dev,
             const int* __restrict__ skip_flag) {
  if (skip_flag != nullptr && *skip_flag != 0) return;
  const float gs = grad_scale * (grad_scale_dev != nullptr ? *grad_scale_dev : 1.f);
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const float decay = 1.f - lr * weight_decay;
  const long long nvec = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float4 p = reinterpret_cast<float4*>(master)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float g[4];
    if constexpr (GRAD_F32) {
      const float4 gg = reinterpret_cast<const float4*>(grad)[i];
      g[0] = gg.x; g[1] = gg.y; g[2] = gg.z; g[3] = gg.w;
    } else {
      const uint2 gg = reinterpret_cast<const uint2*>(grad)[i];
      const float2 a = unpack_bf16x2(gg.x), b = unpack_bf16x2(gg.y);
      g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y;
    }
    float pa[4] = {p.x, p.y, p.z, p.w}, ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float ge = g[e] * gs;
      ma[e] = beta1 * ma[e] + (1.f - beta1) * ge;
      va[e] = beta2 * va[e] + (1.f - beta2) * ge * ge;
      const float denom = sqrtf(va[e]) * inv_sqrt_bc2 + eps;
      pa[e] = pa[e] * decay - step_size * (ma[e] / denom);
    }
    reinterpret_cast<float4*>(master)[i] = make_float4(pa[0], pa[1], pa[2], pa[3]);
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if (param_out != nullptr) {
      uint2 o;
      o.x = pack_bf16x2(pa[0], pa[1]);
      o.y = pack_bf16x2(pa[2], pa[3]);
      reinterpret_cast<uint2*>(param_out)[i] = o;
    }
  }
  // scalar tail (n % 4)
  const long long tail0 = nvec << 2;
  for (long long i = tail0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float ge = (GRAD_F32 ? reinterpret_cast<const float*>(grad)[i]
                         : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(grad)[i])) * gs;
    float mi = beta1 * m[i] + (1.f - beta1) * ge;
    float vi = beta2 * v[i] + (1.f - beta2) * ge * ge;
    float pi = master[i] * decay - step_size * (mi / (sqrtf(vi) * inv_sqrt_bc2 + eps));
    m[i] = mi; v[i] = vi; master[i] = pi;
    if (param_out != nullptr) param_out[i] = __float2bfloat16_rn(pi);
  }
}

// stage 1: per-block partial of sum(g^2) (+ non-finite detection); stage 2 folds partials.
template <bool F32>
__global__ void __launch_bounds__(OPT_THREADS)
sumsq_partial_kernel(const void* __restrict__ g, long long n, float* __restrict__ partials, int* __restrict__ nonfinite) {
  __shared__ float red[32];
  float acc = 0.f;
  bool bad = false;
  if constexpr (F32) {
    const long long nvec = n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
      const float4 x = reinterpret_cast<const float4*>(g)[i];
      acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
    }
    for (long long i = (nvec << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
      const float x = reinterpret_cast<const float*>(g)[i];
      acc += x * x;
    }
  } else {
    const long long nvec = n >> 3;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
      const uint4 u = reinterpret_cast<const uint4*>(g)[i];
      float2 t;
      t = unpack_bf16x2(u.x); acc += t.x * t.x + t.y * t.y;
      t = unpack_bf16x2(u.y); acc += t.x * t.x + t.y * t.y;
      t = unpack_bf16x2(u.z); acc += t.x * t.x + t.y * t.y;
      t = unpack_bf16x2(u.w); acc += t.x * t.x + t.y * t.y;
    }
    for (long long i = (nvec << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
      const float x = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[i]);
      acc += x * x;
    }
  }
  // a NaN/Inf anywhere poisons the sum of squares, so finiteness is checked on the partial
  const float tot = block_sum(acc, red);
  bad = !isfinite(tot);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = tot;
    if (bad && nonfinite != nullptr) atomicOr(nonfinite, 1);
  }
}

__global__ void __launch_bounds__(OPT_THREADS)
sumsq_final_kernel(const float* __restrict__ partials, int P, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < P; i += blockDim.x) acc += partials[i];
  const float tot = block_sum(acc, red);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + tot : tot;
}

// coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)) * extra; max_norm <= 0 disables clipping.
// Also publishes the norm itself (for TrainingMetrics.gradient_norm) and raises the skip
// flag when the norm is not finite.
__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float extra, float* __restrict__ coef,
                                 float* __restrict__ norm_out, int* __restrict__ nonfinite) {
  const float ss = *sumsq;
  const float nrm = sqrtf(ss);
  if (norm_out != nullptr) *norm_out = nrm;
  if (!isfinite(nrm)) {
    if (nonfinite != nullptr) atomicOr(nonfinite, 1);
    *coef = 0.f;
    return;
  }
  float c = 1.f;
  if (max_norm > 0.f) c = fminf(1.f, max_norm / (nrm + 1e-6f));
  *coef = c * extra;
}

__global__ void __launch_bounds__(OPT_THREADS)
cast_f32_to_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  const long long nvec = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    const float4 x = reinterpret_cast<const float4*>(in)[i];
    uint2 o;
    o.x = pack_bf16x2(x.x, x.y);
    o.y = pack_bf16x2(x.z, x.w);
    reinterpret_cast<uint2*>(out)[i] = o;
  }
  for (long long i = (nvec << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16_rn(in[i]);
}

__global__ void __launch_bounds__(OPT_THREADS)
cast_bf16_to_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = __bfloat162float(in[i]);
}

static inline int opt_grid(long long n, int per_thread) {
  long long blocks = (n / per_thread + OPT_THREADS - 1) / OPT_THREADS;
  const long long cap = (long long)device_sm_count() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

int sumsq_launch(const void* g, int is_f32, long long n, float* partials, float* out, int accumulate, int* nonfinite,
                 cudaStream_t st) {
  int P = opt_grid(n, is_f32 ? 4 : 8);
  if (P > MAX_PARTIALS) P = MAX_PARTIALS;
  if (is_f32) sumsq_partial_kernel<true><<<P, OPT_THREADS, 0, st>>>(g, n, partials, nonfinite);
  else sumsq_partial_kernel<false><<<P, OPT_THREADS, 0, st>>>(g, n, partials, nonfinite);
  BZ_CUDA_CHECK(cudaGetLastError());
  sumsq_final_kernel<<<1, OPT_THREADS, 0, st>>>(partials, P, out, accumulate);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace bz

using namespace bz;

extern "C" int bz_adamw(float* master, float* m, float* v, const void* grad, int grad_is_f32, void* param_bf16,
                        long long n, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                        float grad_scale, const float* grad_scale_dev, const int* skip_flag, void* stream) {
  BZ_REQUIRE(step >= 1, "adamw: step must be >= 1 (got %d)", step);
  if (n <= 0) return 0;
  BZ_REQUIRE(((uintptr_t)master & 15) == 0 && ((uintptr_t)m & 15) == 0 && ((uintptr_t)v & 15) == 0 &&
                 ((uintptr_t)grad & 15) == 0 && ((uintptr_t)param_bf16 & 7) == 0,
             "adamw: buffers must be 16-byte aligned");
  const float bc1 = 1.f - powf(beta1, (float)step);
  const float bc2 = 1.f - powf(beta2, (float)step);
  const int grid = opt_grid(n, 4);
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_is_f32)
    adamw_kernel<true><<<grid, OPT_THREADS, 0, st>>>(master, m, v, grad, (__nv_bfloat16*)param_bf16, n, lr, beta1, beta2,
                                                     eps, weight_decay, bc1, bc2, grad_scale, grad_scale_dev, skip_flag);
  else
    adamw_kernel<false><<<grid, OPT_THREADS, 0, st>>>(master, m, v, grad, (__nv_bfloat16*)param_bf16, n, lr, beta1, beta2,
                                                      eps, weight_decay, bc1, bc2, grad_scale, grad_scale_dev, skip_flag);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_sumsq_workspace_floats(void) { return MAX_PARTIALS; }

extern "C" int bz_sumsq(const void* g, int g_is_f32, long long n, float* workspace, float* out, int accumulate,
                        int* nonfinite_flag, void* stream) {
  BZ_REQUIRE(workspace != nullptr && out != nullptr, "sumsq: workspace and out are required");
  BZ_REQUIRE(((uintptr_t)g & 15) == 0, "sumsq: input must be 16-byte aligned");
  if (n <= 0) return 0;
  return sumsq_launch(g, g_is_f32, n, workspace, out, accumulate, nonfinite_flag, (cudaStream_t)stream);
}

extern "C" int bz_clip_coef(const float* sumsq, float max_norm, float extra_scale, float* coef_out, float* norm_out,
                            int* nonfinite_flag, void* stream) {
  clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sumsq, max_norm, extra_scale, coef_out, norm_out, nonfinite_flag);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream) {
  if (n <= 0) return 0;
  cast_f32_to_bf16_kernel<<<opt_grid(n, 4), OPT_THREADS, 0, (cudaStream_t)stream>>>(in, (__nv_bfloat16*)out, n);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_cast_bf16_to_f32(const void* in, float* out, long long n, void* stream) {
  if (n <= 0) return 0;
  cast_bf16_to_f32_kernel<<<opt_grid(n, 1), OPT_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, out, n);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
