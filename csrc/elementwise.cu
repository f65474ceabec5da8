// K5/K7/K8/K10 + cross-entropy: bandwidth-bound kernels of the Llama block.
// All are 128-bit vectorised, keep fp32 statistics, and are launched with grids sized to
// a multiple of the SM count. The reference has no kernels; these replace
// the torch-eager ops a DeepSpeed user script would run (F3, F4, F5, F6, F11).
#include <mutex>

#include "common.cuh"
#include "../include/b200zero.h"

namespace bz {

constexpr int EW_THREADS = 256;

struct alignas(16) bf16x8 {
  uint4 v;
};

__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&f)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 t;
  t = unpack_bf16x2(u.x); f[0] = t.x; f[1] = t.y;
  t = unpack_bf16x2(u.y); f[2] = t.x; f[3] = t.y;
  t = unpack_bf16x2(u.z); f[4] = t.x; f[5] = t.y;
  t = unpack_bf16x2(u.w); f[6] = t.x; f[7] = t.y;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]);
  u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]);
  u.w = pack_bf16x2(f[6], f[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

// ---------------------------------------------------------------------------------
// RMSNorm forward, optionally fused with the residual add that precedes it:
//   xo = x + delta (written if delta != null) ; y = xo * rstd * w ; rstd saved for bwd
// One CTA per row-slot, rows strided over the grid. D % 8 == 0.
// ---------------------------------------------------------------------------------
template <int VPT>  // 8-element vectors cached per thread (D <= 8*VPT*EW_THREADS)
__global__ void __launch_bounds__(EW_THREADS)
rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ delta,
                   const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ xo,
                   __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int T, int D, float eps) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    float vals[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * EW_THREADS;
      if (vi < nvec) {
        load8(x + base + vi * 8, vals[i]);
        if (delta != nullptr) {
          float d[8];
          load8(delta + base + vi * 8, d);
#pragma unroll
          for (int e = 0; e < 8; ++e) vals[i][e] += d[e];
          // round through bf16 so the statistics match what backward will re-read
#pragma unroll
          for (int e = 0; e < 8; ++e) vals[i][e] = __bfloat162float(__float2bfloat16_rn(vals[i][e]));
          store8(xo + base + vi * 8, vals[i]);
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += vals[i][e] * vals[i][e];
      }
    }
    const float tot = block_sum(ss, red);
    const float rstd = rsqrtf(tot / (float)D + eps);
    if (threadIdx.x == 0 && rstd_out != nullptr) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * EW_THREADS;
      if (vi < nvec) {
        float wv[8], o[8];
        load8(w + vi * 8, wv);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = vals[i][e] * rstd * wv[e];
        store8(y + base + vi * 8, o);
      }
    }
  }
}

// ---------------------------------------------------------------------------------
// RMSNorm backward. With xhat = x*rstd, g = dy*w:
//   dx = dres + rstd * (g - xhat * mean(g*xhat))        dw[c] = sum_rows dy*xhat
// Each CTA walks rows with a grid stride and keeps per-column dw partial sums in
// registers; partials land in `dw_partial[gridDim.x, D]` (fp32 scratch supplied by the
// caller) and are folded by rmsnorm_dw_reduce_kernel.
// ---------------------------------------------------------------------------------
template <int VPT>
__global__ void __launch_bounds__(EW_THREADS)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                   const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dy,
                   const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                   float* __restrict__ dw_partial, int T, int D, float eps) {
  __shared__ float red[32];
  const int nvec = D >> 3;
  float dwacc[VPT][8];
  float wv[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int vi = threadIdx.x + i * EW_THREADS;
#pragma unroll
    for (int e = 0; e < 8; ++e) dwacc[i][e] = 0.f;
    if (vi < nvec) load8(w + vi * 8, wv[i]);
  }
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    const size_t base = (size_t)row * D;
    float xv[VPT][8], gv[VPT][8];
    float dot = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * EW_THREADS;
      if (vi < nvec) {
        load8(x + base + vi * 8, xv[i]);
        load8(dy + base + vi * 8, gv[i]);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss += xv[i][e] * xv[i][e];
      }
    }
    float rstd;
    if (rstd_in != nullptr) {
      rstd = rstd_in[row];
    } else {
      rstd = rsqrtf(block_sum(ss, red) / (float)D + eps);
    }
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * EW_THREADS;
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
      const int vi = threadIdx.x + i * EW_THREADS;
      if (vi < nvec) {
        float o[8];
        if (dres != nullptr) load8(dres + base + vi * 8, o);
        else {
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += rstd * (gv[i][e] - xv[i][e] * mean_gx);
        store8(dx + base + vi * 8, o);
      }
    }
  }
  if (dw_partial != nullptr) {
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int vi = threadIdx.x + i * EW_THREADS;
      if (vi < nvec) {
        float* dst = dw_partial + (size_t)blockIdx.x * D + vi * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
      }
    }
  }
}

// dw[c] (+)= sum_p partial[p, c]; output bf16 or fp32.
__global__ void rmsnorm_dw_reduce_kernel(const float* __restrict__ partial, int P, int D, void* __restrict__ dw,
                                         int dw_is_f32, int accumulate) {
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
// RoPE (rotate-half convention, Llama): for each rotated head and pair (i, i+hd/2)
//   out[i]      = a*cos - b*sin        out[i+hd/2] = b*cos + a*sin      (inverse: sin -> -sin)
// Applied in place to the first `n_rot_heads` heads of every row of a fused QKV buffer
// (row stride `ld` elements). cos/sin tables are [S, hd/2] fp32; position = row % S.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(EW_THREADS)
rope_kernel(__nv_bfloat16* __restrict__ buf, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
            int T, int S, int n_rot_heads, int hd, int ld, int inverse) {
  const int half = hd >> 1;
  const int vec_per_head = half >> 3;  // 8 pairs per thread-item
  const long long items = (long long)T * n_rot_heads * vec_per_head;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items;
       it += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(it % vec_per_head);
    const long long r = it / vec_per_head;
    const int h = (int)(r % n_rot_heads);
    const int t = (int)(r / n_rot_heads);
    const int pos = t % S;
    __nv_bfloat16* p = buf + (size_t)t * ld + (size_t)h * hd + v * 8;
    float a[8], b[8];
    load8(p, a);
    load8(p + half, b);
    const float4 c0 = *reinterpret_cast<const float4*>(cos_t + (size_t)pos * half + v * 8);
    const float4 c1 = *reinterpret_cast<const float4*>(cos_t + (size_t)pos * half + v * 8 + 4);
    const float4 s0 = *reinterpret_cast<const float4*>(sin_t + (size_t)pos * half + v * 8);
    const float4 s1 = *reinterpret_cast<const float4*>(sin_t + (size_t)pos * half + v * 8 + 4);
    const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    float sn[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    float oa[8], ob[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float s = inverse ? -sn[e] : sn[e];
      oa[e] = a[e] * cs[e] - b[e] * s;
      ob[e] = b[e] * cs[e] + a[e] * s;
    }
    store8(p, oa);
    store8(p + half, ob);
  }
}

// ---------------------------------------------------------------------------------
// SwiGLU on a fused [T, 2F] gate|up buffer.
//   fwd: act = silu(gate) * up
//   bwd: dgate = dact * up * sig * (1 + gate*(1-sig)) ; dup = dact * silu(gate)
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(EW_THREADS)
swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act, long long T, int F) {
  const int vecs = F >> 3;
  const long long items = T * vecs;
  const long long stride = (long long)gridDim.x * blockDim.x;
  constexpr int UNR = 4;  // 8 independent 128-bit loads in flight per thread
  for (long long it0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; it0 < items; it0 += stride * UNR) {
    float g[UNR][8], u[UNR][8];
#pragma unroll
    for (int k = 0; k < UNR; ++k) {
      const long long it = it0 + k * stride;
      if (it < items) {
        const long long t = it / vecs;
        const int v = (int)(it - t * vecs);
        load8(gu + t * 2 * F + v * 8, g[k]);
        load8(gu + t * 2 * F + F + v * 8, u[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < UNR; ++k) {
      const long long it = it0 + k * stride;
      if (it < items) {
        const long long t = it / vecs;
        const int v = (int)(it - t * vecs);
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float sig = 1.f / (1.f + __expf(-g[k][e]));
          o[e] = g[k][e] * sig * u[k][e];
        }
        store8(act + t * F + v * 8, o);
      }
    }
  }
}
__global__ void __launch_bounds__(EW_THREADS)
swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ gu, const __nv_bfloat16* __restrict__ dact,
                  __nv_bfloat16* __restrict__ dgu, long long T, int F) {
  const int vecs = F >> 3;
  const long long items = T * vecs;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items;
       it += (long long)gridDim.x * blockDim.x) {
    const long long t = it / vecs;
    const int v = (int)(it - t * vecs);
    float g[8], u[8], d[8], dg[8], du[8];
    load8(gu + t * 2 * F + v * 8, g);
    load8(gu + t * 2 * F + F + v * 8, u);
    load8(dact + t * F + v * 8, d);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float sig = 1.f / (1.f + __expf(-g[e]));
      dg[e] = d[e] * u[e] * sig * (1.f + g[e] * (1.f - sig));
      du[e] = d[e] * g[e] * sig;
    }
    store8(dgu + t * 2 * F + v * 8, dg);
    store8(dgu + t * 2 * F + F + v * 8, du);
  }
}

// ---------------------------------------------------------------------------------
// Embedding gather / scatter-add (bf16 table, int64 or int32 ids). D % 8 == 0.
// ---------------------------------------------------------------------------------
template <typename IdT>
__global__ void __launch_bounds__(EW_THREADS)
embedding_fwd_kernel(const IdT* __restrict__ ids, const __nv_bfloat16* __restrict__ table,
                     __nv_bfloat16* __restrict__ out, long long T, int D, int V) {
  const int vecs = D >> 3;
  const long long items = T * vecs;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items;
       it += (long long)gridDim.x * blockDim.x) {
    const long long t = it / vecs;
    const int v = (int)(it - t * vecs);
    long long id = (long long)ids[t];
    id = id < 0 ? 0 : (id >= V ? V - 1 : id);
    *reinterpret_cast<uint4*>(out + t * D + v * 8) = *reinterpret_cast<const uint4*>(table + id * D + v * 8);
  }
}
template <typename IdT>
__global__ void __launch_bounds__(EW_THREADS)
embedding_bwd_kernel(const IdT* __restrict__ ids, const __nv_bfloat16* __restrict__ dout, void* __restrict__ dtable,
                     long long T, int D, int V, int dtable_is_f32) {
  const int pairs = D >> 1;
  const long long items = T * pairs;
  for (long long it = (long long)blockIdx.x * blockDim.x + threadIdx.x; it < items;
       it += (long long)gridDim.x * blockDim.x) {
    const long long t = it / pairs;
    const int p = (int)(it - t * pairs);
    long long id = (long long)ids[t];
    if (id < 0 || id >= V) continue;
    const __nv_bfloat162 g = *reinterpret_cast<const __nv_bfloat162*>(dout + t * D + 2 * p);
    if (dtable_is_f32) {
      float* dst = reinterpret_cast<float*>(dtable) + id * D + 2 * p;
      atomicAdd(dst, __low2float(g));
      atomicAdd(dst + 1, __high2float(g));
    } else {
      atomicAdd(reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(dtable) + id * D + 2 * p), g);
    }
  }
}

// ---------------------------------------------------------------------------------
// Cross-entropy on a [T, V] bf16 logits chunk, forward + backward in one launch:
//   loss[t] = logsumexp(z_t) - z_t[y_t]        dz = (softmax(z) - onehot(y)) * grad_scale
// dz overwrites the logits in place (the chunk is dead afterwards), so the full-vocab
// probabilities are never stored separately. One CTA per row-slot; the row is streamed
// three times out of L2 (max, sum, write). targets < 0 are ignored (loss 0, dz 0).
// ---------------------------------------------------------------------------------
template <typename IdT>
__global__ void __launch_bounds__(512)
cross_entropy_kernel(__nv_bfloat16* __restrict__ logits, const IdT* __restrict__ targets, float* __restrict__ loss,
                     float* __restrict__ lse_out, int T, int V, long long ld, float grad_scale, int write_grad) {
  __shared__ float red[32];
  const int nvec = V >> 3;  // V % 8 == 0 enforced by the host wrapper
  for (int row = blockIdx.x; row < T; row += gridDim.x) {
    __nv_bfloat16* z = logits + (size_t)row * ld;
    const long long y = (long long)targets[row];
    float mx = -INFINITY;
    for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
      float f[8];
      load8(z + vi * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) mx = fmaxf(mx, f[e]);
    }
    mx = block_max(mx, red);
    float se = 0.f;
    for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
      float f[8];
      load8(z + vi * 8, f);
#pragma unroll
      for (int e = 0; e < 8; ++e) se += __expf(f[e] - mx);
    }
    se = block_sum(se, red);
    const float lse = mx + __logf(se);
    const bool valid = (y >= 0 && y < V);
    if (threadIdx.x == 0) {
      loss[row] = valid ? (lse - __bfloat162float(z[y])) : 0.f;
      if (lse_out != nullptr) lse_out[row] = lse;
    }
    if (write_grad) {
      __syncthreads();  // z[y] was read above before anyone overwrites it
      const float sc = valid ? grad_scale : 0.f;
      for (int vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
        float f[8];
        load8(z + vi * 8, f);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float p = __expf(f[e] - lse);
          if (vi * 8 + e == y) p -= 1.f;
          f[e] = p * sc;
        }
        store8(z + vi * 8, f);
      }
    }
  }
}

// out = a + b (bf16), used for residual joins that are not followed by a norm.
__global__ void __launch_bounds__(EW_THREADS)
add_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b, __nv_bfloat16* __restrict__ o,
           long long nvec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    float x[8], y[8];
    load8(a + i * 8, x);
    load8(b + i * 8, y);
#pragma unroll
    for (int e = 0; e < 8; ++e) x[e] += y[e];
    store8(o + i * 8, x);
  }
}

static inline int grid_for(long long items, int threads, int waves = 8) {
  const int sms = device_sm_count();
  long long blocks = (items + threads - 1) / threads;
  const long long cap = (long long)sms * waves;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

static inline int vpt_for(int D) {
  const int nvec = D >> 3;
  if (nvec <= EW_THREADS) return 1;
  if (nvec <= 2 * EW_THREADS) return 2;
  if (nvec <= 4 * EW_THREADS) return 4;
  return 0;
}

}  // namespace bz

using namespace bz;

extern "C" int bz_rmsnorm_fwd(const void* x, const void* delta, const void* w, void* x_out, void* y, float* rstd,
                              int T, int D, float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "rmsnorm: D=%d must be a multiple of 8", D);
  BZ_REQUIRE(delta == nullptr || x_out != nullptr, "rmsnorm: delta given without x_out");
  const int vpt = vpt_for(D);
  BZ_REQUIRE(vpt != 0, "rmsnorm: D=%d too large (max %d)", D, 8 * 4 * EW_THREADS);
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = min(T, device_sm_count() * 8);
  auto X = (const __nv_bfloat16*)x; auto DL = (const __nv_bfloat16*)delta; auto W = (const __nv_bfloat16*)w;
  auto XO = (__nv_bfloat16*)x_out; auto Y = (__nv_bfloat16*)y;
  if (vpt == 1) rmsnorm_fwd_kernel<1><<<grid, EW_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  else if (vpt == 2) rmsnorm_fwd_kernel<2><<<grid, EW_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  else rmsnorm_fwd_kernel<4><<<grid, EW_THREADS, 0, st>>>(X, DL, W, XO, Y, rstd, T, D, eps);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_rmsnorm_bwd_workspace_floats(int T, int D) {
  const int grid = min(T, device_sm_count() * 2);
  return grid * D;
}

extern "C" int bz_rmsnorm_bwd(const void* x, const void* w, const float* rstd, const void* dy, const void* dres,
                              void* dx, void* dw, int dw_is_f32, int dw_accumulate, float* workspace, int T, int D,
                              float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "rmsnorm_bwd: D=%d must be a multiple of 8", D);
  const int vpt = vpt_for(D);
  BZ_REQUIRE(vpt != 0, "rmsnorm_bwd: D=%d too large", D);
  BZ_REQUIRE(dw == nullptr || workspace != nullptr, "rmsnorm_bwd: dw requested without workspace");
  if (T <= 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = min(T, device_sm_count() * 2);
  float* part = dw ? workspace : nullptr;
  auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w; auto DY = (const __nv_bfloat16*)dy;
  auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
  if (vpt == 1) rmsnorm_bwd_kernel<1><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else if (vpt == 2) rmsnorm_bwd_kernel<2><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else rmsnorm_bwd_kernel<4><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  BZ_CUDA_CHECK(cudaGetLastError());
  if (dw) {
    rmsnorm_dw_reduce_kernel<<<(D + 255) / 256, 256, 0, st>>>(part, grid, D, dw, dw_is_f32, dw_accumulate);
    BZ_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

extern "C" int bz_rope(void* buf, const float* cos_t, const float* sin_t, int T, int S, int n_rot_heads, int head_dim,
                       int ld, int inverse, void* stream) {
  BZ_REQUIRE(head_dim % 16 == 0, "rope: head_dim=%d must be a multiple of 16", head_dim);
  BZ_REQUIRE(ld % 8 == 0, "rope: row stride must be a multiple of 8");
  if (T <= 0) return 0;
  const long long items = (long long)T * n_rot_heads * (head_dim / 16);
  rope_kernel<<<grid_for(items, EW_THREADS), EW_THREADS, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)buf, cos_t, sin_t, T, S, n_rot_heads, head_dim, ld, inverse);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_swiglu_fwd(const void* gate_up, void* act, long long T, int F, void* stream) {
  BZ_REQUIRE(F % 8 == 0, "swiglu: F=%d must be a multiple of 8", F);
  if (T <= 0) return 0;
  swiglu_fwd_kernel<<<grid_for(T * (F / 8), EW_THREADS), EW_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gate_up, (__nv_bfloat16*)act, T, F);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_swiglu_bwd(const void* gate_up, const void* dact, void* dgate_up, long long T, int F, void* stream) {
  BZ_REQUIRE(F % 8 == 0, "swiglu: F=%d must be a multiple of 8", F);
  if (T <= 0) return 0;
  swiglu_bwd_kernel<<<grid_for(T * (F / 8), EW_THREADS), EW_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gate_up, (const __nv_bfloat16*)dact, (__nv_bfloat16*)dgate_up, T, F);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_embedding_fwd(const void* ids, int ids_are_i64, const void* table, void* out, long long T, int D,
                                int V, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "embedding: D=%d must be a multiple of 8", D);
  if (T <= 0) return 0;
  const int grid = grid_for(T * (D / 8), EW_THREADS);
  if (ids_are_i64)
    embedding_fwd_kernel<long long><<<grid, EW_THREADS, 0, (cudaStream_t)stream>>>(
        (const long long*)ids, (const __nv_bfloat16*)table, (__nv_bfloat16*)out, T, D, V);
  else
    embedding_fwd_kernel<int><<<grid, EW_THREADS, 0, (cudaStream_t)stream>>>(
        (const int*)ids, (const __nv_bfloat16*)table, (__nv_bfloat16*)out, T, D, V);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_embedding_bwd(const void* ids, int ids_are_i64, const void* dout, void* dtable, int dtable_is_f32,
                                long long T, int D, int V, void* stream) {
  BZ_REQUIRE(D % 2 == 0, "embedding_bwd: D must be even");
  if (T <= 0) return 0;
  const int grid = grid_for(T * (D / 2), EW_THREADS);
  if (ids_are_i64)
    embedding_bwd_kernel<long long><<<grid, EW_THREADS, 0, (cudaStream_t)stream>>>(
        (const long long*)ids, (const __nv_bfloat16*)dout, dtable, T, D, V, dtable_is_f32);
  else
    embedding_bwd_kernel<int><<<grid, EW_THREADS, 0, (cudaStream_t)stream>>>(
        (const int*)ids, (const __nv_bfloat16*)dout, dtable, T, D, V, dtable_is_f32);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_cross_entropy(void* logits, const void* targets, int targets_are_i64, float* loss, float* lse,
                                int T, int V, long long ld, float grad_scale, int write_grad, void* stream) {
  BZ_REQUIRE(V % 8 == 0 && ld % 8 == 0, "cross_entropy: V=%d and ld must be multiples of 8", V);
  if (T <= 0) return 0;
  const int grid = min(T, device_sm_count() * 4);
  if (targets_are_i64)
    cross_entropy_kernel<long long><<<grid, 512, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16*)logits, (const long long*)targets, loss, lse, T, V, ld, grad_scale, write_grad);
  else
    cross_entropy_kernel<int><<<grid, 512, 0, (cudaStream_t)stream>>>(
        (__nv_bfloat16*)logits, (const int*)targets, loss, lse, T, V, ld, grad_scale, write_grad);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

extern "C" int bz_add_bf16(const void* a, const void* b, void* out, long long n, void* stream) {
  BZ_REQUIRE(n % 8 == 0, "add: n must be a multiple of 8");
  if (n <= 0) return 0;
  add_kernel<<<grid_for(n / 8, EW_THREADS), EW_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)out, n / 8);
  BZ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

// =====================================================================================
// RMSNorm backward without caller-provided scratch.
//
// Who this is for: callers that are not allowed to allocate device memory when they need the
// backward pass. Three real cases: (1) a step loop that must not stall — cudaMalloc and cudaFree
// synchronise the whole device, so an allocation in the middle of a step serialises every stream
// that is overlapping copies and collectives with compute; (2) a job whose HBM is planned to the
// last block (ZeRO-3 at 70B sizing): there is no slack to allocate from, and a failed allocation
// mid-step would kill the run; (3) plain C/C++ hosts that link this library without PyTorch and
// therefore have no caching allocator to carve a workspace from.
//
// bz_rmsnorm_bwd needs `grid x D` floats of scratch for the per-CTA dw partial sums. Here that
// scratch lives in the library itself: a file-scope device array of
// BZ_RMSNORM_BWD_STATIC_SCRATCH_FLOATS floats (5 MiB), present on every GPU that loads the
// library until the process exits. Launch geometry is identical to bz_rmsnorm_bwd, so the two
// produce the same bits.
//
// Rules for the caller:
//  * Problems whose partial sums do not fit (min(T, 2*SMs) * D floats > capacity) are refused
//    with an error; use bz_rmsnorm_bwd with your own workspace for those.
//  * The FIRST call on each device looks up the array's device address
//    (cudaGetSymbolAddress, with that device current). Make that first call during set-up,
//    before any stream capture begins. Later calls on that device only launch kernels.
//  * All calls on a device share the one array. Calls whose kernels can overlap on the GPU
//    (different streams without an ordering between them) race on it; keep them on one stream
//    or order them with events. The array keeps whatever the previous call left in it; that is
//    harmless here because every slot the reduction pass reads is written by the partial pass of
//    the same call — nothing relies on the scratch starting at zero.
//  * No cudaMalloc / cudaFree / allocating library routine is called, on any call.
// =====================================================================================
namespace bz {

constexpr int kStaticScratchFloats = BZ_RMSNORM_BWD_STATIC_SCRATCH_FLOATS;
static __device__ __align__(256) float g_rmsnorm_bwd_scratch[kStaticScratchFloats];

// Device address of g_rmsnorm_bwd_scratch per device ordinal. In host code the symbol name is only
// a stand-in; the address a kernel may dereference has to come from cudaGetSymbolAddress.
static float* g_scratch_addr[64] = {nullptr};
static std::mutex g_scratch_mu;      // two host threads may make their first call on a device at the same time

static int static_scratch_for_current_device(float** out) {
  int dev = -1;
  BZ_CUDA_CHECK(cudaGetDevice(&dev));
  BZ_REQUIRE(dev >= 0 && dev < 64, "rmsnorm_bwd_static_scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    BZ_CUDA_CHECK(cudaGetSymbolAddress(&p, g_rmsnorm_bwd_scratch));
    BZ_REQUIRE(p != nullptr, "rmsnorm_bwd_static_scratch: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace bz

extern "C" int bz_rmsnorm_bwd_static_scratch(const void* x, const void* w, const float* rstd, const void* dy,
                                             const void* dres, void* dx, void* dw, int dw_is_f32, int dw_accumulate,
                                             int T, int D, float eps, void* stream) {
  BZ_REQUIRE(D % 8 == 0, "rmsnorm_bwd_static_scratch: D=%d must be a multiple of 8", D);
  const int vpt = vpt_for(D);
  BZ_REQUIRE(vpt != 0, "rmsnorm_bwd_static_scratch: D=%d too large", D);
  if (T <= 0) return 0;
  const int grid = min(T, device_sm_count() * 2);  // same geometry as bz_rmsnorm_bwd
  const long long need = (long long)grid * D;
  BZ_REQUIRE(dw == nullptr || need <= (long long)kStaticScratchFloats,
             "rmsnorm_bwd_static_scratch: T=%d D=%d needs %lld floats of scratch but the built-in scratch holds %d; "
             "call bz_rmsnorm_bwd with a caller-provided workspace instead",
             T, D, need, kStaticScratchFloats);
  float* part = nullptr;
  if (dw != nullptr) {
    if (int rc = static_scratch_for_current_device(&part)) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  auto X = (const __nv_bfloat16*)x; auto W = (const __nv_bfloat16*)w; auto DY = (const __nv_bfloat16*)dy;
  auto DR = (const __nv_bfloat16*)dres; auto DX = (__nv_bfloat16*)dx;
  if (vpt == 1) rmsnorm_bwd_kernel<1><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else if (vpt == 2) rmsnorm_bwd_kernel<2><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  else rmsnorm_bwd_kernel<4><<<grid, EW_THREADS, 0, st>>>(X, W, rstd, DY, DR, DX, part, T, D, eps);
  BZ_CUDA_CHECK(cudaGetLastError());
  if (dw != nullptr) {
    rmsnorm_dw_reduce_kernel<<<(D + 255) / 256, 256, 0, st>>>(part, grid, D, dw, dw_is_f32, dw_accumulate);
    BZ_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}
