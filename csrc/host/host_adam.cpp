// H1: AdamW on the host for ZeRO-Offload. fp32 master / moments live in pinned host
// memory; gradients arrive from the GPU in bf16 or fp32 chunks, updated weights leave as bf16.
// Vectorised by the compiler for AVX-512 / AVX2 through function multiversioning (runtime
// dispatch), parallelised over a persistent pool of worker threads. Same arithmetic as
// csrc/optim.cu and torch.optim.AdamW. Replaces DeepSpeed's cpu_adam.cpp, which the reference
// selects by default with offload_optimizer.device = "cpu" (ref ai_engine/deepspeed_launcher.py:39,197-203).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdint>
#include <cstring>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace {

inline float bf16_to_f32(uint16_t h) {
  uint32_t u = (uint32_t)h << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN stays NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                           // round to nearest even
  return (uint16_t)(u >> 16);
}

struct Args {
  float* master; float* m; float* v; const void* grad; int grad_is_bf16; uint16_t* out;
  float lr, b1, b2, eps, decay, step_size, inv_sqrt_bc2, gs;
};

__attribute__((target_clones("avx512f", "avx2", "default")))
void adam_range(const Args& a, long long lo, long long hi) {
  float* __restrict__ p = a.master; float* __restrict__ m = a.m; float* __restrict__ v = a.v;
  const float b1 = a.b1, b2 = a.b2, ob1 = 1.f - a.b1, ob2 = 1.f - a.b2;
  if (a.grad_is_bf16) {
    const uint16_t* __restrict__ g = (const uint16_t*)a.grad;
    for (long long i = lo; i < hi; ++i) {
      const float ge = bf16_to_f32(g[i]) * a.gs;
      const float mi = b1 * m[i] + ob1 * ge;
      const float vi = b2 * v[i] + ob2 * ge * ge;
      const float pi = p[i] * a.decay - a.step_size * (mi / (std::sqrt(vi) * a.inv_sqrt_bc2 + a.eps));
      m[i] = mi; v[i] = vi; p[i] = pi;
    }
  } else {
    const float* __restrict__ g = (const float*)a.grad;
    for (long long i = lo; i < hi; ++i) {
      const float ge = g[i] * a.gs;
      const float mi = b1 * m[i] + ob1 * ge;
      const float vi = b2 * v[i] + ob2 * ge * ge;
      const float pi = p[i] * a.decay - a.step_size * (mi / (std::sqrt(vi) * a.inv_sqrt_bc2 + a.eps));
      m[i] = mi; v[i] = vi; p[i] = pi;
    }
  }
  if (a.out != nullptr)
    for (long long i = lo; i < hi; ++i) a.out[i] = f32_to_bf16(p[i]);
}

class Pool {
 public:
  explicit Pool(int n) : stop_(false), pending_(0) {
    for (int i = 0; i < n; ++i) workers_.emplace_back([this, i] { loop(i); });
  }
  ~Pool() {
    { std::lock_guard<std::mutex> l(mu_); stop_ = true; ++gen_; }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  int size() const { return (int)workers_.size(); }
  void run(const std::function<void(int)>& fn) {
    std::unique_lock<std::mutex> l(mu_);
    fn_ = &fn; pending_ = (int)workers_.size(); ++gen_;
    cv_.notify_all();
    done_.wait(l, [this] { return pending_ == 0; });
  }
 private:
  void loop(int id) {
    unsigned long long seen = 0;
    for (;;) {
      const std::function<void(int)>* fn;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [&] { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
        fn = fn_;
      }
      (*fn)(id);
      { std::lock_guard<std::mutex> l(mu_); if (--pending_ == 0) done_.notify_all(); }
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_;
  std::condition_variable cv_, done_;
  const std::function<void(int)>* fn_ = nullptr;
  unsigned long long gen_ = 0;
  bool stop_;
  int pending_;
};

Pool* g_pool = nullptr;
std::mutex g_pool_mu;

Pool& pool(int nthreads) {
  std::lock_guard<std::mutex> l(g_pool_mu);
  if (g_pool == nullptr || g_pool->size() != nthreads) {
    delete g_pool;
    g_pool = new Pool(nthreads);
  }
  return *g_pool;
}

}  // namespace

extern "C" {

// In-place AdamW over [0, n). step is 1-based. out (bf16 bits) may be NULL. nthreads <= 0 -> all cores / 2.
int bz_host_adamw(float* master, float* m, float* v, const void* grad, int grad_is_bf16, uint16_t* out, long long n,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int step, float grad_scale,
                  int nthreads) {
  if (n <= 0) return 0;
  if (step < 1) return -1;
  Args a;
  a.master = master; a.m = m; a.v = v; a.grad = grad; a.grad_is_bf16 = grad_is_bf16; a.out = out;
  a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps;
  const float bc1 = 1.f - std::pow(beta1, (float)step), bc2 = 1.f - std::pow(beta2, (float)step);
  a.decay = 1.f - lr * weight_decay;
  a.step_size = lr / bc1;
  a.inv_sqrt_bc2 = 1.f / std::sqrt(bc2);
  a.gs = grad_scale;
  if (nthreads <= 0) nthreads = std::max(1u, std::thread::hardware_concurrency() / 2);
  if (n < (1 << 16) || nthreads == 1) {
    adam_range(a, 0, n);
    return 0;
  }
  Pool& p = pool(nthreads);
  const long long per = ((n + nthreads - 1) / nthreads + 15) & ~15LL;
  p.run([&](int id) {
    const long long lo = std::min<long long>(n, (long long)id * per), hi = std::min<long long>(n, lo + per);
    if (lo < hi) adam_range(a, lo, hi);
  });
  return 0;
}

int bz_host_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
