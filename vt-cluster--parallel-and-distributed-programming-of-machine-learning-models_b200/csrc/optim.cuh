// Optimiser step of the adapter parameters: global gradient-norm clip + AdamW over ALL adapter tensors in two
// launches (HF Trainer defaults the reference runs with, finetune_lora_distilgpt2.py:130-145: AdamW betas
// (0.9, 0.999), eps 1e-8, weight_decay 0, max_grad_norm 1.0; DeepSpeed AdamW + gradient_clipping 1.0 at
// finetune_lora_opt_pp.py:193-197).  torch.optim.AdamW + clip_grad_norm_ issue some thirty tiny kernels for the
// 36-384 adapter tensors; here a device-side table lists the tensors and each 256-thread block owns one chunk.
//   sumsq_kernel : per-chunk sums of g^2; the last block to finish folds them in a fixed order (bit-reproducible)
//                  into norm = sqrt(sum) and coef = min(1, max_norm / (norm + 1e-6))   (torch's clip_grad_norm_)
//   adamw_kernel : g *= coef (stored back, as clip_grad_norm_ does); p *= 1 - lr wd; m, v updated; p -= step
// HBM-bound: 32 bytes per parameter (read and write g, p, m, v).
#pragma once

#include "common.cuh"

namespace lb {
namespace optim {

constexpr int THREADS = 256;
constexpr int CHUNK = 4096;  // elements per block

struct TensorEntry {  // one adapter tensor (all fp32, contiguous)
  float* p;
  float* g;
  float* m;
  float* v;
  long long n;
  long long first_chunk;  // prefix sum of ceil(n / CHUNK) over the tensors before this one
  int* step;              // this tensor's own step count (torch.optim.AdamW keeps one per parameter: a tensor that
                          // had no gradient in some step lags behind); advanced here, on the device
};

// scratch layout (floats): [0] norm, [1] coef, [2] arrival counter (as uint), [4 ...] per-chunk partial sums
__device__ __forceinline__ int find_tensor(const TensorEntry* tab, int n_tensors, long long chunk) {
  int lo = 0, hi = n_tensors - 1;
  while (lo < hi) {  // last entry whose first_chunk <= chunk
    const int mid = (lo + hi + 1) >> 1;
    if (tab[mid].first_chunk <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void __launch_bounds__(THREADS)
    sumsq_kernel(const TensorEntry* __restrict__ tab, int n_tensors, long long n_chunks, float max_norm,
                 float* __restrict__ scratch) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[THREADS / 32];
  __shared__ int s_last;
  const long long chunk = blockIdx.x;
  const TensorEntry e = tab[find_tensor(tab, n_tensors, chunk)];
  const long long base = (chunk - e.first_chunk) * CHUNK;
  const long long end = min(e.n, base + CHUNK);
  float acc = 0.f;
  for (long long i = base + threadIdx.x; i < end; i += THREADS) {
    const float g = e.g[i];
    acc = fmaf(g, g, acc);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < THREADS / 32; ++w) s += red[w];
    scratch[4 + chunk] = s;
    __threadfence();
    const unsigned int prev = atomicInc(reinterpret_cast<unsigned int*>(scratch + 2), static_cast<unsigned int>(n_chunks - 1));
    s_last = prev == static_cast<unsigned int>(n_chunks - 1);  // the counter wraps to zero: ready for the next step
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    // fixed order: thread t sums chunks t, t + 256, ...; then the same shuffle tree and warp-order fold as above
    float s = 0.f;
    for (long long c = threadIdx.x; c < n_chunks; c += THREADS) s += __ldcg(scratch + 4 + c);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
      float tot = 0.f;
      for (int w = 0; w < THREADS / 32; ++w) tot += red[w];
      const float norm = sqrtf(tot);
      scratch[0] = norm;
      scratch[1] = max_norm > 0.f ? fminf(1.f, max_norm / (norm + 1e-6f)) : 1.f;
    }
    for (int t = threadIdx.x; t < n_tensors; t += THREADS) *tab[t].step += 1;  // read by adamw_kernel (next launch)
  }
}

__global__ void __launch_bounds__(THREADS)
    adamw_kernel(const TensorEntry* __restrict__ tab, int n_tensors, const float* __restrict__ scratch, float lr,
                 float beta1, float beta2, float eps, float weight_decay) {
  pdl_launch_dependents();
  pdl_wait();
  const long long chunk = blockIdx.x;
  const TensorEntry e = tab[find_tensor(tab, n_tensors, chunk)];
  const long long base = (chunk - e.first_chunk) * CHUNK;
  const long long end = min(e.n, base + CHUNK);
  const float coef = scratch[1];
  const float t = static_cast<float>(*e.step);  // >= 1: sumsq_kernel advanced it
  const float bias_c1 = 1.f - powf(beta1, t), bias_c2_sqrt = sqrtf(1.f - powf(beta2, t));
  const float step = lr / bias_c1;
  const float decay = 1.f - lr * weight_decay;
  for (long long i = base + threadIdx.x; i < end; i += THREADS) {
    const float g = e.g[i] * coef;
    float p = e.p[i] * decay;
    const float m = beta1 * e.m[i] + (1.f - beta1) * g;
    const float v = beta2 * e.v[i] + (1.f - beta2) * g * g;
    p -= step * m / (sqrtf(v) / bias_c2_sqrt + eps);
    e.g[i] = g;
    e.p[i] = p;
    e.m[i] = m;
    e.v[i] = v;
  }
}

}  // namespace optim
}  // namespace lb
