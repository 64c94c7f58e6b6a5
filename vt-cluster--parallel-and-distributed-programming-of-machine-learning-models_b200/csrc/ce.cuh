// Token-level cross-entropy straight from 16-bit logits: what the reference
// gets from DataCollatorForLanguageModeling(mlm=False) labels + the HF causal-LM loss
// (finetune_lora_distilgpt2.py:114) and from nn.CrossEntropyLoss over shifted logits
// (finetune_lora_opt_pp.py:143-153).  HBM-bound: forward reads the logits once (log-sum-exp and the
// label's logit per row, one pass with an online maximum), backward reads them once more and writes dlogits; no fp32 copy of the
// [tokens, vocab] matrix is ever made.  The causal shift is the caller's: labels[i] is the target
// of row i (ignore_index where there is none), so logits are never sliced or copied.
#pragma once

#include "common.cuh"

namespace lb {
namespace ce {

constexpr int THREADS = 256;

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One CTA per row: lse[row] = log sum_v exp(x[row, v]);  loss[row] = lse - x[row, label] (0 if ignored).
// ONE pass over the row (online softmax: running maximum m and sum s of exp(x - m), rescaled when m
// grows): rows in flight on all SMs add up to about the size of L2, so a second pass would come from HBM.
__global__ void __launch_bounds__(THREADS)
    ce_fwd_kernel(const uint16_t* __restrict__ logits, long long ld, const long long* __restrict__ labels,
                  float* __restrict__ lse, float* __restrict__ loss, int V, long long ignore_index, int dtype) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr float L2E = 1.4426950408889634f;
  const long long row = blockIdx.x;
  const uint16_t* x = logits + row * ld;
  const int V8 = V & ~7;
  float m = -INFINITY, s = 0.f;  // this thread's running maximum and sum of exp(x - m)
  for (int v = threadIdx.x * 8; v < V8; v += THREADS * 8) {
    const uint4 w = __ldg(reinterpret_cast<const uint4*>(x + v));
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
    float f[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 t = unpack2(ws[e], dtype);
      f[2 * e] = t.x;
      f[2 * e + 1] = t.y;
    }
    float vm = f[0];
#pragma unroll
    for (int e = 1; e < 8; ++e) vm = fmaxf(vm, f[e]);
    if (vm > m) {
      s *= exp2f((m - vm) * L2E);  // m = -inf on the first visit: exp2(-inf) = 0 and s is 0
      m = vm;
    }
    const float ml2 = m * L2E;
#pragma unroll
    for (int e = 0; e < 8; ++e) s += exp2f(fmaf(f[e], L2E, -ml2));
  }
  for (int v = V8 + threadIdx.x; v < V; v += THREADS) {
    const float xv = to_float16bits(__ldg(x + v), dtype);
    if (xv > m) {
      s *= exp2f((m - xv) * L2E);
      m = xv;
    }
    s += exp2f((xv - m) * L2E);
  }
  // combine (m, s) pairs: warp, then the 8 warps in fixed order (bit-reproducible)
  const float wm = warp_max(m);
  s = (m == -INFINITY) ? 0.f : s * exp2f((m - wm) * L2E);  // a thread that saw nothing contributes nothing
  s = warp_sum(s);
  __shared__ float red_m[THREADS / 32], red_s[THREADS / 32];
  if ((threadIdx.x & 31) == 0) {
    red_m[threadIdx.x >> 5] = wm;
    red_s[threadIdx.x >> 5] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = red_m[0];
    for (int i = 1; i < THREADS / 32; ++i) M = fmaxf(M, red_m[i]);
    float t = 0.f;
    for (int i = 0; i < THREADS / 32; ++i)
      if (red_m[i] != -INFINITY) t += red_s[i] * exp2f((red_m[i] - M) * L2E);
    const float l = M + logf(t);
    lse[row] = l;
    const long long y = labels[row];
    loss[row] = (y == ignore_index || y < 0 || y >= V) ? 0.f : l - to_float16bits(__ldg(x + y), dtype);
  }
}

// dlogits[row, v] = (exp(x - lse) - [v == label]) * grad_scale   (zeros for ignored rows), 16-bit.
__global__ void __launch_bounds__(THREADS)
    ce_bwd_kernel(const uint16_t* __restrict__ logits, long long ld, const long long* __restrict__ labels,
                  const float* __restrict__ lse, const float* __restrict__ grad_scale, uint16_t* __restrict__ dlogits,
                  long long ldd, int V, long long ignore_index, int dtype) {
  pdl_launch_dependents();
  pdl_wait();
  const long long row = blockIdx.x;
  const uint16_t* x = logits + row * ld;
  uint16_t* d = dlogits + row * ldd;
  const long long y = labels[row];
  const bool live = !(y == ignore_index || y < 0 || y >= V);
  const float g = live ? __ldg(grad_scale) : 0.f;
  const float ll2 = __ldg(lse + row) * 1.4426950408889634f;
  const int V8 = V & ~7;
  for (int v = threadIdx.x * 8; v < V8; v += THREADS * 8) {
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (live) {
      const uint4 w = __ldg(reinterpret_cast<const uint4*>(x + v));
      const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
      uint32_t os[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack2(ws[e], dtype);
        float p0 = exp2f(fmaf(f.x, 1.4426950408889634f, -ll2));
        float p1 = exp2f(fmaf(f.y, 1.4426950408889634f, -ll2));
        if (v + 2 * e == y) p0 -= 1.f;
        if (v + 2 * e + 1 == y) p1 -= 1.f;
        os[e] = pack2(p0 * g, p1 * g, dtype);
      }
      o = make_uint4(os[0], os[1], os[2], os[3]);
    }
    *reinterpret_cast<uint4*>(d + v) = o;
  }
  for (int v = V8 + threadIdx.x; v < V; v += THREADS) {
    float pv = 0.f;
    if (live) {
      pv = exp2f(fmaf(to_float16bits(__ldg(x + v), dtype), 1.4426950408889634f, -ll2));
      if (v == y) pv -= 1.f;
      pv *= g;
    }
    d[v] = static_cast<uint16_t>(pack2(pv, 0.f, dtype) & 0xFFFFu);
  }
}

}  // namespace ce
}  // namespace lb
