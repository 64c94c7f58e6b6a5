// Backward of the activation that the GEMM's activation epilogue applied:
//   gelu_new:  dZ = dY * gelu_new'(Z)   (Z = the stored pre-activation; GPT-2 mlp.c_fc -> act)
//   relu:      dZ = dY * [Y > 0]        (Y = the stored activation;     OPT fc1 -> relu)
// HBM-bound: two 16-bit reads and one 16-bit write per element, 16 bytes per thread per access.
#pragma once

#include "common.cuh"
#include "gemm_sm100.cuh"

namespace lb {
namespace act {

constexpr int THREADS = 256;

__global__ void __launch_bounds__(THREADS)
    act_bwd_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ zy, uint4* __restrict__ dz, long long n8,
                   int act, int dtype) {
  pdl_launch_dependents();
  pdl_wait();
  for (long long i = blockIdx.x * static_cast<long long>(THREADS) + threadIdx.x; i < n8;
       i += static_cast<long long>(gridDim.x) * THREADS) {
    const uint4 g = __ldg(dy + i);
    const uint4 s = __ldg(zy + i);
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, sw[4] = {s.x, s.y, s.z, s.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 gf = unpack2(gw[e], dtype), sf = unpack2(sw[e], dtype);
      float a, b;
      if (act == gemm::kActGeluNew) {
        a = gf.x * gemm::gelu_new_grad(sf.x);
        b = gf.y * gemm::gelu_new_grad(sf.y);
      } else {
        a = sf.x > 0.f ? gf.x : 0.f;
        b = sf.y > 0.f ? gf.y : 0.f;
      }
      o[e] = pack2(a, b, dtype);
    }
    dz[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Merge / unmerge (peft lora.Linear.merge): W <- W + scale * B A with scale = +-s, in place on
// the 16-bit weight, the rank-r product formed in fp32 and added before the single rounding.  One thread per 8
// consecutive elements of a weight row; HBM-bound (one read + one write of W).  W is [N, K] (nn.Linear) or, with
// w_is_KN, [K, N] (GPT-2 Conv1D); A [r, K], B [N, r] fp32.
__global__ void __launch_bounds__(THREADS)
    merge_kernel(uint16_t* __restrict__ W, const float* __restrict__ A, const float* __restrict__ B, long long N, long long K,
                 int r, float scale, int w_is_KN, int dtype) {
  const long long rows = w_is_KN ? K : N, cols = w_is_KN ? N : K;  // cols % 8 == 0 (host check)
  const long long groups = rows * (cols / 8);
  for (long long g = blockIdx.x * static_cast<long long>(THREADS) + threadIdx.x; g < groups;
       g += static_cast<long long>(gridDim.x) * THREADS) {
    const long long row = g / (cols / 8), c0 = (g % (cols / 8)) * 8;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int j = 0; j < r; ++j) {
      if (!w_is_KN) {  // row = n, columns = k: delta[n, k] = sum_j B[n, j] A[j, k]
        const float b = __ldg(B + row * r + j);
        const float4 a0 = __ldg(reinterpret_cast<const float4*>(A + static_cast<long long>(j) * K + c0));
        const float4 a1 = __ldg(reinterpret_cast<const float4*>(A + static_cast<long long>(j) * K + c0 + 4));
        acc[0] = fmaf(b, a0.x, acc[0]), acc[1] = fmaf(b, a0.y, acc[1]), acc[2] = fmaf(b, a0.z, acc[2]), acc[3] = fmaf(b, a0.w, acc[3]);
        acc[4] = fmaf(b, a1.x, acc[4]), acc[5] = fmaf(b, a1.y, acc[5]), acc[6] = fmaf(b, a1.z, acc[6]), acc[7] = fmaf(b, a1.w, acc[7]);
      } else {  // row = k, columns = n: delta[k, n] = sum_j A[j, k] B[n, j]
        const float av = __ldg(A + static_cast<long long>(j) * K + row);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = fmaf(av, __ldg(B + (c0 + e) * r + j), acc[e]);
      }
    }
    uint4* wp = reinterpret_cast<uint4*>(W + row * cols + c0);
    const uint4 w = *wp;
    const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = unpack2(ws[e], dtype);
      o[e] = pack2(fmaf(scale, acc[2 * e], f.x), fmaf(scale, acc[2 * e + 1], f.y), dtype);
    }
    *wp = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace act
}  // namespace lb
