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

}  // namespace act
}  // namespace lb
