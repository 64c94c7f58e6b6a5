// Persistent, warp-specialised tcgen05 GEMM for the frozen projection of a
// LoRA-adapted linear layer, with the rank-r update folded in as one extra
// K-block and the bias / masked low-rank term applied in the epilogue.
//
//   D[M,N] = A1[M,K1] * B1^T  (+ A2[M,rp] * B2[N,rp]^T)  (+ bias[N])
//            (+ keep[M,N] * inv_keep * (dh[M,:r] * a16[:r,N]))
//
// forward  (peft lora.Linear.forward):
//     A1 = X, B1 = W, A2 = H = drop(X) A^T, B2 = s*B, bias = b
// backward (autograd of the same):
//     A1 = dY, B1 = W (other major-ness), p == 0: A2 = dH, B2 = A^T
//                                         p  > 0: masked epilogue with dh, a16, keep bits
//
// Roles per CTA (256 threads): warp 0 = TMA producer, warp 1 = MMA issuer
// (one thread), warp 2 = TMEM allocator, warps 4..7 = epilogue (TMEM -> regs
// -> swizzled smem -> TMA store).  Two 256-column fp32 accumulators in TMEM
// let the epilogue of tile t overlap the main loop of tile t+1.
#pragma once

#include "common.cuh"

namespace lb {
namespace gemm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 x 2 B = one 128-byte swizzle row
constexpr int STAGES = 4;
constexpr int A_TILE_BYTES = BM * BK * 2;                 // 16 KiB
constexpr int B_TILE_BYTES = BN * BK * 2;                 // 32 KiB
constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;  // 48 KiB
constexpr int EPI_WARPS = 4;
constexpr int EPI_BUF_BYTES = 32 * 128;  // 32 rows x 64 cols x 2 B
constexpr int EPI_BYTES = EPI_WARPS * 2 * EPI_BUF_BYTES;
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;
constexpr int THREADS = 256;
constexpr int TMEM_COLS = 512;
constexpr int GROUP_M = 16;  // m-blocks per raster group (L2 reuse of B panels)

struct Params {
  int M, N, K1;
  int r_pad;       // 0: no low-rank K-block; else 16 / 32 / 64
  int b_mn_major;  // 0: B1 stored [N,K1] (K-major); 1: stored [K1,N]
  int dtype;       // kDtypeBf16 / kDtypeFp16
  const void* bias;
  // masked low-rank epilogue (backward with dropout)
  int masked_r;  // 0 = off, else the true rank r (<= 64)
  const void* dh;  // [M, dh_ld] dtype
  int dh_ld;
  const void* a16;      // [>= masked_r, N] dtype, row stride N
  const uint8_t* keep;  // [M, N/8] keep bits
  float inv_keep;
};

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int group_sz = GROUP_M * num_n;
  const int g = tile / group_sz;
  const int first_m = g * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int in_g = tile - g * group_sz;
  m_blk = first_m + in_g % gm;
  n_blk = in_g / gm;
}

__global__ void __launch_bounds__(THREADS, 1)
    gemm_lora_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                     const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                     const __grid_constant__ CUtensorMap tmD, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint8_t* epi_base = smem + STAGES * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EPI_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    tma_prefetch_desc(&tmD);
    if (p.r_pad > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], EPI_WARPS);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_m = (p.M + BM - 1) / BM;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int nk_main = (p.K1 + BK - 1) / BK;
  const int nk = nk_main + (p.r_pad > 0 ? 1 : 0);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const uint32_t lora_bytes = static_cast<uint32_t>((BM + BN) * p.r_pad * 2);
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = stage_base + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_TILE_BYTES;
          if (kb < nk_main) {
            mbar_arrive_expect_tx(&full[stage], STAGE_BYTES);
            tma_load_2d(sa, &tmA1, &full[stage], kb * BK, m_blk * BM);
            if (!p.b_mn_major) {
              tma_load_2d(sb, &tmB1, &full[stage], kb * BK, n_blk * BN);
            } else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                tma_load_2d(sb + j * (BK * 128), &tmB1, &full[stage], n_blk * BN + j * 64, kb * BK);
            }
          } else {
            mbar_arrive_expect_tx(&full[stage], lora_bytes);
            tma_load_2d(sa, &tmA2, &full[stage], 0, m_blk * BM);
            tma_load_2d(sb, &tmB2, &full[stage], 0, n_blk * BN);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const int fmt = (p.dtype == kDtypeBf16) ? 1 : 0;
      const uint32_t idesc_main = make_idesc_f16(fmt, BM, BN, 0, p.b_mn_major);
      const uint32_t idesc_lora = make_idesc_f16(fmt, BM, BN, 0, 0);
      const uint32_t lora_sw = static_cast<uint32_t>(p.r_pad) * 2u;  // swizzle span = row bytes
      const uint32_t lora_layout = (lora_sw == 32u) ? 6u : (lora_sw == 64u ? 4u : 2u);
      const uint32_t lora_sbo = 8u * lora_sw;
      const int lora_ksteps = p.r_pad / 16;
      uint32_t stage = 0, phase = 0;
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const uint32_t buf = it & 1u;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(&tmem_empty[buf], acc_phase ^ 1u, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint32_t sa = smem_u32(stage_base + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_TILE_BYTES;
          if (kb < nk_main) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              const uint64_t da = make_smem_desc(sa + k * 32, 16, 1024, 2);
              const uint64_t db = p.b_mn_major ? make_smem_desc(sb + k * 2048, BK * 128, 1024, 2)
                                               : make_smem_desc(sb + k * 32, 16, 1024, 2);
              umma_f16(d_tmem, da, db, idesc_main, (kb | k) != 0 ? 1u : 0u);
            }
          } else {
            for (int k = 0; k < lora_ksteps; ++k) {
              const uint64_t da = make_smem_desc(sa + k * 32, 16, lora_sbo, lora_layout);
              const uint64_t db = make_smem_desc(sb + k * 32, 16, lora_sbo, lora_layout);
              umma_f16(d_tmem, da, db, idesc_lora, 1u);
            }
          }
          umma_commit(&empty[stage]);  // frees the smem slot once these MMAs retire
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[buf]);  // accumulator complete -> epilogue
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int ew = warp - 4;  // == warp % 4: the TMEM lane quadrant this warp may touch
    uint8_t* my_stage = epi_base + ew * 2 * EPI_BUF_BYTES;
    uint32_t sbuf = 0;
    uint32_t it = 0;
    const int dtype = p.dtype;
    const int keep_ld = p.N >> 3;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, m_blk, n_blk);
      const uint32_t buf = it & 1u;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int row = m_blk * BM + ew * 32 + lane;
      const bool row_ok = row < p.M;

      // Per-row operand of the masked low-rank term, fetched while the main
      // loop of this tile is still running.
      float dhv[16];
      if (p.masked_r > 0 && p.masked_r <= 16) {
#pragma unroll
        for (int j = 0; j < 16; ++j) dhv[j] = 0.f;
        if (row_ok) {
          const uint16_t* dhp = reinterpret_cast<const uint16_t*>(p.dh) + static_cast<size_t>(row) * p.dh_ld;
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j < p.masked_r) dhv[j] = to_float16bits(dhp[j], dtype) * p.inv_keep;
        }
      }

      mbar_wait(&tmem_full[buf], acc_phase, 4);
      tc_fence_after();

      const bool warp_rows_ok = m_blk * BM + ew * 32 < p.M;  // else this warp's 32 rows are all padding
      for (int cb = 0; cb < BN / 64 && warp_rows_ok; ++cb) {
        const int col0 = n_blk * BN + cb * 64;
        if (col0 >= p.N) break;  // warp-uniform
        if (lane == 0) bulk_wait_group_read<1>();  // the buffer we are about to overwrite has been read
        __syncwarp();
        uint8_t* sdst = my_stage + sbuf * EPI_BUF_BYTES + lane * 128;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + buf * BN + cb * 64 + h * 32, v);
          tmem_ld_wait();
          const int c = col0 + h * 32;
          if (p.bias != nullptr) {
            const uint4* bp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.bias) + c);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (c + q * 8 < p.N) {
                const uint4 bw = __ldg(bp + q);
                const uint32_t w4[4] = {bw.x, bw.y, bw.z, bw.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 f = unpack2(w4[e], dtype);
                  v[q * 8 + e * 2] = __float_as_uint(__uint_as_float(v[q * 8 + e * 2]) + f.x);
                  v[q * 8 + e * 2 + 1] = __float_as_uint(__uint_as_float(v[q * 8 + e * 2 + 1]) + f.y);
                }
              }
            }
          }
          if (p.masked_r > 0) {
            // low[i] = sum_j dh[row,j] * a16[j, c+i]; added where the dropout mask kept X[row, c+i].
            float low[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) low[i] = 0.f;
            if (p.masked_r <= 16) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (j < p.masked_r) {
                  const uint4* ap = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.a16) +
                                                                   static_cast<size_t>(j) * p.N + c);
#pragma unroll
                  for (int q = 0; q < 4; ++q) {
                    if (c + q * 8 < p.N) {
                      const uint4 aw = __ldg(ap + q);
                      const uint32_t w4[4] = {aw.x, aw.y, aw.z, aw.w};
#pragma unroll
                      for (int e = 0; e < 4; ++e) {
                        const float2 f = unpack2(w4[e], dtype);
                        low[q * 8 + e * 2] = fmaf(dhv[j], f.x, low[q * 8 + e * 2]);
                        low[q * 8 + e * 2 + 1] = fmaf(dhv[j], f.y, low[q * 8 + e * 2 + 1]);
                      }
                    }
                  }
                }
              }
            } else {
              // Wide ranks: dh row is re-read from L1/L2 instead of living in registers.
              const uint16_t* dhp = reinterpret_cast<const uint16_t*>(p.dh) + static_cast<size_t>(row) * p.dh_ld;
              for (int j = 0; j < p.masked_r; ++j) {
                const float d = row_ok ? to_float16bits(dhp[j], dtype) * p.inv_keep : 0.f;
                const uint4* ap = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.a16) +
                                                                 static_cast<size_t>(j) * p.N + c);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  if (c + q * 8 < p.N) {
                    const uint4 aw = __ldg(ap + q);
                    const uint32_t w4[4] = {aw.x, aw.y, aw.z, aw.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                      const float2 f = unpack2(w4[e], dtype);
                      low[q * 8 + e * 2] = fmaf(d, f.x, low[q * 8 + e * 2]);
                      low[q * 8 + e * 2 + 1] = fmaf(d, f.y, low[q * 8 + e * 2 + 1]);
                    }
                  }
                }
              }
            }
            uint32_t kbits = 0;
            if (row_ok) {
              const uint8_t* kp = p.keep + static_cast<size_t>(row) * keep_ld + (c >> 3);
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (c + q * 8 < p.N) kbits |= static_cast<uint32_t>(__ldg(kp + q)) << (8 * q);
            }
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if ((kbits >> i) & 1u) v[i] = __float_as_uint(__uint_as_float(v[i]) + low[i]);
          }
          // fp32 -> 16-bit, 128B-swizzled staging (chunk index XOR row-in-atom)
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            o.x = pack2(__uint_as_float(v[q * 8 + 0]), __uint_as_float(v[q * 8 + 1]), dtype);
            o.y = pack2(__uint_as_float(v[q * 8 + 2]), __uint_as_float(v[q * 8 + 3]), dtype);
            o.z = pack2(__uint_as_float(v[q * 8 + 4]), __uint_as_float(v[q * 8 + 5]), dtype);
            o.w = pack2(__uint_as_float(v[q * 8 + 6]), __uint_as_float(v[q * 8 + 7]), dtype);
            const int chunk = h * 4 + q;
            *reinterpret_cast<uint4*>(sdst + ((chunk ^ (lane & 7)) << 4)) = o;
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmD, my_stage + sbuf * EPI_BUF_BYTES, col0, m_blk * BM + ew * 32);
          bulk_commit_group();
        }
        sbuf ^= 1u;
      }
      // All TMEM reads of this accumulator are done: hand it back to the MMA warp.
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[buf]);
    }
    if (lane == 0) bulk_wait_group_read<0>();
    __syncwarp();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace gemm
}  // namespace lb
