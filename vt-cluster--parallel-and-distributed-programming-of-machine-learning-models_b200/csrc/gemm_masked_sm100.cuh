// Backward dX GEMM of a LoRA-adapted projection WITH dropout:
//
//   dX[M,N] = dY[M,K1] * W  +  keep[M,N] / (1-p) .* ( dH[M,rp] * A[rp,N] )          (N = in_features here)
//
// The mask is per output element, so the rank-r product T = dH A cannot ride the main accumulator the way
// the forward's H (sB)^T does.  Here T gets an accumulator of its own in tensor memory: tiles are
// 256 x 128 (2-CTA pair, tcgen05.mma.cta_group::2, UMMA 256x128x16), the 512 TMEM columns hold
// acc[2] and T[2] (both double-buffered, 128 columns each), T is ONE extra K-block (K = rp) issued
// ahead of the main loop, and the epilogue is two tcgen05.ld per 32-column chunk plus one select + FMA per
// element, whatever the rank.  The other flavour (gemm_sm100.cuh, kMasked) forms T in the epilogue warps -- a
// CUDA-core loop over the rank, or warp-level MMAs -- which keeps the faster 256-column main loop but makes the
// epilogue the bound for short contractions and wide ranks; the host picks per shape (lora_b200.cu,
// use_masked_tmem, with the measurements).  Same warp roles, rings and watchdogs as gemm_lora_kernel; pair mode
// only (M > 128).  Every tensor map covers this GPU's memory.
#pragma once

#include "gemm_sm100.cuh"

namespace lb {
namespace gemm {

constexpr int MBN = 128;                                      // tile width
constexpr int M_B_ROWS = MBN / 2;                             // rows of the B tile one CTA of the pair stages
constexpr int M_B_TILE_BYTES = M_B_ROWS * BK * 2;             // 8 KiB
// A stage holds TWO 64-deep K sub-blocks (A0 | A1 | B0 | B1): with 128-column tiles an MMA takes half as long as in
// the 256-column kernel, and one barrier round trip per 64-deep block (measured: 244 us against 152 us at
// 2560 -> 2560) leaves the single issuing thread behind the tensor pipe; 128-deep stages restore 8 MMAs per wait.
constexpr int M_KSUB = 2;
constexpr int M_STAGE_BYTES = M_KSUB * (A_TILE_BYTES + M_B_TILE_BYTES);  // 48 KiB
constexpr int M_STAGES = 4;
static_assert(M_STAGES * M_STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 <= SMEM_BYTES, "masked GEMM smem budget");

struct MaskedParams {
  int M, N, K1;
  int r_pad;       // 16 / 32 / 64: K of the rank-r block
  int b_mn_major;  // 0: B1 stored [N,K1] (K-major); 1: stored [K1,N]
  int dtype;
  const uint8_t* keep;  // [M, N/8] keep bits
  float inv_keep;
};

__global__ void __launch_bounds__(THREADS, 1)
    gemm_masked_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                       const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                       const __grid_constant__ CUtensorMap tmD, const MaskedParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint8_t* epi_base = smem + M_STAGES * M_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EPI_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + M_STAGES;
  uint64_t* tmem_full = bars + 2 * M_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();  // rank 0 is the leader (issues every MMA, owns full / tmem_empty)
  const bool is_leader = cta_rank == 0;

  if (warp == PRODUCER_WARP && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    tma_prefetch_desc(&tmA2);
    tma_prefetch_desc(&tmB2);
    tma_prefetch_desc(&tmD);
  }
  if (warp == MMA_WARP && lane == 0) {
    for (int i = 0; i < M_STAGES; ++i) {
      mbar_init(&full[i], 2);   // leader's expect_tx arrive + the peer's remote arrive
      mbar_init(&empty[i], 1);  // one multicast tcgen05.commit per use
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], EPI_WARPS * 2);  // both CTAs' epilogues release it
    }
    fence_barrier_init();
  }
  if (warp == ALLOC_WARP) {
    tmem_alloc_2sm(tmem_slot, TMEM_COLS);
    tmem_relinquish_2sm();
  }
  pdl_launch_dependents();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  constexpr int TILE_M = 2 * BM;
  const int num_m = (p.M + TILE_M - 1) / TILE_M;
  const int num_n = (p.N + MBN - 1) / MBN;
  const int num_tiles = num_m * num_n;
  const int nk_main = (p.K1 + M_KSUB * BK - 1) / (M_KSUB * BK);
  const int nk = nk_main + 1;  // block 0 = the rank-r product
  const int tile0 = static_cast<int>(blockIdx.x) / 2;
  const int tile_step = static_cast<int>(gridDim.x) / 2;

  if (warp == PRODUCER_WARP) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const uint32_t lora_bytes = static_cast<uint32_t>((BM + M_B_ROWS) * p.r_pad * 2) * 2;
      for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        const int row0 = m_blk * TILE_M + static_cast<int>(cta_rank) * BM;       // this CTA's A rows
        const int col0 = n_blk * MBN + static_cast<int>(cta_rank) * M_B_ROWS;    // this CTA's slice of B
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 31);
          uint8_t* sa = stage_base + stage * M_STAGE_BYTES;
          uint8_t* sb = sa + M_KSUB * A_TILE_BYTES;
          if (is_leader) mbar_arrive_expect_tx(&full[stage], kb == 0 ? lora_bytes : M_STAGE_BYTES * 2);
          else mbar_arrive_cluster(&full[stage], 0);
          if (kb == 0) {
            tma_load_2d_2sm(sa, &tmA2, &full[stage], 0, row0);
            tma_load_2d_2sm(sb, &tmB2, &full[stage], 0, col0);
          } else {
#pragma unroll
            for (int sub = 0; sub < M_KSUB; ++sub) {  // a sub-block beyond K1 arrives as zeros (TMA out-of-bounds fill)
              const int k0 = ((kb - 1) * M_KSUB + sub) * BK;
              tma_load_2d_2sm(sa + sub * A_TILE_BYTES, &tmA1, &full[stage], k0, row0);
              if (!p.b_mn_major) tma_load_2d_2sm(sb + sub * M_B_TILE_BYTES, &tmB1, &full[stage], k0, col0);
              else tma_load_2d_2sm(sb + sub * M_B_TILE_BYTES, &tmB1, &full[stage], col0, k0);
            }
          }
          if (++stage == M_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer =====================
    if (lane == 0 && is_leader) {
      const int fmt = (p.dtype == kDtypeBf16) ? 1 : 0;
      const uint32_t idesc_main = make_idesc_f16(fmt, TILE_M, MBN, 0, p.b_mn_major);
      const uint32_t idesc_lora = make_idesc_f16(fmt, TILE_M, MBN, 0, 0);
      const uint32_t lora_sw = static_cast<uint32_t>(p.r_pad) * 2u;  // swizzle span = row bytes
      const uint32_t lora_layout = (lora_sw == 32u) ? 6u : (lora_sw == 64u ? 4u : 2u);
      const uint32_t lora_sbo = 8u * lora_sw;
      const int lora_ksteps = p.r_pad / 16;
      const uint32_t smem0 = smem_u32(stage_base);
      const uint64_t da_main0 = make_smem_desc(smem0, 16, 1024, 2);
      const uint64_t db_main0 = p.b_mn_major ? make_smem_desc(smem0 + M_KSUB * A_TILE_BYTES, BK * 128, 1024, 2)
                                             : make_smem_desc(smem0 + M_KSUB * A_TILE_BYTES, 16, 1024, 2);
      const uint64_t da_lora0 = make_smem_desc(smem0, 16, lora_sbo, lora_layout);
      const uint64_t db_lora0 = make_smem_desc(smem0 + M_KSUB * A_TILE_BYTES, 16, lora_sbo, lora_layout);
      const uint32_t b_kstep = p.b_mn_major ? (2048u >> 4) : (32u >> 4);  // K advance of 16 elements
      constexpr uint32_t kStageUnits = M_STAGE_BYTES >> 4;
      uint32_t stage = 0, phase = 0;
      uint32_t it = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
        const uint32_t buf = it & 1u;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(&tmem_empty[buf], acc_phase ^ 1u, 32);
        tc_fence_after();
        const uint32_t d_acc = tmem_base + buf * MBN;
        const uint32_t d_t = tmem_base + 2 * MBN + buf * MBN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&full[stage], phase, 33);
          tc_fence_after();
          const uint32_t soff = stage * kStageUnits;
          if (kb == 0) {
            const uint64_t da = da_lora0 + soff, db = db_lora0 + soff;
            for (int k = 0; k < lora_ksteps; ++k) umma_f16_2sm(d_t, da + k * 2u, db + k * 2u, idesc_lora, k != 0 ? 1u : 0u);
          } else {
#pragma unroll
            for (uint32_t sub = 0; sub < M_KSUB; ++sub) {
              const uint64_t da = da_main0 + soff + sub * (A_TILE_BYTES >> 4);
              const uint64_t db = db_main0 + soff + sub * (M_B_TILE_BYTES >> 4);
#pragma unroll
              for (uint32_t k = 0; k < BK / 16; ++k)
                umma_f16_2sm(d_acc, da + k * 2u, db + k * b_kstep, idesc_main, (kb != 1 || sub != 0 || k != 0) ? 1u : 0u);
            }
          }
          umma_commit_2sm(&empty[stage], 0x3);  // frees the smem slot in both CTAs once these MMAs retire
          if (++stage == M_STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_2sm(&tmem_full[buf], 0x3);  // both accumulators complete -> epilogue (of both CTAs)
      }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0 && warp < EPI_WARP0 + EPI_WARPS) {
    // ===================== epilogue =====================
    const int ew = warp - EPI_WARP0;  // == warp % 4: the TMEM lane quadrant this warp may touch
    uint8_t* my_stage = epi_base + ew * 2 * EPI_BUF_BYTES;
    uint32_t sbuf = 0;
    uint32_t it = 0;
    const int dtype = p.dtype;
    const int keep_ld = p.N >> 3;
    const float inv_keep = p.inv_keep;
    const bool no_main = p.K1 == 0;  // K1 = 0: the masked rank-r term alone, D = keep/(1-p) .* (dH A)
    for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, m_blk, n_blk);
      const uint32_t buf = it & 1u;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int rbase = m_blk * TILE_M + static_cast<int>(cta_rank) * BM + ew * 32;  // first row of this warp
      const int row = rbase + lane;
      // This row's keep-bits for the tile's 4 x 32 columns, requested while the main loop is still running.
      uint32_t kb4[4] = {0u, 0u, 0u, 0u};
      if (row < p.M) {
        const uint8_t* kp0 = p.keep + static_cast<size_t>(row) * keep_ld + ((n_blk * MBN) >> 3);
        const bool fast = ((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep))) & 3) == 0;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int c = n_blk * MBN + q4 * 32;
          if (fast && c + 32 <= p.N) {
            kb4[q4] = __ldg(reinterpret_cast<const uint32_t*>(kp0 + q4 * 4));
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (c + q * 8 < p.N) kb4[q4] |= static_cast<uint32_t>(__ldg(kp0 + q4 * 4 + q)) << (8 * q);
          }
        }
      }
      mbar_wait(&tmem_full[buf], acc_phase, 34);
      tc_fence_after();

      if (rbase < p.M) {  // else this warp's 32 rows are all padding
        const int ncols = min(MBN, p.N - n_blk * MBN);
        const uint32_t tacc = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + buf * MBN;
        const uint32_t tlow = tacc + 2 * MBN;
        // dX = acc + keep ? T / (1-p) : 0, sixteen columns at a time to 16-bit, 128B-swizzled staging
        auto emit = [&](const uint32_t (&v)[32], const uint32_t (&t)[32], uint32_t kbits, int h, uint8_t* sdst) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float o[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int i = q * 8 + e;
              const float w = ((kbits >> i) & 1u) ? inv_keep : 0.f;
              o[e] = fmaf(__uint_as_float(t[i]), w, no_main ? 0.f : __uint_as_float(v[i]));
            }
            uint4 pk;
            pk.x = pack2(o[0], o[1], dtype);
            pk.y = pack2(o[2], o[3], dtype);
            pk.z = pack2(o[4], o[5], dtype);
            pk.w = pack2(o[6], o[7], dtype);
            *reinterpret_cast<uint4*>(sdst + (((h * 4 + q) ^ (lane & 7)) << 4)) = pk;
          }
        };
        uint32_t va[32], ta[32], vb[32], tb[32];
        tmem_ld_32x32b_x32(tacc, va);
        tmem_ld_32x32b_x32(tlow, ta);
#pragma unroll
        for (int cb = 0; cb < MBN / 64; ++cb) {
          if (cb * 64 >= ncols) break;  // warp-uniform
          // the staging buffer about to be overwritten must have been read by its TMA store
          if (lane == 0) bulk_wait_group_read<1>();
          __syncwarp();
          uint8_t* sdst = my_stage + sbuf * EPI_BUF_BYTES + lane * 128;
          const bool has_b = cb * 64 + 32 < ncols;
          const bool has_next = (cb + 1) * 64 < ncols;
          tmem_ld_wait();
          if (has_b) {
            tmem_ld_32x32b_x32(tacc + cb * 64 + 32, vb);
            tmem_ld_32x32b_x32(tlow + cb * 64 + 32, tb);
          }
          emit(va, ta, kb4[cb * 2], 0, sdst);
          if (has_b) {
            tmem_ld_wait();
            if (has_next) {
              tmem_ld_32x32b_x32(tacc + (cb + 1) * 64, va);
              tmem_ld_32x32b_x32(tlow + (cb + 1) * 64, ta);
            }
            emit(vb, tb, kb4[cb * 2 + 1], 1, sdst);
          } else if (has_next) {  // cannot happen (has_next implies has_b); keeps the pipeline well formed
            tmem_ld_32x32b_x32(tacc + (cb + 1) * 64, va);
            tmem_ld_32x32b_x32(tlow + (cb + 1) * 64, ta);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmD, my_stage + sbuf * EPI_BUF_BYTES, n_blk * MBN + cb * 64, rbase);
            bulk_commit_group();
          }
          sbuf ^= 1u;
        }
        tmem_ld_wait();
      }
      // All TMEM reads of these accumulators are done: hand them back to the MMA warp.
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tmem_empty[buf], 0);
    }
    if (lane == 0) bulk_wait_group_read<0>();
    __syncwarp();
  }

  tc_fence_before();
  cluster_sync_all();  // the peer's smem / TMEM / barriers stay alive until both CTAs are done
  if (warp == ALLOC_WARP) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
}

}  // namespace gemm
}  // namespace lb
