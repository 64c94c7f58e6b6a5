// Persistent, warp-specialised tcgen05 GEMM for the frozen projection of a
// LoRA-adapted linear layer, with the rank-r update folded in as one extra
// K-block and the bias / masked low-rank term applied in the epilogue.
//
//   D[M,N] = A1[M,K1] * B1^T  (+ A2[M,rp] * B2[N,rp]^T)  (+ bias[N])
//            (+ keep[M,N] * inv_keep * (dh[M,:rp] * a16[N,:rp]^T))
//
// forward  (peft lora.Linear.forward):
//     A1 = X, B1 = W, A2 = H = drop(X) A^T, B2 = s*B, bias = b
// backward (autograd of the same):
//     A1 = dY, B1 = W (other major-ness), p == 0: A2 = dH, B2 = A^T
//                                         p  > 0: masked epilogue with dh, A^T, keep bits
//
// Two launch shapes share this code (template kPair): a lone CTA per 128x256 tile, or a 2-CTA
// cluster per 256x256 tile driving tcgen05.mma.cta_group::2 (see Cfg below).
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
constexpr int A_TILE_BYTES = BM * BK * 2;  // 16 KiB: the CTA's own 128 rows of A1
// kPair = false: one CTA per 128x256 tile, B tile = 256 rows (32 KiB), 4 stages of 48 KiB.
// kPair = true : a 2-CTA cluster per 256x256 tile (tcgen05 cta_group::2, UMMA M = 256); each
//                CTA stages its own 128 rows of A1 and HALF of the B tile (128 rows, 16 KiB),
//                so L2 -> SMEM traffic per flop halves and 6 stages of 32 KiB fit.
// kMasked trades one ring stage for a 32 KiB scratch region (the epilogue's staged A slice or
// the warp-MMA transposition buffers), so that output staging stays double-buffered.
constexpr int MASK_SCRATCH_BYTES = 32 * 1024;
template <bool kPair, bool kMasked = false>
struct Cfg {
  static constexpr int STAGES = (kPair ? 6 : 4) - (kMasked ? 1 : 0);
  static constexpr int B_ROWS = kPair ? BN / 2 : BN;  // rows of the B tile held by one CTA
  static constexpr int B_TILE_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_TILE_BYTES + B_TILE_BYTES;
  static constexpr int CTAS = kPair ? 2 : 1;
};
constexpr int EPI_WARPS = 4;
// Warp roles.  The four epilogue warps must cover the four TMEM lane quadrants (warp % 4); which of the
// two warps of a scheduler is the epilogue warp is free.  LB_EPI_FIRST = 1 gives the epilogue the LOW
// warp ids (0..3) and the producer / MMA issuer / TMEM allocator ids 4 / 5 / 6, so that on the two
// schedulers they share, the single-thread producer and issuer are the higher-numbered warps.
#ifndef LB_EPI_FIRST
#define LB_EPI_FIRST 0
#endif
constexpr int EPI_WARP0 = LB_EPI_FIRST ? 0 : 4;
constexpr int PRODUCER_WARP = LB_EPI_FIRST ? 4 : 0;
constexpr int MMA_WARP = LB_EPI_FIRST ? 5 : 1;
constexpr int ALLOC_WARP = LB_EPI_FIRST ? 6 : 2;
constexpr int EPI_BUF_BYTES = 32 * 128;  // 32 rows x 64 cols x 2 B
constexpr int EPI_BYTES = EPI_WARPS * 2 * EPI_BUF_BYTES;
constexpr int BAR_BYTES = 256;
constexpr int BIAS_BYTES = BN * 4;  // the tile's bias slice as fp32 (unmasked kernels)
constexpr int SMEM_BYTES = 4 * (A_TILE_BYTES + BN * BK * 2) + EPI_BYTES + BAR_BYTES + BIAS_BYTES + 1024;  // same for both
constexpr int THREADS = 256;
constexpr int TMEM_COLS = 512;
#ifndef LB_EXP_GROUP_M
#define LB_EXP_GROUP_M 16
#endif
constexpr int GROUP_M = LB_EXP_GROUP_M;  // m-blocks (of one CTA or one pair) per raster group (L2 reuse of B panels)

struct Params {
  int M, N, K1;
  int r_pad;       // 0: no low-rank K-block; else 16 / 32 / 64
  int b_mn_major;  // 0: B1 stored [N,K1] (K-major); 1: stored [K1,N]
  int dtype;       // kDtypeBf16 / kDtypeFp16
  const void* bias;
  // masked low-rank epilogue (backward with dropout)
  int masked_r;  // 0 = off, else the true rank r (<= 64); selects the kMasked instantiation
  const void* dh;  // [M, dh_ld] dtype, dh_ld = padded rank (16/32/64), padding columns zero
  int dh_ld;
  const void* a16;      // A^T as [N, dh_ld] dtype (rank index contiguous), padding columns zero
  const void* a16_rows; // A as [dh_ld, N] dtype (column index contiguous), padding rows zero
  const uint8_t* keep;  // [M, N/8] keep bits
  float inv_keep;
  // activation epilogue (kAct instantiation, un-adapted neighbours of the LoRA path): D = act(acc + bias)
  int act;      // kActNone / kActRelu / kActGeluNew
  int store_z;  // also store the pre-activation acc + bias through tmZ (gelu_new backward needs it)
  // residual epilogue (plain kernel): D = acc + bias + R, R [M, res_ld] of the activation dtype; R may alias D
  // (each tile is read by the CTA that later stores it)
  const void* residual;
  long long res_ld;
};

constexpr int kActNone = 0, kActRelu = 1, kActGeluNew = 2;

// gelu_new of HF GPT-2 (transformers NewGELUActivation): 0.5 x (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3))).
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gelu_new(float x) {
  const float u = 0.7978845608028654f * fmaf(0.044715f * x * x, x, x);
  return 0.5f * x * (1.f + tanh_approx(u));
}
__device__ __forceinline__ float gelu_new_grad(float x) {
  const float x2 = x * x;
  const float u = 0.7978845608028654f * fmaf(0.044715f * x2, x, x);
  const float t = tanh_approx(u);
  const float du = 0.7978845608028654f * fmaf(3.f * 0.044715f, x2, 1.f);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

__device__ __forceinline__ void tile_coords(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int group_sz = GROUP_M * num_n;
  const int g = tile / group_sz;
  const int first_m = g * GROUP_M;
  const int gm = min(GROUP_M, num_m - first_m);
  const int in_g = tile - g * group_sz;
  m_blk = first_m + in_g % gm;
  n_blk = in_g / gm;
}

// Warp-level m16n8k16 MMA used by the masked epilogue for the rank-r term.
__device__ __forceinline__ void mma_16816_bf16(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma_16816_f16(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// Eight consecutive 16-bit bias values; `left` = columns remaining up to N (ragged N: the tail is
// read element by element so that nothing beyond bias[N-1] is touched).
__device__ __forceinline__ uint4 load_bias8(const uint4* p16, int left) {
  if (left >= 8) return __ldg(p16);
  const uint16_t* s = reinterpret_cast<const uint16_t*>(p16);
  uint32_t w[4] = {0u, 0u, 0u, 0u};
  for (int e = 0; e < left; ++e) w[e >> 1] |= static_cast<uint32_t>(__ldg(s + e)) << ((e & 1) * 16);
  return make_uint4(w[0], w[1], w[2], w[3]);
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// kMasked: backward with dropout -- the rank-r term keep/(1-p) .* (dH A) is formed in the
// epilogue (warp-level MMA + keep-bits) instead of riding the main loop as a K-block.
// kAct (unmasked only): the epilogue applies relu / gelu_new after the bias and can store the
// pre-activation as a second output (tmZ); used for the frozen, un-adapted layers next to the path.
// kFma (masked only): padded rank 16 -- the rank-r term runs on the CUDA cores; its own instantiation so that the
// warp-MMA path's fragments (64+ registers) do not count against it.
template <bool kMasked, bool kPair, bool kAct = false, bool kFma = false>
__global__ void __launch_bounds__(THREADS, 1)
    gemm_lora_kernel(const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmB1,
                     const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2,
                     const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmZ, const Params p) {
  static_assert(!(kMasked && kAct), "the activation epilogue belongs to the unmasked kernel");
  static_assert(kMasked || !kFma, "kFma selects a masked-epilogue flavour");
  using C = Cfg<kPair, kMasked>;
  constexpr int STAGES = C::STAGES;
  constexpr int STAGE_BYTES = C::STAGE_BYTES;
  constexpr int B_ROWS = C::B_ROWS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint8_t* epi_base = smem + STAGES * STAGE_BYTES;
  uint8_t* mask_scratch = epi_base + EPI_BYTES;  // kMasked only
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_base + EPI_BYTES + (kMasked ? MASK_SCRATCH_BYTES : 0));
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* bias_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + BAR_BYTES);  // [BN]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // Pair mode: rank 0 is the leader (issues every MMA, owns the full / tmem_empty barriers).
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const bool is_leader = cta_rank == 0;

  if (warp == PRODUCER_WARP && lane == 0) {
    tma_prefetch_desc(&tmA1);
    tma_prefetch_desc(&tmB1);
    tma_prefetch_desc(&tmD);
    if (p.r_pad > 0) {
      tma_prefetch_desc(&tmA2);
      tma_prefetch_desc(&tmB2);
    }
  }
  if (warp == MMA_WARP && lane == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], C::CTAS);  // pair: leader's expect_tx arrive + the peer's remote arrive
      mbar_init(&empty[i], 1);       // one (multicast) tcgen05.commit per use
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], EPI_WARPS * C::CTAS);  // pair: both CTAs' epilogues release it
    }
    fence_barrier_init();
  }
  if (warp == ALLOC_WARP) {
    if constexpr (kPair) {
      tmem_alloc_2sm(tmem_slot, TMEM_COLS);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(tmem_slot, TMEM_COLS);
      tmem_relinquish();
    }
  }
  pdl_launch_dependents();
  tc_fence_before();
  if constexpr (kPair) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();  // everything above overlapped the previous kernel's tail; its outputs are needed from here on

  // Tiles are BM*CTAS rows tall; a cluster (or lone CTA) walks them persistently.
  constexpr int TILE_M = BM * C::CTAS;
  const int num_m = (p.M + TILE_M - 1) / TILE_M;
  const int num_n = (p.N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int nk_main = (p.K1 + BK - 1) / BK;
  const int nk = nk_main + (p.r_pad > 0 ? 1 : 0);
  const int tile0 = static_cast<int>(blockIdx.x) / C::CTAS;
  const int tile_step = static_cast<int>(gridDim.x) / C::CTAS;

  if (warp == PRODUCER_WARP) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      const uint32_t lora_bytes = static_cast<uint32_t>((BM + B_ROWS) * p.r_pad * 2) * C::CTAS;
      for (int tile = tile0; tile < num_tiles; tile += tile_step) {
        int m_blk, n_blk;
        tile_coords(tile, num_m, num_n, m_blk, n_blk);
        const int row0 = m_blk * TILE_M + static_cast<int>(cta_rank) * BM;       // this CTA's A rows
        const int col0 = n_blk * BN + static_cast<int>(cta_rank) * B_ROWS;       // this CTA's slice of B
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1, 1);
          uint8_t* sa = stage_base + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_TILE_BYTES;
          const bool main_blk = kb < nk_main;
          if (is_leader) mbar_arrive_expect_tx(&full[stage], main_blk ? STAGE_BYTES * C::CTAS : lora_bytes);
          else mbar_arrive_cluster(&full[stage], 0);
          // In pair mode every load credits the LEADER's barrier (tma_load_2d_2sm).
          auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
            if constexpr (kPair) tma_load_2d_2sm(dst, tm, &full[stage], c0, c1);
            else tma_load_2d(dst, tm, &full[stage], c0, c1);
          };
          if (main_blk) {
            load(sa, &tmA1, kb * BK, row0);
            if (!p.b_mn_major) {
              load(sb, &tmB1, kb * BK, col0);
            } else {
#pragma unroll
              for (int j = 0; j < B_ROWS / 64; ++j) load(sb + j * (BK * 128), &tmB1, col0 + j * 64, kb * BK);
            }
          } else {
            load(sa, &tmA2, 0, row0);
            load(sb, &tmB2, 0, col0);
          }
          if (++stage == STAGES) {
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
      const uint32_t idesc_main = make_idesc_f16(fmt, TILE_M, BN, 0, p.b_mn_major);
      const uint32_t idesc_lora = make_idesc_f16(fmt, TILE_M, BN, 0, 0);
      const uint32_t lora_sw = static_cast<uint32_t>(p.r_pad) * 2u;  // swizzle span = row bytes
      const uint32_t lora_layout = (lora_sw == 32u) ? 6u : (lora_sw == 64u ? 4u : 2u);
      const uint32_t lora_sbo = 8u * lora_sw;
      const int lora_ksteps = p.r_pad / 16;
      auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
        if constexpr (kPair) umma_f16_2sm(d, da, db, idesc, acc); else umma_f16(d, da, db, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) {
        if constexpr (kPair) umma_commit_2sm(bar, 0x3); else umma_commit(bar);  // pair: arrive in both CTAs
      };
      // The issuing thread is a serial instruction stream racing a 128-cycle MMA: descriptors are
      // built once (stage 0, k-step 0) and advanced by adding 16-byte units to their low word
      // (the 14-bit start-address field cannot overflow: smem is < 256 KiB).
      const uint32_t smem0 = smem_u32(stage_base);
      const uint64_t da_main0 = make_smem_desc(smem0, 16, 1024, 2);
      const uint64_t db_main0 = p.b_mn_major ? make_smem_desc(smem0 + A_TILE_BYTES, BK * 128, 1024, 2)
                                             : make_smem_desc(smem0 + A_TILE_BYTES, 16, 1024, 2);
      const uint64_t da_lora0 = make_smem_desc(smem0, 16, lora_sbo, lora_layout);
      const uint64_t db_lora0 = make_smem_desc(smem0 + A_TILE_BYTES, 16, lora_sbo, lora_layout);
      const uint32_t b_kstep = p.b_mn_major ? (2048u >> 4) : (32u >> 4);  // K advance of 16 elements
      constexpr uint32_t kStageUnits = STAGE_BYTES >> 4;
      uint32_t stage = 0, phase = 0;
      uint32_t it = 0;
      for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
        const uint32_t buf = it & 1u;
        const uint32_t acc_phase = (it >> 1) & 1u;
        mbar_wait(&tmem_empty[buf], acc_phase ^ 1u, 2);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = 0; kb < nk; ++kb) {
          mbar_wait(&full[stage], phase, 3);
          tc_fence_after();
          const uint32_t soff = stage * kStageUnits;
          if (kb < nk_main) {
            const uint64_t da = da_main0 + soff, db = db_main0 + soff;
            mma(d_tmem, da, db, idesc_main, kb != 0 ? 1u : 0u);
#pragma unroll
            for (uint32_t k = 1; k < BK / 16; ++k) mma(d_tmem, da + k * 2u, db + k * b_kstep, idesc_main, 1u);
          } else {
            const uint64_t da = da_lora0 + soff, db = db_lora0 + soff;
            for (int k = 0; k < lora_ksteps; ++k) mma(d_tmem, da + k * 2u, db + k * 2u, idesc_lora, 1u);
          }
          commit(&empty[stage]);  // frees the smem slot (in both CTAs) once these MMAs retire
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        commit(&tmem_full[buf]);  // accumulator complete -> epilogue (of both CTAs)
      }
    }
    __syncwarp();
  } else if (warp >= EPI_WARP0 && warp < EPI_WARP0 + EPI_WARPS) {
    // ===================== epilogue =====================
    const int ew = warp - EPI_WARP0;  // == warp % 4: the TMEM lane quadrant this warp may touch
    uint8_t* my_stage = epi_base + ew * 2 * EPI_BUF_BYTES;
    // Masked mode, wide ranks: a 32x32 fp32 scratch per warp through which the warp-level MMA
    // result is transposed to one-row-per-thread.
    float* scratch = reinterpret_cast<float*>(mask_scratch + ew * EPI_BUF_BYTES);
    uint32_t sbuf = 0;
    uint32_t it = 0;
    const int dtype = p.dtype;
    const int keep_ld = p.N >> 3;
    const int mk_steps = kMasked ? p.dh_ld / 16 : 0;  // k16 steps of the rank-r product (1, 2 or 4)
    // Masked mode: dH.A for a 32x32 chunk is a handful of warp-level MMAs.  Their B-fragments
    // (A^T, rank index contiguous) come from the tile's 256-column slice of a16t, which the four
    // epilogue warps copy into smem once per tile while the main loop runs (one contiguous block of
    // 256*rp*2 bytes); only the widest rank (rp = 64, slice too big) reads them from L2 per chunk.
    const bool slice_in_smem = kMasked && !kFma && p.dh_ld == 32;
    // Padded rank 16 (every adapter of the reference): the warp-level MMAs would queue behind the
    // UMMA stream on the one tensor pipe, so dH.A is formed on the CUDA cores instead, in a compact
    // runtime loop over the rank: A's 16 x 256 slice as fp32 in smem (row reads are warp-uniform
    // broadcasts), this thread's dH row in a per-warp smem column (scratch + 16 KiB).
    constexpr bool use_fma = kMasked && kFma;
    const float* fslice = reinterpret_cast<const float*>(mask_scratch);                      // [16 j][256 cols]
    float* dh_col = reinterpret_cast<float*>(mask_scratch + 4 * EPI_BUF_BYTES) + ew * 16 * 32;  // [16 j][32 lanes]
    const uint32_t* aslice = reinterpret_cast<const uint32_t*>(mask_scratch + 4 * EPI_BUF_BYTES);  // [256][rp] 16-bit
    for (int tile = tile0; tile < num_tiles; tile += tile_step, ++it) {
      int m_blk, n_blk;
      tile_coords(tile, num_m, num_n, m_blk, n_blk);
      const uint32_t buf = it & 1u;
      const uint32_t acc_phase = (it >> 1) & 1u;
      const int rbase = m_blk * TILE_M + static_cast<int>(cta_rank) * BM + ew * 32;  // first row of this warp
      const int row = rbase + lane;
      const bool row_ok = row < p.M;

      // dH rows of this warp as mma.sync A-fragments (2 m16 blocks x up to 4 k16 steps),
      // fetched while the main loop of this tile is still running.
      uint32_t dh_frag[2][4][4];
      uint32_t kb8[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};  // kFma: this row's keep-bits for the tile's 8 x 32 columns
      if constexpr (kMasked) {
        if (use_fma) {
          // requested before the accumulator is ready: an L2 round trip per 32-column chunk inside the epilogue
          // would sit on the critical path of the CTA's last tile, which nothing overlaps
          if (row_ok) {
            const uint8_t* kp0 = p.keep + static_cast<size_t>(row) * keep_ld + ((n_blk * BN) >> 3);
            const bool fast = ((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep))) & 3) == 0;
#pragma unroll
            for (int q8 = 0; q8 < 8; ++q8) {
              const int c = n_blk * BN + q8 * 32;
              if (fast && c + 32 <= p.N) {
                kb8[q8] = __ldg(reinterpret_cast<const uint32_t*>(kp0 + q8 * 4));
              } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  if (c + q * 8 < p.N) kb8[q8] |= static_cast<uint32_t>(__ldg(kp0 + q8 * 4 + q)) << (8 * q);
              }
            }
          }
          asm volatile("bar.sync 2, 128;" ::: "memory");  // every warp is done with the previous tile's slice
          const int t128 = ew * 32 + lane;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int g = t128 + 128 * i;  // 512 groups of 8 columns: j = g / 32, column group g % 32
            const int j = g >> 5, cg = g & 31;
            const int col = n_blk * BN + cg * 8;
            uint4 raw = make_uint4(0u, 0u, 0u, 0u);
            if (col < p.N)
              raw = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.a16_rows) +
                                                        static_cast<size_t>(j) * p.N + col));
            const float2 f0 = unpack2(raw.x, dtype), f1 = unpack2(raw.y, dtype), f2 = unpack2(raw.z, dtype),
                         f3 = unpack2(raw.w, dtype);
            float* dstA = reinterpret_cast<float*>(mask_scratch) + j * BN + cg * 8;
            *reinterpret_cast<float4*>(dstA) = make_float4(f0.x, f0.y, f1.x, f1.y);
            *reinterpret_cast<float4*>(dstA + 4) = make_float4(f2.x, f2.y, f3.x, f3.y);
          }
          uint4 d0 = make_uint4(0u, 0u, 0u, 0u), d1 = d0;
          if (row_ok) {
            const uint4* dhp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.dh) +
                                                              static_cast<size_t>(row) * 16);
            d0 = __ldg(dhp);
            d1 = __ldg(dhp + 1);
          }
          const uint32_t w8[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float2 f = unpack2(w8[e], dtype);
            dh_col[(2 * e) * 32 + lane] = f.x * p.inv_keep;
            dh_col[(2 * e + 1) * 32 + lane] = f.y * p.inv_keep;
          }
          asm volatile("bar.sync 2, 128;" ::: "memory");  // slice (and every warp's dH column) complete
        }
        if (slice_in_smem) {
          asm volatile("bar.sync 2, 128;" ::: "memory");  // every warp is done with the previous tile's slice
          const int ncol = min(BN, p.N - n_blk * BN);
          const int n16 = ncol * p.dh_ld / 8;  // 16-byte units to copy (contiguous in a16t)
          const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.a16) +
                                                            static_cast<size_t>(n_blk) * BN * p.dh_ld);
          uint4* dst = reinterpret_cast<uint4*>(mask_scratch + 4 * EPI_BUF_BYTES);
          for (int i = ew * 32 + lane; i < BN * p.dh_ld / 8; i += 128)
            dst[i] = i < n16 ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
          asm volatile("bar.sync 2, 128;" ::: "memory");  // slice complete
        }
      }
      if constexpr (kMasked) if (!use_fma) {
        const uint32_t* dh32 = reinterpret_cast<const uint32_t*>(p.dh);
#pragma unroll
        for (int mb = 0; mb < 2; ++mb)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            const int r0 = rbase + mb * 16 + (lane >> 2);
            const int k0 = ks * 16 + (lane & 3) * 2;
            const bool ks_ok = ks < mk_steps;
            const size_t o0 = (static_cast<size_t>(r0) * p.dh_ld + k0) >> 1;
            const size_t o1 = (static_cast<size_t>(r0 + 8) * p.dh_ld + k0) >> 1;
            dh_frag[mb][ks][0] = (ks_ok && r0 < p.M) ? __ldg(dh32 + o0) : 0u;
            dh_frag[mb][ks][1] = (ks_ok && r0 + 8 < p.M) ? __ldg(dh32 + o1) : 0u;
            dh_frag[mb][ks][2] = (ks_ok && r0 < p.M) ? __ldg(dh32 + o0 + 4) : 0u;
            dh_frag[mb][ks][3] = (ks_ok && r0 + 8 < p.M) ? __ldg(dh32 + o1 + 4) : 0u;
          }
      }

      if constexpr (!kMasked) {
        // The tile's 256 bias values go to smem (fp32) while the main loop is still running: with
        // ~225 KiB of the SM's 256 KiB configured as shared memory there is no L1 to speak of, and a
        // per-chunk __ldg would stall every chunk for an L2 round trip.
        if (p.bias != nullptr) {
          asm volatile("bar.sync 2, 128;" ::: "memory");  // every warp is done with the previous tile's slice
          const int t128 = ew * 32 + lane;
          const int col = n_blk * BN + 2 * t128;
          const uint16_t* bp = reinterpret_cast<const uint16_t*>(p.bias) + col;
          uint32_t w = 0u;
          if (col + 1 < p.N) w = __ldg(reinterpret_cast<const uint32_t*>(bp));
          else if (col < p.N) w = __ldg(bp);
          const float2 f = unpack2(w, dtype);
          *reinterpret_cast<float2*>(bias_s + 2 * t128) = f;
          asm volatile("bar.sync 2, 128;" ::: "memory");
        }
      }

      mbar_wait(&tmem_full[buf], acc_phase, 4);
      tc_fence_after();

      const bool warp_rows_ok = rbase < p.M;  // else this warp's 32 rows are all padding
      if constexpr (kAct) {
        // Activation epilogue, software-pipelined like the plain one.  With the pre-activation asked
        // for, each 64-column group is staged twice (Z in buffer 0, act(Z) in buffer 1) and leaves
        // through two TMA stores; without it the two buffers alternate.
        if (warp_rows_ok) {
          const int nchunks = (min(BN, p.N - n_blk * BN) + 31) / 32;  // warp-uniform, >= 1
          const uint32_t tbase = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + buf * BN;
          const bool store_z = p.store_z != 0;
          const int act = p.act;
          auto stage = [&](const uint32_t (&v)[32], int h, uint8_t* dst) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack2(__uint_as_float(v[q * 8 + 0]), __uint_as_float(v[q * 8 + 1]), dtype);
              o.y = pack2(__uint_as_float(v[q * 8 + 2]), __uint_as_float(v[q * 8 + 3]), dtype);
              o.z = pack2(__uint_as_float(v[q * 8 + 4]), __uint_as_float(v[q * 8 + 5]), dtype);
              o.w = pack2(__uint_as_float(v[q * 8 + 6]), __uint_as_float(v[q * 8 + 7]), dtype);
              *reinterpret_cast<uint4*>(dst + (((h * 4 + q) ^ (lane & 7)) << 4)) = o;
            }
          };
          auto emit = [&](uint32_t (&v)[32], int chunk, int h, uint8_t* zdst, uint8_t* ydst) {
            if (p.bias != nullptr) {
              const float4* bs = reinterpret_cast<const float4*>(bias_s + chunk * 32);
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 b4 = bs[q];  // warp-uniform address: a broadcast read
                v[q * 4 + 0] = __float_as_uint(__uint_as_float(v[q * 4 + 0]) + b4.x);
                v[q * 4 + 1] = __float_as_uint(__uint_as_float(v[q * 4 + 1]) + b4.y);
                v[q * 4 + 2] = __float_as_uint(__uint_as_float(v[q * 4 + 2]) + b4.z);
                v[q * 4 + 3] = __float_as_uint(__uint_as_float(v[q * 4 + 3]) + b4.w);
              }
            }
            if (store_z) stage(v, h, zdst);
            if (act == kActGeluNew) {
#pragma unroll
              for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(gelu_new(__uint_as_float(v[e])));
            } else {
#pragma unroll
              for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(fmaxf(__uint_as_float(v[e]), 0.f));
            }
            stage(v, h, ydst);
          };
          uint32_t va[32], vb[32];
          tmem_ld_32x32b_x32(tbase, va);
          for (int ci = 0; ci < nchunks; ci += 2) {
            const int col0 = n_blk * BN + ci * 32;
            const bool has_b = ci + 1 < nchunks;
            // the staging buffer(s) about to be overwritten must have been read by their TMA stores
            if (lane == 0) {
              if (store_z) bulk_wait_group_read<0>(); else bulk_wait_group_read<1>();
            }
            __syncwarp();
            uint8_t* zbuf = my_stage;
            uint8_t* ybuf = my_stage + (store_z ? 1u : sbuf) * EPI_BUF_BYTES;
            tmem_ld_wait();
            if (has_b) tmem_ld_32x32b_x32(tbase + (ci + 1) * 32, vb);
            emit(va, ci, 0, zbuf + lane * 128, ybuf + lane * 128);
            if (has_b) {
              tmem_ld_wait();
              if (ci + 2 < nchunks) tmem_ld_32x32b_x32(tbase + (ci + 2) * 32, va);
              emit(vb, ci + 1, 1, zbuf + lane * 128, ybuf + lane * 128);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (store_z) tma_store_2d(&tmZ, zbuf, col0, rbase);
              tma_store_2d(&tmD, ybuf, col0, rbase);
              bulk_commit_group();
            }
            sbuf ^= 1u;
          }
        }
      } else if constexpr (!kMasked) {
        // Plain epilogue, software-pipelined: the TMEM load of the next 32-column chunk is in flight
        // while the current one is biased, converted and staged (two register sets alternate).
        if (warp_rows_ok) {
          const int nchunks = (min(BN, p.N - n_blk * BN) + 31) / 32;  // warp-uniform, >= 1
          const uint32_t tbase = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + buf * BN;
          // Residual rows come straight from global memory, 64 bytes per thread and chunk, requested one pair of
          // chunks ahead (two register sets) so that an L2 round trip is covered by the chunk in between.
          const bool has_res = p.residual != nullptr;
          const uint16_t* rrow = reinterpret_cast<const uint16_t*>(p.residual) + static_cast<long long>(row) * p.res_ld;
          auto load_res = [&](uint4 (&r)[4], int c) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              r[q] = (row < p.M && c + q * 8 < p.N) ? *reinterpret_cast<const uint4*>(rrow + c + q * 8)  // R may be D: no ld.nc
                                                     : make_uint4(0u, 0u, 0u, 0u);
          };
          auto add_res = [&](uint32_t (&v)[32], const uint4 (&r)[4]) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const uint32_t w[4] = {r[q].x, r[q].y, r[q].z, r[q].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = unpack2(w[e], dtype);
                v[q * 8 + 2 * e] = __float_as_uint(__uint_as_float(v[q * 8 + 2 * e]) + f.x);
                v[q * 8 + 2 * e + 1] = __float_as_uint(__uint_as_float(v[q * 8 + 2 * e + 1]) + f.y);
              }
            }
          };
          uint4 ra[4], rb[4];
          if (has_res) {
            load_res(ra, n_blk * BN);
            load_res(rb, n_blk * BN + 32);
          }
          auto emit = [&](uint32_t (&v)[32], int c, int h, uint8_t* sdst) {
            if (p.bias != nullptr) {
              const float4* bs = reinterpret_cast<const float4*>(bias_s + (c - n_blk * BN));
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 b4 = bs[q];  // warp-uniform address: a broadcast read
                v[q * 4 + 0] = __float_as_uint(__uint_as_float(v[q * 4 + 0]) + b4.x);
                v[q * 4 + 1] = __float_as_uint(__uint_as_float(v[q * 4 + 1]) + b4.y);
                v[q * 4 + 2] = __float_as_uint(__uint_as_float(v[q * 4 + 2]) + b4.z);
                v[q * 4 + 3] = __float_as_uint(__uint_as_float(v[q * 4 + 3]) + b4.w);
              }
            }
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
          };
          uint32_t va[32], vb[32];
          tmem_ld_32x32b_x32(tbase, va);
          for (int ci = 0; ci < nchunks; ci += 2) {
            // the staging buffer we are about to overwrite must have been read by its TMA store
            if (lane == 0) bulk_wait_group_read<1>();
            __syncwarp();
            uint8_t* sdst = my_stage + sbuf * EPI_BUF_BYTES + lane * 128;
            const int col0 = n_blk * BN + ci * 32;
            const bool has_b = ci + 1 < nchunks;
            tmem_ld_wait();
            if (has_b) tmem_ld_32x32b_x32(tbase + (ci + 1) * 32, vb);
            if (has_res) {
              add_res(va, ra);
              if (ci + 2 < nchunks) load_res(ra, col0 + 64);
            }
            emit(va, col0, 0, sdst);
            if (has_b) {
              tmem_ld_wait();
              if (ci + 2 < nchunks) tmem_ld_32x32b_x32(tbase + (ci + 2) * 32, va);
              if (has_res) {
                add_res(vb, rb);
                if (ci + 3 < nchunks) load_res(rb, col0 + 96);
              }
              emit(vb, col0 + 32, 1, sdst);
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(&tmD, my_stage + sbuf * EPI_BUF_BYTES, col0, rbase);
              bulk_commit_group();
            }
            sbuf ^= 1u;
          }
        }
      }
      for (int cb = 0; cb < BN / 64 && warp_rows_ok && kMasked; ++cb) {
        const int col0 = n_blk * BN + cb * 64;
        if (col0 >= p.N) break;  // warp-uniform
        // the staging buffer we are about to overwrite must have been read by its TMA store
        if (lane == 0) bulk_wait_group_read<1>();
        __syncwarp();
        uint8_t* sdst = my_stage + sbuf * EPI_BUF_BYTES + lane * 128;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + buf * BN + cb * 64 + h * 32, v);
          const int c = col0 + h * 32;
          uint32_t kbits = 0;
          if constexpr (kMasked && kFma) kbits = kb8[h];  // rotated by two per 64-column group (static indices only)
          if constexpr (kMasked && !kFma) {
            // keep-bits of this thread's row for the 32 columns: requested now, consumed after the MMA
            if (row_ok) {
              const uint8_t* kp = p.keep + static_cast<size_t>(row) * keep_ld + (c >> 3);
              if (((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep))) & 3) == 0 && c + 32 <= p.N) {
                kbits = __ldg(reinterpret_cast<const uint32_t*>(kp));
              } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                  if (c + q * 8 < p.N) kbits |= static_cast<uint32_t>(__ldg(kp + q)) << (8 * q);
              }
            }
          }
          if constexpr (kMasked) if (!use_fma) {
            // T[32 rows, 32 cols] = dH[32, rp] * A[rp, c..c+31] on the warp's tensor-core path,
            // overlapping the TMEM load above.  B-fragments come from a16t [N, rp] (j contiguous).
            float t[2][4][4];
#pragma unroll
            for (int mb = 0; mb < 2; ++mb)
#pragma unroll
              for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                for (int e = 0; e < 4; ++e) t[mb][nb][e] = 0.f;
            const uint32_t* at32 = reinterpret_cast<const uint32_t*>(p.a16);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
              if (ks < mk_steps) {
                uint32_t bfr[4][2];
#pragma unroll
                for (int nb = 0; nb < 4; ++nb) {
                  const int col = c + nb * 8 + (lane >> 2);
                  if (slice_in_smem) {
                    const int o = ((col - n_blk * BN) * p.dh_ld + ks * 16 + (lane & 3) * 2) >> 1;
                    bfr[nb][0] = aslice[o];
                    bfr[nb][1] = aslice[o + 4];
                  } else {
                    const size_t o = (static_cast<size_t>(col) * p.dh_ld + ks * 16 + (lane & 3) * 2) >> 1;
                    bfr[nb][0] = col < p.N ? __ldg(at32 + o) : 0u;
                    bfr[nb][1] = col < p.N ? __ldg(at32 + o + 4) : 0u;
                  }
                }
#pragma unroll
                for (int mb = 0; mb < 2; ++mb)
#pragma unroll
                  for (int nb = 0; nb < 4; ++nb) {
                    if (dtype == kDtypeBf16) mma_16816_bf16(t[mb][nb], dh_frag[mb][ks], bfr[nb]);
                    else mma_16816_f16(t[mb][nb], dh_frag[mb][ks], bfr[nb]);
                  }
              }
            }
            // fragment (row = mb*16 + lane/4 [+8], col = nb*8 + (lane%4)*2 [+1]) -> scratch,
            // float4 slots XOR-swizzled by row so that both sides are bank-conflict free.
            __syncwarp();  // previous chunk's readers are done with the scratch
#pragma unroll
            for (int mb = 0; mb < 2; ++mb)
#pragma unroll
              for (int nb = 0; nb < 4; ++nb)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh) {
                  const int rr = mb * 16 + (lane >> 2) + hh * 8;
                  const int cc = nb * 8 + (lane & 3) * 2;
                  float2 val = make_float2(t[mb][nb][hh * 2], t[mb][nb][hh * 2 + 1]);
                  *reinterpret_cast<float2*>(scratch + rr * 32 + (((cc >> 2) ^ (rr & 7)) << 2) + (cc & 3)) = val;
                }
            __syncwarp();
          }
          tmem_ld_wait();
          if (p.bias != nullptr) {
            const uint4* bp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.bias) + c);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              if (c + q * 8 < p.N) {
                const uint4 bw = load_bias8(bp + q, p.N - (c + q * 8));
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
          if constexpr (kMasked) if (use_fma) {
            // low[i] = sum_j dH'[row, j] * A[j, c + i], as packed FP32 pairs (FFMA2).  Measured on the headline
            // shape (scripts/masked_probe.py): the kernel pays ~1.45 us per rank iteration of this loop with scalar
            // FFMA, ~1.2 us with FFMA2, whatever the unroll factor; halving the LDS count (16-bit slice, widened in
            // registers) made it worse and yielding (nanosleep) inside the loop did not help -- the cost follows the
            // number of instructions the epilogue warps issue while the main loop of the next tile runs.
            float2 low2[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) low2[i] = make_float2(0.f, 0.f);
            const float* Ac = fslice + cb * 64 + h * 32;
#pragma unroll 2
            for (int j = 0; j < p.masked_r; ++j) {
              const float d = dh_col[j * 32 + lane];
              const float2 d2 = make_float2(d, d);
              const float4* Ar = reinterpret_cast<const float4*>(Ac + j * BN);
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 a4 = Ar[q];  // warp-uniform address: a broadcast read
                ffma2(low2[q * 2 + 0], d2, make_float2(a4.x, a4.y));
                ffma2(low2[q * 2 + 1], d2, make_float2(a4.z, a4.w));
              }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              if ((kbits >> (2 * i)) & 1u) v[2 * i] = __float_as_uint(__uint_as_float(v[2 * i]) + low2[i].x);
              if ((kbits >> (2 * i + 1)) & 1u) v[2 * i + 1] = __float_as_uint(__uint_as_float(v[2 * i + 1]) + low2[i].y);
            }
          }
          if constexpr (kMasked) if (!use_fma) {
#pragma unroll
            for (int f = 0; f < 8; ++f) {
              const float4 t4 = *reinterpret_cast<const float4*>(scratch + lane * 32 + ((f ^ (lane & 7)) << 2));
              const float tv[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if ((kbits >> (f * 4 + e)) & 1u)
                  v[f * 4 + e] = __float_as_uint(fmaf(tv[e], p.inv_keep, __uint_as_float(v[f * 4 + e])));
            }
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
        if constexpr (kMasked && kFma) {
#pragma unroll
          for (int i = 0; i < 6; ++i) kb8[i] = kb8[i + 2];
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmD, my_stage + sbuf * EPI_BUF_BYTES, col0, rbase);
          bulk_commit_group();
        }
        sbuf ^= 1u;
      }
      // All TMEM reads of this accumulator are done: hand it back to the MMA warp.
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (kPair) mbar_arrive_cluster(&tmem_empty[buf], 0); else mbar_arrive(&tmem_empty[buf]);
      }
    }
    if (lane == 0) bulk_wait_group_read<0>();
    __syncwarp();
  }

  tc_fence_before();
  if constexpr (kPair) {
    cluster_sync_all();  // the peer's smem / TMEM / barriers stay alive until both CTAs are done
    if (warp == ALLOC_WARP) tmem_dealloc_2sm(tmem_base, TMEM_COLS);
  } else {
    __syncthreads();
    if (warp == ALLOC_WARP) tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

}  // namespace gemm
}  // namespace lb
