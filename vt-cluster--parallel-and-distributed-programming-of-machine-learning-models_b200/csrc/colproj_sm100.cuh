// Column projections of the backward on the tcgen05 tensor cores (every padded rank: 16 / 32 / 64):
//     dB = s dY^T H        dA = dH^T drop(X)          (torch autograd over peft's lora_B / lora_A linears)
// i.e. out[c, j] = sum_m keep(m, c) Act[m, c] Small[m, j] with Act = dY or X ([M, C], 16-bit), Small = H or dH
// ([M, r_pad], 16-bit).  One pass over dY and one over X: HBM-bound by design.  The reduction over the rows m is ONE
// accumulator in tensor memory per CTA:  D[128 columns c, r_pad] (fp32, r_pad TMEM columns) += Act_tile^T
// [128 c x 128 m] Small_tile [128 m x r_pad], eight UMMA 128 x r_pad x 16 per 128-row tile.  Both operands have
// the reduction dimension m as their SLOW dimension in memory, so both are "MN-major" UMMA operands straight out of
// the TMA boxes:
//     Act  : two boxes of 64 columns x 128 rows, 128-byte swizzle  (LBO = one box, SBO = 8 rows)
//     Small: one box of r_pad columns x 128 rows, swizzle = row bytes (32 / 64 / 128)
// (The warp-mma.sync kernel this replaced in round 2 was HBM-bound at padded rank 16 but ran at 12 % of the legacy
// HMMA rate at rank 64: 128 accumulator registers per thread, spilled, eight warps per SM)
// Warp roles (320 threads, one CTA per SM, 4-slot ring of 36 / 40 / 48 KiB):
//     warp 0   TMA producer
//     warp 1   TMEM allocation + the single MMA-issuing thread
//     warps 2-9  dropout (dA only): AND the keep-bits into the Act tile in shared memory before the MMA reads it
//               (half a row = 64 columns per thread; chunks whose eight elements all survived -- two thirds at
//               p = 0.05 -- are left alone)
//     warps 2-5  then the epilogue: TMEM -> registers -> the split's partial sums in global memory (or, with a
//               single row split, straight to the gradient)
// grid = (column tiles of 128, row splits, problems).  The last CTA of a column tile to finish (self-resetting
// atomicInc arrival counter) folds the row splits in split order: no reduction launch, bit-reproducible.
// Every tensor map covers this GPU's memory.
#pragma once

#include "skinny.cuh"

namespace lb {
namespace skinny {

constexpr int CP_BM = 128;                 // rows m per tile
constexpr int CP_BOX_BYTES = CP_BM * 128;  // 16 KiB: one TMA box of 128 rows x 64 columns
constexpr int CU_COLS = 128;               // columns per CTA = UMMA M

struct ColProjProblem {
  float* partial;          // [splits][C][kRPad] split partial sums (scratch)
  unsigned int* counters;  // [c_tiles], zero on entry; self-resetting (atomicInc wraps)
  const uint8_t* keep;     // [M, C/8] or null
  float* out;              // out[c*ld_c + j*ld_j], j < r
  long long ld_c, ld_j;
  float scale;
  int accumulate;
  int r;
  int M, C;
  int rows_per_split;  // multiple of CP_BM
  int splits, c_tiles;
};
// One launch carries up to two independent problems (dB and dA of one backward): blockIdx.z.
struct ColProjParams {
  ColProjProblem prob[2];
};

constexpr int CU_STAGES = 4;
constexpr int CU_THREADS = 320;
constexpr int CU_MASK_WARP0 = 2, CU_MASK_WARPS = 8, CU_EPI_WARPS = 4;
template <int kRPad>
__host__ __device__ constexpr int cu_slot_bytes() {
  return 2 * CP_BOX_BYTES + CP_BM * kRPad * 2;
}
template <int kRPad>
__host__ __device__ constexpr int cu_smem_bytes() {
  return CU_STAGES * cu_slot_bytes<kRPad>() + 256 + 1024;
}

__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, const uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// prmt.b32 in its default mode: selector nibble = byte index | 8 replicates that byte's SIGN over the result byte
// (the __byte_perm intrinsic documents three selector bits only, hence the PTX)
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}

template <int kRPad>
__global__ void __launch_bounds__(CU_THREADS, 1)
    colproj_umma_kernel(const __grid_constant__ CUtensorMap tmAct0, const __grid_constant__ CUtensorMap tmSmall0,
                        const __grid_constant__ CUtensorMap tmAct1, const __grid_constant__ CUtensorMap tmSmall1,
                        const ColProjParams pp, const int dtype) {
  static_assert(kRPad == 16 || kRPad == 32 || kRPad == 64, "padded ranks of the library");
  constexpr int TCOLS = kRPad < 32 ? 32 : kRPad;  // TMEM columns are allocated in powers of two >= 32
  const ColProjProblem& p = pp.prob[blockIdx.z];
  if (static_cast<int>(blockIdx.x) >= p.c_tiles || static_cast<int>(blockIdx.y) >= p.splits) return;  // CTA-uniform
  const CUtensorMap* tmAct = blockIdx.z == 0 ? &tmAct0 : &tmAct1;
  const CUtensorMap* tmSmall = blockIdx.z == 0 ? &tmSmall0 : &tmSmall1;
  constexpr int SLOT = cu_slot_bytes<kRPad>();
  constexpr int SROW = kRPad * 2;  // bytes per row of the Small tile = its swizzle span
  constexpr int SMALL_OFF = 2 * CP_BOX_BYTES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + CU_STAGES * SLOT);
  uint64_t* empty = full + CU_STAGES;
  uint64_t* masked = empty + CU_STAGES;
  uint64_t* acc_full = masked + CU_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * CU_COLS;
  const int m_begin = blockIdx.y * p.rows_per_split;
  const int m_end = min(p.M, m_begin + p.rows_per_split);
  const int nt = (m_end > m_begin) ? (m_end - m_begin + CP_BM - 1) / CP_BM : 0;
  const bool has_keep = p.keep != nullptr;

  if (threadIdx.x == 0) {
    for (int i = 0; i < CU_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
      mbar_init(&masked[i], CU_MASK_WARPS);
    }
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TCOLS);
    tmem_relinquish();
  }
  pdl_launch_dependents();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      for (int t = 0; t < nt; ++t) {
        const int st = t % CU_STAGES;
        mbar_wait(&empty[st], ((t / CU_STAGES) & 1) ^ 1, 31);
        uint8_t* act = smem + st * SLOT;
        mbar_arrive_expect_tx(&full[st], SLOT);
        const int row = m_begin + t * CP_BM;
        tma_load_2d(act, tmAct, &full[st], c0, row);  // columns / rows past the tensor arrive as zeros
        tma_load_2d(act + CP_BOX_BYTES, tmAct, &full[st], c0 + 64, row);
        tma_load_2d(act + SMALL_OFF, tmSmall, &full[st], 0, row);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_f16(dtype == kDtypeBf16 ? 1 : 0, CU_COLS, kRPad, 1, 1);
      const uint32_t ring_u = smem_u32(smem);
      const uint32_t small_layout = SROW == 128 ? 2u : (SROW == 64 ? 4u : 6u);
      for (int t = 0; t < nt; ++t) {
        const int st = t % CU_STAGES;
        mbar_wait(has_keep ? &masked[st] : &full[st], (t / CU_STAGES) & 1, 32);
        tc_fence_after();
        const uint32_t base = ring_u + st * SLOT;
        // MN-major operands: LBO = stride between 64-element groups of the M / N dimension, SBO = 8 rows of m
        const uint64_t da = make_smem_desc(base, CP_BOX_BYTES, 1024, 2);
        const uint64_t db = make_smem_desc(base + SMALL_OFF, 8 * SROW, 8 * SROW, small_layout);
#pragma unroll
        for (uint32_t k = 0; k < CP_BM / 16; ++k)  // 16 rows of m per instruction
          umma_f16(tmem_base, da + k * ((16u * 128u) >> 4), db + k * ((16u * SROW) >> 4), idesc, (t | k) != 0 ? 1u : 0u);
        umma_commit(&empty[st]);
      }
      umma_commit(acc_full);
    }
    __syncwarp();
  } else {
    // ===================== dropout mask, then epilogue =====================
    const int mt = (warp - CU_MASK_WARP0) * 32 + lane;
    const int mrow = mt & (CP_BM - 1), mbox = mt >> 7;  // row of the tile and 64-column box this thread masks
    if (has_keep) {
      const int keep_ld = p.C >> 3;
      const int kcol = (c0 >> 3) + mbox * 8;  // first of this thread's eight keep-bytes in a row
      const bool kfast = (((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep))) & 7) == 0) && kcol + 8 <= keep_ld;
      auto load_keep = [&](int t) -> uint2 {  // the 8 keep-bytes (64 columns) of this thread's half row of tile t
        uint2 k = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
        const int m = m_begin + t * CP_BM + mrow;
        if (t < nt && m < p.M) {
          const uint8_t* src = p.keep + static_cast<size_t>(m) * keep_ld + kcol;
          if (kfast) {
            k = __ldg(reinterpret_cast<const uint2*>(src));
          } else {
            uint32_t w[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
            for (int c = 0; c < 8; ++c)
              if (kcol + c < keep_ld)
                w[c >> 2] = (w[c >> 2] & ~(0xFFu << (8 * (c & 3)))) | (static_cast<uint32_t>(__ldg(src + c)) << (8 * (c & 3)));
            k = make_uint2(w[0], w[1]);
          }
        }
        return k;
      };
      // The keep-bytes come from HBM (the step streams far more than L2 between the forward that wrote them and
      // this kernel) and a tile lasts about one DRAM round trip: three tiles of them are kept in flight.
      uint2 kq0 = load_keep(0), kq1 = load_keep(1), kq2 = load_keep(2);
      for (int t = 0; t < nt; ++t) {
        const int st = t % CU_STAGES;
        const uint2 kcur = kq0;
        kq0 = kq1;
        kq1 = kq2;
        kq2 = load_keep(t + 3);
        mbar_wait(&full[st], (t / CU_STAGES) & 1, 33);
        const uint32_t row_u = smem_u32(smem + st * SLOT) + mbox * CP_BOX_BYTES + mrow * 128;
        const uint32_t kw[2] = {kcur.x, kcur.y};
        // Only chunks that lost an element are touched (a third of them at p = 0.05): shared-memory bandwidth is
        // what this kernel runs on -- TMA writes and tensor-core reads already move 96 KiB per tile.  All loads
        // first (one round trip, predicated, no branch), then the stores.
        uint4 v[8];
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {  // 16-byte chunk ch of the half row (XOR-swizzled by row)
          const uint32_t b = (kw[ch >> 2] >> (8 * (ch & 3))) & 0xFFu;
          if (b != 0xFFu) v[ch] = lds128(row_u + ((ch ^ (mrow & 7)) << 4));
        }
#pragma unroll
        for (int ch = 0; ch < 8; ++ch) {
          const uint32_t b = (kw[ch >> 2] >> (8 * (ch & 3))) & 0xFFu;
          if (b != 0xFFu) {
            // keep-bit k -> the sign bit of byte k (the products' bit positions j + 7m never collide), then PRMT
            // in sign-replication mode spreads each sign over the two bytes of its element
            const uint32_t zl = ((b & 0xFu) * 0x10204080u) & 0x80808080u, zh = ((b >> 4) * 0x10204080u) & 0x80808080u;
            sts128(row_u + ((ch ^ (mrow & 7)) << 4),
                   make_uint4(v[ch].x & prmt(zl, 0u, 0x9988u), v[ch].y & prmt(zl, 0u, 0xBBAAu),
                              v[ch].z & prmt(zh, 0u, 0x9988u), v[ch].w & prmt(zh, 0u, 0xBBAAu)));
          }
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(&masked[st]);
      }
    }
    // epilogue: a warp may touch TMEM lanes 32*(warp % 4) .. +31 = columns c of the tile
    if (warp < CU_MASK_WARP0 + CU_EPI_WARPS) {
      const int q = warp & 3;
      const int cl = q * 32 + lane;
      float* dst = p.partial + (static_cast<size_t>(blockIdx.y) * p.C + c0 + cl) * kRPad;
      float* o = p.out + static_cast<long long>(c0 + cl) * p.ld_c;
      const bool live = c0 + cl < p.C;
      if (nt > 0) {
        mbar_wait(acc_full, 0, 34);
        tc_fence_after();
      }
#pragma unroll
      for (int j0 = 0; j0 < kRPad; j0 += 32) {
        uint32_t v[32];
        if (nt > 0) {
          tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + j0, v);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0u;
        }
        if (!live) continue;
        if (p.splits == 1) {  // nothing to fold: scale and write the gradient (0 + v: what the fold computes)
#pragma unroll
          for (int i = 0; i < (kRPad < 32 ? kRPad : 32); ++i)
            if (j0 + i < p.r) {
              float* e = o + static_cast<long long>(j0 + i) * p.ld_j;
              const float val = p.scale * (0.f + __uint_as_float(v[i]));
              *e = p.accumulate ? (*e + val) : val;
            }
        } else {
#pragma unroll
          for (int i = 0; i < (kRPad < 32 ? kRPad : 32); i += 4)
            *reinterpret_cast<uint4*>(dst + j0 + i) = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
      }
    }
  }

  // The last CTA of this column tile to finish folds the splits, always in split order 0,1,2,... (the arrival order
  // only decides WHO adds), applies the scale and writes the gradient.
  __shared__ int s_is_last;
  tc_fence_before();
  __threadfence();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TCOLS);
  if (p.splits == 1) return;  // CTA-uniform
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicInc(p.counters + blockIdx.x, static_cast<unsigned int>(p.splits - 1));
    s_is_last = (prev == static_cast<unsigned int>(p.splits - 1));
  }
  __syncthreads();
  if (s_is_last) {
    __threadfence();
    const int ncols = min(CU_COLS, p.C - c0);
    const size_t sstride = static_cast<size_t>(p.C) * kRPad;
    // four consecutive ranks per thread and trip (16-byte loads, up to four splits in flight): a scalar loop
    // over 128 x r outputs is a chain of L2 round trips as long as the streaming itself
    constexpr int G = kRPad / 4;
    for (int i = threadIdx.x; i < ncols * G; i += CU_THREADS) {
      const int cl = i / G, j = (i - cl * G) * 4;
      if (j >= p.r) continue;
      const float4* src = reinterpret_cast<const float4*>(p.partial + static_cast<size_t>(c0 + cl) * kRPad + j);
      const size_t sv = sstride / 4;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      auto add = [&](const float4 v) {
        acc.x += v.x;
        acc.y += v.y;
        acc.z += v.z;
        acc.w += v.w;
      };
      int sp = 0;
      for (; sp + 4 <= p.splits; sp += 4) {
        const float4 v0 = __ldcg(src + (sp + 0) * sv), v1 = __ldcg(src + (sp + 1) * sv);
        const float4 v2 = __ldcg(src + (sp + 2) * sv), v3 = __ldcg(src + (sp + 3) * sv);
        add(v0);
        add(v1);
        add(v2);
        add(v3);
      }
      for (; sp < p.splits; ++sp) add(__ldcg(src + sp * sv));
      const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (j + e < p.r) {
          float* o = p.out + static_cast<long long>(c0 + cl) * p.ld_c + static_cast<long long>(j + e) * p.ld_j;
          *o = p.accumulate ? (*o + p.scale * a4[e]) : (p.scale * a4[e]);
        }
    }
  }
}

}  // namespace skinny
}  // namespace lb
