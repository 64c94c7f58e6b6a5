// HBM-bound companions of the tcgen05 GEMM: everything whose arithmetic
// intensity is ~r flop/byte: the rank-r projections
//     H  = drop(X) A^T          dH = dY (sB)            (rowproj)
//     dB = s dY^T H             dA = dH^T drop(X)       (colproj + reduce)
// plus the fp32 -> 16-bit adapter packing.  They stream the activation once
// through a TMA + mbarrier ring and use warp-level mma.sync on the tiles:
// the tensor pipe is nowhere near the limit here, HBM is.
//
// Follows peft lora.Linear semantics as used at
// finetune_lora_distilgpt2.py:98-104 / finetune_lora_opt_pp.py:104-112:
// dropout acts on the adapter input only, scaling = lora_alpha / r.
#pragma once

#include <type_traits>

#include "common.cuh"

namespace lb {
namespace skinny {

// ---------------------------------------------------------------------------
// mma.sync / ldmatrix wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2_t(uint32_t addr, uint32_t (&r)[2]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];"
               : "=r"(r[0]), "=r"(r[1])
               : "r"(addr));
}
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
  }
}
// ---------------------------------------------------------------------------
// pack: fp32 masters A[r,K], B[N,r] -> 16-bit operands in the four layouts the
// kernels read (zero padded to r_pad; the scaling s is folded into B):
//   a16 [r_pad,K]   a16t [K,r_pad]   sb16 [N,r_pad]   sb16t [r_pad,N]
// Runs once per forward; the backward reads the forward's block (autograd saved-tensor semantics).
// ---------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T cvt16(float v);
template <>
__device__ __forceinline__ __nv_bfloat16 cvt16<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <>
__device__ __forceinline__ __half cvt16<__half>(float v) {
  return __float2half_rn(v);
}

template <typename T>
__device__ __forceinline__ void pack_adapters_body(const float* __restrict__ A, const float* __restrict__ B, T* __restrict__ a16,
                                                   T* __restrict__ a16t, T* __restrict__ sb16, T* __restrict__ sb16t, int K,
                                                   int N, int r, int r_pad, float s, long long first, long long stride) {
  const long long na = static_cast<long long>(r_pad) * K;
  const long long nb = static_cast<long long>(N) * r_pad;
  for (long long i = first; i < na + nb; i += stride) {
    if (i < na) {
      const int j = static_cast<int>(i / K), k = static_cast<int>(i % K);
      const T v = cvt16<T>(j < r ? A[static_cast<long long>(j) * K + k] : 0.f);
      a16[i] = v;
      a16t[static_cast<long long>(k) * r_pad + j] = v;
    } else {
      const long long t = i - na;
      const int n = static_cast<int>(t / r_pad), j = static_cast<int>(t % r_pad);
      const T v = cvt16<T>(j < r ? s * B[static_cast<long long>(n) * r + j] : 0.f);
      sb16[t] = v;
      sb16t[static_cast<long long>(j) * N + n] = v;
    }
  }
}

template <typename T>
__global__ void pack_adapters_kernel(const float* __restrict__ A, const float* __restrict__ B, T* __restrict__ a16,
                                     T* __restrict__ a16t, T* __restrict__ sb16, T* __restrict__ sb16t, int K, int N,
                                     int r, int r_pad, float s) {
  pdl_launch_dependents();
  pdl_wait();  // the block may still be read by kernels of an earlier call on this stream
  pack_adapters_body<T>(A, B, a16, a16t, sb16, sb16t, K, N, r, r_pad, s,
                        blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x,
                        static_cast<long long>(gridDim.x) * blockDim.x);
}

// ---------------------------------------------------------------------------
// rowproj: Out[M, kRPad] = out_scale * (keep .* Act)[M, C] * P[kRPad, C]^T
// One CTA per 16*SLABS rows.  The last warp is the TMA producer; consumer warps w and w+SLABS share
// the 16-row slab w%SLABS and split every 64-column tile between them (two k16 steps each); their two
// partial accumulators are summed through shared memory at the end.  With dropout each lane draws ONE
// Philox4x32-10 call per k16 step whose 128 bits are exactly the keep decisions of the eight elements
// the lane holds in its A fragment (common.cuh, "layout of the dropout stream"): the fragment is
// masked in registers, nothing is written back to shared memory and no barrier separates the draw
// from the MMAs.  The draw does not depend on the data, so it is issued BEFORE the wait for the tile.
// The keep-bits leave as bytes keep[M, C/8]: a quad (four lanes that share two rows) ORs its 2-bit
// pieces together with two shuffles and stores one 32-bit word per row and tile.
// ---------------------------------------------------------------------------
// 16-row slabs per CTA.  Measured on B200 (M = 16384, C = 2560, r_pad 16, round 1): 4 slabs (64 rows,
// 256 CTAs) beat 1 slab (1024 small CTAs: every CTA re-reads the P tiles from L2 and 2 KiB TMA boxes
// are inefficient) and 7 slabs (147 CTAs, one per SM: too few warps per SM).
__host__ __device__ constexpr int rp_slabs(int rpad) { return 4; }
__host__ __device__ constexpr int rp_rows(int rpad) { return 16 * rp_slabs(rpad); }
__host__ __device__ constexpr int rp_threads(int rpad) { return (2 * rp_slabs(rpad) + 1) * 32; }
#ifndef LB_EXP_RP_MIN_CTAS
#define LB_EXP_RP_MIN_CTAS 2  // measured: capping registers for a third CTA (72, spills) 64 us, uncapped 39 us
#endif
#ifndef LB_EXP_RP_STAGES
#define LB_EXP_RP_STAGES 6
#endif
__host__ __device__ constexpr int rp_min_ctas(int rpad) { return rpad <= 16 ? LB_EXP_RP_MIN_CTAS : (rpad <= 32 ? 2 : 1); }
constexpr int RP_BK = 64;
constexpr int RP_STAGES = LB_EXP_RP_STAGES;
template <int kRPad>
__host__ __device__ constexpr int rp_act_bytes() {
  return ((rp_rows(kRPad) * 128 + 1023) / 1024) * 1024;
}
template <int kRPad>
__host__ __device__ constexpr int rp_slot_bytes() {
  return rp_act_bytes<kRPad>() + ((kRPad * 128 + 1023) / 1024) * 1024;
}
template <int kRPad>
__host__ __device__ constexpr int rp_smem_bytes() {
  return RP_STAGES * rp_slot_bytes<kRPad>() + rp_rows(kRPad) * kRPad * 4 + 1024 + 256;
}

struct RowProjParams {
  void* out;  // [M, kRPad] 16-bit
  uint8_t* keep_out;  // [M, C/8] (dropout only)
  int M, C;
  float out_scale;
  int dropout;  // 0/1
  uint64_t seed, offset;
  PhiloxKeys rk;  // philox_round_keys(seed), filled by the host
  const unsigned long long* offset_dev;  // optional: *offset_dev is added to `offset` (graph-replayable RNG state)
  uint32_t thresh14;  // round(p * 16384)
  unsigned int* zero_words;  // optional: words CTA 0 clears (the column projections' arrival counters)
  int n_zero_words;
  // kDirectA (forward, padded rank 16): the projection matrix is read straight from the fp32 master A [r, C]
  // and the kernel also does the pack (A, B -> the four 16-bit layouts) as a side job for the kernels that follow,
  // so the forward needs no pack launch.
  const float* A32;
  const float* B32;  // [N, r]
  int r, N;
  float s;           // scaling folded into the packed B
  void *a16, *a16t, *sb16, *sb16t;
};

template <bool kBf16, int kRPad, bool kDirectA = false, int kNBE = 2>
__global__ void __launch_bounds__(rp_threads(kRPad), rp_min_ctas(kRPad))
    rowproj_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmP,
                   const RowProjParams p) {
  static_assert(!kDirectA || kRPad == 16, "the direct-A path holds its B fragments in registers: padded rank 16 only");
  constexpr int SLOT = rp_slot_bytes<kRPad>();
  constexpr int NB = kRPad / 8;
  constexpr int SLABS = rp_slabs(kRPad);
  constexpr int RP_BM = rp_rows(kRPad);
  constexpr int RP_CONSUMERS = 2 * SLABS;
  constexpr int RP_ACT_BYTES = rp_act_bytes<kRPad>();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* sred = reinterpret_cast<float*>(smem + RP_STAGES * SLOT);  // [64 rows][kRPad] pair reduction
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + RP_STAGES * SLOT + RP_BM * kRPad * 4);
  uint64_t* empty = full + RP_STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * RP_BM;
  const int nk = (p.C + RP_BK - 1) / RP_BK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < RP_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], RP_CONSUMERS);
    }
    fence_barrier_init();
  }
  pdl_launch_dependents();
  __syncthreads();
  pdl_wait();
  // Arrival counters of the column projections that follow in the same call: they reset themselves,
  // this only covers a workspace that has never been used (or a call that was cut short).
  if (blockIdx.x == 0 && p.zero_words != nullptr)
    for (int i = threadIdx.x; i < p.n_zero_words; i += blockDim.x) p.zero_words[i] = 0u;
  if constexpr (kDirectA) {
    // the pack, spread over every thread of the grid (a few elements each); nobody in THIS kernel reads it
    using T16 = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    pack_adapters_body<T16>(p.A32, p.B32, static_cast<T16*>(p.a16), static_cast<T16*>(p.a16t), static_cast<T16*>(p.sb16),
                            static_cast<T16*>(p.sb16t), p.C, p.N, p.r, kRPad, p.s,
                            blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x,
                            static_cast<long long>(gridDim.x) * blockDim.x);
  }

  if (warp == RP_CONSUMERS) {
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int kb = 0; kb < nk; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1, 11);
        uint8_t* act = smem + stage * SLOT;
        mbar_arrive_expect_tx(&full[stage], RP_BM * 128 + (kDirectA ? 0 : kRPad * 128));
        tma_load_2d(act, &tmAct, &full[stage], kb * RP_BK, m0);
        if constexpr (!kDirectA) tma_load_2d(act + RP_ACT_BYTES, &tmP, &full[stage], kb * RP_BK, 0);
        if (++stage == RP_STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
    __syncwarp();
    return;
  }

  const int slab = warp % SLABS;  // 16-row slab of the tile
  const int half = warp / SLABS;  // which two k16 steps of every tile this warp takes
  const int g8 = lane >> 2, q = lane & 3;
  float acc[NB][4];
#pragma unroll
  for (int n = 0; n < NB; ++n)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[n][e] = 0.f;

  // Per-lane constants of the dropout draw.
  const int keep_ld = p.C >> 3;
  const PhiloxKeys& rk = p.rk;
  const uint32_t thresh2 = dropout_thresh_half2(p.thresh14);
  const uint64_t offset = p.offset + ((p.dropout && p.offset_dev != nullptr) ? *p.offset_dev : 0ull);
  const uint32_t off_lo = static_cast<uint32_t>(offset), off_hi = static_cast<uint32_t>(offset >> 32);
  const uint32_t blocks_per_row = static_cast<uint32_t>((p.C + 15) >> 4);
  uint64_t gbase = 0;
  if (p.dropout) {
    gbase = dropout_counter(static_cast<uint32_t>((m0 >> 4) + slab), static_cast<uint32_t>(half * 2), blocks_per_row,
                            static_cast<uint32_t>(lane));
  }
  // keep-bytes: lane q == 0 stores the four bytes of row g8 for this warp's half tile, q == 1 those of row g8 + 8
  const int krow = m0 + slab * 16 + g8 + (q & 1) * 8;
  uint8_t* kdst = p.dropout ? p.keep_out + static_cast<size_t>(krow) * keep_ld + half * 4 : nullptr;
  const bool kstore = p.dropout && q < 2 && krow < p.M;
  const bool kfast = ((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep_out))) & 3) == 0;

  // ldmatrix addresses inside a slot (128-byte rows, 16-byte chunks XOR-swizzled by row & 7)
  const int a_row = slab * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
  const uint32_t a_off = static_cast<uint32_t>(a_row * 128);
  const int a_sw = a_row & 7, a_hi = lane >> 4;
  const int b_row = lane & 7, b_hi = (lane >> 3) & 1;

  uint32_t stage = 0, phase = 0;
  // Two tiles per trip: their four Philox draws are independent dependency chains (10 rounds of a wide multiply
  // feeding a 3-input XOR each), issued back to back so that they hide one another's latency; with ~4 warps per
  // scheduler a single chain per warp leaves the issue slots idle (ncu: "wait" and "math pipe throttle" stalls).
  constexpr int NBE_MAX = kDirectA ? kNBE : NB;  // kNBE = 1: r <= 8 (the reference), half the fragment registers
  const int nbe = kDirectA ? min((p.r + 7) >> 3, kNBE) : NB;  // n-blocks (8 rank columns each) that are not all padding
  // kDirectA: B fragments straight from the fp32 master, fetched one trip ahead (the loads of trip t+1 are issued
  // after trip t's MMAs and packed to 16 bits at the top of trip t+1, so their latency hides under the Philox draws
  // and only eight registers stay live across them): lane (g8, q) holds A[n*8 + g8][c .. c+1] and [c+8 .. c+9],
  // c = 16-column step + 2q.
  float2 braw[2][2][NBE_MAX][2];
  auto load_braw = [&](int kb_first) {
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int kk = 0; kk < 2; ++kk)
#pragma unroll
        for (int n = 0; n < NBE_MAX; ++n)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int j = n * 8 + g8, c = (kb_first + u) * RP_BK + (half * 2 + kk) * 16 + h * 8 + 2 * q;
            braw[u][kk][n][h] = (n < nbe && j < p.r && c < p.C)
                                    ? __ldg(reinterpret_cast<const float2*>(p.A32 + static_cast<size_t>(j) * p.C + c))
                                    : make_float2(0.f, 0.f);
          }
  };
  if constexpr (kDirectA) load_braw(0);
  for (int kb0 = 0; kb0 < nk; kb0 += 2) {
    uint32_t bpk[2][2][NBE_MAX][2];
    if constexpr (kDirectA) {
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int kk = 0; kk < 2; ++kk)
#pragma unroll
          for (int n = 0; n < NBE_MAX; ++n)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              bpk[u][kk][n][h] = pack2(braw[u][kk][n][h].x, braw[u][kk][n][h].y, kBf16 ? kDtypeBf16 : kDtypeFp16);
    }
    uint32_t kmask[2][2][4];
    if (p.dropout) {
      uint32_t kw[2][2];
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint64_t g = gbase + static_cast<uint64_t>((kb0 + u) * 4 + kk) * 32ull;
          const uint32_t b = dropout_fragment_masks(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), off_lo,
                                                    off_hi, rk, thresh2, kmask[u][kk]);
          // 2-bit field of register w -> byte w of a word, at bit 2q; the quad's four pieces OR into whole bytes:
          // byte 0 = (row g8, chunk 0), byte 1 = (row g8+8, chunk 0), byte 2 = (g8, chunk 1), byte 3 = (g8+8, chunk 1)
          uint32_t t = (b * 0x1001u) & 0x000F000Fu;
          t = ((t * 0x41u) & 0x03030303u) << (2 * q);
          t |= __shfl_xor_sync(0xFFFFFFFFu, t, 1);
          t |= __shfl_xor_sync(0xFFFFFFFFu, t, 2);
          kw[u][kk] = t;
        }
      if (kstore) {
        // this lane's row: chunk bytes (kk 0: c0, c1; kk 1: c0, c1) of the warp's half tile
        const uint32_t sel = (q & 1) ? 0x7531u : 0x6420u;
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int kb = kb0 + u;
          const uint32_t word = __byte_perm(kw[u][0], kw[u][1], sel);
          const int c8 = kb * 8 + half * 4;  // first chunk column of the word
          uint8_t* dst = kdst + kb * 8;
          if (kfast && c8 + 4 <= keep_ld) {
            *reinterpret_cast<uint32_t*>(dst) = word;
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (c8 + i < keep_ld) dst[i] = static_cast<uint8_t>(word >> (8 * i));
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (kb0 + u >= nk) break;
      mbar_wait(&full[stage], phase, 12);
      const uint32_t act_u = smem_u32(smem + stage * SLOT), pt_u = act_u + RP_ACT_BYTES;
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        const int ks = half * 2 + kk;
        uint32_t a[4];
        ldsm_x4(act_u + a_off + (((ks * 2 + a_hi) ^ a_sw) << 4), a);
        if (p.dropout) {
          a[0] &= kmask[u][kk][0]; a[1] &= kmask[u][kk][1]; a[2] &= kmask[u][kk][2]; a[3] &= kmask[u][kk][3];
        }
#pragma unroll
        for (int n = 0; n < NB; ++n) {
          uint32_t b[2];
          if constexpr (kDirectA) {
            if (n >= nbe || n >= NBE_MAX) continue;  // warp-uniform: all-padding rank columns contribute nothing
            b[0] = bpk[u][kk][n < NBE_MAX ? n : 0][0];
            b[1] = bpk[u][kk][n < NBE_MAX ? n : 0][1];
          } else {
            const int prow = n * 8 + b_row;
            ldsm_x2(pt_u + prow * 128 + (((ks * 2 + b_hi) ^ (prow & 7)) << 4), b);
          }
          mma_16816<kBf16>(acc[n], a, b);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[stage]);
      if (++stage == RP_STAGES) {
        stage = 0;
        phase ^= 1;
      }
    }
    if constexpr (kDirectA) load_braw(kb0 + 2);  // columns beyond C come back as zeros
  }

  // acc fragment: (row = lane/4 [+8], col = n*8 + (lane%4)*2 [+1]); half 1 hands its partial
  // sums to half 0 through smem (fixed order -> bit-reproducible).
  const int rloc = slab * 16 + g8;
  if (half == 1) {
#pragma unroll
    for (int n = 0; n < NB; ++n) {
      const int col = n * 8 + q * 2;
      *reinterpret_cast<float2*>(sred + rloc * kRPad + col) = make_float2(acc[n][0], acc[n][1]);
      *reinterpret_cast<float2*>(sred + (rloc + 8) * kRPad + col) = make_float2(acc[n][2], acc[n][3]);
    }
  }
  asm volatile("bar.sync %0, 64;" ::"r"(1 + slab) : "memory");
  if (half == 0) {
    uint32_t* out = reinterpret_cast<uint32_t*>(p.out);
    const int r0 = m0 + rloc;
#pragma unroll
    for (int n = 0; n < NB; ++n) {
      const int col = n * 8 + q * 2;
      const float2 o0 = *reinterpret_cast<const float2*>(sred + rloc * kRPad + col);
      const float2 o1 = *reinterpret_cast<const float2*>(sred + (rloc + 8) * kRPad + col);
      if (r0 < p.M)
        out[(static_cast<size_t>(r0) * kRPad + col) >> 1] =
            pack2((acc[n][0] + o0.x) * p.out_scale, (acc[n][1] + o0.y) * p.out_scale, kBf16 ? kDtypeBf16 : kDtypeFp16);
      if (r0 + 8 < p.M)
        out[(static_cast<size_t>(r0 + 8) * kRPad + col) >> 1] =
            pack2((acc[n][2] + o1.x) * p.out_scale, (acc[n][3] + o1.y) * p.out_scale, kBf16 ? kDtypeBf16 : kDtypeFp16);
    }
  }
}

// ---------------------------------------------------------------------------
// colproj: partial[split][C, kRPad] = sum over the split's rows m of
//          (keep .* Act)[m, c] * Small[m, j]        (fp32)
// grid = (ceil(C / (64*kCB)), splits).  A tile is 128 rows x kCB boxes of 64 columns (each box
// one 128-byte-swizzled TMA load): kCB = 2 makes every row a 256-byte DRAM segment.  Warp 8 is
// the TMA producer; consumer warp w owns rows 16w..16w+15 of every tile and all its columns;
// partial sums of the 8 warps meet in shared memory; the last CTA of a column tile folds the row splits.
// Padded rank 32 / 64 ("wide"): 64 / 128 accumulator registers per thread.  ptxas sizes the register file for
// the thread count rounded up to four warps, so nine warps leave 168 registers and the rank-64 accumulators
// spilled (2.3 M local loads per launch, measured); the wide form therefore runs EIGHT warps -- lane 0 of warp 0
// also issues the TMA loads, one tile behind its own consumption -- and has 255 registers.  Its eight partial
// accumulators meet in the tile ring (idle by then) and all threads add them at once, in warp order, instead of
// eight serial turns of read-modify-write (which cost as much as the main loop at 19 tiles per CTA).
// ---------------------------------------------------------------------------
constexpr int CP_BM = 128;
constexpr int CP_BOX_BYTES = CP_BM * 128;  // 16 KiB: 128 rows x 64 columns
constexpr int CP_CONSUMERS = 8;
constexpr int CP_THREADS = (CP_CONSUMERS + 1) * 32;
__host__ __device__ constexpr bool cp_wide(int rpad) { return rpad >= 32; }
__host__ __device__ constexpr int cp_threads(int rpad) { return cp_wide(rpad) ? CP_CONSUMERS * 32 : CP_THREADS; }
// 64-column boxes per tile / ring depth, by padded rank (register and smem budget)
__host__ __device__ constexpr int cp_boxes(int rpad) { return 1; }  // 2 boxes cost the second resident CTA (smem)
__host__ __device__ constexpr int cp_cols(int rpad) { return 64 * cp_boxes(rpad); }
#ifndef LB_EXP_CP_STAGES
#define LB_EXP_CP_STAGES 4
#endif
__host__ __device__ constexpr int cp_stages(int rpad) { return LB_EXP_CP_STAGES; }
template <int kRPad>
__host__ __device__ constexpr int cp_slot_bytes() {
  return cp_boxes(kRPad) * CP_BOX_BYTES + ((CP_BM * kRPad * 2 + 1023) / 1024) * 1024;
}
constexpr int CP_KEEP_STAGES = 4;  // keep-byte ring of the dA problem: 4 tiles x 128 rows x 8 bytes
constexpr int CP_KEEP_BYTES = CP_KEEP_STAGES * CP_BM * 8;
template <int kRPad>
__host__ __device__ constexpr int cp_smem_bytes() {
  return cp_stages(kRPad) * cp_slot_bytes<kRPad>() + (cp_wide(kRPad) ? 0 : cp_cols(kRPad) * kRPad * 4) + CP_KEEP_BYTES +
         1024 + 256;
}
__host__ __device__ constexpr int cp_min_ctas(int rpad) { return rpad <= 16 ? 2 : 1; }

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

template <bool kBf16, int kRPad>
__global__ void __launch_bounds__(cp_threads(kRPad), cp_min_ctas(kRPad))
    colproj_kernel(const __grid_constant__ CUtensorMap tmAct0, const __grid_constant__ CUtensorMap tmSmall0,
                   const __grid_constant__ CUtensorMap tmAct1, const __grid_constant__ CUtensorMap tmSmall1,
                   const ColProjParams pp) {
  const ColProjProblem& p = pp.prob[blockIdx.z];
  if (static_cast<int>(blockIdx.x) >= p.c_tiles || static_cast<int>(blockIdx.y) >= p.splits) return;  // CTA-uniform
  const CUtensorMap* tmAct = blockIdx.z == 0 ? &tmAct0 : &tmAct1;
  const CUtensorMap* tmSmall = blockIdx.z == 0 ? &tmSmall0 : &tmSmall1;
  constexpr int SLOT = cp_slot_bytes<kRPad>();
  constexpr int NB = kRPad / 8;
  constexpr int SROW = kRPad * 2;  // bytes per row of the Small tile
  constexpr int kCB = cp_boxes(kRPad);
  constexpr int COLS = 64 * kCB;
  constexpr int STAGES = cp_stages(kRPad);
  constexpr int ACT_BYTES = kCB * CP_BOX_BYTES;
  constexpr int NCB = 4 * kCB;  // 16-column blocks per tile
  constexpr bool kWide = cp_wide(kRPad);
  constexpr int THREADS = cp_threads(kRPad);
  constexpr int SOUT_BYTES = kWide ? 0 : COLS * kRPad * 4;
  static_assert(!kWide || STAGES * SLOT >= CP_CONSUMERS * COLS * kRPad * 4, "the tile ring holds the eight partial accumulators");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* sout = reinterpret_cast<float*>(smem + STAGES * SLOT);
  uint8_t* keep_s = smem + STAGES * SLOT + SOUT_BYTES;  // [CP_KEEP_STAGES][128 rows][8 bytes]
  uint64_t* full = reinterpret_cast<uint64_t*>(keep_s + CP_KEEP_BYTES);
  uint64_t* empty = full + STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.x * COLS;
  const int m_begin = blockIdx.y * p.rows_per_split;
  const int m_end = min(p.M, m_begin + p.rows_per_split);
  const int nt = (m_end > m_begin) ? (m_end - m_begin + CP_BM - 1) / CP_BM : 0;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], CP_CONSUMERS);
    }
    fence_barrier_init();
  }
  if constexpr (!kWide)
    for (int i = threadIdx.x; i < COLS * kRPad; i += THREADS) sout[i] = 0.f;
  pdl_launch_dependents();
  __syncthreads();
  pdl_wait();

  // tile t -> ring slot t % STAGES, once every consumer warp has released the slot's previous tile
  auto produce = [&](int t) {
    const int st = t % STAGES;
    mbar_wait(&empty[st], ((t / STAGES) & 1) ^ 1, 21);
    uint8_t* act = smem + st * SLOT;
    mbar_arrive_expect_tx(&full[st], ACT_BYTES + CP_BM * SROW);
#pragma unroll
    for (int bx = 0; bx < kCB; ++bx)
      tma_load_2d(act + bx * CP_BOX_BYTES, tmAct, &full[st], c0 + bx * 64, m_begin + t * CP_BM);
    tma_load_2d(act + ACT_BYTES, tmSmall, &full[st], 0, m_begin + t * CP_BM);
  };

  if (!kWide && warp == CP_CONSUMERS) {
    if (lane == 0)
      for (int t = 0; t < nt; ++t) produce(t);
    __syncwarp();
  } else {
    if (kWide && warp == 0 && lane == 0)
      for (int t = 0; t < min(nt, STAGES); ++t) produce(t);
    float acc[NCB][NB][4];
#pragma unroll
    for (int cbk = 0; cbk < NCB; ++cbk)
#pragma unroll
      for (int n = 0; n < NB; ++n)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[cbk][n][e] = 0.f;

    const int keep_ld = p.C >> 3;
    // Dropout mask of the activation (dA = dH^T drop(X)): applied to the transposed ldmatrix fragments in
    // registers.  Register a[j] of a 16x16 block holds the two elements (rows 2q, 2q+1 [+8], one column g) of
    // one 16-byte chunk; their keep-bits are bit g of that chunk's keep-byte in rows 2q and 2q+1.  A warp
    // needs the 8 keep-bytes (64 columns) of each of its 16 rows per tile: lanes 0..15 copy them, 8 bytes
    // each, into a small ring in shared memory with cp.async three tiles ahead (the loads come from HBM -- the
    // step streams far more than L2 between the forward that wrote them and this kernel -- and no register is
    // held while they are in flight); the ring slice of a warp is private to it, so a warp-level sync suffices.
    static_assert(kCB == 1, "the register masking below assumes 64-column tiles");
    constexpr int PF = CP_KEEP_STAGES - 1;
    const int g8 = lane >> 2, q4 = lane & 3;
    const bool has_keep = p.keep != nullptr;
    const bool kfast = has_keep && (((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep))) & 7) == 0) &&
                       c0 + 64 <= p.C;
    const uint32_t keep_u = smem_u32(keep_s) + warp * 16 * 8;
    auto issue_keep = [&](int t) {  // keep-bytes of tile t -> ring slot t % CP_KEEP_STAGES (lanes 0..15: one row each)
      if (t < nt && lane < 16) {
        const int m = m_begin + t * CP_BM + warp * 16 + lane;
        const uint32_t dst = keep_u + (t % CP_KEEP_STAGES) * (CP_BM * 8) + lane * 8;
        const uint8_t* src = p.keep + static_cast<size_t>(m) * keep_ld + (c0 >> 3);
        if (m < p.M && kfast) {
          cp_async8(dst, src);
        } else {
          uint32_t w[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
          if (m < p.M) {
#pragma unroll
            for (int c = 0; c < 8; ++c)
              if ((c0 >> 3) + c < keep_ld)
                w[c >> 2] = (w[c >> 2] & ~(0xFFu << (8 * (c & 3)))) | (static_cast<uint32_t>(__ldg(src + c)) << (8 * (c & 3)));
          }
          sts64(dst, make_uint2(w[0], w[1]));
        }
      }
      cp_async_commit();  // one group per tile, also when this lane copied nothing: the counting stays uniform
    };
    if (has_keep)
      for (int t = 0; t < PF; ++t) issue_keep(t);
    uint32_t stage = 0, phase = 0;
    for (int t = 0; t < nt; ++t) {
      const int mt = m_begin + t * CP_BM;
      // wide form: the slot of tile t - 1 is refilled here, when the other warps have (nearly) left it too
      if (kWide && warp == 0 && lane == 0 && t >= 1 && t - 1 + STAGES < nt) produce(t - 1 + STAGES);
      // per row and 32-bit half: byte c = 0xFF iff bit g8 of chunk c's keep-byte is set
      uint32_t ksel[4][2];
      if (has_keep) {
        issue_keep(t + PF);
        cp_async_wait<PF>();  // this lane's copy for tile t has landed
        __syncwarp();         // ... and so have the other lanes' (the slice is this warp's own)
        const uint32_t src = keep_u + (t % CP_KEEP_STAGES) * (CP_BM * 8);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint2 kr = lds64(src + ((j >> 1) * 8 + 2 * q4 + (j & 1)) * 8);
          ksel[j][0] = ((kr.x >> g8) & 0x01010101u) * 0xFFu;
          ksel[j][1] = ((kr.y >> g8) & 0x01010101u) * 0xFFu;
        }
      }
      mbar_wait(&full[stage], phase, 22);
      uint8_t* act = smem + stage * SLOT;
      // Rows of the last tile of a split that belong to the next split (or lie
      // beyond M, where TMA filled zeros) must not contribute.
      const int valid_rows = m_end - mt;  // > 0
      const uint32_t act_u = smem_u32(act), sm_u = act_u + ACT_BYTES;
      const bool warp_live = warp * 16 < valid_rows;
      if (warp_live) {
        uint32_t b[NB][2];
#pragma unroll
        for (int n = 0; n < NB; ++n) {
          const int mrow = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
          // rows of 64 / 128 bytes are stored TMA-swizzled (16-byte chunk index XOR row bits): unswizzled, the eight
          // row addresses of an ldmatrix fall into the same banks (8-way conflict at r_pad = 64)
          const int chunk = kRPad == 64 ? (n ^ (mrow & 7)) : (kRPad == 32 ? (n ^ ((mrow >> 1) & 3)) : n);
          ldsm_x2_t(sm_u + mrow * SROW + chunk * 16, b[n]);
        }
        if (warp * 16 + 16 > valid_rows) {
          // Partially valid 16-row slab: zero B-fragment entries of dead rows.
          // b[n][0] holds k = (lane%4)*2 + {0,1}, b[n][1] holds k + 8.
#pragma unroll
          for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int k = warp * 16 + h * 8 + (lane & 3) * 2;
              uint32_t msk = 0;
              if (k < valid_rows) msk |= 0x0000FFFFu;
              if (k + 1 < valid_rows) msk |= 0xFFFF0000u;
              b[n][h] &= msk;
            }
        }
#pragma unroll
        for (int cbk = 0; cbk < NCB; ++cbk) {
          uint32_t a[4];
          const int q = lane >> 3, i = lane & 7;
          const int mrow = warp * 16 + (q >> 1) * 8 + i;
          const int ch = (cbk & 3) * 2 + (q & 1);
          ldsm_x4_t(act_u + (cbk >> 2) * CP_BOX_BYTES + mrow * 128 + ((ch ^ (mrow & 7)) << 4), a);
          if (has_keep) {
            // a[0]: rows (2q, 2q+1), chunk 2*cbk; a[1]: same rows, chunk 2*cbk + 1; a[2], a[3]: rows + 8
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int chunk = 2 * cbk + (j & 1), rows = (j >> 1) * 2, c = chunk & 3;
              const uint32_t sel = static_cast<uint32_t>(((4 + c) << 12) | ((4 + c) << 8) | (c << 4) | c);
              a[j] &= __byte_perm(ksel[rows][chunk >> 2], ksel[rows + 1][chunk >> 2], sel);
            }
          }
#pragma unroll
          for (int n = 0; n < NB; ++n) mma_16816<kBf16>(acc[cbk][n], a, b[n]);
        }
      }
      __syncwarp();  // also: every lane has read its keep rows before the ring slot is refilled (next trip's issue)
      if (lane == 0) mbar_arrive(&empty[stage]);
      if (++stage == STAGES) {
        stage = 0;
        phase ^= 1;
      }
    }
    // acc fragment: (row c = cbk*16 + lane/4 [+8], col j = n*8 + (lane%4)*2 [+1]).
    // The eight warps add their partial sums in warp order (named barrier 1 between turns):
    // a fixed summation order keeps dA/dB bit-reproducible, which gradient-checkpoint
    // recompute and the DDP equality tests rely on.
    if constexpr (kWide) {
      // every warp has left the ring: park the eight accumulators there, [warp][c][j ^ swz(c)] (the XOR spreads
      // the eight rows of a store over the banks), then all threads add the eight copies in warp order
      __syncthreads();
      float* red = reinterpret_cast<float*>(smem);
#pragma unroll
      for (int cbk = 0; cbk < NCB; ++cbk)
#pragma unroll
        for (int n = 0; n < NB; ++n)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int c = cbk * 16 + (lane >> 2) + h * 8;
            const int j = (n * 8 + (lane & 3) * 2) ^ ((c & 7) * (kRPad / 8));
            *reinterpret_cast<float2*>(red + (warp * COLS + c) * kRPad + j) = make_float2(acc[cbk][n][2 * h], acc[cbk][n][2 * h + 1]);
          }
    } else
    for (int turn = 0; turn < CP_CONSUMERS; ++turn) {
      if (warp == turn) {
#pragma unroll
        for (int cbk = 0; cbk < NCB; ++cbk)
#pragma unroll
          for (int n = 0; n < NB; ++n)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = cbk * 16 + (lane >> 2) + ((e >> 1) ? 8 : 0);
              const int j = n * 8 + (lane & 3) * 2 + (e & 1);
              sout[c * kRPad + j] += acc[cbk][n][e];
            }
      }
      asm volatile("bar.sync 1, %0;" ::"n"(CP_CONSUMERS * 32) : "memory");
    }
  }
  __syncthreads();
  float* dst = p.partial + (static_cast<size_t>(blockIdx.y) * p.C + c0) * kRPad;
  const int ncols = min(COLS, p.C - c0);
  if constexpr (kWide) {
    const float* red = reinterpret_cast<const float*>(smem);
    for (int i = threadIdx.x; i < ncols * kRPad; i += THREADS) {
      const int c = i / kRPad, jj = i % kRPad;
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < CP_CONSUMERS; ++w) s += red[(w * COLS + c) * kRPad + jj];
      dst[c * kRPad + (jj ^ ((c & 7) * (kRPad / 8)))] = s;
    }
  } else {
    for (int i = threadIdx.x; i < ncols * kRPad; i += THREADS) dst[i] = sout[i];
  }

  // The last CTA of this column tile to finish folds the splits, always in split order 0,1,2,...
  // (the arrival order only decides WHO adds, never the order of the additions), applies the
  // scale and writes the gradient: no separate reduction launch.
  __shared__ int s_is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicInc(p.counters + blockIdx.x, static_cast<unsigned int>(p.splits - 1));
    s_is_last = (prev == static_cast<unsigned int>(p.splits - 1));
  }
  __syncthreads();
  if (s_is_last) {
    __threadfence();
    for (int i = threadIdx.x; i < ncols * p.r; i += THREADS) {
      const int cl = i / p.r, j = i - cl * p.r;
      const float* src = p.partial + static_cast<size_t>(c0 + cl) * kRPad + j;
      // split order 0,1,2,... as before, but four loads in flight at a time: with r = 64 and a dozen splits the
      // serial chain of L2 round trips made this tail the longest part of the kernel (110 us for 768 -> 768)
      const size_t sstride = static_cast<size_t>(p.C) * kRPad;
      float acc = 0.f;
      int sp = 0;
      for (; sp + 4 <= p.splits; sp += 4) {
        const float v0 = __ldcg(src + (sp + 0) * sstride), v1 = __ldcg(src + (sp + 1) * sstride);
        const float v2 = __ldcg(src + (sp + 2) * sstride), v3 = __ldcg(src + (sp + 3) * sstride);
        acc += v0;
        acc += v1;
        acc += v2;
        acc += v3;
      }
      for (; sp < p.splits; ++sp) acc += __ldcg(src + sp * sstride);
      float* o = p.out + static_cast<long long>(c0 + cl) * p.ld_c + static_cast<long long>(j) * p.ld_j;
      *o = p.accumulate ? (*o + p.scale * acc) : (p.scale * acc);
    }
  }
}

}  // namespace skinny
}  // namespace lb
