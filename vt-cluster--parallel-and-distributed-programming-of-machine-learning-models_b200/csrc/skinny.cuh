// HBM-bound companions of the tcgen05 GEMM: the row projections, whose arithmetic
// intensity is ~r flop/byte,
//     H  = drop(X) A^T          dH = dY (sB)            (rowproj)
// plus the fp32 -> 16-bit adapter packing.  They stream the activation once
// through a TMA + mbarrier ring and use warp-level mma.sync on the tiles:
// the tensor pipe is nowhere near the limit here, HBM is.  (The column
// projections dB = s dY^T H, dA = dH^T drop(X) are in colproj_sm100.cuh.)
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
__device__ __forceinline__ uint2 lds64(uint32_t saddr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
  return v;
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
// kSlabs is a template parameter; the library runs 4 (few rows go to rowproj_small.cuh: one slab per CTA, eight warps
// splitting the columns).  Measured on the way, H with dropout at K = 2560: M = 4096 as 64 four-slab CTAs 35.7 us (the
// hardware packs small CTAs several to an SM, so the generator ran on 32-64 SMs), as 128 two-slab CTAs with an SM
// each 18.8 us; seven-slab CTAs, one per SM, at M = 16384: 35.2 against 36.2 us.
__host__ __device__ constexpr int rp_rows(int slabs) { return 16 * slabs; }
constexpr int RP_ONE_PER_SM_BYTES = 116 * 1024;  // more than half of an SM's 227 KiB: at most one CTA per SM
// Padded rank 64 needs ~123 registers per thread (32 accumulators, 16 B-fragment words).  ptxas sizes the register
// file for the warp count rounded up to an even number, so nine warps cap a 2-CTA/SM kernel at 96; with EIGHT warps
// (lane 0 of warp 0 also issues the TMA loads, one tile behind its own consumption) the cap is 128, two CTAs fit
// per SM and the 256 CTAs of M = 16384 make one wave instead of two.
#ifndef LB_EXP_RP_INLINE_ALL
#define LB_EXP_RP_INLINE_ALL 0  // measured (scripts/gpu_r2_rp_variants2.sh): at padded rank 16 the inline producer LOSES
#endif                          // (H with dropout 46.3 against 41.5 us, dH 17.0 / 15.7): the dedicated warp stays there
__host__ __device__ constexpr bool rp_inline_producer(int rpad) { return LB_EXP_RP_INLINE_ALL || rpad == 64; }
__host__ __device__ constexpr int rp_threads(int rpad, int slabs) { return (2 * slabs + (rp_inline_producer(rpad) ? 0 : 1)) * 32; }
#ifndef LB_EXP_RP_MIN_CTAS
#define LB_EXP_RP_MIN_CTAS 2  // measured: capping registers for a third CTA (72, spills) 64 us, uncapped 39 us
#endif
#ifndef LB_EXP_RP_STAGES
#define LB_EXP_RP_STAGES 6
#endif
__host__ __device__ constexpr int rp_min_ctas(int rpad) { return rpad <= 16 ? LB_EXP_RP_MIN_CTAS : 2; }
__host__ __device__ constexpr int rp_stages(int rpad) { return rpad == 64 ? 5 : LB_EXP_RP_STAGES; }  // 64: 2 x 98 KiB per SM
constexpr int RP_BK = 64;
template <int kSlabs>
__host__ __device__ constexpr int rp_act_bytes() {
  return ((rp_rows(kSlabs) * 128 + 1023) / 1024) * 1024;
}
template <int kRPad, int kSlabs>
__host__ __device__ constexpr int rp_slot_bytes() {
  // the projection-matrix region: kRPad rows x 128 bytes of 16-bit P, or (padded rank 16, direct-A) two boxes of
  // 8 / 16 rows x 32 fp32 columns straight from the master: 4 KiB covers both
  return rp_act_bytes<kSlabs>() + (kRPad == 16 ? 4096 : ((kRPad * 128 + 1023) / 1024) * 1024);
}
template <int kRPad, int kSlabs>
__host__ __device__ constexpr int rp_smem_bytes() {
  return rp_stages(kRPad) * rp_slot_bytes<kRPad, kSlabs>() + rp_rows(kSlabs) * kRPad * 4 + 1024 + 256;
}
template <int kRPad, int kSlabs>
__host__ __device__ constexpr int rp_smem_optin() {  // the launch may ask for RP_ONE_PER_SM_BYTES instead
  return rp_smem_bytes<kRPad, kSlabs>() > RP_ONE_PER_SM_BYTES ? rp_smem_bytes<kRPad, kSlabs>() : RP_ONE_PER_SM_BYTES;
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
  // kDirectA (forward, padded rank 16): the projection matrix comes straight from the fp32 master A [r, C] (tmP is a
  // tensor map over it: two 128-byte-swizzled boxes of 32 columns ride every stage, rows beyond r and columns beyond
  // C arrive as zeros) and the kernel also does the pack (A, B -> the four 16-bit layouts) as a side job for the
  // kernels that follow, so the forward needs no pack launch.
  const float* A32;
  const float* B32;  // [N, r]
  int r, N;
  float s;           // scaling folded into the packed B
  void *a16, *a16t, *sb16, *sb16t;
};

template <bool kBf16, int kRPad, bool kDirectA = false, int kNBE = 2, int kSlabs = 4>
__global__ void __launch_bounds__(rp_threads(kRPad, kSlabs), rp_min_ctas(kRPad))
    rowproj_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmP,
                   const RowProjParams p) {
  static_assert(!kDirectA || kRPad == 16, "the direct-A path stages 8 / 16 fp32 rows per box: padded rank 16 only");
  constexpr int A32_BOX_BYTES = 8 * kNBE * 128;  // direct-A: kNBE = 1 (r <= 8, the reference) loads 8 rows, else 16
  constexpr int SLOT = rp_slot_bytes<kRPad, kSlabs>();
  constexpr int NB = kRPad / 8;
  constexpr int SLABS = kSlabs;
  constexpr int RP_BM = rp_rows(kSlabs);
  constexpr int RP_CONSUMERS = 2 * SLABS;
  constexpr int RP_STAGES = rp_stages(kRPad);
  constexpr bool kInline = rp_inline_producer(kRPad);
  constexpr int RP_ACT_BYTES = rp_act_bytes<kSlabs>();
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

  // column tile kb -> ring slot kb % RP_STAGES, once every consumer warp has released the slot's previous tile
  auto produce = [&](int kb) {
    const int st = kb % RP_STAGES;
    mbar_wait(&empty[st], ((kb / RP_STAGES) & 1) ^ 1, 11);
    uint8_t* act = smem + st * SLOT;
    mbar_arrive_expect_tx(&full[st], RP_BM * 128 + (kDirectA ? 2 * A32_BOX_BYTES : kRPad * 128));
    tma_load_2d(act, &tmAct, &full[st], kb * RP_BK, m0);
    if constexpr (!kDirectA) {
      tma_load_2d(act + RP_ACT_BYTES, &tmP, &full[st], kb * RP_BK, 0);
    } else {
      tma_load_2d(act + RP_ACT_BYTES, &tmP, &full[st], kb * RP_BK, 0);
      tma_load_2d(act + RP_ACT_BYTES + A32_BOX_BYTES, &tmP, &full[st], kb * RP_BK + 32, 0);
    }
  };
  if (!kInline && warp == RP_CONSUMERS) {
    if (lane == 0)
      for (int kb = 0; kb < nk; ++kb) produce(kb);
    __syncwarp();
    return;
  }
  if (kInline && warp == 0 && lane == 0)
    for (int kb = 0; kb < min(nk, RP_STAGES); ++kb) produce(kb);

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
  // kDirectA: B fragments from the fp32 tile of A in the stage: lane (g8, q) needs A[n*8 + g8][c .. c+1] and
  // [c+8 .. c+9], c = 16-column step + 2q -- two 8-byte shared loads (the 128-byte swizzle spreads the eight rows of
  // a load over the banks) packed to 16 bits.  (First written with the fragments fetched from global memory into
  // registers one trip ahead: the L2 round trip sat on the pack, 11 % of the kernel's stall samples at M = 16384
  // and most of the time of a lone warp at small M.)
  for (int kb0 = 0; kb0 < nk; kb0 += 2) {
    uint32_t kmask[2][2][4];
    if (p.dropout) {
      uint32_t kw[2][2];
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint64_t g = gbase + static_cast<uint64_t>((kb0 + u) * 4 + kk) * 32ull;
          kw[u][kk] = dropout_fragment_masks(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), off_lo, off_hi, rk,
                                             thresh2, kmask[u][kk]);
        }
      // The four draws above are pure arithmetic and independent; the shuffles come only now, all together: a
      // shuffle between two draws keeps the compiler from interleaving their dependency chains.
      // 2-bit field of register w -> byte w of a word, at bit 2q; the quad's four pieces OR into whole bytes:
      // byte 0 = (row g8, chunk 0), byte 1 = (row g8+8, chunk 0), byte 2 = (g8, chunk 1), byte 3 = (g8+8, chunk 1)
#pragma unroll
      for (int u = 0; u < 2; ++u)
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          uint32_t t = (kw[u][kk] * 0x1001u) & 0x000F000Fu;
          kw[u][kk] = ((t * 0x41u) & 0x03030303u) << (2 * q);
        }
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1)
#pragma unroll
        for (int u = 0; u < 2; ++u)
#pragma unroll
          for (int kk = 0; kk < 2; ++kk) kw[u][kk] |= __shfl_xor_sync(0xFFFFFFFFu, kw[u][kk], o);
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
      // inline producer: the slot of the tile before this one is refilled here, when the other warps have (nearly) left it
      if (kInline && warp == 0 && lane == 0 && kb0 + u >= 1 && kb0 + u - 1 + RP_STAGES < nk) produce(kb0 + u - 1 + RP_STAGES);
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
            if (n >= kNBE) continue;  // all-padding rank columns contribute nothing
            const int arow = n * 8 + g8;
            const uint32_t abox = pt_u + (ks >> 1) * A32_BOX_BYTES + arow * 128 + (q & 1) * 8;
            const int ch = (ks & 1) * 4 + (q >> 1);  // 16-byte chunk of columns c .. c+1; columns c+8 .. c+9 are two chunks on
            const uint2 lo = lds64(abox + ((ch ^ (arow & 7)) << 4)), hi = lds64(abox + (((ch + 2) ^ (arow & 7)) << 4));
            b[0] = pack2(__uint_as_float(lo.x), __uint_as_float(lo.y), kBf16 ? kDtypeBf16 : kDtypeFp16);
            b[1] = pack2(__uint_as_float(hi.x), __uint_as_float(hi.y), kBf16 ? kDtypeBf16 : kDtypeFp16);
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

}  // namespace skinny
}  // namespace lb
