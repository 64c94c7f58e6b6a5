// Shared device helpers for the sm_100a LoRA-linear kernels: mbarrier, TMA,
// tcgen05/TMEM wrappers (inline PTX), Philox4x32-10, small packing utilities.
// Nothing here has a counterpart in the reference (it has no native code,
// SURVEY.md section 2); the arithmetic these helpers serve is peft's
// lora.Linear as called from finetune_lora_distilgpt2.py:98-104 and
// finetune_lora_opt_pp.py:104-112.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace lb {

// ---------------------------------------------------------------------------
// dtype tags used across the C-ABI (include/lora_b200.h)
// ---------------------------------------------------------------------------
constexpr int kDtypeBf16 = 0;
constexpr int kDtypeFp16 = 1;

// A waiter that sees no progress for this many SM cycles traps instead of
// hanging the GPU (a hung box costs the whole pool a reset).
#ifndef LB_WATCHDOG_CYCLES
#define LB_WATCHDOG_CYCLES (1ll << 33)
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------------------
// Programmatic dependent launch: every kernel of a forward/backward call is launched with the
// programmatic-stream-serialization attribute, signals `launch_dependents` at once (the next
// kernel's CTAs may take over SMs as ours retire and run their set-up: barrier init, TMEM
// allocation, tensor-map prefetch) and executes `wait` before its first access to memory that an
// earlier kernel of the stream wrote (wait returns when the preceding grid has fully completed
// and its writes are visible).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Arrive on the barrier at the same smem offset in CTA `cta` of this cluster.
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __noinline__ void mbar_watchdog_trap(int tag, uint32_t parity) {
  printf("[lora_b200] mbarrier watchdog: block %d thread %d tag %d parity %u\n", (int)blockIdx.x,
         (int)threadIdx.x, tag, parity);
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, int tag = 0) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (((++spins) & 0xfffu) == 0 && (clock64() - t0) > LB_WATCHDOG_CYCLES) mbar_watchdog_trap(tag, parity);
  }
}

// ---------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int32_t c0,
                                            int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// 2-CTA variant: data lands in this CTA's smem, bytes are credited to the
// barrier at the same offset in the pair's leader CTA (peer bit cleared).
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int32_t c0,
                                                int32_t c1) {
  const uint32_t bar_addr = smem_u32(bar) & 0xFEFFFFFFu;
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
      "%4}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(tm), "r"(bar_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, const void* smem_src, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tm),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_group_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
// Generic-proxy smem writes -> visible to the async proxy (TMA store, UMMA).
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------------------
// tcgen05 / TMEM
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread for the whole CTA.
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// CTA-pair form: issued by the leader CTA only, M spans both CTAs.
__device__ __forceinline__ void umma_f16_2sm(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on `bar` once every MMA issued so far by this thread has retired.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// Pair form: arrives on the barrier at this offset in every CTA of `cta_mask`.
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(cta_mask)
      : "memory");
}

// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread.
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// UMMA shared-memory matrix descriptor (sm_100 "version 1").
//   layout_type: 0 none, 2 = 128B swizzle, 4 = 64B, 6 = 32B.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                                   uint32_t layout_type) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(layout_type & 7u) << 61;
  return d;
}
// kind::f16 instruction descriptor: fp32 accumulate, A/B in bf16 (fmt=1) or fp16 (fmt=0).
__host__ __device__ __forceinline__ uint32_t make_idesc_f16(int fmt, int umma_m, int umma_n, int a_mn_major,
                                                            int b_mn_major) {
  return (1u << 4) | (static_cast<uint32_t>(fmt) << 7) | (static_cast<uint32_t>(fmt) << 10) |
         (static_cast<uint32_t>(a_mn_major) << 15) | (static_cast<uint32_t>(b_mn_major) << 16) |
         (static_cast<uint32_t>(umma_n >> 3) << 17) | (static_cast<uint32_t>(umma_m >> 4) << 24);
}

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) and the layout of the dropout stream.
//
// The [M,K] adapter input is cut into 16x16 element blocks (block (R,C) = rows 16R.., columns
// 16C..).  Each block takes 32 Philox calls, one per lane of the warp that masks it; call t draws
// for exactly the eight elements lane t holds in the A-fragment of a warp-level m16n8k16 MMA over
// that block, so the row projection masks its fragments with its own draw and nothing is exchanged
// between lanes:
//     element (i, j) of the block  ->  call t = 4*(i%8) + (j%8)/2,
//                                      16-bit lane e = 4*(j/8) + 2*(i/8) + j%2 of the 128 random bits
// counter = (g_lo, g_hi, offset_lo, offset_hi) with g = (R * ceil(K/16) + C) * 32 + t, key = seed.
// An element survives iff the low 14 bits of its lane are >= round(p * 16384).
// What is STORED is layout-neutral: keep[M, K/8] bytes, bit e of byte (m, k/8) = element
// (m, 8*(k/8) + e).  The same mapping is restated in numpy in oracle/philox.py so that CPU and GPU
// share masks bit for bit.
// ---------------------------------------------------------------------------
__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                       uint32_t k0, uint32_t k1, uint32_t (&out)[4]) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint64_t p0 = static_cast<uint64_t>(M0) * c0;
    const uint64_t p1 = static_cast<uint64_t>(M1) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = static_cast<uint32_t>(p1);
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = static_cast<uint32_t>(p0);
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// Philox counter of call `t` for block (row16, col16); blocks_per_row = ceil(K / 16).
__host__ __device__ __forceinline__ uint64_t dropout_counter(uint32_t row16, uint32_t col16, uint32_t blocks_per_row,
                                                             uint32_t t) {
  return (static_cast<uint64_t>(row16) * blocks_per_row + col16) * 32ull + t;
}
// Keep-byte of row m, columns 8*c8 .. 8*c8+7; plain reference form (four calls, two bits from each).
__host__ __device__ __forceinline__ uint32_t dropout_keep_byte(uint32_t m, uint32_t c8, uint32_t blocks_per_row,
                                                               uint64_t seed, uint64_t offset, uint32_t thresh14) {
  const uint32_t i = m & 15u, w = (i >> 3) + 2u * (c8 & 1u);  // 32-bit word of the draw that holds this row's pair
  uint32_t bits = 0;
#pragma unroll
  for (uint32_t q = 0; q < 4; ++q) {
    const uint64_t g = dropout_counter(m >> 4, c8 >> 1, blocks_per_row, 4u * (i & 7u) + q);
    uint32_t r[4];
    philox4x32_10(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), static_cast<uint32_t>(offset),
                  static_cast<uint32_t>(offset >> 32), static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32),
                  r);
    const uint32_t lo = r[w] & 0x3FFFu, hi = (r[w] >> 16) & 0x3FFFu;
    bits |= (lo >= thresh14 ? 1u : 0u) << (2u * q);
    bits |= (hi >= thresh14 ? 1u : 0u) << (2u * q + 1u);
  }
  return bits;
}
// Round keys of Philox4x32-10 for a fixed seed: hoisted out of the per-call rounds (the seed is a
// kernel constant), which removes 20 adds from every draw.  The host fills them into the kernel's
// parameter block, so the rounds can take them as constant-bank operands and no register holds a key.
struct PhiloxKeys {
  uint32_t k[10][2];
};
__host__ __device__ __forceinline__ PhiloxKeys philox_round_keys(uint64_t seed) {
  PhiloxKeys rk;
  uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    rk.k[i][0] = k0;
    rk.k[i][1] = k1;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return rk;
}
#ifdef __CUDACC__
// thresh14 -> the packed-half constant dropout_fragment_masks compares against
__host__ __device__ __forceinline__ uint32_t dropout_thresh_half2(uint32_t thresh14) {
  return thresh14 * 0x00010001u;
}
// One draw for the calling lane: mask[w] is the AND-mask (0xFFFF per surviving 16-bit element) of
// A-fragment register w of an m16n8k16 MMA over the block whose counter base is (g_lo, g_hi) + lane.
// The 14-bit lanes are non-negative finite fp16 bit patterns (0x0000..0x3FFF: zero, subnormals, normals below 2),
// whose floating-point order equals their integer order, and set.ge.f16x2 without .ftz honours subnormals:
// one HSET2 per word gives both masks.
// Returns the eight keep-bits of the draw, bit 2w + h = element h of register w.
__device__ __forceinline__ uint32_t dropout_fragment_masks(uint32_t g_lo, uint32_t g_hi, uint32_t off_lo,
                                                           uint32_t off_hi, const PhiloxKeys& rk, uint32_t thresh_h2,
                                                           uint32_t (&mask)[4]) {
  uint32_t c0 = g_lo, c1 = g_hi, c2 = off_lo, c3 = off_hi;
#ifndef LB_EXP_PHILOX_ROUNDS
#define LB_EXP_PHILOX_ROUNDS 10  // experiment builds only (timing); the shipped library always runs 10 rounds
#endif
#pragma unroll
  for (int i = 0; i < LB_EXP_PHILOX_ROUNDS; ++i) {
#ifdef LB_EXP_PHILOX_MULHI
    const uint32_t p0l = 0xD2511F53u * c0, p0h = __umulhi(0xD2511F53u, c0);
    const uint32_t p1l = 0xCD9E8D57u * c2, p1h = __umulhi(0xCD9E8D57u, c2);
    const uint32_t n0 = p1h ^ c1 ^ rk.k[i][0];
    const uint32_t n2 = p0h ^ c3 ^ rk.k[i][1];
    c1 = p1l;
    c3 = p0l;
#else
    const uint64_t p0 = static_cast<uint64_t>(0xD2511F53u) * c0;
    const uint64_t p1 = static_cast<uint64_t>(0xCD9E8D57u) * c2;
    const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c1 ^ rk.k[i][0];
    const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c3 ^ rk.k[i][1];
    c1 = static_cast<uint32_t>(p1);
    c3 = static_cast<uint32_t>(p0);
#endif
    c0 = n0;
    c2 = n2;
  }
  const uint32_t r[4] = {c0, c1, c2, c3};
  const __half2 th = *reinterpret_cast<const __half2*>(&thresh_h2);
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const uint32_t x = r[w] & 0x3FFF3FFFu;
    mask[w] = __hge2_mask(*reinterpret_cast<const __half2*>(&x), th);
  }
  // byte 0 of each half-word mask -> one bit each: (m0.lo, m0.hi, m1.lo, m1.hi | m2.lo, ...)
  const uint32_t w01 = __byte_perm(mask[0], mask[1], 0x6420);
  const uint32_t w23 = __byte_perm(mask[2], mask[3], 0x6420);
  const uint32_t spread = (w01 & 0x08040201u) | (w23 & 0x80402010u);
  return (spread * 0x01010101u) >> 24;
}
#endif

// ---------------------------------------------------------------------------
// 16-bit float helpers (dtype is a runtime tag in the tcgen05 kernel)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack2(float a, float b, int dtype) {
  if (dtype == kDtypeBf16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
// Packed FP32 FMA (FFMA2 on sm_100a): two fused multiply-adds per issued instruction.  d = a * b + d, pairwise.
__device__ __forceinline__ void ffma2(float2& d, const float2 a, const float2 b) {
  unsigned long long dd = *reinterpret_cast<unsigned long long*>(&d);
  asm("fma.rn.f32x2 %0, %1, %2, %0;"
      : "+l"(dd)
      : "l"(*reinterpret_cast<const unsigned long long*>(&a)), "l"(*reinterpret_cast<const unsigned long long*>(&b)));
  d = *reinterpret_cast<float2*>(&dd);
}

__device__ __forceinline__ float2 unpack2(uint32_t w, int dtype) {
  if (dtype == kDtypeBf16) {
    return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&w));
  } else {
    return __half22float2(*reinterpret_cast<__half2*>(&w));
  }
}
__device__ __forceinline__ float to_float16bits(uint16_t h, int dtype) {
  if (dtype == kDtypeBf16) return __uint_as_float(static_cast<uint32_t>(h) << 16);
  return __half2float(__ushort_as_half(h));
}

}  // namespace lb
