// Row projection for FEW rows (M <= 4736: pipeline micro-batches of 2048 tokens, short sequences):
//     Out[M, kRPad] = out_scale * (keep .* Act)[M, C] * P[kRPad, C]^T        (H = drop(X) A^T, dH = dY (sB))
// skinny::rowproj_kernel gives a 16-row slab to two warps, each of which walks the whole of C; with few rows that
// walk IS the kernel's time (M = 4096, C = 10240: 69 us with every SM holding a CTA, where a quarter of the
// M = 16384 time would be 35 us).  Here a CTA is ONE slab and EIGHT consumer warps split C between them: warp w
// takes k16 step (w & 3) of the 64-column tiles of parity (w >> 2), i.e. an eighth of the walk, one Philox draw per
// tile, four tiles (= four independent draws) per trip.  Same dropout stream, same keep-byte layout, same operands
// as the large-M kernel (common.cuh "layout of the dropout stream"); the eight partial accumulators are summed in
// warp order through shared memory (bit-reproducible).  Ring: 16 (8) slots; a slot always carries tiles of one parity,
// so its `empty` barrier counts the four warps of that parity.
#pragma once

#include "skinny.cuh"

namespace lb {
namespace skinny {

// ring depth (even: slot s <-> tiles of parity s & 1).  A slot is only 6 KiB (10 at rank 64), and with one or two
// CTAs per SM the bytes in flight are what the ring holds: 16 slots where they fit twice into an SM
__host__ __device__ constexpr int rs_stages(int rpad) { return rpad == 64 ? 8 : 16; }
constexpr int RS_CONSUMERS = 8;
constexpr int RS_THREADS = (RS_CONSUMERS + 1) * 32;
constexpr int RS_ACT_BYTES = 2048;  // 16 rows x 128 bytes
template <int kRPad>
__host__ __device__ constexpr int rs_slot_bytes() {
  return RS_ACT_BYTES + (kRPad == 16 ? 4096 : ((kRPad * 128 + 1023) / 1024) * 1024);
}
template <int kRPad>
__host__ __device__ constexpr int rs_smem_bytes() {
  return rs_stages(kRPad) * rs_slot_bytes<kRPad>() + (RS_CONSUMERS - 1) * 16 * kRPad * 4 + 1024 + 256;
}

template <bool kBf16, int kRPad, bool kDirectA = false, int kNBE = 2>
__global__ void __launch_bounds__(RS_THREADS, kRPad == 64 ? 1 : 2)
    rowproj_small_kernel(const __grid_constant__ CUtensorMap tmAct, const __grid_constant__ CUtensorMap tmP,
                         const RowProjParams p) {
  static_assert(!kDirectA || kRPad == 16, "the direct-A path stages 8 / 16 fp32 rows per box: padded rank 16 only");
  constexpr int A32_BOX_BYTES = 8 * kNBE * 128;
  constexpr int SLOT = rs_slot_bytes<kRPad>();
  constexpr int NB = kRPad / 8;
  constexpr int RS_STAGES = rs_stages(kRPad);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* sred = reinterpret_cast<float*>(smem + RS_STAGES * SLOT);  // [7 warps][16 rows][kRPad]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + RS_STAGES * SLOT + (RS_CONSUMERS - 1) * 16 * kRPad * 4);
  uint64_t* empty = full + RS_STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * 16;
  const int nk = (p.C + RP_BK - 1) / RP_BK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < RS_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], RS_CONSUMERS / 2);
    }
    fence_barrier_init();
  }
  pdl_launch_dependents();
  __syncthreads();
  pdl_wait();
  if (blockIdx.x == 0 && p.zero_words != nullptr)
    for (int i = threadIdx.x; i < p.n_zero_words; i += blockDim.x) p.zero_words[i] = 0u;
  if constexpr (kDirectA) {
    using T16 = typename std::conditional<kBf16, __nv_bfloat16, __half>::type;
    pack_adapters_body<T16>(p.A32, p.B32, static_cast<T16*>(p.a16), static_cast<T16*>(p.a16t), static_cast<T16*>(p.sb16),
                            static_cast<T16*>(p.sb16t), p.C, p.N, p.r, kRPad, p.s,
                            blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x,
                            static_cast<long long>(gridDim.x) * blockDim.x);
  }

  if (warp == RS_CONSUMERS) {
    // ===================== TMA producer =====================
    // (a second issuing lane for the projection matrix was tried: slower, 54 against 50 us at M = 2048, C = 10240;
    // so was a ring of 8 instead of 16 slots: the same)
    if (lane == 0) {
      for (int kb = 0; kb < nk; ++kb) {
        const int st = kb % RS_STAGES;
        mbar_wait(&empty[st], ((kb / RS_STAGES) & 1) ^ 1, 13);
        uint8_t* act = smem + st * SLOT;
        mbar_arrive_expect_tx(&full[st], 16 * 128 + (kDirectA ? 2 * A32_BOX_BYTES : kRPad * 128));
        tma_load_2d(act, &tmAct, &full[st], kb * RP_BK, m0);
        tma_load_2d(act + RS_ACT_BYTES, &tmP, &full[st], kb * RP_BK, 0);
        if constexpr (kDirectA) tma_load_2d(act + RS_ACT_BYTES + A32_BOX_BYTES, &tmP, &full[st], kb * RP_BK + 32, 0);
      }
    }
    __syncwarp();
    return;
  }

  const int ks = warp & 3;    // k16 step of every tile this warp takes
  const int par = warp >> 2;  // ... of the tiles of this parity
  const int g8 = lane >> 2, q = lane & 3;
  float acc[NB][4];
#pragma unroll
  for (int n = 0; n < NB; ++n)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[n][e] = 0.f;

  const int keep_ld = p.C >> 3;
  const PhiloxKeys& rk = p.rk;
  const uint32_t thresh2 = dropout_thresh_half2(p.thresh14);
  const uint64_t offset = p.offset + ((p.dropout && p.offset_dev != nullptr) ? *p.offset_dev : 0ull);
  const uint32_t off_lo = static_cast<uint32_t>(offset), off_hi = static_cast<uint32_t>(offset >> 32);
  const uint32_t blocks_per_row = static_cast<uint32_t>((p.C + 15) >> 4);
  const uint64_t gbase = dropout_counter(static_cast<uint32_t>(m0 >> 4), static_cast<uint32_t>(ks), blocks_per_row,
                                         static_cast<uint32_t>(lane));
  // keep-bytes: lane q == 0 stores the two bytes (16 columns) of row g8 for this warp's k16 step, q == 1 those of g8 + 8
  const int krow = m0 + g8 + (q & 1) * 8;
  uint8_t* kdst = p.dropout ? p.keep_out + static_cast<size_t>(krow) * keep_ld + ks * 2 : nullptr;
  const bool kstore = p.dropout && q < 2 && krow < p.M;
  const bool kfast = ((keep_ld | static_cast<int>(reinterpret_cast<uintptr_t>(p.keep_out))) & 1) == 0;

  const int a_row = (lane & 7) + ((lane >> 3) & 1) * 8;
  const uint32_t a_off = static_cast<uint32_t>(a_row * 128);
  const int a_sw = a_row & 7, a_hi = lane >> 4;
  const int b_row = lane & 7, b_hi = (lane >> 3) & 1;

  for (int kbt = par; kbt < nk; kbt += 8) {  // a trip: own tiles kbt, kbt + 2, kbt + 4, kbt + 6
    uint32_t kmask[4][4];
    if (p.dropout) {
      uint32_t kw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // four independent draws, before any shuffle
        const uint64_t g = gbase + static_cast<uint64_t>((kbt + 2 * u) * 4) * 32ull;
        kw[u] = dropout_fragment_masks(static_cast<uint32_t>(g), static_cast<uint32_t>(g >> 32), off_lo, off_hi, rk, thresh2,
                                       kmask[u]);
      }
      // 2-bit field of register w -> byte w of a word, at bit 2q; the quad's four pieces OR into whole bytes:
      // byte 0 = (row g8, chunk 0), byte 1 = (row g8+8, chunk 0), byte 2 = (g8, chunk 1), byte 3 = (g8+8, chunk 1)
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t t = (kw[u] * 0x1001u) & 0x000F000Fu;
        kw[u] = ((t * 0x41u) & 0x03030303u) << (2 * q);
      }
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1)
#pragma unroll
        for (int u = 0; u < 4; ++u) kw[u] |= __shfl_xor_sync(0xFFFFFFFFu, kw[u], o);
      if (kstore) {
        const uint32_t sel = (q & 1) ? 0x4431u : 0x4420u;  // this lane's row: (chunk 0, chunk 1) of the k16 step
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int kb = kbt + 2 * u;
          if (kb >= nk) break;
          const uint32_t two = __byte_perm(kw[u], 0u, sel);
          const int c8 = kb * 8 + ks * 2;  // first chunk column of the pair
          uint8_t* dst = kdst + kb * 8;
          if (kfast && c8 + 2 <= keep_ld) {
            *reinterpret_cast<uint16_t*>(dst) = static_cast<uint16_t>(two);
          } else {
            if (c8 < keep_ld) dst[0] = static_cast<uint8_t>(two);
            if (c8 + 1 < keep_ld) dst[1] = static_cast<uint8_t>(two >> 8);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int kb = kbt + 2 * u;
      if (kb >= nk) break;
      const int st = kb % RS_STAGES;
      mbar_wait(&full[st], (kb / RS_STAGES) & 1, 14);
      const uint32_t act_u = smem_u32(smem + st * SLOT), pt_u = act_u + RS_ACT_BYTES;
      uint32_t a[4];
      ldsm_x4(act_u + a_off + (((ks * 2 + a_hi) ^ a_sw) << 4), a);
      if (p.dropout) {
        a[0] &= kmask[u][0]; a[1] &= kmask[u][1]; a[2] &= kmask[u][2]; a[3] &= kmask[u][3];
      }
#pragma unroll
      for (int n = 0; n < NB; ++n) {
        uint32_t b[2];
        if constexpr (kDirectA) {
          if (n >= kNBE) continue;  // all-padding rank columns contribute nothing
          const int arow = n * 8 + g8;
          const uint32_t abox = pt_u + (ks >> 1) * A32_BOX_BYTES + arow * 128 + (q & 1) * 8;
          const int ch = (ks & 1) * 4 + (q >> 1);
          const uint2 lo = lds64(abox + ((ch ^ (arow & 7)) << 4)), hi = lds64(abox + (((ch + 2) ^ (arow & 7)) << 4));
          b[0] = pack2(__uint_as_float(lo.x), __uint_as_float(lo.y), kBf16 ? kDtypeBf16 : kDtypeFp16);
          b[1] = pack2(__uint_as_float(hi.x), __uint_as_float(hi.y), kBf16 ? kDtypeBf16 : kDtypeFp16);
        } else {
          const int prow = n * 8 + b_row;
          ldsm_x2(pt_u + prow * 128 + (((ks * 2 + b_hi) ^ (prow & 7)) << 4), b);
        }
        mma_16816<kBf16>(acc[n], a, b);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[st]);
    }
  }

  // acc fragment: (row = lane/4 [+8], col = n*8 + (lane%4)*2 [+1]); warps 1..7 hand their partial sums to warp 0
  // through smem, which adds them in warp order (fixed order -> bit-reproducible).
  if (warp > 0) {
    float* mine = sred + (warp - 1) * 16 * kRPad;
#pragma unroll
    for (int n = 0; n < NB; ++n) {
      const int col = n * 8 + q * 2;
      *reinterpret_cast<float2*>(mine + g8 * kRPad + col) = make_float2(acc[n][0], acc[n][1]);
      *reinterpret_cast<float2*>(mine + (g8 + 8) * kRPad + col) = make_float2(acc[n][2], acc[n][3]);
    }
  }
  asm volatile("bar.sync 1, %0;" ::"n"(RS_CONSUMERS * 32) : "memory");
  if (warp == 0) {
    uint32_t* out = reinterpret_cast<uint32_t*>(p.out);
    const int r0 = m0 + g8;
#pragma unroll
    for (int n = 0; n < NB; ++n) {
      const int col = n * 8 + q * 2;
#pragma unroll
      for (int w = 0; w < RS_CONSUMERS - 1; ++w) {
        const float2 o0 = *reinterpret_cast<const float2*>(sred + w * 16 * kRPad + g8 * kRPad + col);
        const float2 o1 = *reinterpret_cast<const float2*>(sred + w * 16 * kRPad + (g8 + 8) * kRPad + col);
        acc[n][0] += o0.x;
        acc[n][1] += o0.y;
        acc[n][2] += o1.x;
        acc[n][3] += o1.y;
      }
      if (r0 < p.M)
        out[(static_cast<size_t>(r0) * kRPad + col) >> 1] =
            pack2(acc[n][0] * p.out_scale, acc[n][1] * p.out_scale, kBf16 ? kDtypeBf16 : kDtypeFp16);
      if (r0 + 8 < p.M)
        out[(static_cast<size_t>(r0 + 8) * kRPad + col) >> 1] =
            pack2(acc[n][2] * p.out_scale, acc[n][3] * p.out_scale, kBf16 ? kDtypeBf16 : kDtypeFp16);
    }
  }
}

}  // namespace skinny
}  // namespace lb
