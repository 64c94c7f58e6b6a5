// The path's one exchange step under DDP -- averaging the flat adapter-gradient bucket across the
// GPUs of one NVSwitch box -- as ONE in-stream kernel over peer memory instead of a library
// collective on a side stream (finetune_lora_distilgpt2.py:142 leaves this to torch DDP over gloo;
// SURVEY.md 8e).  Every rank owns a "symmetric" buffer that all peers have mapped (CUDA IPC):
//
//   [ Ctrl | data[2][cap] | result[2][cap] ]        (fp32; [2] = parity of the call sequence number)
//
// one-shot (small buckets: the 164 KB of one layer, distilgpt2's 1.6 MB):
//   1. copy the local bucket into data[parity]                      (local HBM)
//   2. last CTA to finish publishes flagA[parity][rank] = seq in EVERY peer's Ctrl (st.release.sys over
//      NVLink); every CTA then polls its OWN Ctrl until all `world` flags carry seq
//   3. bucket[i] = (sum over ranks q = 0..world-1 of peer q's data[parity][i]) / world, P2P loads
//      through NVSwitch, fixed summation order => every rank computes the bit-identical average
// two-shot (large buckets: OPT's 10-47 MB): step 3 only over this rank's 1/world slice, the averaged
//   slice is pushed into every peer's result[parity] (P2P stores), a second flag round (flagB), then a
//   local copy result[parity] -> bucket.  Per rank (world-1)/world * 2n values cross NVLink instead of
//   (world-1) * n.
//
// The sequence number of a call lives in Ctrl (starts at 0, the published count): no host-side state,
// so the launch can sit in a CUDA graph that is replayed.
//
// Why two parities are enough with one flag round per call: a rank overwrites data[parity] two calls
// later, after it has passed the flag wait of the call in between -- which every peer enters only after
// finishing its reads of this call (kernels of one stream run in order).  Flags compare with ==, and a
// peer cannot publish seq + 2 before this rank has published seq + 1, i.e. has consumed seq.
//
// Nothing here allocates or synchronises the host; a peer that never arrives trips a clock64 watchdog
// (PEER_WATCHDOG_CYCLES, minutes) and the kernel traps instead of spinning for ever.
#pragma once

#include "common.cuh"

namespace lb {
namespace p2p {

constexpr int MAX_WORLD = 8;
// 256-thread CTAs (60 registers each) fit beside a resident CTA of the forward GEMM (256 x 139 registers, all of
// the SM's shared memory), so an exchange that runs on a side stream under the next step's forward does not keep
// the GEMM's persistent CTAs off the SMs it sits on.
constexpr int THREADS = 256;
constexpr int MAX_CTAS = 128;  // all CTAs must become resident (they wait for each other): below 148 SMs
constexpr int CTRL_BYTES = 1024;
// A collective must tolerate a rank that is late for ordinary reasons (first-iteration library set-up, a slow input
// pipeline, a garbage-collection pause): the peers' wait gives up only after 2^40 cycles (~9 minutes at 1.96 GHz),
// far beyond the 2^33 cycles (~4 s) of the intra-kernel mbarrier watchdog.
constexpr long long PEER_WATCHDOG_CYCLES = 1ll << 40;

struct Ctrl {
  uint32_t flag_a[2][MAX_WORLD];  // [parity][source rank] = seq: source's data[parity] is complete
  uint32_t flag_b[2][MAX_WORLD];  // [parity][source rank] = seq: source's slice sits in my result[parity]
  uint32_t arrive[2][2];          // [parity][round]: CTAs of THIS rank that finished the round (self-resetting)
  uint32_t seq;                   // number of exchanges this rank has published (advances on the device, so a
                                  // captured CUDA graph can be replayed); identical on all ranks by construction
};
static_assert(sizeof(Ctrl) <= CTRL_BYTES, "control block");

struct Peers {
  void* sym[MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer data is read with system-scope relaxed loads: they cannot be served from a stale L1 line.
__device__ __forceinline__ float4 ld_peer_v4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ld_peer_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// Grid-wide "everyone of this rank is done" + publish to all peers + wait for all peers.
__device__ __forceinline__ void publish_and_wait(const Peers& peers, int rank, int world, int parity, int round, uint32_t seq) {
  Ctrl* mine = reinterpret_cast<Ctrl*>(peers.sym[rank]);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // this CTA's writes (local copy / slice pushed over NVLink) precede its arrival
    const uint32_t prev = atomicAdd(&mine->arrive[parity][round], 1u);
    if (prev == gridDim.x - 1) {
      mine->arrive[parity][round] = 0u;  // next use is two calls away
      if (round == 0) mine->seq = seq;   // every CTA of this launch has read the old value by now
      __threadfence_system();
      for (int q = 0; q < world; ++q) {
        Ctrl* c = reinterpret_cast<Ctrl*>(peers.sym[q]);
        st_release_sys(round == 0 ? &c->flag_a[parity][rank] : &c->flag_b[parity][rank], seq);
      }
    }
  }
  if (threadIdx.x < world) {
    const uint32_t* f = round == 0 ? &mine->flag_a[parity][threadIdx.x] : &mine->flag_b[parity][threadIdx.x];
    const long long t0 = clock64();
    uint32_t spins = 0;
    while (ld_acquire_sys(f) != seq) {
      if (((++spins) & 0x3ffu) == 0 && (clock64() - t0) > PEER_WATCHDOG_CYCLES) {
        printf("lora_b200: peer rank %d never arrived at exchange %u (round %d); aborting\n", (int)threadIdx.x, seq, round);
        __trap();
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(THREADS)
    bucket_allreduce_kernel(float* __restrict__ bucket, long long n, const Peers peers, int rank, int world,
                            long long cap, int two_shot) {
  pdl_wait();  // the bucket is written by the kernels before this one in the stream
  const uint32_t seq = *reinterpret_cast<volatile const uint32_t*>(&reinterpret_cast<const Ctrl*>(peers.sym[rank])->seq) + 1u;
  const int parity = static_cast<int>(seq & 1u);
  const long long n4 = n >> 2;
  const long long tid = blockIdx.x * static_cast<long long>(THREADS) + threadIdx.x;
  const long long nthr = static_cast<long long>(gridDim.x) * THREADS;
  auto data_of = [&](int q) { return reinterpret_cast<float*>(static_cast<uint8_t*>(peers.sym[q]) + CTRL_BYTES) + parity * cap; };
  auto result_of = [&](int q) {
    return reinterpret_cast<float*>(static_cast<uint8_t*>(peers.sym[q]) + CTRL_BYTES) + (2 + parity) * cap;
  };
  const float inv = 1.f / static_cast<float>(world);

  // 1. local bucket -> my data[parity]
  {
    float4* dst = reinterpret_cast<float4*>(data_of(rank));
    const float4* src = reinterpret_cast<const float4*>(bucket);
    for (long long i = tid; i < n4; i += nthr) dst[i] = src[i];
    for (long long i = (n4 << 2) + tid; i < n; i += nthr) data_of(rank)[i] = bucket[i];
  }
  publish_and_wait(peers, rank, world, parity, 0, seq);

  if (!two_shot) {
    // 3. every rank reduces the whole bucket from all peers, same order everywhere
    float4* out = reinterpret_cast<float4*>(bucket);
    for (long long i = tid; i < n4; i += nthr) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int q = 0; q < MAX_WORLD; ++q) {
        if (q < world) {
          const float4 v = ld_peer_v4(reinterpret_cast<const float4*>(data_of(q)) + i);
          acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
        }
      }
      out[i] = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
    }
    for (long long i = (n4 << 2) + tid; i < n; i += nthr) {
      float acc = 0.f;
      for (int q = 0; q < world; ++q) acc += ld_peer_f32(data_of(q) + i);
      bucket[i] = acc * inv;
    }
    return;
  }

  // two-shot: my slice of float4s [lo, hi) is reduced here and pushed to everyone (n % 4 == 0 is the host's check)
  const long long lo = n4 * rank / world, hi = n4 * (rank + 1) / world;
  for (long long i = lo + tid; i < hi; i += nthr) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int q = 0; q < MAX_WORLD; ++q) {
      if (q < world) {
        const float4 v = ld_peer_v4(reinterpret_cast<const float4*>(data_of(q)) + i);
        acc.x += v.x, acc.y += v.y, acc.z += v.z, acc.w += v.w;
      }
    }
    acc = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
#pragma unroll
    for (int q = 0; q < MAX_WORLD; ++q)
      if (q < world) reinterpret_cast<float4*>(result_of(q))[i] = acc;
  }
  publish_and_wait(peers, rank, world, parity, 1, seq);
  {
    const float4* src = reinterpret_cast<const float4*>(result_of(rank));
    float4* dst = reinterpret_cast<float4*>(bucket);
    for (long long i = tid; i < n4; i += nthr) dst[i] = __ldcg(src + i);  // written by peers: not through L1
  }
}

}  // namespace p2p
}  // namespace lb
