// C-ABI entry points of liblora_b200.so (declared in include/lora_b200.h) and
// the host-side launch logic: tensor-map encoding, workspace carving, kernel
// sequencing.  Replaces the peft lora.Linear forward and its autograd backward
// that the reference reaches through get_peft_model
// (finetune_lora_distilgpt2.py:98-104, finetune_lora_opt_pp.py:104-112).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <utility>

#include "act.cuh"
#include "ce.cuh"
#include "colproj_sm100.cuh"
#include "p2p.cuh"
#include "gemm_sm100.cuh"
#include "gemm_masked_sm100.cuh"
#include "lora_b200.h"
#include "optim.cuh"
#include "rowproj_small.cuh"
#include "skinny.cuh"

namespace {

using namespace lb;

thread_local std::string g_err;

int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

#define CUDA_TRY(expr)                                                                      \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess) return fail("%s failed: %s", #expr, cudaGetErrorString(e__));  \
  } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

// Host-side cache of encoded tensor maps (SURVEY.md 8b allows one): a training step asks for the same
// (pointer, shape, stride, box, swizzle, dtype) descriptors again and again -- W, bias-free operands, the
// packed adapters, activation buffers that the caching allocator hands back at the same address -- and
// cuTensorMapEncodeTiled costs about a microsecond each, 17-19 times per layer forward+backward.
// One small direct-mapped table per host thread (autograd's backward thread gets its own), no locks.
// A descriptor depends on nothing but its key, so a hit can never be stale.
struct TmapKey {
  const void* ptr;
  uint64_t inner, outer, stride;
  uint32_t box_inner, box_outer;
  int32_t dtype, swizzle;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && inner == o.inner && outer == o.outer && stride == o.stride && box_inner == o.box_inner &&
           box_outer == o.box_outer && dtype == o.dtype && swizzle == o.swizzle;
  }
};
struct TmapSlot {
  TmapKey key{};
  bool valid = false;
  CUtensorMap map;
};
constexpr int kTmapSlots = 1024;  // power of two
thread_local std::unique_ptr<TmapSlot[]> g_tmap_cache;

inline size_t tmap_hash(const TmapKey& k) {
  uint64_t h = reinterpret_cast<uintptr_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
  h ^= (k.inner * 0xC2B2AE3D27D4EB4Full) ^ (k.outer * 0x165667B19E3779F9ull) ^ (k.stride << 7) ^
       (static_cast<uint64_t>(k.box_inner) << 40) ^ (static_cast<uint64_t>(k.box_outer) << 52) ^
       (static_cast<uint64_t>(k.dtype) << 3) ^ static_cast<uint64_t>(k.swizzle);
  h ^= h >> 29;
  return static_cast<size_t>(h) & (kTmapSlots - 1);
}

// 2-D row-major tensor [outer, inner] of 16-bit elements (dtype 0 / 1) or fp32 (kTmapF32), row stride in elements.
constexpr int kTmapF32 = 2;
int make_tmap(CUtensorMap* tm, const void* ptr, int dtype, uint64_t inner, uint64_t outer, uint64_t row_stride_elems,
              uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle sw, const char* what) {
  const TmapKey key{ptr, inner, outer, row_stride_elems, box_inner, box_outer, dtype, static_cast<int32_t>(sw)};
  if (!g_tmap_cache) g_tmap_cache.reset(new TmapSlot[kTmapSlots]);
  TmapSlot& slot = g_tmap_cache[tmap_hash(key)];
  if (slot.valid && slot.key == key) {
    *tm = slot.map;
    return 0;
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) return fail("cuTensorMapEncodeTiled not available from the driver");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15u) != 0) return fail("%s: pointer not 16-byte aligned", what);
  const uint64_t esz = dtype == kTmapF32 ? 4 : 2;
  if ((row_stride_elems * esz) % 16 != 0) return fail("%s: row stride not a multiple of 16 bytes", what);
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {row_stride_elems * esz};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapDataType dt = dtype == kTmapF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                 : (dtype == kDtypeBf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
  CUresult r = fn(tm, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail("cuTensorMapEncodeTiled(%s) failed with CUresult %d", what, (int)r);
  slot.key = key;
  slot.map = *tm;
  slot.valid = true;
  return 0;
}

CUtensorMapSwizzle swizzle_for_row_bytes(int bytes) {
  return bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : (bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
}

// Profiling hook (lora_debug_set_phase_mask): which kernels a call launches.
enum : int { kPhasePack = 1, kPhaseRowProj = 2, kPhaseGemm = 4, kPhaseColProj = 8, kPhaseAll = 15 };
std::atomic<int> g_phase_mask{kPhaseAll};
inline bool phase_on(int ph) { return (g_phase_mask.load(std::memory_order_relaxed) & ph) != 0; }

// Launch with programmatic stream serialization (see common.cuh: pdl_wait / pdl_launch_dependents) and an
// optional cluster width.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster,
                       Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[n].val.programmaticStreamSerializationAllowed = 1;
  ++n;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

constexpr int kMaxDevices = 64;
struct DeviceInfo {
  std::atomic<int> ready{0};
  int sm_count = 0;
};
DeviceInfo g_dev[kMaxDevices];

template <typename K>
cudaError_t opt_in_smem(K kernel, int bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

// Per-device one-time set-up (SM count, >48 KiB dynamic smem opt-in).  Safe to race.
int device_setup(int* sm_count) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail("device ordinal %d out of range", dev);
  DeviceInfo& d = g_dev[dev];
  if (!d.ready.load(std::memory_order_acquire)) {
    int major = 0, sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major != 10) return fail("this library contains sm_100a code only; device %d is compute capability %d.x", dev, major);
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<false, false>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<true, false>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<false, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<true, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<true, false, false, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<true, true, false, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<false, false, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_lora_kernel<false, true, true>, gemm::SMEM_BYTES));
    CUDA_TRY(opt_in_smem(gemm::gemm_masked_kernel, gemm::SMEM_BYTES));
#define OPT_IN_SKINNY(BF, RP)                                                                                       \
  CUDA_TRY(opt_in_smem(skinny::rowproj_kernel<BF, RP, false, 2, 4>, skinny::rp_smem_optin<RP, 4>()));               \
  CUDA_TRY(opt_in_smem(skinny::rowproj_small_kernel<BF, RP, false, 2>, skinny::RP_ONE_PER_SM_BYTES));
#define OPT_IN_DIRECT(BF, NBE)                                                                                      \
  CUDA_TRY(opt_in_smem(skinny::rowproj_kernel<BF, 16, true, NBE, 4>, skinny::rp_smem_optin<16, 4>()));              \
  CUDA_TRY(opt_in_smem(skinny::rowproj_small_kernel<BF, 16, true, NBE>, skinny::RP_ONE_PER_SM_BYTES));
    CUDA_TRY(opt_in_smem(skinny::colproj_umma_kernel<16>, skinny::cu_smem_bytes<16>()));
    CUDA_TRY(opt_in_smem(skinny::colproj_umma_kernel<32>, skinny::cu_smem_bytes<32>()));
    CUDA_TRY(opt_in_smem(skinny::colproj_umma_kernel<64>, skinny::cu_smem_bytes<64>()));
    OPT_IN_DIRECT(true, 1) OPT_IN_DIRECT(false, 1) OPT_IN_DIRECT(true, 2) OPT_IN_DIRECT(false, 2)
    OPT_IN_SKINNY(true, 16) OPT_IN_SKINNY(true, 32) OPT_IN_SKINNY(true, 64)
    OPT_IN_SKINNY(false, 16) OPT_IN_SKINNY(false, 32) OPT_IN_SKINNY(false, 64)
#undef OPT_IN_SKINNY
#undef OPT_IN_DIRECT
    d.sm_count = sms;
    d.ready.store(1, std::memory_order_release);
  }
  *sm_count = d.sm_count;
  return 0;
}

int rpad_of(int r) { return r <= 16 ? 16 : (r <= 32 ? 32 : 64); }
size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

struct Split {
  int splits;
  int rows_per_split;
};
// Row splits of the column projections (colproj_sm100.cuh); sized for 148 SMs on purpose so that the workspace size
// does not depend on the device queried.  The kernel runs one CTA per SM and is HBM-bound from ~100 CTAs on (three
// tiles of loads in flight each), so a launch is planned as ONE wave, never two: with a single problem the most
// splits that fit; with two problems (dB and dA of one backward) the pair of split counts whose longest CTA is
// shortest, a tile of a dropout-masked problem counting kMaskedTileCost tiles of a plain one (measured: the mask pass
// shares the shared-memory bandwidth the tensor core reads with); ties go to the fewer CTAs.  More column tiles than
// SMs: one split each.
constexpr double kMaskedTileCost = 1.25;
int64_t colproj_tiles(int64_t C) { return (C + skinny::CU_COLS - 1) / skinny::CU_COLS; }
void colproj_plan(int64_t M, int njobs, const int64_t* tiles, const bool* masked, Split* out) {
  const int64_t m_tiles = (M + skinny::CP_BM - 1) / skinny::CP_BM;
  int64_t want[2] = {1, 1};
  if (njobs == 1) {
    want[0] = std::max<int64_t>(1, std::min<int64_t>(m_tiles, 148 / tiles[0]));
  } else if (tiles[0] + tiles[1] <= 148) {
    const int64_t smax = std::min<int64_t>(m_tiles, 32);
    double best = 1e30;
    int64_t best_ctas = 0;
    for (int64_t a = 1; a <= smax; ++a)
      for (int64_t b = 1; b <= smax; ++b) {
        const int64_t ctas = tiles[0] * a + tiles[1] * b;
        if (ctas > 148) break;
        const double ca = static_cast<double>((m_tiles + a - 1) / a) * (masked[0] ? kMaskedTileCost : 1.0);
        const double cb = static_cast<double>((m_tiles + b - 1) / b) * (masked[1] ? kMaskedTileCost : 1.0);
        const double cost = std::max(ca, cb);
        if (cost < best - 1e-9 || (cost < best + 1e-9 && ctas < best_ctas)) {
          best = cost;
          best_ctas = ctas;
          want[0] = a;
          want[1] = b;
        }
      }
  }
  for (int i = 0; i < njobs; ++i) {
    const int64_t tiles_per = (m_tiles + want[i] - 1) / want[i];
    out[i].rows_per_split = static_cast<int>(tiles_per * skinny::CP_BM);
    out[i].splits = static_cast<int>((M + out[i].rows_per_split - 1) / out[i].rows_per_split);
  }
}

constexpr int kMaxColTiles = 4096;  // arrival counters per column projection (C <= 4096*64 columns)
// Packed adapter operands (written by the forward or by lora_linear_pack, read by the backward):
//   a16 [rp,K]   a16t [K,rp]   sb16 [N,rp] (= s*B)   sb16t [rp,N]      16-bit, rank padded to rp
struct Packed {
  size_t a16, a16t, sb16, sb16t, total;
};
Packed carve_packed(int64_t K, int64_t N, int r) {
  const int rp = rpad_of(r);
  Packed w;
  size_t off = 0;
  w.a16 = off;   off += align256(static_cast<size_t>(rp) * K * 2);
  w.a16t = off;  off += align256(static_cast<size_t>(rp) * K * 2);
  w.sb16 = off;  off += align256(static_cast<size_t>(rp) * N * 2);
  w.sb16t = off; off += align256(static_cast<size_t>(rp) * N * 2);
  w.total = off;
  return w;
}
// Scratch of the backward: dH, the column projections' arrival counters and split partial sums.
struct Workspace {
  size_t dh, counters, partial, partial2, total;
};
Workspace carve(int64_t M, int64_t K, int64_t N, int r) {
  const int rp = rpad_of(r);
  Workspace w;
  size_t off = 0;
  w.dh = off;    off += align256(static_cast<size_t>(M) * rp * 2);
  w.counters = off; off += align256(2 * kMaxColTiles * sizeof(unsigned int));
  // the backward's launch: problem 0 = dB over N columns, problem 1 = dA over K columns (masked when p > 0);
  // sized for the larger split count of the two cases so that the size does not depend on p
  const int64_t tiles[2] = {colproj_tiles(N), colproj_tiles(K)};
  const bool with_mask[2] = {false, true}, without[2] = {false, false};
  Split pa[2], pb[2];
  colproj_plan(M, 2, tiles, with_mask, pa);
  colproj_plan(M, 2, tiles, without, pb);
  const size_t pn = static_cast<size_t>(std::max(pa[0].splits, pb[0].splits)) * N * rp * 4;
  const size_t pk = static_cast<size_t>(std::max(pa[1].splits, pb[1].splits)) * K * rp * 4;
  w.partial = off;  off += align256(pn);   // dB partial sums
  w.partial2 = off; off += align256(pk);   // dA partial sums (both column projections share one launch)
  w.total = off;
  return w;
}

int check_common(int64_t M, int64_t K, int64_t N, int r, int dtype) {
  if (M <= 0 || K <= 0 || N <= 0) return fail("M, K, N must be positive (got %lld, %lld, %lld)", (long long)M, (long long)K, (long long)N);
  if (M > INT32_MAX || K > INT32_MAX || N > INT32_MAX) return fail("dimension exceeds 2^31-1");
  if (K % 8 != 0 || N % 8 != 0) return fail("K and N must be multiples of 8 (16-byte rows for TMA); got K=%lld N=%lld", (long long)K, (long long)N);
  if (r < 1 || r > 64) return fail("rank r must be in [1, 64]; got %d", r);
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  return 0;
}

int launch_pack(const float* A, const float* B, uint8_t* pk, const Packed& w, int64_t K, int64_t N, int r, float s,
                int dtype, cudaStream_t stream) {
  const int rp = rpad_of(r);
  const long long total = static_cast<long long>(rp) * (K + N);
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 2048));
  if (dtype == kDtypeBf16)
    CUDA_TRY(launch_pdl(skinny::pack_adapters_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, stream, 1, A, B,
                        (__nv_bfloat16*)(pk + w.a16), (__nv_bfloat16*)(pk + w.a16t), (__nv_bfloat16*)(pk + w.sb16),
                        (__nv_bfloat16*)(pk + w.sb16t), (int)K, (int)N, r, rp, s));
  else
    CUDA_TRY(launch_pdl(skinny::pack_adapters_kernel<__half>, dim3(grid), dim3(256), 0, stream, 1, A, B,
                        (__half*)(pk + w.a16), (__half*)(pk + w.a16t), (__half*)(pk + w.sb16), (__half*)(pk + w.sb16t),
                        (int)K, (int)N, r, rp, s));
  CUDA_TRY(cudaGetLastError());
  return 0;
}

template <bool BF, int RP, bool DIRECT, int NBE, int SLABS>
int launch_rowproj_s(const CUtensorMap& tmAct, const CUtensorMap& tmP, const skinny::RowProjParams& p,
                     cudaStream_t stream) {
  const int grid = (p.M + skinny::rp_rows(SLABS) - 1) / skinny::rp_rows(SLABS);
  // Fewer CTAs than SMs: ask for more than half an SM's shared memory so that every CTA gets an SM of its own.  The
  // hardware places small CTAs several to an SM before it moves on, and co-resident CTAs of this kernel slow one
  // another (measured, H with dropout at K = 2560: M = 8192 as 128 four-slab CTAs 35.9 us packed, 24.0 us one per SM;
  // M = 4096 as 128 two-slab CTAs 29.9 / 18.8 us).
  size_t smem = skinny::rp_smem_bytes<RP, SLABS>();
  if (grid <= 148) smem = std::max<size_t>(smem, skinny::RP_ONE_PER_SM_BYTES);
  CUDA_TRY(launch_pdl(skinny::rowproj_kernel<BF, RP, DIRECT, NBE, SLABS>, dim3(grid), dim3(skinny::rp_threads(RP, SLABS)),
                      smem, stream, 1, tmAct, tmP, p));
  return 0;
}
// Few rows (rowproj_small.cuh): one 16-row slab per CTA, eight warps splitting the columns.  At most 296 CTAs; the
// shared-memory request keeps them to one (<= 148 CTAs) or two per SM.
constexpr int64_t kRowprojSmallMaxM = 296 * 16;
template <bool BF, int RP, bool DIRECT, int NBE>
int launch_rowproj_small(const CUtensorMap& tmAct, const CUtensorMap& tmP, const skinny::RowProjParams& p,
                         cudaStream_t stream) {
  const int grid = (p.M + 15) / 16;
  size_t smem = skinny::rs_smem_bytes<RP>();
  smem = std::max<size_t>(smem, grid <= 148 ? skinny::RP_ONE_PER_SM_BYTES : 80 * 1024);
  CUDA_TRY(launch_pdl(skinny::rowproj_small_kernel<BF, RP, DIRECT, NBE>, dim3(grid), dim3(skinny::RS_THREADS), smem, stream, 1,
                      tmAct, tmP, p));
  return 0;
}
// The tensor map of the activation was built for the same box (rowproj_box_rows).
template <bool BF, int RP, bool DIRECT = false, int NBE = 2>
int launch_rowproj_t(const CUtensorMap& tmAct, const CUtensorMap& tmP, const skinny::RowProjParams& p,
                     cudaStream_t stream) {
  if (p.M <= kRowprojSmallMaxM) return launch_rowproj_small<BF, RP, DIRECT, NBE>(tmAct, tmP, p, stream);
  return launch_rowproj_s<BF, RP, DIRECT, NBE, 4>(tmAct, tmP, p, stream);
}
int rowproj_box_rows(int64_t M) { return M <= kRowprojSmallMaxM ? 16 : skinny::rp_rows(4); }

// Forward with padded rank 16 and the fp32 masters at hand: the row projection reads A directly and packs A, B
// for the kernels after it (no separate pack launch).
struct DirectPack {
  const float* A = nullptr;
  const float* B = nullptr;
  int r = 0;
  int64_t N = 0;
  float s = 0.f;
  uint8_t* pk = nullptr;
  const Packed* w = nullptr;
};

// out[M, rp] = out_scale * (keep .* act)[M, C] * P[rp, C]^T; with p > 0 the keep-bits are drawn here and stored.
// zero_words: optional block of words the kernel clears (arrival counters of the column projections).
int launch_rowproj(const void* act, const void* P, void* out, uint8_t* keep_out, int64_t M, int64_t C, int rp, int dtype,
                   float out_scale, float p, uint64_t seed, uint64_t offset, cudaStream_t stream,
                   const uint64_t* offset_dev = nullptr, unsigned int* zero_words = nullptr, int n_zero_words = 0,
                   const DirectPack* direct = nullptr) {
  CUtensorMap tmAct, tmP;
  if (make_tmap(&tmAct, act, dtype, C, M, C, skinny::RP_BK, rowproj_box_rows(M), CU_TENSOR_MAP_SWIZZLE_128B,
                "rowproj act"))
    return 1;
  if (direct) {  // the fp32 master A [r, C]: boxes of 32 columns (one 128-byte swizzle row) x 8 or 16 rank rows
    if (make_tmap(&tmP, direct->A, kTmapF32, C, direct->r, C, 32, direct->r <= 8 ? 8 : 16, CU_TENSOR_MAP_SWIZZLE_128B, "rowproj A (fp32)"))
      return 1;
  } else if (make_tmap(&tmP, P, dtype, C, rp, C, skinny::RP_BK, rp, CU_TENSOR_MAP_SWIZZLE_128B, "rowproj P")) {
    return 1;
  }
  skinny::RowProjParams rp_p{};
  rp_p.out = out;
  rp_p.keep_out = keep_out;
  rp_p.M = (int)M;
  rp_p.C = (int)C;
  rp_p.out_scale = out_scale;
  rp_p.dropout = p > 0.f ? 1 : 0;
  rp_p.seed = seed;
  rp_p.rk = philox_round_keys(seed);
  rp_p.offset = offset;
  rp_p.offset_dev = reinterpret_cast<const unsigned long long*>(offset_dev);
  rp_p.thresh14 = static_cast<uint32_t>(p * 16384.0f + 0.5f);
  rp_p.zero_words = zero_words;
  rp_p.n_zero_words = n_zero_words;
  const bool bf = dtype == kDtypeBf16;
  if (direct) {
    rp_p.A32 = direct->A;
    rp_p.B32 = direct->B;
    rp_p.r = direct->r;
    rp_p.N = (int)direct->N;
    rp_p.s = direct->s;
    rp_p.a16 = direct->pk + direct->w->a16;
    rp_p.a16t = direct->pk + direct->w->a16t;
    rp_p.sb16 = direct->pk + direct->w->sb16;
    rp_p.sb16t = direct->pk + direct->w->sb16t;
    if (direct->r <= 8)
      return bf ? launch_rowproj_t<true, 16, true, 1>(tmAct, tmP, rp_p, stream)
                : launch_rowproj_t<false, 16, true, 1>(tmAct, tmP, rp_p, stream);
    return bf ? launch_rowproj_t<true, 16, true, 2>(tmAct, tmP, rp_p, stream)
              : launch_rowproj_t<false, 16, true, 2>(tmAct, tmP, rp_p, stream);
  }
#define RP_CASE(BF, RP) return launch_rowproj_t<BF, RP>(tmAct, tmP, rp_p, stream);
  if (bf) {
    if (rp == 16) RP_CASE(true, 16) else if (rp == 32) RP_CASE(true, 32) else RP_CASE(true, 64)
  } else {
    if (rp == 16) RP_CASE(false, 16) else if (rp == 32) RP_CASE(false, 32) else RP_CASE(false, 64)
  }
#undef RP_CASE
}

// One column projection: out[c*ld_c + j*ld_j] (=|+=) scale * sum_m keep(m,c) act[m,c] small[m,j],  c < C, j < r
struct ColProjJob {
  const void* act;
  const void* small;
  const uint8_t* keep;
  float* partial;
  unsigned int* counters;
  float* out;
  int64_t C;
  long long ld_c, ld_j;
  float scale;
};

template <int RP>
int launch_colproj_umma_t(const CUtensorMap (&tmAct)[2], const CUtensorMap (&tmSmall)[2], const skinny::ColProjParams& pp,
                          int njobs, int dtype, cudaStream_t stream) {
  int gx = 0, gy = 0;
  for (int i = 0; i < njobs; ++i) {
    gx = std::max(gx, pp.prob[i].c_tiles);
    gy = std::max(gy, pp.prob[i].splits);
  }
  CUDA_TRY(launch_pdl(skinny::colproj_umma_kernel<RP>, dim3(gx, gy, njobs), dim3(skinny::CU_THREADS),
                      skinny::cu_smem_bytes<RP>(), stream, 1, tmAct[0], tmSmall[0], tmAct[1], tmSmall[1], pp, dtype));
  return 0;
}

// Up to two column projections over the same M rows in ONE launch; each folds its own row splits
// (last CTA per column tile), so no reduction kernel follows.
int launch_colproj(const ColProjJob* jobs, int njobs, int64_t M, int r, int rp, int dtype, int accumulate,
                   cudaStream_t stream, const Split* given_plan = nullptr) {
  CUtensorMap tmAct[2], tmSmall[2];
  skinny::ColProjParams pp{};
  int64_t tiles[2] = {0, 0};
  bool masked[2] = {false, false};
  for (int i = 0; i < njobs; ++i) {
    tiles[i] = colproj_tiles(jobs[i].C);
    masked[i] = jobs[i].keep != nullptr;
  }
  Split plan[2];
  if (given_plan) plan[0] = plan[1] = *given_plan;  // single-problem callers that sized their scratch themselves
  else colproj_plan(M, njobs, tiles, masked, plan);
  for (int i = 0; i < 2; ++i) {
    const ColProjJob& j = jobs[i < njobs ? i : 0];
    const Split& sp = plan[i < njobs ? i : 0];
    if (make_tmap(&tmAct[i], j.act, dtype, j.C, M, j.C, 64, skinny::CP_BM, CU_TENSOR_MAP_SWIZZLE_128B, "colproj act")) return 1;
    // the Small tile's rows are r_pad elements long: swizzle span = row bytes
    const CUtensorMapSwizzle ssw = rp == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (rp == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    if (make_tmap(&tmSmall[i], j.small, dtype, rp, M, rp, rp, skinny::CP_BM, ssw, "colproj small")) return 1;
    skinny::ColProjProblem& q = pp.prob[i];
    q.partial = j.partial;
    q.counters = j.counters;
    q.keep = j.keep;
    q.out = j.out;
    q.ld_c = j.ld_c;
    q.ld_j = j.ld_j;
    q.scale = j.scale;
    q.accumulate = accumulate;
    q.r = r;
    q.M = (int)M;
    q.C = (int)j.C;
    q.rows_per_split = sp.rows_per_split;
    q.splits = sp.splits;
    q.c_tiles = static_cast<int>(colproj_tiles(j.C));
    if (q.c_tiles > kMaxColTiles) return fail("column projection over %lld columns exceeds the %d-tile counter table", (long long)j.C, kMaxColTiles);
  }
  return rp == 16 ? launch_colproj_umma_t<16>(tmAct, tmSmall, pp, njobs, dtype, stream)
                  : (rp == 32 ? launch_colproj_umma_t<32>(tmAct, tmSmall, pp, njobs, dtype, stream)
                              : launch_colproj_umma_t<64>(tmAct, tmSmall, pp, njobs, dtype, stream));
}

struct GemmArgs {
  const void* A1;  // [M, K1]
  const void* B1;  // [N, K1] or [K1, N]
  int64_t lda1 = 0;  // row stride of A1 in elements (0: K1)
  int64_t ldd = 0;   // row stride of D in elements (0: N)
  int b_mn_major;
  const void* A2;  // [M, rp] or null
  const void* B2;  // [N, rp] or null
  int rp;
  void* D;  // [M, N]
  const void* bias;
  int masked_r;
  const void* dh;
  const void* a16;
  const void* a16_rows;
  const uint8_t* keep;
  float inv_keep;
  int act = 0;          // gemm::kActRelu / kActGeluNew: activation epilogue (frozen neighbours only)
  void* Z = nullptr;    // pre-activation output [M, N] (optional, act != 0)
  int64_t ldz = 0;
  const void* residual = nullptr;  // [M, ldr] added in the plain epilogue (may be D itself)
  int64_t ldr = 0;
};

// LORA_B200_GEMM=1cta forces the single-CTA tile shape (A/B testing); default: CTA pairs whenever
// the problem has more than one 128-row block.
bool pair_mode_allowed() {
  static const bool allowed = [] {
    const char* e = getenv("LORA_B200_GEMM");
    return !(e && strcmp(e, "1cta") == 0);
  }();
  return allowed;
}

template <bool kMasked, bool kPair, bool kAct = false, bool kFma = false>
int launch_gemm_t(const CUtensorMap& tmA1, const CUtensorMap& tmB1, const CUtensorMap& tmA2, const CUtensorMap& tmB2,
                  const CUtensorMap& tmD, const CUtensorMap& tmZ, const gemm::Params& p, int grid, cudaStream_t stream) {
  CUDA_TRY(launch_pdl(gemm::gemm_lora_kernel<kMasked, kPair, kAct, kFma>, dim3(grid), dim3(gemm::THREADS), gemm::SMEM_BYTES,
                      stream, kPair ? 2 : 1, tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p));
  return 0;
}

// Backward with dropout, pair mode: 256 x 128 tiles with the rank-r product in its own TMEM accumulator
// (gemm_masked_sm100.cuh).
// term_only: no main loop, D = keep/(1-p) .* (dH A) (first pass of the two-pass form below).
int launch_gemm_masked_pair(const GemmArgs& g, int64_t M, int64_t N, int64_t K1, int dtype, int sm_count, cudaStream_t stream,
                            bool term_only = false) {
  const int rp = rpad_of(g.masked_r);
  CUtensorMap tmA1, tmB1, tmA2, tmB2, tmD;
  if (make_tmap(&tmA1, g.A1, dtype, K1, M, g.lda1 ? g.lda1 : K1, gemm::BK, gemm::BM, CU_TENSOR_MAP_SWIZZLE_128B, "masked gemm A1")) return 1;
  if (!g.b_mn_major) {
    if (make_tmap(&tmB1, g.B1, dtype, K1, N, K1, gemm::BK, gemm::M_B_ROWS, CU_TENSOR_MAP_SWIZZLE_128B, "masked gemm B1 (K-major)")) return 1;
  } else {
    if (make_tmap(&tmB1, g.B1, dtype, N, K1, N, 64, gemm::BK, CU_TENSOR_MAP_SWIZZLE_128B, "masked gemm B1 (MN-major)")) return 1;
  }
  const CUtensorMapSwizzle sw = swizzle_for_row_bytes(rp * 2);
  if (make_tmap(&tmA2, g.dh, dtype, rp, M, rp, rp, gemm::BM, sw, "masked gemm dH")) return 1;
  if (make_tmap(&tmB2, g.a16, dtype, rp, N, rp, rp, gemm::M_B_ROWS, sw, "masked gemm A^T")) return 1;
  if (make_tmap(&tmD, g.D, dtype, N, M, g.ldd ? g.ldd : N, 64, 32, CU_TENSOR_MAP_SWIZZLE_128B, "masked gemm D")) return 1;
  gemm::MaskedParams p;
  p.M = (int)M;
  p.N = (int)N;
  p.K1 = term_only ? 0 : (int)K1;
  p.r_pad = rp;
  p.b_mn_major = g.b_mn_major;
  p.dtype = dtype;
  p.keep = g.keep;
  p.inv_keep = g.inv_keep;
  const long long tiles = ((M + 2 * gemm::BM - 1) / (2 * gemm::BM)) * ((N + gemm::MBN - 1) / gemm::MBN);
  const int grid = 2 * static_cast<int>(std::min<long long>(tiles, sm_count / 2));
  CUDA_TRY(launch_pdl(gemm::gemm_masked_kernel, dim3(grid), dim3(gemm::THREADS), gemm::SMEM_BYTES, stream, 2, tmA1, tmB1, tmA2,
                      tmB2, tmD, p));
  return 0;
}

// Which masked flavour a pair-mode backward takes.  Measured (M = 16384, r02): the 128-column TMEM flavour costs the
// same ~194 us at 2560 -> 2560 whatever the rank (its main loop runs at ~3/4 of the 256-column kernel's rate), the
// epilogue flavour 165 / 180 / 288 us at r = 8 / 16 / 64; with a short contraction the epilogue flavour is bound by
// its epilogue (dX of 3072 -> 768 over K1 = 768: 86 us against 64 us).  So: wide ranks and short contractions go
// to the TMEM flavour.  LORA_B200_MASKED=epilogue|tmem forces one for A/B timing; both are exact implementations
// of the same formula.
bool use_masked_tmem(int rp, int64_t K1) {
  static const int forced = [] {
    const char* e = getenv("LORA_B200_MASKED");
    return !e ? 0 : (strcmp(e, "epilogue") == 0 ? 1 : (strcmp(e, "tmem") == 0 ? 2 : 0));
  }();
  if (forced) return forced == 2;
  return rp >= 32 || K1 <= 1024;
}

// Wide ranks over a long contraction: the TMEM flavour's 128-column main loop costs more than a second pass over
// the output.  Then the masked rank-r term is written first (TMEM kernel, K1 = 0) and the 256-column kernel adds
// dY W on top of it through its residual epilogue.  Measured, r = 64, M = 16384, both passes together against the
// single launch: dX of 2560 -> 10240 (K1 = 10240) 638 against 822 us, 2560 -> 2560 184 / 193, 10240 -> 2560 739 / 747,
// 768 -> 2304 63 / 62; forced on at K1 = 768 (3072 -> 768) 129 against 85, hence the threshold.
// LORA_B200_MASKED2=off|on forces the choice for A/B timing.
bool masked_two_pass(int rp, int64_t K1) {
  static const int forced = [] {
    const char* e = getenv("LORA_B200_MASKED2");
    return !e ? 0 : (strcmp(e, "off") == 0 ? 1 : (strcmp(e, "on") == 0 ? 2 : 0));
  }();
  if (forced) return forced == 2;
  return rp >= 32 && K1 >= 2048;
}

int launch_gemm(const GemmArgs& g, int64_t M, int64_t N, int64_t K1, int dtype, int sm_count, cudaStream_t stream) {
  const bool pair = pair_mode_allowed() && M > gemm::BM && sm_count >= 2;
  if (pair && g.masked_r > 0 && !g.bias && !g.act && use_masked_tmem(rpad_of(g.masked_r), K1)) {
    if (!masked_two_pass(rpad_of(g.masked_r), K1)) return launch_gemm_masked_pair(g, M, N, K1, dtype, sm_count, stream);
    // pass 1: D = keep/(1-p) .* (dH A) (the TMEM kernel without a main loop); pass 2: D = A1 B1 + D, 256-column tiles
    if (launch_gemm_masked_pair(g, M, N, K1, dtype, sm_count, stream, true)) return 1;
    GemmArgs g2{};
    g2.A1 = g.A1;
    g2.B1 = g.B1;
    g2.lda1 = g.lda1;
    g2.ldd = g.ldd;
    g2.b_mn_major = g.b_mn_major;
    g2.D = g.D;
    g2.inv_keep = 1.f;
    g2.residual = g.D;
    g2.ldr = g.ldd ? g.ldd : N;
    return launch_gemm(g2, M, N, K1, dtype, sm_count, stream);
  }
  const uint32_t b_rows = pair ? gemm::BN / 2 : gemm::BN;  // rows of the B tile one CTA stages
  CUtensorMap tmA1, tmB1, tmA2, tmB2, tmD, tmZ;
  if (make_tmap(&tmA1, g.A1, dtype, K1, M, g.lda1 ? g.lda1 : K1, gemm::BK, gemm::BM, CU_TENSOR_MAP_SWIZZLE_128B, "gemm A1")) return 1;
  if (!g.b_mn_major) {
    if (make_tmap(&tmB1, g.B1, dtype, K1, N, K1, gemm::BK, b_rows, CU_TENSOR_MAP_SWIZZLE_128B, "gemm B1 (K-major)")) return 1;
  } else {
    if (make_tmap(&tmB1, g.B1, dtype, N, K1, N, 64, gemm::BK, CU_TENSOR_MAP_SWIZZLE_128B, "gemm B1 (MN-major)")) return 1;
  }
  if (g.rp > 0) {
    const CUtensorMapSwizzle sw = swizzle_for_row_bytes(g.rp * 2);
    if (make_tmap(&tmA2, g.A2, dtype, g.rp, M, g.rp, g.rp, gemm::BM, sw, "gemm A2")) return 1;
    if (make_tmap(&tmB2, g.B2, dtype, g.rp, N, g.rp, g.rp, b_rows, sw, "gemm B2")) return 1;
  } else {
    tmA2 = tmA1;
    tmB2 = tmB1;
  }
  if (make_tmap(&tmD, g.D, dtype, N, M, g.ldd ? g.ldd : N, 64, 32, CU_TENSOR_MAP_SWIZZLE_128B, "gemm D")) return 1;
  tmZ = tmD;
  if (g.Z && make_tmap(&tmZ, g.Z, dtype, N, M, g.ldz ? g.ldz : N, 64, 32, CU_TENSOR_MAP_SWIZZLE_128B, "gemm Z")) return 1;

  gemm::Params p;
  p.M = (int)M;
  p.N = (int)N;
  p.K1 = (int)K1;
  p.r_pad = g.rp;
  p.b_mn_major = g.b_mn_major;
  p.dtype = dtype;
  p.bias = g.bias;
  p.masked_r = g.masked_r;
  p.dh = g.dh;
  p.dh_ld = g.rp > 0 ? g.rp : rpad_of(g.masked_r > 0 ? g.masked_r : 1);
  p.a16 = g.a16;
  p.a16_rows = g.a16_rows;
  p.keep = g.keep;
  p.inv_keep = g.inv_keep;
  p.act = g.act;
  p.store_z = g.Z != nullptr;
  p.residual = g.residual;
  p.res_ld = g.ldr;
  if (g.residual && (g.masked_r > 0 || g.act)) return fail("internal: the residual epilogue belongs to the plain kernel");
  if (g.residual && ((g.ldr & 7) || (reinterpret_cast<uintptr_t>(g.residual) & 15)))
    return fail("residual rows must be 16-byte aligned (row stride a multiple of 8 elements)");

  const int tile_m = pair ? 2 * gemm::BM : gemm::BM;
  const long long tiles = ((M + tile_m - 1) / tile_m) * ((N + gemm::BN - 1) / gemm::BN);
  const bool masked = p.masked_r > 0;
  if (masked && p.act) return fail("internal: activation epilogue with the masked kernel");
#ifdef LORA_B200_DEBUG
  // Timing experiments only (results are wrong; never compiled into the shipped library):
  // LORA_B200_DEBUG_MASKED_R=j runs the masked kernel with the rank-r loop cut to j iterations.
  static const int dbg_r = [] {
    const char* e = std::getenv("LORA_B200_DEBUG_MASKED_R");
    return e ? std::atoi(e) : -1;
  }();
  if (masked && dbg_r >= 0) p.masked_r = dbg_r;
#endif
  if (pair) {
    const int grid = 2 * static_cast<int>(std::min<long long>(tiles, sm_count / 2));
    if (p.act) return launch_gemm_t<false, true, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
    if (masked && p.dh_ld == 16) return launch_gemm_t<true, true, false, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
    return masked ? launch_gemm_t<true, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream)
                  : launch_gemm_t<false, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
  }
  const int grid = static_cast<int>(std::min<long long>(tiles, sm_count));
  if (p.act) return launch_gemm_t<false, false, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
  if (masked && p.dh_ld == 16) return launch_gemm_t<true, false, false, true>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
  return masked ? launch_gemm_t<true, false>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream)
                : launch_gemm_t<false, false>(tmA1, tmB1, tmA2, tmB2, tmD, tmZ, p, grid, stream);
}

// ---------------------------------------------------------------------------------------------
// Scratch for lora_linear_grad_b_static(): a fixed block of device memory that is part of the
// library image itself, so the entry point never has to ask the driver for memory.  Every GPU
// that loads the library carries these LORA_B200_STATIC_SCRATCH_BYTES until the process exits.
// In host code `g_static_scratch` is only a handle for the symbol, not a device address: the
// real address is looked up once per device with cudaGetSymbolAddress (static_scratch_ptr below)
// and handed to the kernels as an ordinary pointer.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[LORA_B200_STATIC_SCRATCH_BYTES];

struct StaticScratchSlot {
  std::atomic<int> ready{0};
  void* ptr = nullptr;
};
StaticScratchSlot g_static_slot[kMaxDevices];

// Device address of g_static_scratch on the CURRENT device; resolved on the first call per device.
int static_scratch_ptr(void** out) {
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail("device ordinal %d out of range", dev);
  StaticScratchSlot& s = g_static_slot[dev];
  if (!s.ready.load(std::memory_order_acquire)) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail("cudaGetSymbolAddress(static scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    s.ptr = p;  // racing first calls resolve the same address; the store order below publishes it
    s.ready.store(1, std::memory_order_release);
  }
  *out = s.ptr;
  return 0;
}

__global__ void keep_mask_kernel(uint8_t* keep, int M, int keep_ld, uint32_t blocks_per_row, uint64_t seed,
                                 uint64_t offset, uint32_t thresh14) {
  const long long total = static_cast<long long>(M) * keep_ld;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint32_t m = static_cast<uint32_t>(i / keep_ld), c8 = static_cast<uint32_t>(i % keep_ld);
    keep[i] = static_cast<uint8_t>(lb::dropout_keep_byte(m, c8, blocks_per_row, seed, offset, thresh14));
  }
}

}  // namespace

extern "C" {

const char* lora_last_error(void) { return g_err.c_str(); }

int32_t lora_linear_rpad(int32_t r) { return rpad_of(r); }

size_t lora_linear_packed_bytes(int64_t K, int64_t N, int32_t r) {
  if (K <= 0 || N <= 0 || r < 1 || r > 64) return 0;
  return carve_packed(K, N, r).total;
}

size_t lora_linear_workspace_bytes(int64_t M, int64_t K, int64_t N, int32_t r) {
  if (M <= 0 || K <= 0 || N <= 0 || r < 1 || r > 64) return 0;
  return carve(M, K, N, r).total;
}

int lora_linear_pack(const float* A, const float* B, void* packed, size_t packed_bytes, int64_t K, int64_t N, int32_t r,
                     float scaling, int32_t dtype, cudaStream_t stream) {
  g_err.clear();
  if (check_common(1, K, N, r, dtype)) return 1;
  if (!A || !B || !packed) return fail("lora_linear_pack: null pointer argument");
  const Packed w = carve_packed(K, N, r);
  if (packed_bytes < w.total) return fail("packed-operand block too small: need %zu bytes, got %zu", w.total, packed_bytes);
  if (reinterpret_cast<uintptr_t>(packed) & 255u) return fail("lora_linear_pack: packed must be 256-byte aligned");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  return launch_pack(A, B, static_cast<uint8_t*>(packed), w, K, N, r, scaling, dtype, stream);
}

int lora_linear_fwd(const void* X, const void* W, const void* bias, const float* A, const float* B, void* Y, void* H,
                    uint8_t* keep_mask, int64_t M, int64_t K, int64_t N, int32_t r, float scaling, float p, uint64_t seed,
                    uint64_t offset, int32_t dtype, int32_t w_is_KN, void* packed, size_t packed_bytes,
                    cudaStream_t stream) {
  return lora_linear_fwd_devrng(X, W, bias, A, B, Y, H, keep_mask, M, K, N, r, scaling, p, seed, offset, nullptr, dtype,
                                w_is_KN, packed, packed_bytes, stream);
}

int lora_linear_fwd_devrng(const void* X, const void* W, const void* bias, const float* A, const float* B, void* Y, void* H,
                           uint8_t* keep_mask, int64_t M, int64_t K, int64_t N, int32_t r, float scaling, float p,
                           uint64_t seed, uint64_t offset, const uint64_t* offset_dev, int32_t dtype, int32_t w_is_KN,
                           void* packed, size_t packed_bytes, cudaStream_t stream) {
  g_err.clear();
  if (check_common(M, K, N, r, dtype)) return 1;
  if (!X || !W || !Y || !H || !packed) return fail("lora_linear_fwd: null pointer argument");
  if ((A == nullptr) != (B == nullptr)) return fail("lora_linear_fwd: pass both A and B, or neither (packed operands already valid)");
  if (!(p >= 0.f && p < 1.f)) return fail("dropout p must be in [0,1); got %f", (double)p);
  if (p > 0.f && !keep_mask) return fail("lora_linear_fwd: keep_mask is required when p > 0");
  const Packed w = carve_packed(K, N, r);
  if (packed_bytes < w.total) return fail("packed-operand block too small: need %zu bytes, got %zu", w.total, packed_bytes);
  if (reinterpret_cast<uintptr_t>(packed) & 255u) return fail("lora_linear_fwd: packed must be 256-byte aligned");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  uint8_t* pk = static_cast<uint8_t*>(packed);
  const int rp = rpad_of(r);
  const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;

  // Padded rank 16 (every adapter of the reference) with the masters at hand: the row projection reads A itself and
  // does the pack as a side job -- two launches per forward.  Wider ranks (or A = B = NULL) keep the pack launch.
  const bool direct = A != nullptr && rp == 16 && phase_on(kPhaseRowProj) && phase_on(kPhasePack) &&
                      (reinterpret_cast<uintptr_t>(A) & 15u) == 0;  // TMA reads the master: 16-byte aligned base
  if (A != nullptr && !direct && phase_on(kPhasePack) && launch_pack(A, B, pk, w, K, N, r, scaling, dtype, stream)) return 1;
  if (phase_on(kPhaseRowProj)) {
    DirectPack dp;
    dp.A = A;
    dp.B = B;
    dp.r = r;
    dp.N = N;
    dp.s = scaling;
    dp.pk = pk;
    dp.w = &w;
    if (launch_rowproj(X, pk + w.a16, H, keep_mask, M, K, rp, dtype, inv_keep, p, seed, offset, stream, offset_dev, nullptr, 0,
                       direct ? &dp : nullptr))
      return 1;
  }
  if (!phase_on(kPhaseGemm)) return 0;
  GemmArgs g{};
  g.A1 = X;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 1 : 0;
  g.A2 = H;
  g.B2 = pk + w.sb16;
  g.rp = rp;
  g.D = Y;
  g.bias = bias;
  g.inv_keep = 1.f;
  return launch_gemm(g, M, N, K, dtype, sms, stream);
}

int lora_linear_bwd(const void* dY, const void* X, const void* W, const void* packed, const void* H,
                    const uint8_t* keep_mask, void* dX, float* dA, float* dB, int64_t M, int64_t K, int64_t N,
                    int32_t r, float scaling, float p, int32_t dtype, int32_t w_is_KN, int32_t accumulate,
                    void* workspace, size_t workspace_bytes, cudaStream_t stream) {
  g_err.clear();
  if (check_common(M, K, N, r, dtype)) return 1;
  if (!dY || !X || !W || !packed || !H || !dA || !dB || !workspace) return fail("lora_linear_bwd: null pointer argument");
  if (!(p >= 0.f && p < 1.f)) return fail("dropout p must be in [0,1); got %f", (double)p);
  if (p > 0.f && !keep_mask) return fail("lora_linear_bwd: keep_mask is required when p > 0");
  if (reinterpret_cast<uintptr_t>(packed) & 255u) return fail("lora_linear_bwd: packed must be 256-byte aligned");
  const Packed pw = carve_packed(K, N, r);
  const Workspace w = carve(M, K, N, r);
  if (workspace_bytes < w.total) return fail("workspace too small: need %zu bytes, got %zu", w.total, workspace_bytes);
  int sms = 0;
  if (device_setup(&sms)) return 1;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  const uint8_t* pk = static_cast<const uint8_t*>(packed);
  const int rp = rpad_of(r);
  const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const uint8_t* keep = p > 0.f ? keep_mask : nullptr;
  void* dH = ws + w.dh;
  float* partial = reinterpret_cast<float*>(ws + w.partial);
  unsigned int* counters = reinterpret_cast<unsigned int*>(ws + w.counters);

  // dH[M, rp] = dY (sB); the same launch clears the arrival counters of the column projections below
  if (phase_on(kPhaseRowProj) &&
      launch_rowproj(dY, pk + pw.sb16t, dH, nullptr, M, N, rp, dtype, 1.f, 0.f, 0, 0, stream, nullptr, counters,
                     2 * kMaxColTiles))
    return 1;
  if (dX && phase_on(kPhaseGemm)) {
    GemmArgs g{};
    g.A1 = dY;
    g.B1 = W;
    g.b_mn_major = w_is_KN ? 0 : 1;  // contraction runs over N here
    g.D = dX;
    g.inv_keep = inv_keep;
    if (keep) {
      g.rp = 0;
      g.masked_r = r;
      g.dh = dH;
      g.a16 = pk + pw.a16t;      // [K, rp]: rank index contiguous, as the warp-MMA B-fragments want
      g.a16_rows = pk + pw.a16;  // [rp, K]: column index contiguous, for the staged fp32 slice (rp == 16)
      g.keep = keep;
    } else {
      g.A2 = dH;
      g.B2 = pk + pw.a16t;
      g.rp = rp;
    }
    if (launch_gemm(g, M, K, N, dtype, sms, stream)) return 1;
    // launch_gemm derives dh_ld from rp/masked_r; both give rp here.
  }
  if (!phase_on(kPhaseColProj)) return 0;
  // dB[N, r] = s * dY^T H (H already carries 1/(1-p))  and  dA[r, K] = dH^T (keep .* X) / (1-p), one launch
  ColProjJob jobs[2];
  jobs[0] = ColProjJob{dY, H, nullptr, partial, counters, dB, N, (long long)r, 1ll, scaling};
  jobs[1] = ColProjJob{X, dH, keep, reinterpret_cast<float*>(ws + w.partial2), counters + kMaxColTiles, dA, K, 1ll,
                       (long long)K, inv_keep};
  return launch_colproj(jobs, 2, M, r, rp, dtype, accumulate, stream);
}

int lora_linear_base(const void* X, const void* W, const void* bias, void* Y, int64_t M, int64_t K, int64_t N,
                     int32_t dtype, int32_t w_is_KN, cudaStream_t stream) {
  g_err.clear();
  if (check_common(M, K, N, 1, dtype)) return 1;
  if (!X || !W || !Y) return fail("lora_linear_base: null pointer argument");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  GemmArgs g{};
  g.A1 = X;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 1 : 0;
  g.D = Y;
  g.bias = bias;
  g.inv_keep = 1.f;
  return launch_gemm(g, M, N, K, dtype, sms, stream);
}

int lora_linear_grad_b_static(const void* dY, const void* H, float* dB, int64_t M, int64_t N, int32_t r,
                               float scaling, int32_t dtype, int32_t accumulate, cudaStream_t stream) {
  g_err.clear();
  if (M <= 0 || N <= 0 || M > INT32_MAX || N > INT32_MAX)
    return fail("lora_linear_grad_b_static: M and N must be positive and below 2^31 (got %lld, %lld)", (long long)M, (long long)N);
  if (N % 8 != 0) return fail("lora_linear_grad_b_static: N must be a multiple of 8 (16-byte rows of dY); got N=%lld", (long long)N);
  if (r < 1 || r > 64) return fail("lora_linear_grad_b_static: rank r must be in [1, 64]; got %d", r);
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("lora_linear_grad_b_static: dtype must be 0 (bf16) or 1 (fp16)");
  if (!dY || !H || !dB) return fail("lora_linear_grad_b_static: null pointer argument");
  const int rp = rpad_of(r);
  const int64_t tiles = colproj_tiles(N);
  const bool no_mask = false;
  Split sp;
  colproj_plan(M, 1, &tiles, &no_mask, &sp);
  const size_t counter_bytes = align256(kMaxColTiles * sizeof(unsigned int));
  // fewer row splits than planned when the planned number does not fit the built-in block (longer CTAs, same result
  // up to the order of the fp32 additions); refused only when a single split does not fit
  const size_t per_split = static_cast<size_t>(N) * rp * sizeof(float);
  const size_t fit = (static_cast<size_t>(LORA_B200_STATIC_SCRATCH_BYTES) - counter_bytes) / per_split;
  if (fit >= 1 && static_cast<size_t>(sp.splits) > fit) {
    const int64_t m_tiles = (M + skinny::CP_BM - 1) / skinny::CP_BM;
    const int64_t tiles_per = (m_tiles + static_cast<int64_t>(fit) - 1) / static_cast<int64_t>(fit);
    sp.rows_per_split = static_cast<int>(tiles_per * skinny::CP_BM);
    sp.splits = static_cast<int>((M + sp.rows_per_split - 1) / sp.rows_per_split);
  }
  const size_t need = counter_bytes + static_cast<size_t>(sp.splits) * per_split;
  if (need > static_cast<size_t>(LORA_B200_STATIC_SCRATCH_BYTES))
    return fail("lora_linear_grad_b_static: this shape needs %zu bytes of partial sums (M=%lld, N=%lld, padded rank %d) "
                "but the built-in scratch holds %d; use lora_linear_bwd with a caller-provided workspace instead",
                need, (long long)M, (long long)N, rp, (int)LORA_B200_STATIC_SCRATCH_BYTES);
  int sms = 0;
  if (device_setup(&sms)) return 1;
  void* scratch = nullptr;
  if (static_scratch_ptr(&scratch)) return 1;
  // What must start at zero are the arrival counters at the head of the block: they are zero when
  // the library is loaded and every call leaves them at zero again (each counter wraps after the
  // last row split of its column tile has arrived).  The partial sums behind them need no clearing:
  // every one that is read was written earlier in the same call.
  ColProjJob job{dY, H, nullptr, reinterpret_cast<float*>(static_cast<unsigned char*>(scratch) + counter_bytes),
                 static_cast<unsigned int*>(scratch), dB, N, (long long)r, 1ll, scaling};
  return launch_colproj(&job, 1, M, r, rp, dtype, accumulate, stream, &sp);
}

int lora_frozen_linear_fwd(const void* X, int64_t ldx, const void* W, const void* bias, void* Y, int64_t ldy, int64_t M,
                           int64_t K, int64_t N, int32_t dtype, int32_t w_is_KN, cudaStream_t stream) {
  g_err.clear();
  if (M <= 0 || K <= 0 || N <= 0 || M > INT32_MAX || K > INT32_MAX || N > INT32_MAX) return fail("lora_frozen_linear_fwd: bad shape");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if (!X || !W || !Y) return fail("lora_frozen_linear_fwd: null pointer argument");
  if (ldx < K || ldy < N || ldx % 8 || ldy % 8) return fail("lora_frozen_linear_fwd: ldx/ldy must be >= K/N and multiples of 8 elements");
  if ((w_is_KN ? N : K) % 8) return fail("lora_frozen_linear_fwd: the weight's row length (%s) must be a multiple of 8", w_is_KN ? "N" : "K");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  GemmArgs g{};
  g.A1 = X;
  g.lda1 = ldx;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 1 : 0;
  g.D = Y;
  g.ldd = ldy;
  g.bias = bias;
  g.inv_keep = 1.f;
  return launch_gemm(g, M, N, K, dtype, sms, stream);
}

int lora_frozen_linear_add_fwd(const void* X, int64_t ldx, const void* W, const void* bias, const void* R, int64_t ldr,
                               void* Y, int64_t ldy, int64_t M, int64_t K, int64_t N, int32_t dtype, int32_t w_is_KN,
                               cudaStream_t stream) {
  g_err.clear();
  if (M <= 0 || K <= 0 || N <= 0 || M > INT32_MAX || K > INT32_MAX || N > INT32_MAX) return fail("lora_frozen_linear_add_fwd: bad shape");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if (!X || !W || !Y || !R) return fail("lora_frozen_linear_add_fwd: null pointer argument");
  if (ldx < K || ldy < N || ldr < N || ldx % 8 || ldy % 8 || ldr % 8)
    return fail("lora_frozen_linear_add_fwd: ldx/ldr/ldy must be >= K/N/N and multiples of 8 elements");
  if ((w_is_KN ? N : K) % 8) return fail("lora_frozen_linear_add_fwd: the weight's row length (%s) must be a multiple of 8", w_is_KN ? "N" : "K");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  GemmArgs g{};
  g.A1 = X;
  g.lda1 = ldx;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 1 : 0;
  g.D = Y;
  g.ldd = ldy;
  g.bias = bias;
  g.inv_keep = 1.f;
  g.residual = R;
  g.ldr = ldr;
  return launch_gemm(g, M, N, K, dtype, sms, stream);
}

int lora_frozen_linear_act_fwd(const void* X, int64_t ldx, const void* W, const void* bias, void* Y, int64_t ldy, void* Z,
                               int64_t ldz, int64_t M, int64_t K, int64_t N, int32_t act, int32_t dtype, int32_t w_is_KN,
                               cudaStream_t stream) {
  g_err.clear();
  if (M <= 0 || K <= 0 || N <= 0 || M > INT32_MAX || K > INT32_MAX || N > INT32_MAX) return fail("lora_frozen_linear_act_fwd: bad shape");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if (!X || !W || !Y) return fail("lora_frozen_linear_act_fwd: null pointer argument");
  if (act != gemm::kActRelu && act != gemm::kActGeluNew) return fail("lora_frozen_linear_act_fwd: act must be 1 (relu) or 2 (gelu_new); got %d", act);
  if (ldx < K || ldy < N || ldx % 8 || ldy % 8 || (Z && (ldz < N || ldz % 8)))
    return fail("lora_frozen_linear_act_fwd: ldx/ldy/ldz must be >= K/N/N and multiples of 8 elements");
  if ((w_is_KN ? N : K) % 8) return fail("lora_frozen_linear_act_fwd: the weight's row length (%s) must be a multiple of 8", w_is_KN ? "N" : "K");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  GemmArgs g{};
  g.A1 = X;
  g.lda1 = ldx;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 1 : 0;
  g.D = Y;
  g.ldd = ldy;
  g.bias = bias;
  g.inv_keep = 1.f;
  g.act = act;
  g.Z = Z;
  g.ldz = ldz;
  return launch_gemm(g, M, N, K, dtype, sms, stream);
}

int lora_act_bwd(const void* dY, const void* ZorY, void* dZ, int64_t n_elems, int32_t act, int32_t dtype, cudaStream_t stream) {
  g_err.clear();
  if (!dY || !ZorY || !dZ) return fail("lora_act_bwd: null pointer argument");
  if (n_elems <= 0 || n_elems % 8) return fail("lora_act_bwd: element count must be a positive multiple of 8");
  if (act != gemm::kActRelu && act != gemm::kActGeluNew) return fail("lora_act_bwd: act must be 1 (relu) or 2 (gelu_new); got %d", act);
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if ((reinterpret_cast<uintptr_t>(dY) | reinterpret_cast<uintptr_t>(ZorY) | reinterpret_cast<uintptr_t>(dZ)) & 15)
    return fail("lora_act_bwd: pointers must be 16-byte aligned");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  const long long n8 = n_elems / 8;
  const int grid = static_cast<int>(std::min<long long>((n8 + act::THREADS - 1) / act::THREADS, 8LL * sms));
  CUDA_TRY(launch_pdl(act::act_bwd_kernel, dim3(grid), dim3(act::THREADS), 0, stream, 1, static_cast<const uint4*>(dY),
                      static_cast<const uint4*>(ZorY), static_cast<uint4*>(dZ), n8, (int)act, (int)dtype));
  return 0;
}

int lora_merge_weight(void* W, const float* A, const float* B, int64_t K, int64_t N, int32_t r, float scale, int32_t dtype,
                      int32_t w_is_KN, cudaStream_t stream) {
  g_err.clear();
  if (!W || !A || !B) return fail("lora_merge_weight: null pointer argument");
  if (K <= 0 || N <= 0 || r < 1 || r > 64) return fail("lora_merge_weight: need K, N > 0 and 1 <= r <= 64");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if ((w_is_KN ? N : K) % 8) return fail("lora_merge_weight: the weight's row length (%s) must be a multiple of 8", w_is_KN ? "N" : "K");
  if ((reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(A)) & 15) return fail("lora_merge_weight: W and A must be 16-byte aligned");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  const long long groups = static_cast<long long>(K) * N / 8;
  const int grid = static_cast<int>(std::max<long long>(1, std::min<long long>((groups + act::THREADS - 1) / act::THREADS, 16LL * sms)));
  act::merge_kernel<<<grid, act::THREADS, 0, stream>>>(static_cast<uint16_t*>(W), A, B, (long long)N, (long long)K, (int)r, scale,
                                                     (int)w_is_KN, (int)dtype);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int lora_frozen_linear_bwd(const void* dY, int64_t lddy, const void* W, void* dX, int64_t lddx, int64_t M, int64_t K,
                           int64_t N, int32_t dtype, int32_t w_is_KN, cudaStream_t stream) {
  g_err.clear();
  if (M <= 0 || K <= 0 || N <= 0 || M > INT32_MAX || K > INT32_MAX || N > INT32_MAX) return fail("lora_frozen_linear_bwd: bad shape");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  if (!dY || !W || !dX) return fail("lora_frozen_linear_bwd: null pointer argument");
  if (lddy < N || lddx < K || lddy % 8 || lddx % 8) return fail("lora_frozen_linear_bwd: lddy/lddx must be >= N/K and multiples of 8 elements");
  if ((w_is_KN ? N : K) % 8) return fail("lora_frozen_linear_bwd: the weight's row length (%s) must be a multiple of 8", w_is_KN ? "N" : "K");
  int sms = 0;
  if (device_setup(&sms)) return 1;
  GemmArgs g{};
  g.A1 = dY;
  g.lda1 = lddy;
  g.B1 = W;
  g.b_mn_major = w_is_KN ? 0 : 1;  // contraction runs over N
  g.D = dX;
  g.ldd = lddx;
  g.inv_keep = 1.f;
  return launch_gemm(g, M, K, N, dtype, sms, stream);
}

int lora_ce_fwd(const void* logits, int64_t ld, const int64_t* labels, float* lse, float* loss_rows, int64_t M, int64_t V,
                int64_t ignore_index, int32_t dtype, cudaStream_t stream) {
  g_err.clear();
  if (!logits || !labels || !lse || !loss_rows) return fail("lora_ce_fwd: null pointer argument");
  if (M <= 0 || V <= 0 || M > INT32_MAX || V > INT32_MAX || ld < V || ld % 8) return fail("lora_ce_fwd: need M, V > 0 and ld >= V, ld %% 8 == 0");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  CUDA_TRY(launch_pdl(ce::ce_fwd_kernel, dim3((unsigned)M), dim3(ce::THREADS), 0, stream, 1,
                      static_cast<const uint16_t*>(logits), (long long)ld, reinterpret_cast<const long long*>(labels), lse,
                      loss_rows, (int)V, (long long)ignore_index, (int)dtype));
  return 0;
}

int lora_ce_bwd(const void* logits, int64_t ld, const int64_t* labels, const float* lse, const float* grad_scale,
                void* dlogits, int64_t ldd, int64_t M, int64_t V, int64_t ignore_index, int32_t dtype, cudaStream_t stream) {
  g_err.clear();
  if (!logits || !labels || !lse || !grad_scale || !dlogits) return fail("lora_ce_bwd: null pointer argument");
  if (M <= 0 || V <= 0 || M > INT32_MAX || V > INT32_MAX || ld < V || ld % 8 || ldd < V || ldd % 8)
    return fail("lora_ce_bwd: need M, V > 0 and ld, ldd >= V, multiples of 8");
  if (dtype != kDtypeBf16 && dtype != kDtypeFp16) return fail("dtype must be 0 (bf16) or 1 (fp16); fp32 is not supported");
  CUDA_TRY(launch_pdl(ce::ce_bwd_kernel, dim3((unsigned)M), dim3(ce::THREADS), 0, stream, 1,
                      static_cast<const uint16_t*>(logits), (long long)ld, reinterpret_cast<const long long*>(labels), lse,
                      grad_scale, static_cast<uint16_t*>(dlogits), (long long)ldd, (int)V, (long long)ignore_index, (int)dtype));
  return 0;
}

// ---- peer-memory exchange of the adapter-gradient bucket (p2p.cuh) ----
size_t lora_p2p_buffer_bytes(int64_t capacity_floats) {
  if (capacity_floats <= 0) return 0;
  const size_t cap = (static_cast<size_t>(capacity_floats) + 63) & ~size_t(63);
  return p2p::CTRL_BYTES + 4 * cap * sizeof(float);
}

int lora_p2p_alloc(int64_t capacity_floats, void** sym, void* ipc_handle_out) {
  g_err.clear();
  if (!sym || !ipc_handle_out || capacity_floats <= 0) return fail("lora_p2p_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == LORA_P2P_HANDLE_BYTES, "handle size");
  void* p = nullptr;
  const size_t bytes = lora_p2p_buffer_bytes(capacity_floats);
  CUDA_TRY(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);  // flags, arrival counters and the sequence counter start at zero
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(p);
    return fail("lora_p2p_alloc: %s", cudaGetErrorString(e));
  }
  std::memcpy(ipc_handle_out, &h, sizeof(h));
  *sym = p;
  return 0;
}

int lora_p2p_open(const void* ipc_handle, void** peer_sym) {
  g_err.clear();
  if (!ipc_handle || !peer_sym) return fail("lora_p2p_open: null argument");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, ipc_handle, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail("lora_p2p_open: cudaIpcOpenMemHandle: %s (peer GPUs must be P2P-capable, one process per GPU)", cudaGetErrorString(e));
  *peer_sym = p;
  return 0;
}

int lora_p2p_close(void* peer_sym) {
  g_err.clear();
  if (peer_sym) CUDA_TRY(cudaIpcCloseMemHandle(peer_sym));
  return 0;
}

int lora_p2p_free(void* sym) {
  g_err.clear();
  if (sym) CUDA_TRY(cudaFree(sym));
  return 0;
}

int lora_bucket_allreduce_avg(float* bucket, int64_t n, void* const* syms, int32_t rank, int32_t world,
                              int64_t capacity_floats, int32_t mode, cudaStream_t stream) {
  g_err.clear();
  if (!bucket || !syms) return fail("lora_bucket_allreduce_avg: null argument");
  if (world < 1 || world > p2p::MAX_WORLD || rank < 0 || rank >= world) return fail("lora_bucket_allreduce_avg: world must be 1..%d and 0 <= rank < world", p2p::MAX_WORLD);
  if (n <= 0 || n > capacity_floats) return fail("lora_bucket_allreduce_avg: %lld values do not fit the exchange buffers (%lld)", (long long)n, (long long)capacity_floats);
  if (reinterpret_cast<uintptr_t>(bucket) & 15) return fail("lora_bucket_allreduce_avg: bucket must be 16-byte aligned");
  if (mode < 0 || mode > 2) return fail("lora_bucket_allreduce_avg: mode is 0 (auto), 1 (one-shot) or 2 (two-shot)");
  if (world == 1) return 0;
  int sms = 0;
  if (device_setup(&sms)) return 1;
  p2p::Peers peers{};
  for (int q = 0; q < world; ++q) {
    if (!syms[q]) return fail("lora_bucket_allreduce_avg: syms[%d] is null", q);
    peers.sym[q] = syms[q];
  }
  const long long cap = (static_cast<long long>(capacity_floats) + 63) & ~63LL;
  int two_shot = mode == 2 || (mode == 0 && world > 2 && n * 4 > (1 << 20));
  if (n % 4) two_shot = 0;
  const long long work4 = two_shot ? (n / 4 + world - 1) / world : (n + 3) / 4;
  const int grid = static_cast<int>(std::max<long long>(1, std::min<long long>((work4 + p2p::THREADS - 1) / p2p::THREADS, p2p::MAX_CTAS)));
  CUDA_TRY(launch_pdl(p2p::bucket_allreduce_kernel, dim3(grid), dim3(p2p::THREADS), 0, stream, 1, bucket, (long long)n, peers,
                      (int)rank, (int)world, cap, two_shot));
  return 0;
}

int lora_adamw_clip_step(const void* table, int32_t n_tensors, int64_t n_chunks, float lr, float beta1, float beta2,
                         float eps, float weight_decay, float max_grad_norm, float* scratch, cudaStream_t stream) {
  g_err.clear();
  static_assert(sizeof(optim::TensorEntry) == 7 * sizeof(int64_t), "table row = seven 64-bit words");
  static_assert(optim::CHUNK == LORA_ADAMW_CHUNK, "chunk size");
  if (!table || !scratch) return fail("lora_adamw_clip_step: null pointer argument");
  if (n_tensors < 1 || n_chunks < n_tensors || n_chunks > INT32_MAX) return fail("lora_adamw_clip_step: bad tensor / chunk counts");
  if (!(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f)) return fail("lora_adamw_clip_step: betas must be in [0, 1)");
  const optim::TensorEntry* tab = static_cast<const optim::TensorEntry*>(table);
  CUDA_TRY(launch_pdl(optim::sumsq_kernel, dim3((unsigned)n_chunks), dim3(optim::THREADS), 0, stream, 1, tab, (int)n_tensors,
                      (long long)n_chunks, max_grad_norm, scratch));
  CUDA_TRY(launch_pdl(optim::adamw_kernel, dim3((unsigned)n_chunks), dim3(optim::THREADS), 0, stream, 1, tab, (int)n_tensors,
                      (const float*)scratch, lr, beta1, beta2, eps, weight_decay));
  return 0;
}

void lora_debug_set_phase_mask(int32_t mask) { g_phase_mask.store(mask & kPhaseAll, std::memory_order_relaxed); }

int lora_dropout_keep_mask(uint8_t* keep_mask, int64_t M, int64_t K, float p, uint64_t seed, uint64_t offset,
                           cudaStream_t stream) {
  g_err.clear();
  if (!keep_mask) return fail("lora_dropout_keep_mask: null pointer argument");
  if (M <= 0 || K <= 0 || K % 8 != 0 || M > INT32_MAX || K > INT32_MAX) return fail("lora_dropout_keep_mask: need M > 0, K > 0, K %% 8 == 0");
  if (!(p >= 0.f && p < 1.f)) return fail("dropout p must be in [0,1); got %f", (double)p);
  const long long total = static_cast<long long>(M) * (K / 8);
  const int grid = static_cast<int>(std::min<long long>((total + 255) / 256, 4096));
  keep_mask_kernel<<<grid, 256, 0, stream>>>(keep_mask, (int)M, (int)(K / 8), static_cast<uint32_t>((K + 15) / 16), seed,
                                             offset, static_cast<uint32_t>(p * 16384.0f + 0.5f));
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // extern "C"
