/*
 * lora_b200.h -- C ABI of the B200 (sm_100a) fused LoRA-linear library.
 *
 * The reference (SalahElshafey/VT-Cluster...) has no native boundary of its
 * own: its LoRA arithmetic is third-party peft `lora.Linear` running on torch
 * CPU, reached through
 *     LoraConfig(...) / get_peft_model(...)
 *         Cluster/Project 1 - Fine Tuning Distilgpt2/finetune_lora_distilgpt2.py:98-104
 *         Cluster/Project 2 - Course Project/finetune_lora_opt_pp.py:104-112
 * Each entry point below names the peft/torch step it replaces.  All pointers
 * are DEVICE pointers unless stated; every buffer (workspace included) is owned
 * by the caller; calls are asynchronous on `stream`; nothing is retained after
 * return.  Functions return 0 on success, non-zero on error with a
 * thread-local message available from lora_last_error().  There is no CPU
 * path and no fallback: unsupported input is an error.
 *
 * Shapes: X [M,K], Y [M,N], W [N,K] (nn.Linear) or [K,N] (GPT-2 Conv1D,
 * w_is_KN=1), bias [N] or NULL, A [r,K] fp32, B [N,r] fp32, H [M,r_pad],
 * keep_mask [M,K/8] bytes (bit e of byte (m,k/8) = element (m, 8*(k/8)+e)
 * survived dropout).  Activations, W and bias are bf16 (dtype 0) or fp16
 * (dtype 1); accumulation is fp32.  K and N must be multiples of 8,
 * 1 <= r <= 64.
 */
#ifndef LORA_B200_H_
#define LORA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_API_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#define LORA_B200_DTYPE_BF16 0
#define LORA_B200_DTYPE_FP16 1

/* Message for the last failing call made by this host thread ("" if none). */
const char* lora_last_error(void);

/* Rank padded to the tensor-core K granule the kernels use: 16, 32 or 64. */
int32_t lora_linear_rpad(int32_t r);

/*
 * Bytes of the packed-operand block of one adapted layer (see lora_linear_fwd): the fp32 adapter
 * masters A, B as 16-bit tensor-core operands in the four layouts the kernels read.  No
 * counterpart in the reference (peft casts nothing; torch's CPU GEMMs read fp32).
 */
size_t lora_linear_packed_bytes(int64_t K, int64_t N, int32_t r);

/* Bytes of caller-owned scratch that lora_linear_bwd needs (the forward needs none). */
size_t lora_linear_workspace_bytes(int64_t M, int64_t K, int64_t N, int32_t r);

/*
 * A [r,K], B [N,r] (fp32 masters) -> packed [lora_linear_packed_bytes], 256-byte aligned.
 * lora_linear_fwd does this itself when handed A and B; the separate entry point is for callers
 * that run a backward without a forward of their own, or that pack once per optimiser step and
 * pass A = B = NULL to the forward.  No counterpart in the reference.
 */
int lora_linear_pack(const float* A, const float* B, void* packed, size_t packed_bytes, int64_t K, int64_t N, int32_t r,
                     float scaling, int32_t dtype, cudaStream_t stream);

/*
 * Forward of one LoRA-adapted projection; replaces peft lora.Linear.forward
 * (base_layer(x) + lora_B(lora_A(lora_dropout(x))) * scaling) as invoked for
 * every module matched by target_modules (finetune_lora_distilgpt2.py:100,
 * finetune_lora_opt_pp.py:109):
 *     H = dropout_p(X) A^T          (written to H, 16-bit, saved for backward)
 *     Y = X W^T + bias + scaling * H B^T
 * p > 0 draws the dropout mask from Philox4x32-10 keyed by (seed, offset) (p is
 * resolved to 1/16384: an element is dropped with probability round(p*16384)/16384) and
 * writes the keep-bits to keep_mask (needed by backward); p == 0 ignores
 * keep_mask (may be NULL).
 * `packed` (caller-owned, lora_linear_packed_bytes, 256-byte aligned) receives the 16-bit operand
 * forms of A and B; hand the same block to lora_linear_bwd, which therefore differentiates with the
 * adapter values the forward saw (what autograd's saved tensors give the reference).  A = B = NULL:
 * `packed` already holds them (lora_linear_pack).  What the backward needs from the forward is
 * X itself, H, keep_mask and packed -- no masked copy of X is written.
 */
int lora_linear_fwd(const void* X, const void* W, const void* bias, const float* A, const float* B, void* Y, void* H,
                    uint8_t* keep_mask, int64_t M, int64_t K, int64_t N, int32_t r, float scaling, float p, uint64_t seed,
                    uint64_t offset, int32_t dtype, int32_t w_is_KN, void* packed, size_t packed_bytes,
                    cudaStream_t stream);

/*
 * lora_linear_fwd whose dropout offset also has a device-side part: the Philox offset used is
 * offset + *offset_dev (offset_dev: device pointer to one uint64, NULL = 0), read when the kernel runs.
 * A training step captured in a CUDA graph advances *offset_dev inside the graph and so draws a fresh
 * mask on every replay -- the role of torch's graph-safe generator state for torch's own dropout
 * (the reference's lora_dropout is torch.nn.Dropout inside peft's lora.Linear).  No other difference.
 */
int lora_linear_fwd_devrng(const void* X, const void* W, const void* bias, const float* A, const float* B, void* Y, void* H,
                           uint8_t* keep_mask, int64_t M, int64_t K, int64_t N, int32_t r, float scaling, float p,
                           uint64_t seed, uint64_t offset, const uint64_t* offset_dev, int32_t dtype, int32_t w_is_KN,
                           void* packed, size_t packed_bytes, cudaStream_t stream);

/*
 * Backward of the same projection; replaces what torch autograd derives from
 * peft's forward (no custom backward exists in the reference):
 *     dH = scaling * dY B                 dX = dY W + keep/(1-p) .* (dH A)
 *     dB = scaling * dY^T H               dA = dH^T dropout_p(X)
 * No dW, no dbias (base layer frozen, LoraConfig bias="none").  dX may be NULL
 * (input needs no gradient).  dA [r,K] and dB [N,r] are fp32 and are
 * overwritten, or added to when accumulate != 0.  X is the forward's input (the mask is re-applied
 * from keep_mask), `packed` the block the forward (or lora_linear_pack) filled, `workspace` scratch of
 * lora_linear_workspace_bytes that must not be shared by calls that may overlap on the GPU.
 */
int lora_linear_bwd(const void* dY, const void* X, const void* W, const void* packed, const void* H,
                    const uint8_t* keep_mask, void* dX, float* dA, float* dB, int64_t M, int64_t K, int64_t N,
                    int32_t r, float scaling, float p, int32_t dtype, int32_t w_is_KN, int32_t accumulate,
                    void* workspace, size_t workspace_bytes, cudaStream_t stream);

/*
 * Frozen projection alone, Y = X W^T + bias: the base_layer call of peft's
 * forward (torch addmm for Conv1D, F.linear for nn.Linear) and what a merged
 * or disabled adapter reduces to.  Same kernel with the low-rank K-block off.
 */
int lora_linear_base(const void* X, const void* W, const void* bias, void* Y, int64_t M, int64_t K, int64_t N,
                     int32_t dtype, int32_t w_is_KN, cudaStream_t stream);

/*
 * Dropout keep-bits alone for an [M,K] input (K % 8 == 0): the mask
 * lora_linear_fwd would draw for the same (p, seed, offset).  Test hook for
 * checking the device generator against oracle/philox.py.
 */
int lora_dropout_keep_mask(uint8_t* keep_mask, int64_t M, int64_t K, float p, uint64_t seed, uint64_t offset,
                           cudaStream_t stream);

/*
 * Adapter gradient dB = scaling * dY^T H without any memory handed in or asked for.
 * No counterpart in the reference.
 *
 * For callers that are in no position to allocate device memory or to size a workspace: for
 * example a plug-in hosted by a runtime whose own pool already owns the GPU's memory (a cudaMalloc
 * from inside the plug-in may fail, or fragment that pool), or a latency-sensitive loop that must
 * not hit the driver's allocator (cudaMalloc / cudaFree synchronise the device).  This entry point
 * calls no cudaMalloc, cudaFree or other allocating routine (see the note on the first call below).  The split partial sums that
 * lora_linear_bwd keeps in the caller's workspace live here in a block of
 * LORA_B200_STATIC_SCRATCH_BYTES that is compiled into the library (every GPU that loads the
 * library holds it until the process exits).  The caller still passes the inputs dY [M,N] and
 * H [M, lora_linear_rpad(r)] (as written by lora_linear_fwd) and the output dB [N,r] fp32
 * (overwritten, or added to when accumulate != 0).
 *
 *  - Shapes whose partial sums do not fit the block are rejected with an error message that
 *    states the need and the limit (roughly: row-splits * N * padded-rank * 4 bytes; all the
 *    reference's shapes at r <= 16 fit).  Use lora_linear_bwd for those.
 *  - No allocation after the first call on a device.  That first call does the library's one-time
 *    set-up for the device: kernel attributes (cudaFuncSetAttribute), the block's device address
 *    (one cudaGetSymbolAddress) and, under CUDA's lazy loading, the module load during which the
 *    driver reserves the block itself.  Make it at start-up, before any stream capture on that
 *    device and outside any region in which the allocator must stay untouched.
 *  - All calls on a device share the one block and it keeps whatever the last call left.  The
 *    only part that must start at zero -- a small table of arrival counters at its head -- is
 *    zero when the library loads and is left at zero by every completed call; the partial sums
 *    need no clearing (each call overwrites every one it later reads).  Two calls whose kernels
 *    overlap on the GPU (different streams, or different host threads on different streams)
 *    race on the block: keep calls on one stream, or order them with events.  A call that is
 *    aborted mid-kernel (device reset, trap) may leave counters non-zero: reload the library.
 *  - Result: the dB that lora_linear_bwd writes for the same dY and H, up to fp32 rounding
 *    (the two paths may cut the rows into a different number of partial sums).
 */
#define LORA_B200_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int lora_linear_grad_b_static(const void* dY, const void* H, float* dB, int64_t M, int64_t N, int32_t r,
                              float scaling, int32_t dtype, int32_t accumulate, cudaStream_t stream);

/*
 * Un-adapted (frozen, no LoRA) projection through the same tcgen05 kernel, with leading dimensions
 * so that outputs whose width is not a multiple of 8 can live in a padded buffer (SURVEY.md 8f row
 * N3).  Replaces the torch F.linear / addmm calls HF makes for modules the reference does not
 * target -- above all lm_head (768 x 50257 in distilgpt2: cuBLAS has no aligned kernel for that
 * width).  fwd: Y[M,N] = X[M,K] W^T + bias;  bwd: dX[M,K] = dY[M,N] W (no dW: frozen).
 * ld* are row strides in elements, multiples of 8; the weight's own row length (K for [N,K], N for
 * w_is_KN) must be a multiple of 8; N (resp. K) itself may be ragged.
 */
int lora_frozen_linear_fwd(const void* X, int64_t ldx, const void* W, const void* bias, void* Y, int64_t ldy, int64_t M,
                           int64_t K, int64_t N, int32_t dtype, int32_t w_is_KN, cudaStream_t stream);
int lora_frozen_linear_bwd(const void* dY, int64_t lddy, const void* W, void* dX, int64_t lddx, int64_t M, int64_t K,
                           int64_t N, int32_t dtype, int32_t w_is_KN, cudaStream_t stream);

/*
 * Frozen projection with a residual added in the GEMM epilogue:
 *     Y[M,N] = X[M,K] W^T + bias + R[M,N]
 * which is what the HF blocks compute as projection followed by `residual + hidden_states`
 * (GPT2Block, OPTDecoderLayer) whenever no dropout sits between the two (eval, or dropout = 0).
 * R has the activation dtype and row stride ldr (multiple of 8 elements, 16-byte aligned base); R may
 * be Y itself (in-place accumulation).  The backward of the projection is lora_frozen_linear_bwd; the
 * residual's gradient is dY itself.
 */
int lora_frozen_linear_add_fwd(const void* X, int64_t ldx, const void* W, const void* bias, const void* R, int64_t ldr,
                               void* Y, int64_t ldy, int64_t M, int64_t K, int64_t N, int32_t dtype, int32_t w_is_KN,
                               cudaStream_t stream);

/*
 * Frozen projection with its activation applied in the GEMM epilogue: the
 * un-adapted `mlp.c_fc` + gelu_new of GPT-2 blocks (c_fc is no LoRA target at
 * finetune_lora_distilgpt2.py:100; HF runs addmm followed by ~8 elementwise kernels of
 * NewGELUActivation) and, when fc1 is not targeted (BASELINE configs[3], q_proj/v_proj), OPT's
 * fc1 + relu.  Y = act(X W^T + bias) with act = LORA_ACT_RELU | LORA_ACT_GELU_NEW; Z (optional,
 * may be NULL) receives the pre-activation X W^T + bias, which gelu_new's backward needs.
 * lora_act_bwd: dZ = dY * act'(.) elementwise over n_elems (multiple of 8) contiguous elements;
 * the second operand is Z for gelu_new and Y (or Z) for relu.  dX then comes from
 * lora_frozen_linear_bwd(dZ, ...).
 */
#define LORA_ACT_RELU 1
#define LORA_ACT_GELU_NEW 2
int lora_frozen_linear_act_fwd(const void* X, int64_t ldx, const void* W, const void* bias, void* Y, int64_t ldy, void* Z,
                               int64_t ldz, int64_t M, int64_t K, int64_t N, int32_t act, int32_t dtype, int32_t w_is_KN,
                               cudaStream_t stream);
int lora_act_bwd(const void* dY, const void* ZorY, void* dZ, int64_t n_elems, int32_t act, int32_t dtype,
                 cudaStream_t stream);

/*
 * Optimiser step of the adapter parameters in two launches: global gradient-norm clip + AdamW over every adapter
 * tensor.  Replaces torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW.step as HF Trainer / DeepSpeed run them for
 * the reference (finetune_lora_distilgpt2.py:130-145: lr 5e-5, betas (0.9, 0.999), eps 1e-8, weight_decay 0,
 * max_grad_norm 1.0; finetune_lora_opt_pp.py:193-197), same formulas: coef = min(1, max_norm / (norm + 1e-6)),
 * g *= coef (stored back), p *= 1 - lr wd, m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2,
 * p -= lr / (1 - b1^step) * m / (sqrt(v) / sqrt(1 - b2^step) + eps).
 * `table` is a DEVICE array of n_tensors rows of seven 64-bit words {param ptr, grad ptr, exp_avg ptr, exp_avg_sq ptr,
 * element count, first chunk, ptr to the tensor's int32 step counter} (all tensors fp32, contiguous; first chunk =
 * sum over earlier rows of ceil(count / LORA_ADAMW_CHUNK)); n_chunks is that sum over all rows.  Step counters live
 * on the device, start at 0 and are advanced by the call, one per tensor as in torch.optim.AdamW (a tensor left out
 * of a call keeps its count).  `scratch`: 4 + n_chunks floats, zeroed once by the caller; after the call scratch[0]
 * holds the gradient norm before clipping, scratch[1] the clip factor.  max_grad_norm <= 0 disables clipping.
 */
#define LORA_ADAMW_CHUNK 4096
int lora_adamw_clip_step(const void* table, int32_t n_tensors, int64_t n_chunks, float lr, float beta1, float beta2,
                         float eps, float weight_decay, float max_grad_norm, float* scratch, cudaStream_t stream);

/*
 * The path's one exchange step under DDP: average the flat fp32 adapter-gradient bucket (dA|dB of every
 * adapted layer) over the GPUs of one NVLink/NVSwitch box, in the caller's stream, through peer memory.
 * Replaces the gloo/NCCL all-reduce torch DDP issues for the reference
 * (finetune_lora_distilgpt2.py:142, ddp_find_unused_parameters=False).  One process per GPU.
 *   lora_p2p_alloc   allocates this rank's exchange buffer (lora_p2p_buffer_bytes(capacity)), zeroes it and
 *                    returns its 64-byte CUDA IPC handle; the host exchanges the handles (any side channel)
 *   lora_p2p_open    maps a peer's buffer;  lora_p2p_close / lora_p2p_free undo the two
 *   lora_bucket_allreduce_avg  bucket[i] <- mean over ranks, in place.  syms[q] (host array, q < world) is
 *                    rank q's buffer as mapped in THIS process (syms[rank] = own).  Calls are collective (every
 *                    rank makes the same sequence of calls; the call counter lives in the buffers, on the
 *                    device) and stream-ordered, no host synchronisation.  Every rank sums in rank order, so all ranks
 *                    get the bit-identical result.  mode 0 picks one-shot (small buckets) or two-shot
 *                    (reduce my slice, push it to every peer) by size.
 * A rank that never calls trips the device watchdog on its peers (trap after 2^40 cycles, about nine minutes).
 */
#define LORA_P2P_HANDLE_BYTES 64
#define LORA_P2P_MAX_WORLD 8
size_t lora_p2p_buffer_bytes(int64_t capacity_floats);
int lora_p2p_alloc(int64_t capacity_floats, void** sym, void* ipc_handle_out);
int lora_p2p_open(const void* ipc_handle, void** peer_sym);
int lora_p2p_close(void* peer_sym);
int lora_p2p_free(void* sym);
int lora_bucket_allreduce_avg(float* bucket, int64_t n, void* const* syms, int32_t rank, int32_t world,
                              int64_t capacity_floats, int32_t mode, cudaStream_t stream);

/*
 * Merge / unmerge of one adapter into its frozen 16-bit weight, in place (peft's
 * lora.Linear.merge()/unmerge(), reached after training through merge_and_unload-style use of the adapters the
 * reference saves at finetune_lora_distilgpt2.py:167): W <- W + scale * B A, scale = +lora_alpha/r to merge,
 * -lora_alpha/r to unmerge.  W is [N, K], or [K, N] with w_is_KN; A [r, K], B [N, r] fp32.
 */
int lora_merge_weight(void* W, const float* A, const float* B, int64_t K, int64_t N, int32_t r, float scale, int32_t dtype,
                      int32_t w_is_KN, cudaStream_t stream);

/*
 * Token cross-entropy from 16-bit logits; replaces the fp32 logits copy +
 * log_softmax + nll_loss that the HF causal-LM loss runs on DataCollatorForLanguageModeling labels
 * (finetune_lora_distilgpt2.py:114) and nn.CrossEntropyLoss in finetune_lora_opt_pp.py:143-153.
 * labels[i] (int64) is the target of row i, ignore_index where there is none -- the causal shift is
 * applied to the labels by the caller, never to the logits.
 * fwd: lse[i] = log sum_v exp(logits[i,v]);  loss_rows[i] = lse[i] - logits[i, labels[i]] (0 if ignored).
 * bwd: dlogits[i,v] = (exp(logits[i,v] - lse[i]) - [v == labels[i]]) * *grad_scale (zero rows if ignored);
 *      grad_scale is a device scalar (upstream gradient / number of counted tokens).
 */
int lora_ce_fwd(const void* logits, int64_t ld, const int64_t* labels, float* lse, float* loss_rows, int64_t M, int64_t V,
                int64_t ignore_index, int32_t dtype, cudaStream_t stream);
int lora_ce_bwd(const void* logits, int64_t ld, const int64_t* labels, const float* lse, const float* grad_scale,
                void* dlogits, int64_t ldd, int64_t M, int64_t V, int64_t ignore_index, int32_t dtype, cudaStream_t stream);

/*
 * Profiling hook, no counterpart in the reference: restricts which kernels the
 * calls above launch (bit 0 adapter packing, bit 1 rank-r row projections
 * H/dH, bit 2 the tcgen05 GEMM, bit 3 dA/dB column projections + reduce;
 * default 15 = all).  Process-wide; bench.py uses it to time the dominant
 * kernel alone on data left in the workspace by a preceding full call.
 * Results of a masked call are NOT a valid forward/backward.
 */
void lora_debug_set_phase_mask(int32_t mask);

#ifdef __cplusplus
}
#endif
#endif /* LORA_B200_H_ */
