#!/bin/bash
# Runs each GPU parity test function in its own process under a timeout, so that a trapped
# kernel (dead CUDA context) in one does not hide the results of the others.
# Usage (on the GPU box): bash scripts/gpu_diag.sh [extra pytest args]
mkdir -p gpurun_out
LOG=gpurun_out/diag.log
: > $LOG
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $LOG 2>&1
for t in test_keep_mask_bit_exact test_base_gemm test_fwd_bwd_parity test_step0_known_answers \
         test_merged_weight_equivalence test_accumulate_flag test_deterministic_replay \
         test_rejects_unsupported test_full_size_against_fp32_matmul; do
  echo "=== $t" >> $LOG
  timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "$t" --tb=short -p no:cacheprovider "$@" >> $LOG 2>&1
  echo "=== $t exit $?" >> $LOG
done
grep -E "^=== |passed|failed|error" $LOG | tail -40
