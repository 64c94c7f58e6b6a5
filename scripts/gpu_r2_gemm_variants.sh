#!/bin/bash
# A/B of build variants (_build.build_variant): raster group height of the GEMM (-DLB_EXP_GROUP_M=4|8|32) and the
# column projections launched before the dX GEMM (a source edit at the time; both measured within noise)
mkdir -p gpurun_out
PKG=$PWD/vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
for v in "" _gm4 _gm8 _gm32 _cpfirst ""; do
  LORA_B200_LIB=$PKG/liblora_b200$v.so timeout 300 python scripts/kernel_probe.py 2>> gpurun_out/gv_probe.err | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print(d['lib'], {k: d[k] for k in ('fwd_gemm_us_p0.05', 'bwd_gemm_us_p0.05', 'fwd_all_us_p0.05', 'bwd_all_us_p0.05', 'step_us_p0.05')})" | tee -a gpurun_out/gv_probe.txt
done
