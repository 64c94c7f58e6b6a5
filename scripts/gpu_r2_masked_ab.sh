#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2j_pytest.log
tail -3 gpurun_out/r2j_pytest.log
for shape in "16384 2560 2560 8" "16384 2560 2560 16" "16384 2560 2560 64" "16384 768 3072 8 --conv1d" "16384 2304 768 8 --conv1d" "16384 10240 2560 16" "4096 2560 2560 8"; do
  for mode in tmem epilogue; do
    LORA_B200_MASKED=$mode timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/r2j_probe.err | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$mode', d['shape'], {k: d[k] for k in ('fwd_gemm_us_p0.05', 'bwd_gemm_us_p0.05', 'bwd_gemm_us_p0.0', 'step_us_p0.05')})" | tee -a gpurun_out/r2j_probe.txt
  done
done
