#!/bin/bash
# quick A/B of the HBM-bound kernels at rank 64 / 32 (three shapes)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "fwd_bwd" > gpurun_out/r64c_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r64c_pytest.log
for shape in "16384 2560 2560 64" "16384 2560 10240 64" "16384 10240 2560 64" "16384 2560 2560 32" "16384 768 2304 64 --conv1d"; do
  PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/r64c_probe.err | tee -a gpurun_out/r64c_probe.jsonl
done
tail -5 gpurun_out/r64c_probe.err
