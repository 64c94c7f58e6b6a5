#!/bin/bash
# quick A/B: parity of the changed kernels, the frozen-projection tests, skinny probes at the headline shape and at small M
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_frozen.py tests/test_zz_no_alloc.py tests/test_gpu_model.py -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/q_pytest.log
for shape in "16384 2560 2560 8" "4096 2560 2560 8" "4096 10240 2560 8" "2048 2560 2560 8" "2048 10240 2560 8" "2048 2560 10240 64"; do
  PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/q_probe.err | tee -a gpurun_out/q_probe.jsonl
done
tail -5 gpurun_out/q_probe.err
