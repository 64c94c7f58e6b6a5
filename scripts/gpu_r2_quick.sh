#!/bin/bash
# quick A/B: parity of the changed kernels, the frozen-projection tests, skinny probes at the headline shape
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_frozen.py -x -q > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/q_pytest.log
for shape in "16384 2560 2560 8" "16384 2560 2560 16" "16384 768 2304 8 --conv1d"; do
  PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/q_probe.err | tee -a gpurun_out/q_probe.jsonl
done
tail -5 gpurun_out/q_probe.err
