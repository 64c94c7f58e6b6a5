#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2m_pytest.log
for shape in "16384 768 768 64 --conv1d" "16384 2560 2560 64" "16384 2560 10240 64" "16384 2560 2560 32" "16384 2560 2560 8"; do
  PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/r2m_probe.err | tee -a gpurun_out/r2m_probe.txt
done
