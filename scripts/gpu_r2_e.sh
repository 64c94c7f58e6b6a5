#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2e_pytest.log
tail -4 gpurun_out/r2e_pytest.log
timeout 300 python scripts/kernel_probe.py 2> gpurun_out/r2e_probe.err | tee gpurun_out/r2e_probe.json
