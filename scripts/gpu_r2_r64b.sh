#!/bin/bash
# after a change to the rank-32/64 kernels: parity tests, then the per-kernel probe at rank 64 and 32
# (LORA_B200_COLPROJ=hmma: the warp-MMA column projection instead of the tensor-memory one, for A/B)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q > gpurun_out/r64b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r64b_pytest.log
timeout 300 python scripts/kernel_probe.py 16384 2560 2560 64 2>> gpurun_out/r64b_probe.err | tee gpurun_out/r64b_probe.json
timeout 300 python scripts/kernel_probe.py 16384 2560 10240 64 2>> gpurun_out/r64b_probe.err | tee gpurun_out/r64b_probe_fc1.json
timeout 300 python scripts/kernel_probe.py 16384 2560 2560 32 2>> gpurun_out/r64b_probe.err | tee gpurun_out/r64b_probe_r32.json
PROBE_ONLY=skinny LORA_B200_COLPROJ=hmma timeout 300 python scripts/kernel_probe.py 16384 2560 2560 64 2>> gpurun_out/r64b_probe.err | tee gpurun_out/r64b_probe_hmma.json
tail -5 gpurun_out/r64b_probe.err
PROBE_ONLY=once timeout 300 python scripts/kernel_probe.py 16384 2560 2560 64 > gpurun_out/r64b_once.log 2>&1; echo "once rc=$?"
PROBE_ONLY=once timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:colproj -c 2 \
  -o gpurun_out/r64b_colproj -f python scripts/kernel_probe.py 16384 2560 2560 64 > gpurun_out/r64b_ncu.log 2>&1; echo "ncu rc=$?"
