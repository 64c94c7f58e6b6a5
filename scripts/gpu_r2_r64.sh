#!/bin/bash
# r = 64 diagnosis: per-kernel probe at the OPT q_proj shape with rank 64, then an ncu --set full capture of the
# column projection (after the same command has run clean without ncu); also the rank-8 skinny probe.
mkdir -p gpurun_out
PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py 2>> gpurun_out/r64_probe.err | tee gpurun_out/r64_probe_r8.json
timeout 300 python scripts/kernel_probe.py 16384 2560 2560 64 2>> gpurun_out/r64_probe.err | tee gpurun_out/r64_probe.json
timeout 300 python scripts/kernel_probe.py 16384 2560 10240 64 2>> gpurun_out/r64_probe.err | tee gpurun_out/r64_probe_fc1.json
PROBE_ONLY=once timeout 300 python scripts/kernel_probe.py 16384 2560 2560 64 > gpurun_out/r64_once.log 2>&1; echo "once rc=$?"
PROBE_ONLY=once timeout 600 ncu --set full --clock-control none --import-source on --kernel-name regex:colproj -c 2 \
  -o gpurun_out/r64_colproj -f python scripts/kernel_probe.py 16384 2560 2560 64 > gpurun_out/r64_ncu.log 2>&1; echo "ncu rc=$?"
