#!/bin/bash
mkdir -p gpurun_out
PKG=$PWD/vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
for v in "" _cp3 _cp3b _cps6; do
  LORA_B200_LIB=$PKG/liblora_b200$v.so PROBE_ONLY=skinny timeout 200 python scripts/kernel_probe.py 2>> gpurun_out/r2d_probe.err | tee -a gpurun_out/r2d_probe.jsonl
done
PROBE_ONLY=skinny timeout 200 python scripts/kernel_probe.py > gpurun_out/r2d_plain.log 2>&1 && \
PROBE_ONLY=skinny timeout 600 ncu --set full --clock-control none --import-source on -k regex:'rowproj|colproj' -c 3 -f -o gpurun_out/r2d_skinny python scripts/kernel_probe.py > gpurun_out/r2d_ncu.log 2>&1
echo "ncu rc=$?"
