#!/bin/bash
# round-2 GPU run B: rowproj timing variants, ncu --set full of the skinny kernels, tests
mkdir -p gpurun_out
PKG=$PWD/vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
for v in "" _r0 _r5 _mulhi _cta2 _st4; do
  LORA_B200_LIB=$PKG/liblora_b200$v.so PROBE_ONLY=skinny timeout 200 python scripts/kernel_probe.py 2>> gpurun_out/r2b_probe.err | tee -a gpurun_out/r2b_probe.jsonl
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2b_pytest.log
tail -8 gpurun_out/r2b_pytest.log
PROBE_ONLY=skinny timeout 200 python scripts/kernel_probe.py > gpurun_out/r2b_plain.log 2>&1 && \
PROBE_ONLY=skinny timeout 600 ncu --set full --clock-control none --import-source on -k regex:'rowproj|colproj' -c 3 -f -o gpurun_out/r2b_skinny python scripts/kernel_probe.py > gpurun_out/r2b_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2b_ncu.log
