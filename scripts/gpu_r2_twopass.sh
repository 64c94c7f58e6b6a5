#!/bin/bash
# wide-rank masked backward: one launch of the TMEM flavour (LORA_B200_MASKED2=off) against the two-pass form
# (masked rank-r term, then the 256-column kernel with the residual epilogue)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/tp_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/tp_pytest.log
for shape in "16384 2560 2560 64" "16384 2560 10240 64" "16384 10240 2560 64" "16384 2560 2560 32" "16384 768 2304 64 --conv1d" "16384 3072 768 64 --conv1d"; do
  for mode in off on; do
    LORA_B200_MASKED2=$mode timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/tp_probe.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$mode', d['shape'], {k:v for k,v in d.items() if 'p0.05' in k})" | tee -a gpurun_out/tp_probe.txt
  done
done
timeout 300 python scripts/kernel_probe.py 2>> gpurun_out/tp_probe.err | tee -a gpurun_out/tp_probe.txt
tail -5 gpurun_out/tp_probe.err
