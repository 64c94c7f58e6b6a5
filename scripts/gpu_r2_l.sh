#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2l_pytest.log
for shape in "16384 768 768 64 --conv1d" "16384 2560 2560 64" "16384 2560 10240 64"; do
  PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/r2l_probe.err | tee -a gpurun_out/r2l_probe.txt
done
timeout 900 python bench.py --steps 20 --warmup 5 --no-finetune > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2l_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r2l_bench.json')); print(d['ms_per_step'], d['e2e'])"
