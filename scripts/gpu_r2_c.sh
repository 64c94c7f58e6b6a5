#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2c_pytest.log
tail -5 gpurun_out/r2c_pytest.log
timeout 300 python scripts/kernel_probe.py 2> gpurun_out/r2c_probe.err | tee gpurun_out/r2c_probe.json
timeout 600 python bench.py --steps 20 --warmup 5 --no-finetune > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2c_bench.json')); print(d['ms_per_step'], d['roofline']['step_tflops'], [ (k['kernel'][:24], round(k['kernel_ms']*1e3,1), round(k.get('frac',0),3)) for k in d['roofline']['kernels']], d['gpu_unfused'])"
