#!/bin/bash
# quick single-GPU check: all GPU tests, the per-kernel probe, the layer bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/chk_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/chk_pytest.log
timeout 300 python scripts/kernel_probe.py 2>> gpurun_out/chk_probe.err | tee gpurun_out/chk_probe.json
timeout 900 python bench.py --steps 20 --warmup 5 --no-finetune > gpurun_out/chk_bench.json 2> gpurun_out/chk_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/chk_bench.err
python -c "
import json; d=json.load(open('gpurun_out/chk_bench.json')); print(d['ms_per_step'], d['roofline']['step_tflops'], d['gpu_launches'], [(k['kernel'][:22], round(k['kernel_ms']*1e3,1)) for k in d['roofline']['kernels']])"
