#!/bin/bash
# A/B on the GPU box: parity diag (per-function processes), then bench with 2-CTA pairs vs single CTA.
bash scripts/gpu_diag.sh
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_2cta.json 2> gpurun_out/bench_2cta.err; echo "bench 2cta exit $?"
python -c "import json;d=json.load(open('gpurun_out/bench_2cta.json'));print('2cta ms/step',d['ms_per_step'],'gemm_ms',d['roofline']['kernel_ms'],'TF',d['roofline']['achieved'],'step TF',d['roofline']['step_tflops'])"
LORA_B200_GEMM=1cta timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_1cta.json 2> gpurun_out/bench_1cta.err; echo "bench 1cta exit $?"
python -c "import json;d=json.load(open('gpurun_out/bench_1cta.json'));print('1cta ms/step',d['ms_per_step'],'gemm_ms',d['roofline']['kernel_ms'],'TF',d['roofline']['achieved'],'step TF',d['roofline']['step_tflops'])"
tail -3 gpurun_out/bench_2cta.err
