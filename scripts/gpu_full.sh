#!/bin/bash
# GPU box: parity diag (per function), model-level tests, bench, ncu launch list.
bash scripts/gpu_diag.sh
timeout 900 python -m pytest tests/test_gpu_model.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/model_tests.log 2>&1
echo "model tests exit $?"; tail -25 gpurun_out/model_tests.log | cut -c1-220
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; python -c "import json;d=json.load(open('gpurun_out/bench.json'));print('ms/step',d['ms_per_step'],'gemm_ms',d['roofline']['kernel_ms'],'TF',d['roofline']['achieved'],'step TF',d['roofline']['step_tflops'], d['clocks'])"; tail -3 gpurun_out/bench.err
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $SHORT > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
