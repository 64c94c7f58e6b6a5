#!/bin/bash
# GPU box: frozen lm_head / fused CE tests, then model-level tokens/s with and without them.
timeout 900 python -m pytest tests/test_gpu_frozen.py -q -m gpu --tb=short -p no:cacheprovider > gpurun_out/frozen_tests.log 2>&1
echo "frozen tests exit $?"; tail -30 gpurun_out/frozen_tests.log | cut -c1-250
for wl in distilgpt2 opt27b; do
  for flag in "" "--stock-head"; do
    timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline $flag > gpurun_out/model_${wl}${flag}.json 2> gpurun_out/model_${wl}${flag}.err
    echo "$wl $flag exit $?"; python -c "import json,sys;d=json.load(open(sys.argv[1]));print(d['value'],d['unit'],d['ms_per_step'],'e2e',d['e2e']['value'])" gpurun_out/model_${wl}${flag}.json; tail -2 gpurun_out/model_${wl}${flag}.err
  done
done
