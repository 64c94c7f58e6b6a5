#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2g_pytest.log
tail -4 gpurun_out/r2g_pytest.log
for wl in distilgpt2 opt27b; do
  for mode in "" "--head materialised" "--head materialised --no-plain-neighbours"; do
    timeout 600 python bench.py --workload $wl --steps 8 --warmup 3 $mode 2>> gpurun_out/r2g_model.err | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$wl', '$mode', round(d['ms_per_step'], 2), round(d['value']), d['clocks']['sm_mhz'], d['clocks']['reasons'])
" | tee -a gpurun_out/r2g_model.txt
  done
done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench.json')); print(d['ms_per_step'], d['roofline']['step_tflops'], [ (k['kernel'][:24], round(k['kernel_ms']*1e3,1), round(k.get('frac',0),3)) for k in d['roofline']['kernels']], d['gpu_unfused'], {k:(v['value'],v['ms_per_step']) for k,v in d['finetune'].items()}, d['e2e'])"
