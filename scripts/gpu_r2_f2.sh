#!/bin/bash
# 2-GPU validation: multi-GPU tests, layer bench with the overlapped exchange + exchange_check, pipeline workload
mkdir -p gpurun_out
N=${1:-2}
timeout 1200 python -m pytest tests/test_gpu_p2p.py tests/test_gpu_pipeline.py tests/test_gpu_parity.py::test_operands_on_another_device_are_refused -m gpu -q > gpurun_out/r2f_pytest_n$N.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2f_pytest_n$N.log
tail -6 gpurun_out/r2f_pytest_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-finetune > gpurun_out/r2f_layer_n$N.json 2> gpurun_out/r2f_layer_n$N.err; echo "layer rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2f_layer_n$N.json') if l.startswith('{')][-1])
print(d['ms_per_step'], d['value'], d.get('exchange_check'), d['run'])
PY
timeout 900 $TR bench.py --gpus $N --workload opt27b_pp --micro 8,16 --steps 3 --warmup 1 > gpurun_out/r2f_pp_n$N.json 2> gpurun_out/r2f_pp_n$N.err; echo "pp rc=$?"
tail -2 gpurun_out/r2f_pp_n$N.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2f_pp_n$N.json') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'])
for r in d['pipeline']['runs']: print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items()})
print({k:v for k,v in d['pipeline'].items() if k!='runs'})
PY
