#!/bin/bash
# 2-GPU check of the multi-rank paths on the final kernels: multi-GPU tests, the DDP layer bench (with
# exchange_check), a short distilgpt2 DDP step and a 2-stage pipeline run
mkdir -p gpurun_out
export LORA_B200_DIST_TIMEOUT_S=180
timeout -k 10 600 python -m pytest tests/test_gpu_p2p.py tests/test_gpu_pipeline.py -m gpu -q > gpurun_out/r2n2_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2n2_pytest.log
tail -4 gpurun_out/r2n2_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713"
timeout -k 10 400 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2n2_layer.json 2> gpurun_out/r2n2_layer.err; echo "layer rc=$?"
python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2n2_layer.json') if l.startswith('{')][-1]); print(d['ms_per_step'], d['value'], d.get('exchange_check'), {k:(v.get('value'), v.get('ms_per_step')) for k,v in d.get('finetune',{}).items()})"
timeout -k 10 400 $TR bench.py --gpus 2 --workload opt27b_pp --micro 8 --steps 2 --warmup 1 > gpurun_out/r2n2_pp.json 2> gpurun_out/r2n2_pp.err; echo "pp rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2n2_pp.json') if l.startswith('{')][-1])
for r in d['pipeline']['runs']: print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items()})
PY
