#!/bin/bash
# 8-GPU evidence run: multi-rank tests, layer scaling (20 and 200 steps), scale diagnostic, pipeline at 2/4/8 stages
mkdir -p gpurun_out
export LORA_B200_DIST_TIMEOUT_S=240
TR() { n=$1; port=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port "$@"; }
timeout -k 10 900 python -m pytest tests/test_gpu_p2p.py tests/test_gpu_pipeline.py -m gpu -q -v > gpurun_out/r2n8_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2n8_pytest.log
tail -9 gpurun_out/r2n8_pytest.log
# pipeline at 2 and 4 stages side by side on disjoint GPUs, then 8 stages
(CUDA_VISIBLE_DEVICES=0,1,2,3 timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 4 --workload opt27b_pp --micro 8,32,96 --steps 2 --warmup 1 > gpurun_out/r2n8_pp4.json 2> gpurun_out/r2n8_pp4.err; echo "pp4 rc=$?") &
(CUDA_VISIBLE_DEVICES=4,5 timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29722 bench.py --gpus 2 --workload opt27b_pp --micro 8,32,96 --steps 2 --warmup 1 > gpurun_out/r2n8_pp2.json 2> gpurun_out/r2n8_pp2.err; echo "pp2 rc=$?") &
wait
timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29723 bench.py --gpus 8 --workload opt27b_pp --micro 8,32,96 --steps 2 --warmup 1 > gpurun_out/r2n8_pp8.json 2> gpurun_out/r2n8_pp8.err; echo "pp8 rc=$?"
# layer step + fine-tune runs at N = 8 (the driver's line), then 200 steps, then the diagnostic
timeout -k 10 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29724 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2n8_layer20.json 2> gpurun_out/r2n8_layer20.err; echo "layer20 rc=$?"
timeout -k 10 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29725 bench.py --gpus 8 --steps 200 --warmup 10 --no-finetune > gpurun_out/r2n8_layer200.json 2> gpurun_out/r2n8_layer200.err; echo "layer200 rc=$?"
timeout -k 10 300 python bench.py --steps 200 --warmup 10 --no-finetune --no-cpu-baseline > gpurun_out/r2n8_layer200_n1.json 2> gpurun_out/r2n8_layer200_n1.err; echo "layer200 n1 rc=$?"
timeout -k 10 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29726 scripts/scale_diag.py > gpurun_out/r2n8_scale_diag.jsonl 2> gpurun_out/r2n8_scale_diag.err; echo "diag rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2n8_*.json')):
    try:
        d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    except Exception as e:
        print(f, 'NO LINE', e); continue
    print(f, round(d['ms_per_step'], 4), round(d['value']), d.get('exchange_check'), {k: round(v['value']) for k, v in (d.get('finetune') or {}).items()})
    if 'pipeline' in d:
        for r in d['pipeline']['runs']:
            print('   ', {k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items() if k in ('m', 'overlap', 'ms_per_step', 'tokens_per_s', 'bubble_ideal', 'bubble_measured')})
        print('   ', {k: v for k, v in d['pipeline'].items() if k != 'runs'})
PY
cat gpurun_out/r2n8_scale_diag.jsonl
