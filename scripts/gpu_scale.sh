#!/bin/bash
# Multi-GPU box: 1 -> N scaling of the layer bench and the whole-model DDP workloads (bench.py contract launch).
NG=${1:-8}
run() {  # n, extra args, tag
  local n=$1; shift; local tag=$1; shift
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/scale_${tag}_n1.json 2> gpurun_out/scale_${tag}_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/scale_${tag}_n$n.json 2> gpurun_out/scale_${tag}_n$n.err
  fi
  echo "$tag n=$n exit $?"
  python -c "import json,sys;d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]);print('  ',d['value'],d['unit'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d.get('clocks'))" gpurun_out/scale_${tag}_n$n.json || tail -5 gpurun_out/scale_${tag}_n$n.err
}
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; nproc
for n in 1 $NG; do run $n layer; done
for n in 1 $NG; do run $n distilgpt2 --workload distilgpt2; done
for n in 1 $NG; do run $n opt27b --workload opt27b; done
if [ "$NG" -ge 8 ]; then for n in 2 4; do run $n layer; run $n distilgpt2 --workload distilgpt2; done; fi
