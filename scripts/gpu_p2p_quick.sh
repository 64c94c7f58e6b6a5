#!/bin/bash
# N-GPU box: bench.py contract launch for the layer workload and distilgpt2 (default exchange), short.
NG=${1:-2}; STEPS=${2:-10}
source <(sed -n '/^run() {/,/^}/p' scripts/gpu_p2p.sh)
run() {
  local n=$1; shift; local tag=$1; shift
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps $STEPS --warmup 3 --no-cpu-baseline "$@" > gpurun_out/q_${tag}_n1.json 2> gpurun_out/q_${tag}_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus $n --steps $STEPS --warmup 3 --no-cpu-baseline "$@" > gpurun_out/q_${tag}_n$n.json 2> gpurun_out/q_${tag}_n$n.err
  fi
  echo "$tag n=$n exit $?"
  python -c "import json,sys;d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]);print('  ',d['value'],d['unit'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],d['clocks'])" gpurun_out/q_${tag}_n$n.json || tail -5 gpurun_out/q_${tag}_n$n.err
}
run $NG layer
run $NG distil --workload distilgpt2
if [ -n "$3" ]; then run $NG opt --workload opt27b; fi
