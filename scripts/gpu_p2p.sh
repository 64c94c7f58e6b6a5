#!/bin/bash
# >= 2-GPU box: peer-memory exchange test, then the layer bench and distilgpt2 with both exchange kinds.
NG=${1:-2}
timeout 600 python -m pytest tests/test_gpu_p2p.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -15
run() {
  local n=$1; shift; local tag=$1; shift
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/p2p_${tag}_n1.json 2> gpurun_out/p2p_${tag}_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/p2p_${tag}_n$n.json 2> gpurun_out/p2p_${tag}_n$n.err
  fi
  echo "$tag n=$n exit $?"
  python -c "import json,sys;d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]);print('  ',d['value'],d['unit'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'])" gpurun_out/p2p_${tag}_n$n.json || tail -5 gpurun_out/p2p_${tag}_n$n.err
}
run 1 layer
run $NG layer_p2p --exchange p2p
run $NG layer_nccl --exchange nccl
run 1 distil --workload distilgpt2
run $NG distil_p2p --workload distilgpt2 --exchange p2p
run $NG distil_nccl --workload distilgpt2 --exchange nccl
