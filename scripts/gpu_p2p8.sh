#!/bin/bash
# 8-GPU box: exchange test at world 8, layer bench and model workloads with the in-stream exchange vs NCCL.
source <(sed -n '/^run() {/,/^}/p' scripts/gpu_p2p.sh)
timeout 600 python -m pytest tests/test_gpu_p2p.py -x -q -m gpu -p no:cacheprovider 2>&1 | tail -15
run 1 layer
run 8 layer_p2p --exchange p2p
run 8 layer_nccl --exchange nccl
run 4 layer_p2p --exchange p2p
run 2 layer_p2p --exchange p2p
run 8 distil_p2p --workload distilgpt2 --exchange p2p
run 8 opt_p2p --workload opt27b --exchange p2p
