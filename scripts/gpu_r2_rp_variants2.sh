#!/bin/bash
# row projection A/B by build variant (LORA_B200_LIB): ring depth 8 against the shipped 6
mkdir -p gpurun_out
P=vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
for lib in liblora_b200.so liblora_b200_st8.so; do
  for shape in "16384 2560 2560 8" "4096 2560 2560 8" "16384 2560 10240 8" "16384 768 2304 8 --conv1d" "4096 10240 2560 8"; do
    LORA_B200_LIB=$PWD/$P/$lib PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/rpv2.err | tee -a gpurun_out/rpv2.jsonl
  done
done
tail -3 gpurun_out/rpv2.err
