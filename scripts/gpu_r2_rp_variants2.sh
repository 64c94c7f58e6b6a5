#!/bin/bash
# row projection A/B: eight-warp CTAs with the producer inline for every rank (build variant) against the shipped nine-warp form
mkdir -p gpurun_out
P=vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
for lib in liblora_b200.so liblora_b200_inl2.so; do
  for shape in "16384 2560 2560 8" "16384 2560 2560 16" "16384 768 2304 8 --conv1d"; do
    LORA_B200_LIB=$PWD/$P/$lib PROBE_ONLY=skinny timeout 300 python scripts/kernel_probe.py $shape 2>> gpurun_out/rpv2.err | tee -a gpurun_out/rpv2.jsonl
  done
done
tail -3 gpurun_out/rpv2.err
