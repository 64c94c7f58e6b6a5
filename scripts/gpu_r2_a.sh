#!/bin/bash
# round-2 GPU run A: tests, kernel probe (shipped build vs epilogue-first warp roles), bench line
mkdir -p gpurun_out
PKG=vt-cluster--parallel-and-distributed-programming-of-machine-learning-models_b200
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest.log
tail -15 gpurun_out/r2a_pytest.log
timeout 300 python scripts/kernel_probe.py > gpurun_out/r2a_probe_base.json 2> gpurun_out/r2a_probe_base.err; echo "probe rc=$?"
LORA_B200_LIB=$PWD/$PKG/liblora_b200_epifirst.so timeout 300 python scripts/kernel_probe.py > gpurun_out/r2a_probe_epifirst.json 2> gpurun_out/r2a_probe_epifirst.err; echo "probe2 rc=$?"
cat gpurun_out/r2a_probe_base.json gpurun_out/r2a_probe_epifirst.json
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2a_bench.err
head -c 3000 gpurun_out/r2a_bench.json
