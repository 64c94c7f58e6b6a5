#!/bin/bash
# round-2 evidence run: tests, bench line, ncu --set full of the five kernels of a step, ncu launch list of the bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2h_pytest.log
tail -4 gpurun_out/r2h_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo "bench rc=$?"
PROBE_ONLY=once timeout 200 python scripts/kernel_probe.py > gpurun_out/r2h_plain1.log 2>&1 && \
PROBE_ONLY=once timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_lora|rowproj|colproj' -s 5 -c 5 -f -o gpurun_out/r02_kernels python scripts/kernel_probe.py > gpurun_out/r2h_ncu1.log 2>&1
echo "ncu full rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-finetune --no-cpu-baseline --no-graph > gpurun_out/r2h_plain2.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-finetune --no-cpu-baseline --no-graph > gpurun_out/r2h_ncu2.log 2>&1
echo "ncu launches rc=$?"
