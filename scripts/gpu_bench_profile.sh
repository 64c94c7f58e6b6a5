#!/bin/bash
# GPU box: full gpu test-suite, the bench line, then (only if the plain runs exited 0) the ncu
# launch list and one full capture of the dominant kernel.  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph"
timeout 300 $SHORT > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 $SHORT > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_lora -s 4 -c 2 \
    -o gpurun_out/prof_gemm -f $SHORT > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la gpurun_out
