#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/model_profile.py --workload distilgpt2 > gpurun_out/r2k_prof_distilgpt2.txt 2>&1; echo "prof1 rc=$?"
timeout 400 python scripts/model_profile.py --workload opt27b > gpurun_out/r2k_prof_opt27b.txt 2>&1; echo "prof2 rc=$?"
timeout 1400 python bench.py --sweep --steps 10 --warmup 3 > gpurun_out/r02_sweep.jsonl 2> gpurun_out/r02_sweep.err; echo "sweep rc=$?"; wc -l gpurun_out/r02_sweep.jsonl
